// Offline stand-in for <tbb/parallel_sort.h> (included by include/vstore/b_tree.h:15, unused).
#pragma once
#include <algorithm>
namespace tbb { template <typename It, typename C> void parallel_sort(It a, It b, C c) { std::sort(a, b, c); } }
