// Offline stand-in for <tbb/concurrent_vector.h> (included by include/vstore/b_tree.h:14, unused).
#pragma once
#include <vector>
namespace tbb { template <typename T> using concurrent_vector = std::vector<T>; }
