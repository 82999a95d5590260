// tools/register_cost.cu — would DMA straight from the tree's leaf blocks beat copying them into pinned staging buffers?
// The reference allocates every 64 KiB leaf separately (DramBlockAllocator::New, include/vstore/valen_buffer.h:330-335), so
// "straight from them" = one cudaHostRegister + one 64 KiB cudaMemcpyAsync per leaf. Measures both on N separately
// allocated 64 KiB blocks, against memcpy into one pinned buffer + one large copy (what the library does).
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char **argv) {
    const size_t n = argc > 1 ? atol(argv[1]) : 16384, ls = 65536;
    std::vector<char *> blk(n);
    for (size_t i = 0; i < n; ++i) {
        blk[i] = (char *) aligned_alloc(4096, ls + 4096 * (i % 3));      // not contiguous
        memset(blk[i], (int) i, ls);
    }
    char *dev, *pinned;
    CK(cudaMalloc(&dev, n * ls));
    CK(cudaHostAlloc(&pinned, n * ls, cudaHostAllocDefault));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    double t0 = now();
    for (size_t i = 0; i < n; ++i) memcpy(pinned + i * ls, blk[i], ls);
    double t1 = now();
    CK(cudaMemcpyAsync(dev, pinned, n * ls, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    double t2 = now();
    printf("staging (1 thread): memcpy %.3f s (%.1f GB/s) + one H2D %.3f s (%.1f GB/s)\n", t1 - t0, n * ls / (t1 - t0) / 1e9, t2 - t1,
           n * ls / (t2 - t1) / 1e9);
    t0 = now();
    for (size_t i = 0; i < n; ++i) CK(cudaMemcpyAsync(dev + i * ls, blk[i], ls, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    t1 = now();
    printf("pageable, one H2D per block: %.3f s (%.1f GB/s, %.1f us per block)\n", t1 - t0, n * ls / (t1 - t0) / 1e9, (t1 - t0) / n * 1e6);
    t0 = now();
    for (size_t i = 0; i < n; ++i) CK(cudaHostRegister(blk[i], ls, cudaHostRegisterDefault));
    t1 = now();
    printf("cudaHostRegister of %zu blocks: %.3f s (%.1f us per block = %.1f GB/s of leaf images)\n", n, t1 - t0, (t1 - t0) / n * 1e6,
           n * ls / (t1 - t0) / 1e9);
    for (int rep = 0; rep < 2; ++rep) {
        t0 = now();
        for (size_t i = 0; i < n; ++i) CK(cudaMemcpyAsync(dev + i * ls, blk[i], ls, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        t1 = now();
        printf("registered, one H2D per block: %.3f s (%.1f GB/s, %.1f us per block)\n", t1 - t0, n * ls / (t1 - t0) / 1e9, (t1 - t0) / n * 1e6);
    }
    t0 = now();
    for (size_t i = 0; i < n; ++i) CK(cudaHostUnregister(blk[i]));
    t1 = now();
    printf("cudaHostUnregister: %.3f s (%.1f us per block)\n", t1 - t0, (t1 - t0) / n * 1e6);
    return 0;
}
