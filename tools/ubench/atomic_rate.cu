// Rate of single-lane returning atomics from many warps to ONE address (the list counters of the day pipeline: every warp
// reserves its list entries with one atomicAdd on a global counter). Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
//   same1   : every warp of the grid, lane 0: atomicAdd(&ctr[0], n) with the result used
//   same3   : three counters in one 64-byte line, back to back (list_reserve3)
//   lines3  : three counters 256 bytes apart
//   spread  : every warp its own address (128 bytes apart)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void k(unsigned int* ctr, int mode, int iters, unsigned int* out) {
    const unsigned int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    unsigned int acc = 0;
    for (int i = 0; i < iters; ++i) {
        unsigned int r = 0;
        if (lane == 0) {
            if (mode == 0) r = atomicAdd(&ctr[0], 3u);
            else if (mode == 1) { r = atomicAdd(&ctr[0], 3u); r += atomicAdd(&ctr[1], 1u); r += atomicAdd(&ctr[2], 1u); }
            else if (mode == 2) { r = atomicAdd(&ctr[0], 3u); r += atomicAdd(&ctr[64], 1u); r += atomicAdd(&ctr[128], 1u); }
            else r = atomicAdd(&ctr[1024 + warp * 32], 3u);
        }
        r = __shfl_sync(0xffffffffu, r, 0);
        acc += r;
    }
    if (acc == 0x12345678u) out[0] = acc;
}

int main() {
    unsigned int *ctr, *out;
    CK(cudaMalloc(&ctr, 64 << 20)); CK(cudaMemset(ctr, 0, 64 << 20)); CK(cudaMalloc(&out, 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char* names[4] = { "same1", "same3", "lines3", "spread" };
    for (int grid : { 148, 1184 }) for (int mode = 0; mode < 4; ++mode) {
        const int iters = 200; float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0)); k<<<grid, 256>>>(ctr, mode, iters, out); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = best < ms ? best : ms;
        }
        const double ops = (double) grid * 8 * iters * (mode == 1 || mode == 2 ? 3 : 1);
        printf("grid %5d %-7s %9.1f us  %7.2f ns per atomic  %7.1f M atomics/s\n", grid, names[mode], best * 1e3, best * 1e6 / ops, ops / (best * 1e-3) * 1e-6);
    }
    return 0;
}
