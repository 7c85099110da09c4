// Microbenchmark (development aid): how fast can an SM evaluate Philox4x32-10 blocks? Sets the floor for k_visit / k_trials.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o philox_rate philox_rate.cu
// Measured on B200 (round 2): 270-450 G blocks/s depending on warps per SM (8 -> 64) and chains per thread (1 -> 4), i.e. the
// 2.7 M blocks of one k_visit launch cost 6-10 us of pure arithmetic: 80-130 issue cycles per warp-block per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ uint4 philox(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned long long m0 = (unsigned long long) c.x * 0xD2511F53ull, m1 = (unsigned long long) c.z * 0xCD9E8D57ull;
        c = make_uint4((unsigned int) (m1 >> 32) ^ c.y ^ k.x, (unsigned int) m1, (unsigned int) (m0 >> 32) ^ c.w ^ k.y, (unsigned int) m0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
template <int ILP>
__global__ void k(unsigned int n_per_thread, uint2 key, unsigned int* out) {
    unsigned int acc = 0;
    const unsigned int tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (unsigned int i = 0; i < n_per_thread; i += ILP) {
        uint4 x[ILP];
#pragma unroll
        for (int a = 0; a < ILP; ++a) x[a] = philox(make_uint4(0u, tid * 131u + i + a, 0u, 77u), key);
#pragma unroll
        for (int a = 0; a < ILP; ++a) acc += x[a].x ^ x[a].y ^ x[a].z ^ x[a].w;
    }
    if (acc == 0x12345678u) out[0] = acc;
}
template <int ILP> void run(int bps, int tpb) {
    int sms = 148; unsigned int* d; cudaMalloc(&d, 4);
    const unsigned int per = 64;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<ILP><<<sms * bps, tpb>>>(per, make_uint2(1, 2), d);
    cudaEventRecord(e0);
    for (int r = 0; r < 10; ++r) k<ILP><<<sms * bps, tpb>>>(per, make_uint2(1, 2), d);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
    const double blocks = (double) sms * bps * tpb * per;
    printf("ILP %d  warps/SM %2d : %.3f ms for %.0f Philox blocks = %.2f G blocks/s  (2.7M blocks = %.2f us)\n", ILP, bps * tpb / 32, ms, blocks,
           blocks / ms / 1e6, 2.7e6 / (blocks / ms / 1e6) / 1e3);
    cudaFree(d);
}
int main() {
    for (int bps : {1, 2, 3, 4, 8}) { run<1>(bps, 256); run<2>(bps, 256); run<4>(bps, 256); }
    return 0;
}
