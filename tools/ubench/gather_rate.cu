// Random-gather rate of HBM on this GPU: how many scattered 32-byte sectors per second, and what co-locating an agent's
// fields buys. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_rate gather_rate.cu ; run: ./gather_rate
//   soa6   : six separate arrays (16,16,4,4,20->32,1 bytes per element), one element of each per index  (the v5 layout of k_active)
//   aos64  : one 64-byte record per index (two sectors of one line)
//   sec32  : one 32-byte sector per index
// Indices: M random distinct-ish ids out of N (sorted ascending, as k_scan leaves the active list), N = 21.5 M.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void k_soa6(const unsigned int* __restrict__ idx, unsigned int m, const uint4* a, const uint4* b, const int* c, const int* d, const uint4* e, const unsigned char* f, unsigned int* out) {
    unsigned int acc = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const unsigned int p = idx[i];
        const uint4 va = a[p], vb = b[p], ve = e[2 * (size_t) p]; const int vc = c[p], vd = d[p]; const unsigned char vf = f[p];
        acc += va.x ^ vb.y ^ ve.z ^ (unsigned int) vc ^ (unsigned int) vd ^ vf;
    }
    if (acc == 0x12345678u) out[0] = acc;
}
__global__ void k_aos64(const unsigned int* __restrict__ idx, unsigned int m, const uint4* a, unsigned int* out) {
    unsigned int acc = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const size_t p = idx[i];
        const uint4 v0 = a[4 * p], v1 = a[4 * p + 1], v2 = a[4 * p + 2], v3 = a[4 * p + 3];
        acc += v0.x ^ v1.y ^ v2.z ^ v3.w;
    }
    if (acc == 0x12345678u) out[0] = acc;
}
__global__ void k_sec32(const unsigned int* __restrict__ idx, unsigned int m, const uint4* a, unsigned int* out) {
    unsigned int acc = 0;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const size_t p = idx[i];
        const uint4 v0 = a[2 * p], v1 = a[2 * p + 1];
        acc += v0.x ^ v1.y;
    }
    if (acc == 0x12345678u) out[0] = acc;
}

int main() {
    const unsigned int N = 21500000;
    uint4 *a, *b, *e, *big; int *c, *d; unsigned char* f; unsigned int *idx, *out, *flush;
    CK(cudaMalloc(&a, (size_t) N * 16)); CK(cudaMalloc(&b, (size_t) N * 16)); CK(cudaMalloc(&e, (size_t) N * 32));
    CK(cudaMalloc(&c, (size_t) N * 4)); CK(cudaMalloc(&d, (size_t) N * 4)); CK(cudaMalloc(&f, N)); CK(cudaMalloc(&big, (size_t) N * 64));
    CK(cudaMalloc(&out, 4)); CK(cudaMalloc(&flush, 512u << 20));
    CK(cudaMemset(a, 1, (size_t) N * 16)); CK(cudaMemset(b, 1, (size_t) N * 16)); CK(cudaMemset(e, 1, (size_t) N * 32)); CK(cudaMemset(c, 1, (size_t) N * 4));
    CK(cudaMemset(d, 1, (size_t) N * 4)); CK(cudaMemset(f, 1, N)); CK(cudaMemset(big, 1, (size_t) N * 64));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (unsigned int m : { 50000u, 500000u, 2000000u }) {
        std::vector<unsigned int> h(m);
        unsigned long long s = 88172645463325252ull;
        for (auto& x : h) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; x = (unsigned int) (s % N); }
        std::sort(h.begin(), h.end());
        CK(cudaMalloc(&idx, (size_t) m * 4)); CK(cudaMemcpy(idx, h.data(), (size_t) m * 4, cudaMemcpyHostToDevice));
        for (int which = 0; which < 3; ++which) {
            float best = 1e30f;
            for (int rep = 0; rep < 5; ++rep) {
                CK(cudaMemset(flush, rep, 512u << 20));
                CK(cudaEventRecord(e0));
                if (which == 0) k_soa6<<<1184, 256>>>(idx, m, a, b, c, d, e, f, out);
                else if (which == 1) k_aos64<<<1184, 256>>>(idx, m, big, out);
                else k_sec32<<<1184, 256>>>(idx, m, big, out);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = std::min(best, ms);
            }
            const double sectors = which == 0 ? 6.0 : which == 1 ? 2.0 : 1.0;
            printf("m=%8u %-6s %8.1f us  %6.2f ns/index  %6.1f G sectors/s\n", m, which == 0 ? "soa6" : which == 1 ? "aos64" : "sec32", best * 1e3, best * 1e6 / m, sectors * m / (best * 1e-3) * 1e-9);
        }
        CK(cudaFree(idx));
    }
    return 0;
}
