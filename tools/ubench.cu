// Micro-benchmarks that decide kernel configurations (run on the GPU box, not part of the library):
//   1. split_reduce_kernel launch-geometry sweep on 2^30 f32 (C2);
//   2. atomic throughput: RED.global into an L2-resident 2^20-bin table vs ATOMS into a
//      per-CTA 32K-bin shared-memory table (decides whether a partition + shared-memory
//      histogram can beat one global RED per key for C4).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr \
//        -diag-suppress 128 -o tools/ubench tools/ubench.cu dace_b200/csrc/api.cu ... (see tools/Makefile)
#include <cstdio>
#include <vector>
#include "../dace_b200/csrc/reduce_full.cu"

using namespace db200;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__global__ void fill_uniform(float* p, size_t n, unsigned seed) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t st = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += st) {
        unsigned x = unsigned(i) * 2654435761u ^ seed;
        x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
        p[i] = (x >> 8) * (1.0f / 16777216.0f);
    }
}
__global__ void fill_keys(int* p, size_t n, unsigned mask, unsigned seed) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t st = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += st) {
        unsigned x = unsigned(i) * 2654435761u ^ seed;
        x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
        p[i] = int(x & mask);
    }
}

template <int VB, int U, int T>
static float time_split(const float* x, float* out, size_t n, unsigned ctas, unsigned cluster, Launch& l, int reps) {
    SplitArgs a{x, out, n, 1, 0, 0, ctas, cluster, nullptr, 0};
    float zero = 0.f;
    a.identity = &zero;
    using TT = Types<float, float, float>;
    for (int i = 0; i < 3; ++i) launch_split_cfg<TT, OpSum<float>, VB, U, T>(a, l);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, l.stream);
    for (int i = 0; i < reps; ++i) launch_split_cfg<TT, OpSum<float>, VB, U, T>(a, l);
    cudaEventRecord(e1, l.stream);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) return -1.f;
    return ms / reps;
}

// ---- atomics ----
template <int U>
__global__ void __launch_bounds__(512) red_global(const int* __restrict__ keys, size_t n, int* __restrict__ bins) {
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, nthr = size_t(gridDim.x) * blockDim.x;
    const size_t nvec = n / 4;
    for (size_t v = tid; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
        Words<16> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg_stream16(keys + (v + size_t(u) * nthr) * 4);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) atomicAdd(bins + w[u].w[j], 1);
    }
}

template <int U, class CT>
__global__ void __launch_bounds__(1024) atoms_smem(const int* __restrict__ keys, size_t n, int* __restrict__ bins, int nb) {
    extern __shared__ __align__(16) unsigned char raw[];
    CT* h = reinterpret_cast<CT*>(raw);
    for (int i = threadIdx.x; i < nb; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, nthr = size_t(gridDim.x) * blockDim.x;
    const size_t nvec = n / 4;
    for (size_t v = tid; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
        Words<16> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg_stream16(keys + (v + size_t(u) * nthr) * 4);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) atomicAdd(h + (w[u].w[j] & (nb - 1)), CT(1));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nb; i += blockDim.x) if (h[i]) atomicAdd(bins + i, int(h[i]));
}

template <class K>
static float time_kernel(K&& launch, int reps) {
    for (int i = 0; i < 2; ++i) launch();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) return -1.f;
    return ms / reps;
}

int main() {
    const size_t n = size_t(1) << 30;
    float* x; float* out; void* ws;
    CK(cudaMalloc(&x, n * 4)); CK(cudaMalloc(&out, 4)); CK(cudaMalloc(&ws, kWorkspaceBytes));
    CK(cudaMemset(ws, 0, kCounterBytes));
    fill_uniform<<<148 * 8, 256>>>(x, n, 7u);
    CK(cudaDeviceSynchronize());
    Launch l; l.stream = 0; l.ws.counters = (unsigned*)ws; l.ws.partials = (char*)ws + kCounterBytes;
    const int sms = 148;
    printf("# split_reduce sweep, 2^30 f32, GB/s = 4.294967296e9 / t\n");
#define RUN(VB, U, T, PER_SM, CL) { float ms = time_split<VB, U, T>(x, out, n, sms * PER_SM, CL, l, 20); \
    printf("split VB=%d U=%d T=%d ctas/SM=%d cluster=%d : %.4f ms  %.1f GB/s\n", VB, U, T, PER_SM, CL, ms, 4.294967296 / ms * 1e3); }
    RUN(32, 4, 512, 2, 2) RUN(32, 4, 512, 1, 1) RUN(32, 4, 512, 2, 1) RUN(32, 4, 512, 3, 1) RUN(32, 4, 512, 4, 2)
    RUN(32, 8, 512, 1, 1) RUN(32, 8, 512, 2, 2) RUN(32, 2, 512, 2, 2) RUN(32, 2, 512, 4, 2)
    RUN(32, 4, 1024, 1, 1) RUN(32, 4, 1024, 2, 2) RUN(32, 2, 1024, 2, 2) RUN(32, 8, 1024, 1, 1)
    RUN(32, 4, 256, 4, 2) RUN(32, 4, 256, 8, 2) RUN(32, 8, 256, 4, 2) RUN(32, 8, 256, 2, 2) RUN(32, 16, 256, 2, 2)
    RUN(16, 8, 512, 2, 2) RUN(16, 4, 512, 4, 2) RUN(16, 8, 1024, 2, 2) RUN(32, 16, 128, 4, 2) RUN(32, 8, 128, 8, 2)
    float h = 0; CK(cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost));
    printf("sum = %.1f (expect ~%.1f)\n", h, n * 0.5);

    // ---- atomics ----
    const size_t nk = size_t(1) << 28;
    int* keys = reinterpret_cast<int*>(x);
    int* bins; CK(cudaMalloc(&bins, (size_t(1) << 20) * 4));
    printf("# atomics, 2^28 int32 keys\n");
    for (int bits : {20, 15, 13, 10}) {
        fill_keys<<<148 * 8, 256>>>(keys, nk, (1u << bits) - 1, 11u);
        CK(cudaDeviceSynchronize());
        for (int per_sm : {2, 4}) {
            float ms = time_kernel([&] { red_global<4><<<148 * per_sm, 512>>>(keys, nk, bins); }, 5);
            printf("RED.global bins=2^%d ctas/SM=%d : %.4f ms  %.1f Gkeys/s\n", bits, per_sm, ms, nk / ms * 1e-6);
        }
    }
    fill_keys<<<148 * 8, 256>>>(keys, nk, (1u << 20) - 1, 11u);
    CK(cudaDeviceSynchronize());
    CK(cudaFuncSetAttribute(atoms_smem<4, int>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10));
    CK(cudaFuncSetAttribute(atoms_smem<8, int>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10));
    for (int nb : {1 << 15, 1 << 13, 1 << 10}) {
        for (int threads : {512, 1024}) {
            float ms = time_kernel([&] { atoms_smem<4, int><<<148, threads, nb * 4>>>(keys, nk, bins, nb); }, 5);
            printf("ATOMS.smem int32 bins/CTA=%d threads=%d ctas/SM=1 : %.4f ms  %.1f Gkeys/s\n", nb, threads, ms, nk / ms * 1e-6);
        }
        float ms = time_kernel([&] { atoms_smem<8, int><<<148, 1024, nb * 4>>>(keys, nk, bins, nb); }, 5);
        printf("ATOMS.smem int32 U=8 bins/CTA=%d threads=1024 : %.4f ms  %.1f Gkeys/s\n", nb, ms, nk / ms * 1e-6);
    }
    {
        int nb = 1 << 13;
        float ms = time_kernel([&] { atoms_smem<4, int><<<148 * 4, 512, nb * 4>>>(keys, nk, bins, nb); }, 5);
        printf("ATOMS.smem int32 bins/CTA=%d threads=512 ctas/SM=4 : %.4f ms  %.1f Gkeys/s\n", nb, ms, nk / ms * 1e-6);
    }
    printf("done\n");
    return 0;
}
