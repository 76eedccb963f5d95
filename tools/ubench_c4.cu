// Micro-benchmarks behind the C4 (2^28 int32 keys -> 2^20 int32 bins) design decisions of round 2
// (run on the GPU box; not part of the library):
//   1. DSMEM atomics: a cluster of 8 / 16 CTAs owns ONE privatised table spread over its CTAs' shared
//      memory (32768 bins = 128 KiB per CTA), every key is one red.shared::cluster.add.u32 to the
//      owning CTA - VERDICT r1 item 4(ii).  Measures keys/s against plain local shared atomics.
//   2. Ranking inside (tile, bucket) for the radix partition: shared-memory atomics with return
//      (round 1) vs warp ballots (5 VOTE + logic per key round, running counts in lane registers).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench_c4 tools/ubench_c4.cu
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__global__ void fill_keys(unsigned* p, size_t n, unsigned mask, unsigned seed) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t st = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += st) {
        unsigned x = unsigned(i) * 2654435761u ^ seed;
        x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
        p[i] = x & mask;
    }
}

__device__ __forceinline__ uint4 ldg16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

constexpr int kSliceBits = 15;                 // 32768 bins (128 KiB) per CTA
constexpr int kSliceBins = 1 << kSliceBits;

__device__ __forceinline__ void red_cluster(unsigned local_smem_addr, unsigned rank, unsigned v) {
    unsigned remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_smem_addr), "r"(rank));
    asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(remote), "r"(v) : "memory");
}

// mode 0: every key -> DSMEM atomic to the owning CTA (key >> 15 = rank in cluster);
// mode 1: same loads, local shared atomic only (key & 32767): the no-network baseline.
template <int U, int MODE>
__global__ void __launch_bounds__(1024, 1) dsmem_hist(const unsigned* __restrict__ keys, size_t n, unsigned* __restrict__ bins,
                                                       unsigned table_bins) {
    extern __shared__ __align__(16) unsigned char raw[];
    unsigned* h = reinterpret_cast<unsigned*>(raw);
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks(), crank = cluster.block_rank();
    for (int i = threadIdx.x; i < kSliceBins; i += blockDim.x) h[i] = 0;
    cluster.sync();
    const unsigned base = (unsigned)__cvta_generic_to_shared(h);
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x, nthr = size_t(gridDim.x) * blockDim.x;
    const size_t nvec = n / 4;
    for (size_t v = tid; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
        uint4 w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg16(keys + (v + size_t(u) * nthr) * 4);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned k[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (MODE == 0) red_cluster(base + (k[j] & (kSliceBins - 1)) * 4, (k[j] >> kSliceBits) & (csize - 1), 1u);
                else atomicAdd(h + (k[j] & (kSliceBins - 1)), 1u);
            }
        }
    }
    cluster.sync();
    const unsigned off = (MODE == 0 ? crank : 0) * kSliceBins;
    for (int i = threadIdx.x; i < kSliceBins; i += blockDim.x)
        if (h[i] && off + i < table_bins) atomicAdd(bins + off + i, h[i]);
}

// ---- ranking inside (tile, bucket), 32 buckets = key >> 15 of 2^20-bin keys ----
constexpr int kGroupThreads = 256, kKeysPerThread = 8;

// round 1: shared atomics with return on 33 counters per 256-thread group
__global__ void __launch_bounds__(1024, 1) rank_atomics(const unsigned* __restrict__ keys, size_t n, unsigned* __restrict__ out) {
    __shared__ unsigned cnt[4][33];
    const unsigned g = threadIdx.x / kGroupThreads, gt = threadIdx.x % kGroupThreads;
    if (gt < 33) cnt[g][gt] = 0;
    __syncthreads();
    unsigned acc = 0;
    const size_t ntiles = n / (kGroupThreads * kKeysPerThread);
    for (size_t t = size_t(blockIdx.x) * 4 + g; t < ntiles; t += size_t(gridDim.x) * 4) {
        const unsigned* base = keys + t * kGroupThreads * kKeysPerThread;
        unsigned k[kKeysPerThread];
#pragma unroll
        for (int u = 0; u < 2; ++u) { uint4 w = ldg16(base + (gt + u * kGroupThreads) * 4); k[u * 4] = w.x; k[u * 4 + 1] = w.y; k[u * 4 + 2] = w.z; k[u * 4 + 3] = w.w; }
#pragma unroll
        for (int j = 0; j < kKeysPerThread; ++j) acc += atomicAdd(&cnt[g][k[j] >> 15], 1u);
    }
    if (acc == 0xffffffffu) out[0] = acc;
}

// ballots: per key round 5 votes; lane b keeps the warp's running count of bucket b in a register
__global__ void __launch_bounds__(1024, 1) rank_ballot(const unsigned* __restrict__ keys, size_t n, unsigned* __restrict__ out) {
    const unsigned g = threadIdx.x / kGroupThreads, gt = threadIdx.x % kGroupThreads;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    unsigned acc = 0;
    const size_t ntiles = n / (kGroupThreads * kKeysPerThread);
    for (size_t t = size_t(blockIdx.x) * 4 + g; t < ntiles; t += size_t(gridDim.x) * 4) {
        const unsigned* base = keys + t * kGroupThreads * kKeysPerThread;
        unsigned k[kKeysPerThread];
#pragma unroll
        for (int u = 0; u < 2; ++u) { uint4 w = ldg16(base + (gt + u * kGroupThreads) * 4); k[u * 4] = w.x; k[u * 4 + 1] = w.y; k[u * 4 + 2] = w.z; k[u * 4 + 3] = w.w; }
        unsigned run = 0;   // lane b: keys of bucket b seen by this warp in this tile
#pragma unroll
        for (int j = 0; j < kKeysPerThread; ++j) {
            const unsigned b = k[j] >> 15;
            unsigned mine = 0xffffffffu, forlane = 0xffffffffu;
#pragma unroll
            for (int bit = 0; bit < 5; ++bit) {
                const unsigned v = __ballot_sync(0xffffffffu, (b >> bit) & 1u);
                mine &= ((b >> bit) & 1u) ? v : ~v;
                forlane &= ((lane >> bit) & 1u) ? v : ~v;
            }
            const unsigned before = __shfl_sync(0xffffffffu, run, b);     // running count of my bucket
            const unsigned rank = before + __popc(mine & lt);
            run += __popc(forlane);
            acc += rank;
        }
        acc += run;
    }
    if (acc == 0xffffffffu) out[0] = acc;
}

// match.any variant of the same
__global__ void __launch_bounds__(1024, 1) rank_match(const unsigned* __restrict__ keys, size_t n, unsigned* __restrict__ out) {
    const unsigned g = threadIdx.x / kGroupThreads, gt = threadIdx.x % kGroupThreads;
    const unsigned lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    unsigned acc = 0;
    const size_t ntiles = n / (kGroupThreads * kKeysPerThread);
    for (size_t t = size_t(blockIdx.x) * 4 + g; t < ntiles; t += size_t(gridDim.x) * 4) {
        const unsigned* base = keys + t * kGroupThreads * kKeysPerThread;
        unsigned k[kKeysPerThread];
#pragma unroll
        for (int u = 0; u < 2; ++u) { uint4 w = ldg16(base + (gt + u * kGroupThreads) * 4); k[u * 4] = w.x; k[u * 4 + 1] = w.y; k[u * 4 + 2] = w.z; k[u * 4 + 3] = w.w; }
#pragma unroll
        for (int j = 0; j < kKeysPerThread; ++j) {
            const unsigned peers = __match_any_sync(0xffffffffu, k[j] >> 15);
            acc += __popc(peers & lt);
        }
    }
    if (acc == 0xffffffffu) out[0] = acc;
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
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("  (error: %s)\n", cudaGetErrorString(e)); return -1.f; }
    return ms / reps;
}

template <int U, int MODE>
static float run_dsmem(const unsigned* keys, size_t n, unsigned* bins, unsigned table_bins, int cluster, int ctas) {
    auto kern = dsmem_hist<U, MODE>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSliceBins * 4);
    if (cluster > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = kSliceBins * 4;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int maxc = 0;
    cudaOccupancyMaxActiveClusters(&maxc, kern, &cfg);
    printf("  [cluster %d: max active clusters %d]\n", cluster, maxc);
    return time_kernel([&] { cudaLaunchKernelEx(&cfg, kern, keys, n, bins, table_bins); }, 5);
}

int main() {
    const size_t nk = size_t(1) << 28;
    unsigned* keys; unsigned* bins; unsigned* out;
    CK(cudaMalloc(&keys, nk * 4)); CK(cudaMalloc(&bins, (size_t(1) << 20) * 4)); CK(cudaMalloc(&out, 256));
    CK(cudaMemset(bins, 0, (size_t(1) << 20) * 4));
    printf("# DSMEM atomics, 2^28 uint32 keys, 32768 bins (128 KiB) per CTA, 1024 threads, 1 CTA/SM\n");
    for (int cluster : {1, 2, 4, 8, 16}) {
        const unsigned table = unsigned(cluster) * kSliceBins;
        fill_keys<<<148 * 8, 256>>>(keys, nk, table - 1, 11u);
        CK(cudaDeviceSynchronize());
        const int ctas = (148 / cluster) * cluster;
        float ms = run_dsmem<4, 0>(keys, nk, bins, table, cluster, ctas);
        printf("red.shared::cluster cluster=%2d table=%7u bins ctas=%3d : %.4f ms  %.1f Gkeys/s\n", cluster, table, ctas, ms, nk / ms * 1e-6);
        ms = run_dsmem<4, 1>(keys, nk, bins, table, cluster, ctas);
        printf("local ATOMS (same loads) cluster=%2d                ctas=%3d : %.4f ms  %.1f Gkeys/s\n", cluster, ctas, ms, nk / ms * 1e-6);
    }
    // verify the cluster-16 histogram once
    {
        const int cluster = 16;
        const unsigned table = cluster * kSliceBins;
        fill_keys<<<148 * 8, 256>>>(keys, nk, table - 1, 11u);
        CK(cudaMemset(bins, 0, (size_t(1) << 20) * 4));
        auto kern = dsmem_hist<4, 0>;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(128); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = kSliceBins * 4;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, kern, (const unsigned*)keys, nk, bins, table);
        CK(cudaDeviceSynchronize());
        unsigned* hb = new unsigned[table];
        CK(cudaMemcpy(hb, bins, table * 4, cudaMemcpyDeviceToHost));
        unsigned long long tot = 0;
        for (unsigned i = 0; i < table; ++i) tot += hb[i];
        printf("cluster-16 histogram checksum %llu (expect %zu)\n", tot, nk);
        delete[] hb;
    }
    printf("# ranking inside (tile, bucket): 2^28 keys, 32 buckets, 4 groups x 256 threads per CTA, 1 CTA/SM\n");
    fill_keys<<<148 * 8, 256>>>(keys, nk, (1u << 20) - 1, 11u);
    CK(cudaDeviceSynchronize());
    float ms = time_kernel([&] { rank_atomics<<<148, 1024>>>(keys, nk, out); }, 5);
    printf("rank by shared atomics with return : %.4f ms  %.1f Gkeys/s\n", ms, nk / ms * 1e-6);
    ms = time_kernel([&] { rank_ballot<<<148, 1024>>>(keys, nk, out); }, 5);
    printf("rank by 5 ballots + lane registers : %.4f ms  %.1f Gkeys/s\n", ms, nk / ms * 1e-6);
    ms = time_kernel([&] { rank_match<<<148, 1024>>>(keys, nk, out); }, 5);
    printf("rank by match.any                  : %.4f ms  %.1f Gkeys/s\n", ms, nk / ms * 1e-6);
    printf("done\n");
    return 0;
}
