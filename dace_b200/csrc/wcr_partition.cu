// Histogram / scatter-add into bin arrays that do NOT fit in shared memory
// (BASELINE config C4: 2^28 int32 keys -> 2^20 int32 bins), sm_100a.
//
// One global atomic per key - what the reference's generated WCR kernel does
// (reference dace/codegen/targets/cpu.py:965-1013, dace/runtime/include/dace/reduction.h:170-173)
// and what wcr_scatter.cu's global path does - is bound by L2 atomic throughput:
// 173 G keys/s measured on B200 for 2^20 uniform bins, while shared-memory atomics
// sustain > 1500 G keys/s, i.e. HBM speed (profiles/r1_ubench_split_sweep_and_atomics.txt).
// So: radix-partition the keys by their high bits into B = nbins / 32768 buckets,
// then histogram each bucket in a 128 KiB shared-memory table.
//
// ONE persistent cooperative kernel, one 1024-thread CTA per SM, keys processed in
// chunks of 2^23 so the partitioned copy (double-buffered, 2 x 48 MiB) stays in
// the 126 MB L2 instead of round-tripping through HBM:
//   phase A (all CTAs): 8192-key tiles; rank inside the tile by shared-memory
//            atomics with return, bucket-sort the tile in shared memory, reserve space
//            per bucket with one global atomic per (tile, bucket), coalesced copy-out;
//   grid.sync();
//   phase B (CTA = bucket x slice): stream the bucket slice with 128-bit L2 loads and
//            atomically add into the CTA's persistent shared-memory table.
// Tables are flushed once at the end with RED.global.  Keys that do not fit the
// bucket's region (skewed distributions) take a warp-aggregated global atomic.
// Integer results are bit-exact (integer addition is associative).
#include <mutex>
#include "common.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kPartThreads = 1024;
constexpr int kTileKeys = 8192;                 // 8 keys per thread
constexpr int kKeysPerThread = kTileKeys / kPartThreads;
constexpr int kBucketBits = 15;                 // 32768 bins per bucket = 128 KiB of 4-byte counters
constexpr int kBucketBins = 1 << kBucketBits;
constexpr int kMaxBuckets = 64;
constexpr size_t kChunkKeys = size_t(1) << 23;  // 32 MiB of keys per chunk
constexpr size_t kMaxChunks = 4096;

template <class V>
__device__ __forceinline__ void smem_add_v(V* p, V v) { atomicAdd(p, v); }

template <class V>
__global__ void __launch_bounds__(kPartThreads, 1)
scatter_partition_kernel(const unsigned* __restrict__ keys, size_t n, V* __restrict__ bins, unsigned nbins, V cval,
                         unsigned* buf, unsigned* gcur, size_t chunk, unsigned cap, int B, int S, int aligned) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    V* hist = reinterpret_cast<V*>(smem_raw);                                  // [kBucketBins]
    unsigned* stage = reinterpret_cast<unsigned*>(smem_raw + size_t(kBucketBins) * sizeof(V));   // [kTileKeys]
    __shared__ unsigned cnt[kMaxBuckets], off[kMaxBuckets], gbase[kMaxBuckets];
    __shared__ unsigned tile_total;

    cg::grid_group grid = cg::this_grid();
    const unsigned tid = threadIdx.x;
    const unsigned lane = tid & 31;
    for (unsigned i = tid; i < kBucketBins; i += kPartThreads) hist[i] = V(0);
    __syncthreads();

    const size_t nchunks = (n + chunk - 1) / chunk;
    for (size_t c = 0; c < nchunks; ++c) {
        // ------------------------------ phase A: partition ------------------------------
        const size_t chunk_lo = c * chunk;
        const size_t chunk_n = (n - chunk_lo < chunk) ? (n - chunk_lo) : chunk;
        const size_t ntiles = (chunk_n + kTileKeys - 1) / kTileKeys;
        unsigned* bufc = buf + size_t(c & 1) * size_t(B) * cap;
        unsigned* cur = gcur + c * size_t(B);
        const unsigned* ck = keys + chunk_lo;

        unsigned knext[kKeysPerThread];
        auto load_tile = [&](size_t t, unsigned (&k)[kKeysPerThread]) {
            const size_t base = t * kTileKeys;
            if (aligned && base + kTileKeys <= chunk_n) {
#pragma unroll
                for (int u = 0; u < kKeysPerThread / 4; ++u) {
                    Words<16> w = ldg_stream16(ck + base + (size_t(tid) + size_t(u) * kPartThreads) * 4);
#pragma unroll
                    for (int j = 0; j < 4; ++j) k[u * 4 + j] = w.w[j];
                }
            } else {
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) {
                    const size_t i = base + tid + size_t(j) * kPartThreads;
                    k[j] = i < chunk_n ? __ldg(ck + i) : 0xffffffffu;   // 0xffffffff is never a valid bin
                }
            }
        };
        size_t t = blockIdx.x;
        if (t < ntiles) load_tile(t, knext);
        for (; t < ntiles; t += gridDim.x) {
            unsigned k[kKeysPerThread], rank[kKeysPerThread];
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j) k[j] = knext[j];
            if (t + gridDim.x < ntiles) load_tile(t + gridDim.x, knext);   // software prefetch of the next tile
            if (tid < kMaxBuckets) cnt[tid] = 0;
            __syncthreads();
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j)
                rank[j] = k[j] < nbins ? atomicAdd(&cnt[k[j] >> kBucketBits], 1u) : 0u;
            __syncthreads();
            if (tid < 32) {
                // exclusive scan over <= 64 buckets + one global reservation per non-empty bucket
                const unsigned c0 = cnt[lane], c1 = cnt[lane + 32];
                unsigned i0 = c0, i1 = c1;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    unsigned a0 = __shfl_up_sync(kFullMask, i0, d), a1 = __shfl_up_sync(kFullMask, i1, d);
                    if (lane >= (unsigned)d) { i0 += a0; i1 += a1; }
                }
                const unsigned tot0 = __shfl_sync(kFullMask, i0, 31);
                off[lane] = i0 - c0;
                off[lane + 32] = tot0 + i1 - c1;
                if (lane == 31) tile_total = tot0 + i1;
                if (c0) gbase[lane] = atomicAdd(&cur[lane], c0);
                if (c1) gbase[lane + 32] = atomicAdd(&cur[lane + 32], c1);
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j)
                if (k[j] < nbins) stage[off[k[j] >> kBucketBits] + rank[j]] = k[j];
            __syncthreads();
            const unsigned total = tile_total;
            for (unsigned sb = 0; sb < total; sb += kPartThreads) {   // warp-uniform trip count
                const unsigned slot = sb + tid;
                const bool valid = slot < total;
                unsigned key = 0, dst = 0, b = 0;
                if (valid) {
                    key = stage[slot];
                    b = key >> kBucketBits;
                    dst = gbase[b] + (slot - off[b]);
                }
                const bool fits = valid && dst < cap;
                if (fits) bufc[size_t(b) * cap + dst] = key;
                const unsigned ovf = __ballot_sync(kFullMask, valid && !fits);
                if (ovf && valid && !fits) {
                    // bucket region full (skew): warp-aggregated atomic straight into the bins
                    const unsigned peers = __match_any_sync(ovf, key);
                    if (lane == unsigned(__ffs(peers) - 1)) atomicAdd(bins + key, cval * V(__popc(peers)));
                }
            }
            __syncthreads();
        }
        grid.sync();
        // ------------------------------ phase B: histogram ------------------------------
        if (blockIdx.x < unsigned(B * S)) {
            const unsigned b = blockIdx.x / S, s = blockIdx.x % S;
            unsigned cntb = __ldcg(cur + b);
            if (cntb > cap) cntb = cap;
            const unsigned lo = unsigned((size_t(cntb) * s / S)) & ~3u;
            const unsigned hi = (s == unsigned(S - 1)) ? cntb : (unsigned((size_t(cntb) * (s + 1) / S)) & ~3u);
            const unsigned* src = bufc + size_t(b) * cap;
            const unsigned nvec = (hi - lo) / 4;
            const uint4* v4 = reinterpret_cast<const uint4*>(src + lo);
            unsigned v = tid;
            for (; v + 3 * kPartThreads < nvec; v += 4 * kPartThreads) {
                uint4 w[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) w[u] = __ldcg(v4 + v + u * kPartThreads);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    smem_add_v(hist + (w[u].x & (kBucketBins - 1)), cval);
                    smem_add_v(hist + (w[u].y & (kBucketBins - 1)), cval);
                    smem_add_v(hist + (w[u].z & (kBucketBins - 1)), cval);
                    smem_add_v(hist + (w[u].w & (kBucketBins - 1)), cval);
                }
            }
            for (; v < nvec; v += kPartThreads) {
                uint4 w = __ldcg(v4 + v);
                smem_add_v(hist + (w.x & (kBucketBins - 1)), cval);
                smem_add_v(hist + (w.y & (kBucketBins - 1)), cval);
                smem_add_v(hist + (w.z & (kBucketBins - 1)), cval);
                smem_add_v(hist + (w.w & (kBucketBins - 1)), cval);
            }
            for (unsigned i = lo + nvec * 4 + tid; i < hi; i += kPartThreads)
                smem_add_v(hist + (__ldcg(src + i) & (kBucketBins - 1)), cval);
        }
        // no second grid-wide barrier: the next chunk writes the other buffer and its own counters
    }
    __syncthreads();
    if (blockIdx.x < unsigned(B * S)) {
        const unsigned b = blockIdx.x / S;
        for (unsigned i = tid; i < kBucketBins; i += kPartThreads) {
            const V h = hist[i];
            const unsigned bin = b * kBucketBins + i;
            if (h != V(0) && bin < nbins) atomicAdd(bins + bin, h);
        }
    }
}

// Per-device scratch: the double-buffered partitioned keys and the per-chunk
// bucket cursors.  Uses from different streams are serialised with an event.
struct PartScratch {
    unsigned* buf = nullptr;
    unsigned* gcur = nullptr;
    size_t buf_words = 0;
    cudaEvent_t last = nullptr;
    int coop = -1;
};
static std::mutex g_part_mutex;
static PartScratch g_part[64];

template <class V>
static cudaError_t launch_partitioned_t(const ScatterArgs& a, V cval, const Launch& l) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorNotSupported;
    const int B = int((a.nbins + kBucketBins - 1) / kBucketBins);
    const int sms = sm_count();
    if (B > kMaxBuckets || B > sms) return cudaErrorNotSupported;
    const int S = sms / B;
    const size_t chunk = kChunkKeys;
    const size_t nchunks = (a.n + chunk - 1) / chunk;
    if (nchunks > kMaxChunks) return cudaErrorNotSupported;
    // 1.5x the uniform share per bucket, multiple of 4 words so every bucket base stays 16-byte aligned
    const unsigned cap = unsigned(((chunk * 3 / 2) / B + 3) & ~size_t(3));
    const size_t smem = size_t(kBucketBins) * sizeof(V) + size_t(kTileKeys) * sizeof(unsigned);
    auto kern = scatter_partition_kernel<V>;

    std::lock_guard<std::mutex> lock(g_part_mutex);
    PartScratch& ps = g_part[dev];
    if (ps.coop < 0) {
        int coop = 0, occ = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (coop) {
            if (cudaFuncSetAttribute(scatter_partition_kernel<int>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
                cudaFuncSetAttribute(scatter_partition_kernel<unsigned>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
                cudaFuncSetAttribute(scatter_partition_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
                coop = 0;
            else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kPartThreads, smem) != cudaSuccess || occ < 1)
                coop = 0;
        }
        cudaGetLastError();
        ps.coop = coop;
    }
    if (!ps.coop) return cudaErrorNotSupported;
    const size_t buf_words = 2 * size_t(B) * cap;
    if (ps.buf == nullptr || ps.buf_words < buf_words) {
        if (ps.buf) { cudaFree(ps.buf); ps.buf = nullptr; }
        // 2 * 1.5 * chunk words (+ rounding slack for any B) = ~96 MiB
        const size_t alloc_words = 2 * (chunk * 3 / 2 + 4 * size_t(kMaxBuckets));
        e = cudaMalloc(&ps.buf, alloc_words * sizeof(unsigned));
        if (e != cudaSuccess) { ps.buf = nullptr; return e; }
        ps.buf_words = alloc_words;
    }
    if (ps.gcur == nullptr) {
        e = cudaMalloc(&ps.gcur, kMaxChunks * size_t(kMaxBuckets) * sizeof(unsigned));
        if (e != cudaSuccess) { ps.gcur = nullptr; return e; }
    }
    if (ps.last == nullptr) {
        e = cudaEventCreateWithFlags(&ps.last, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    } else {
        e = cudaStreamWaitEvent(l.stream, ps.last, 0);   // scratch is shared: order uses across streams
        if (e != cudaSuccess) return e;
    }
    e = cudaMemsetAsync(ps.gcur, 0, nchunks * size_t(B) * sizeof(unsigned), l.stream);
    if (e != cudaSuccess) return e;

    const unsigned* keys = reinterpret_cast<const unsigned*>(a.idx);
    size_t n = a.n;
    V* bins = reinterpret_cast<V*>(a.bins);
    unsigned nbins = unsigned(a.nbins);
    unsigned* buf = ps.buf;
    unsigned* gcur = ps.gcur;
    size_t chunk_arg = chunk;
    unsigned cap_arg = cap;
    int B_arg = B, S_arg = S;
    int aligned = (reinterpret_cast<size_t>(a.idx) & 15) == 0;
    void* args[] = {&keys, &n, &bins, &nbins, &cval, &buf, &gcur, &chunk_arg, &cap_arg, &B_arg, &S_arg, &aligned};
    e = cudaLaunchCooperativeKernel((void*)kern, dim3(sms), dim3(kPartThreads), args, smem, l.stream);
    if (e != cudaSuccess) { cudaGetLastError(); return cudaErrorNotSupported; }
    return cudaEventRecord(ps.last, l.stream);
}

// Returns cudaErrorNotSupported when the partitioned path does not apply (caller falls back
// to the global-atomic kernel - still a GPU path).
cudaError_t launch_scatter_partitioned(int dtidx, int dtval, const ScatterArgs& a, const Launch& l) {
    if (a.values != nullptr) return cudaErrorNotSupported;
    if (dtidx != DT_I32 && dtidx != DT_U32) return cudaErrorNotSupported;
    if (a.n < (size_t(1) << 22)) return cudaErrorNotSupported;
    if (a.nbins * 4 <= (size_t(96) << 10) || a.nbins > size_t(kMaxBuckets) * kBucketBins) return cudaErrorNotSupported;
    switch (dtval) {
        case DT_I32: { int c = 1; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<int>(a, c, l); }
        case DT_U32: { unsigned c = 1; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<unsigned>(a, c, l); }
        case DT_F32: { float c = 1.f; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<float>(a, c, l); }
        default: return cudaErrorNotSupported;
    }
}

}  // namespace db200
