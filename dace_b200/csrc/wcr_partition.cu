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
// chunks of ~2^24 (a whole number of tiles per group, so the grid barrier waits for nobody);
// the partitioned copy is double-buffered and mostly lives in the 126 MB L2:
//   phase A (4 independent 256-thread groups per CTA, named barriers), 2048-key tiles,
//            register-prefetched one tile ahead: rank inside (tile, bucket) by shared-memory
//            atomics with return, bucket-sort the tile in shared memory, reserve space per
//            bucket with one global atomic per (tile, bucket) - its L2 round trip overlaps
//            the sort - then a coalesced copy-out of ~256-byte runs;
//   grid.sync();
//   phase B (CTA = bucket x slice): stream the bucket slice with 128-bit L2 loads and
//            atomically add into the CTA's persistent shared-memory table.
// Tables are flushed once at the end with RED.global.  Keys that do not fit the
// bucket's region (skewed distributions) take a warp-aggregated global atomic.
// Integer results are bit-exact (integer addition is associative and commutative).
//
// Where the time goes (B200, C4, profiles/r1_c4_partition_notes.md): both phases are bound by
// the shared-memory pipeline - random 4-byte accesses hit ~3.5-way bank conflicts, and a key
// costs one ATOMS with return + one scattered STS in phase A and one ATOMS in phase B
// (~18 cycles per 32 keys per SM, ~0.5 ms for 2^28 keys) - not by HBM (0.15 ms) or L2.
// Tried and measured slower or equal: TMA bulk loads into an in-place ring, L2 prefetch,
// fixed-slot staging with a warp-per-bucket copy-out, carrying phase-B loads through phase A.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include "common.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kPartThreads = 1024;
constexpr int kGroups = 4;                       // independent quarters of the CTA in phase A (named barriers)
constexpr int kGroupThreads = kPartThreads / kGroups;
constexpr int kKeysPerThread = 8;
constexpr int kTileKeys = kGroupThreads * kKeysPerThread;   // 2048 keys per group tile
constexpr int kBucketBits = 15;                 // 32768 bins per bucket = 128 KiB of 4-byte counters
constexpr int kBucketBins = 1 << kBucketBits;
constexpr int kMaxBuckets = 64;
constexpr int kCntSlots = kMaxBuckets + 1;      // + the bucket that collects out-of-range keys
constexpr size_t kChunkKeysDefault = size_t(1) << 24;  // ~64 MiB of keys per chunk (measured best of 2^22..2^26)
constexpr size_t kMaxChunks = 4096;
constexpr int kDivertDefault = 0;             // 1 tile in N straight to global atomics (0 = none)
constexpr int kHintDefault = 0;               // L2 residency hints (bits 1 | 2, see below); DACE_B200_PART_HINT
constexpr int kPhaseBUnroll = 8;              // 128-bit L2 loads in flight per thread in phase B

template <class V>
__device__ __forceinline__ void smem_add_v(V* p, V v) { atomicAdd(p, v); }

// Bins may be wider than the shared-memory table: 64-bit bins (i64 / u64 / f64) are COUNTED in 32-bit tables
// (a call has fewer than 2^32 keys) and the counts are scaled by the constant increment when they are flushed.
template <class W>
__device__ __forceinline__ void bins_add(W* p, W v) {
    if constexpr (std::is_same<W, long long>::value)
        atomicAdd(reinterpret_cast<unsigned long long*>(p), static_cast<unsigned long long>(v));
    else
        atomicAdd(p, v);
}

// L2 residency control for the partition buffer (hint bits of the kernel, DACE_B200_PART_HINT):
//   1: the key stream is read with an evict-first policy (it is dead after one use);
//   2: the partitioned keys are stored with an evict-last policy (phase B reads them back a moment later);
// Measured (profiles/r2_c4_hint_sweep.txt): no change in time for chunks of 2^22 .. 2^24 keys (0.785 / 0.84 / 1.03 ms
// with or without) - the kernel is bound by the shared-memory pipeline, not by the buffer's trip through
// DRAM.  A third variant, `discard.global.L2` of every consumed 128-byte line in phase B (so that dirty lines
// never reach DRAM), was 47 % SLOWER (1.155 ms) and raced with the next use of the buffer at small chunks
// (wrong counts at 2^22): removed.
__device__ __forceinline__ unsigned long long policy_evict_first() {
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ unsigned long long policy_evict_last() {
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ Words<16> ldg_policy16(const void* p, unsigned long long pol) {
    Words<16> r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ void stg_policy(unsigned* p, unsigned v, unsigned long long pol) {
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(pol) : "memory");
}


// Barrier among the kGroupThreads threads of one group (ids 1..kGroups; 0 is __syncthreads / grid.sync).
__device__ __forceinline__ void group_sync(unsigned g) {
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kGroupThreads) : "memory");
}

// KW: 32-bit words per key (1: int32 / uint32 keys, 2: int64 keys - DaCe's default index type; the high word must
// be zero for a key to be in range, the partition itself moves 32-bit keys either way).
template <class V, class W, int KW>
__global__ void __launch_bounds__(kPartThreads, 1)
scatter_partition_kernel(const unsigned* __restrict__ keys, size_t n, W* __restrict__ bins, unsigned nbins, V cval, W scale,
                         unsigned* buf, unsigned* gcur, size_t chunk, unsigned cap, int B, int S, int aligned,
                         const int* __restrict__ gate, int divert, int hint, int bits) {
    // `bits`: log2 of the bins per bucket - kBucketBits for large bin arrays, fewer for mid-sized ones so that there
    // are still ~32 buckets to rank into (two buckets meant 16-way conflicts on the rank counters: 2^16 bins ran at
    // 206 Gkeys/s where 2^20 bins reach 306).
    const unsigned bucket_bins = 1u << bits, bucket_mask = bucket_bins - 1;
    if (*gate != 0) return;   // clustered / skewed keys: one of the kernels behind this one does the work (grid-uniform)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    V* hist = reinterpret_cast<V*>(smem_raw);                                  // [kBucketBins]
    unsigned* stage_all = reinterpret_cast<unsigned*>(smem_raw + size_t(kBucketBins) * sizeof(V));   // [kGroups][kTileKeys]
    __shared__ unsigned s_cnt[kGroups][kCntSlots], s_off[kGroups][kCntSlots], s_delta[kGroups][kCntSlots],
        s_gbase[kGroups][kCntSlots];
    __shared__ unsigned s_ovf[kGroups];

    cg::grid_group grid = cg::this_grid();
    const unsigned tid = threadIdx.x;
    const unsigned lane = tid & 31;
    const unsigned g = tid / kGroupThreads, gtid = tid % kGroupThreads;
    unsigned* stage = stage_all + g * kTileKeys;
    unsigned* cnt = s_cnt[g];
    unsigned* off = s_off[g];
    unsigned* delta = s_delta[g];
    unsigned* gbase = s_gbase[g];
    const unsigned uB = unsigned(B);

    for (unsigned i = tid; i < bucket_bins; i += kPartThreads) hist[i] = V(0);
    if (gtid < kCntSlots) cnt[gtid] = 0;
    __syncthreads();
    const unsigned long long pol_first = policy_evict_first(), pol_last = policy_evict_last();
    const bool keys_first = hint & 1, buf_last = hint & 2;

    const size_t nchunks = (n + chunk - 1) / chunk;
    for (size_t c = 0; c < nchunks; ++c) {
        // ------------------------------ phase A: partition ------------------------------
        const size_t chunk_lo = c * chunk;
        const size_t chunk_n = (n - chunk_lo < chunk) ? (n - chunk_lo) : chunk;
        const unsigned ntiles = unsigned((chunk_n + kTileKeys - 1) / kTileKeys);
        unsigned* bufc = buf + size_t(c & 1) * size_t(B) * cap;
        unsigned* cur = gcur + c * size_t(B);
        const unsigned* ck = keys + chunk_lo * KW;

        unsigned knext[kKeysPerThread];
        auto load_tile = [&](unsigned t, unsigned (&k)[kKeysPerThread]) {
            const size_t base = size_t(t) * kTileKeys;
            if (aligned && base + kTileKeys <= chunk_n) {
                constexpr int kPerLoad = 4 / KW;                        // keys per 16-byte load
#pragma unroll
                for (int u = 0; u < kKeysPerThread / kPerLoad; ++u) {
                    const unsigned* src = ck + (base + (size_t(gtid) + size_t(u) * kGroupThreads) * kPerLoad) * KW;
                    Words<16> w = keys_first ? ldg_policy16(src, pol_first) : ldg_stream16(src);
#pragma unroll
                    for (int j = 0; j < kPerLoad; ++j)
                        k[u * kPerLoad + j] = (KW == 1 || w.w[j * KW + KW - 1] == 0u) ? w.w[j * KW] : 0xffffffffu;
                }
            } else {
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) {
                    const size_t i = base + gtid + size_t(j) * kGroupThreads;
                    unsigned key = 0xffffffffu;                         // 0xffffffff is never a valid bin
                    if (i < chunk_n && (KW == 1 || __ldg(ck + i * KW + KW - 1) == 0u)) key = __ldg(ck + i * KW);
                    k[j] = key;
                }
            }
        };
        const unsigned tstride = gridDim.x * kGroups;
        unsigned t = blockIdx.x * kGroups + g;
        if (t < ntiles) load_tile(t, knext);
        for (unsigned it = g; t < ntiles; t += tstride, ++it) {
            unsigned k[kKeysPerThread], rank[kKeysPerThread];
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j) k[j] = knext[j];
            if (t + tstride < ntiles) load_tile(t + tstride, knext);   // software prefetch of the next tile
            if (divert && (it % unsigned(divert)) == 0) {
                // Hybrid: one tile in `divert` bypasses the partition and goes straight to the bins with
                // RED.global.  The L2 atomic units (~190 G keys/s) are idle otherwise, the shared-memory
                // pipeline that bounds the rest of this kernel is not touched, and the groups of a CTA
                // take their turn at different iterations (it starts at g), so the two paths overlap.
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j)
                    if (k[j] < nbins) bins_add(bins + k[j], W(cval) * scale);
                continue;
            }
            // rank inside (tile, bucket); keys >= nbins land in bucket index >= B and are dropped
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j)
                rank[j] = atomicAdd(&cnt[min(k[j]  >> bits, uB)], 1u);
            group_sync(g);
            unsigned c0 = 0, c1 = 0, o0 = 0, o1 = 0, gb0 = 0, gb1 = 0;
            if (gtid < 32) {
                // exclusive scan over the 64 bucket counters, one global reservation per non-empty bucket
                c0 = cnt[lane];
                c1 = cnt[lane + 32];
                cnt[lane] = 0;
                cnt[lane + 32] = 0;
                if (lane == 0) cnt[kMaxBuckets] = 0;
                unsigned i0 = c0, i1 = c1;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    unsigned a0 = __shfl_up_sync(kFullMask, i0, d), a1 = __shfl_up_sync(kFullMask, i1, d);
                    if (lane >= (unsigned)d) { i0 += a0; i1 += a1; }
                }
                const unsigned tot0 = __shfl_sync(kFullMask, i0, 31);
                o0 = i0 - c0;
                o1 = tot0 + i1 - c1;
                off[lane] = o0;
                off[lane + 32] = o1;
                if (lane == 31) off[kMaxBuckets] = tot0 + i1;
                // the L2 round trip of the reservations overlaps the staging below; results are used after it
                if (lane < uB && c0) gb0 = atomicAdd(&cur[lane], c0);
                if (lane + 32 < uB && c1) gb1 = atomicAdd(&cur[lane + 32], c1);
            }
            group_sync(g);
#pragma unroll
            for (int j = 0; j < kKeysPerThread; ++j)
                stage[off[min(k[j]  >> bits, uB)] + rank[j]] = k[j];
            if (gtid < 32) {
                gbase[lane] = gb0;
                gbase[lane + 32] = gb1;
                delta[lane] = lane * cap + gb0 - o0;
                delta[lane + 32] = (lane + 32) * cap + gb1 - o1;
                const bool over = __any_sync(kFullMask, (gb0 + c0 > cap) || (gb1 + c1 > cap));
                if (lane == 0) s_ovf[g] = over;
            }
            group_sync(g);
            const unsigned total = off[uB];   // valid keys of the tile (the out-of-range bucket sorts last)
            const unsigned ovf_tile = s_ovf[g];
            if (total == unsigned(kTileKeys) && !ovf_tile) {
                unsigned key[kKeysPerThread], d[kKeysPerThread];
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) key[j] = stage[j * kGroupThreads + gtid];
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) d[j] = delta[key[j]  >> bits];
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) {
                    unsigned* dstp = bufc + (j * kGroupThreads + gtid + d[j]);
                    if (buf_last) stg_policy(dstp, key[j], pol_last); else *dstp = key[j];
                }
            } else if (!ovf_tile) {
#pragma unroll
                for (int j = 0; j < kKeysPerThread; ++j) {
                    const unsigned slot = j * kGroupThreads + gtid;
                    if (slot < total) {
                        const unsigned key = stage[slot];
                        bufc[slot + delta[key  >> bits]] = key;
                    }
                }
            } else {
                // a bucket region is full (skewed keys): what does not fit goes straight to the bins
                for (unsigned sb = 0; sb < total; sb += kGroupThreads) {   // warp-uniform trip count
                    const unsigned slot = sb + gtid;
                    const bool valid = slot < total;
                    unsigned key = 0, dst = 0, b = 0;
                    if (valid) {
                        key = stage[slot];
                        b = key  >> bits;
                        dst = gbase[b] + (slot - off[b]);
                    }
                    const bool fits = valid && dst < cap;
                    if (fits) bufc[size_t(b) * cap + dst] = key;
                    const unsigned ovf = __ballot_sync(kFullMask, valid && !fits);
                    if (valid && !fits) {
                        const unsigned peers = __match_any_sync(ovf, key);
                        if (lane == unsigned(__ffs(peers) - 1) && key < nbins) bins_add(bins + key, W(cval * V(__popc(peers))) * scale);
                    }
                }
            }
            // no barrier here: the next tile touches cnt only (zeroed above) until its first group_sync
        }
        grid.sync();
        // ------------------------------ phase B: histogram ------------------------------
        if (blockIdx.x < unsigned(B * S)) {
            const unsigned b = blockIdx.x / S, s = blockIdx.x % S;
            unsigned cntb = __ldcg(cur + b);
            if (cntb > cap) cntb = cap;
            // slices start on 128-byte lines (bucket bases are 128-byte aligned: cap is a multiple of 32 words)
            const unsigned lo = unsigned((size_t(cntb) * s / S)) & ~31u;
            const unsigned hi = (s == unsigned(S - 1)) ? cntb : (unsigned((size_t(cntb) * (s + 1) / S)) & ~31u);
            const unsigned* src = bufc + size_t(b) * cap;
            const unsigned nvec = (hi - lo) / 4;
            const uint4* v4 = reinterpret_cast<const uint4*>(src + lo);
            unsigned v = tid;
            for (; v + (kPhaseBUnroll - 1) * kPartThreads < nvec; v += kPhaseBUnroll * kPartThreads) {
                uint4 w[kPhaseBUnroll];
#pragma unroll
                for (int u = 0; u < kPhaseBUnroll; ++u) w[u] = __ldcg(v4 + v + u * kPartThreads);
#pragma unroll
                for (int u = 0; u < kPhaseBUnroll; ++u) {
                    smem_add_v(hist + (w[u].x & bucket_mask), cval);
                    smem_add_v(hist + (w[u].y & bucket_mask), cval);
                    smem_add_v(hist + (w[u].z & bucket_mask), cval);
                    smem_add_v(hist + (w[u].w & bucket_mask), cval);
                }
            }
            for (; v < nvec; v += kPartThreads) {
                uint4 w = __ldcg(v4 + v);
                smem_add_v(hist + (w.x & bucket_mask), cval);
                smem_add_v(hist + (w.y & bucket_mask), cval);
                smem_add_v(hist + (w.z & bucket_mask), cval);
                smem_add_v(hist + (w.w & bucket_mask), cval);
            }
            for (unsigned i = lo + nvec * 4 + tid; i < hi; i += kPartThreads)
                smem_add_v(hist + (__ldcg(src + i) & bucket_mask), cval);
        }
        // no second grid-wide barrier: the next chunk writes the other buffer and its own counters
    }
    __syncthreads();
    if (blockIdx.x < unsigned(B * S)) {
        const unsigned b = blockIdx.x / S;
        for (unsigned i = tid; i < bucket_bins; i += kPartThreads) {
            const V h = hist[i];
            const unsigned bin = b * bucket_bins + i;
            if (h != V(0) && bin < nbins) bins_add(bins + bin, W(h) * scale);
        }
    }
}

// Key sampler (one CTA, same stream, ~3 us): decides ON THE DEVICE which of three kernels serves the call.
//   * 4096 keys spread over the input go into a shared-memory hash set; keys that were already there are
//     counted.  Uniform keys over nbins bins repeat ~S^2 / (2 nbins) times in S samples; many more repeats
//     = a few hot keys carry much of the mass (Zipf, all keys equal, keys confined to a small range)
//     -> flag 2: the hot-key cache kernel (wcr_scatter.cu), which absorbs them in shared memory.
//   * 1024 adjacent pairs: sorted input / long runs of one key (>= half of the pairs equal) are cheap for
//     warp-merged global atomics and unbalanced for the partition (one bucket gets a whole chunk)
//     -> flag 1: the global-atomic kernel.
//   * otherwise flag 0: the radix partition below.
constexpr int kSampleKeys = 4096;
constexpr int kSampleSlots = 8192;

__global__ void __launch_bounds__(1024) scatter_sample_kernel(const unsigned* __restrict__ keys, size_t n, unsigned nbins,
                                                              int* flag, int kw) {
    __shared__ unsigned table[kSampleSlots];
    __shared__ unsigned dups;
    for (int i = threadIdx.x; i < kSampleSlots; i += 1024) table[i] = 0xffffffffu;
    if (threadIdx.x == 0) dups = 0;
    __syncthreads();
    const size_t stride = (n - 1) / kSampleKeys;
    unsigned mine = 0;
    int equal = 0;
#pragma unroll
    for (int j = 0; j < kSampleKeys / 1024; ++j) {
        const size_t p = (size_t(j) * 1024 + threadIdx.x) * stride;
        const unsigned k = __ldg(keys + p * kw);
        const unsigned hi = kw == 2 ? __ldg(keys + p * kw + 1) : 0u;
        if (j == 0) equal = k == __ldg(keys + (p + 1) * kw) && (kw == 1 || hi == __ldg(keys + (p + 1) * kw + 1));
        if (k >= nbins || hi != 0u) continue;
        unsigned slot = (k * 2654435761u) >> 19;          // top 13 bits
        for (int probe = 0; probe < 16; ++probe) {
            const unsigned old = atomicCAS(&table[slot], 0xffffffffu, k);
            if (old == 0xffffffffu) break;
            if (old == k) { ++mine; break; }
            slot = (slot + 1) & (kSampleSlots - 1);
        }
    }
    if (mine) atomicAdd(&dups, mine);
    const int cnt = __syncthreads_count(equal);
    if (threadIdx.x == 0) {
        // repeats expected from uniform keys: S - nbins * (1 - exp(-S / nbins)) ~ S^2 / (2 nbins) for nbins >> S
        const float S = float(kSampleKeys), nb = float(nbins);
        const float expect = S - nb * (1.0f - __expf(-S / nb));
        int f = 0;
        if (float(dups) > 2.0f * expect + 128.0f) f = 2;
        else if (cnt >= 512) f = 1;
        *flag = f;
    }
}

// Geometry + scratch layout of one partitioned call.  The scratch is the caller's
// (dace_b200_wcr_scatter_workspace_bytes() / the `workspace` argument of dace_b200_wcr_scatter):
//     [ flag: 256 B ][ per-chunk bucket cursors: nchunks x B words ][ double buffer: 2 x B x cap words ]
struct PartPlan {
    int B, S;
    size_t chunk, nchunks;
    unsigned cap;
    size_t smem;
    size_t gcur_off, buf_off, total;
    int bits;
};

static bool part_plan(int dtidx, int dtval, bool has_values, size_t n, size_t nbins, PartPlan& p) {
    if (has_values) return false;
    if (dtidx != DT_I32 && dtidx != DT_U32 && dtidx != DT_I64) return false;
    if (dtval != DT_I32 && dtval != DT_U32 && dtval != DT_F32 && dtval != DT_I64 && dtval != DT_U64 && dtval != DT_F64)
        return false;
    if (n < (size_t(1) << 22)) return false;
    if (nbins * 4 <= (size_t(96) << 10) || nbins > size_t(kMaxBuckets) * kBucketBins) return false;
    const int sms = sm_count();
    // bins per bucket: 2^15 (what fits shared memory) for large arrays, otherwise whatever leaves 17 ... 32 buckets
    p.bits = kBucketBits;
    while (p.bits > 10 && ((nbins + (size_t(1) << (p.bits - 1)) - 1) >> (p.bits - 1)) <= 32) --p.bits;
    p.B = int((nbins + (size_t(1) << p.bits) - 1) >> p.bits);
    if (p.B > kMaxBuckets || p.B > sms) return false;
    p.S = sms / p.B;
    // every group of every CTA gets the same number of tiles per chunk (the grid barrier waits for the slowest)
    const size_t wave = size_t(sms) * kPartThreads * kKeysPerThread;
    p.chunk = wave * std::max<size_t>(1, (kChunkKeysDefault + wave / 2) / wave);
    const int lg = tuning().part_chunk_log2;   // tuning knob for experiments (read at load)
    if (lg >= 20 && lg <= 26) p.chunk = wave * std::max<size_t>(1, ((size_t(1) << lg) + wave / 2) / wave);
    p.nchunks = (n + p.chunk - 1) / p.chunk;
    if (p.nchunks > kMaxChunks) return false;
    // 1.5x the uniform share per bucket, multiple of 32 words so every bucket base stays 128-byte aligned
    p.cap = unsigned(((p.chunk * 3 / 2) / p.B + 31) & ~size_t(31));   // 128-byte lines: phase B discards whole lines
    p.smem = size_t(kBucketBins) * 4 + size_t(kGroups) * kTileKeys * sizeof(unsigned);
    p.gcur_off = 256;
    p.buf_off = (p.gcur_off + p.nchunks * size_t(p.B) * sizeof(unsigned) + 255) & ~size_t(255);
    p.total = p.buf_off + (2 * size_t(p.B) * p.cap + 8 * size_t(kMaxBuckets)) * sizeof(unsigned);
    return true;
}

// Per-element values: no partition, but the key sampler still picks between the global atomics and the hot-key
// cache (skewed keys), so the call wants the 256-byte flag.
static bool valued_selection_applies(int dtidx, int dtval, size_t n, size_t nbins) {
    if (dtidx != DT_I32 && dtidx != DT_U32) return false;
    if (dtval != DT_I32 && dtval != DT_U32 && dtval != DT_F32 && dtval != DT_I64 && dtval != DT_U64 && dtval != DT_F64)
        return false;
    return n >= (size_t(1) << 22) && nbins * dtype_size(dtval) > (size_t(96) << 10) && nbins <= 0xffffffffull;
}

size_t scatter_partition_bytes(int dtidx, int dtval, bool has_values, size_t n, size_t nbins) {
    PartPlan p;
    if (part_plan(dtidx, dtval, has_values, n, nbins, p)) return p.total;
    return (has_values && valued_selection_applies(dtidx, dtval, n, nbins)) ? 256 : 0;
}

// Only the sampler (flag at the start of the caller's scratch): for scatters with per-element values.
cudaError_t launch_scatter_sample(int dtidx, int dtval, const ScatterArgs& a, const Launch& l, const int** gate) {
    if (a.values == nullptr || !valued_selection_applies(dtidx, dtval, a.n, a.nbins)) return cudaErrorNotSupported;
    if (l.extra == nullptr || l.extra_bytes < 256 || (reinterpret_cast<size_t>(l.extra) & 15) != 0) return cudaErrorNotSupported;
    int* flag = reinterpret_cast<int*>(l.extra);
    scatter_sample_kernel<<<1, 1024, 0, l.stream>>>(reinterpret_cast<const unsigned*>(a.idx), a.n, unsigned(a.nbins), flag,
                                                    dtidx == DT_I64 ? 2 : 1);
    *gate = flag;
    return cudaGetLastError();
}

// Cooperative launch + 192 KiB of dynamic shared memory: checked once per device (-1 unknown, 0 no, 1 yes).
static int g_part_coop[64];
static bool g_part_coop_init = [] { for (int& v : g_part_coop) v = -1; return true; }();

static bool partition_supported(size_t smem) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
    if (g_part_coop[dev] < 0) {
        int coop = 0, occ = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        if (coop) {
            const int sm = (int)smem;
            const cudaFuncAttribute at = cudaFuncAttributeMaxDynamicSharedMemorySize;
            auto set = [&](auto kernel) { return cudaFuncSetAttribute(kernel, at, sm) == cudaSuccess; };
            if (!set(scatter_partition_kernel<int, int, 1>) || !set(scatter_partition_kernel<unsigned, unsigned, 1>) ||
                !set(scatter_partition_kernel<float, float, 1>) || !set(scatter_partition_kernel<unsigned, long long, 1>) ||
                !set(scatter_partition_kernel<unsigned, unsigned long long, 1>) || !set(scatter_partition_kernel<unsigned, double, 1>) ||
                !set(scatter_partition_kernel<int, int, 2>) || !set(scatter_partition_kernel<unsigned, unsigned, 2>) ||
                !set(scatter_partition_kernel<float, float, 2>) || !set(scatter_partition_kernel<unsigned, long long, 2>) ||
                !set(scatter_partition_kernel<unsigned, unsigned long long, 2>) || !set(scatter_partition_kernel<unsigned, double, 2>))
                coop = 0;
            else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter_partition_kernel<int, int, 2>, kPartThreads, smem) != cudaSuccess || occ < 1)
                coop = 0;
        }
        cudaGetLastError();
        g_part_coop[dev] = coop ? 1 : 0;
    }
    return g_part_coop[dev] == 1;
}

template <class V, class W, int KW>
static cudaError_t launch_partitioned_kw(const ScatterArgs& a, V cval, W scale, const PartPlan& p, const Launch& l,
                                         const int** gate_out) {
    unsigned char* base = static_cast<unsigned char*>(l.extra);
    int* flag = reinterpret_cast<int*>(base);
    unsigned* gcur = reinterpret_cast<unsigned*>(base + p.gcur_off);
    unsigned* buf = reinterpret_cast<unsigned*>(base + p.buf_off);
    cudaError_t e = cudaMemsetAsync(gcur, 0, p.nchunks * size_t(p.B) * sizeof(unsigned), l.stream);
    if (e != cudaSuccess) return e;
    const unsigned* keys = reinterpret_cast<const unsigned*>(a.idx);
    size_t n = a.n;
    scatter_sample_kernel<<<1, 1024, 0, l.stream>>>(keys, n, unsigned(a.nbins), flag, KW);
    const int* gate = flag;
    *gate_out = gate;
    W* bins = reinterpret_cast<W*>(a.bins);
    unsigned nbins = unsigned(a.nbins);
    size_t chunk_arg = p.chunk;
    unsigned cap_arg = p.cap;
    int B_arg = p.B, S_arg = p.S;
    int aligned = (reinterpret_cast<size_t>(a.idx) & 15) == 0;
    int divert = tuning().part_divert >= 0 ? tuning().part_divert : kDivertDefault;   // tuning knob (0 = off)
    int hint = tuning().part_hint >= 0 ? tuning().part_hint : kHintDefault;
    int bits_arg = p.bits;
    void* args[] = {&keys, &n, &bins, &nbins, &cval, &scale, &buf, &gcur, &chunk_arg, &cap_arg, &B_arg, &S_arg, &aligned, &gate, &divert, &hint, &bits_arg};
    e = cudaLaunchCooperativeKernel((void*)scatter_partition_kernel<V, W, KW>, dim3(sm_count()), dim3(kPartThreads), args,
                                    p.smem, l.stream);
    if (e != cudaSuccess) { cudaGetLastError(); return cudaErrorNotSupported; }
    return cudaSuccess;
}

template <class V, class W>
static cudaError_t launch_partitioned_t(const ScatterArgs& a, V cval, W scale, const PartPlan& p, const Launch& l,
                                        const int** gate_out, bool keys64) {
    return keys64 ? launch_partitioned_kw<V, W, 2>(a, cval, scale, p, l, gate_out)
                  : launch_partitioned_kw<V, W, 1>(a, cval, scale, p, l, gate_out);
}

// Returns cudaErrorNotSupported when the partitioned path does not apply or the caller's scratch is too
// small (caller falls back to the global-atomic kernel - still a GPU path).
cudaError_t launch_scatter_partitioned(int dtidx, int dtval, const ScatterArgs& a, const Launch& l, const int** gate) {
    PartPlan p;
    if (!part_plan(dtidx, dtval, a.values != nullptr, a.n, a.nbins, p)) return cudaErrorNotSupported;
    if (l.extra == nullptr || l.extra_bytes < p.total || (reinterpret_cast<size_t>(l.extra) & 15) != 0)
        return cudaErrorNotSupported;
    if (!partition_supported(p.smem)) return cudaErrorNotSupported;
    const bool k64 = dtidx == DT_I64;
    switch (dtval) {
        case DT_I32: { int c = 1; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<int, int>(a, c, 1, p, l, gate, k64); }
        case DT_U32: { unsigned c = 1; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<unsigned, unsigned>(a, c, 1u, p, l, gate, k64); }
        case DT_F32: { float c = 1.f; if (a.const_value) memcpy(&c, a.const_value, 4); return launch_partitioned_t<float, float>(a, c, 1.f, p, l, gate, k64); }
        // 64-bit bins: occurrences are counted in 32-bit tables, the counts scaled by the increment at the flush
        case DT_I64: { long long c = 1; if (a.const_value) memcpy(&c, a.const_value, 8); return launch_partitioned_t<unsigned, long long>(a, 1u, c, p, l, gate, k64); }
        case DT_U64: { unsigned long long c = 1; if (a.const_value) memcpy(&c, a.const_value, 8); return launch_partitioned_t<unsigned, unsigned long long>(a, 1u, c, p, l, gate, k64); }
        case DT_F64: { double c = 1.0; if (a.const_value) memcpy(&c, a.const_value, 8); return launch_partitioned_t<unsigned, double>(a, 1u, c, p, l, gate, k64); }
        default: return cudaErrorNotSupported;
    }
}

}  // namespace db200
