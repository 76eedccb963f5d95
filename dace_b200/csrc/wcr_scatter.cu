// Write-conflict-resolution scatter (scatter-add / histogram) for sm_100a:
//     bins[idx[i]] = op(bins[idx[i]], values ? values[i] : const_value)
//
// Replaces the code DaCe generates for a GPU map with a dynamic WCR memlet:
// one thread per element, 32-thread blocks, one global atomic per element via
// dace::wcr_fixed<...>::reduce_atomic (reference dace/codegen/targets/cpu.py:965-1013,
// reference dace/runtime/include/dace/reduction.h:167-340; generated kernel shape
// in SURVEY.md A.5), and the only privatisation primitive the reference has,
// SharedToGlobal1D::Accum (reference dace/runtime/include/dace/cuda/copy.cuh:789-811).
//
// Two paths:
//   * bins fit in shared memory (Sum): per-CTA privatised sub-histograms,
//     replicated per warp group for small bin counts, warp-aggregated updates
//     (match.any) when collisions inside a warp are likely, one flush of the
//     non-zero bins with global reductions (RED, no return value);
//   * large bin arrays with constant increments: radix partition + shared-memory
//     histogram in one cooperative kernel (wcr_partition.cu);
//   * otherwise: persistent grid, 128-bit index loads, U vectors in flight,
//     RED.global straight into the (L2-resident) bin array.
// Out-of-range indices are skipped.  Integer results are bit-exact by
// construction (integer addition is associative); float sums are atomics, so
// order- and run-dependent in the last ulps like the reference's own.
#include <cstdlib>
#include <cstring>
#include "common.cuh"
#include "kernels.h"

namespace db200 {

// ---- atomic "reduce into memory" for every (op, type) we support ----------
template <class V> __device__ __forceinline__ void cas_apply_product(V* addr, V val) {
    if constexpr (sizeof(V) == 4) {
        unsigned* a = reinterpret_cast<unsigned*>(addr);
        unsigned old = *a, assumed;
        do {
            assumed = old;
            V cur = *reinterpret_cast<V*>(&assumed);
            V nv = cur * val;
            old = atomicCAS(a, assumed, *reinterpret_cast<unsigned*>(&nv));
        } while (old != assumed);
    } else {
        unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
        unsigned long long old = *a, assumed;
        do {
            assumed = old;
            V cur = *reinterpret_cast<V*>(&assumed);
            V nv = cur * val;
            old = atomicCAS(a, assumed, *reinterpret_cast<unsigned long long*>(&nv));
        } while (old != assumed);
    }
}

template <int RED, class V>
__device__ __forceinline__ void atomic_reduce(V* addr, V val) {
    if constexpr (RED == RT_Sum) {
        if constexpr (std::is_same<V, long long>::value)
            atomicAdd(reinterpret_cast<unsigned long long*>(addr), static_cast<unsigned long long>(val));
        else
            atomicAdd(addr, val);
    } else if constexpr (RED == RT_Product) {
        cas_apply_product<V>(addr, val);
    } else if constexpr (RED == RT_Min || RED == RT_Max) {
        if constexpr (std::is_same<V, float>::value) {
            // NaN-free ordering trick: non-negative floats order like ints,
            // negative floats order inversely like unsigned ints.
            const bool nonneg = !(__float_as_uint(val) >> 31);
            if (RED == RT_Max) {
                if (nonneg) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(val));
                else atomicMin(reinterpret_cast<unsigned*>(addr), __float_as_uint(val));
            } else {
                if (nonneg) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(val));
                else atomicMax(reinterpret_cast<unsigned*>(addr), __float_as_uint(val));
            }
        } else if constexpr (std::is_same<V, double>::value) {
            const bool nonneg = !(static_cast<unsigned long long>(__double_as_longlong(val)) >> 63);
            if (RED == RT_Max) {
                if (nonneg) atomicMax(reinterpret_cast<long long*>(addr), __double_as_longlong(val));
                else atomicMin(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(val));
            } else {
                if (nonneg) atomicMin(reinterpret_cast<long long*>(addr), __double_as_longlong(val));
                else atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(val));
            }
        } else {
            if (RED == RT_Max) atomicMax(addr, val); else atomicMin(addr, val);
        }
    } else if constexpr (std::is_integral<V>::value) {
        using U = typename std::conditional<sizeof(V) == 4, unsigned, unsigned long long>::type;
        U* a = reinterpret_cast<U*>(addr);
        U u = static_cast<U>(val);
        if constexpr (RED == RT_BAnd) atomicAnd(a, u);
        else if constexpr (RED == RT_BOr) atomicOr(a, u);
        else atomicXor(a, u);
    }
}

template <class I> __device__ __forceinline__ long long idx_at(const Words<16>& w, int j) {
    return (long long)vec_elem<I, I, 16>(w, j);
}

// ---- global path ----------------------------------------------------------
template <int RED, class I, class V, bool HAS_VALUES, int U>
__global__ void __launch_bounds__(512)
scatter_global_kernel(const I* __restrict__ idx, const V* __restrict__ values, V cval, size_t n,
                      V* __restrict__ bins, size_t nbins, const int* __restrict__ gate, int gate_run_if) {
    if (gate != nullptr && (gate_run_if >= 0 ? *gate != gate_run_if : *gate == -gate_run_if)) return;   // another kernel took this call
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t nthr = size_t(gridDim.x) * blockDim.x;
    if constexpr (HAS_VALUES) {
        for (size_t i = tid; i < n; i += nthr) {
            long long k = (long long)ldg_scalar(idx + i);
            V v = ldg_scalar(values + i);
            if (k >= 0 && (size_t)k < nbins) atomic_reduce<RED, V>(bins + k, v);
        }
    } else {
        constexpr int N = 16 / int(sizeof(I));
        const size_t addr = reinterpret_cast<size_t>(idx);
        size_t h = ((16 - (addr & 15)) & 15) / sizeof(I);
        if (h > n) h = n;
        const char* pv = reinterpret_cast<const char*>(idx + h);
        const size_t nvec = (n - h) / N;
        // Warp-uniform trip count (every lane of a warp runs the same iterations, out-of-range lanes
        // carry an invalid key), so the warp can vote.  Every 16th iteration the warp samples one key
        // per lane: if a quarter of the lanes see a duplicate, the keys are skewed and the next
        // iterations merge equal keys inside the warp (match.any + one atomic per distinct key)
        // instead of hammering one L2 address 32 times.  Uniform keys never pay for the merge.
        constexpr bool kCanMerge = RED == RT_Sum || RED == RT_Min || RED == RT_Max;
        const unsigned lane = threadIdx.x & 31;
        bool merge = false;
        unsigned it = 0;
        for (size_t vb = tid - lane; vb < nvec; vb += size_t(U) * nthr, ++it) {
            Words<16> w[U];
            bool have[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t vi = vb + lane + size_t(u) * nthr;
                have[u] = vi < nvec;
                if (have[u]) w[u] = ldg_stream16(pv + vi * 16);
            }
            if constexpr (kCanMerge) {
                if ((it & 15u) == 0) {
                    const long long k0 = have[0] ? idx_at<I>(w[0], 0) : -1 - (long long)lane;
                    const unsigned same = __match_any_sync(kFullMask, k0);
                    merge = __popc(__ballot_sync(kFullMask, __popc(same) > 1)) >= 8;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    const long long k = have[u] ? idx_at<I>(w[u], j) : -1;
                    const bool ok = k >= 0 && (size_t)k < nbins;
                    if (kCanMerge && merge) {
                        const unsigned act = __ballot_sync(kFullMask, ok);
                        if (ok) {
                            const unsigned peers = __match_any_sync(act, k);
                            if (lane == unsigned(__ffs(peers) - 1))
                                atomic_reduce<RED, V>(bins + k, RED == RT_Sum ? V(cval * V(__popc(peers))) : cval);
                        }
                    } else if (ok) {
                        atomic_reduce<RED, V>(bins + k, cval);
                    }
                }
        }
        for (size_t i = tid; i < h; i += nthr) {
            long long k = (long long)ldg_scalar(idx + i);
            if (k >= 0 && (size_t)k < nbins) atomic_reduce<RED, V>(bins + k, cval);
        }
        for (size_t i = h + nvec * N + tid; i < n; i += nthr) {
            long long k = (long long)ldg_scalar(idx + i);
            if (k >= 0 && (size_t)k < nbins) atomic_reduce<RED, V>(bins + k, cval);
        }
    }
}

// ---- shared-memory privatised path (Sum, Min, Max; bins that fit 96 KiB) ----------------------------------
// Every warp of a CTA owns one of up to 16 private copies of the bins (seeded with the operator's neutral
// element), so two lanes only meet on an address when they hold the same key.  Keys come in as 16-byte vectors,
// U of them in flight per thread (per-element values: U scalar key / value pairs); the first version of this
// kernel had ONE 4-byte load in flight per thread and ran 2^26 keys into 256 bins at 70-240 Gkeys/s, and Min / Max
// did not have it at all (global atomics on 256 addresses: 4.4 Gkeys/s).  Constant increments merge equal keys
// inside the warp (match.any) only while a periodic sample says that the keys are skewed, as the global kernel does.
// At the end the copies are folded in a fixed order and each bin that received something goes to the global
// bins with one atomic.
template <int RED, class V>
__device__ __forceinline__ V scatter_neutral() {
    if constexpr (RED == RT_Sum) return V(0);
    else if constexpr (std::is_same<V, float>::value) return __uint_as_float(RED == RT_Min ? 0x7f800000u : 0xff800000u);
    else if constexpr (std::is_same<V, double>::value)
        return __longlong_as_double(RED == RT_Min ? 0x7ff0000000000000ll : (long long)0xfff0000000000000ull);
    else return RED == RT_Min ? std::numeric_limits<V>::max() : std::numeric_limits<V>::lowest();
}
template <int RED, class V>
__device__ __forceinline__ V scatter_apply(V a, V b) {
    if constexpr (RED == RT_Sum) return a + b;
    else if constexpr (RED == RT_Min) return b < a ? b : a;
    else return b > a ? b : a;
}

template <int RED, class I, class V, bool HAS_VALUES, int U>
__global__ void __launch_bounds__(512)
scatter_smem_kernel(const I* __restrict__ idx, const V* __restrict__ values, V cval, size_t n,
                    V* __restrict__ bins, size_t nbins, int copies) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    V* hist = reinterpret_cast<V*>(smem_raw);
    const size_t total = nbins * copies;
    const V neutral = scatter_neutral<RED, V>();
    for (size_t i = threadIdx.x; i < total; i += blockDim.x) hist[i] = neutral;
    __syncthreads();
    V* mine = hist + size_t((threadIdx.x >> 5) % copies) * nbins;
    const size_t tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t nthr = size_t(gridDim.x) * blockDim.x;
    if constexpr (HAS_VALUES) {
        for (size_t i0 = tid; i0 < n; i0 += size_t(U) * nthr) {
            long long k[U];
            V v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + size_t(u) * nthr;
                k[u] = -1;
                if (i < n) { k[u] = (long long)ldg_scalar(idx + i); v[u] = ldg_scalar(values + i); }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (k[u] >= 0 && (size_t)k[u] < nbins) atomic_reduce<RED, V>(mine + k[u], v[u]);
        }
    } else {
        constexpr int N = 16 / int(sizeof(I));
        const size_t addr = reinterpret_cast<size_t>(idx);
        size_t h = ((16 - (addr & 15)) & 15) / sizeof(I);
        if (h > n) h = n;
        const char* pv = reinterpret_cast<const char*>(idx + h);
        const size_t nvec = (n - h) / N;
        const unsigned lane = threadIdx.x & 31;
        bool merge = false;
        unsigned it = 0;
        for (size_t vb = tid - lane; vb < nvec; vb += size_t(U) * nthr, ++it) {     // warp-uniform trip count
            Words<16> w[U];
            bool have[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t vi = vb + lane + size_t(u) * nthr;
                have[u] = vi < nvec;
                if (have[u]) w[u] = ldg_stream16(pv + vi * 16);
            }
            if ((it & 15u) == 0) {
                const long long k0 = have[0] ? idx_at<I>(w[0], 0) : -1 - (long long)lane;
                const unsigned same = __match_any_sync(kFullMask, k0);
                merge = __popc(__ballot_sync(kFullMask, __popc(same) > 1)) >= 16;     // half of the lanes share a key
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    const long long k = have[u] ? idx_at<I>(w[u], j) : -1;
                    const bool ok = k >= 0 && (size_t)k < nbins;
                    if (merge) {
                        const unsigned act = __ballot_sync(kFullMask, ok);
                        if (ok) {
                            const unsigned peers = __match_any_sync(act, k);
                            if (lane == unsigned(__ffs(peers) - 1))
                                atomic_reduce<RED, V>(mine + k, RED == RT_Sum ? V(cval * V(__popc(peers))) : cval);
                        }
                    } else if (ok) {
                        atomic_reduce<RED, V>(mine + k, cval);
                    }
                }
        }
        for (size_t i = tid; i < h; i += nthr) {
            const long long k = (long long)ldg_scalar(idx + i);
            if (k >= 0 && (size_t)k < nbins) atomic_reduce<RED, V>(mine + k, cval);
        }
        for (size_t i = h + nvec * N + tid; i < n; i += nthr) {
            const long long k = (long long)ldg_scalar(idx + i);
            if (k >= 0 && (size_t)k < nbins) atomic_reduce<RED, V>(mine + k, cval);
        }
    }
    __syncthreads();
    for (size_t b = threadIdx.x; b < nbins; b += blockDim.x) {
        V s = hist[b];
        for (int c = 1; c < copies; ++c) s = scatter_apply<RED, V>(s, hist[size_t(c) * nbins + b]);
        if (s != neutral) atomic_reduce<RED, V>(bins + b, s);
    }
}

// ---- hot-key cache path (Sum, constant increment, skewed keys) ---------------------------------
// Skewed key distributions (Zipf, all keys equal, keys confined to a small range) put much of the mass on
// a few bins; one atomic per key on those bins serialises in L2 (Zipf(1.1) over 2^20 bins: 21 ms, all keys
// equal: 5.8 ms for 2^28 keys with the kernels above).  Here every CTA keeps a direct-mapped cache of
// (bin, count) pairs in shared memory: a key whose slot is free claims it (compare-and-swap on the tag),
// a key that finds its own tag adds to the slot's count with a shared-memory atomic, and only keys whose
// slot belongs to another bin go to the global bins.  Hot keys show up early and often, so they own their
// slots; what reaches L2 is the cold tail.  Counts are flushed once at the end.  A warp whose 32 keys
// are all equal (runs, constant input) adds 32 with one atomic.  Integer results are bit-exact.
constexpr int kCacheSlots = 16384;                 // 64 KiB of tags + 64 KiB of counts
constexpr int kCacheThreads = 1024;

template <class V>
__device__ __forceinline__ void add_any(V* p, V v) {
    if constexpr (std::is_same<V, long long>::value)
        atomicAdd(reinterpret_cast<unsigned long long*>(p), static_cast<unsigned long long>(v));
    else
        atomicAdd(p, v);
}

template <class V, int U>
__global__ void __launch_bounds__(kCacheThreads, 1)
scatter_cache_kernel(const unsigned* __restrict__ keys, const V* __restrict__ values, size_t n, V* __restrict__ bins,
                     unsigned nbins, V cval, const int* __restrict__ gate, int gate_run_if) {
    if (gate != nullptr && *gate != gate_run_if) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned* tags = reinterpret_cast<unsigned*>(smem_raw);
    V* counts = reinterpret_cast<V*>(smem_raw + size_t(kCacheSlots) * sizeof(unsigned));
    for (int i = threadIdx.x; i < kCacheSlots; i += kCacheThreads) { tags[i] = 0xffffffffu; counts[i] = V(0); }
    __syncthreads();

    auto add_key = [&](unsigned k, V v) {
        if (k >= nbins) return;
        const unsigned slot = (k * 2654435761u) >> 18;            // top 14 bits
        unsigned tag = tags[slot];
        if (tag == 0xffffffffu) {
            const unsigned old = atomicCAS(&tags[slot], 0xffffffffu, k);
            tag = (old == 0xffffffffu) ? k : old;
        }
        if (tag == k) add_any(&counts[slot], v);
        else add_any(bins + k, v);
    };
    if (values != nullptr) {
        // per-element values: scalar coalesced loads of (key, value), U pairs in flight per thread
        const size_t tid = size_t(blockIdx.x) * kCacheThreads + threadIdx.x, nthr = size_t(gridDim.x) * kCacheThreads;
        size_t i = tid;
        for (; i + size_t(2 * U - 1) * nthr < n; i += size_t(2 * U) * nthr) {
            unsigned k[2 * U];
            V v[2 * U];
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) { k[u] = __ldg(keys + i + size_t(u) * nthr); v[u] = ldg_scalar(values + i + size_t(u) * nthr); }
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) add_key(k[u], v[u]);
        }
        for (; i < n; i += nthr) add_key(__ldg(keys + i), ldg_scalar(values + i));
        __syncthreads();
        for (int s = threadIdx.x; s < kCacheSlots; s += kCacheThreads) {
            const V c = counts[s];
            if (c != V(0)) add_any(bins + tags[s], c);
        }
        return;
    }
    const size_t addr = reinterpret_cast<size_t>(keys);
    size_t h = ((16 - (addr & 15)) & 15) / 4;
    if (h > n) h = n;
    const char* pv = reinterpret_cast<const char*>(keys + h);
    const size_t nvec = (n - h) / 4;
    const size_t tid = size_t(blockIdx.x) * kCacheThreads + threadIdx.x, nthr = size_t(gridDim.x) * kCacheThreads;
    const unsigned lane = threadIdx.x & 31;
    // warp-uniform trip count: every lane of a warp runs the same iterations (lanes past the end carry invalid keys)
    for (size_t vb = tid - lane; vb < nvec; vb += size_t(U) * nthr) {
        Words<16> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t vi = vb + lane + size_t(u) * nthr;
            if (vi < nvec) w[u] = ldg_stream16(pv + vi * 16);
            else w[u].w[0] = w[u].w[1] = w[u].w[2] = w[u].w[3] = 0xffffffffu;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            // all 128 keys of the warp equal (constant input, long runs): one atomic for the lot
            const unsigned k0 = __shfl_sync(kFullMask, w[u].w[0], 0);
            const bool same = w[u].w[0] == k0 && w[u].w[1] == k0 && w[u].w[2] == k0 && w[u].w[3] == k0;
            if (__all_sync(kFullMask, same)) {
                if (lane == 0) add_key(k0, cval * V(128));
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) add_key(w[u].w[j], cval);
            }
        }
    }
    for (size_t i = tid; i < h; i += nthr) add_key(__ldg(keys + i), cval);
    for (size_t i = h + nvec * 4 + tid; i < n; i += nthr) add_key(__ldg(keys + i), cval);
    __syncthreads();
    for (int i = threadIdx.x; i < kCacheSlots; i += kCacheThreads) {
        const V c = counts[i];
        if (c != V(0)) add_any(bins + tags[i], c);
    }
}

// Launched behind the partitioned kernel, gated on the sampler's flag (2 = skewed keys).
cudaError_t launch_scatter_cache(int dtval, const ScatterArgs& a, const Launch& l) {
    const unsigned grid = unsigned(sm_count());
#define DB200_CACHE_LAUNCH(V)                                                                                   \
    {                                                                                                           \
        V cval = static_cast<V>(1);                                                                             \
        if (a.const_value) memcpy(&cval, a.const_value, sizeof(V));                                             \
        const size_t smem = size_t(kCacheSlots) * (sizeof(unsigned) + sizeof(V));                               \
        static PerDeviceOnce once;                                                                              \
        if (bool* done = once.slot(); !*done) {                                                                 \
            cudaError_t e = cudaFuncSetAttribute(scatter_cache_kernel<V, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (e != cudaSuccess) return e;                                                                     \
            *done = true;                                                                                       \
        }                                                                                                       \
        scatter_cache_kernel<V, 4><<<grid, kCacheThreads, smem, l.stream>>>(                                    \
            (const unsigned*)a.idx, (const V*)a.values, a.n, (V*)a.bins, (unsigned)a.nbins, cval, a.gate, a.gate_run_if); \
        return cudaGetLastError();                                                                              \
    }
    switch (dtval) {
        case DT_I32: DB200_CACHE_LAUNCH(int)
        case DT_U32: DB200_CACHE_LAUNCH(unsigned)
        case DT_F32: DB200_CACHE_LAUNCH(float)
        case DT_I64: DB200_CACHE_LAUNCH(long long)
        case DT_U64: DB200_CACHE_LAUNCH(unsigned long long)
        case DT_F64: DB200_CACHE_LAUNCH(double)
        default: return cudaErrorNotSupported;
    }
#undef DB200_CACHE_LAUNCH
}

template <int RED, class I, class V>
static cudaError_t launch_scatter_t(const ScatterArgs& a, const Launch& l) {
    const int sms = sm_count();
    V cval = V(1);
    if (a.const_value) memcpy(&cval, a.const_value, sizeof(V));
    if (a.n == 0) return cudaSuccess;
    const size_t hist_bytes = a.nbins * sizeof(V);
    if constexpr (RED == RT_Sum || RED == RT_Min || RED == RT_Max) {
        if (hist_bytes <= (size_t(96) << 10) && a.n >= (size_t(1) << 14) && a.gate == nullptr) {
            int copies = int((size_t(96) << 10) / hist_bytes);
            if (copies > 16) copies = 16;
            if (copies < 1) copies = 1;
            const size_t smem = hist_bytes * copies;
            const size_t need = (a.n + 511) / 512;
            cudaError_t e;
#define DB200_SMEM_LAUNCH(HV)                                                                                  \
            {                                                                                                  \
                static PerDeviceOnce attr_set;                                                                 \
                bool* done = attr_set.slot();                                                                  \
                if (!*done) {                                                                                  \
                    e = cudaFuncSetAttribute(scatter_smem_kernel<RED, I, V, HV, 4>,                            \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(size_t(96) << 10)); \
                    if (e != cudaSuccess) return e;                                                            \
                    *done = true;                                                                              \
                }                                                                                              \
                size_t grid = size_t(sms) * 2;                                                                 \
                if (grid > need) grid = need;                                                                  \
                scatter_smem_kernel<RED, I, V, HV, 4><<<(unsigned)grid, 512, smem, l.stream>>>(                \
                    (const I*)a.idx, (const V*)a.values, cval, a.n, (V*)a.bins, a.nbins, copies);              \
            }
            if (a.values) DB200_SMEM_LAUNCH(true) else DB200_SMEM_LAUNCH(false)
#undef DB200_SMEM_LAUNCH
            return cudaGetLastError();
        }
    }
    // a persistent grid of exactly one wave (what fits depends on the instantiation's registers)
    size_t grid = size_t(sms) * (a.values ? resident_ctas<scatter_global_kernel<RED, I, V, true, 4>>(512)
                                          : resident_ctas<scatter_global_kernel<RED, I, V, false, 4>>(512));
    const size_t need = (a.n + 511) / 512;
    if (grid > need) grid = need;
    if (a.values)
        scatter_global_kernel<RED, I, V, true, 4><<<(unsigned)grid, 512, 0, l.stream>>>(
            (const I*)a.idx, (const V*)a.values, cval, a.n, (V*)a.bins, a.nbins, a.gate, a.gate_run_if);
    else
        scatter_global_kernel<RED, I, V, false, 4><<<(unsigned)grid, 512, 0, l.stream>>>(
            (const I*)a.idx, (const V*)a.values, cval, a.n, (V*)a.bins, a.nbins, a.gate, a.gate_run_if);
    return cudaGetLastError();
}

template <class I, class V>
static cudaError_t launch_scatter_op(int red, const ScatterArgs& a, const Launch& l) {
    switch (red) {
        case RT_Sum: return launch_scatter_t<RT_Sum, I, V>(a, l);
        case RT_Min: return launch_scatter_t<RT_Min, I, V>(a, l);
        case RT_Max: return launch_scatter_t<RT_Max, I, V>(a, l);
        case RT_Product: return launch_scatter_t<RT_Product, I, V>(a, l);
        default: break;
    }
    if constexpr (std::is_integral<V>::value) {
        switch (red) {
            case RT_BAnd: return launch_scatter_t<RT_BAnd, I, V>(a, l);
            case RT_BOr: return launch_scatter_t<RT_BOr, I, V>(a, l);
            case RT_BXor: return launch_scatter_t<RT_BXor, I, V>(a, l);
            default: break;
        }
    }
    return cudaErrorNotSupported;
}

template <class I>
static cudaError_t launch_scatter_val(int red, int dtval, const ScatterArgs& a, const Launch& l) {
    switch (dtval) {
        case DT_I32: return launch_scatter_op<I, int>(red, a, l);
        case DT_U32: return launch_scatter_op<I, unsigned>(red, a, l);
        case DT_I64: return launch_scatter_op<I, long long>(red, a, l);
        case DT_U64: return launch_scatter_op<I, unsigned long long>(red, a, l);
        case DT_F32: return launch_scatter_op<I, float>(red, a, l);
        case DT_F64: return launch_scatter_op<I, double>(red, a, l);
        default: return cudaErrorNotSupported;
    }
}

cudaError_t launch_scatter(int red, int dtidx, int dtval, const ScatterArgs& a, const Launch& l) {
    if (red == RT_Sum && !tuning().no_partition) {
        const int* gate = nullptr;
        cudaError_t e = launch_scatter_partitioned(dtidx, dtval, a, l, &gate);
        if (e == cudaSuccess) {
            // Keys that arrive sorted / in long runs are served better by the warp-merged global
            // atomics (0.36 ms vs 2.7 ms for sorted C4 keys), skewed keys by the hot-key cache: the
            // sampler decided on the device, exactly one of the three kernels does the work, the
            // others return at once.
            ScatterArgs g = a;
            g.gate = gate;
            g.gate_run_if = 1;
            if (dtidx == DT_I64) {
                // 64-bit keys: the hot-key cache takes 32-bit keys, so skewed keys (flag 2) go to the warp-merging
                // global kernel as well
                e = launch_scatter_val<long long>(red, dtval, g, l);
                if (e != cudaSuccess) return e;
                g.gate_run_if = 2;
                return launch_scatter_val<long long>(red, dtval, g, l);
            }
            e = (dtidx == DT_I32) ? launch_scatter_val<int>(red, dtval, g, l) : launch_scatter_val<unsigned>(red, dtval, g, l);
            if (e != cudaSuccess) return e;
            g.gate_run_if = 2;
            return launch_scatter_cache(dtval, g, l);
        }
        if (e != cudaErrorNotSupported) return e;
        // per-element values (weighted histograms) into a large bin array: the sampler still guards against skewed
        // keys - hot bins are absorbed by the shared-memory cache instead of serialising in L2
        e = launch_scatter_sample(dtidx, dtval, a, l, &gate);
        if (e == cudaSuccess) {
            ScatterArgs g = a;
            g.gate = gate;
            g.gate_run_if = -2;       // anything but "skewed"
            e = (dtidx == DT_I32) ? launch_scatter_val<int>(red, dtval, g, l) : launch_scatter_val<unsigned>(red, dtval, g, l);
            if (e != cudaSuccess) return e;
            g.gate_run_if = 2;
            return launch_scatter_cache(dtval, g, l);
        }
        if (e != cudaErrorNotSupported) return e;
    }
    switch (dtidx) {
        case DT_I32: return launch_scatter_val<int>(red, dtval, a, l);
        case DT_U32: return launch_scatter_val<unsigned>(red, dtval, a, l);
        case DT_I64: return launch_scatter_val<long long>(red, dtval, a, l);
        case DT_U8: return launch_scatter_val<unsigned char>(red, dtval, a, l);
        default: return cudaErrorNotSupported;
    }
}

}  // namespace db200
