/* dace_b200_shared_accum.cuh - privatised shared-memory tile -> global accumulate / reset, for
 * DaCe-generated CUDA kernels on sm_100.
 *
 * Replaces the reference's only "privatise in shared memory, then flush with atomics" primitive,
 *
 *     dace::ResetShared<T, BW, BH, BD, LEN, 1, ASYNC>::Reset(smem)
 *     dace::SharedToGlobal1D<T, BW, BH, BD, LEN, ASYNC>::template Accum<REDTYPE>(smem, sstride, gptr, gstride)
 *                     (reference dace/runtime/include/dace/cuda/copy.cuh:706-812)
 *
 * which the reference's CUDA code generator emits for a GPU_Shared transient with `setzero` and for a
 * shared -> global copy whose memlet carries a WCR (reference dace/codegen/targets/cuda.py:744-749,
 * :1379-1417) - the code shape AccumulateTransient (reference
 * dace/transformation/dataflow/stream_transient.py:136) produces when its transient lives in shared
 * memory.  The reference issues one atomic per tile element, neutral or not, float min / max / product
 * through compare-and-swap loops, scalar shared loads.  Here:
 *   - entries that still hold the operator's neutral element are skipped (a privatised histogram or
 *     scatter tile is mostly untouched: no global traffic at all for those);
 *   - the flush is `red.global` (no return value); float min / max use the ordered-integer atomics
 *     instead of CAS loops; contiguous fp32 sums go out as `red.global.add.v4.f32` (one L2 transaction
 *     per four bins) from 128-bit shared loads;
 *   - the reset writes 128-bit words when the tile allows it, and can seed any value (the reference
 *     can only zero, which is the neutral element of Sum / Or / Xor only).
 * Drop-in: when DaCe's runtime header is present (generated .cu files include <dace/dace.h> first),
 * this header specialises dace::SharedToGlobal1D / dace::ResetShared for the arithmetic element types,
 * so UNMODIFIED generated kernels pick the functions below up - the generated call sites stay as they
 * are.  Stand-alone use (tests/native/shared_accum_harness.cu): dace_b200::shared_reset / shared_accum.
 *
 * Contract (same as the reference's): every thread of the block calls; NTHREADS = threads per block
 * (x * y * z) at compile time, `ltid` the x-fastest linear thread id; LEN tile elements; strides in
 * elements; ASYNC == false ends with __syncthreads().  Results are exact for integer operators
 * (atomics commute); float sums depend on arrival order in the last ulps like the reference's own.
 */
#ifndef DACE_B200_SHARED_ACCUM_CUH
#define DACE_B200_SHARED_ACCUM_CUH

#if defined(__CUDACC__)

#include <stdint.h>

#include "dace_b200_block_reduce.cuh" /* operator codes kMin ... kBitwiseXor */

namespace dace_b200 {

namespace detail {

template <int BYTES> struct word_of { using type = unsigned; };
template <> struct word_of<8> { using type = unsigned long long; };

template <typename T> struct is_fp { static constexpr bool value = false; };
template <> struct is_fp<float> { static constexpr bool value = true; };
template <> struct is_fp<double> { static constexpr bool value = true; };

/* Neutral element of RED for T (what a freshly reset tile entry holds). */
template <int RED, typename T>
__device__ __forceinline__ T accum_neutral() {
    if constexpr (RED == kSum || RED == kBitwiseOr || RED == kBitwiseXor || RED == kLogicalOr || RED == kLogicalXor)
        return T(0);
    else if constexpr (RED == kProduct || RED == kLogicalAnd) return T(1);
    else if constexpr (RED == kBitwiseAnd) return (T)(~T(0));
    else if constexpr (RED == kMin) {
        if constexpr (sizeof(T) == 4 && is_fp<T>::value) return __int_as_float(0x7f800000);
        else if constexpr (sizeof(T) == 8 && is_fp<T>::value) return __longlong_as_double(0x7ff0000000000000LL);
        else if constexpr (T(-1) < T(0)) return (T)((~(unsigned long long)0) >> (65 - 8 * sizeof(T)));
        else return (T)(~T(0));
    } else {
        if constexpr (sizeof(T) == 4 && is_fp<T>::value) return __int_as_float(0xff800000);
        else if constexpr (sizeof(T) == 8 && is_fp<T>::value) return __longlong_as_double(0xfff0000000000000LL);
        else if constexpr (T(-1) < T(0)) return (T)(-(T)((~(unsigned long long)0) >> (65 - 8 * sizeof(T))) - 1);
        else return T(0);
    }
}

/* One `red.global`: fold `v` into *p with RED (arithmetic of dace::_wcr_fixed::reduce_atomic, reference
 * dace/runtime/include/dace/reduction.h:167-340, without the returned old value). */
template <int RED, typename T>
__device__ __forceinline__ void red_global(T* p, T v) {
    if constexpr (RED == kSum) {
        if constexpr (sizeof(T) == 8 && !is_fp<T>::value) atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
        else atomicAdd(p, v);
    } else if constexpr (RED == kMin || RED == kMax) {
        if constexpr (is_fp<T>::value && sizeof(T) == 4) {
            /* NaN-free ordering: non-negative floats order like ints, negative ones inversely like unsigned ints */
            const bool nonneg = !(__float_as_uint(v) >> 31);
            if (RED == kMax) { if (nonneg) atomicMax(reinterpret_cast<int*>(p), __float_as_int(v)); else atomicMin(reinterpret_cast<unsigned*>(p), __float_as_uint(v)); }
            else { if (nonneg) atomicMin(reinterpret_cast<int*>(p), __float_as_int(v)); else atomicMax(reinterpret_cast<unsigned*>(p), __float_as_uint(v)); }
        } else if constexpr (is_fp<T>::value) {
            const bool nonneg = !((unsigned long long)__double_as_longlong(v) >> 63);
            if (RED == kMax) { if (nonneg) atomicMax(reinterpret_cast<long long*>(p), __double_as_longlong(v)); else atomicMin(reinterpret_cast<unsigned long long*>(p), (unsigned long long)__double_as_longlong(v)); }
            else { if (nonneg) atomicMin(reinterpret_cast<long long*>(p), __double_as_longlong(v)); else atomicMax(reinterpret_cast<unsigned long long*>(p), (unsigned long long)__double_as_longlong(v)); }
        } else {
            if (RED == kMax) atomicMax(p, v); else atomicMin(p, v);
        }
    } else if constexpr (RED == kBitwiseAnd || RED == kBitwiseOr || RED == kBitwiseXor) {
        if constexpr (sizeof(T) == 4) {
            unsigned* a = reinterpret_cast<unsigned*>(p); const unsigned u = (unsigned)v;
            if (RED == kBitwiseAnd) atomicAnd(a, u); else if (RED == kBitwiseOr) atomicOr(a, u); else atomicXor(a, u);
        } else {
            unsigned long long* a = reinterpret_cast<unsigned long long*>(p); const unsigned long long u = (unsigned long long)v;
            if (RED == kBitwiseAnd) atomicAnd(a, u); else if (RED == kBitwiseOr) atomicOr(a, u); else atomicXor(a, u);
        }
    } else {
        /* Product and the logical operators: compare-and-swap on the element's bits */
        using W = typename ::dace_b200::detail::word_of<sizeof(T)>::type;
        W* a = reinterpret_cast<W*>(p);
        W old = *a, assumed;
        do {
            assumed = old;
            T cur;
            __builtin_memcpy(&cur, &assumed, sizeof(T));
            const T nv = op<RED, T>()(cur, v);
            W nw;
            __builtin_memcpy(&nw, &nv, sizeof(T));
            old = atomicCAS(a, assumed, nw);
        } while (old != assumed);
    }
}

}  // namespace detail

/* x-fastest linear thread id of a (BW, BH, BD) block (dace::GetLinearTID). */
template <int BW, int BH, int BD>
__device__ __forceinline__ int linear_tid() {
    if constexpr (BH == 1 && BD == 1) return threadIdx.x;
    else if constexpr (BD == 1) return threadIdx.x + threadIdx.y * BW;
    else return threadIdx.x + (threadIdx.y + threadIdx.z * BH) * BW;
}

/* smem[0 : LEN] = value (default 0), block-cooperative. */
template <typename T, int NTHREADS, int LEN, bool ASYNC = false>
__device__ __forceinline__ void shared_reset(T* smem, int ltid, T value = T(0)) {
    constexpr int kPerVec = 16 / (int)sizeof(T);
    if constexpr (sizeof(T) <= 8 && (LEN % kPerVec) == 0 && LEN >= kPerVec * 32) {
        if ((reinterpret_cast<uintptr_t>(smem) & 15) == 0) {
            T fill[kPerVec];
#pragma unroll
            for (int j = 0; j < kPerVec; ++j) fill[j] = value;
            uint4 w;
            __builtin_memcpy(&w, fill, 16);
            uint4* v = reinterpret_cast<uint4*>(smem);
            for (int i = ltid; i < LEN / kPerVec; i += NTHREADS) v[i] = w;
            if (!ASYNC) __syncthreads();
            return;
        }
    }
    for (int i = ltid; i < LEN; i += NTHREADS) smem[i] = value;
    if (!ASYNC) __syncthreads();
}

/* gptr[i * gstride] = RED(gptr[i * gstride], smem[i * sstride]) for i in [0, LEN), skipping entries that
 * hold RED's neutral element.  Every thread of the block calls.  SYNC_BEFORE (default) starts with a
 * barrier so that the other threads' updates of the tile are visible: the reference's generated code
 * goes from the thread-block scope that fills the tile straight into the flush without one. */
template <int RED, typename T, int NTHREADS, int LEN, bool ASYNC = false, bool SYNC_BEFORE = true>
__device__ __forceinline__ void shared_accum(const T* smem, int sstride, T* gptr, int gstride, int ltid) {
    if (SYNC_BEFORE) __syncthreads();
    const T neutral = detail::accum_neutral<RED, T>();
    if constexpr (RED == kSum && sizeof(T) == 4 && detail::is_fp<T>::value && (LEN % 4) == 0) {
        /* contiguous fp32 tile into contiguous bins: 128-bit shared loads, one vector reduction per four bins */
        if (sstride == 1 && gstride == 1 && ((reinterpret_cast<uintptr_t>(smem) | reinterpret_cast<uintptr_t>(gptr)) & 15) == 0) {
            const float4* s4 = reinterpret_cast<const float4*>(smem);
            for (int i = ltid; i < LEN / 4; i += NTHREADS) {
                const float4 v = s4[i];
                if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f)
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gptr + 4 * i), "f"(v.x), "f"(v.y),
                                 "f"(v.z), "f"(v.w) : "memory");
            }
            if (!ASYNC) __syncthreads();
            return;
        }
    }
    for (int i = ltid; i < LEN; i += NTHREADS) {
        const T v = smem[(long long)i * sstride];
        if (!(v == neutral)) detail::red_global<RED, T>(gptr + (long long)i * gstride, v);
    }
    if (!ASYNC) __syncthreads();
}

}  // namespace dace_b200

/* ---- drop-in for DaCe-generated kernels ------------------------------------------------------
 * The generated .cu includes <dace/dace.h> before the environment's headers, so the primary
 * templates exist here; the specialisations below are chosen for every later instantiation in the
 * file.  Copy and the custom-lambda Accum keep the reference's behaviour. */
#if defined(__DACE_CUDACOPY_CUH)
namespace dace {

#define DACE_B200_SPECIALISE_SHARED_TILE(T)                                                                      \
    template <int BW, int BH, int BD, int LEN, bool ASYNC>                                                       \
    struct SharedToGlobal1D<T, BW, BH, BD, LEN, ASYNC> {                                                         \
        static constexpr int BLOCK_SIZE = BW * BH * BD;                                                          \
        static constexpr int kDaceB200 = 1; /* lets a static_assert prove this specialisation is the one in use */ \
        static DACE_DFI void Copy(const T* smem, int src_xstride, T* ptr, int dst_xstride) {                     \
            const int ltid = ::dace_b200::linear_tid<BW, BH, BD>();                                              \
            for (int i = ltid; i < LEN; i += BLOCK_SIZE) ptr[(long long)i * dst_xstride] = smem[(long long)i * src_xstride]; \
            if (!ASYNC) __syncthreads();                                                                         \
        }                                                                                                        \
        template <typename WCR>                                                                                  \
        static DACE_DFI void Accum(const T* smem, int src_xstride, T* ptr, int dst_xstride, WCR wcr) {          \
            const int ltid = ::dace_b200::linear_tid<BW, BH, BD>();                                              \
            for (int i = ltid; i < LEN; i += BLOCK_SIZE)                                                         \
                wcr_custom<T>::template reduce_atomic<WCR>(wcr, ptr + (long long)i * dst_xstride,               \
                                                           smem[(long long)i * src_xstride]);                   \
            if (!ASYNC) __syncthreads();                                                                         \
        }                                                                                                        \
        template <ReductionType REDTYPE>                                                                         \
        static DACE_DFI void Accum(const T* smem, int src_xstride, T* ptr, int dst_xstride) {                   \
            ::dace_b200::shared_accum<(int)REDTYPE, T, BLOCK_SIZE, LEN, ASYNC>(                                  \
                smem, src_xstride, ptr, dst_xstride, ::dace_b200::linear_tid<BW, BH, BD>());                     \
        }                                                                                                        \
    };                                                                                                           \
    template <int BW, int BH, int BD, int LEN, bool ASYNC>                                                       \
    struct ResetShared<T, BW, BH, BD, LEN, 1, ASYNC> {                                                           \
        static constexpr int kDaceB200 = 1;                                                                      \
        static DACE_DFI void Reset(T* smem) {                                                                    \
            ::dace_b200::shared_reset<T, BW * BH * BD, LEN, ASYNC>(smem, ::dace_b200::linear_tid<BW, BH, BD>()); \
        }                                                                                                        \
    };

DACE_B200_SPECIALISE_SHARED_TILE(float)
DACE_B200_SPECIALISE_SHARED_TILE(double)
DACE_B200_SPECIALISE_SHARED_TILE(int)
DACE_B200_SPECIALISE_SHARED_TILE(unsigned int)
DACE_B200_SPECIALISE_SHARED_TILE(long long)
DACE_B200_SPECIALISE_SHARED_TILE(unsigned long long)
#undef DACE_B200_SPECIALISE_SHARED_TILE

}  // namespace dace
#endif /* __DACE_CUDACOPY_CUH */

#endif /* __CUDACC__ */
#endif /* DACE_B200_SHARED_ACCUM_CUH */
