/* dace_b200_block_reduce.cuh - in-kernel (thread-block / warp) reductions for DaCe-generated
 * CUDA kernels on sm_100.
 *
 * Device-side counterpart of dace_b200_reduce.h.  A reduction that happens INSIDE a generated
 * kernel cannot cross a C ABI, so - exactly like the code it replaces - this boundary is a C++
 * header that DaCe's own nvcc pass compiles into the program's .cu file.  It replaces what the
 * reference's `CUDA (block)` / `CUDA (block allreduce)` expansions emit,
 *
 *     typedef cub::BlockReduce<T, NTHREADS> BlockReduce_<id>;
 *     __shared__ typename BlockReduce_<id>::TempStorage temp_storage_<id>;
 *     out = BlockReduce_<id>(temp_storage_<id>).Reduce(in, op);
 *                     (reference dace/libraries/standard/nodes/reduce.py:698-712)
 *
 * and the warp tree of the reference runtime (dace::warpReduce<..>::reduce, reference
 * dace/runtime/include/dace/reduction.h:624-660).  Contract kept from CUB's BlockReduce, which
 * is what generated code relies on:
 *   - every thread of the block calls (the call contains a barrier);
 *   - NTHREADS is the compile-time number of threads of the block (x*y*z), any value <= 1024;
 *     threads are ranked x-fastest like CUB's linear thread id;
 *   - `in` is a per-thread register value or register array, all of it is folded;
 *   - block_reduce: the result is defined in linear thread 0 only (other threads get a partial);
 *     block_allreduce: defined in every thread;
 *   - any associative + commutative functor `T op(const T&, const T&)` works (custom WCR lambdas
 *     included, reference reduce.py:676-689); the built-in functors below additionally map
 *     32-bit integer Sum/Min/Max/And/Or/Xor to one `redux.sync` per warp.
 * Shared memory: one element of T per warp (+1) per call site; UID (any integer unique per call
 * site inside one kernel, the expansion derives it from the node id) keeps the scratch of
 * different call sites apart, and the call ends with a barrier, so neither back-to-back
 * reductions nor a reduction inside a loop need an extra __syncthreads().
 *
 * No dependency on DaCe's or CUB's headers.
 */
#ifndef DACE_B200_BLOCK_REDUCE_CUH
#define DACE_B200_BLOCK_REDUCE_CUH

#if defined(__CUDACC__)

#include <stdint.h>

namespace dace_b200 {

/* Operator codes = dace::ReductionType (reference dace/runtime/include/dace/types.h:301-317). */
enum : int {
    kMin = 1, kMax = 2, kSum = 3, kProduct = 4, kLogicalAnd = 5, kBitwiseAnd = 6, kLogicalOr = 7,
    kBitwiseOr = 8, kLogicalXor = 9, kBitwiseXor = 10
};

/* Built-in functors; arithmetic of dace::_wcr_fixed (reference reduction.h:167-517, math.h:51-55). */
template <int RED, typename T>
struct op {
    static constexpr int kRed = RED;
    __device__ __forceinline__ T operator()(const T& a, const T& b) const {
        if constexpr (RED == kMin) return (a < b) ? a : b;
        else if constexpr (RED == kMax) return (a > b) ? a : b;
        else if constexpr (RED == kSum) return a + b;
        else if constexpr (RED == kProduct) return a * b;
        else if constexpr (RED == kLogicalAnd) return (T)(a && b);
        else if constexpr (RED == kBitwiseAnd) return a & b;
        else if constexpr (RED == kLogicalOr) return (T)(a || b);
        else if constexpr (RED == kBitwiseOr) return a | b;
        else if constexpr (RED == kLogicalXor) return (T)((a != T(0)) != (b != T(0)));
        else return a ^ b;
    }
};

namespace detail {

template <typename T> struct is_int32 { static constexpr bool value = false; };
template <> struct is_int32<int> { static constexpr bool value = true; };
template <> struct is_int32<unsigned> { static constexpr bool value = true; };

template <typename Op> struct redux_code { static constexpr int value = 0; };
template <int RED, typename T>
struct redux_code<op<RED, T>> {
    static constexpr int value =
        (is_int32<T>::value && (RED == kMin || RED == kMax || RED == kSum || RED == kBitwiseAnd ||
                                RED == kBitwiseOr || RED == kBitwiseXor)) ? RED : 0;
};

template <typename T>
__device__ __forceinline__ T shfl_down(unsigned mask, const T& v, int delta) {
    static_assert(sizeof(T) % 4 == 0 || sizeof(T) < 4, "unsupported element size");
    if constexpr (sizeof(T) <= 4) {
        uint32_t w = 0;
        __builtin_memcpy(&w, &v, sizeof(T));
        w = __shfl_down_sync(mask, w, delta);
        T r;
        __builtin_memcpy(&r, &w, sizeof(T));
        return r;
    } else {  /* 8, 12, 16... bytes (double, int64, value/index pairs): one shuffle per word */
        constexpr int W = sizeof(T) / 4;
        uint32_t w[W];
        __builtin_memcpy(w, &v, sizeof(T));
#pragma unroll
        for (int i = 0; i < W; ++i) w[i] = __shfl_down_sync(mask, w[i], delta);
        T r;
        __builtin_memcpy(&r, w, sizeof(T));
        return r;
    }
}

template <typename T>
__device__ __forceinline__ T shfl_lane0(unsigned mask, const T& v) {
    constexpr int W = (sizeof(T) + 3) / 4;
    uint32_t w[W] = {};
    __builtin_memcpy(w, &v, sizeof(T));
#pragma unroll
    for (int i = 0; i < W; ++i) w[i] = __shfl_sync(mask, w[i], 0);
    T r;
    __builtin_memcpy(&r, w, sizeof(T));
    return r;
}

/* Tree over the first LANES lanes of a warp (1 <= LANES <= 32); only those lanes call.
 * Result in lane 0. */
template <int LANES, typename T, typename Op>
__device__ __forceinline__ T warp_fold(T v, Op f, int lane) {
    constexpr unsigned mask = LANES == 32 ? 0xffffffffu : ((1u << LANES) - 1u);
    constexpr int R = redux_code<Op>::value;
    if constexpr (R == kSum) return (T)__reduce_add_sync(mask, v);
    else if constexpr (R == kMin) return (T)__reduce_min_sync(mask, v);
    else if constexpr (R == kMax) return (T)__reduce_max_sync(mask, v);
    else if constexpr (R == kBitwiseAnd) return (T)__reduce_and_sync(mask, (unsigned)v);
    else if constexpr (R == kBitwiseOr) return (T)__reduce_or_sync(mask, (unsigned)v);
    else if constexpr (R == kBitwiseXor) return (T)__reduce_xor_sync(mask, (unsigned)v);
    else {
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            if (d < LANES) {
                T o = shfl_down(mask, v, d);
                if (LANES == 32 || lane + d < LANES) v = f(v, o);
            }
        }
        return v;
    }
}

__device__ __forceinline__ int linear_tid() {
    return threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
}

template <typename T, int NTHREADS, int UID, typename Op>
__device__ __forceinline__ T block_fold(T v, Op f, bool broadcast) {
    static_assert(NTHREADS >= 1 && NTHREADS <= 1024, "block size out of range");
    constexpr int NWARPS = (NTHREADS + 31) / 32;
    constexpr int TAIL = NTHREADS - (NWARPS - 1) * 32; /* threads of the last warp */
    const int tid = linear_tid();
    const int lane = tid & 31, warp = tid >> 5;
    if constexpr (NWARPS == 1) {
        v = warp_fold<TAIL>(v, f, lane);
        if (broadcast) v = shfl_lane0(TAIL == 32 ? 0xffffffffu : ((1u << TAIL) - 1u), v);
        return v;
    } else {
        __shared__ T scratch[NWARPS + 1];
        if (TAIL != 32 && warp == NWARPS - 1) v = warp_fold<TAIL>(v, f, lane);
        else v = warp_fold<32>(v, f, lane);
        if (lane == 0) scratch[warp] = v;
        __syncthreads();
        if (warp == 0) {
            if (lane < NWARPS) {
                T x = warp_fold<NWARPS>(scratch[lane], f, lane);
                if (lane == 0) { v = x; if (broadcast) scratch[NWARPS] = x; }
            }
        }
        /* Second barrier: publishes the result for the broadcast and makes the scratch reusable, so
         * a call site inside a loop needs no barrier of its own (CUB asks the caller for one). */
        __syncthreads();
        if (broadcast) v = scratch[NWARPS];
        return v;
    }
}

}  // namespace detail

/* ---- public entry points -------------------------------------------------------------- */

/* Warp-wide fold over all 32 lanes, result in every lane (replaces dace::warpReduce<..>::reduce). */
template <typename T, typename Op>
__device__ __forceinline__ T warp_allreduce(T v, Op f) {
    constexpr int R = detail::redux_code<Op>::value;
    if constexpr (R != 0) return detail::warp_fold<32>(v, f, 0);
    else {
#pragma unroll
        for (int m = 16; m >= 1; m >>= 1) {
            T o;
            if constexpr (sizeof(T) <= 4) {
                uint32_t w = 0;
                __builtin_memcpy(&w, &v, sizeof(T));
                w = __shfl_xor_sync(0xffffffffu, w, m);
                __builtin_memcpy(&o, &w, sizeof(T));
            } else {
                constexpr int W = sizeof(T) / 4;
                uint32_t w[W];
                __builtin_memcpy(w, &v, sizeof(T));
#pragma unroll
                for (int i = 0; i < W; ++i) w[i] = __shfl_xor_sync(0xffffffffu, w[i], m);
                __builtin_memcpy(&o, w, sizeof(T));
            }
            v = f(v, o);
        }
        return v;
    }
}

/* Block-wide fold of one value per thread; result in linear thread 0. */
template <typename T, int NTHREADS, int UID, typename Op>
__device__ __forceinline__ T block_reduce(const T& in, Op f) {
    return detail::block_fold<T, NTHREADS, UID>(in, f, false);
}

/* Block-wide fold of ITEMS register values per thread; result in linear thread 0. */
template <typename T, int NTHREADS, int UID, int ITEMS, typename Op>
__device__ __forceinline__ T block_reduce(const T (&in)[ITEMS], Op f) {
    T v = in[0];
#pragma unroll
    for (int i = 1; i < ITEMS; ++i) v = f(v, in[i]);
    return detail::block_fold<T, NTHREADS, UID>(v, f, false);
}

/* Same, result in every thread of the block. */
template <typename T, int NTHREADS, int UID, typename Op>
__device__ __forceinline__ T block_allreduce(const T& in, Op f) {
    return detail::block_fold<T, NTHREADS, UID>(in, f, true);
}

template <typename T, int NTHREADS, int UID, int ITEMS, typename Op>
__device__ __forceinline__ T block_allreduce(const T (&in)[ITEMS], Op f) {
    T v = in[0];
#pragma unroll
    for (int i = 1; i < ITEMS; ++i) v = f(v, in[i]);
    return detail::block_fold<T, NTHREADS, UID>(v, f, true);
}

}  // namespace dace_b200

#endif /* __CUDACC__ */
#endif /* DACE_B200_BLOCK_REDUCE_CUH */
