/* dace_b200_fused_reduce.cuh - map -> reduce in ONE kernel, for DaCe-generated programs on sm_100.
 *
 * A DaCe program such as `s[:] = np.sum(np.exp(A - mx[:, None]), axis=1)` is a Map that writes a
 * transient, followed by a Reduce node that reads it back: three passes over HBM (read A, write
 * T, read T).  The reference's answer is MapReduceFusion (reference
 * dace/transformation/dataflow/mapreduce.py:18-146), which turns the pair into a WCR map - one
 * atomic per element on a GPU (reference dace/codegen/targets/cpu.py:965-1013).  Here the
 * `FuseMapIntoReduce` rewrite (dace_b200/fused_expansion.py) generates a LOAD FUNCTOR from the
 * map's tasklet and instantiates the streaming kernels below with it inside the program's own
 * .cu file: the elementwise producer runs in registers, the transient never exists, one pass.
 *
 * Like dace_b200_block_reduce.cuh this part of the boundary is a device header (the functor is
 * generated code, so it cannot sit behind a precompiled C ABI) - the same mechanism by which the
 * reference's CUB expansion instantiates cub::DeviceReduce with a generated functor (reference
 * dace/libraries/standard/nodes/reduce.py:453-463,574-588).
 *
 *   out[r * out_stride] = fold(op, identity, load(r, 0), ..., load(r, n - 1))      r in [0, rows)
 *
 * `load(row, col)` is any device functor returning a value convertible to the accumulator A.
 * Lanes walk `col`, so a functor that reads x[row * ld + col] is coalesced.  Long rows are split
 * over several CTAs; the per-split partials go through the caller's workspace (the partials area
 * of the dace_b200 workspace, see dace_b200_reduce.h) and a second tiny kernel folds them in index
 * order: deterministic results, no atomics.  `identity == nullptr` folds into the value already in
 * `out` (reference dace/frontend/operations.py:251).
 */
#ifndef DACE_B200_FUSED_REDUCE_CUH
#define DACE_B200_FUSED_REDUCE_CUH

#if defined(__CUDACC__)

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "dace_b200_block_reduce.cuh"

namespace dace_b200 {

/* Neutral element of the built-in operators (seed of every thread's private accumulator). */
template <int RED, typename A>
__host__ __device__ __forceinline__ A neutral() {
    if constexpr (RED == kSum || RED == kBitwiseOr || RED == kBitwiseXor || RED == kLogicalOr || RED == kLogicalXor)
        return A(0);
    else if constexpr (RED == kProduct || RED == kLogicalAnd) return A(1);
    else if constexpr (RED == kBitwiseAnd) return (A)(~A(0));
    else if constexpr (RED == kMin) {
        if constexpr (sizeof(A) == 4 && A(0.5) != A(0)) { const uint32_t b = 0x7f800000u; A r; __builtin_memcpy(&r, &b, 4); return r; }
        else if constexpr (sizeof(A) == 8 && A(0.5) != A(0)) { const uint64_t b = 0x7ff0000000000000ull; A r; __builtin_memcpy(&r, &b, 8); return r; }
        else if constexpr (A(-1) < A(0)) return (A)((~(unsigned long long)0) >> (65 - 8 * sizeof(A)));
        else return (A)(~A(0));
    } else {  /* kMax */
        if constexpr (sizeof(A) == 4 && A(0.5) != A(0)) { const uint32_t b = 0xff800000u; A r; __builtin_memcpy(&r, &b, 4); return r; }
        else if constexpr (sizeof(A) == 8 && A(0.5) != A(0)) { const uint64_t b = 0xfff0000000000000ull; A r; __builtin_memcpy(&r, &b, 8); return r; }
        else if constexpr (A(-1) < A(0)) return (A)(-(A)((~(unsigned long long)0) >> (65 - 8 * sizeof(A))) - 1);
        else return A(0);
    }
}

/* What the kernels below need to know about an operator: the accumulator type, the functor, the seed of
 * a thread's private accumulator, and the conversions at the two ends.  `builtin_ops` = the operators of
 * dace::ReductionType with their neutral elements.  `custom_ops` = ANY associative + commutative functor
 * `T f(const T&, const T&)` (custom WCR lambdas, reference dace/libraries/standard/nodes/reduce.py:453-463,
 * dace/runtime/include/dace/reduction.h:56-164): such an operator has no known neutral element, so the
 * accumulator carries a "holds a value" flag and empty accumulators are skipped when two are combined. */
template <int RED, typename A_>
struct builtin_ops {
    using A = A_;
    using functor = op<RED, A_>;
    __host__ __device__ static __forceinline__ A seed() { return neutral<RED, A_>(); }
    template <typename V> __device__ static __forceinline__ A from_value(const V& v) { return (A)v; }
    template <typename S> __device__ static __forceinline__ S to_output(const A& a) { return (S)a; }
    __host__ static A from_identity(const A_* identity) { return identity ? *identity : A_(0); }
};

template <typename T>
struct maybe {   /* value + "holds a value" */
    T v;
    int has;
};

template <typename F, typename T>
struct custom_ops {
    using A = maybe<T>;
    struct functor {
        __device__ __forceinline__ A operator()(const A& a, const A& b) const {
            if (!a.has) return b;
            if (!b.has) return a;
            return A{F()(a.v, b.v), 1};
        }
    };
    __host__ __device__ static __forceinline__ A seed() { return A{T(), 0}; }
    template <typename V> __device__ static __forceinline__ A from_value(const V& v) { return A{(T)v, 1}; }
    template <typename S> __device__ static __forceinline__ S to_output(const A& a) { return (S)a.v; }
    __host__ static A from_identity(const T* identity) { return identity ? A{*identity, 1} : A{T(), 0}; }
};

/* VECTOR LOADS.  A load functor may, next to operator(), provide
 *
 *     static constexpr int kVec = VMAX;                                              (a power of two, 2 ... 8)
 *     using value_type = ...;
 *     template <int V> __device__ void vec(size_t row, size_t col, value_type (&v)[V]) const;
 *                                                                  v[k] = load(row, col + k), col a multiple of V
 *
 * for V = VMAX and VMAX / 2; the call site passes `max_vec` = the widest V that is legal for THIS call (n a
 * multiple of V and every operand the functor reads with a vector access aligned to V elements).  The kernels then
 * hand V adjacent columns to one thread, so a functor over contiguous operands issues 16/32-byte loads
 * (`stream_load` below) instead of V scalar ones.  dace_b200/fused_expansion.py generates vec() whenever an
 * operand of the fused map is read at the map's own flat index. */
/* V consecutive elements at `p` (aligned to min(V * sizeof(T), 32) bytes), streamed past L1. */
template <typename T, int V>
__device__ __forceinline__ void stream_load(const T* p, T (&out)[V]) {
    constexpr int kBytes = int(sizeof(T)) * V;
    if constexpr (kBytes < 4) {                       /* two one-byte elements ...: nothing to gain, plain loads */
#pragma unroll
        for (int k = 0; k < V; ++k) out[k] = p[k];
        return;
    } else if constexpr (kBytes > 32) {               /* in 32-byte pieces */
        constexpr int kPer = 32 / int(sizeof(T));
#pragma unroll
        for (int k = 0; k < V; k += kPer) stream_load<T, kPer>(p + k, reinterpret_cast<T(&)[kPer]>(out[k]));
        return;
    }
    static_assert(kBytes < 4 || kBytes > 32 || kBytes == 4 || kBytes == 8 || kBytes == 16 || kBytes == 32,
                  "stream_load: a power-of-two number of bytes");
    uint32_t w[kBytes >= 4 && kBytes <= 32 ? kBytes / 4 : 1];
    if constexpr (kBytes == 4)
        asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w[0]) : "l"(p));
    else if constexpr (kBytes == 8)
        asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
    else if constexpr (kBytes == 16)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
    else
        asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
    __builtin_memcpy(out, w, kBytes);
}

namespace detail {

template <typename L, typename = void>
struct load_vec_width { static constexpr int value = 1; };
template <typename L>
struct load_vec_width<L, decltype((void)L::kVec)> { static constexpr int value = L::kVec; };

constexpr int kFusedThreads = 256;
#ifndef DACE_B200_FUSED_UNROLL
#define DACE_B200_FUSED_UNROLL 8
#endif
constexpr int kFusedUnroll = DACE_B200_FUSED_UNROLL;   /* independent loads in flight per thread */

template <typename P, typename S>
__device__ __forceinline__ void fused_store(S* out, typename P::A v, typename P::A identity, int has_identity) {
    const typename P::A base = has_identity ? identity : P::from_value(*out);
    *out = P::template to_output<S>(typename P::functor()(base, v));
}

/* A CTA per (row, split) work item, CTAs loop over the items (persistent grid): threads stride the columns of
 * the split with kFusedUnroll independent accumulators (that many loads in flight per thread), then a block tree. */
template <typename P, typename S, typename Load>
__global__ void __launch_bounds__(kFusedThreads)
fused_rows_cta_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, unsigned splits,
                      typename P::A* __restrict__ partials, typename P::A identity, int has_identity) {
    using A = typename P::A;
    const typename P::functor f;
    const size_t seg = (n + splits - 1) / splits;
    const size_t items = rows * splits;
    for (size_t item = blockIdx.x; item < items; item += gridDim.x) {
        const size_t row = item / splits;
        const unsigned split = unsigned(item % splits);
        const size_t lo = size_t(split) * seg;
        const size_t hi = (lo + seg < n) ? lo + seg : n;
        A acc[kFusedUnroll];
#pragma unroll
        for (int u = 0; u < kFusedUnroll; ++u) acc[u] = P::seed();
        size_t j = lo + threadIdx.x;
        for (; j + size_t(kFusedUnroll - 1) * kFusedThreads < hi; j += size_t(kFusedUnroll) * kFusedThreads) {
            A v[kFusedUnroll];
#pragma unroll
            for (int u = 0; u < kFusedUnroll; ++u) v[u] = P::from_value(load(row, j + size_t(u) * kFusedThreads));
#pragma unroll
            for (int u = 0; u < kFusedUnroll; ++u) acc[u] = f(acc[u], v[u]);
        }
        for (; j < hi; j += kFusedThreads) acc[0] = f(acc[0], P::from_value(load(row, j)));
        A r = block_reduce<A, kFusedThreads, 0x7f01>(acc, f);      /* ends with a barrier: safe inside the loop */
        if (threadIdx.x == 0) {
            if (splits == 1) fused_store<P, S>(out + (ptrdiff_t)row * out_stride, r, identity, has_identity);
            else partials[item] = r;
        }
    }
}

/* Short rows: a warp takes two rows at a time, four columns-of-32 each: eight loads in flight per lane. */
template <typename P, typename S, typename Load>
__global__ void __launch_bounds__(kFusedThreads)
fused_rows_warp_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A identity,
                       int has_identity) {
    const unsigned lane = threadIdx.x & 31;
    const size_t warps_per_grid = size_t(gridDim.x) * (kFusedThreads / 32);
    using A = typename P::A;
    const typename P::functor f;
    for (size_t row = 2 * (size_t(blockIdx.x) * (kFusedThreads / 32) + (threadIdx.x >> 5)); row < rows;
         row += 2 * warps_per_grid) {
        const bool two = row + 1 < rows;
        const size_t row1 = two ? row + 1 : row;
        A a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { a[u] = P::seed(); b[u] = P::seed(); }
        size_t j = lane;
        for (; j + 96 < n; j += 128) {
            A va[4], vb[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) va[u] = P::from_value(load(row, j + 32 * u));
#pragma unroll
            for (int u = 0; u < 4; ++u) vb[u] = P::from_value(load(row1, j + 32 * u));
#pragma unroll
            for (int u = 0; u < 4; ++u) { a[u] = f(a[u], va[u]); b[u] = f(b[u], vb[u]); }
        }
        for (; j < n; j += 32) {
            const A va = P::from_value(load(row, j)), vb = P::from_value(load(row1, j));
            a[0] = f(a[0], va);
            b[0] = f(b[0], vb);
        }
        const A ra = warp_allreduce(f(f(a[0], a[1]), f(a[2], a[3])), f);
        const A rb = warp_allreduce(f(f(b[0], b[1]), f(b[2], b[3])), f);
        if (lane == 0) {
            fused_store<P, S>(out + (ptrdiff_t)row * out_stride, ra, identity, has_identity);
            if (two) fused_store<P, S>(out + (ptrdiff_t)row1 * out_stride, rb, identity, has_identity);
        }
    }
}

template <typename P, typename S>
__global__ void fused_finish_kernel(const typename P::A* __restrict__ partials, S* __restrict__ out, ptrdiff_t out_stride, size_t rows,
                                    unsigned splits, typename P::A identity, int has_identity) {
    const size_t row = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    using A = typename P::A;
    const typename P::functor f;
    A r = partials[row * splits];
    for (unsigned s = 1; s < splits; ++s) r = f(r, partials[row * splits + s]);   /* fixed order */
    fused_store<P, S>(out + (ptrdiff_t)row * out_stride, r, identity, has_identity);
}

/* Column fold: out[c] = fold over r of load(r, c).  A CTA owns 32 adjacent columns (one per lane, so the
 * functor's reads of x[r * ld + c] are coalesced) and one slice of the rows; its 8 warps take rows
 * warp, warp + 8, ... with kFusedColUnroll rows in flight per warp. */
constexpr int kFusedColUnroll = 8;

template <typename P, typename S, typename Load>
__global__ void __launch_bounds__(kFusedThreads)
fused_cols_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, unsigned splits,
                  typename P::A* __restrict__ partials, typename P::A identity, int has_identity) {
    using A = typename P::A;
    constexpr int kWarps = kFusedThreads / 32;
    __shared__ A tile[kWarps][33];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t col = size_t(blockIdx.x) * 32 + lane;
    const unsigned split = blockIdx.y;
    const size_t seg = (rows + splits - 1) / splits;
    const size_t lo = size_t(split) * seg;
    const size_t hi = (lo + seg < rows) ? lo + seg : rows;
    const typename P::functor f;
    A acc = P::seed();
    if (col < n) {
        size_t r = lo + warp;
        for (; r + size_t(kFusedColUnroll - 1) * kWarps < hi; r += size_t(kFusedColUnroll) * kWarps) {
            A v[kFusedColUnroll];
#pragma unroll
            for (int u = 0; u < kFusedColUnroll; ++u) v[u] = P::from_value(load(r + size_t(u) * kWarps, col));
#pragma unroll
            for (int u = 0; u < kFusedColUnroll; ++u) acc = f(acc, v[u]);
        }
        for (; r < hi; r += kWarps) acc = f(acc, P::from_value(load(r, col)));
    }
    tile[warp][lane] = acc;
    __syncthreads();
    if (warp == 0 && col < n) {
        A r = tile[0][lane];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) r = f(r, tile[w][lane]);
        if (splits == 1) fused_store<P, S>(out + (ptrdiff_t)col * out_stride, r, identity, has_identity);
        else partials[size_t(split) * n + col] = r;
    }
}

/* Few rows, many columns: a thread per column, all rows of the slice in registers' reach. */
template <typename P, typename S, typename Load>
__global__ void __launch_bounds__(kFusedThreads)
fused_cols_thread_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A identity,
                         int has_identity) {
    const size_t col = size_t(blockIdx.x) * kFusedThreads + threadIdx.x;
    if (col >= n) return;
    using A = typename P::A;
    const typename P::functor f;
    A acc = P::seed();
    size_t r = 0;
    for (; r + kFusedColUnroll <= rows; r += kFusedColUnroll) {
        A v[kFusedColUnroll];
#pragma unroll
        for (int u = 0; u < kFusedColUnroll; ++u) v[u] = P::from_value(load(r + u, col));
#pragma unroll
        for (int u = 0; u < kFusedColUnroll; ++u) acc = f(acc, v[u]);
    }
    for (; r < rows; ++r) acc = f(acc, P::from_value(load(r, col)));
    fused_store<P, S>(out + (ptrdiff_t)col * out_stride, acc, identity, has_identity);
}

template <typename P, typename S>
__global__ void fused_cols_finish_kernel(const typename P::A* __restrict__ partials, S* __restrict__ out, ptrdiff_t out_stride, size_t n,
                                         unsigned splits, typename P::A identity, int has_identity) {
    const size_t col = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (col >= n) return;
    using A = typename P::A;
    const typename P::functor f;
    A r = partials[col];
    for (unsigned s = 1; s < splits; ++s) r = f(r, partials[size_t(s) * n + col]);   /* fixed order, coalesced */
    fused_store<P, S>(out + (ptrdiff_t)col * out_stride, r, identity, has_identity);
}

/* ---- vector variants: a thread owns V adjacent columns per access (see VECTOR LOADS above) ---- */
constexpr int kFusedVecUnroll = 4;      /* vector accesses in flight per thread */

template <typename P, typename S, typename Load, int V>
__global__ void __launch_bounds__(kFusedThreads)
fused_rows_cta_vec_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, unsigned splits,
                          typename P::A* __restrict__ partials, typename P::A identity, int has_identity) {
    using A = typename P::A;
    using T = typename Load::value_type;
    const typename P::functor f;
    size_t seg = (n + splits - 1) / splits;
    seg = (seg + V - 1) / V * V;                  /* splits start on a vector boundary (n % V == 0) */
    const size_t items = rows * splits;
    constexpr size_t kStep = size_t(kFusedThreads) * V;
    for (size_t item = blockIdx.x; item < items; item += gridDim.x) {
        const size_t row = item / splits;
        const unsigned split = unsigned(item % splits);
        const size_t lo = size_t(split) * seg < n ? size_t(split) * seg : n;
        const size_t hi = (lo + seg < n) ? lo + seg : n;
        A acc[kFusedVecUnroll];
#pragma unroll
        for (int u = 0; u < kFusedVecUnroll; ++u) acc[u] = P::seed();
        size_t j = lo + size_t(threadIdx.x) * V;
        for (; j + (kFusedVecUnroll - 1) * kStep < hi; j += kFusedVecUnroll * kStep) {
            T v[kFusedVecUnroll][V];
#pragma unroll
            for (int u = 0; u < kFusedVecUnroll; ++u) load.template vec<V>(row, j + u * kStep, v[u]);
#pragma unroll
            for (int k = 0; k < V; ++k)
#pragma unroll
                for (int u = 0; u < kFusedVecUnroll; ++u) acc[u] = f(acc[u], P::from_value(v[u][k]));
        }
        for (; j < hi; j += kStep) {
            T v[V];
            load.template vec<V>(row, j, v);
#pragma unroll
            for (int k = 0; k < V; ++k) acc[0] = f(acc[0], P::from_value(v[k]));
        }
        A r = block_reduce<A, kFusedThreads, 0x7f01>(acc, f);
        if (threadIdx.x == 0) {
            if (splits == 1) fused_store<P, S>(out + (ptrdiff_t)row * out_stride, r, identity, has_identity);
            else partials[item] = r;
        }
    }
}

/* Short rows: a warp takes R rows at a time, one vector per lane and row in flight (R of them per lane). */
template <typename P, typename S, typename Load, int V, int R>
__global__ void __launch_bounds__(kFusedThreads)
fused_rows_warp_vec_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A identity,
                           int has_identity) {
    using A = typename P::A;
    using T = typename Load::value_type;
    const typename P::functor f;
    const unsigned lane = threadIdx.x & 31;
    const size_t warps_per_grid = size_t(gridDim.x) * (kFusedThreads / 32);
    for (size_t row0 = R * (size_t(blockIdx.x) * (kFusedThreads / 32) + (threadIdx.x >> 5)); row0 < rows;
         row0 += R * warps_per_grid) {
        A acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = P::seed();
        for (size_t j = size_t(lane) * V; j < n; j += 32 * V) {
            T v[R][V];
#pragma unroll
            for (int r = 0; r < R; ++r) load.template vec<V>(row0 + r < rows ? row0 + r : rows - 1, j, v[r]);
#pragma unroll
            for (int k = 0; k < V; ++k)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = f(acc[r], P::from_value(v[r][k]));
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const A res = warp_allreduce(acc[r], f);
            if (lane == 0 && row0 + r < rows) fused_store<P, S>(out + (ptrdiff_t)(row0 + r) * out_stride, res, identity, has_identity);
        }
    }
}

/* Column fold, V adjacent columns per lane: a CTA owns 32 * V columns and one slice of the rows. */
template <typename P, typename S, typename Load, int V>
__global__ void __launch_bounds__(kFusedThreads)
fused_cols_vec_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, unsigned splits,
                      typename P::A* __restrict__ partials, typename P::A identity, int has_identity) {
    using A = typename P::A;
    using T = typename Load::value_type;
    constexpr int kWarps = kFusedThreads / 32;
    constexpr int kRowsInFlight = kFusedVecUnroll;
    __shared__ A tile[kWarps][32 * V + 1];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t col0 = (size_t(blockIdx.x) * 32 + lane) * V;
    const unsigned split = blockIdx.y;
    const size_t seg = (rows + splits - 1) / splits;
    const size_t lo = size_t(split) * seg;
    const size_t hi = (lo + seg < rows) ? lo + seg : rows;
    const typename P::functor f;
    A acc[V];
#pragma unroll
    for (int k = 0; k < V; ++k) acc[k] = P::seed();
    if (col0 < n) {
        size_t r = lo + warp;
        for (; r + size_t(kRowsInFlight - 1) * kWarps < hi; r += size_t(kRowsInFlight) * kWarps) {
            T v[kRowsInFlight][V];
#pragma unroll
            for (int u = 0; u < kRowsInFlight; ++u) load.template vec<V>(r + size_t(u) * kWarps, col0, v[u]);
#pragma unroll
            for (int u = 0; u < kRowsInFlight; ++u)
#pragma unroll
                for (int k = 0; k < V; ++k) acc[k] = f(acc[k], P::from_value(v[u][k]));
        }
        for (; r < hi; r += kWarps) {
            T v[V];
            load.template vec<V>(r, col0, v);
#pragma unroll
            for (int k = 0; k < V; ++k) acc[k] = f(acc[k], P::from_value(v[k]));
        }
    }
#pragma unroll
    for (int k = 0; k < V; ++k) tile[warp][k * 32 + lane] = acc[k];      /* conflict-free: lanes are adjacent */
    __syncthreads();
    if (warp == 0 && col0 < n) {
#pragma unroll
        for (int k = 0; k < V; ++k) {
            A r = tile[0][k * 32 + lane];
#pragma unroll
            for (int w = 1; w < kWarps; ++w) r = f(r, tile[w][k * 32 + lane]);
            if (splits == 1) fused_store<P, S>(out + (ptrdiff_t)(col0 + k) * out_stride, r, identity, has_identity);
            else partials[size_t(split) * n + col0 + k] = r;
        }
    }
}

/* Few rows, many columns: a thread per V adjacent columns. */
template <typename P, typename S, typename Load, int V>
__global__ void __launch_bounds__(kFusedThreads)
fused_cols_thread_vec_kernel(Load load, S* __restrict__ out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A identity,
                             int has_identity) {
    using A = typename P::A;
    using T = typename Load::value_type;
    const size_t col0 = (size_t(blockIdx.x) * kFusedThreads + threadIdx.x) * V;
    if (col0 >= n) return;
    const typename P::functor f;
    A acc[V];
#pragma unroll
    for (int k = 0; k < V; ++k) acc[k] = P::seed();
    size_t r = 0;
    for (; r + kFusedVecUnroll <= rows; r += kFusedVecUnroll) {
        T v[kFusedVecUnroll][V];
#pragma unroll
        for (int u = 0; u < kFusedVecUnroll; ++u) load.template vec<V>(r + u, col0, v[u]);
#pragma unroll
        for (int u = 0; u < kFusedVecUnroll; ++u)
#pragma unroll
            for (int k = 0; k < V; ++k) acc[k] = f(acc[k], P::from_value(v[u][k]));
    }
    for (; r < rows; ++r) {
        T v[V];
        load.template vec<V>(r, col0, v);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = f(acc[k], P::from_value(v[k]));
    }
#pragma unroll
    for (int k = 0; k < V; ++k) fused_store<P, S>(out + (ptrdiff_t)(col0 + k) * out_stride, acc[k], identity, has_identity);
}

inline int fused_sm_count() {
    static int cached = 0;
    if (cached == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached = n;
    }
    return cached;
}

/* CTAs of `Kernel` (kFusedThreads threads, static shared memory only) resident per SM: grids are sized in whole
 * waves of sms * this (a grid of 8 CTAs per SM for a kernel that fits 6 runs 1.33 waves: measured 5.0 instead of
 * 5.8 TB/s on the [65536, 8192] column fold). */
template <auto Kernel>
inline size_t fused_slots() {
    static int occ = 0;
    if (occ == 0) {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, Kernel, kFusedThreads, 0) != cudaSuccess || v < 1) v = 4;
        occ = v;
    }
    return size_t(occ) * size_t(fused_sm_count());
}

/* Pieces per tile so that tiles * pieces CTAs fill whole waves of `slots` as well as possible (at most two waves;
 * ties go to fewer pieces). */
inline size_t fused_pick_splits(size_t tiles, size_t slots, size_t max_splits) {
    if (tiles == 0 || tiles >= slots || max_splits <= 1) return 1;
    size_t limit = 2 * slots / tiles;
    if (limit > max_splits) limit = max_splits;
    size_t best = 1;
    double best_eff = 0.0;
    for (size_t s = 1; s <= limit; ++s) {
        const size_t total = tiles * s, waves = (total + slots - 1) / slots;
        const double eff = double(total) / double(waves * slots);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
    }
    return best;
}

}  // namespace detail

/* Bytes at the start of the dace_b200 workspace that belong to the library's ticket counters
 * (dace_b200/csrc/common.cuh kCounterBytes); the partials area behind them is free between calls. */
constexpr size_t kWorkspaceCounterBytes = 16384;

namespace detail {

template <typename A>
inline A* fused_partials(void* workspace, size_t workspace_bytes, size_t count) {
    if (workspace == nullptr || workspace_bytes < kWorkspaceCounterBytes + count * sizeof(A)) return nullptr;
    return reinterpret_cast<A*>(static_cast<char*>(workspace) + kWorkspaceCounterBytes);
}

/* Rows, a thread owning V adjacent columns per access (V = 1: the scalar kernels). */
template <typename P, typename S, typename Load, int V>
inline cudaError_t fused_rows_launch(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A ident,
                                     int has, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    using A = typename P::A;
    const size_t sms = (size_t)fused_sm_count();
    constexpr size_t kWarps = kFusedThreads / 32;
    if (n < 2048 && rows >= 2 * sms) {                 /* short rows: warps take whole rows */
        if constexpr (V == 1) {
            size_t grid = ((rows + 1) / 2 + kWarps - 1) / kWarps;
            const size_t slots = fused_slots<fused_rows_warp_kernel<P, S, Load>>();
            if (grid > slots) grid = slots;
            fused_rows_warp_kernel<P, S, Load><<<(unsigned)grid, kFusedThreads, 0, stream>>>(load, out, out_stride, rows, n,
                                                                                          ident, has);
        } else if (n <= 64 * V) {                      /* one or two vectors per lane and row: 8 rows at a time */
            size_t grid = ((rows + 7) / 8 + kWarps - 1) / kWarps;
            const size_t slots = fused_slots<fused_rows_warp_vec_kernel<P, S, Load, V, 8>>();
            if (grid > slots) grid = slots;
            fused_rows_warp_vec_kernel<P, S, Load, V, 8><<<(unsigned)grid, kFusedThreads, 0, stream>>>(
                load, out, out_stride, rows, n, ident, has);
        } else {
            size_t grid = ((rows + 3) / 4 + kWarps - 1) / kWarps;
            const size_t slots = fused_slots<fused_rows_warp_vec_kernel<P, S, Load, V, 4>>();
            if (grid > slots) grid = slots;
            fused_rows_warp_vec_kernel<P, S, Load, V, 4><<<(unsigned)grid, kFusedThreads, 0, stream>>>(
                load, out, out_stride, rows, n, ident, has);
        }
        return cudaGetLastError();
    }
    /* a CTA per (row, split), at least 4096 * V elements per split */
    size_t slots;
    if constexpr (V == 1) slots = fused_slots<fused_rows_cta_kernel<P, S, Load>>();
    else slots = fused_slots<fused_rows_cta_vec_kernel<P, S, Load, V>>();
    size_t splits = fused_pick_splits(rows, slots, (n + 4096 * V - 1) / (4096 * V));
    A* partials = nullptr;
    if (splits > 1) {
        partials = fused_partials<A>(workspace, workspace_bytes, rows * splits);
        if (partials == nullptr) splits = 1;               /* still correct, just fewer CTAs */
    }
    /* Many items: a persistent grid of one wave saves the CTA turnover ([65536, 8192] f32: 5.0 -> 5.6 TB/s).  Few,
     * long items: one CTA each, so that the hardware hands out work as CTAs retire - a static split of 8192 items
     * over one wave leaves some CTAs with 7 and some with 6 (measured 6.5 -> 6.0 TB/s). */
    size_t grid = rows * splits;
    if (grid > sms * 64) grid = slots;
    if constexpr (V == 1)
        fused_rows_cta_kernel<P, S, Load><<<(unsigned)grid, kFusedThreads, 0, stream>>>(
            load, out, out_stride, rows, n, (unsigned)splits, partials, ident, has);
    else
        fused_rows_cta_vec_kernel<P, S, Load, V><<<(unsigned)grid, kFusedThreads, 0, stream>>>(
            load, out, out_stride, rows, n, (unsigned)splits, partials, ident, has);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || splits == 1) return e;
    fused_finish_kernel<P, S><<<(unsigned)((rows + 127) / 128), 128, 0, stream>>>(partials, out, out_stride, rows,
                                                                               (unsigned)splits, ident, has);
    return cudaGetLastError();
}

/* Columns, a lane owning V adjacent columns (V = 1: the scalar kernels). */
template <typename P, typename S, typename Load, int V>
inline cudaError_t fused_cols_launch(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A ident,
                                     int has, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    using A = typename P::A;
    const size_t sms = (size_t)fused_sm_count();
    if (rows <= 64 && n >= sms * kFusedThreads * 4 * V) {        /* few rows: a thread per column (group) */
        const unsigned grid = (unsigned)((n / V + kFusedThreads - 1) / kFusedThreads);
        if constexpr (V == 1)
            fused_cols_thread_kernel<P, S, Load><<<grid, kFusedThreads, 0, stream>>>(load, out, out_stride, rows, n, ident, has);
        else
            fused_cols_thread_vec_kernel<P, S, Load, V><<<grid, kFusedThreads, 0, stream>>>(load, out, out_stride, rows, n, ident, has);
        return cudaGetLastError();
    }
    const size_t col_tiles = (n / V + 31) / 32;
    size_t slots;
    if constexpr (V == 1) slots = fused_slots<fused_cols_kernel<P, S, Load>>();
    else slots = fused_slots<fused_cols_vec_kernel<P, S, Load, V>>();
    size_t max_splits = (rows + 63) / 64;                        /* at least 64 rows per slice */
    if (max_splits > 65535) max_splits = 65535;
    const size_t avail = (workspace != nullptr && workspace_bytes > kWorkspaceCounterBytes)
                             ? (workspace_bytes - kWorkspaceCounterBytes) / (n * sizeof(A)) : 0;
    if (max_splits > avail) max_splits = avail;
    const size_t splits = fused_pick_splits(col_tiles, slots, max_splits);
    A* partials = splits > 1 ? reinterpret_cast<A*>(static_cast<char*>(workspace) + kWorkspaceCounterBytes) : nullptr;
    const dim3 grid((unsigned)col_tiles, (unsigned)splits);
    if constexpr (V == 1)
        fused_cols_kernel<P, S, Load><<<grid, kFusedThreads, 0, stream>>>(load, out, out_stride, rows, n, (unsigned)splits,
                                                                       partials, ident, has);
    else
        fused_cols_vec_kernel<P, S, Load, V><<<grid, kFusedThreads, 0, stream>>>(load, out, out_stride, rows, n,
                                                                              (unsigned)splits, partials, ident, has);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || splits == 1) return e;
    fused_cols_finish_kernel<P, S><<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(partials, out, out_stride, n,
                                                                                 (unsigned)splits, ident, has);
    return cudaGetLastError();
}

}  // namespace detail

/* Host launchers.  `identity` is a HOST pointer (or nullptr: fold into `out`).  `workspace` /
 * `workspace_bytes`: a dace_b200 workspace owned by the call site; only needed when rows are split.
 * `max_vec`: the widest vector access (in elements, a power of two <= Load::kVec) that is legal for this call's
 * operands, 0 or 1 for none; ignored for functors without vec().  The launchers use Load::kVec-wide or half as wide
 * accesses (32 / 16 bytes for the generated functors), whichever suits the shape (measured, f32 exp(x - m[r]):
 * rows [524288, 1024] 4.4 scalar / 5.9 half / 7.0 TB/s full; rows of 128 elements 3.3 / 4.2 / 3.2; column folds of
 * 1024-wide arrays 5.5 / 6.4 / 5.4, of 8192-wide arrays 5.0 / 4.1 / 5.8). */
template <typename P, typename S, typename Load>
inline cudaError_t fused_rows_ops(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A ident,
                                  int has, void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using namespace detail;
    if (rows == 0) return cudaSuccess;
    constexpr int VH = load_vec_width<Load>::value, VL = VH / 2;
    if constexpr (VH > 1) {
        if (max_vec >= VH && n % VH == 0 && n >= 64 * VH)
            return fused_rows_launch<P, S, Load, VH>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
        if constexpr (VL > 1) {
            if (max_vec >= VL && n % VL == 0 && n >= 32 * VL)
                return fused_rows_launch<P, S, Load, VL>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
        }
    }
    (void)max_vec;
    return fused_rows_launch<P, S, Load, 1>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
}

/* out[c * out_stride] = fold(op, identity, load(0, c), ..., load(rows - 1, c))      c in [0, n) */
template <typename P, typename S, typename Load>
inline cudaError_t fused_cols_ops(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, typename P::A ident,
                                  int has, void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using namespace detail;
    if (n == 0) return cudaSuccess;
    constexpr int VH = load_vec_width<Load>::value, VL = VH / 2;
    if constexpr (VH > 1) {
        const size_t sms = (size_t)fused_sm_count();
        const bool few_rows = rows <= 64;
        /* wide accesses only for wide arrays (a CTA spans 32 * V columns) and not in the thread-per-column kernel */
        if (!few_rows && max_vec >= VH && n % VH == 0 && n >= 1024 * VH)
            return fused_cols_launch<P, S, Load, VH>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
        if constexpr (VL > 1) {
            const size_t tiles = (n / VL + 31) / 32, slices = (rows + 63) / 64;
            if (max_vec >= VL && n % VL == 0 && n >= 32 * VL && tiles * slices >= 2 * sms)
                return fused_cols_launch<P, S, Load, VL>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
        }
    }
    (void)max_vec;
    return fused_cols_launch<P, S, Load, 1>(load, out, out_stride, rows, n, ident, has, workspace, workspace_bytes, stream);
}

/* Built-in operators (RED = dace::ReductionType value). */
template <int RED, typename A, typename S, typename Load>
inline cudaError_t fused_rows(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, const A* identity,
                              void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using P = builtin_ops<RED, A>;
    return fused_rows_ops<P, S, Load>(load, out, out_stride, rows, n, P::from_identity(identity), identity ? 1 : 0,
                                      workspace, workspace_bytes, stream, max_vec);
}
template <int RED, typename A, typename S, typename Load>
inline cudaError_t fused_cols(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, const A* identity,
                              void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using P = builtin_ops<RED, A>;
    return fused_cols_ops<P, S, Load>(load, out, out_stride, rows, n, P::from_identity(identity), identity ? 1 : 0,
                                      workspace, workspace_bytes, stream, max_vec);
}

/* Custom WCR functor F (default-constructible, `T operator()(const T&, const T&) const`, associative and
 * commutative like every DaCe WCR): what the `B200` expansion instantiates for ReductionType.Custom Reduce
 * nodes, in place of cub::DeviceReduce::Reduce(..., __reduce_<id>(), identity) (reference reduce.py:453-468). */
template <typename F, typename T, typename S, typename Load>
inline cudaError_t fused_rows_custom(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, const T* identity,
                                     void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using P = custom_ops<F, T>;
    return fused_rows_ops<P, S, Load>(load, out, out_stride, rows, n, P::from_identity(identity), identity ? 1 : 0,
                                      workspace, workspace_bytes, stream, max_vec);
}
template <typename F, typename T, typename S, typename Load>
inline cudaError_t fused_cols_custom(Load load, S* out, ptrdiff_t out_stride, size_t rows, size_t n, const T* identity,
                                     void* workspace, size_t workspace_bytes, cudaStream_t stream, int max_vec = 0) {
    using P = custom_ops<F, T>;
    return fused_cols_ops<P, S, Load>(load, out, out_stride, rows, n, P::from_identity(identity), identity ? 1 : 0,
                                      workspace, workspace_bytes, stream, max_vec);
}

}  // namespace dace_b200

#endif /* __CUDACC__ */
#endif /* DACE_B200_FUSED_REDUCE_CUH */
