// Contiguous-span accumulation used by the row kernels: a group of `nthr`
// threads (a sub-warp, a warp or a CTA) folds `n` contiguous elements.
// The span may start at any element-aligned address (memlet offsets land the
// `_in` pointer anywhere, reference dace/libraries/standard/nodes/reduce.py:599-602):
// a scalar head brings the pointer to vector alignment, the body issues U
// independent 16/32-byte streaming loads per thread, a scalar tail finishes.
#pragma once
#include "common.cuh"

namespace db200 {

template <class L, int VB> struct VecInfo {
    static constexpr int N = VB / int(sizeof(L));      // elements per vector
    static constexpr int NA = N < 8 ? N : 8;           // independent accumulators per thread
};

template <class L, class A, class Op, int VB>
__device__ __forceinline__ void fold_vec(A (&acc)[VecInfo<L, VB>::NA], const Words<VB>& w) {
    constexpr int N = VecInfo<L, VB>::N;
    constexpr int NA = VecInfo<L, VB>::NA;
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j % NA] = Op::apply(acc[j % NA], vec_elem<L, A, VB>(w, j));
}

template <class A, class Op, int NA>
__device__ __forceinline__ A fold_acc(const A (&acc)[NA]) {
    A v = acc[0];
#pragma unroll
    for (int j = 1; j < NA; ++j) v = Op::apply(v, acc[j]);
    return v;
}

// SHORT: the caller guarantees at most U vectors per thread (rows shorter than U * 32 bytes * group width): they are
// loaded as one predicated batch, no loop.
template <class L, class A, class Op, int VB, int U, bool SHORT = false>
__device__ __forceinline__ A span_accumulate(const L* __restrict__ p, size_t n, unsigned tid, unsigned nthr) {
    constexpr int N = VecInfo<L, VB>::N;
    constexpr int NA = VecInfo<L, VB>::NA;
    A acc[NA];
#pragma unroll
    for (int j = 0; j < NA; ++j) acc[j] = Op::neutral();

    const size_t addr = reinterpret_cast<size_t>(p);
    size_t h = ((size_t(VB) - (addr & size_t(VB - 1))) & size_t(VB - 1)) / sizeof(L);
    if (h > n) h = n;
    const char* pv = reinterpret_cast<const char*>(p + h);
    const size_t nvec = (n - h) / N;
    size_t v = tid;
    if constexpr (SHORT) {
        // everything in one round trip: the first head and tail elements are requested before the vectors and
        // folded after them (rows that are not whole vectors: head, vectors, tail were three dependent trips)
        const size_t t0 = h + nvec * N;
        const bool has_head = tid < h, has_tail = t0 + tid < n;
        L he{}, te{};
        if (has_head) he = ldg_scalar(p + tid);
        if (has_tail) te = ldg_scalar(p + t0 + tid);
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec) fold_vec<L, A, Op, VB>(acc, w[u]);
        if (has_head) acc[0] = Op::apply(acc[0], to_acc<A, L>(he));
        if (has_tail) acc[0] = Op::apply(acc[0], to_acc<A, L>(te));
        for (size_t i = size_t(tid) + nthr; i < h; i += nthr) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
        for (size_t i = t0 + tid + nthr; i < n; i += nthr) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
        return fold_acc<A, Op, NA>(acc);
    }
    for (size_t i = tid; i < h; i += nthr) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
    for (; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u) fold_vec<L, A, Op, VB>(acc, w[u]);
    }
    // (the few vectors left here are walked one at a time: issuing them as a predicated batch is faster for rows of
    // ~U vectors per thread, but it raises the register count of most instantiations past 64 - three CTAs per SM
    // instead of four - and bf16 [65536, 8192] max went from 0.153 to 0.191 ms; short and ragged rows have the SHORT
    // path above instead)
    for (; v < nvec; v += nthr) {
        Words<VB> w = ldg_stream<VB>(pv + v * VB);
        fold_vec<L, A, Op, VB>(acc, w);
    }
    for (size_t i = h + nvec * N + tid; i < n; i += nthr) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
    return fold_acc<A, Op, NA>(acc);
}

}  // namespace db200
