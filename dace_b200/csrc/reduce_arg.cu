// argmin / argmax (ReductionType Min_Location / Max_Location) for sm_100a.
//
// What it replaces.  The reference frontend lowers np.argmax / np.argmin to a map that
// folds struct values {idx, val} through the custom WCR
//     lambda x, y: _val_and_idx(val=max(x.val, y.val), idx=(y.idx if x.val < y.val else x.idx))
// (reference dace/frontend/python/replacements/reduction.py:261-348).  On a GPU every
// element becomes one compare-and-swap loop on the struct (dace::wcr_custom<T>::reduce_atomic,
// reference dace/runtime/include/dace/reduction.h:56-75); the CUB expansion maps
// Min_Location / Max_Location to cub::DeviceReduce::ArgMin / ArgMax
// (reference dace/libraries/standard/nodes/reduce.py:399-402).
//
// Semantics kept from that lambda, evaluated in index order:
//   * the candidate replaces the running pair only when STRICTLY better, so the first
//     occurrence of the extremum wins (NumPy's rule as well);
//   * the running pair starts from an initial pair (init value, index -1) - the frontend
//     writes (min_value(dtype), -1) for argmax, (max_value(dtype), -1) for argmin - and an
//     element merely EQUAL to the initial value never replaces it, so e.g. argmax over a
//     row of -inf yields index -1 and value -finfo.max, exactly like the reference;
//   * with no initial value given the call folds into the pair already stored in the outputs.
//
// Layout: in[o][r][i] with element strides (so, sr, si); outputs are indexed o * inner + i with
// byte strides, so they may be the fields of the reference's interleaved struct array.
//   * sr == 1 (reduced axis contiguous): G lanes per row for short rows, CTAs (+ row splits,
//     workspace partials and a ticket) for long ones; 256-bit streaming loads, the running
//     maximum of a whole vector is taken first and the position is only searched when it beats
//     the thread's best (O(log n) times per thread);
//   * otherwise (si == 1 is the coalesced case): a thread owns adjacent columns, warps stride
//     the rows, row splits meet in the workspace.
#include "common.cuh"
#include "span.cuh"
#include "kernels.h"

namespace db200 {

constexpr long long kNoIndex = 0x7fffffffffffffffLL;

template <class A, bool IsMax>
struct ArgOp {
    static __device__ __forceinline__ A neutral() { return IsMax ? OpMax<A>::neutral() : OpMin<A>::neutral(); }
    // strictly better than the current best
    static __device__ __forceinline__ bool better(A cand, A best) { return IsMax ? (cand > best) : (cand < best); }
    static __device__ __forceinline__ A pick(A a, A b) { return IsMax ? OpMax<A>::apply(a, b) : OpMin<A>::apply(a, b); }
    // lexicographic combine of two partial results: better value, then lower index
    static __device__ __forceinline__ void combine(A& v, long long& i, A ov, long long oi) {
        if (better(ov, v) || (ov == v && oi < i)) { v = ov; i = oi; }
    }
};

template <class A, bool IsMax, int WIDTH = 32>
__device__ __forceinline__ void warp_arg(A& v, long long& i) {
#pragma unroll
    for (int m = WIDTH / 2; m >= 1; m >>= 1) {
        A ov = shfl_xor(v, m);
        long long oi = shfl_xor(i, m);
        ArgOp<A, IsMax>::combine(v, i, ov, oi);
    }
}

// Where the final pair goes.  idx/val are byte pointers to output 0; strides in bytes.
struct ArgOut {
    char* idx; char* val;
    long long idx_stride, val_stride;
    int idx_bytes;     // 4 or 8
    int has_init;
};

template <class L, class A, bool IsMax>
__device__ __forceinline__ void arg_finish(const ArgOut& o, size_t pos, A v, long long i, A init) {
    A cur_v;
    long long cur_i;
    char* pi = o.idx + (long long)pos * o.idx_stride;
    L* pv = o.val ? reinterpret_cast<L*>(o.val + (long long)pos * o.val_stride) : nullptr;
    if (o.has_init) { cur_v = init; cur_i = -1; }
    else {
        cur_v = to_acc<A, L>(*pv);
        cur_i = o.idx_bytes == 8 ? *reinterpret_cast<long long*>(pi) : (long long)*reinterpret_cast<int*>(pi);
    }
    if (i != kNoIndex && ArgOp<A, IsMax>::better(v, cur_v)) { cur_v = v; cur_i = i; }
    if (o.idx_bytes == 8) *reinterpret_cast<long long*>(pi) = cur_i;
    else *reinterpret_cast<int*>(pi) = (int)cur_i;
    if (pv) *pv = from_acc<L, A>(cur_v);
}

// One vector: running extremum first, position only when it wins.
template <class L, class A, bool IsMax, int VB>
__device__ __forceinline__ void arg_fold_vec(A& best, long long& bidx, const Words<VB>& w, long long first) {
    constexpr int N = VecInfo<L, VB>::N;
    A m = vec_elem<L, A, VB>(w, 0);
#pragma unroll
    for (int j = 1; j < N; ++j) m = ArgOp<A, IsMax>::pick(m, vec_elem<L, A, VB>(w, j));
    if (ArgOp<A, IsMax>::better(m, best)) {
        int at = N - 1;
        best = vec_elem<L, A, VB>(w, N - 1);
#pragma unroll
        for (int j = N - 2; j >= 0; --j) {
            const A e = vec_elem<L, A, VB>(w, j);
            if (e == m) { at = j; best = e; }      // the value follows the index (-0.0 vs +0.0)
        }
        bidx = first + at;
    }
}

// A group of nthr threads scans n contiguous elements starting at p (element index `base` of
// the row); each thread visits its elements in increasing index order.
// SHORT: the caller guarantees at most U vectors per thread (rows of a sub-warp group): they go out as ONE
// predicated batch - the unrolled loop below only starts with U whole vectors per thread, and what it leaves is
// walked one round trip at a time (rows of 128 bytes, two lanes per row: two dependent trips).
template <class L, class A, bool IsMax, int VB, int U, bool SHORT = false>
__device__ __forceinline__ void span_arg(const L* __restrict__ p, size_t n, long long base, unsigned tid, unsigned nthr,
                                         A& best, long long& bidx) {
    constexpr int N = VecInfo<L, VB>::N;
    const size_t addr = reinterpret_cast<size_t>(p);
    size_t h = ((size_t(VB) - (addr & size_t(VB - 1))) & size_t(VB - 1)) / sizeof(L);
    if (h > n) h = n;
    for (size_t i = tid; i < h; i += nthr) {
        A x = to_acc<A, L>(ldg_scalar(p + i));
        if (ArgOp<A, IsMax>::better(x, best)) { best = x; bidx = base + (long long)i; }
    }
    const char* pv = reinterpret_cast<const char*>(p + h);
    const size_t nvec = (n - h) / N;
    const long long vbase = base + (long long)h;
    size_t v = tid;
    if constexpr (SHORT) {
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec)
                arg_fold_vec<L, A, IsMax, VB>(best, bidx, w[u], vbase + (long long)((v + size_t(u) * nthr) * N));
    } else {
        for (; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
            Words<VB> w[U];
#pragma unroll
            for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
            for (int u = 0; u < U; ++u)
                arg_fold_vec<L, A, IsMax, VB>(best, bidx, w[u], vbase + (long long)((v + size_t(u) * nthr) * N));
        }
        for (; v < nvec; v += nthr) {
            Words<VB> w = ldg_stream<VB>(pv + v * VB);
            arg_fold_vec<L, A, IsMax, VB>(best, bidx, w, vbase + (long long)(v * N));
        }
    }
    for (size_t i = h + nvec * N + tid; i < n; i += nthr) {
        A x = to_acc<A, L>(ldg_scalar(p + i));
        if (ArgOp<A, IsMax>::better(x, best)) { best = x; bidx = base + (long long)i; }
    }
}

constexpr int kArgThreads = 256;

// ---- contiguous reduced axis, short rows: G lanes per row -------------------------------------
template <class L, class A, bool IsMax, bool SHORT>
__global__ void __launch_bounds__(kArgThreads)
arg_rows_group_kernel(const L* __restrict__ in, size_t rows, size_t n, size_t inner, long long so, long long si,
                      ArgOut out, A init, int G) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned g = lane & (G - 1);
    const unsigned sub = lane / G;
    const unsigned rows_per_warp = 32 / G;
    const size_t warp = (size_t(blockIdx.x) * kArgThreads + threadIdx.x) >> 5;
    const size_t nwarps = (size_t(gridDim.x) * kArgThreads) >> 5;
    for (size_t rb = warp * rows_per_warp; rb < rows; rb += nwarps * rows_per_warp) {
        const size_t r = rb + sub;
        const bool active = r < rows;
        A v = ArgOp<A, IsMax>::neutral();
        long long i = kNoIndex;
        if (active) {
            const L* p = in + (long long)(r / inner) * so + (long long)(r % inner) * si;
            span_arg<L, A, IsMax, 32, 4, SHORT>(p, n, 0, g, G, v, i);
        }
        for (int m = G >> 1; m >= 1; m >>= 1) {
            A ov = shfl_xor(v, m);
            long long oi = shfl_xor(i, m);
            ArgOp<A, IsMax>::combine(v, i, ov, oi);
        }
        if (active && g == 0) arg_finish<L, A, IsMax>(out, r, v, i, init);
    }
}

template <class A, bool IsMax>
__device__ __forceinline__ void block_arg(A& v, long long& i, A* sv, long long* si) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    warp_arg<A, IsMax>(v, i);
    if (lane == 0) { sv[warp] = v; si[warp] = i; }
    __syncthreads();
    if (warp == 0) {
        A x = lane < nwarps ? sv[lane] : ArgOp<A, IsMax>::neutral();
        long long xi = lane < nwarps ? si[lane] : kNoIndex;
        warp_arg<A, IsMax>(x, xi);
        v = x; i = xi;
    }
}

// ---- contiguous reduced axis, long rows: gridDim.x CTAs share a row ------------------------------
template <class L, class A, bool IsMax, int THREADS, int U>
__global__ void __launch_bounds__(THREADS)
arg_rows_split_kernel(const L* __restrict__ in, size_t rows, size_t n, size_t inner, long long so, long long si,
                      ArgOut out, A init, unsigned* __restrict__ counters, char* __restrict__ partials) {
    __shared__ A sv[32];
    __shared__ long long sidx[32];
    __shared__ int is_last;
    const unsigned splits = gridDim.x;
    for (size_t r = blockIdx.y; r < rows; r += gridDim.y) {
        const L* p = in + (long long)(r / inner) * so + (long long)(r % inner) * si;
        // contiguous chunk per CTA, boundaries on 4 KiB multiples of elements
        size_t chunk = (n + splits - 1) / splits;
        chunk = (chunk + 1023) & ~size_t(1023);
        size_t lo = size_t(blockIdx.x) * chunk;
        size_t hi = lo + chunk < n ? lo + chunk : n;
        A v = ArgOp<A, IsMax>::neutral();
        long long i = kNoIndex;
        if (lo < n) span_arg<L, A, IsMax, 32, U>(p + lo, hi - lo, (long long)lo, threadIdx.x, THREADS, v, i);
        block_arg<A, IsMax>(v, i, sv, sidx);
        if (splits == 1) {
            if (threadIdx.x == 0) arg_finish<L, A, IsMax>(out, r, v, i, init);
            __syncthreads();
            continue;
        }
        A* pv = reinterpret_cast<A*>(partials) + r * splits;
        long long* pi = reinterpret_cast<long long*>(partials + (size_t(gridDim.y) * splits * sizeof(A) + 7) / 8 * 8) + r * splits;
        if (threadIdx.x == 0) {
            pv[blockIdx.x] = v;
            pi[blockIdx.x] = i;
            __threadfence();
            is_last = atomicAdd(&counters[r], 1u) == splits - 1;
        }
        __syncthreads();
        if (is_last && threadIdx.x < 32) {
            __threadfence();
            A x = ArgOp<A, IsMax>::neutral();
            long long xi = kNoIndex;
            for (unsigned s = threadIdx.x; s < splits; s += 32) ArgOp<A, IsMax>::combine(x, xi, __ldcg(pv + s), __ldcg(pi + s));
            warp_arg<A, IsMax>(x, xi);
            if (threadIdx.x == 0) { arg_finish<L, A, IsMax>(out, r, x, xi, init); counters[r] = 0; }
        }
        __syncthreads();
    }
}

// ---- strided reduced axis: thread owns NV adjacent columns, warps stride the rows ------------------
template <class L, class A, bool IsMax, int NV, int U>
__global__ void __launch_bounds__(kArgThreads)
arg_cols_kernel(const L* __restrict__ in, size_t reduced, size_t inner, long long so, long long sr, long long si,
                ArgOut out, A init, unsigned* __restrict__ counters, char* __restrict__ partials, size_t ncolblocks) {
    constexpr int kWarps = kArgThreads / 32;
    constexpr int VB = NV * int(sizeof(L));
    __shared__ A sv[kWarps][32 * NV];
    __shared__ long long sidx[kWarps][32 * NV];
    __shared__ int is_last;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t cb = blockIdx.x % ncolblocks, o = blockIdx.x / ncolblocks;
    const unsigned splits = gridDim.y;
    const size_t col0 = (cb * 32 + lane) * NV;
    const bool active = col0 < inner;      // NV > 1 only when inner % NV == 0
    size_t chunk = (reduced + splits - 1) / splits;
    const size_t r0 = size_t(blockIdx.y) * chunk;
    const size_t r1 = r0 + chunk < reduced ? r0 + chunk : reduced;

    A v[NV];
    long long ix[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) { v[j] = ArgOp<A, IsMax>::neutral(); ix[j] = kNoIndex; }
    if (active) {
        const L* p = in + (long long)o * so + (long long)col0 * si;
        size_t r = r0 + warp;
        if constexpr (NV > 1) {
            for (; r + size_t(U - 1) * kWarps < r1; r += size_t(U) * kWarps) {
                Words<VB> w[U];
#pragma unroll
                for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(p + (long long)(r + size_t(u) * kWarps) * sr);
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int j = 0; j < NV; ++j) {
                        A x = vec_elem<L, A, VB>(w[u], j);
                        if (ArgOp<A, IsMax>::better(x, v[j])) { v[j] = x; ix[j] = (long long)(r + size_t(u) * kWarps); }
                    }
            }
            for (; r < r1; r += kWarps) {
                Words<VB> w = ldg_stream<VB>(p + (long long)r * sr);
#pragma unroll
                for (int j = 0; j < NV; ++j) {
                    A x = vec_elem<L, A, VB>(w, j);
                    if (ArgOp<A, IsMax>::better(x, v[j])) { v[j] = x; ix[j] = (long long)r; }
                }
            }
        } else {
            for (; r + size_t(U - 1) * kWarps < r1; r += size_t(U) * kWarps) {
                L w[U];
#pragma unroll
                for (int u = 0; u < U; ++u) w[u] = ldg_scalar(p + (long long)(r + size_t(u) * kWarps) * sr);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    A x = to_acc<A, L>(w[u]);
                    if (ArgOp<A, IsMax>::better(x, v[0])) { v[0] = x; ix[0] = (long long)(r + size_t(u) * kWarps); }
                }
            }
            for (; r < r1; r += kWarps) {
                A x = to_acc<A, L>(ldg_scalar(p + (long long)r * sr));
                if (ArgOp<A, IsMax>::better(x, v[0])) { v[0] = x; ix[0] = (long long)r; }
            }
        }
    }
    // across the warps of the CTA (different rows, same columns)
#pragma unroll
    for (int j = 0; j < NV; ++j) { sv[warp][lane * NV + j] = v[j]; sidx[warp][lane * NV + j] = ix[j]; }
    __syncthreads();
    const size_t ncols_cta = 32 * NV;
    for (unsigned c = threadIdx.x; c < ncols_cta; c += kArgThreads) {
        A x = sv[0][c];
        long long xi = sidx[0][c];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) ArgOp<A, IsMax>::combine(x, xi, sv[w][c], sidx[w][c]);
        const size_t col = cb * ncols_cta + c;
        if (col < inner) {
            if (splits == 1) arg_finish<L, A, IsMax>(out, o * inner + col, x, xi, init);
            else {
                const size_t slot = (size_t(blockIdx.x) * splits + blockIdx.y) * ncols_cta + c;
                reinterpret_cast<A*>(partials)[slot] = x;
                reinterpret_cast<long long*>(partials + (size_t(gridDim.x) * splits * ncols_cta * sizeof(A) + 7) / 8 * 8)[slot] = xi;
            }
        }
    }
    if (splits == 1) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&counters[blockIdx.x], 1u) == splits - 1;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    const A* pv = reinterpret_cast<const A*>(partials);
    const long long* pi = reinterpret_cast<const long long*>(partials + (size_t(gridDim.x) * splits * ncols_cta * sizeof(A) + 7) / 8 * 8);
    for (unsigned c = threadIdx.x; c < ncols_cta; c += kArgThreads) {
        const size_t col = cb * ncols_cta + c;
        if (col >= inner) continue;
        A x = ArgOp<A, IsMax>::neutral();
        long long xi = kNoIndex;
        for (unsigned s = 0; s < splits; ++s) {
            const size_t slot = (size_t(blockIdx.x) * splits + s) * ncols_cta + c;
            ArgOp<A, IsMax>::combine(x, xi, __ldcg(pv + slot), __ldcg(pi + slot));
        }
        arg_finish<L, A, IsMax>(out, o * inner + col, x, xi, init);
    }
    if (threadIdx.x == 0) counters[blockIdx.x] = 0;
}

// Few rows, many columns ([R <= 64, K], argmax over the R axis): a thread owns NV adjacent columns and walks all
// rows itself, CTAs walk column chunks - no shared memory, no cross-warp combine, nothing to split.  The kernel
// above would launch one CTA per 32 * NV columns with one row per warp (measured 1.1 TB/s at [8, 33M] f32).
template <class L, class A, bool IsMax, int NV, int U>
__global__ void __launch_bounds__(kArgThreads)
arg_cols_few_rows_kernel(const L* __restrict__ in, size_t outer, size_t reduced, size_t inner, long long so, long long sr,
                         ArgOut out, A init) {
    constexpr int VB = NV * int(sizeof(L));
    const size_t nvec = (inner + NV - 1) / NV;           // NV > 1 only when inner % NV == 0
    const size_t per_o = (nvec + kArgThreads - 1) / kArgThreads;      // CTA-sized chunks per outer slice
    for (size_t chunk = blockIdx.x; chunk < per_o * outer; chunk += gridDim.x) {
        const size_t o = chunk / per_o, cv = (chunk % per_o) * kArgThreads + threadIdx.x;
        if (cv >= nvec) continue;
        const size_t col0 = cv * NV;
        const L* p = in + (long long)o * so + (long long)col0;
        A v[NV];
        long long ix[NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) { v[j] = ArgOp<A, IsMax>::neutral(); ix[j] = kNoIndex; }
        size_t r = 0;
        for (; r + U <= reduced; r += U) {
            if constexpr (NV > 1) {
                Words<VB> w[U];
#pragma unroll
                for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(p + (long long)(r + u) * sr);
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int j = 0; j < NV; ++j) {
                        const A x = vec_elem<L, A, VB>(w[u], j);
                        if (ArgOp<A, IsMax>::better(x, v[j])) { v[j] = x; ix[j] = (long long)(r + u); }
                    }
            } else {
                L w[U];
#pragma unroll
                for (int u = 0; u < U; ++u) w[u] = ldg_scalar(p + (long long)(r + u) * sr);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const A x = to_acc<A, L>(w[u]);
                    if (ArgOp<A, IsMax>::better(x, v[0])) { v[0] = x; ix[0] = (long long)(r + u); }
                }
            }
        }
        for (; r < reduced; ++r) {
            if constexpr (NV > 1) {
                Words<VB> w = ldg_stream<VB>(p + (long long)r * sr);
#pragma unroll
                for (int j = 0; j < NV; ++j) {
                    const A x = vec_elem<L, A, VB>(w, j);
                    if (ArgOp<A, IsMax>::better(x, v[j])) { v[j] = x; ix[j] = (long long)r; }
                }
            } else {
                const A x = to_acc<A, L>(ldg_scalar(p + (long long)r * sr));
                if (ArgOp<A, IsMax>::better(x, v[0])) { v[0] = x; ix[0] = (long long)r; }
            }
        }
#pragma unroll
        for (int j = 0; j < NV; ++j)
            if (col0 + j < inner) arg_finish<L, A, IsMax>(out, o * inner + col0 + j, v[j], ix[j], init);
    }
}

template <class L, class A, bool IsMax>
static cudaError_t launch_arg_t(const ArgArgs& a_in, const Launch& l) {
    ArgOut out;
    out.idx = (char*)a_in.out_idx;
    out.val = (char*)a_in.out_val;
    out.idx_bytes = a_in.idx_bytes;
    out.idx_stride = a_in.out_idx_stride_bytes ? a_in.out_idx_stride_bytes : a_in.idx_bytes;
    out.val_stride = a_in.out_val_stride_bytes ? a_in.out_val_stride_bytes : (long long)sizeof(L);
    out.has_init = a_in.init != nullptr;
    Finish<A> f = make_finish<A, L>(a_in.init);
    const A init = f.identity;
    const size_t sms = sm_count();
    ArgArgs a = a_in;
    if (a.inner == 1 && a.outer > 1 && a.in_stride_r != 1) {
        // [outer, reduced] with a strided reduced axis (e.g. a transposed view): the kept axis is the
        // column axis of the strided kernel; output positions o * inner + i are unchanged.
        a.inner = a.outer; a.in_stride_i = a.in_stride_o;
        a.outer = 1; a.in_stride_o = 0;
    }
    const size_t nout = a.outer * a.inner;

    if (a.in_stride_r == 1 || a.reduced <= 1) {
        const size_t rows = nout, n = a.reduced;
        const size_t row_bytes = n * sizeof(L);
        // a warp per row (or less) also for rows up to 256 KiB when there are enough rows for every warp of the
        // machine: a CTA per 32 KiB row left each thread with two loads and a block-wide combine per row
        // (f64 [32768, 4096]: 3.2 -> 6.x TB/s)
        if (row_bytes < (size_t(32) << 10) || (row_bytes <= (size_t(256) << 10) && rows >= sms * 32)) {
            int G = 1;
            const size_t per_lane = tuning().rows_lane_bytes > 0 ? size_t(tuning().rows_lane_bytes) : 128;   // as in reduce_rows.cu
            while (G < 32 && size_t(G) * per_lane < row_bytes) G <<= 1;
            const size_t rows_per_cta = size_t(kArgThreads / 32) * (32 / G);
            size_t grid = (rows + rows_per_cta - 1) / rows_per_cta;
            if (row_bytes / 32 <= size_t(G) * 4) {            // at most four vectors per lane: one predicated batch
                size_t cap = size_t(sms) * resident_ctas<arg_rows_group_kernel<L, A, IsMax, true>>(kArgThreads);
                if (grid > cap) grid = cap;
                if (grid == 0) grid = 1;
                arg_rows_group_kernel<L, A, IsMax, true><<<(unsigned)grid, kArgThreads, 0, l.stream>>>(
                    (const L*)a.in, rows, n, a.inner, a.in_stride_o, a.in_stride_i, out, init, G);
                return cudaGetLastError();
            }
            const size_t cap = size_t(sms) * resident_ctas<arg_rows_group_kernel<L, A, IsMax, false>>(kArgThreads);
            if (grid > cap) grid = cap;
            if (grid == 0) grid = 1;
            arg_rows_group_kernel<L, A, IsMax, false><<<(unsigned)grid, kArgThreads, 0, l.stream>>>(
                (const L*)a.in, rows, n, a.inner, a.in_stride_o, a.in_stride_i, out, init, G);
            return cudaGetLastError();
        }
        // long rows: enough CTAs to fill the machine, at least 64 KiB per CTA
        size_t gy = rows < 65535 ? rows : 65535;
        size_t splits = 1;
        if (rows < 2 * sms && rows <= kCounters) {
            splits = (2 * sms + rows - 1) / rows;
            size_t most = (row_bytes + (size_t(64) << 10) - 1) / (size_t(64) << 10);
            if (splits > most) splits = most;
            while (splits > 1 && rows * splits * 16 + 8 > kPartialBytes) --splits;
            if (splits < 1) splits = 1;
        }
        if (row_bytes * rows >= (size_t(256) << 20) && splits * rows <= sms) {
            // very large: one 1024-thread CTA per SM like the split sum kernel
            arg_rows_split_kernel<L, A, IsMax, 1024, 8><<<dim3((unsigned)splits, (unsigned)gy), 1024, 0, l.stream>>>(
                (const L*)a.in, rows, n, a.inner, a.in_stride_o, a.in_stride_i, out, init, l.ws.counters,
                (char*)l.ws.partials);
        } else {
            arg_rows_split_kernel<L, A, IsMax, 512, 4><<<dim3((unsigned)splits, (unsigned)gy), 512, 0, l.stream>>>(
                (const L*)a.in, rows, n, a.inner, a.in_stride_o, a.in_stride_i, out, init, l.ws.counters,
                (char*)l.ws.partials);
        }
        return cudaGetLastError();
    }

    // strided reduced axis
    constexpr int NVmax = sizeof(L) == 1 ? 8 : 16 / int(sizeof(L));   // 8- or 16-byte loads
    constexpr size_t kVecBytes = size_t(NVmax) * sizeof(L);
    const bool vec = a.in_stride_i == 1 && a.inner % NVmax == 0 && a.in_stride_r % NVmax == 0 &&
                     a.in_stride_o % NVmax == 0 && (reinterpret_cast<size_t>(a.in) & (kVecBytes - 1)) == 0;
    const size_t nv = vec ? NVmax : 1;
    if (a.reduced <= 64 && a.in_stride_i == 1 && a.inner >= sms * kArgThreads) {
        const size_t chunks = ((a.inner + nv - 1) / nv + kArgThreads - 1) / kArgThreads * a.outer;
        if (vec) {
            size_t grid = sms * resident_ctas<arg_cols_few_rows_kernel<L, A, IsMax, NVmax, 4>>(kArgThreads);
            if (grid > chunks) grid = chunks;
            arg_cols_few_rows_kernel<L, A, IsMax, NVmax, 4><<<(unsigned)grid, kArgThreads, 0, l.stream>>>(
                (const L*)a.in, a.outer, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, out, init);
        } else {
            size_t grid = sms * resident_ctas<arg_cols_few_rows_kernel<L, A, IsMax, 1, 8>>(kArgThreads);
            if (grid > chunks) grid = chunks;
            arg_cols_few_rows_kernel<L, A, IsMax, 1, 8><<<(unsigned)grid, kArgThreads, 0, l.stream>>>(
                (const L*)a.in, a.outer, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, out, init);
        }
        return cudaGetLastError();
    }
    const size_t ncolblocks = (a.inner + 32 * nv - 1) / (32 * nv);
    const size_t gx = ncolblocks * a.outer;
    if (gx == 0) return cudaSuccess;
    if (gx > 0x7fffffffULL) return cudaErrorNotSupported;
    size_t splits = 1;
    const size_t slots = sms * size_t(vec ? resident_ctas<arg_cols_kernel<L, A, IsMax, NVmax, 4>>(kArgThreads)
                                          : resident_ctas<arg_cols_kernel<L, A, IsMax, 1, 8>>(kArgThreads));
    if (gx < slots && gx <= kCounters) {
        // CTAs of equal work: gx * splits should fill whole waves of what fits the machine (19 splits of 32 column
        // blocks were 608 CTAs on 592 slots - a second wave for 16 of them: [65536, 4096] f32 at 0.68 of the peak)
        size_t most = (a.reduced + 63) / 64;        // at least 64 rows per CTA
        if (most > 65535) most = 65535;
        while (most > 1 && gx * most * 32 * nv * 16 + 8 > kPartialBytes) --most;
        double best_eff = 0.0;
        for (size_t sp = 1; sp <= most && gx * sp <= 2 * slots; ++sp) {
            const size_t total = gx * sp, waves = (total + slots - 1) / slots;
            const double eff = double(total) / double(waves * slots);
            if (eff > best_eff + 1e-9) { best_eff = eff; splits = sp; }
        }
    }
    dim3 grid((unsigned)gx, (unsigned)splits);
    if (vec)
        arg_cols_kernel<L, A, IsMax, NVmax, 4><<<grid, kArgThreads, 0, l.stream>>>(
            (const L*)a.in, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.in_stride_i, out, init, l.ws.counters,
            (char*)l.ws.partials, ncolblocks);
    else
        arg_cols_kernel<L, A, IsMax, 1, 8><<<grid, kArgThreads, 0, l.stream>>>(
            (const L*)a.in, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.in_stride_i, out, init, l.ws.counters,
            (char*)l.ws.partials, ncolblocks);
    return cudaGetLastError();
}

cudaError_t launch_arg(int red, int dtin, const ArgArgs& a, const Launch& l) {
    if (red != RT_MinLoc && red != RT_MaxLoc) return cudaErrorNotSupported;
    if (a.idx_bytes != 4 && a.idx_bytes != 8) return cudaErrorNotSupported;
#define DB200_ARG(DT, L, A) \
    if (dtin == DT) return red == RT_MaxLoc ? launch_arg_t<L, A, true>(a, l) : launch_arg_t<L, A, false>(a, l);
    DB200_ARG(DT_F32, float, float)
    DB200_ARG(DT_F64, double, double)
    DB200_ARG(DT_I32, int, int)
    DB200_ARG(DT_I64, long long, long long)
    DB200_ARG(DT_U32, unsigned, unsigned)
    DB200_ARG(DT_U64, unsigned long long, unsigned long long)
    DB200_ARG(DT_BF16, __nv_bfloat16, float)
    DB200_ARG(DT_F16, __half, float)
    DB200_ARG(DT_I8, signed char, int)
    DB200_ARG(DT_U8, unsigned char, unsigned)
    DB200_ARG(DT_I16, short, int)
    DB200_ARG(DT_U16, unsigned short, unsigned)
#undef DB200_ARG
    return cudaErrorNotSupported;
}

}  // namespace db200
