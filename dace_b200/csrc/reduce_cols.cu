// Column (strided-axis) reduction for sm_100a: in[o][r][i] with i contiguous,
// reduce over r, out[o][i].
//
// Replaces the GPUAuto strided schedule - (16,32) blocks, 32 output columns per
// block, shared-memory WCR, grid = ceil(C/32) blocks only
// (reference dace/libraries/standard/nodes/reduce.py:1141-1332,
// reference dace/libraries/standard/reduction_planner.py:269-284); the CUB
// expansion cannot do this layout at all and falls back to a sequential CPU-style
// loop (reference dace/libraries/standard/nodes/reduce.py:493-496).
//
// Each thread owns one 256-bit vector of adjacent columns and strides the rows
// with U independent loads in flight; P lanes span the columns, 32/P lanes take
// different rows (so narrow matrices still issue full-warp coalesced requests).
// Rows are split over gridDim.x; the CTAs of a cluster (<= 8 along x) combine in
// DSMEM, clusters combine through workspace partials + a ticket, folded in a
// fixed order (deterministic, no float atomics, no second kernel).
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kColWarps = 8;

template <class L, class A, int N>
__device__ __forceinline__ void load_cols(const L* __restrict__ p, A (&x)[N]) {
    if constexpr (N == 1) {
        x[0] = to_acc<A, L>(ldg_scalar(p));
    } else {
        constexpr int VB = N * int(sizeof(L));
        Words<VB> w = ldg_stream<VB>(p);
#pragma unroll
        for (int j = 0; j < N; ++j) x[j] = vec_elem<L, A, VB>(w, j);
    }
}

template <class L, class A, class S, class Op, int N, int U>
__global__ void __launch_bounds__(32 * kColWarps)
cols_kernel(const L* __restrict__ in, S* __restrict__ out, size_t outer, size_t reduced, size_t inner,
            size_t so, size_t sr, ptrdiff_t oo, ptrdiff_t oi, Finish<A> fin,
            unsigned* __restrict__ counters, A* __restrict__ partials,
            int P, size_t colgroups, size_t rows_per_split) {
    __shared__ A sm[kColWarps][32 * N];

    const unsigned lane = threadIdx.x;
    const unsigned warp = threadIdx.y;
    const size_t yz = size_t(blockIdx.z) * gridDim.y + blockIdx.y;
    if (yz >= colgroups * outer) return;  // uniform over the cluster (cluster spans x only)
    const size_t o = yz / colgroups;
    const size_t cgi = yz % colgroups;
    const unsigned cl = lane & (P - 1);            // column slot
    const unsigned rs = lane / P;                  // row slot inside the warp
    const unsigned rpw = 32 / P;
    const size_t c0 = (cgi * P + cl) * N;
    const bool col_ok = c0 < inner;

    const size_t r_begin = size_t(blockIdx.x) * rows_per_split;
    size_t r_end = r_begin + rows_per_split;
    if (r_end > reduced) r_end = reduced;
    const size_t step = size_t(kColWarps) * rpw;

    A acc[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = Op::neutral();

    if (col_ok) {
        const L* base = in + o * so + c0;
        size_t r = r_begin + size_t(warp) * rpw + rs;
        for (; r + size_t(U - 1) * step < r_end; r += size_t(U) * step) {
            A x[U][N];
#pragma unroll
            for (int u = 0; u < U; ++u) load_cols<L, A, N>(base + (r + size_t(u) * step) * sr, x[u]);
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], x[u][j]);
        }
        for (; r < r_end; r += step) {
            A x[N];
            load_cols<L, A, N>(base + r * sr, x);
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], x[j]);
        }
    }
    // Lanes that share a column slot (different row slots) fold through shuffles.
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = warp_reduce_strided<Op, A>(acc[j], P);

    if (lane < (unsigned)P) {
#pragma unroll
        for (int j = 0; j < N; ++j) sm[warp][lane * N + j] = acc[j];
    }
    __syncthreads();
    const bool owner = (warp == 0) && (lane < (unsigned)P);
    if (owner) {
        for (int w = 1; w < kColWarps; ++w)
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], sm[w][lane * N + j]);
    }

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks();
    const unsigned crank = cluster.block_rank();
    if (csize > 1) {
        cluster.sync();  // every CTA is done reading its own sm[][] -> slots may be overwritten
        if (owner && crank != 0) {
            A* dst = cluster.map_shared_rank(&sm[0][0], 0) + size_t(crank) * (32 * N);
#pragma unroll
            for (int j = 0; j < N; ++j) dst[lane * N + j] = acc[j];
        }
        cluster.sync();
        if (crank != 0) return;
        if (owner) {
            for (unsigned k = 1; k < csize; ++k)
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], sm[k][lane * N + j]);
        }
    }
    if (warp != 0) return;

    const unsigned nclusters = gridDim.x / csize;
    S* obase = out + ptrdiff_t(o) * oo;
    if (nclusters == 1) {
        if (owner && col_ok) {
#pragma unroll
            for (int j = 0; j < N; ++j) write_result<Op, A, S>(obase + ptrdiff_t(c0 + j) * oi, acc[j], fin);
        }
        return;
    }
    // Cross-cluster: partials[yz][cluster][P*N] + ticket, last cluster folds in order.
    const size_t slot = size_t(P) * N;
    A* my = partials + (yz * nclusters + blockIdx.x / csize) * slot;
    if (owner) {
#pragma unroll
        for (int j = 0; j < N; ++j) my[lane * N + j] = acc[j];
        __threadfence();
    }
    __syncwarp();
    unsigned ticket = 0;
    if (lane == 0) ticket = atomicAdd(&counters[yz], 1u);
    ticket = __shfl_sync(kFullMask, ticket, 0);
    if (ticket != nclusters - 1) return;
    __threadfence();
    if (owner && col_ok) {
        const A* first = partials + yz * nclusters * slot;
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] = __ldcg(first + lane * N + j);
        for (unsigned k = 1; k < nclusters; ++k)
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], __ldcg(first + k * slot + lane * N + j));
#pragma unroll
        for (int j = 0; j < N; ++j) write_result<Op, A, S>(obase + ptrdiff_t(c0 + j) * oi, acc[j], fin);
    }
    if (lane == 0) counters[yz] = 0;
}

// ---------------------------------------------------------------------------
// Wide rows (>= 4 KiB of contiguous columns, >= 1024 rows; BASELINE C5): one 1024-thread CTA per SM.
// WC = 4 warps lie side by side along the columns (a warp covers 32 vectors = 1 KiB, a CTA 4 KiB
// of each row), the other WR = 8 warps take consecutive rows, and every thread keeps U = 8 rows
// in flight: 64 rows x 4 KiB = 256 KiB per SM outstanding in 4 KiB contiguous pieces (the split
// kernel's recipe; 1 KiB pieces in cols_kernel reach 6.7 TB/s, these 7.15).  Row splits go over gridDim.x; their partials meet in the workspace and the last CTA
// of a column group (ticket) folds them in split order - deterministic, one launch.
// ---------------------------------------------------------------------------
constexpr int kWideThreads = 1024;
constexpr int kWideMaxGroups = 16;

// op-fold of `count` values `stride` apart, in index order, eight L2 loads in flight.
template <class A, class Op>
__device__ __forceinline__ A fold_in_order(const A* p, size_t stride, unsigned count) {
    A v = __ldcg(p);
    unsigned k = 1;
    for (; k + 8 <= count; k += 8) {
        A t[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) t[q] = __ldcg(p + size_t(k + q) * stride);
#pragma unroll
        for (int q = 0; q < 8; ++q) v = Op::apply(v, t[q]);
    }
    for (; k < count; ++k) v = Op::apply(v, __ldcg(p + size_t(k) * stride));
    return v;
}

constexpr int kWideWarps = kWideThreads / 32;

template <class L, class A, class S, class Op, int N, int U>
__global__ void __launch_bounds__(kWideThreads, 1)
cols_wide_kernel(const L* __restrict__ in, S* __restrict__ out, size_t outer, size_t reduced, size_t inner,
                 size_t so, size_t sr, ptrdiff_t oo, ptrdiff_t oi, Finish<A> fin,
                 unsigned* __restrict__ counters, A* __restrict__ partials,
                 int WC, size_t colgroups, size_t rows_per_split, unsigned gsize) {
    extern __shared__ __align__(16) unsigned char wide_smem[];
    A* sm = reinterpret_cast<A*>(wide_smem);   // [kWideWarps][32 * N] when WR > 1
    __shared__ unsigned s_ticket;

    const unsigned tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    const unsigned WR = kWideWarps / unsigned(WC);
    const unsigned wc = warp % unsigned(WC), wr = warp / unsigned(WC);
    const size_t yz = size_t(blockIdx.z) * gridDim.y + blockIdx.y;
    if (yz >= colgroups * outer) return;
    const size_t o = yz / colgroups, cgi = yz % colgroups;
    const size_t c0 = ((cgi * WC + wc) * 32 + lane) * N;
    const bool col_ok = c0 < inner;

    const size_t r_begin = size_t(blockIdx.x) * rows_per_split;
    size_t r_end = r_begin + rows_per_split;
    if (r_end > reduced) r_end = reduced;

    A acc[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = Op::neutral();
    if (col_ok) {
        const L* base = in + o * so + c0;
        size_t r = r_begin + wr;
        for (; r + size_t(U - 1) * WR < r_end; r += size_t(U) * WR) {
            A x[U][N];
#pragma unroll
            for (int u = 0; u < U; ++u) load_cols<L, A, N>(base + (r + size_t(u) * WR) * sr, x[u]);
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], x[u][j]);
        }
        for (; r < r_end; r += WR) {
            A x[N];
            load_cols<L, A, N>(base + r * sr, x);
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], x[j]);
        }
    }
    // warps that share a column chunk (different row slots) meet in shared memory
    if (WR > 1) {
#pragma unroll
        for (int j = 0; j < N; ++j) sm[(warp * 32 + lane) * N + j] = acc[j];
        __syncthreads();
        if (wr == 0) {
            for (unsigned k = 1; k < WR; ++k)
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], sm[((k * WC + wc) * 32 + lane) * N + j]);
        }
    }
    const bool owner = wr == 0;
    const unsigned nsplits = gridDim.x;
    S* obase = out + ptrdiff_t(o) * oo;
    if (nsplits == 1) {
        if (owner && col_ok) {
#pragma unroll
            for (int j = 0; j < N; ++j) write_result<Op, A, S>(obase + ptrdiff_t(c0 + j) * oi, acc[j], fin);
        }
        return;
    }
    // Two-level, fixed-order combine through the workspace: splits form groups of `gsize`; the last
    // CTA of a group (ticket) folds the group's partials in split order into a group partial, the
    // last group folds the group partials in group order.  One CTA folding all ~148 partials would
    // be bound by a single SM's L2 bandwidth and latency (~45 us measured on C5); two levels of ~12
    // cost a few microseconds.
    const size_t slot = size_t(WC) * 32 * N;
    const unsigned ngroups = (nsplits + gsize - 1) / gsize;
    const unsigned grp = blockIdx.x / gsize;
    const unsigned in_grp = min(gsize, nsplits - grp * gsize);
    A* level1 = partials + yz * (size_t(nsplits) + ngroups) * slot;   // [nsplits][slot]
    A* level2 = level1 + size_t(nsplits) * slot;                        // [ngroups][slot]
    unsigned* ctr = counters + yz * (1 + kWideMaxGroups);               // [0] top, [1 + g] group g
    if (owner) {
        A* my = level1 + size_t(blockIdx.x) * slot + (size_t(wc) * 32 + lane) * N;
#pragma unroll
        for (int j = 0; j < N; ++j) my[j] = acc[j];
        __threadfence();
    }
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(&ctr[1 + grp], 1u);
    __syncthreads();
    if (s_ticket != in_grp - 1) return;
    __threadfence();
    const bool single_group = ngroups == 1;
    const size_t cbase = cgi * WC * 32 * N;
    for (size_t idx = tid; idx < slot; idx += kWideThreads) {
        const A v = fold_in_order<A, Op>(level1 + size_t(grp) * gsize * slot + idx, slot, in_grp);
        if (single_group) {
            if (cbase + idx < inner) write_result<Op, A, S>(obase + ptrdiff_t(cbase + idx) * oi, v, fin);
        } else {
            level2[size_t(grp) * slot + idx] = v;
        }
    }
    if (tid == 0) ctr[1 + grp] = 0;
    if (single_group) return;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(&ctr[0], 1u);
    __syncthreads();
    if (s_ticket != ngroups - 1) return;
    __threadfence();
    for (size_t idx = tid; idx < slot; idx += kWideThreads) {
        if (cbase + idx >= inner) continue;
        const A v = fold_in_order<A, Op>(level2 + idx, slot, ngroups);
        write_result<Op, A, S>(obase + ptrdiff_t(cbase + idx) * oi, v, fin);
    }
    if (tid == 0) ctr[0] = 0;
}

static inline size_t ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }

template <class T, class Op, int N>
static cudaError_t launch_cols_n(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    constexpr int U = 4;
    const size_t colvecs = ceil_div(a.inner, N);
    int P = 1;
    while (P < 32 && size_t(P) < colvecs) P <<= 1;
    const size_t colgroups = ceil_div(colvecs, P);
    const size_t total = colgroups * a.outer;
    const size_t rows_per_step = size_t(kColWarps) * (32 / P);

    // Row splits: fill ~8 CTAs (2048 threads) per SM, keep >= 2 unrolled steps per CTA.
    const size_t target = size_t(sm_count()) * 8;
    size_t splits = ceil_div(target, total);
    const size_t max_by_rows = a.reduced / (rows_per_step * U * 2);
    if (splits > max_by_rows) splits = max_by_rows;
    if (splits < 1) splits = 1;
    unsigned cluster = 1;
    if (splits > 8) {
        // Needs the workspace: bounded by ticket counters and the partials area.
        const size_t slot_bytes = size_t(P) * N * sizeof(A);
        if (total > kCounters) splits = 8;
        else {
            size_t max_clusters = kPartialBytes / (slot_bytes * total);
            if (max_clusters < 1) max_clusters = 1;
            if (splits > max_clusters * 8) splits = max_clusters * 8;
        }
        cluster = 8;
        splits = (splits / 8) * 8;
        if (splits < 8) splits = 8;
    } else {
        size_t p2 = 1;
        while (p2 * 2 <= splits) p2 <<= 1;
        splits = p2;                 // 1, 2, 4 or 8: the whole split dimension is one cluster
        cluster = (unsigned)splits;
    }
    const size_t rows_per_split = ceil_div(a.reduced, splits);

    size_t gy = total, gz = 1;
    if (gy > 65535) { gz = ceil_div(total, 65535); gy = 65535; }
    if (gz > 65535) return cudaErrorInvalidValue;

    auto kern = cols_kernel<L, A, S, Op, N, U>;
    if (cluster > 8) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)splits, (unsigned)gy, (unsigned)gz);
    cfg.blockDim = dim3(32, kColWarps, 1);
    cfg.stream = l.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    Finish<A> fin = make_finish<A, S>(a.identity);
    return cudaLaunchKernelEx(&cfg, kern, (const L*)a.in, (S*)a.out, a.outer, a.reduced, a.inner,
                              a.in_stride_o, a.in_stride_r, a.out_stride_o, a.out_stride_i, fin,
                              l.ws.counters, (A*)l.ws.partials, P, colgroups, rows_per_split);
}

// Returns cudaErrorNotSupported when the shape is not a wide-row one (caller uses cols_kernel).
template <class T, class Op, int N>
static cudaError_t launch_cols_wide(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    constexpr int U = 8;
    if (tuning().no_wide_cols) return cudaErrorNotSupported;
    const size_t colvecs = ceil_div(a.inner, N);
    if (colvecs < 128) return cudaErrorNotSupported;                 // rows narrower than 4 KiB
    // 4 warps (4 KiB) side by side measured best or within 1 % of best from 16 MiB to 8 GiB (tools/cols_sweep.py):
    // wider CTAs only make the partials bigger.
    int wc_max = 4;
    if (tuning().wide_wc > 0) wc_max = tuning().wide_wc;   // tuning knob for experiments (read at load)
    int WC = 4;
    while (WC < wc_max && size_t(WC) * 32 < colvecs) WC <<= 1;
    const unsigned WR = kWideWarps / WC;
    const size_t colgroups = ceil_div(colvecs, size_t(WC) * 32);
    const size_t total = colgroups * a.outer;
    if (a.reduced < 1024 || a.reduced < size_t(WR) * U * 4) return cudaErrorNotSupported;  // too few rows to stream
    size_t splits = size_t(sm_count()) / total;   // one wave: never more CTAs than SMs
    const size_t max_by_rows = a.reduced / (size_t(WR) * U * 2);
    if (splits > max_by_rows) splits = max_by_rows;
    if (splits < 1) splits = 1;
    const size_t slot_bytes = size_t(WC) * 32 * N * sizeof(A);
    if (splits > 256) splits = 256;
    if (splits > 1) {
        if (total * (1 + kWideMaxGroups) > kCounters) splits = 1;
        else {
            // level-1 partials (splits) + level-2 partials (<= 16 groups) per column group
            const size_t fit = kPartialBytes / (slot_bytes * total);
            if (fit < 2 + kWideMaxGroups) return cudaErrorNotSupported;
            if (splits > fit - kWideMaxGroups) splits = fit - kWideMaxGroups;
        }
    }
    const size_t rows_per_split = ceil_div(a.reduced, splits);
    splits = ceil_div(a.reduced, rows_per_split);
    unsigned gsize = 1;
    while (size_t(gsize) * gsize < splits) ++gsize;          // ceil(sqrt(splits)): two balanced levels
    size_t gy = total, gz = 1;
    if (gy > 65535) { gz = ceil_div(total, 65535); gy = 65535; }
    if (gz > 65535) return cudaErrorNotSupported;
    auto kern = cols_wide_kernel<L, A, S, Op, N, U>;
    const size_t smem = WR > 1 ? size_t(kWideThreads) * N * sizeof(A) : 0;
    static PerDeviceOnce attr_set;   // per instantiation and per device (function attributes are per device)
    if (bool* done = attr_set.slot(); !*done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             int(size_t(kWideThreads) * N * sizeof(A)));
        if (e != cudaSuccess) { cudaGetLastError(); return cudaErrorNotSupported; }
        *done = true;
    }
    Finish<A> fin = make_finish<A, S>(a.identity);
    kern<<<dim3((unsigned)splits, (unsigned)gy, (unsigned)gz), kWideThreads, smem, l.stream>>>(
        (const L*)a.in, (S*)a.out, a.outer, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.out_stride_o,
        a.out_stride_i, fin, l.ws.counters, (A*)l.ws.partials, WC, colgroups, rows_per_split, gsize);
    return cudaGetLastError();
}

// Few rows, many columns ([R <= 64, K]): a thread owns a column (adjacent threads adjacent columns, any
// alignment) and four of them kFewStride apart; all rows of up to four columns are in flight at once.
// The general kernels above would launch one tiny CTA per 32 columns here (measured 0.44 TB/s at [7, 38M]).
constexpr int kFewThreads = 256;
constexpr int kFewCols = 4;
constexpr int kFewMaxRows = 64;

template <class L, class A, class S, class Op>
__global__ void __launch_bounds__(kFewThreads)
cols_few_rows_kernel(const L* __restrict__ in, S* __restrict__ out, size_t reduced, size_t inner, size_t in_stride_o,
                     size_t in_stride_r, ptrdiff_t out_stride_o, ptrdiff_t out_stride_i, Finish<A> fin) {
    // A CTA walks chunks of kFewThreads * kFewCols adjacent columns; a thread's columns are kFewThreads apart, so
    // their addresses differ by compile-time offsets from one row pointer (the first version of this kernel spent
    // 123 IMAD + 70 ISETP on 16 loads: ALU pipe 62 % busy; [7, 38M] f32 0.76 -> 0.83 of the HBM peak counting the
    // input only, 0.95 with the 1/7 of output it also writes).  A variant with 16-byte loads per thread was slower
    // (0.77): its four results per thread are stored 16 bytes apart across the warp.
    constexpr size_t kChunk = size_t(kFewThreads) * kFewCols;
    const L* base = in + size_t(blockIdx.y) * in_stride_o;
    S* obase = out + ptrdiff_t(blockIdx.y) * out_stride_o;
    for (size_t chunk = size_t(blockIdx.x) * kChunk; chunk < inner; chunk += size_t(gridDim.x) * kChunk) {
        const size_t c0 = chunk + threadIdx.x;
        A acc[kFewCols];
#pragma unroll
        for (int u = 0; u < kFewCols; ++u) acc[u] = Op::neutral();
        if (chunk + kChunk <= inner) {                       // CTA-uniform: every column of the chunk exists
            const L* pr = base + c0;
            size_t r = 0;
            for (; r + 4 <= reduced; r += 4, pr += 4 * in_stride_r) {
                A v[4][kFewCols];
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int u = 0; u < kFewCols; ++u) v[q][u] = to_acc<A, L>(ldg_scalar(pr + q * in_stride_r + u * kFewThreads));
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int u = 0; u < kFewCols; ++u) acc[u] = Op::apply(acc[u], v[q][u]);
            }
            for (; r < reduced; ++r, pr += in_stride_r) {
                A v[kFewCols];
#pragma unroll
                for (int u = 0; u < kFewCols; ++u) v[u] = to_acc<A, L>(ldg_scalar(pr + u * kFewThreads));
#pragma unroll
                for (int u = 0; u < kFewCols; ++u) acc[u] = Op::apply(acc[u], v[u]);
            }
#pragma unroll
            for (int u = 0; u < kFewCols; ++u)
                write_result<Op, A, S>(obase + ptrdiff_t(c0 + size_t(u) * kFewThreads) * out_stride_i, acc[u], fin);
        } else {                                             // the ragged last chunk
            for (size_t r = 0; r < reduced; ++r)
#pragma unroll
                for (int u = 0; u < kFewCols; ++u) {
                    const size_t c = c0 + size_t(u) * kFewThreads;
                    if (c < inner) acc[u] = Op::apply(acc[u], to_acc<A, L>(ldg_scalar(base + r * in_stride_r + c)));
                }
#pragma unroll
            for (int u = 0; u < kFewCols; ++u) {
                const size_t c = c0 + size_t(u) * kFewThreads;
                if (c < inner) write_result<Op, A, S>(obase + ptrdiff_t(c) * out_stride_i, acc[u], fin);
            }
        }
    }
}

template <class T, class Op>
static cudaError_t launch_cols_few_rows(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    if (a.reduced > size_t(kFewMaxRows) || a.outer > 65535 || a.inner < size_t(sm_count()) * kFewThreads * 2 ||
        tuning().no_few_rows)
        return cudaErrorNotSupported;
    size_t grid = ceil_div(a.inner, size_t(kFewThreads) * kFewCols);
    // (grid, outer) CTAs in whole waves
    const size_t cap = std::max<size_t>(1, size_t(sm_count()) * resident_ctas<cols_few_rows_kernel<L, A, S, Op>>(kFewThreads) / a.outer);
    if (grid > cap) grid = cap;
    cols_few_rows_kernel<L, A, S, Op><<<dim3((unsigned)grid, (unsigned)a.outer), kFewThreads, 0, l.stream>>>(
        (const L*)a.in, (S*)a.out, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.out_stride_o, a.out_stride_i,
        make_finish<A, S>(a.identity));
    return cudaGetLastError();
}

// Narrow dense matrices ([R, K] with K <= 1024 columns, e.g. [N, 3] coordinates): the array is a flat stream
// in which element i belongs to column i % K.  CTAs of B = floor(256 / K) * K threads (1024 for K > 256) walk it with a grid stride
// that is a multiple of K, so a thread's column never changes (column = tid % K): fully coalesced scalar loads,
// one accumulator chain per thread, 8 loads in flight.  In-CTA combine in shared memory (fixed order), one partial
// row per CTA in the workspace, and a warp per column folds the CTA partials (fixed order): deterministic.
// The general kernel spends most lanes on padding here (measured 0.35-0.45 of the peak for K = 3 ... 100).
constexpr int kNarrowMaxK = 1024;
constexpr int kNarrowUnroll = 8;

// N = adjacent columns per thread: 1, or as many 1- / 2-byte elements as fill a 32-bit word (a warp of scalar
// 2-byte loads moves 64 bytes per instruction: [4M, 128] bf16 ran at 0.60 of the peak, uint8 at 0.35).  Then the
// stream is one of K / N "word columns" and everything above holds for those.
template <class L, class A, class Op, int N>
__global__ void __launch_bounds__(1024)
cols_narrow_kernel(const L* __restrict__ in, size_t total, unsigned K, A* __restrict__ partials) {
    __shared__ A sm[N][1024];
    const unsigned B = blockDim.x;
    const unsigned KV = K / N;
    const size_t stride = size_t(gridDim.x) * B;
    const size_t totalv = total / N;                  // K % N == 0, so total % N == 0
    A acc[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = Op::neutral();
    auto load = [&](size_t iv, A (&x)[N]) {
        if constexpr (N == 1) {
            x[0] = to_acc<A, L>(ldg_scalar(in + iv));
        } else {
            Words<4> w;
            w.w[0] = __ldg(reinterpret_cast<const unsigned*>(in) + iv);
#pragma unroll
            for (int j = 0; j < N; ++j) x[j] = vec_elem<L, A, 4>(w, j);
        }
    };
    size_t i = size_t(blockIdx.x) * B + threadIdx.x;
    for (; i + size_t(kNarrowUnroll - 1) * stride < totalv; i += size_t(kNarrowUnroll) * stride) {
        A v[kNarrowUnroll][N];
#pragma unroll
        for (int u = 0; u < kNarrowUnroll; ++u) load(i + size_t(u) * stride, v[u]);
#pragma unroll
        for (int u = 0; u < kNarrowUnroll; ++u)
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], v[u][j]);
    }
    for (; i < totalv; i += stride) {
        A v[N];
        load(i, v);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] = Op::apply(acc[j], v[j]);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) sm[j][threadIdx.x] = acc[j];
    __syncthreads();
    if (threadIdx.x < KV) {
#pragma unroll
        for (int j = 0; j < N; ++j) {
            A r = sm[j][threadIdx.x];
            for (unsigned t = threadIdx.x + KV; t < B; t += KV) r = Op::apply(r, sm[j][t]);
            partials[size_t(blockIdx.x) * K + threadIdx.x * N + j] = r;
        }
    }
}

template <class A, class S, class Op>
__global__ void __launch_bounds__(256)
cols_narrow_finish_kernel(const A* __restrict__ partials, unsigned nctas, unsigned K, S* __restrict__ out,
                          ptrdiff_t out_stride, Finish<A> fin) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned col = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (col >= K) return;
    A r = Op::neutral();
    for (unsigned c = lane; c < nctas; c += 32) r = Op::apply(r, partials[size_t(c) * K + col]);
    r = warp_reduce<Op, A>(r);
    if (lane == 0) write_result<Op, A, S>(out + ptrdiff_t(col) * out_stride, r, fin);
}

template <class T, class Op>
static cudaError_t launch_cols_narrow(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    if (a.outer != 1 || a.inner > size_t(kNarrowMaxK) || a.inner < 2 || a.in_stride_r != a.inner ||
        a.reduced * a.inner < (size_t(1) << 22) || l.ws.partials == nullptr || tuning().no_narrow_cols)
        return cudaErrorNotSupported;
    const unsigned K = unsigned(a.inner);
    // 1- and 2-byte elements: a 32-bit word of adjacent columns per thread when width and base allow
    constexpr int NW = sizeof(L) < 4 ? 4 / int(sizeof(L)) : 1;
    const bool words = NW > 1 && K % NW == 0 && K / NW >= 2 && (reinterpret_cast<size_t>(a.in) & 3) == 0;
    const unsigned KV = words ? K / NW : K;
    const unsigned tmax = KV <= 256 ? 256 : 1024;         // CTA size: the largest multiple of KV that fits
    const unsigned B = (tmax / KV) * KV;
    const size_t total = a.reduced * a.inner;
    size_t grid = size_t(sm_count()) * (2048 / tmax);
    const size_t need = ceil_div(total / (words ? NW : 1), size_t(B) * kNarrowUnroll);
    if (grid > need) grid = need;
    if (grid == 0) grid = 1;
    if (grid * K * sizeof(A) > kPartialBytes) return cudaErrorNotSupported;
    A* partials = reinterpret_cast<A*>(l.ws.partials);
    if (words)
        cols_narrow_kernel<L, A, Op, NW><<<(unsigned)grid, B, 0, l.stream>>>((const L*)a.in, total, K, partials);
    else
        cols_narrow_kernel<L, A, Op, 1><<<(unsigned)grid, B, 0, l.stream>>>((const L*)a.in, total, K, partials);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    cols_narrow_finish_kernel<A, S, Op><<<(K + 7) / 8, 256, 0, l.stream>>>(partials, (unsigned)grid, K, (S*)a.out,
                                                                         a.out_stride_i, make_finish<A, S>(a.identity));
    return cudaGetLastError();
}

// Many small dense blocks ([O, R, K] with K a power of two <= 32 and R * K elements of a few KiB: the first pass
// of an inner-first plan over [..., R3, K4] with a tiny innermost kept axis, per-pixel channel statistics, ...).
// A warp streams one block as a flat run: lane l reads elements l, l + 32, ... whose column is l % K for good, so a
// lane needs one accumulator; lanes with the same column meet through xor-shuffles.  Fully coalesced 128-byte
// requests, 8 loads in flight per lane, warps walk consecutive blocks.  The general kernel would launch one
// 256-thread CTA per block with most warps idle (measured 0.07 of the HBM peak at [64^4, 4] axes (1, 3)).
constexpr int kSmallThreads = 256;
constexpr int kSmallUnroll = 8;
constexpr size_t kSmallMaxBlock = 16384;   // elements per block

template <class L, class A, class S, class Op>
__global__ void __launch_bounds__(kSmallThreads)
cols_small_kernel(const L* __restrict__ in, S* __restrict__ out, size_t outer, unsigned n, unsigned K, size_t so,
                  ptrdiff_t oo, ptrdiff_t oi, Finish<A> fin) {
    const unsigned lane = threadIdx.x & 31;
    const size_t warp = (size_t(blockIdx.x) * kSmallThreads + threadIdx.x) >> 5;
    const size_t nwarps = (size_t(gridDim.x) * kSmallThreads) >> 5;
    for (size_t o = warp; o < outer; o += nwarps) {
        const L* p = in + o * so;
        A acc = Op::neutral();
        unsigned e = lane;
        for (; e + (kSmallUnroll - 1) * 32 < n; e += kSmallUnroll * 32) {
            A v[kSmallUnroll];
#pragma unroll
            for (int u = 0; u < kSmallUnroll; ++u) v[u] = to_acc<A, L>(ldg_scalar(p + e + u * 32));
#pragma unroll
            for (int u = 0; u < kSmallUnroll; ++u) acc = Op::apply(acc, v[u]);
        }
        for (; e < n; e += 32) acc = Op::apply(acc, to_acc<A, L>(ldg_scalar(p + e)));
        for (unsigned m = 16; m >= K; m >>= 1) acc = Op::apply(acc, shfl_xor(acc, int(m)));
        if (lane < K) write_result<Op, A, S>(out + ptrdiff_t(o) * oo + ptrdiff_t(lane) * oi, acc, fin);
    }
}

template <class T, class Op>
static cudaError_t launch_cols_small(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    const size_t K = a.inner;
    if (K < 2 || K > 32 || (K & (K - 1)) != 0 || a.in_stride_r != K || a.reduced * K > kSmallMaxBlock ||
        a.outer < size_t(sm_count()) * 64 || tuning().no_tiled)
        return cudaErrorNotSupported;
    const size_t warps_per_cta = kSmallThreads / 32;
    size_t grid = (a.outer + warps_per_cta - 1) / warps_per_cta;
    const size_t cap = size_t(sm_count()) * resident_ctas<cols_small_kernel<L, A, S, Op>>(kSmallThreads);
    if (grid > cap) grid = cap;
    cols_small_kernel<L, A, S, Op><<<(unsigned)grid, kSmallThreads, 0, l.stream>>>(
        (const L*)a.in, (S*)a.out, a.outer, unsigned(a.reduced * K), unsigned(K), a.in_stride_o, a.out_stride_o,
        a.out_stride_i, make_finish<A, S>(a.identity));
    return cudaGetLastError();
}

// A short reduced axis between MANY outer slices ([N, 3, M] channel folds, [1M, 4, 64]): a thread owns one output
// (o, c) - adjacent threads adjacent columns - and folds its R <= 64 values itself, two outputs and four rows at a
// time in flight.  The general kernel launches a CTA per (slice, 256 columns) for R rows each ([1M, 4, 64] f32:
// 0.48 TB/s, [4096, 4, 16384]: 1.9), the few-rows kernel above wants one huge inner axis.
template <class L, class A, class S, class Op>
__global__ void __launch_bounds__(kFewThreads)
cols_short_axis_kernel(const L* __restrict__ in, S* __restrict__ out, size_t outer, size_t reduced, size_t inner,
                       size_t in_stride_o, size_t in_stride_r, ptrdiff_t out_stride_o, ptrdiff_t out_stride_i, Finish<A> fin) {
    const size_t total = outer * inner;
    const size_t nthr = size_t(gridDim.x) * kFewThreads;
    for (size_t t0 = size_t(blockIdx.x) * kFewThreads + threadIdx.x; t0 < total; t0 += 2 * nthr) {
        const size_t t1 = t0 + nthr;
        const bool two = t1 < total;
        const size_t o0 = t0 / inner, c0 = t0 - o0 * inner;
        const size_t o1 = two ? t1 / inner : o0, c1 = two ? t1 - o1 * inner : c0;
        const L* p0 = in + o0 * in_stride_o + c0;
        const L* p1 = in + o1 * in_stride_o + c1;
        A a0 = Op::neutral(), a1 = Op::neutral();
        size_t r = 0;
        for (; r + 4 <= reduced; r += 4, p0 += 4 * in_stride_r, p1 += 4 * in_stride_r) {
            A v0[4], v1[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { v0[q] = to_acc<A, L>(ldg_scalar(p0 + q * in_stride_r)); v1[q] = to_acc<A, L>(ldg_scalar(p1 + q * in_stride_r)); }
#pragma unroll
            for (int q = 0; q < 4; ++q) { a0 = Op::apply(a0, v0[q]); a1 = Op::apply(a1, v1[q]); }
        }
        for (; r < reduced; ++r, p0 += in_stride_r, p1 += in_stride_r) {
            a0 = Op::apply(a0, to_acc<A, L>(ldg_scalar(p0)));
            a1 = Op::apply(a1, to_acc<A, L>(ldg_scalar(p1)));
        }
        write_result<Op, A, S>(out + ptrdiff_t(o0) * out_stride_o + ptrdiff_t(c0) * out_stride_i, a0, fin);
        if (two) write_result<Op, A, S>(out + ptrdiff_t(o1) * out_stride_o + ptrdiff_t(c1) * out_stride_i, a1, fin);
    }
}

// The same with NV adjacent outputs per thread (one 16-byte load per row): R = 3 or 4 rows of scalar loads left a
// thread with 32 bytes in flight ([4096, 4, 16384] f32: 3.6 TB/s).
template <class L, class A, class S, class Op, int NV>
__global__ void __launch_bounds__(kFewThreads)
cols_short_axis_vec_kernel(const L* __restrict__ in, S* __restrict__ out, size_t outer, size_t reduced, size_t inner,
                           size_t in_stride_o, size_t in_stride_r, ptrdiff_t out_stride_o, ptrdiff_t out_stride_i, Finish<A> fin) {
    const size_t innerv = inner / NV, total = outer * innerv;
    const size_t nthr = size_t(gridDim.x) * kFewThreads;
    for (size_t t0 = size_t(blockIdx.x) * kFewThreads + threadIdx.x; t0 < total; t0 += 2 * nthr) {
        const size_t t1 = t0 + nthr;
        const bool two = t1 < total;
        const size_t o0 = t0 / innerv, c0 = (t0 - o0 * innerv) * NV;
        const size_t o1 = two ? t1 / innerv : o0, c1 = two ? (t1 - o1 * innerv) * NV : c0;
        const L* p0 = in + o0 * in_stride_o + c0;
        const L* p1 = in + o1 * in_stride_o + c1;
        A a0[NV], a1[NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) { a0[j] = Op::neutral(); a1[j] = Op::neutral(); }
        size_t r = 0;
        for (; r + 4 <= reduced; r += 4, p0 += 4 * in_stride_r, p1 += 4 * in_stride_r) {
            A v0[4][NV], v1[4][NV];
#pragma unroll
            for (int q = 0; q < 4; ++q) { load_cols<L, A, NV>(p0 + q * in_stride_r, v0[q]); load_cols<L, A, NV>(p1 + q * in_stride_r, v1[q]); }
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int j = 0; j < NV; ++j) { a0[j] = Op::apply(a0[j], v0[q][j]); a1[j] = Op::apply(a1[j], v1[q][j]); }
        }
        for (; r < reduced; ++r, p0 += in_stride_r, p1 += in_stride_r) {
            A v0[NV], v1[NV];
            load_cols<L, A, NV>(p0, v0);
            load_cols<L, A, NV>(p1, v1);
#pragma unroll
            for (int j = 0; j < NV; ++j) { a0[j] = Op::apply(a0[j], v0[j]); a1[j] = Op::apply(a1[j], v1[j]); }
        }
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            write_result<Op, A, S>(out + ptrdiff_t(o0) * out_stride_o + ptrdiff_t(c0 + j) * out_stride_i, a0[j], fin);
            if (two) write_result<Op, A, S>(out + ptrdiff_t(o1) * out_stride_o + ptrdiff_t(c1 + j) * out_stride_i, a1[j], fin);
        }
    }
}

template <class T, class Op>
static cudaError_t launch_cols_short_axis(const ColsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    // many slices of few rows; an inner run of at least 32 bytes so that the scalar loads of a warp stay in sectors
    if (a.reduced > 16 || a.outer < 64 || a.inner * sizeof(L) < 32 || a.outer * a.inner < size_t(sm_count()) * kFewThreads * 4 ||
        tuning().no_few_rows)
        return cudaErrorNotSupported;
    if constexpr (sizeof(L) == 4 || sizeof(L) == 8) {
        constexpr int NV = 16 / int(sizeof(L));
        if (a.inner % NV == 0 && a.in_stride_o % NV == 0 && a.in_stride_r % NV == 0 && (reinterpret_cast<size_t>(a.in) & 15) == 0 &&
            a.outer * (a.inner / NV) >= size_t(sm_count()) * kFewThreads * 2) {
            size_t grid = ceil_div(a.outer * (a.inner / NV), size_t(kFewThreads) * 2);
            const size_t cap = size_t(sm_count()) * resident_ctas<cols_short_axis_vec_kernel<L, A, S, Op, NV>>(kFewThreads);
            if (grid > cap) grid = cap;
            cols_short_axis_vec_kernel<L, A, S, Op, NV><<<(unsigned)grid, kFewThreads, 0, l.stream>>>(
                (const L*)a.in, (S*)a.out, a.outer, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.out_stride_o,
                a.out_stride_i, make_finish<A, S>(a.identity));
            return cudaGetLastError();
        }
    }
    size_t grid = ceil_div(a.outer * a.inner, size_t(kFewThreads) * 2);
    const size_t cap = size_t(sm_count()) * resident_ctas<cols_short_axis_kernel<L, A, S, Op>>(kFewThreads);
    if (grid > cap) grid = cap;
    cols_short_axis_kernel<L, A, S, Op><<<(unsigned)grid, kFewThreads, 0, l.stream>>>(
        (const L*)a.in, (S*)a.out, a.outer, a.reduced, a.inner, a.in_stride_o, a.in_stride_r, a.out_stride_o, a.out_stride_i,
        make_finish<A, S>(a.identity));
    return cudaGetLastError();
}

template <class T, class Op>
static cudaError_t launch_cols_t(const ColsArgs& a, const Launch& l) {
    using L = typename T::L;
    {
        cudaError_t e = launch_cols_small<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    {
        cudaError_t e = launch_cols_narrow<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    {
        cudaError_t e = launch_cols_few_rows<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    {
        cudaError_t e = launch_cols_short_axis<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    constexpr int NV = 32 / int(sizeof(L));
    const size_t vb = 32;
    const bool vec_ok = (reinterpret_cast<size_t>(a.in) % vb == 0) && (a.inner % NV == 0) &&
                        ((a.in_stride_o * sizeof(L)) % vb == 0) && ((a.in_stride_r * sizeof(L)) % vb == 0);
    if (vec_ok) {
        cudaError_t e = launch_cols_wide<T, Op, NV>(a, l);
        if (e != cudaErrorNotSupported) return e;
        return launch_cols_n<T, Op, NV>(a, l);
    }
    return launch_cols_n<T, Op, 1>(a, l);
}

cudaError_t launch_cols(int red, int dtin, int dtout, const ColsArgs& a, const Launch& l) {
    return dispatch_types(dtin, dtout, [&](auto t) -> cudaError_t {
        using T = decltype(t);
        using A = typename T::A;
        return dispatch_op<A>(red, [&](auto op) -> cudaError_t {
            using Op = decltype(op);
            return launch_cols_t<T, Op>(a, l);
        });
    });
}

}  // namespace db200
