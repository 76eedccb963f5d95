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

template <class T, class Op>
static cudaError_t launch_cols_t(const ColsArgs& a, const Launch& l) {
    using L = typename T::L;
    constexpr int NV = 32 / int(sizeof(L));
    const size_t vb = 32;
    const bool vec_ok = (reinterpret_cast<size_t>(a.in) % vb == 0) && (a.inner % NV == 0) &&
                        ((a.in_stride_o * sizeof(L)) % vb == 0) && ((a.in_stride_r * sizeof(L)) % vb == 0);
    if (vec_ok) return launch_cols_n<T, Op, NV>(a, l);
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
