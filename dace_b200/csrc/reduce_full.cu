// Full reduction (and "few, very long rows") kernel for sm_100a.
//
// Replaces the reference's CUB call site cub::DeviceReduce::Reduce
// (reference dace/libraries/standard/nodes/reduce.py:520-525,574-588) and the
// GPUAuto 1-D schedule where 2^20 32-thread blocks each atomically add into
// _out[0] (reference dace/libraries/standard/nodes/reduce.py:993-1004,1068-1076).
//
// One launch, no second pass:
//   * persistent grid, one or two CTAs per SM, each CTA walks 64/256 KiB tiles with
//     U = 4/8 independent 256-bit streaming loads per thread (LDG.E.256, L1 no-allocate,
//     L2 evict-first);
//   * thread -> warp (shuffle / redux.sync) -> CTA (shared memory);
//   * CTAs of a thread-block cluster combine through distributed shared memory;
//   * cluster leaders publish one partial each and take a ticket; the last
//     cluster folds the partials in a fixed order, so floating-point results are
//     run-to-run deterministic (no float atomics).
// gridDim.y indexes independent rows (row_stride apart), so the same kernel
// serves segmented reductions with few long segments.
#include <cstdio>
#include <mutex>
#include "common.cuh"
#include "span.cuh"
#include "kernels.h"

namespace db200 {

template <class L, class A, class S, class Op, int VB, int U, int THREADS>
__global__ void __launch_bounds__(THREADS)
split_reduce_kernel(const L* __restrict__ in, S* __restrict__ out, size_t n, size_t row_stride,
                    ptrdiff_t out_stride, Finish<A> fin, unsigned* __restrict__ counters,
                    A* __restrict__ partials) {
    constexpr int N = VecInfo<L, VB>::N;
    constexpr int NA = VecInfo<L, VB>::NA;
    __shared__ A scratch[32];
    __shared__ A cluster_part[8];
    __shared__ int is_last;

    const unsigned tid = threadIdx.x;
    const size_t row = blockIdx.y;
    const L* p = in + row * row_stride;

    // DSMEM may only be touched once every CTA of the cluster is running: arrive
    // now, wait right before the remote store (the streaming loop hides it).
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks();
    const unsigned crank = cluster.block_rank();
    if (csize > 1) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");

    A acc[NA];
#pragma unroll
    for (int j = 0; j < NA; ++j) acc[j] = Op::neutral();

    const size_t addr = reinterpret_cast<size_t>(p);
    size_t h = ((size_t(VB) - (addr & size_t(VB - 1))) & size_t(VB - 1)) / sizeof(L);
    if (h > n) h = n;
    const char* pv = reinterpret_cast<const char*>(p + h);
    const size_t nvec = (n - h) / N;
    constexpr size_t kTile = size_t(U) * THREADS;
    const size_t ntiles = nvec / kTile;

    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const char* base = pv + (t * kTile + tid) * VB;
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(base + size_t(u) * THREADS * VB);
#pragma unroll
        for (int u = 0; u < U; ++u) fold_vec<L, A, Op, VB>(acc, w[u]);
    }
    // Vectors that do not fill a whole tile, spread over the row's CTAs.
    for (size_t v = ntiles * kTile + size_t(blockIdx.x) * THREADS + tid; v < nvec; v += size_t(gridDim.x) * THREADS) {
        Words<VB> w = ldg_stream<VB>(pv + v * VB);
        fold_vec<L, A, Op, VB>(acc, w);
    }
    // Unaligned head and sub-vector tail: scalar, first CTA of the row only.
    if (blockIdx.x == 0) {
        for (size_t i = tid; i < h; i += THREADS) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
        for (size_t i = h + nvec * N + tid; i < n; i += THREADS) acc[0] = Op::apply(acc[0], to_acc<A, L>(ldg_scalar(p + i)));
    }

    A v = fold_acc<A, Op, NA>(acc);
    v = block_reduce<Op, A>(v, scratch);

    // Cluster combine through DSMEM: every CTA drops its partial into rank 0's slot array.
    if (csize > 1) {
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (tid == 0) {
            A* dst = cluster.map_shared_rank(cluster_part, 0);
            dst[crank] = v;
        }
        cluster.sync();
        if (crank != 0) return;
        if (tid == 0) {
            v = cluster_part[0];
            for (unsigned i = 1; i < csize; ++i) v = Op::apply(v, cluster_part[i]);
        }
    }

    const unsigned nclusters = gridDim.x / csize;
    S* o = out + ptrdiff_t(row) * out_stride;
    if (nclusters == 1) {
        if (tid == 0) write_result<Op, A, S>(o, v, fin);
        return;
    }
    // Publish + ticket; the last cluster leader folds all partials in index order.
    A* row_partials = partials + row * nclusters;
    if (tid == 0) {
        row_partials[blockIdx.x / csize] = v;
        __threadfence();
        unsigned ticket = atomicAdd(&counters[row], 1u);
        is_last = (ticket == nclusters - 1);
    }
    __syncthreads();
    if (!is_last) return;
    if (tid < 32) {
        __threadfence();
        A x = Op::neutral();
        for (unsigned i = tid; i < nclusters; i += 32) x = Op::apply(x, __ldcg(row_partials + i));
        x = warp_reduce<Op, A>(x);
        if (tid == 0) {
            write_result<Op, A, S>(o, x, fin);
            counters[row] = 0;  // self-reset for the next launch on this workspace
        }
    }
}

template <class T, class Op, int VB, int U, int THREADS>
static cudaError_t launch_split_cfg(const SplitArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    auto kern = split_reduce_kernel<L, A, S, Op, VB, U, THREADS>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.ctas_per_row, (unsigned)a.rows, 1);
    cfg.blockDim = dim3(THREADS, 1, 1);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = l.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = a.cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    Finish<A> fin = make_finish<A, S>(a.identity);
    return cudaLaunchKernelEx(&cfg, kern, (const L*)a.in, (S*)a.out, a.n, a.row_stride, a.out_stride, fin,
                              l.ws.counters, (A*)l.ws.partials);
}

cudaError_t launch_split(int red, int dtin, int dtout, const SplitArgs& a, const Launch& l) {
    return dispatch_types(dtin, dtout, [&](auto t) -> cudaError_t {
        using T = decltype(t);
        using A = typename T::A;
        return dispatch_op<A>(red, [&](auto op) -> cudaError_t {
            using Op = decltype(op);
            // Geometry from the sweep in profiles/r1_ubench_split_sweep_and_atomics.txt: full occupancy
            // (1024-thread CTAs) and 8 x 256-bit loads in flight per thread is worth +4 % on 4 GiB.
            if (a.big) return launch_split_cfg<T, Op, 32, 8, 1024>(a, l);
            return launch_split_cfg<T, Op, 32, 4, 512>(a, l);
        });
    });
}


// ---------------------------------------------------------------------------
// Allocation-free variant (dace_b200_reduce_rows_noalloc in include/dace_b200_reduce.h): the
// split kernel above with its tickets and partials in a file-scope device array instead of a
// caller / library workspace, for callers that may not call cudaMalloc (stream capture, serving
// loops).  The array lives in every GPU that loads this library until the process exits.  In host
// code `g_noalloc_scratch` is only a stand-in for the symbol: the device address is looked up
// once per device with cudaGetSymbolAddress (that device current) and handed to the kernel.
// ---------------------------------------------------------------------------
constexpr size_t kNoAllocMaxRows = 256;                                   // ticket counters
constexpr size_t kNoAllocMaxPartials = 8192;                              // rows x clusters, 8 bytes each
constexpr size_t kNoAllocCounterBytes = kNoAllocMaxRows * sizeof(unsigned);               // 1 KiB
constexpr size_t kNoAllocScratchBytes = kNoAllocCounterBytes + kNoAllocMaxPartials * 8;   // 65 KiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

static std::mutex g_noalloc_mutex;
static void* g_noalloc_ptr[64] = {nullptr};

static cudaError_t noalloc_scratch(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // device `dev` is current
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return cudaSuccess;
}

cudaError_t reduce_rows_noalloc(int red, int dtin, int dtout, const void* in, void* out, size_t rows, size_t n,
                                size_t row_stride, const void* identity, cudaStream_t stream) {
    if (red < RT_Min || red > RT_BXor) return cudaErrorNotSupported;
    if (dispatch_types(dtin, dtout, [](auto) -> cudaError_t { return cudaSuccess; }) != cudaSuccess)
        return cudaErrorNotSupported;
    if (rows == 0) return cudaSuccess;
    if (rows > kNoAllocMaxRows) {
        fprintf(stderr,
                "dace_b200_reduce_rows_noalloc: %zu segments exceed the maximum of %zu that the static device "
                "scratch (%zu bytes) serves; use dace_b200_reduce_rows with a workspace\n",
                rows, kNoAllocMaxRows, kNoAllocScratchBytes);
        return cudaErrorInvalidValue;
    }
    if (!out || (n != 0 && !in)) return cudaErrorInvalidValue;
    void* scratch = nullptr;
    cudaError_t e = noalloc_scratch(&scratch);
    if (e != cudaSuccess) return e;
    Launch l;
    l.stream = stream;
    l.ws.counters = reinterpret_cast<unsigned*>(scratch);
    l.ws.partials = static_cast<unsigned char*>(scratch) + kNoAllocCounterBytes;
    if (n == 0) {   // empty range: outputs become the identity, or stay as they are without one
        if (!identity) return cudaSuccess;
        const long long ks[1] = {(long long)rows}, ko[1] = {1};
        return launch_fill(dtout, out, 1, ks, ko, identity, l);
    }
    // The array is shared by every call and keeps the last call's contents: the tickets must start at zero.
    e = cudaMemsetAsync(scratch, 0, kNoAllocCounterBytes, stream);
    if (e != cudaSuccess) return e;
    const size_t row_bytes = n * size_t(dtype_size(dtin));
    const size_t sms = size_t(sm_count());
    const int big = row_bytes * rows >= (size_t(256) << 20) && row_bytes >= (size_t(2) << 20);
    const size_t tile = big ? (size_t(256) << 10) : (size_t(64) << 10);
    size_t want = (row_bytes + tile - 1) / tile;
    const size_t cap = ((big ? 1 : 2) * sms + rows - 1) / rows;     // CTAs resident at a time, per row
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    unsigned cluster = 1;
    if (want >= 2) { cluster = 2; want = (want / 2) * 2; }
    while (want > cluster && rows * (want / cluster) > kNoAllocMaxPartials) want -= cluster;
    SplitArgs a{in, out, n, rows, row_stride, 1, (unsigned)want, cluster, identity, big};
    return launch_split(red, dtin, dtout, a, l);
}

}  // namespace db200
