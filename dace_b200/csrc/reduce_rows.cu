// Segmented (row) reductions for sm_100a: the reduced axes are the trailing,
// contiguous ones, so every output owns one contiguous span of the input.
//
// Replaces cub::DeviceSegmentedReduce::Reduce with dace::stridedIterator offsets
// (reference dace/libraries/standard/nodes/reduce.py:526-544,
// reference dace/runtime/include/dace/reduction.h:594-622) and the GPUAuto
// contiguous schedule, one 32-thread block per output
// (reference dace/libraries/standard/nodes/reduce.py:984-1139).
//
// Work decomposition by row length:
//   * G in {1,2,4,8,16,32} lanes per row (sub-warp groups, persistent grid of
//     full CTAs: no 1-warp blocks) - each lane issues U independent 256-bit loads;
//   * one CTA per row for rows of tens of KiB and up;
//   * rows of MiB size and few of them go to the split kernel (reduce_full.cu).
#include "common.cuh"
#include "span.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kRowThreads = 256;

// G lanes per row; 32/G rows per warp per step.  G is a run-time power of two.
template <class L, class A, class S, class Op, int VB, int U>
__global__ void __launch_bounds__(kRowThreads)
rows_group_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, size_t n, size_t row_stride,
                  ptrdiff_t out_stride, Finish<A> fin, int G) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned g = lane & (G - 1);            // lane within the row group
    const unsigned sub = lane / G;                // row slot within the warp
    const unsigned rows_per_warp = 32 / G;
    // Warps are numbered CTA-interleaved (warp w of CTA c is global warp w * gridDim.x + c): when the rows do not
    // fill the last CTAs, every CTA - and with a grid that is a multiple of the SM count every SM - is short of
    // the same number of rows, instead of some SMs holding one whole CTA more than others (C1: 4096 rows of
    // 16 KiB were 512 full CTAs on 148 SMs, 4 on some and 3 on others: 16 % imbalance on a one-wave kernel).
    const size_t warp = size_t(threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    const size_t nwarps = (size_t(gridDim.x) * kRowThreads) >> 5;

    for (size_t rb = warp * rows_per_warp; rb < rows; rb += nwarps * rows_per_warp) {
        const size_t r = rb + sub;
        const bool active = r < rows;
        A v = Op::neutral();
        if (active) v = span_accumulate<L, A, Op, VB, U>(in + r * row_stride, n, g, G);
        for (int m = G >> 1; m >= 1; m >>= 1) v = Op::apply(v, shfl_xor(v, m));
        if (active && g == 0) write_result<Op, A, S>(out + ptrdiff_t(r) * out_stride, v, fin);
    }
}

// One CTA per row (grid-stride over rows).
template <class L, class A, class S, class Op, int VB, int U>
__global__ void __launch_bounds__(kRowThreads)
rows_block_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, size_t n, size_t row_stride,
                  ptrdiff_t out_stride, Finish<A> fin) {
    __shared__ A scratch[32];
    for (size_t r = blockIdx.x; r < rows; r += gridDim.x) {
        A v = span_accumulate<L, A, Op, VB, U>(in + r * row_stride, n, threadIdx.x, kRowThreads);
        v = block_reduce<Op, A>(v, scratch);
        if (threadIdx.x == 0) write_result<Op, A, S>(out + ptrdiff_t(r) * out_stride, v, fin);
        __syncthreads();  // scratch is reused by the next row
    }
}

static inline int pick_group(size_t row_bytes) {
    // Aim for 128 bytes (U = 4 256-bit vectors, all in flight at once) per lane before widening the group: with 64
    // a lane of a 1 KiB row had two loads in flight and then a shuffle tree (ncu, [1M, 256] f32: 0.90 of the peak,
    // warps waiting on L1TEX 60 % of the time at 50 % occupancy).
    int G = 1;
    const size_t per_lane = tuning().rows_lane_bytes > 0 ? size_t(tuning().rows_lane_bytes) : 128;
    while (G < 32 && size_t(G) * per_lane < row_bytes) G <<= 1;
    return G;
}

// Tiny dense rows ([R, n <= 128], rows back to back in memory, e.g. [N, 33] feature rows): a row is shorter than
// one coalesced warp request, so the CTA treats its rows as one flat tile - 256 threads load kTinyTile elements with
// scalar coalesced loads, every one of them issued before the first use (16 loads in flight per thread), into
// shared memory; then warps fold rows out of shared memory, WIDTH lanes per row (conflict-free: a row is read by
// consecutive lanes).  The lanes-per-row kernel above reaches 0.4-0.5 of the peak on such rows.
constexpr int kTinyTile = 4096;            // elements per CTA tile
constexpr int kTinyMaxN = 128;

template <class L, class A, class S, class Op, int WIDTH>
__global__ void __launch_bounds__(kRowThreads)
rows_tiny_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, unsigned n, unsigned rows_per_tile,
                 ptrdiff_t out_stride, Finish<A> fin) {
    __shared__ L tile[kTinyTile];
    constexpr int kPer = kTinyTile / kRowThreads;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int W = WIDTH < 2 ? 2 : WIDTH;
    const unsigned seg = lane / W, pos = lane % W;
    constexpr unsigned kSegs = 32 / W;
    (void)seg; (void)pos; (void)warp;
    const size_t ntiles = (rows + rows_per_tile - 1) / rows_per_tile;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const size_t row0 = t * rows_per_tile;
        const unsigned nrows = (rows - row0 < rows_per_tile) ? unsigned(rows - row0) : rows_per_tile;
        const unsigned count = nrows * n;
        const L* src = in + row0 * n;
        L v[kPer];
#pragma unroll
        for (int u = 0; u < kPer; ++u) {
            const unsigned i = threadIdx.x + unsigned(u) * kRowThreads;
            if (i < count) v[u] = ldg_scalar(src + i);
        }
#pragma unroll
        for (int u = 0; u < kPer; ++u) {
            const unsigned i = threadIdx.x + unsigned(u) * kRowThreads;
            if (i < count) tile[i] = v[u];
        }
        __syncthreads();
        if constexpr (WIDTH == 1) {
            // a thread per row: n shared-memory reads at stride n elements - conflict-free for odd n.  For even n
            // every thread starts at column (row mod n) and wraps around: neighbouring threads are then n + 1
            // elements apart, an odd stride again (a fold does not care where it starts).
            const bool rotate = (n & 1u) == 0 && sizeof(L) >= 4;
            for (unsigned r = threadIdx.x; r < nrows; r += kRowThreads) {
                A a0 = Op::neutral(), a1 = Op::neutral();
                const L* p = tile + r * n;
                unsigned c = rotate ? r % n : 0;
                unsigned j = 0;
                for (; j + 1 < n; j += 2) {
                    const unsigned c1 = (c + 1 == n) ? 0 : c + 1;
                    a0 = Op::apply(a0, to_acc<A, L>(p[c]));
                    a1 = Op::apply(a1, to_acc<A, L>(p[c1]));
                    c = (c1 + 1 == n) ? 0 : c1 + 1;
                }
                if (j < n) a0 = Op::apply(a0, to_acc<A, L>(p[c]));
                write_result<Op, A, S>(out + ptrdiff_t(row0 + r) * out_stride, Op::apply(a0, a1), fin);
            }
        } else {
            // warp-uniform trip count: the shuffles inside warp_reduce need all 32 lanes, also those whose row is
            // past the end of the tile
            for (unsigned rb = warp * kSegs; rb < nrows; rb += (kRowThreads / 32) * kSegs) {
                const unsigned r = rb + seg;
                const bool valid = r < nrows;
                A acc = Op::neutral();
                if (valid)
                    for (unsigned c = pos; c < n; c += WIDTH) acc = Op::apply(acc, to_acc<A, L>(tile[r * n + c]));
                acc = warp_reduce<Op, A, WIDTH>(acc);
                if (valid && pos == 0) write_result<Op, A, S>(out + ptrdiff_t(row0 + r) * out_stride, acc, fin);
            }
        }
        __syncthreads();
    }
}

template <class T, class Op>
static cudaError_t launch_rows_tiny(const RowsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    if (a.n > size_t(kTinyMaxN) || a.n < 2 || a.row_stride != a.n || a.rows * a.n < (size_t(1) << 22) ||
        tuning().no_tiny_rows)
        return cudaErrorNotSupported;
    const unsigned n = unsigned(a.n);
    const unsigned rows_per_tile = kTinyTile / n;
    const size_t ntiles = (a.rows + rows_per_tile - 1) / rows_per_tile;
    Finish<A> fin = make_finish<A, S>(a.identity);
#define DB200_TINY(W)                                                                                                   \
    do {                                                                                                                \
        size_t grid = size_t(sm_count()) * resident_ctas<rows_tiny_kernel<L, A, S, Op, W>>(kRowThreads);                 \
        if (grid > ntiles) grid = ntiles;                                                                               \
        rows_tiny_kernel<L, A, S, Op, W><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(                                  \
            (const L*)a.in, (S*)a.out, a.rows, n, rows_per_tile, a.out_stride, fin);                                    \
    } while (0)
    // lanes per row: a thread per row while a tile holds at least 64 rows (n <= 64: measured faster than two lanes
    // per row, [8M, 33] 0.63 vs 0.57 of the peak); for longer rows as many lanes as keep the 256 threads busy in the
    // fold ([2.7M, 100]: 40 rows per tile, 4 lanes each, 0.54 -> 0.61)
    unsigned width = 1;
    if (rows_per_tile <= 64)
        while (width < 32 && rows_per_tile * width * 2 <= unsigned(kRowThreads)) width <<= 1;
    switch (width) {
        case 1: DB200_TINY(1); break;
        case 2: DB200_TINY(2); break;
        case 4: DB200_TINY(4); break;
        case 8: DB200_TINY(8); break;
        case 16: DB200_TINY(16); break;
        default: DB200_TINY(32); break;
    }
#undef DB200_TINY
    return cudaGetLastError();
}

template <class T, class Op>
static cudaError_t launch_rows_t(const RowsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    constexpr int VB = 32;
    {
        cudaError_t e = launch_rows_tiny<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    Finish<A> fin = make_finish<A, S>(a.identity);
    const size_t row_bytes = a.n * sizeof(L);
    const int sms = sm_count();
    if (row_bytes >= (size_t(32) << 10)) {
        const size_t slots = size_t(sms) * resident_ctas<rows_block_kernel<L, A, S, Op, VB, 4>>(kRowThreads);
        size_t grid = a.rows < slots ? a.rows : slots;
        rows_block_kernel<L, A, S, Op, VB, 4><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, a.rows, a.n, a.row_stride, a.out_stride, fin);
        return cudaGetLastError();
    }
    const int G = pick_group(row_bytes);
    const size_t rows_per_cta = size_t(kRowThreads / 32) * (32 / G);
    size_t grid = (a.rows + rows_per_cta - 1) / rows_per_cta;
    const size_t cap = size_t(sms) * resident_ctas<rows_group_kernel<L, A, S, Op, VB, 4>>(kRowThreads);   // one full wave
    if (grid > cap) grid = cap;
    else if (grid > size_t(sms)) grid = ((grid + sms - 1) / sms) * sms;   // whole CTAs per SM (see the kernel)
    if (grid == 0) grid = 1;
    rows_group_kernel<L, A, S, Op, VB, 4><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(
        (const L*)a.in, (S*)a.out, a.rows, a.n, a.row_stride, a.out_stride, fin, G);
    return cudaGetLastError();
}

cudaError_t launch_rows(int red, int dtin, int dtout, const RowsArgs& a, const Launch& l) {
    return dispatch_types(dtin, dtout, [&](auto t) -> cudaError_t {
        using T = decltype(t);
        using A = typename T::A;
        return dispatch_op<A>(red, [&](auto op) -> cudaError_t {
            using Op = decltype(op);
            return launch_rows_t<T, Op>(a, l);
        });
    });
}

// ---------------------------------------------------------------------------
// Fused row max + row sum: one read of the input for both outputs (softmax-
// style pair, reference tests/softmax_test.py:8-31 runs them as two Reduce
// nodes, i.e. two full passes).  fp32 accumulation for every input type.
// ---------------------------------------------------------------------------
template <class L, int VB, int U>
__device__ __forceinline__ void span_max_sum(const L* __restrict__ p, size_t n, unsigned tid, unsigned nthr,
                                             float& mx, float& sm) {
    constexpr int N = VecInfo<L, VB>::N;
    constexpr int NA = 4;
    float amax[NA], asum[NA];
#pragma unroll
    for (int j = 0; j < NA; ++j) { amax[j] = OpMax<float>::neutral(); asum[j] = 0.f; }
    const size_t addr = reinterpret_cast<size_t>(p);
    size_t h = ((size_t(VB) - (addr & size_t(VB - 1))) & size_t(VB - 1)) / sizeof(L);
    if (h > n) h = n;
    for (size_t i = tid; i < h; i += nthr) {
        float x = to_acc<float, L>(ldg_scalar(p + i));
        amax[0] = fmaxf(amax[0], x); asum[0] += x;
    }
    const char* pv = reinterpret_cast<const char*>(p + h);
    const size_t nvec = (n - h) / N;
    size_t v = tid;
    for (; v + size_t(U - 1) * nthr < nvec; v += size_t(U) * nthr) {
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < N; ++j) {
                float x = vec_elem<L, float, VB>(w[u], j);
                amax[j % NA] = fmaxf(amax[j % NA], x);
                asum[j % NA] += x;
            }
    }
    for (; v < nvec; v += nthr) {
        Words<VB> w = ldg_stream<VB>(pv + v * VB);
#pragma unroll
        for (int j = 0; j < N; ++j) {
            float x = vec_elem<L, float, VB>(w, j);
            amax[j % NA] = fmaxf(amax[j % NA], x);
            asum[j % NA] += x;
        }
    }
    for (size_t i = h + nvec * N + tid; i < n; i += nthr) {
        float x = to_acc<float, L>(ldg_scalar(p + i));
        amax[0] = fmaxf(amax[0], x); asum[0] += x;
    }
    mx = fmaxf(fmaxf(amax[0], amax[1]), fmaxf(amax[2], amax[3]));
    sm = (asum[0] + asum[1]) + (asum[2] + asum[3]);
}

template <class L, int VB, int U>
__global__ void __launch_bounds__(kRowThreads)
rows_max_sum_kernel(const L* __restrict__ in, float* __restrict__ out_max, float* __restrict__ out_sum,
                    size_t rows, size_t n, size_t row_stride, int G) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned g = lane & (G - 1);
    const unsigned sub = lane / G;
    const unsigned rows_per_warp = 32 / G;
    const size_t warp = (size_t(blockIdx.x) * kRowThreads + threadIdx.x) >> 5;
    const size_t nwarps = (size_t(gridDim.x) * kRowThreads) >> 5;
    for (size_t rb = warp * rows_per_warp; rb < rows; rb += nwarps * rows_per_warp) {
        const size_t r = rb + sub;
        const bool active = r < rows;
        float mx = OpMax<float>::neutral(), sm = 0.f;
        if (active) span_max_sum<L, VB, U>(in + r * row_stride, n, g, G, mx, sm);
        for (int m = G >> 1; m >= 1; m >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(kFullMask, mx, m));
            sm += __shfl_xor_sync(kFullMask, sm, m);
        }
        if (active && g == 0) { out_max[r] = mx; out_sum[r] = sm; }
    }
}

template <class L>
static cudaError_t launch_rows_max_sum_t(const RowsMaxSumArgs& a, const Launch& l) {
    const int sms = sm_count();
    const int G = pick_group(a.n * sizeof(L));
    const size_t rows_per_cta = size_t(kRowThreads / 32) * (32 / G);
    size_t grid = (a.rows + rows_per_cta - 1) / rows_per_cta;
    const size_t cap = size_t(sms) * resident_ctas<rows_max_sum_kernel<L, 32, 4>>(kRowThreads);
    if (grid > cap) grid = cap;
    if (grid == 0) grid = 1;
    rows_max_sum_kernel<L, 32, 4><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(
        (const L*)a.in, a.out_max, a.out_sum, a.rows, a.n, a.row_stride, G);
    return cudaGetLastError();
}

cudaError_t launch_rows_max_sum(int dtin, const RowsMaxSumArgs& a, const Launch& l) {
    switch (dtin) {
        case DT_F32: return launch_rows_max_sum_t<float>(a, l);
        case DT_BF16: return launch_rows_max_sum_t<__nv_bfloat16>(a, l);
        case DT_F16: return launch_rows_max_sum_t<__half>(a, l);
        default: return cudaErrorNotSupported;
    }
}

}  // namespace db200
