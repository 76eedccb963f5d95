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
#include <cooperative_groups.h>
#include "span.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kRowThreads = 256;

// G lanes per row; 32/G rows per warp per step.  G is a run-time power of two.
template <class L, class A, class S, class Op, int VB, int U, bool SHORT>
__global__ void __launch_bounds__(kRowThreads)
rows_group_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, size_t n, size_t row_stride,
                  ptrdiff_t out_stride, Finish<A> fin, int G, size_t hi_in_stride, ptrdiff_t hi_out_stride) {
    in += size_t(blockIdx.y) * hi_in_stride;          // outer level of rows (RowsArgs::hi_*), one slice per blockIdx.y
    out += ptrdiff_t(blockIdx.y) * hi_out_stride;
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
        if (active) v = span_accumulate<L, A, Op, VB, U, SHORT>(in + r * row_stride, n, g, G);
        for (int m = G >> 1; m >= 1; m >>= 1) v = Op::apply(v, shfl_xor(v, m));
        if (active && g == 0) write_result<Op, A, S>(out + ptrdiff_t(r) * out_stride, v, fin);
    }
}

// One CTA per row (grid-stride over rows).
template <class L, class A, class S, class Op, int VB, int U>
__global__ void __launch_bounds__(kRowThreads)
rows_block_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, size_t n, size_t row_stride,
                  ptrdiff_t out_stride, Finish<A> fin, size_t hi_in_stride, ptrdiff_t hi_out_stride) {
    __shared__ A scratch[32];
    in += size_t(blockIdx.y) * hi_in_stride;
    out += ptrdiff_t(blockIdx.y) * hi_out_stride;
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
    // never a whole multi-vector row per lane: 32 lanes reading 32 rows 128 bytes apart touch 32 lines per load
    // (bf16 [8M, 64]: 2.4 TB/s with one lane per 128-byte row, 6.4 with two)
    if (G == 1 && row_bytes >= 128 && tuning().rows_lane_bytes == 0) G = 2;
    return G;
}

// Tiny dense rows ([R, n <= 128], rows back to back in memory, e.g. [N, 33] feature rows): a row is shorter than
// one coalesced warp request, so the CTA treats its rows as one flat tile - 256 threads load kTinyTile elements with
// scalar coalesced loads, every one of them issued before the first use (16 loads in flight per thread), into
// shared memory; then warps fold rows out of shared memory, WIDTH lanes per row (conflict-free: a row is read by
// consecutive lanes).  The lanes-per-row kernel above reaches 0.4-0.5 of the peak on such rows.
constexpr int kTinyTile = 4096;            // elements per CTA tile
constexpr int kTinyMaxN = 128;
constexpr int kTinyPadMax = 1024;          // room for the per-row padding below

// Padding (elements) between the rows of a shared-memory tile read by `width` lanes per row, so that the lanes of a
// warp - 32 / width rows at a time - land on different banks: the row pitch in bytes must be an odd multiple of the
// bytes one row's lanes span per step (at least a word).  n = 64 floats, 4 lanes per row: pitch 256 B put all 8 rows
// of a warp on the same four banks (measured 2.9 TB/s where n = 100 runs at 5.1); pitch 272 B does not.
static inline unsigned tiny_row_pad(unsigned n, unsigned width, unsigned elem_bytes, unsigned rows_per_tile) {
    unsigned span = width * elem_bytes;
    if (span < 4) span = 4;
    if (span > 64) return 0;
    unsigned pad = 0;
    while (((n + pad) * elem_bytes) % (2 * span) != span && pad < 64) ++pad;
    if (pad >= 64 || size_t(pad) * rows_per_tile > size_t(kTinyPadMax)) return 0;
    return pad;
}
static inline unsigned tiny_magic(unsigned n) { return unsigned((0x100000000ull / n) + 1); }   // i / n for i * n < 2^32

// Stage `count` contiguous elements into the CTA's tile, every load issued before the first store.  `vec16`: the
// tile starts on a 16-byte boundary (the host picks rows_per_tile accordingly), so whole 16-byte vectors go
// global -> shared as they are (a warp of 2-byte scalar loads moves 64 bytes per instruction: [5M, 100] bf16 ran at
// 2.6 TB/s).  The padded layout places every element on its own, so it keeps the scalar path.
template <class L, bool PAD>
__device__ __forceinline__ void tiny_stage(L* __restrict__ tile, const L* __restrict__ src, unsigned count, unsigned pad,
                                           unsigned magic, int vec16) {
    if (!PAD && vec16) {
        constexpr int kV = (kTinyTile * int(sizeof(L)) / 16 + kRowThreads - 1) / kRowThreads;
        const unsigned nv = count * unsigned(sizeof(L)) / 16;
        Words<16> w[kV];
#pragma unroll
        for (int u = 0; u < kV; ++u) {
            const unsigned iv = threadIdx.x + unsigned(u) * kRowThreads;
            if (iv < nv) w[u] = ldg_stream16(reinterpret_cast<const char*>(src) + size_t(iv) * 16);
        }
#pragma unroll
        for (int u = 0; u < kV; ++u) {
            const unsigned iv = threadIdx.x + unsigned(u) * kRowThreads;
            if (iv < nv) reinterpret_cast<Words<16>*>(tile)[iv] = w[u];
        }
        for (unsigned i = nv * (16 / unsigned(sizeof(L))) + threadIdx.x; i < count; i += kRowThreads) tile[i] = ldg_scalar(src + i);
        return;
    }
    constexpr int kPer = kTinyTile / kRowThreads;
    L v[kPer];
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
        const unsigned i = threadIdx.x + unsigned(u) * kRowThreads;
        if (i < count) v[u] = ldg_scalar(src + i);
    }
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
        const unsigned i = threadIdx.x + unsigned(u) * kRowThreads;
        if (i < count) tile[PAD ? i + __umulhi(i, magic) * pad : i] = v[u];        // __umulhi(i, magic) == i / n
    }
}

// rows per tile such that every tile starts on a 16-byte boundary (0: not possible / not worth it)
static inline unsigned tiny_rows_per_tile_vec(unsigned n, unsigned elem_bytes, const void* base) {
    if ((reinterpret_cast<size_t>(base) & 15) != 0) return 0;
    unsigned g = n * elem_bytes, b = 16;
    while (b) { const unsigned t = g % b; g = b; b = t; }      // g = gcd(n * elem_bytes, 16)
    const unsigned m = 16 / g;
    const unsigned rpt = (unsigned(kTinyTile) / n) / m * m;
    return rpt;
}

template <class L, class A, class S, class Op, int WIDTH, bool PAD>
__global__ void __launch_bounds__(kRowThreads)
rows_tiny_kernel(const L* __restrict__ in, S* __restrict__ out, size_t rows, unsigned n, unsigned rows_per_tile,
                 ptrdiff_t out_stride, Finish<A> fin, unsigned pad, unsigned magic, int vec16) {
    // rows sit n + pad elements apart in the tile (tiny_row_pad: no two lanes of a warp on one bank)
    // (PAD = false: the plain layout, no index arithmetic on the way in - most row lengths need no padding)
    __shared__ __align__(16) L tile[kTinyTile + (PAD ? kTinyPadMax : 0)];
    const unsigned np = PAD ? n + pad : n;
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
        tiny_stage<L, PAD>(tile, src, count, pad, magic, vec16);
        __syncthreads();
        if constexpr (WIDTH == 1) {
            // a thread per row: n shared-memory reads at stride n elements - conflict-free for odd n.  For even n
            // every thread starts at column (row mod n) and wraps around: neighbouring threads are then n + 1
            // elements apart, an odd stride again (a fold does not care where it starts).
            const bool rotate = (n & 1u) == 0 && sizeof(L) >= 4 && !PAD;
            for (unsigned r = threadIdx.x; r < nrows; r += kRowThreads) {
                A a0 = Op::neutral(), a1 = Op::neutral();
                const L* p = tile + r * np;
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
                    for (unsigned c = pos; c < n; c += WIDTH) acc = Op::apply(acc, to_acc<A, L>(tile[r * np + c]));
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
    // rows that are whole 32-byte vectors stream through the lanes-per-row kernel at full speed since its lanes
    // take 128 bytes each ([4M, 64] f32: 6.1 TB/s there, 4.3 through the tile)
    if ((a.n * sizeof(L)) % 32 == 0 && (reinterpret_cast<size_t>(a.in) & 31) == 0 &&
        !(sizeof(L) >= 4 && a.n * sizeof(L) == 64))          // (64-byte rows of 4-byte elements: 4.4 vs 3.3 TB/s)
        return cudaErrorNotSupported;
    // ... and so do longer odd rows since that kernel fetches head, vectors and tail in one round trip: the tile
    // keeps rows under 192 bytes of 4- and 8-byte elements ([8M, 33] f32 5.1 vs 4.7 TB/s, [32M, 7] 4.1 vs 2.0; but
    // [2.7M, 100] f32 5.6 vs 6.3 through the lanes, bf16 [5M, 100] 3.4 vs 5.2, f64 [4M, 33] 4.0 vs 5.5)
    if (sizeof(L) < 4 || a.n * sizeof(L) >= 192) return cudaErrorNotSupported;
    const unsigned n = unsigned(a.n);
    const unsigned rpt_vec = tiny_rows_per_tile_vec(n, unsigned(sizeof(L)), a.in);
    const int vec16 = rpt_vec > 0;
    const unsigned rows_per_tile = vec16 ? rpt_vec : kTinyTile / n;
    const size_t ntiles = (a.rows + rows_per_tile - 1) / rows_per_tile;
    Finish<A> fin = make_finish<A, S>(a.identity);
#define DB200_TINY(W)                                                                                                   \
    do {                                                                                                                \
        const unsigned pad = tiny_row_pad(n, W, unsigned(sizeof(L)), rows_per_tile);                                    \
        if (pad == 0) {                                                                                                 \
            size_t grid = size_t(sm_count()) * resident_ctas<rows_tiny_kernel<L, A, S, Op, W, false>>(kRowThreads);      \
            if (grid > ntiles) grid = ntiles;                                                                           \
            rows_tiny_kernel<L, A, S, Op, W, false><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(                       \
                (const L*)a.in, (S*)a.out, a.rows, n, rows_per_tile, a.out_stride, fin, 0u, 0u, vec16);                        \
        } else {                                                                                                        \
            size_t grid = size_t(sm_count()) * resident_ctas<rows_tiny_kernel<L, A, S, Op, W, true>>(kRowThreads);       \
            if (grid > ntiles) grid = ntiles;                                                                           \
            rows_tiny_kernel<L, A, S, Op, W, true><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(                        \
                (const L*)a.in, (S*)a.out, a.rows, n, rows_per_tile, a.out_stride, fin, pad, tiny_magic(n), 0);            \
        }                                                                                                               \
    } while (0)
    // rows under 192 bytes: a tile holds at least 80 of them, a thread folds a whole row
    DB200_TINY(1);
#undef DB200_TINY
    return cudaGetLastError();
}

template <class T, class Op>
static cudaError_t launch_rows_t(const RowsArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    constexpr int VB = 32;
    if (a.hi_count > 65535 || a.hi_count == 0) return cudaErrorNotSupported;
    if (a.hi_count == 1) {
        cudaError_t e = launch_rows_tiny<T, Op>(a, l);
        if (e != cudaErrorNotSupported) return e;
    }
    Finish<A> fin = make_finish<A, S>(a.identity);
    const size_t row_bytes = a.n * sizeof(L);
    const int sms = sm_count();
    const unsigned gy = unsigned(a.hi_count);
    if (row_bytes >= (size_t(32) << 10)) {
        const size_t slots = size_t(sms) * resident_ctas<rows_block_kernel<L, A, S, Op, VB, 4>>(kRowThreads);
        size_t grid = a.rows < slots ? a.rows : slots;
        if (gy > 1) grid = std::max<size_t>(1, std::min(grid, slots / gy));
        rows_block_kernel<L, A, S, Op, VB, 4><<<dim3((unsigned)grid, gy), kRowThreads, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, a.rows, a.n, a.row_stride, a.out_stride, fin, a.hi_in_stride, a.hi_out_stride);
        return cudaGetLastError();
    }
    const int G = pick_group(row_bytes);
    const size_t rows_per_cta = size_t(kRowThreads / 32) * (32 / G);
    size_t grid = (a.rows + rows_per_cta - 1) / rows_per_cta;
    // At most four vectors per lane: one batch with the row's head and tail in the same round trip, no loop.  That
    // instantiation needs 72 registers (3 CTAs per SM instead of 4), so rows that are whole aligned vectors stay with
    // the loop kernel, which is better there ([262144, 1000] f32: 6.4 vs 5.9 TB/s, bf16 [8M, 64]: 6.5 vs 5.8).
    const bool whole = row_bytes % 32 == 0 && (a.row_stride * sizeof(L)) % 32 == 0 && (a.hi_in_stride * sizeof(L)) % 32 == 0 &&
                       (reinterpret_cast<size_t>(a.in) & 31) == 0;
    const bool short_rows = row_bytes / 32 <= size_t(G) * 4 && !whole;
    const size_t cap = size_t(sms) * (short_rows ? resident_ctas<rows_group_kernel<L, A, S, Op, VB, 4, true>>(kRowThreads)
                                                 : resident_ctas<rows_group_kernel<L, A, S, Op, VB, 4, false>>(kRowThreads));   // one full wave
    if (gy > 1) {
        const size_t per_slice = std::max<size_t>(1, cap / gy);          // (slices x CTAs) is about one wave
        if (grid > per_slice) grid = per_slice;
    } else if (grid > cap) grid = cap;
    else if (grid > size_t(sms)) grid = ((grid + sms - 1) / sms) * sms;   // whole CTAs per SM (see the kernel)
    if (grid == 0) grid = 1;
    if (short_rows)
        rows_group_kernel<L, A, S, Op, VB, 4, true><<<dim3((unsigned)grid, gy), kRowThreads, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, a.rows, a.n, a.row_stride, a.out_stride, fin, G, a.hi_in_stride, a.hi_out_stride);
    else
        rows_group_kernel<L, A, S, Op, VB, 4, false><<<dim3((unsigned)grid, gy), kRowThreads, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, a.rows, a.n, a.row_stride, a.out_stride, fin, G, a.hi_in_stride, a.hi_out_stride);
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
template <class L, int VB, int U, bool SHORT = false>       // SHORT: at most U vectors per thread, one predicated batch
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
    if constexpr (SHORT) {
        Words<VB> w[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec) w[u] = ldg_stream<VB>(pv + (v + size_t(u) * nthr) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (v + size_t(u) * nthr < nvec) {
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    float x = vec_elem<L, float, VB>(w[u], j);
                    amax[j % NA] = fmaxf(amax[j % NA], x);
                    asum[j % NA] += x;
                }
            }
    } else {
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
    }
    for (size_t i = h + nvec * N + tid; i < n; i += nthr) {
        float x = to_acc<float, L>(ldg_scalar(p + i));
        amax[0] = fmaxf(amax[0], x); asum[0] += x;
    }
    mx = fmaxf(fmaxf(amax[0], amax[1]), fmaxf(amax[2], amax[3]));
    sm = (asum[0] + asum[1]) + (asum[2] + asum[3]);
}

template <class L, int VB, int U, bool SHORT>
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
        if (active) span_max_sum<L, VB, U, SHORT>(in + r * row_stride, n, g, G, mx, sm);
        for (int m = G >> 1; m >= 1; m >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(kFullMask, mx, m));
            sm += __shfl_xor_sync(kFullMask, sm, m);
        }
        if (active && g == 0) { out_max[r] = mx; out_sum[r] = sm; }
    }
}

// Tiny dense rows (softmax over 10 ... 128 classes): the flat-tile scheme of rows_tiny_kernel - the CTA stages
// kTinyTile elements with coalesced loads, then WIDTH lanes fold each row out of shared memory for both results
// (the lanes-per-row kernel above ran [16M, 33] bf16 at 2.6 TB/s, [8M, 64] at 3.5).
template <class L, int WIDTH, bool PAD>
__global__ void __launch_bounds__(kRowThreads)
rows_max_sum_tiny_kernel(const L* __restrict__ in, float* __restrict__ out_max, float* __restrict__ out_sum, size_t rows,
                         unsigned n, unsigned rows_per_tile, unsigned pad, unsigned magic, int vec16) {
    __shared__ __align__(16) L tile[kTinyTile + (PAD ? kTinyPadMax : 0)];
    const unsigned np = PAD ? n + pad : n;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned seg = lane / WIDTH, pos = lane % WIDTH;
    constexpr unsigned kSegs = 32 / WIDTH;
    const size_t ntiles = (rows + rows_per_tile - 1) / rows_per_tile;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const size_t row0 = t * rows_per_tile;
        const unsigned nrows = (rows - row0 < rows_per_tile) ? unsigned(rows - row0) : rows_per_tile;
        const unsigned count = nrows * n;
        const L* src = in + row0 * n;
        tiny_stage<L, PAD>(tile, src, count, pad, magic, vec16);
        __syncthreads();
        for (unsigned rb = warp * kSegs; rb < nrows; rb += (kRowThreads / 32) * kSegs) {      // warp-uniform trip count
            const unsigned r = rb + seg;
            const bool valid = r < nrows;
            float mx = OpMax<float>::neutral(), sm = 0.f;
            if (valid) {
                const L* p = tile + r * np;
                if constexpr (WIDTH == 1) {
                    // a lane per row: the padded pitch keeps the banks apart; without padding an even n starts at
                    // column (row mod n) and wraps, so that neighbouring lanes are n + 1 elements apart
                    unsigned c = (!PAD && (n & 1u) == 0) ? r % n : 0;
                    for (unsigned j = 0; j < n; ++j) {
                        const float x = to_acc<float, L>(p[c]);
                        mx = fmaxf(mx, x); sm += x;
                        c = (c + 1 == n) ? 0 : c + 1;
                    }
                } else {
                    for (unsigned c = pos; c < n; c += WIDTH) {
                        const float x = to_acc<float, L>(p[c]);
                        mx = fmaxf(mx, x); sm += x;
                    }
                }
            }
#pragma unroll
            for (int m = WIDTH / 2; m >= 1; m >>= 1) {
                mx = fmaxf(mx, __shfl_xor_sync(kFullMask, mx, m));
                sm += __shfl_xor_sync(kFullMask, sm, m);
            }
            if (valid && pos == 0) { out_max[row0 + r] = mx; out_sum[row0 + r] = sm; }
        }
        __syncthreads();
    }
}

// Few long rows (softmax of a small batch over a large vocabulary): a cluster of C CTAs per row, each folding one
// segment; the CTAs of a cluster meet in the first one's shared memory (DSMEM) and are folded there in rank order -
// no workspace, deterministic.  A warp per row left [256, 1M] f32 at 1.1 TB/s.
template <class L, int VB, int U>
__global__ void __launch_bounds__(kRowThreads)
rows_max_sum_cluster_kernel(const L* __restrict__ in, float* __restrict__ out_max, float* __restrict__ out_sum,
                            size_t rows, size_t n, size_t row_stride) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned C = cluster.num_blocks(), rank = cluster.block_rank();
    __shared__ float smx[32], ssm[32];
    __shared__ float cmx[16], csm[16];               // written by the CTAs of the cluster, read by rank 0
    constexpr size_t N = VecInfo<L, VB>::N;
    size_t seg = (n + C - 1) / C;
    seg = (seg + N - 1) / N * N;
    const size_t lo = size_t(rank) * seg < n ? size_t(rank) * seg : n;
    const size_t hi = lo + seg < n ? lo + seg : n;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (size_t r = blockIdx.y; r < rows; r += gridDim.y) {
        float mx = OpMax<float>::neutral(), sm = 0.f;
        if (hi > lo) span_max_sum<L, VB, U>(in + r * row_stride + lo, hi - lo, threadIdx.x, kRowThreads, mx, sm);
#pragma unroll
        for (int m = 16; m >= 1; m >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(kFullMask, mx, m));
            sm += __shfl_xor_sync(kFullMask, sm, m);
        }
        if (lane == 0) { smx[warp] = mx; ssm[warp] = sm; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < kRowThreads / 32; ++w) { mx = fmaxf(mx, smx[w]); sm += ssm[w]; }
            float* rmx = cluster.map_shared_rank(cmx, 0);
            float* rsm = cluster.map_shared_rank(csm, 0);
            rmx[rank] = mx;
            rsm[rank] = sm;
        }
        cluster.sync();
        if (rank == 0 && threadIdx.x == 0) {
            float m0 = cmx[0], s0 = csm[0];
            for (unsigned c = 1; c < C; ++c) { m0 = fmaxf(m0, cmx[c]); s0 += csm[c]; }
            out_max[r] = m0;
            out_sum[r] = s0;
        }
        cluster.sync();                              // rank 0 has read the slots before anyone writes the next row's
    }
}

template <class L>
static cudaError_t launch_rows_max_sum_t(const RowsMaxSumArgs& a, const Launch& l) {
    const int sms = sm_count();
    const bool whole_vectors = (a.n * sizeof(L)) % 32 == 0 && (reinterpret_cast<size_t>(a.in) & 31) == 0;
    if (a.n >= 2 && a.n <= size_t(kTinyMaxN) && a.row_stride == a.n && a.rows * a.n >= (size_t(1) << 22) &&
        !tuning().no_tiny_rows && !whole_vectors && sizeof(L) >= 4 && a.n * sizeof(L) < 192) {
        const unsigned n = unsigned(a.n);
        const unsigned rpt_vec = tiny_rows_per_tile_vec(n, unsigned(sizeof(L)), a.in);
        const int vec16 = rpt_vec > 0;
        const unsigned rows_per_tile = vec16 ? rpt_vec : kTinyTile / n;
        const size_t ntiles = (a.rows + rows_per_tile - 1) / rows_per_tile;
#define DB200_MS_TINY(W)                                                                                              \
        do {                                                                                                          \
            const unsigned pad = tiny_row_pad(n, W, unsigned(sizeof(L)), rows_per_tile);                              \
            if (pad == 0) {                                                                                           \
                size_t grid = size_t(sms) * resident_ctas<rows_max_sum_tiny_kernel<L, W, false>>(kRowThreads);        \
                if (grid > ntiles) grid = ntiles;                                                                     \
                rows_max_sum_tiny_kernel<L, W, false><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(                  \
                    (const L*)a.in, a.out_max, a.out_sum, a.rows, n, rows_per_tile, 0u, 0u, vec16);                          \
            } else {                                                                                                  \
                size_t grid = size_t(sms) * resident_ctas<rows_max_sum_tiny_kernel<L, W, true>>(kRowThreads);         \
                if (grid > ntiles) grid = ntiles;                                                                     \
                rows_max_sum_tiny_kernel<L, W, true><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(                   \
                    (const L*)a.in, a.out_max, a.out_sum, a.rows, n, rows_per_tile, pad, tiny_magic(n), 0);              \
            }                                                                                                         \
        } while (0)
        DB200_MS_TINY(1);
#undef DB200_MS_TINY
        return cudaGetLastError();
    }
    const size_t slots = size_t(sms) * resident_ctas<rows_max_sum_cluster_kernel<L, 32, 4>>(kRowThreads);
    if (a.n * sizeof(L) >= (size_t(64) << 10) && a.rows < slots * (kRowThreads / 32)) {
        // fewer rows than warps in a wave: CTAs (clusters of up to 8 along the row) instead of warps
        unsigned C = 1;
        while (C < 8 && a.rows * C < slots && a.n * sizeof(L) / (2 * C) >= (size_t(32) << 10)) C <<= 1;
        size_t gy = a.rows < 65535 ? a.rows : 65535;
        if (gy * C > slots) gy = slots / C;
        if (gy == 0) gy = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(C, (unsigned)gy, 1);
        cfg.blockDim = dim3(kRowThreads, 1, 1);
        cfg.stream = l.stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = C;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, rows_max_sum_cluster_kernel<L, 32, 4>, (const L*)a.in, a.out_max, a.out_sum, a.rows,
                                  a.n, a.row_stride);
    }
    const int G = pick_group(a.n * sizeof(L));
    const size_t rows_per_cta = size_t(kRowThreads / 32) * (32 / G);
    size_t grid = (a.rows + rows_per_cta - 1) / rows_per_cta;
    if (a.n * sizeof(L) / 32 <= size_t(G) * 4) {             // at most four vectors per lane
        const size_t cap = size_t(sms) * resident_ctas<rows_max_sum_kernel<L, 32, 4, true>>(kRowThreads);
        if (grid > cap) grid = cap;
        if (grid == 0) grid = 1;
        rows_max_sum_kernel<L, 32, 4, true><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(
            (const L*)a.in, a.out_max, a.out_sum, a.rows, a.n, a.row_stride, G);
        return cudaGetLastError();
    }
    const size_t cap = size_t(sms) * resident_ctas<rows_max_sum_kernel<L, 32, 4, false>>(kRowThreads);
    if (grid > cap) grid = cap;
    if (grid == 0) grid = 1;
    rows_max_sum_kernel<L, 32, 4, false><<<(unsigned)grid, kRowThreads, 0, l.stream>>>(
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
