// C ABI + host-side planner of the B200 reduction backend (see
// include/dace_b200_reduce.h for the contract).
//
// The planner canonicalises an arbitrary memlet subset + axis set the way the
// reference's reduction planner does - drop unit dims, merge contiguous dims of
// the same kind (reference dace/libraries/standard/reduction_planner.py:44-118) -
// and then picks one of the hand-written kernels instead of emitting a schedule.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include "../../include/dace_b200_reduce.h"
#include "common.cuh"
#include "kernels.h"

namespace db200 {

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static Tuning read_tuning() {
    auto flag = [](const char* n) { return getenv(n) != nullptr; };
    auto num = [](const char* n, int dflt) { const char* v = getenv(n); return v ? atoi(v) : dflt; };
    Tuning t;
    t.no_two_pass = flag("DACE_B200_NO_TWO_PASS");
    t.no_axis_sort = flag("DACE_B200_NO_AXIS_SORT");
    t.no_peel = flag("DACE_B200_NO_PEEL");
    t.no_wide_cols = flag("DACE_B200_NO_WIDE_COLS");
    t.no_few_rows = flag("DACE_B200_NO_FEW_ROWS");
    t.no_narrow_cols = flag("DACE_B200_NO_NARROW_COLS");
    t.no_partition = flag("DACE_B200_NO_PARTITION");
    t.no_tiled = flag("DACE_B200_NO_TILED");
    t.no_tiny_rows = flag("DACE_B200_NO_TINY_ROWS");
    t.no_outer_first = flag("DACE_B200_NO_OUTER_FIRST");
    t.wide_wc = num("DACE_B200_WIDE_WC", 0);
    t.part_chunk_log2 = num("DACE_B200_PART_CHUNK_LOG2", 0);
    t.rows_lane_bytes = num("DACE_B200_ROWS_LANE_BYTES", 0);
    t.part_divert = num("DACE_B200_PART_DIVERT", -1);
    t.part_hint = num("DACE_B200_PART_HINT", -1);
    return t;
}
static const Tuning g_tuning = read_tuning();   // at load time
const Tuning& tuning() { return g_tuning; }

// The workspace is the caller's (contract of include/dace_b200_reduce.h): nothing is allocated here.
static cudaError_t resolve_workspace(void* workspace, size_t bytes, Launch& l) {
    if (workspace == nullptr || bytes < kWorkspaceBytes) return cudaErrorInvalidValue;
    l.ws.counters = reinterpret_cast<unsigned*>(workspace);
    l.ws.partials = reinterpret_cast<char*>(workspace) + kCounterBytes;
    l.extra = reinterpret_cast<char*>(workspace) + kWorkspaceBytes;
    l.extra_bytes = bytes - kWorkspaceBytes;
    return cudaSuccess;
}

struct Dim {
    long long shape, in_stride, out_stride;
    bool reduced;
};

static inline size_t ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }

static bool supported_pair(int dtin, int dtout) {
    return dispatch_types(dtin, dtout, [](auto) -> cudaError_t { return cudaSuccess; }) == cudaSuccess;
}

// Full reduction / few long rows through the split kernel.
static cudaError_t plan_split(int red, int dtin, int dtout, const void* in, void* out, size_t rows, size_t n,
                              size_t row_stride, ptrdiff_t out_stride, const void* identity, const Launch& l) {
    l.note("split");
    if (l.dry_need) return cudaSuccess;
    const size_t row_bytes = n * dtype_size(dtin);
    const size_t sms = sm_count();
    // Large rows: 1 CTA of 1024 threads per SM, 256 KiB tiles; otherwise 2 x 512 threads, 64 KiB tiles.
    const int big = row_bytes * rows >= (size_t(256) << 20) && row_bytes >= (size_t(2) << 20);
    const size_t tile = big ? (size_t(256) << 10) : (size_t(64) << 10);
    size_t want = ceil_div(row_bytes, tile);
    const size_t slots = (big ? 1 : 2) * sms;            // CTAs resident at a time
    size_t cap = ceil_div(slots, rows);
    if (rows > 1 && want > 1) {
        // Several rows: rows x splits CTAs run in waves of `slots`; pick the split count whose last wave is
        // the fullest (3 rows x 50 splits = 150 CTAs would leave 2 CTAs alone in a second wave: measured 0.78
        // of the one-row bandwidth; 3 x 49 = 147 fills one wave).  Even counts only (clusters of 2).
        const size_t smax = std::min(want, std::max<size_t>(cap, 64));
        size_t best = 1;
        double best_eff = 0.0;
        for (size_t sp = 1; sp <= smax; sp = (sp == 1 ? 2 : sp + 2)) {
            const size_t ctas = rows * sp;
            const double eff = double(ctas) / double(ceil_div(ctas, slots) * slots);
            if (eff > best_eff + 1e-9) { best_eff = eff; best = sp; }
        }
        want = best;
    } else if (want > cap) {
        want = cap;
    }
    if (want < 1) want = 1;
    unsigned cluster = 1;
    if (want >= 2) { cluster = 2; want = (want / 2) * 2; }
    // Partials area: rows * clusters accumulators of <= 8 bytes.
    while (want > cluster && rows * (want / cluster) * 8 > kPartialBytes) want -= cluster;
    SplitArgs a{in, out, n, rows, row_stride, out_stride, (unsigned)want, cluster, identity, big};
    return launch_split(red, dtin, dtout, a, l);
}

// [K, R] with contiguous R: few long rows -> split kernel, else the rows kernel.
static cudaError_t plan_rows(int red, int dtin, int dtout, const void* in, void* out, size_t rows, size_t n,
                             size_t row_stride, ptrdiff_t out_stride, const void* identity, const Launch& l,
                             size_t hi_count = 1, size_t hi_in_stride = 0, ptrdiff_t hi_out_stride = 0) {
    if (hi_count > 65535) return cudaErrorNotSupported;
    const size_t row_bytes = n * dtype_size(dtin);
    if (hi_count == 1 && rows < 2 * (size_t)sm_count() && row_bytes >= (size_t(128) << 10) && rows <= kCounters)
        return plan_split(red, dtin, dtout, in, out, rows, n, row_stride, out_stride, identity, l);
    l.note(hi_count > 1 ? "rows2" : "rows");
    if (l.dry_need) return cudaSuccess;
    RowsArgs a{in, out, rows, n, row_stride, out_stride, identity};
    a.hi_count = hi_count; a.hi_in_stride = hi_in_stride; a.hi_out_stride = hi_out_stride;
    return launch_rows(red, dtin, dtout, a, l);
}

static cudaError_t plan_cols(int red, int dtin, int dtout, const ColsArgs& a, const Launch& l) {
    l.note("cols");
    if (l.dry_need) return cudaSuccess;
    return launch_cols(red, dtin, dtout, a, l);
}

// Host copy of Op::neutral() for the accumulator types a two-pass plan stores between its passes.
static bool neutral_bytes(int red, int dt, unsigned char out[8]) {
    memset(out, 0, 8);
    auto put = [&](auto v) { memcpy(out, &v, sizeof(v)); return true; };
    const bool mn = red == RT_Min, mx = red == RT_Max;
    switch (red) {
        case RT_Sum: case RT_BOr: case RT_BXor: case RT_LOr: case RT_LXor: return true;   // all-zero bits
        case RT_Product: case RT_LAnd:
            switch (dt) {
                case DT_F32: return put(1.0f);
                case DT_F64: return put(1.0);
                case DT_I32: case DT_U32: return put((int)1);
                case DT_I64: case DT_U64: return put((long long)1);
                case DT_I16: case DT_U16: return put((short)1);
                case DT_I8: case DT_U8: return put((signed char)1);
                default: return false;
            }
        case RT_BAnd: memset(out, 0xff, 8); return dt != DT_F32 && dt != DT_F64;
        case RT_Min: case RT_Max:
            switch (dt) {
                case DT_F32: return put(mn ? 0x7f800000u : 0xff800000u);
                case DT_F64: return put(mn ? 0x7ff0000000000000ull : 0xfff0000000000000ull);
                case DT_I32: return put(mn ? std::numeric_limits<int>::max() : std::numeric_limits<int>::lowest());
                case DT_U32: return put(mn ? std::numeric_limits<unsigned>::max() : 0u);
                case DT_I64: return put(mn ? std::numeric_limits<long long>::max() : std::numeric_limits<long long>::lowest());
                case DT_U64: return put(mn ? std::numeric_limits<unsigned long long>::max() : 0ull);
                case DT_I16: return put(mn ? std::numeric_limits<short>::max() : std::numeric_limits<short>::lowest());
                case DT_U16: return put(mn ? std::numeric_limits<unsigned short>::max() : (unsigned short)0);
                case DT_I8: return put(mn ? std::numeric_limits<signed char>::max() : std::numeric_limits<signed char>::lowest());
                case DT_U8: return put(mn ? std::numeric_limits<unsigned char>::max() : (unsigned char)0);
                default: return false;
            }
        default: (void)mx; return false;
    }
}

// Scratch between the passes of a two-pass plan: the part of the caller's workspace past the base
// (dace_b200_reduce_nd_workspace_bytes() says how much a shape wants); a workspace without it sends
// the shape to the slower scratch-free kernels.
constexpr size_t kMidMaxBytes = size_t(256) << 20;

static size_t two_pass_bytes(int dtin, int dtout, size_t O, size_t R1, size_t K) {
    const int dtmid = (dtin == DT_BF16 || dtin == DT_F16) ? DT_F32 : dtin;
    const bool half_out = dtmid == DT_F32 && (dtout == DT_BF16 || dtout == DT_F16);
    const size_t mid_bytes = (O * R1 * K * dtype_size(dtmid) + 255) & ~size_t(255);
    return mid_bytes + (half_out ? ((O * K * sizeof(float) + 255) & ~size_t(255)) : 0);
}

// [R1, K, R2] with contiguous R2 (e.g. per-channel statistics of an [N, C, H*W] array: axes (0, 2)):
// pass 1 folds every contiguous R2 run with the row kernels into mid[R1, K] (neutral seed),
// pass 2 folds mid over R1 with the column kernel and applies the caller's identity.  Both passes
// are the streaming kernels; the generic strided kernel it replaces ran one warp per output.
static cudaError_t plan_two_pass(int red, int dtin, int dtout, const void* in, void* out, size_t O, size_t R1, size_t K,
                                 size_t R2, size_t row_stride, ptrdiff_t out_stride_o, ptrdiff_t out_stride_k,
                                 const void* identity, const Launch& l) {
    const int dtmid = (dtin == DT_BF16 || dtin == DT_F16) ? DT_F32 : dtin;
    // 16-bit output: the second pass stays in fp32 (K values behind mid), a third tiny launch folds the
    // identity and rounds once.
    const bool half_out = dtmid == DT_F32 && (dtout == DT_BF16 || dtout == DT_F16);
    unsigned char neutral[8];
    if (!supported_pair(dtin, dtmid) || !(half_out || supported_pair(dtmid, dtout)) || !neutral_bytes(red, dtmid, neutral))
        return cudaErrorNotSupported;
    // O > 1: an outer kept axis ([O, R1, K, R2] dense, e.g. per-sample per-channel statistics): the rows pass
    // covers O * R1 * K runs, the column pass folds R1 for every (o, k).
    const size_t mid_bytes = (O * R1 * K * dtype_size(dtmid) + 255) & ~size_t(255);
    const size_t need = two_pass_bytes(dtin, dtout, O, R1, K);
    if (half_out && O != 1) return cudaErrorNotSupported;
    if (need > kMidMaxBytes) return cudaErrorNotSupported;
    if (l.dry_need) { l.take(need); l.note("two-pass(rows,cols)"); return cudaSuccess; }
    if (need > l.extra_bytes) return cudaErrorNotSupported;
    void* mid = l.extra;
    cudaError_t e;
    e = plan_rows(red, dtin, dtmid, in, mid, O * R1 * K, R2, row_stride, 1, neutral, l);
    if (e != cudaSuccess) return e;
    if (half_out) {
        float* tmp = reinterpret_cast<float*>(static_cast<char*>(mid) + mid_bytes);
        ColsArgs a{mid, tmp, 1, R1, K, 0, K, 0, 1, neutral};
        e = plan_cols(red, DT_F32, DT_F32, a, l);
        if (e != cudaSuccess) return e;
        e = launch_finish_convert(red, dtout, tmp, out, K, out_stride_k, identity, l);
    } else {
        ColsArgs a{mid, out, O, R1, K, R1 * K, K, out_stride_o, out_stride_k, identity};
        e = plan_cols(red, dtmid, dtout, a, l);
    }
    return e;
}

// The same layout, OUTER axis first, for dense [R1, K, R2] with R1 >= R2: pass 1 folds R1 with the column kernels
// over rows of K * R2 contiguous elements into mid[K, R2] (the widest, best-streaming form the array has), pass 2
// folds the R2 runs of that small intermediate.  The intermediate is 1 / R1 of the input instead of 1 / R2, and a
// short innermost axis never meets the tiny-row kernel at full size ([4096, 512, 8] f32: 35 -> ~17 us).
static size_t outer_first_bytes(int dtin, size_t K, size_t R2) {
    const int dtmid = (dtin == DT_BF16 || dtin == DT_F16) ? DT_F32 : dtin;
    return (K * R2 * dtype_size(dtmid) + 255) & ~size_t(255);
}

static cudaError_t plan_outer_first(int red, int dtin, int dtout, const void* in, void* out, size_t R1, size_t K,
                                    size_t R2, ptrdiff_t out_stride_k, const void* identity, const Launch& l) {
    const int dtmid = (dtin == DT_BF16 || dtin == DT_F16) ? DT_F32 : dtin;
    unsigned char neutral[8];
    if (!supported_pair(dtin, dtmid) || !supported_pair(dtmid, dtout) || !neutral_bytes(red, dtmid, neutral))
        return cudaErrorNotSupported;
    const size_t need = outer_first_bytes(dtin, K, R2);
    if (need > kMidMaxBytes) return cudaErrorNotSupported;
    if (l.dry_need) { l.take(need); l.note("outer-first(cols,rows)"); return cudaSuccess; }
    if (need > l.extra_bytes) return cudaErrorNotSupported;
    void* mid = l.extra;
    ColsArgs a{in, mid, 1, R1, K * R2, 0, K * R2, 0, 1, neutral};
    cudaError_t e = plan_cols(red, dtin, dtmid, a, l);
    if (e != cudaSuccess) return e;
    return plan_rows(red, dtmid, dtout, mid, out, K, R2, R2, out_stride_k, identity, l);
}

// The canonical forms that have a streaming kernel; cudaErrorNotSupported = none of them.
static cudaError_t plan_fast(int red, int dtin, int dtout, const void* in, void* out, const Dim* c, int nc, int nr,
                             const void* identity, const Launch& l) {
    if (nr > 0) {
        // [R] contiguous -> full reduction.
        if (nc == 1 && c[0].in_stride == 1)
            return plan_split(red, dtin, dtout, in, out, 1, (size_t)c[0].shape, 0, 0, identity, l);
        // [K, R] with contiguous R -> rows.
        if (nc == 2 && !c[0].reduced && c[1].reduced && c[1].in_stride == 1 && c[0].in_stride >= 0) {
            return plan_rows(red, dtin, dtout, in, out, (size_t)c[0].shape, (size_t)c[1].shape, (size_t)c[0].in_stride,
                             (ptrdiff_t)c[0].out_stride, identity, l);
        }
        // [K1, K2, R]: rows of a window of a larger array (a pitch after every K2 rows) -> rows on two levels.
        if (nc == 3 && !c[0].reduced && !c[1].reduced && c[2].reduced && c[2].in_stride == 1 && c[0].in_stride >= 0 &&
            c[1].in_stride >= 0 && c[0].shape <= 65535) {
            cudaError_t e = plan_rows(red, dtin, dtout, in, out, (size_t)c[1].shape, (size_t)c[2].shape, (size_t)c[1].in_stride,
                                      (ptrdiff_t)c[1].out_stride, identity, l, (size_t)c[0].shape, (size_t)c[0].in_stride,
                                      (ptrdiff_t)c[0].out_stride);
            if (e != cudaErrorNotSupported) return e;
        }
        // [R, K] / [K1, R, K2] with contiguous trailing K -> cols.
        if (nc == 2 && c[0].reduced && !c[1].reduced && c[1].in_stride == 1 && c[0].in_stride >= 0) {
            ColsArgs a{in, out, 1, (size_t)c[0].shape, (size_t)c[1].shape, 0, (size_t)c[0].in_stride,
                       0, (ptrdiff_t)c[1].out_stride, identity};
            return plan_cols(red, dtin, dtout, a, l);
        }
        if (nc == 3 && !c[0].reduced && c[1].reduced && !c[2].reduced && c[2].in_stride == 1 &&
            c[0].in_stride >= 0 && c[1].in_stride >= 0) {
            ColsArgs a{in, out, (size_t)c[0].shape, (size_t)c[1].shape, (size_t)c[2].shape,
                       (size_t)c[0].in_stride, (size_t)c[1].in_stride,
                       (ptrdiff_t)c[0].out_stride, (ptrdiff_t)c[2].out_stride, identity};
            return plan_cols(red, dtin, dtout, a, l);
        }
        // [R, K1, K2] where K1 and K2 do not merge (a window: a pitch after every K2 run; or an output laid out
        // differently) -> the same kernel, K1 as its outer axis: the strides are all it needs.  Only for the slices of a
        // peeled reduced axis ([R0, R, K1, K2] with a pitch on every level: 0.55 -> 3.3 TB/s); on its own the
        // thread-per-output strided kernel serves this form better (5.5 against 4.1 TB/s on [1024, 500, 500]).
        if (l.slice_of_fold && nc == 3 && c[0].reduced && !c[1].reduced && !c[2].reduced && c[2].in_stride == 1 && c[0].in_stride >= 0 &&
            c[1].in_stride >= 0 && c[1].shape <= 65535) {
            ColsArgs a{in, out, (size_t)c[1].shape, (size_t)c[0].shape, (size_t)c[2].shape,
                       (size_t)c[1].in_stride, (size_t)c[0].in_stride,
                       (ptrdiff_t)c[1].out_stride, (ptrdiff_t)c[2].out_stride, identity};
            return plan_cols(red, dtin, dtout, a, l);
        }
    }
    // [R1, K, R2] with contiguous R2 and dense (R1, K): two streaming passes through a small scratch.
    if (nr > 0 && nc == 3 && c[0].reduced && !c[1].reduced && c[2].reduced && c[2].in_stride == 1 &&
        c[1].in_stride >= c[2].shape && c[0].in_stride == c[1].in_stride * c[1].shape && c[2].shape >= 2 &&
        !tuning().no_two_pass) {
        if (c[1].in_stride == c[2].shape && c[0].shape >= c[2].shape && !tuning().no_outer_first) {   // fully dense, long outer axis
            cudaError_t e1 = plan_outer_first(red, dtin, dtout, in, out, (size_t)c[0].shape, (size_t)c[1].shape,
                                              (size_t)c[2].shape, (ptrdiff_t)c[1].out_stride, identity, l);
            if (e1 != cudaErrorNotSupported) return e1;
        }
        cudaError_t e2 = plan_two_pass(red, dtin, dtout, in, out, 1, (size_t)c[0].shape, (size_t)c[1].shape,
                                       (size_t)c[2].shape, (size_t)c[1].in_stride, 0, (ptrdiff_t)c[1].out_stride,
                                       identity, l);
        if (e2 != cudaErrorNotSupported) return e2;
    }
    // [O, R1, K, R2]: the same with an outer kept axis.
    if (nr > 0 && nc == 4 && !c[0].reduced && c[1].reduced && !c[2].reduced && c[3].reduced && c[3].in_stride == 1 &&
        c[2].in_stride >= c[3].shape && c[1].in_stride == c[2].in_stride * c[2].shape &&
        c[0].in_stride == c[1].in_stride * c[1].shape && c[3].shape >= 2 && dtype_size(dtout) >= 4 &&
        !tuning().no_two_pass) {
        cudaError_t e2 = plan_two_pass(red, dtin, dtout, in, out, (size_t)c[0].shape, (size_t)c[1].shape,
                                       (size_t)c[2].shape, (size_t)c[3].shape, (size_t)c[2].in_stride,
                                       (ptrdiff_t)c[0].out_stride, (ptrdiff_t)c[2].out_stride, identity, l);
        if (e2 != cudaErrorNotSupported) return e2;
    }
    return cudaErrorNotSupported;
}

// Merge neighbours of the same kind when stride[i] == stride[i+1] * shape[i+1] (on both sides for kept axes).
static void merge_dims(const Dim* nz, int nnz, Dim* c, int& nc) {
    nc = 0;
    for (int i = 0; i < nnz; ++i) {
        if (nz[i].shape == 1) continue;
        if (nc > 0 && c[nc - 1].reduced == nz[i].reduced &&
            c[nc - 1].in_stride == nz[i].in_stride * nz[i].shape &&
            (nz[i].reduced || c[nc - 1].out_stride == nz[i].out_stride * nz[i].shape)) {
            c[nc - 1].shape *= nz[i].shape;
            c[nc - 1].in_stride = nz[i].in_stride;
            c[nc - 1].out_stride = nz[i].out_stride;
        } else {
            c[nc++] = nz[i];
        }
    }
}

// Gather + canonicalise (shape, strides, axis mask): default output strides are C-contiguous over the kept axes;
// unit dims are dropped, the axes put in memory order (largest input stride first, stable) and neighbours of
// the same kind merged.  A fold does not care in which order the axes are walked and every kept axis carries
// its own output stride, so transposed / permuted views (np.sum(A.T, axis=1), channels-last tensors)
// canonicalise to the same [K, R] / [R, K] / [R1, K, R2] forms as their plain counterparts instead of
// falling through to the strided kernel.
static cudaError_t canonicalise(int ndim, const int64_t* shape, const int64_t* in_strides, const int64_t* out_strides,
                                uint32_t mask, Dim* c, int& nc, size_t& nout, size_t& nred) {
    if (ndim < 0 || ndim > DACE_B200_MAX_DIMS) return cudaErrorInvalidValue;
    if (ndim > 0 && (!shape || !in_strides)) return cudaErrorInvalidValue;
    Dim dims[DACE_B200_MAX_DIMS];
    long long kstride[DACE_B200_MAX_DIMS];
    {
        long long s = 1;
        for (int i = ndim - 1; i >= 0; --i) {
            kstride[i] = 0;
            if (!((mask >> i) & 1u)) { kstride[i] = s; s *= shape[i]; }
        }
    }
    nout = 1; nred = 1;
    int kept_seen = 0;
    for (int i = 0; i < ndim; ++i) {
        if (shape[i] < 0) return cudaErrorInvalidValue;
        dims[i].shape = shape[i];
        dims[i].in_stride = in_strides[i];
        dims[i].reduced = (mask >> i) & 1u;
        if (dims[i].reduced) { dims[i].out_stride = 0; nred *= (size_t)shape[i]; }
        else {
            dims[i].out_stride = out_strides ? out_strides[kept_seen] : kstride[i];
            ++kept_seen;
            nout *= (size_t)shape[i];
        }
    }
    Dim nz[DACE_B200_MAX_DIMS];
    int nnz = 0;
    for (int i = 0; i < ndim; ++i)
        if (dims[i].shape != 1) nz[nnz++] = dims[i];
    if (!tuning().no_axis_sort)
        std::stable_sort(nz, nz + nnz, [](const Dim& a, const Dim& b) { return a.in_stride > b.in_stride; });
    merge_dims(nz, nnz, c, nc);
    return cudaSuccess;
}

static cudaError_t plan_dims(int red, int dtin, int dtout, const void* in, void* out, const Dim* c, int nc,
                             const void* identity, const Launch& l, bool copy_only, int depth);

// Interleaved kept / reduced axes with no direct kernel ([R1, K1, R2, K2], [K0, R1, K2, R3, K4], ...): fold the
// INNERMOST reduced axis first with a streaming kernel, every axis outside it treated as kept - they merge
// into one outer index when they are dense - into a contiguous intermediate of the accumulator type (1 / R of
// the input), then plan the rest on that intermediate (same routine, one reduced axis fewer; the levels'
// intermediates sit one after the other in the caller's scratch).  Every pass reads contiguous memory with
// the column / row kernels; the strided kernel this replaces measured 0.18 - 0.21 of the HBM peak on these
// layouts (profiles/r1h_awkward_shapes_after.txt).  Deterministic like its parts.
static cudaError_t plan_inner_first(int red, int dtin, int dtout, const void* in, void* out, const Dim* c, int nc,
                                    const void* identity, const Launch& l, int depth) {
    if (tuning().no_two_pass || depth >= 4) return cudaErrorNotSupported;
    int j = -1;
    for (int i = 0; i < nc; ++i)
        if (c[i].reduced) j = i;
    if (j < 1 || nc - 1 - j > 1) return cudaErrorNotSupported;          // needs outer axes; at most one kept axis inside
    const bool has_inner = j == nc - 2;
    if (c[nc - 1].in_stride != 1 || c[j].in_stride < 1 || c[j].shape < 2) return cudaErrorNotSupported;
    bool dense = true;
    for (int i = 0; i + 1 < j; ++i)                                        // the outer axes form one dense index ...
        if (c[i].in_stride != c[i + 1].in_stride * c[i + 1].shape) dense = false;
    // ... or there are exactly two of them and the innermost axis is the folded one: rows on two levels (windows of a
    // larger 3-D array: every axis combination that folds the innermost axis ended in the strided kernel - all
    // elements of a [1024, 500, 500] window of a [1024, 512, 512] array with one warp, 1.35 s)
    // (more outer axes are fine as long as they fall into two dense groups: [N, C, H, W] interiors folded over
    // (0, 2, 3) have N and C dense and a pitch after H)
    int brk = -1, nbrk = 0;
    for (int i = 0; i + 1 < j; ++i)
        if (c[i].in_stride != c[i + 1].in_stride * c[i + 1].shape) { brk = i; ++nbrk; }
    size_t hi_count = 1, lo_count = 1;
    for (int i = 0; i < j; ++i) (i <= brk ? hi_count : lo_count) *= (size_t)c[i].shape;
    const bool two_level = !dense && nbrk == 1 && j == nc - 1 && c[brk].in_stride >= 0 && hi_count <= 65535;
    if (!dense && !two_level) return cudaErrorNotSupported;
    if (c[j - 1].in_stride < 0) return cudaErrorNotSupported;
    const int dtmid = (dtin == DT_BF16 || dtin == DT_F16) ? DT_F32 : dtin;
    // 16-bit output behind an fp32 intermediate (bf16 -> bf16, f16 -> f16: no fp32 -> 16-bit kernel pair exists): the
    // rest of the fold stays in fp32 into a small buffer behind the intermediate, one tiny launch folds the
    // caller's identity and rounds once (as in plan_two_pass).  At most one kept axis group may remain.
    const bool half_out = dtmid == DT_F32 && (dtout == DT_BF16 || dtout == DT_F16);
    unsigned char neutral[8];
    if (!supported_pair(dtin, dtmid) || !(half_out || supported_pair(dtmid, dtout)) || !neutral_bytes(red, dtmid, neutral))
        return cudaErrorNotSupported;
    size_t O = 1;
    for (int i = 0; i < j; ++i) O *= (size_t)c[i].shape;
    const size_t K = has_inner ? (size_t)c[nc - 1].shape : 1;
    const size_t R = (size_t)c[j].shape;
    size_t kept_out = 1;
    int kept_groups = 0, kept_at = -1;
    for (int i = 0; i < nc; ++i)
        if (!c[i].reduced) { kept_out *= (size_t)c[i].shape; ++kept_groups; kept_at = i; }
    if (half_out && kept_groups > 1) return cudaErrorNotSupported;
    const size_t fin_bytes = half_out ? ((kept_out * sizeof(float) + 255) & ~size_t(255)) : 0;
    const size_t mid_bytes = ((O * K * dtype_size(dtmid) + 255) & ~size_t(255)) + fin_bytes;
    if (mid_bytes > kMidMaxBytes) return cudaErrorNotSupported;
    if (!l.dry_need && mid_bytes > l.extra_bytes) return cudaErrorNotSupported;
    l.take(mid_bytes);
    l.note("inner-first:");
    void* mid = l.extra;
    cudaError_t e;
    if (has_inner) {
        ColsArgs a{in, mid, O, R, K, (size_t)c[j - 1].in_stride, (size_t)c[j].in_stride, (ptrdiff_t)K, 1, neutral};
        e = plan_cols(red, dtin, dtmid, a, l);
    } else if (two_level) {
        e = plan_rows(red, dtin, dtmid, in, mid, lo_count, R, (size_t)c[j - 1].in_stride, 1, neutral, l,
                      hi_count, (size_t)c[brk].in_stride, (ptrdiff_t)lo_count);
    } else {
        e = plan_rows(red, dtin, dtmid, in, mid, O, R, (size_t)c[j - 1].in_stride, 1, neutral, l);
    }
    if (e != cudaSuccess) return e;
    // the rest of the fold, on the contiguous intermediate [c[0] ... c[j-1], K]
    Dim rest[DACE_B200_MAX_DIMS], merged[DACE_B200_MAX_DIMS];
    int nrest = 0, nmerged = 0;
    long long stride = (long long)K;
    for (int i = j - 1; i >= 0; --i) {
        rest[i] = c[i];
        rest[i].in_stride = stride;
        stride *= c[i].shape;
    }
    nrest = j;
    if (has_inner) { rest[nrest] = c[nc - 1]; rest[nrest].in_stride = 1; ++nrest; }
    merge_dims(rest, nrest, merged, nmerged);
    Launch l2 = l;
    if (!l.dry_need) {
        l2.extra = static_cast<char*>(l.extra) + mid_bytes;
        l2.extra_bytes = l.extra_bytes - mid_bytes;
    }
    l2.dry_base = l.dry_base + mid_bytes;
    if (half_out) {
        float* fin = l.dry_need ? nullptr : reinterpret_cast<float*>(static_cast<char*>(l.extra) + (mid_bytes - fin_bytes));
        for (int i = 0; i < nmerged; ++i)
            if (!merged[i].reduced) merged[i].out_stride = 1;            // the one kept group, dense in `fin`
        e = plan_dims(red, DT_F32, DT_F32, mid, fin, merged, nmerged, neutral, l2, false, depth + 1);
        if (e != cudaSuccess) return e;
        l.note("finish");
        if (l.dry_need) return cudaSuccess;
        return launch_finish_convert(red, dtout, fin, out, kept_out, kept_at >= 0 ? (ptrdiff_t)c[kept_at].out_stride : 1,
                                     identity, l);
    }
    return plan_dims(red, dtmid, dtout, mid, out, merged, nmerged, identity, l2, false, depth + 1);
}

// Everything behind the canonical form: pick a plan and launch it (or, in a dry run, add up its scratch).
static cudaError_t plan_dims(int red, int dtin, int dtout, const void* in, void* out, const Dim* c, int nc,
                             const void* identity, const Launch& l, bool copy_only, int depth) {
    size_t nout = 1, nred = 1;
    int nk = 0, nr = 0;
    for (int i = 0; i < nc; ++i) {
        if (c[i].reduced) { nred *= (size_t)c[i].shape; ++nr; } else { nout *= (size_t)c[i].shape; ++nk; }
    }
    if (nout == 0) return cudaSuccess;
    if (nred == 0) {
        // Empty range: outputs become the identity; without one they stay as they are.
        if (!identity || l.dry_need) return cudaSuccess;
        long long ks[kMaxDims], ko[kMaxDims];
        int n = 0;
        for (int i = 0; i < nc; ++i)
            if (!c[i].reduced) { if (n >= kMaxDims) return cudaErrorNotSupported; ks[n] = c[i].shape; ko[n] = c[i].out_stride; ++n; }
        return launch_fill(dtout, out, n, ks, ko, identity, l);
    }
    {
        cudaError_t ef = plan_fast(red, dtin, dtout, in, out, c, nc, nr, identity, l);
        if (ef != cudaErrorNotSupported) return ef;
    }
    // (also with nothing kept: all elements of a padded view, [R1, R2] with a pitch - the rows go into an
    // intermediate, the intermediate is one contiguous run; the strided kernel folded such a view with ONE warp:
    // 0.96 s for a [16384, 16000] window of a [16384, 16384] array)
    if (nr > 0 && nc >= 2 && (nk > 0 ? nc >= 3 : true)) {
        cudaError_t ei = plan_inner_first(red, dtin, dtout, in, out, c, nc, identity, l, depth);
        if (ei != cudaErrorNotSupported) return ei;
    }
    // Interleaved kept / reduced axes ([R1, K1, R2, K2], [K1, R1, K2, R2], ...): peel the outermost axis on the host
    // when it is short and what remains has a streaming kernel - a reduced axis folds slice after slice into the
    // output (identity for the first slice, "fold into existing" for the rest), a kept axis writes slice by slice.
    if (nr > 0 && nc >= 3 && c[0].shape <= 64 && c[0].in_stride >= 0 && !tuning().no_peel) {
        int sub_nr = 0;
        for (int i = 1; i < nc; ++i) sub_nr += c[i].reduced ? 1 : 0;
        const size_t isz = dtype_size(dtin), osz = dtype_size(dtout);
        size_t slice_elems = 1;
        for (int i = 1; i < nc; ++i) slice_elems *= (size_t)c[i].shape;
        // a reduced outer axis is where the strided kernel is weakest (one thread folds everything for its output);
        // with a kept outer axis it keeps its warp-per-output mapping, so small slices are not worth their launches
        // (unless it folds the innermost axis into few outputs: a warp per output, 16 outputs = 16 warps, 14 GB/s)
        const bool few_outputs = c[nc - 1].reduced && c[nc - 1].in_stride == 1 && nout < size_t(16384);
        bool ok = sub_nr > 0 && slice_elems * isz >= (c[0].reduced || few_outputs ? (size_t(1) << 20) : (size_t(32) << 20));
        Launch ls = l;
        ls.slice_of_fold = c[0].reduced;
        // What would serve one slice?  A dry run of the whole planner on it (multi-pass plans included: a window padded
        // on three levels folded over all its axes is 16 slices of an inner-first plan, not one warp of the strided
        // kernel); slices that would themselves end in the strided kernel are not worth peeling.
        char sub[160] = {0};
        size_t sub_need = 0;
        if (ok) {
            Launch ld = ls;
            ld.dry_need = &sub_need; ld.dry_base = 0; ld.dry_plan = sub; ld.dry_plan_bytes = sizeof(sub);
            const cudaError_t ed = plan_dims(red, dtin, dtout, nullptr, nullptr, c + 1, nc - 1, nullptr, ld, false, depth + 1);
            ok = ed == cudaSuccess && strstr(sub, "generic") == nullptr && (l.dry_need != nullptr || sub_need <= l.extra_bytes);
        }
        if (ok) {
            l.note("peel:");
            l.note(sub);
            l.take(sub_need);
            if (l.dry_need) return cudaSuccess;
        }
        for (long long i = 0; ok && i < c[0].shape; ++i) {
            const char* sin = static_cast<const char*>(in) + size_t(i) * size_t(c[0].in_stride) * isz;
            char* sout = static_cast<char*>(out) + (c[0].reduced ? 0 : ptrdiff_t(i) * c[0].out_stride * ptrdiff_t(osz));
            const void* sid = (c[0].reduced && i > 0) ? nullptr : identity;
            const cudaError_t ep = plan_dims(red, dtin, dtout, sin, sout, c + 1, nc - 1, sid, ls, false, depth + 1);
            if (ep != cudaSuccess) return ep;
        }
        if (ok) return cudaSuccess;
    }
    // Everything else: generic strided kernel.
    if (nk > kMaxDims || nr > kMaxDims) return cudaErrorNotSupported;
    l.note("generic");
    if (l.dry_need) return cudaSuccess;
    GenericArgs g;
    memset(&g, 0, sizeof(g));
    g.in = in; g.out = out; g.identity = identity;
    for (int i = 0; i < nc; ++i) {
        if (c[i].reduced) { g.rshape[g.nr] = c[i].shape; g.rin[g.nr] = c[i].in_stride; ++g.nr; }
        else { g.kshape[g.nk] = c[i].shape; g.kin[g.nk] = c[i].in_stride; g.kout[g.nk] = c[i].out_stride; ++g.nk; }
    }
    const bool last_reduced = nc > 0 && c[nc - 1].reduced;
    g.copy_only = copy_only;
    g.warp_per_output = nr > 0 && (last_reduced || nout < (size_t)sm_count() * 256);
    return launch_generic(red, dtin, dtout, g, l);
}

static size_t reduce_nd_workspace_bytes_impl(int dtin, int dtout, int ndim, const int64_t* shape,
                                             const int64_t* in_strides, uint32_t mask, char* plan = nullptr,
                                             size_t plan_bytes = 0) {
    if (plan && plan_bytes) plan[0] = 0;
    Dim c[DACE_B200_MAX_DIMS];
    int nc = 0;
    size_t nout = 0, nred = 0;
    if (!supported_pair(dtin, dtout) ||
        canonicalise(ndim, shape, in_strides, nullptr, mask, c, nc, nout, nred) != cudaSuccess || nout == 0 || nred == 0)
        return kWorkspaceBytes;
    size_t need = 0;
    Launch l;
    l.stream = nullptr;
    l.ws.counters = nullptr; l.ws.partials = nullptr;
    l.dry_need = &need;
    l.dry_plan = plan; l.dry_plan_bytes = plan_bytes;
    plan_dims(RT_Sum, dtin, dtout, nullptr, nullptr, c, nc, nullptr, l, false, 0);   // the plan does not depend on the operator
    return kWorkspaceBytes + need;
}

static cudaError_t reduce_nd_impl(int red, int dtin, int dtout, const void* in, void* out, int ndim,
                                  const int64_t* shape, const int64_t* in_strides, const int64_t* out_strides,
                                  uint32_t mask, const void* identity, void* workspace, size_t workspace_bytes,
                                  cudaStream_t stream) {
    if (red < RT_Min || red > RT_BXor) return cudaErrorNotSupported;
    if (!supported_pair(dtin, dtout)) return cudaErrorNotSupported;
    Dim c[DACE_B200_MAX_DIMS];
    int nc = 0;
    size_t nout = 0, nred = 0;
    cudaError_t e = canonicalise(ndim, shape, in_strides, out_strides, mask, c, nc, nout, nred);
    if (e != cudaSuccess) return e;
    if (nout == 0) return cudaSuccess;
    if (!out || (nred != 0 && !in)) return cudaErrorInvalidValue;  // empty tensors may carry null pointers
    Launch l;
    l.stream = stream;
    e = resolve_workspace(workspace, workspace_bytes, l);
    if (e != cudaSuccess) return e;
    const bool copy_only = (mask & ((ndim >= 32) ? 0xffffffffu : ((1u << ndim) - 1u))) == 0;
    return plan_dims(red, dtin, dtout, in, out, c, nc, identity, l, copy_only, 0);
}

}  // namespace db200

using namespace db200;

extern "C" {

int dace_b200_version(void) { return 2001; }

const char* dace_b200_status_string(int status) { return cudaGetErrorString((cudaError_t)status); }

size_t dace_b200_workspace_bytes(void) { return kWorkspaceBytes; }

int dace_b200_workspace_init(void* workspace, size_t workspace_bytes, dace_b200_stream_t stream) {
    if (!workspace || workspace_bytes < kWorkspaceBytes) return (int)cudaErrorInvalidValue;
    return (int)cudaMemsetAsync(workspace, 0, kCounterBytes, (cudaStream_t)stream);
}

size_t dace_b200_reduce_nd_workspace_bytes(int dtype_in, int dtype_out, int ndim, const int64_t* shape,
                                           const int64_t* in_strides, uint32_t reduce_mask) {
    return reduce_nd_workspace_bytes_impl(dtype_in, dtype_out, ndim, shape, in_strides, reduce_mask);
}

size_t dace_b200_reduce_nd_plan(int dtype_in, int dtype_out, int ndim, const int64_t* shape, const int64_t* in_strides,
                                uint32_t reduce_mask, char* plan, size_t plan_bytes) {
    return reduce_nd_workspace_bytes_impl(dtype_in, dtype_out, ndim, shape, in_strides, reduce_mask, plan, plan_bytes);
}

int dace_b200_reduce_nd(int redtype, int dtype_in, int dtype_out, const void* in, void* out, int ndim,
                        const int64_t* shape, const int64_t* in_strides, const int64_t* out_strides,
                        uint32_t reduce_mask, const void* identity, void* workspace, size_t workspace_bytes,
                        dace_b200_stream_t stream) {
    return (int)reduce_nd_impl(redtype, dtype_in, dtype_out, in, out, ndim, shape, in_strides, out_strides,
                               reduce_mask, identity, workspace, workspace_bytes, (cudaStream_t)stream);
}

int dace_b200_reduce_full(int redtype, int dtype_in, int dtype_out, const void* in, void* out, size_t num_items,
                          const void* identity, void* workspace, size_t workspace_bytes,
                          dace_b200_stream_t stream) {
    int64_t shape[1] = {(int64_t)num_items}, st[1] = {1};
    return (int)reduce_nd_impl(redtype, dtype_in, dtype_out, in, out, 1, shape, st, nullptr, 1u, identity,
                               workspace, workspace_bytes, (cudaStream_t)stream);
}

int dace_b200_reduce_rows(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                          size_t num_segments, size_t segment_size, size_t segment_stride, const void* identity,
                          void* workspace, size_t workspace_bytes, dace_b200_stream_t stream) {
    int64_t shape[2] = {(int64_t)num_segments, (int64_t)segment_size};
    int64_t st[2] = {(int64_t)segment_stride, 1};
    return (int)reduce_nd_impl(redtype, dtype_in, dtype_out, in, out, 2, shape, st, nullptr, 2u, identity,
                               workspace, workspace_bytes, (cudaStream_t)stream);
}

int dace_b200_reduce_rows_noalloc(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                                  size_t num_segments, size_t segment_size, size_t segment_stride,
                                  const void* identity, dace_b200_stream_t stream) {
    return (int)reduce_rows_noalloc(redtype, dtype_in, dtype_out, in, out, num_segments, segment_size, segment_stride,
                                    identity, (cudaStream_t)stream);
}

int dace_b200_reduce_cols(int redtype, int dtype_in, int dtype_out, const void* in, void* out, size_t outer,
                          size_t reduced, size_t inner, const void* identity, void* workspace,
                          size_t workspace_bytes, dace_b200_stream_t stream) {
    int64_t shape[3] = {(int64_t)outer, (int64_t)reduced, (int64_t)inner};
    int64_t st[3] = {(int64_t)(reduced * inner), (int64_t)inner, 1};
    return (int)reduce_nd_impl(redtype, dtype_in, dtype_out, in, out, 3, shape, st, nullptr, 2u, identity,
                               workspace, workspace_bytes, (cudaStream_t)stream);
}

int dace_b200_reduce_rows_max_sum(int dtype_in, const void* in, float* out_max, float* out_sum,
                                  size_t num_segments, size_t segment_size, size_t segment_stride,
                                  dace_b200_stream_t stream) {
    if (!in || !out_max || !out_sum) return (int)cudaErrorInvalidValue;
    if (num_segments == 0) return (int)cudaSuccess;
    Launch l;
    l.stream = (cudaStream_t)stream;
    l.ws.counters = nullptr; l.ws.partials = nullptr;
    RowsMaxSumArgs a{in, out_max, out_sum, num_segments, segment_size, segment_stride};
    return (int)launch_rows_max_sum(dtype_in, a, l);
}

int dace_b200_reduce_arg(int redtype, int dtype_in, int dtype_idx, const void* in, void* out_idx, void* out_val,
                         size_t outer, size_t reduced, size_t inner, int64_t in_stride_outer,
                         int64_t in_stride_reduced, int64_t in_stride_inner, size_t out_idx_stride_bytes,
                         size_t out_val_stride_bytes, const void* init, void* workspace, size_t workspace_bytes,
                         dace_b200_stream_t stream) {
    if (dtype_idx != DT_I32 && dtype_idx != DT_I64) return (int)cudaErrorNotSupported;
    if (outer == 0 || inner == 0) return (int)cudaSuccess;
    if (!out_idx || (reduced != 0 && !in) || (!init && !out_val)) return (int)cudaErrorInvalidValue;
    if (in_stride_outer < 0 || in_stride_reduced < 0 || in_stride_inner < 0) return (int)cudaErrorNotSupported;
    Launch l;
    l.stream = (cudaStream_t)stream;
    cudaError_t e = resolve_workspace(workspace, workspace_bytes, l);
    if (e != cudaSuccess) return (int)e;
    ArgArgs a{in, out_idx, out_val, outer, reduced, inner, in_stride_outer, in_stride_reduced, in_stride_inner,
              (long long)out_idx_stride_bytes, (long long)out_val_stride_bytes, dtype_idx == DT_I64 ? 8 : 4, init};
    return (int)launch_arg(redtype, dtype_in, a, l);
}

size_t dace_b200_wcr_scatter_workspace_bytes(int redtype, int dtype_idx, int dtype_val, int has_values, size_t n,
                                             size_t nbins) {
    if (redtype != RT_Sum || tuning().no_partition) return 0;
    return scatter_partition_bytes(dtype_idx, dtype_val, has_values != 0, n, nbins);
}

int dace_b200_wcr_scatter(int redtype, int dtype_idx, int dtype_val, const void* idx, const void* values,
                          const void* const_value, size_t n, void* bins, size_t nbins, void* workspace,
                          size_t workspace_bytes, dace_b200_stream_t stream) {
    if (n > 0 && (!idx || !bins)) return (int)cudaErrorInvalidValue;
    if (nbins == 0 || n == 0) return (int)cudaSuccess;
    Launch l;
    l.stream = (cudaStream_t)stream;
    l.ws.counters = nullptr; l.ws.partials = nullptr;
    l.extra = workspace;                               // pure scratch for this entry point (may be NULL)
    l.extra_bytes = workspace ? workspace_bytes : 0;
    ScatterArgs a{idx, values, const_value, n, bins, nbins};
    return (int)launch_scatter(redtype, dtype_idx, dtype_val, a, l);
}

// ---- multi-GPU: exchange buffers in peer memory + one-shot all-reduce (peer_combine.cu) ----

size_t dace_b200_peer_buffer_bytes(int world, size_t cap_bytes) { return peer_buffer_bytes(world, cap_bytes); }

int dace_b200_peer_alloc(size_t bytes, void** ptr) {
    if (!ptr || bytes == 0) return (int)cudaErrorInvalidValue;
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e != cudaSuccess) return (int)e;
    e = cudaMemset(*ptr, 0, bytes);   // flags start at epoch 0
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    return (int)e;
}

int dace_b200_peer_free(void* ptr) { return (int)cudaFree(ptr); }

int dace_b200_peer_export(void* ptr, unsigned char* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    if (!ptr || !handle64) return (int)cudaErrorInvalidValue;
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
    if (e == cudaSuccess) memcpy(handle64, &h, 64);
    return (int)e;
}

int dace_b200_peer_open(const unsigned char* handle64, void** ptr) {
    if (!ptr || !handle64) return (int)cudaErrorInvalidValue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}

int dace_b200_peer_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }

int dace_b200_peer_allreduce(int redtype, int dtype, void* data, size_t nelem, void* const* peer_bases, int rank,
                             int world, unsigned long long epoch, size_t cap_bytes, dace_b200_stream_t stream) {
    if (nelem > 0 && (!data || !peer_bases)) return (int)cudaErrorInvalidValue;
    return (int)launch_peer_allreduce(redtype, dtype, data, nelem, peer_bases, rank, world, epoch, cap_bytes,
                                      (cudaStream_t)stream);
}

}  // extern "C"
