// Internal launch interfaces between the planner (api.cu) and the kernel files.
#pragma once
#include "common.cuh"

namespace db200 {

// out[row] = fold(in[row*row_stride : +n]) with ctas_per_row CTAs per row.
struct SplitArgs {
    const void* in; void* out;
    size_t n; size_t rows; size_t row_stride; ptrdiff_t out_stride;
    unsigned ctas_per_row; unsigned cluster;
    const void* identity;
    int big;  // 1: 1024-thread CTAs, 8 loads in flight (large inputs); 0: 512 threads, 4 loads
};
cudaError_t launch_split(int red, int dtin, int dtout, const SplitArgs& a, const Launch& l);

// reduce_full.cu: the split kernel with tickets / partials in a file-scope __device__ array (no allocation anywhere
// on the call path); rows <= 256.
cudaError_t reduce_rows_noalloc(int red, int dtin, int dtout, const void* in, void* out, size_t rows, size_t n,
                                size_t row_stride, const void* identity, cudaStream_t stream);

// Segmented (row) reduction, one sub-warp / warp / CTA per row.
struct RowsArgs {
    const void* in; void* out;
    size_t rows; size_t n; size_t row_stride; ptrdiff_t out_stride;
    const void* identity;
    // A second, outer level of rows (windows of a larger array: [A, B, n] with a pitch after every B rows): the
    // kernels run `hi_count` slices of `rows` rows along gridDim.y, slice h starting at in + h * hi_in_stride and
    // writing from out + h * hi_out_stride (element strides).  1 / 0 / 0 = one level.
    size_t hi_count = 1; size_t hi_in_stride = 0; ptrdiff_t hi_out_stride = 0;
};
cudaError_t launch_rows(int red, int dtin, int dtout, const RowsArgs& a, const Launch& l);

// Fused row max + row sum in one pass (fp32 accumulate/outputs).
struct RowsMaxSumArgs {
    const void* in; float* out_max; float* out_sum;
    size_t rows; size_t n; size_t row_stride;
};
cudaError_t launch_rows_max_sum(int dtin, const RowsMaxSumArgs& a, const Launch& l);

// Column reduction: in[o][r][i] (i contiguous) -> out[o][i].
struct ColsArgs {
    const void* in; void* out;
    size_t outer, reduced, inner;
    size_t in_stride_o, in_stride_r;       // elements
    ptrdiff_t out_stride_o, out_stride_i;  // elements
    const void* identity;
};
cudaError_t launch_cols(int red, int dtin, int dtout, const ColsArgs& a, const Launch& l);

// Arbitrary strided N-D reduction (collapsed dims, <= kMaxDims each side).
constexpr int kMaxDims = 6;
struct GenericArgs {
    const void* in; void* out;
    int nk, nr;
    long long kshape[kMaxDims], kin[kMaxDims], kout[kMaxDims];
    long long rshape[kMaxDims], rin[kMaxDims];
    const void* identity;
    int warp_per_output;
    int copy_only;  // no reduced axis was requested at all: plain strided copy
};
cudaError_t launch_generic(int red, int dtin, int dtout, const GenericArgs& a, const Launch& l);

// out[j] = identity for all outputs (empty reduction range).
cudaError_t launch_fill(int dtout, void* out, int nk, const long long* kshape, const long long* kout,
                        const void* identity, const Launch& l);

// argmin / argmax: in[o][r][i] (element strides) -> (index along r, value) at output o * inner + i.
struct ArgArgs {
    const void* in; void* out_idx; void* out_val;   // out_val may be null when init is given
    size_t outer, reduced, inner;
    long long in_stride_o, in_stride_r, in_stride_i;              // elements
    long long out_idx_stride_bytes, out_val_stride_bytes;         // 0 = densely packed
    int idx_bytes;                                                // 4 (int32) or 8 (int64)
    const void* init;   // host pointer to the initial value (input dtype) or null = fold into the outputs
};
cudaError_t launch_arg(int red, int dtin, const ArgArgs& a, const Launch& l);

// WCR scatter: bins[idx[i]] = op(bins[idx[i]], values ? values[i] : *const_value).
struct ScatterArgs {
    const void* idx; const void* values; const void* const_value;
    size_t n; void* bins; size_t nbins;
    // Device-side path selection: when `gate` is set, a kernel runs only if *gate == gate_run_if, or for a
    // negative gate_run_if only if *gate != -gate_run_if (written by the key sampler of wcr_partition.cu
    // earlier on the same stream).
    const int* gate = nullptr; int gate_run_if = 0;
};
// reduce_generic.cu: fp32 vector -> bf16 / f16 output with the identity fold (end of a two-pass plan).
cudaError_t launch_finish_convert(int red, int dtout, const float* tmp, void* out, size_t n, ptrdiff_t out_stride,
                                  const void* identity, const Launch& l);
cudaError_t launch_scatter(int red, int dtidx, int dtval, const ScatterArgs& a, const Launch& l);
// Partition + shared-memory histogram path for large bin arrays (wcr_partition.cu);
// cudaErrorNotSupported = does not apply, use the global-atomic kernel.
// peer_combine.cu: one-shot all-reduce of small partials through IPC-mapped peer buffers.
cudaError_t launch_peer_allreduce(int red, int dtype, void* data, size_t nelem, void* const* bases, int rank, int world,
                                  unsigned long long epoch, size_t cap_bytes, cudaStream_t stream);
size_t peer_buffer_bytes(int world, size_t cap_bytes);

// On success *gate is the device flag the partitioned kernel was gated on (0 = it ran; 1 = the sampler
// found clustered keys and it returned at once): the caller then launches the global-atomic kernel
// gated on 1.  Scratch = l.extra (the caller's workspace, scatter_partition_bytes() large).
cudaError_t launch_scatter_partitioned(int dtidx, int dtval, const ScatterArgs& a, const Launch& l, const int** gate);
size_t scatter_partition_bytes(int dtidx, int dtval, bool has_values, size_t n, size_t nbins);
// Scatters with per-element values: only the key sampler runs (flag in l.extra); the caller gates the global-atomic
// kernel on "flag != 2" (gate_run_if = -2) and the hot-key cache kernel on 2.
cudaError_t launch_scatter_sample(int dtidx, int dtval, const ScatterArgs& a, const Launch& l, const int** gate);

}  // namespace db200
