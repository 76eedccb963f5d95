/* dace_b200_reduce.h - C ABI of the B200 (sm_100a) reduction backend for DaCe.
 *
 * This is the drop-in boundary: the functions below are what a DaCe-generated
 * program (host C++ tasklet inside __program_<sdfg>, reference
 * dace/codegen/compiled_sdfg.py:549-592) calls instead of the per-node helper the
 * reference's `CUDA (device)` expansion generates,
 *
 *     cudaError_t __dace_reduce_<sdfg>_<state>_<node>(T_in *input, T_out *output,
 *                                                     size_t num_items | size_t num_segments,
 *                                                     size_t segment_size, cudaStream_t stream);
 *                     (reference dace/libraries/standard/nodes/reduce.py:574-588)
 *
 * and instead of the generated GPU map kernels that resolve write conflicts
 * with one global atomic per element (reference dace/codegen/targets/cpu.py:965-1013).
 *
 * Conventions (same as the reference's, reference dace/runtime/include/dace/cuda/cudacommon.cuh:28-83):
 *   - every entry point returns a cudaError_t value as int (0 = cudaSuccess), so
 *     the call site can stay wrapped in DACE_GPU_CHECK(...);
 *   - work is enqueued on `stream`; the compute entry points never synchronise the stream or the
 *     device and never allocate or free memory (no cudaMalloc / cudaFree / cudaEventSynchronize
 *     and no getenv on any call path: they can be captured into a CUDA graph); every byte of
 *     scratch they use is the caller's `workspace`;
 *   - `in`, `out`, `bins`... are device (or pinned, device-mapped) pointers owned
 *     by the caller, already offset to the memlet start; nothing is retained;
 *   - unsupported (op, dtype) combinations return cudaErrorNotSupported (801),
 *     malformed arguments cudaErrorInvalidValue (1); nothing throws or exits;
 *   - there is no CPU fallback anywhere behind this interface.
 *
 * Plain C, no torch / DaCe types in any signature.
 */
#ifndef DACE_B200_REDUCE_H
#define DACE_B200_REDUCE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define DACE_B200_API __declspec(dllexport)
#else
#define DACE_B200_API __attribute__((visibility("default")))
#endif

/* Same handle type as cudaStream_t (struct CUstream_st *). */
typedef struct CUstream_st* dace_b200_stream_t;

/* Reduction operator; numeric values mirror dace::ReductionType
 * (reference dace/runtime/include/dace/types.h:301-317, dace/dtypes.py:93-113). */
enum dace_b200_redtype {
    DACE_B200_CUSTOM = 0, /* not supported by the precompiled library */
    DACE_B200_MIN = 1,
    DACE_B200_MAX = 2,
    DACE_B200_SUM = 3,
    DACE_B200_PRODUCT = 4,
    DACE_B200_LOGICAL_AND = 5,
    DACE_B200_BITWISE_AND = 6,
    DACE_B200_LOGICAL_OR = 7,
    DACE_B200_BITWISE_OR = 8,
    DACE_B200_LOGICAL_XOR = 9,
    DACE_B200_BITWISE_XOR = 10,
    DACE_B200_MIN_LOCATION = 11, /* dace_b200_reduce_arg only */
    DACE_B200_MAX_LOCATION = 12  /* dace_b200_reduce_arg only */
};

/* Element types. Supported (in -> out) pairs: T -> T for every type below, plus
 * bf16 -> f32 and f16 -> f32 (fp32 accumulation in all four 16-bit cases). */
enum dace_b200_dtype {
    DACE_B200_F32 = 0,
    DACE_B200_F64 = 1,
    DACE_B200_I32 = 2,
    DACE_B200_I64 = 3,
    DACE_B200_U32 = 4,
    DACE_B200_U64 = 5,
    DACE_B200_BF16 = 6,
    DACE_B200_F16 = 7,
    DACE_B200_I8 = 8,
    DACE_B200_U8 = 9,
    DACE_B200_I16 = 10,
    DACE_B200_U16 = 11
};

#define DACE_B200_MAX_DIMS 12

/* Library/ABI version (major * 1000 + minor). */
DACE_B200_API int dace_b200_version(void);

/* Human-readable name of a returned status (cudaGetErrorString). */
DACE_B200_API const char* dace_b200_status_string(int status);

/* ---- Workspace ---------------------------------------------------------
 * Cross-CTA combining uses a device workspace owned by the CALLER: ticket counters + partials
 * (dace_b200_workspace_bytes() = 8 MiB, the "base"), optionally followed by more scratch for the
 * plans that need it.  Mirrors the reference's CUB temp-storage protocol: query the size,
 * cudaMalloc once in __dace_init_cuda, cudaFree in __dace_exit_cuda
 * (reference dace/libraries/standard/nodes/reduce.py:550-570).  The base must be zero-initialised
 * once with dace_b200_workspace_init(); kernels leave it clean.  One workspace serves one stream at
 * a time: calls that may overlap on the GPU (independent nodes on different streams) need separate
 * workspaces - the `B200` expansions allocate one per expanded node.  The library never allocates
 * a workspace itself: workspace == NULL or workspace_bytes < the base is cudaErrorInvalidValue.
 *
 * dace_b200_reduce_nd_workspace_bytes(): bytes a workspace should have for dace_b200_reduce_nd to
 * run this shape at full speed: the base, plus the intermediate of the two-pass plans ([R1, K, R2]
 * and [O, R1, K, R2] forms).  With a smaller (base-only) workspace those shapes still work, through
 * the slower scratch-free kernels.  Pure host arithmetic, no CUDA calls besides a cached device query. */
DACE_B200_API size_t dace_b200_workspace_bytes(void);
DACE_B200_API int dace_b200_workspace_init(void* workspace, size_t workspace_bytes, dace_b200_stream_t stream);
DACE_B200_API size_t dace_b200_reduce_nd_workspace_bytes(int dtype_in, int dtype_out, int ndim, const int64_t* shape,
                                                         const int64_t* in_strides, uint32_t reduce_mask);
/* Diagnostics (no reference counterpart): the same dry run, which also writes WHICH plan the planner picks for the
 * layout into `plan` as a NUL-terminated, space-separated trace ("rows", "cols", "split", "rows2" = rows on two
 * levels, "two-pass(rows,cols)", "outer-first(cols,rows)", "inner-first: <first pass> <rest>", "peel: <slice plan>",
 * "generic" = the strided fallback kernel).  Needs no GPU; returns what dace_b200_reduce_nd_workspace_bytes returns.
 * tests/test_cabi_library.py uses it to keep layouts off the fallback kernel. */
DACE_B200_API size_t dace_b200_reduce_nd_plan(int dtype_in, int dtype_out, int ndim, const int64_t* shape,
                                              const int64_t* in_strides, uint32_t reduce_mask, char* plan,
                                              size_t plan_bytes);

/* ---- Reduce library node -------------------------------------------------
 * General N-D entry, the call the `B200` expansion of
 * dace.libraries.standard.Reduce emits.  Semantics of the node
 * (reference dace/libraries/standard/nodes/reduce.py:1346-1411):
 *     out[kept...] = fold(op, identity, in[kept..., reduced...])
 * `shape`/`in_strides` describe the input memlet subset (ndim entries, strides
 * in elements, any order/steps).  Bit i of `reduce_mask` marks axis i as reduced
 * (all axes: (1<<ndim)-1).  `out_strides` has one entry per KEPT axis in axis
 * order (NULL = C-contiguous).  `identity` is a HOST pointer to one value of
 * dtype_out, or NULL meaning "fold into the value already stored in out"
 * (reference dace/frontend/operations.py:251).  With reduce_mask == 0 (the
 * expansion squeezed every reduced axis away at compile time) the call is a
 * strided copy, identity ignored (reference reduce.py:77-87); a reduced axis of
 * run-time extent 1 folds normally.
 * Replaces: reduce.py:574-588 (CUB), :984-1332 (GPUAuto kernels). */
DACE_B200_API int dace_b200_reduce_nd(int redtype, int dtype_in, int dtype_out, const void* in, void* out, int ndim,
                                      const int64_t* shape, const int64_t* in_strides, const int64_t* out_strides,
                                      uint32_t reduce_mask, const void* identity, void* workspace,
                                      size_t workspace_bytes, dace_b200_stream_t stream);

/* All axes reduced, contiguous input: replaces
 * cub::DeviceReduce::Reduce(storage, ssize, in, out, num_items, op, identity, stream)
 * (reference reduce.py:520-525). */
DACE_B200_API int dace_b200_reduce_full(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                                        size_t num_items, const void* identity, void* workspace,
                                        size_t workspace_bytes, dace_b200_stream_t stream);

/* Trailing axes reduced: num_segments contiguous segments of segment_size
 * elements, segment s starting at in + s*segment_stride (pass segment_size for a
 * dense array).  Replaces cub::DeviceSegmentedReduce::Reduce with
 * dace::stridedIterator offsets (reference reduce.py:526-544). */
DACE_B200_API int dace_b200_reduce_rows(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                                        size_t num_segments, size_t segment_size, size_t segment_stride,
                                        const void* identity, void* workspace, size_t workspace_bytes,
                                        dace_b200_stream_t stream);

/* Allocation-free variant of dace_b200_reduce_rows for few, long segments (num_segments = 1 is a
 * full reduction).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  Same semantics and kernel as dace_b200_reduce_rows, but the cross-CTA
 * tickets and partials live in a file-scope `static __device__` array of 66,560 bytes that every GPU
 * loading the library holds until exit; the call path contains no cudaMalloc / cudaFree / allocating
 * library routine (one cudaMemsetAsync on `stream` clears the tickets).  Limits: num_segments <= 256
 * (more returns cudaErrorInvalidValue and prints the reason to stderr); any segment_size.  Every
 * call on a device shares the one array: calls that may overlap on the GPU (different streams
 * without ordering between them) race - serialise them or use dace_b200_reduce_rows with separate
 * workspaces. */
DACE_B200_API int dace_b200_reduce_rows_noalloc(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                                                size_t num_segments, size_t segment_size, size_t segment_stride,
                                                const void* identity, dace_b200_stream_t stream);

/* Middle/leading axis reduced: in[outer][reduced][inner] (dense, inner
 * contiguous) -> out[outer][inner].  Replaces the GPUAuto strided schedule
 * (reference reduce.py:1141-1332); the CUB expansion has no equivalent
 * (reference reduce.py:493-496). */
DACE_B200_API int dace_b200_reduce_cols(int redtype, int dtype_in, int dtype_out, const void* in, void* out,
                                        size_t outer, size_t reduced, size_t inner, const void* identity,
                                        void* workspace, size_t workspace_bytes, dace_b200_stream_t stream);

/* Fused row max + row sum in a single pass over `in` (two Reduce nodes sharing
 * an input, as in reference tests/softmax_test.py:8-31).  dtype_in: f32, bf16
 * or f16; outputs fp32; identities -inf / 0. */
DACE_B200_API int dace_b200_reduce_rows_max_sum(int dtype_in, const void* in, float* out_max, float* out_sum,
                                                size_t num_segments, size_t segment_size, size_t segment_stride,
                                                dace_b200_stream_t stream);

/* ---- argmin / argmax (ReductionType Min_Location / Max_Location) --------------
 * in[o][r][i] with ELEMENT strides (in_stride_outer, in_stride_reduced, in_stride_inner);
 * for every (o, i) the index along r of the first minimum / maximum and, optionally, its
 * value are written to output position o * inner + i:
 *     *(idx_t*)((char*)out_idx + pos * out_idx_stride_bytes),  idx_t = dtype_idx (i32 or i64)
 *     *(T*)    ((char*)out_val + pos * out_val_stride_bytes),  T = dtype_in
 * (stride 0 = densely packed), so the two outputs may be the fields of the interleaved
 * struct array `_val_and_idx {idx; val}` the reference frontend reduces into
 * (reference dace/frontend/python/replacements/reduction.py:299-334).  Semantics of that
 * map's WCR lambda `val = max(x.val, y.val), idx = y.idx if x.val < y.val else x.idx`
 * evaluated in index order: strictly-better replaces, first occurrence wins.  `init` is a
 * HOST pointer to the initial value (dtype_in); the running pair starts as (*init, -1), and
 * elements equal to *init never replace it (the frontend's init state writes
 * min_value(dtype) / max_value(dtype), reduction.py:303-304).  init == NULL folds into the
 * (val, idx) pairs already stored in the outputs (out_val required then).
 * Replaces: the generated WCR-map kernel with dace::wcr_custom<T>::reduce_atomic CAS loops
 * (reference dace/runtime/include/dace/reduction.h:56-75) and cub::DeviceReduce::ArgMin/ArgMax
 * of the CUB expansion (reference dace/libraries/standard/nodes/reduce.py:399-402). */
DACE_B200_API int dace_b200_reduce_arg(int redtype, int dtype_in, int dtype_idx, const void* in, void* out_idx,
                                       void* out_val, size_t outer, size_t reduced, size_t inner,
                                       int64_t in_stride_outer, int64_t in_stride_reduced, int64_t in_stride_inner,
                                       size_t out_idx_stride_bytes, size_t out_val_stride_bytes, const void* init,
                                       void* workspace, size_t workspace_bytes, dace_b200_stream_t stream);

/* ---- WCR map reductions (scatter-add / histogram) ---------------------------
 *     bins[idx[i]] = op(bins[idx[i]], values ? values[i] : *const_value)   i in [0, n)
 * Replaces the generated one-thread-per-element kernel whose body is
 * dace::wcr_fixed<op, T>::reduce_atomic(bins + k, v)
 * (reference dace/codegen/targets/cpu.py:965-1013,
 *  reference dace/runtime/include/dace/reduction.h:167-340).  `bins` is updated in
 * place (WCR semantics: it is the caller's job to initialise it).  dtype_idx:
 * i32, u32, i64, u8; dtype_val (= bins type): i32, u32, i64, u64, f32, f64.
 * `values` may be NULL, then `const_value` (HOST pointer, dtype_val; NULL = 1) is
 * applied for every index.  Indices outside [0, nbins) are skipped.
 * `workspace` is pure scratch for this call (no initialisation needed, 16-byte aligned):
 * dace_b200_wcr_scatter_workspace_bytes() says how much the radix-partition path wants for the given
 * problem (0 = the path does not apply: small bin arrays, per-element values, 64-bit bins, < 2^22
 * keys).  workspace == NULL or too small selects the global-atomic kernel instead (correct, slower
 * for large bin arrays). */
DACE_B200_API size_t dace_b200_wcr_scatter_workspace_bytes(int redtype, int dtype_idx, int dtype_val, int has_values,
                                                           size_t n, size_t nbins);
DACE_B200_API int dace_b200_wcr_scatter(int redtype, int dtype_idx, int dtype_val, const void* idx,
                                        const void* values, const void* const_value, size_t n, void* bins,
                                        size_t nbins, void* workspace, size_t workspace_bytes,
                                        dace_b200_stream_t stream);

/* ---- Multi-GPU combine of small partials over NVLink peer memory --------------
 * When the sharded axis is the reduced one, every GPU's local call above leaves a partial that
 * must be all-reduced across the GPUs of the box.  The reference's analogue is the MPI Allreduce
 * library node issued after the local computation (reference
 * dace/libraries/mpi/nodes/allreduce.py:9-60, dace/libraries/mpi/environments/mpi.py:6-22).
 * Each rank allocates ONE exchange buffer (`_peer_alloc`, size `_peer_buffer_bytes(world, cap)`),
 * exports a 64-byte CUDA IPC handle (`_peer_export`) that the host side ships to the other ranks
 * (any transport; dace_b200/sharding.py uses torch.distributed), and maps the peers' buffers
 * (`_peer_open`).  `_peer_allreduce` then combines `data[0:nelem]` (f32/f64/i32/u32/i64/u64;
 * nelem * sizeof <= cap_bytes) in place on every rank: push to all peers, release/acquire flags,
 * fold in rank order (bit-identical on all ranks).  `epoch` is a host counter starting at 1 and
 * incremented per call, identical on all ranks.  `peer_bases[r]` is THIS process' mapping of rank
 * r's buffer (its own allocation for r == rank).  Enqueued on `stream`; never synchronises. */
DACE_B200_API size_t dace_b200_peer_buffer_bytes(int world, size_t cap_bytes);
DACE_B200_API int dace_b200_peer_alloc(size_t bytes, void** ptr);
DACE_B200_API int dace_b200_peer_free(void* ptr);
DACE_B200_API int dace_b200_peer_export(void* ptr, unsigned char* handle64);
DACE_B200_API int dace_b200_peer_open(const unsigned char* handle64, void** ptr);
DACE_B200_API int dace_b200_peer_close(void* ptr);
DACE_B200_API int dace_b200_peer_allreduce(int redtype, int dtype, void* data, size_t nelem, void* const* peer_bases,
                                           int rank, int world, unsigned long long epoch, size_t cap_bytes,
                                           dace_b200_stream_t stream);

/* ---- Communicator for generated programs (one process per GPU) ----------------------------
 * What the `B200` expansion of the AllReduceCombine library node (dace_b200/comm_node.py) calls
 * right after a local reduction whose sharded axis is the reduced one.  Plays the role of the
 * reference's MPI environment + Allreduce node (reference dace/libraries/mpi/environments/mpi.py:6-22,
 * dace/libraries/mpi/nodes/allreduce.py:9-60): NCCL over NVLink / NVSwitch, bound at run time
 * (dlopen of libnccl.so.2; DACE_B200_NCCL_LIB overrides), plus the one-shot peer-memory kernel
 * above for payloads <= 64 KiB when CUDA IPC works between the ranks (DACE_B200_NO_PEER=1 turns
 * it off).  The IPC handles travel through NCCL itself (all-gather), so no other transport is
 * needed.
 *   _unique_id / _create : explicit rendezvous, `id128` (128 bytes, made on rank 0) shipped by the
 *                          host over any transport; world == 1 needs neither an id nor NCCL.
 *   _create_from_env     : RANK / WORLD_SIZE from the environment (torchrun, mpirun wrappers);
 *                          the id travels through the file DACE_B200_RENDEZVOUS (default
 *                          /tmp/dace_b200_nccl_<MASTER_ADDR>_<MASTER_PORT>.id).  The CUDA device
 *                          must already be selected (cudaSetDevice / CUDA_VISIBLE_DEVICES).
 *   _default             : process-wide communicator: the one installed with _set_default, else
 *                          created from the environment on first use (generated init code).
 *   _allreduce           : in place on `data[0:nelem]`, enqueued on `stream`, never synchronises.
 *                          Sum / Product / Min / Max for every dtype of this header that NCCL
 *                          has; Logical_And/Or as Min/Max of the 0/1 partials; bitwise operators
 *                          and Logical_Xor only on the peer-memory path (32/64-bit integers,
 *                          <= 64 KiB), else cudaErrorNotSupported.  Results of the peer path are
 *                          folded in rank order: bit-identical on every rank.
 * All calls are collective over the ranks of the communicator. */
DACE_B200_API int dace_b200_comm_unique_id(unsigned char* id128);
DACE_B200_API int dace_b200_comm_create(int rank, int world, const unsigned char* id128, void** comm);
DACE_B200_API int dace_b200_comm_create_from_env(void** comm);
DACE_B200_API int dace_b200_comm_destroy(void* comm);
DACE_B200_API int dace_b200_comm_rank(void* comm);
DACE_B200_API int dace_b200_comm_world(void* comm);
DACE_B200_API int dace_b200_comm_uses_peer_memory(void* comm);
DACE_B200_API int dace_b200_comm_set_default(void* comm);
DACE_B200_API int dace_b200_comm_default(void** comm);
DACE_B200_API int dace_b200_comm_allreduce(void* comm, int redtype, int dtype, void* data, size_t nelem,
                                           dace_b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DACE_B200_REDUCE_H */
