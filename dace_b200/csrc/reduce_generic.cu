// Arbitrary strided N-D reduction for sm_100a: whatever the planner cannot
// canonicalise to full / rows / cols (interleaved reduced and kept axes such as
// [R1, K, R2], stepped memlets, transposed views).  The reference covers these
// with the `pure` nested-map expansion (one WCR per element,
// reference dace/libraries/standard/nodes/reduce.py:27-154); CUB refuses them
// (reference dace/libraries/standard/nodes/reduce.py:493-496).
//
// Two mappings, picked by which axis is contiguous in memory:
//   * warp-per-output: lanes stride the flattened reduced index (innermost
//     reduced axis has stride 1 -> coalesced), xor-shuffle at the end;
//   * thread-per-output: adjacent threads own adjacent outputs (innermost kept
//     axis has stride 1 -> coalesced across the warp), sequential fold.
#include "common.cuh"
#include "kernels.h"

namespace db200 {

struct GenericDims {
    int nk, nr;
    long long kshape[kMaxDims], kin[kMaxDims], kout[kMaxDims];
    long long rshape[kMaxDims], rin[kMaxDims];
};

__device__ __forceinline__ void kept_offsets(const GenericDims& d, long long j, long long& in_off, long long& out_off) {
    in_off = 0; out_off = 0;
    for (int k = d.nk - 1; k >= 0; --k) {
        long long q = j / d.kshape[k];
        long long idx = j - q * d.kshape[k];
        in_off += idx * d.kin[k];
        out_off += idx * d.kout[k];
        j = q;
    }
}

__device__ __forceinline__ long long reduced_offset(const GenericDims& d, long long i) {
    long long off = 0;
    for (int k = d.nr - 1; k >= 0; --k) {
        long long q = i / d.rshape[k];
        off += (i - q * d.rshape[k]) * d.rin[k];
        i = q;
    }
    return off;
}

template <class L, class A, class S, class Op, bool WARP>
__global__ void __launch_bounds__(256)
generic_kernel(const L* __restrict__ in, S* __restrict__ out, GenericDims d, long long nout, long long nred,
               Finish<A> fin) {
    if constexpr (WARP) {
        const unsigned lane = threadIdx.x & 31;
        const long long warp = ((long long)(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
        const long long nwarps = ((long long)(gridDim.x) * blockDim.x) >> 5;
        const long long inner = d.rshape[d.nr - 1];
        const long long inner_stride = d.rin[d.nr - 1];
        for (long long j = warp; j < nout; j += nwarps) {
            long long in_off, out_off;
            kept_offsets(d, j, in_off, out_off);
            A v = Op::neutral();
            if (d.nr == 1) {
                for (long long i = lane; i < nred; i += 32)
                    v = Op::apply(v, to_acc<A, L>(ldg_scalar(in + in_off + i * inner_stride)));
            } else {
                // Walk the innermost reduced axis with the lanes; outer reduced axes sequentially.
                const long long nouter = nred / inner;
                for (long long ob = 0; ob < nouter; ++ob) {
                    long long off = 0, t = ob;
                    for (int k = d.nr - 2; k >= 0; --k) {
                        long long q = t / d.rshape[k];
                        off += (t - q * d.rshape[k]) * d.rin[k];
                        t = q;
                    }
                    const L* p = in + in_off + off;
                    for (long long i = lane; i < inner; i += 32)
                        v = Op::apply(v, to_acc<A, L>(ldg_scalar(p + i * inner_stride)));
                }
            }
            v = warp_reduce<Op, A>(v);
            if (lane == 0) write_result<Op, A, S>(out + out_off, v, fin);
        }
    } else {
        const long long tid = (long long)(blockIdx.x) * blockDim.x + threadIdx.x;
        const long long nthr = (long long)(gridDim.x) * blockDim.x;
        for (long long j = tid; j < nout; j += nthr) {
            long long in_off, out_off;
            kept_offsets(d, j, in_off, out_off);
            A v = Op::neutral();
            if (d.nr == 1) {
                const long long st = d.rin[0];
                for (long long i = 0; i < nred; ++i) v = Op::apply(v, to_acc<A, L>(ldg_scalar(in + in_off + i * st)));
            } else {
                for (long long i = 0; i < nred; ++i)
                    v = Op::apply(v, to_acc<A, L>(ldg_scalar(in + in_off + reduced_offset(d, i))));
            }
            write_result<Op, A, S>(out + out_off, v, fin);
        }
    }
}

template <class T, class Op>
static cudaError_t launch_generic_t(const GenericArgs& a, const Launch& l) {
    using L = typename T::L; using A = typename T::A; using S = typename T::S;
    GenericDims d;
    d.nk = a.nk; d.nr = a.nr;
    long long nout = 1, nred = 1;
    for (int k = 0; k < kMaxDims; ++k) {
        d.kshape[k] = k < a.nk ? a.kshape[k] : 1; d.kin[k] = k < a.nk ? a.kin[k] : 0; d.kout[k] = k < a.nk ? a.kout[k] : 0;
        d.rshape[k] = k < a.nr ? a.rshape[k] : 1; d.rin[k] = k < a.nr ? a.rin[k] : 0;
        nout *= d.kshape[k]; nred *= d.rshape[k];
    }
    Finish<A> fin = make_finish<A, S>(a.identity);
    if (d.nr == 0) { d.nr = 1; d.rshape[0] = 1; d.rin[0] = 0; }
    // Empty axis list: the reference copies input to output and ignores the
    // identity (reference dace/libraries/standard/nodes/reduce.py:77-87).
    if (a.copy_only) fin.has_identity = 2;
    if (a.warp_per_output) {
        const long long cap = (long long)sm_count() * resident_ctas<generic_kernel<L, A, S, Op, true>>(256);
        long long grid = (nout + 7) / 8;
        if (grid > cap) grid = cap;
        if (grid < 1) grid = 1;
        generic_kernel<L, A, S, Op, true><<<(unsigned)grid, 256, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, d, nout, nred, fin);
    } else {
        const long long cap = (long long)sm_count() * resident_ctas<generic_kernel<L, A, S, Op, false>>(256);
        long long grid = (nout + 255) / 256;
        if (grid > cap) grid = cap;
        if (grid < 1) grid = 1;
        generic_kernel<L, A, S, Op, false><<<(unsigned)grid, 256, 0, l.stream>>>(
            (const L*)a.in, (S*)a.out, d, nout, nred, fin);
    }
    return cudaGetLastError();
}

cudaError_t launch_generic(int red, int dtin, int dtout, const GenericArgs& a, const Launch& l) {
    return dispatch_types(dtin, dtout, [&](auto t) -> cudaError_t {
        using T = decltype(t);
        using A = typename T::A;
        return dispatch_op<A>(red, [&](auto op) -> cudaError_t {
            using Op = decltype(op);
            return launch_generic_t<T, Op>(a, l);
        });
    });
}

// ---------------------------------------------------------------------------
// Empty reduction range: every output is the identity (the reference's init
// state runs even when the reduce maps are empty,
// reference dace/libraries/standard/nodes/reduce.py:90-102).
// ---------------------------------------------------------------------------
template <class S>
__global__ void fill_kernel(S* __restrict__ out, GenericDims d, long long nout, S value) {
    const long long tid = (long long)(blockIdx.x) * blockDim.x + threadIdx.x;
    const long long nthr = (long long)(gridDim.x) * blockDim.x;
    for (long long j = tid; j < nout; j += nthr) {
        long long in_off, out_off;
        kept_offsets(d, j, in_off, out_off);
        out[out_off] = value;
    }
}

template <class S>
static cudaError_t launch_fill_t(void* out, int nk, const long long* kshape, const long long* kout,
                                 const void* identity, const Launch& l) {
    GenericDims d;
    d.nk = nk; d.nr = 0;
    long long nout = 1;
    for (int k = 0; k < kMaxDims; ++k) {
        d.kshape[k] = k < nk ? kshape[k] : 1; d.kin[k] = 0; d.kout[k] = k < nk ? kout[k] : 0;
        d.rshape[k] = 1; d.rin[k] = 0;
        nout *= d.kshape[k];
    }
    if (nout == 0) return cudaSuccess;
    S value;
    memcpy(&value, identity, sizeof(S));
    long long grid = (nout + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    if (grid > cap) grid = cap;
    fill_kernel<S><<<(unsigned)grid, 256, 0, l.stream>>>((S*)out, d, nout, value);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Last step of a two-pass plan whose output is a 16-bit float: the fp32 result of the second pass
// is folded with the caller's identity (or with what `out` holds) and rounded ONCE to bf16 / f16.
// ---------------------------------------------------------------------------
template <class Op, class S>
__global__ void finish_convert_kernel(const float* __restrict__ tmp, S* __restrict__ out, size_t n, ptrdiff_t out_stride,
                                      Finish<float> fin) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) write_result<Op, float, S>(out + (ptrdiff_t)i * out_stride, tmp[i], fin);
}

template <class S>
static cudaError_t launch_finish_convert_t(int red, const float* tmp, void* out, size_t n, ptrdiff_t out_stride,
                                           const void* identity, const Launch& l) {
    return dispatch_op<float>(red, [&](auto op) -> cudaError_t {
        using Op = decltype(op);
        if (n == 0) return cudaSuccess;
        finish_convert_kernel<Op, S><<<(unsigned)((n + 255) / 256), 256, 0, l.stream>>>(
            tmp, static_cast<S*>(out), n, out_stride, make_finish<float, S>(identity));
        return cudaGetLastError();
    });
}

cudaError_t launch_finish_convert(int red, int dtout, const float* tmp, void* out, size_t n, ptrdiff_t out_stride,
                                  const void* identity, const Launch& l) {
    if (dtout == DT_BF16) return launch_finish_convert_t<__nv_bfloat16>(red, tmp, out, n, out_stride, identity, l);
    if (dtout == DT_F16) return launch_finish_convert_t<__half>(red, tmp, out, n, out_stride, identity, l);
    return cudaErrorNotSupported;
}

cudaError_t launch_fill(int dtout, void* out, int nk, const long long* kshape, const long long* kout,
                        const void* identity, const Launch& l) {
    switch (dtype_size(dtout)) {
        case 1: return launch_fill_t<uint8_t>(out, nk, kshape, kout, identity, l);
        case 2: return launch_fill_t<uint16_t>(out, nk, kshape, kout, identity, l);
        case 4: return launch_fill_t<uint32_t>(out, nk, kshape, kout, identity, l);
        case 8: return launch_fill_t<unsigned long long>(out, nk, kshape, kout, identity, l);
    }
    return cudaErrorNotSupported;
}

}  // namespace db200
