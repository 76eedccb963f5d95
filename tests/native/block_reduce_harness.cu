// Test harness for include/dace_b200_block_reduce.cuh: one kernel per (type, block size, operator) case,
// launched through a plain C entry point so tests/test_gpu_block_reduce.py can drive it with ctypes.
// Every thread loads ITEMS consecutive values; outputs: block_reduce result (thread 0) per block and the
// block_allreduce result of the FIRST item per thread (must be identical in all threads).
#include <cuda_runtime.h>
#include "dace_b200_block_reduce.cuh"

namespace {

struct ValIdx {
    double v;
    long long i;
};
struct ArgMaxOp {  // custom functor on a 16-byte struct (value/index pairs, reference reduction.h:56-88)
    __device__ ValIdx operator()(const ValIdx& a, const ValIdx& b) const {
        return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
    }
};
struct AffineOp {  // custom lambda a + b + a*b on integers (associative, commutative)
    template <typename T>
    __device__ T operator()(const T& a, const T& b) const { return a + b + a * b; }
};

template <typename T, int NT, int ITEMS, typename Op>
__global__ void __launch_bounds__(NT) case_kernel(const T* in, T* out, T* out_all, int dimx) {
    const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
    T r[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) r[j] = in[((size_t)blockIdx.x * NT + tid) * ITEMS + j];
    T s = dace_b200::block_reduce<T, NT, 1>(r, Op());
    if (tid == 0) out[blockIdx.x] = s;
    // same call site in a loop: the second round must not be disturbed by the first
    T a = r[0];
    for (int round = 0; round < 2; ++round) a = dace_b200::block_allreduce<T, NT, 2>(round == 0 ? r[0] : r[0], Op());
    out_all[(size_t)blockIdx.x * NT + tid] = a;
}

template <int NT>
__global__ void __launch_bounds__(NT) argmax_kernel(const double* in, ValIdx* out, ValIdx* out_all) {
    ValIdx p{in[(size_t)blockIdx.x * NT + threadIdx.x], (long long)threadIdx.x};
    ValIdx r = dace_b200::block_reduce<ValIdx, NT, 3>(p, ArgMaxOp());
    if (threadIdx.x == 0) out[blockIdx.x] = r;
    out_all[(size_t)blockIdx.x * NT + threadIdx.x] = dace_b200::block_allreduce<ValIdx, NT, 4>(p, ArgMaxOp());
}

__global__ void warp_kernel(const int* in, int* out_sum, const float* fin, float* out_max) {
    const int t = blockIdx.x * 32 + threadIdx.x;
    out_sum[t] = dace_b200::warp_allreduce(in[t], dace_b200::op<dace_b200::kSum, int>());
    out_max[t] = dace_b200::warp_allreduce(fin[t], dace_b200::op<dace_b200::kMax, float>());
}

template <typename T, int NT, int ITEMS, typename Op>
cudaError_t run(const void* in, void* out, void* out_all, int nblocks, int dimx, cudaStream_t s) {
    dim3 block(dimx, NT / dimx, 1);
    case_kernel<T, NT, ITEMS, Op><<<nblocks, block, 0, s>>>((const T*)in, (T*)out, (T*)out_all, dimx);
    return cudaGetLastError();
}

}  // namespace

using namespace dace_b200;

// case table (kept in sync with CASES in tests/test_gpu_block_reduce.py)
extern "C" __attribute__((visibility("default"))) int block_reduce_harness_run(int case_id, const void* in, void* out,
                                                                              void* out_all, int nblocks, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    switch (case_id) {
        case 0: return run<float, 32, 2, op<kSum, float>>(in, out, out_all, nblocks, 32, s);
        case 1: return run<float, 128, 1, op<kSum, float>>(in, out, out_all, nblocks, 128, s);
        case 2: return run<int, 96, 3, op<kSum, int>>(in, out, out_all, nblocks, 96, s);
        case 3: return run<int, 50, 2, op<kMax, int>>(in, out, out_all, nblocks, 50, s);
        case 4: return run<double, 1024, 1, op<kMin, double>>(in, out, out_all, nblocks, 1024, s);
        case 5: return run<unsigned, 32, 4, op<kBitwiseXor, unsigned>>(in, out, out_all, nblocks, 32, s);
        case 6: return run<long long, 7, 2, op<kSum, long long>>(in, out, out_all, nblocks, 7, s);
        case 7: return run<long long, 1000, 1, op<kMax, long long>>(in, out, out_all, nblocks, 1000, s);
        case 8: return run<int, 256, 1, op<kMin, int>>(in, out, out_all, nblocks, 16, s);       /* 16 x 16 block */
        case 9: return run<unsigned, 33, 1, op<kBitwiseAnd, unsigned>>(in, out, out_all, nblocks, 33, s);
        case 10: return run<unsigned, 480, 2, op<kBitwiseOr, unsigned>>(in, out, out_all, nblocks, 480, s);
        case 11: return run<double, 65, 3, op<kProduct, double>>(in, out, out_all, nblocks, 65, s);
        case 12: return run<int, 192, 1, op<kLogicalAnd, int>>(in, out, out_all, nblocks, 192, s);
        case 13: return run<int, 192, 1, op<kLogicalOr, int>>(in, out, out_all, nblocks, 192, s);
        case 14: return run<int, 100, 1, op<kLogicalXor, int>>(in, out, out_all, nblocks, 100, s);
        case 15: return run<long long, 72, 2, AffineOp>(in, out, out_all, nblocks, 72, s);
        case 16: return run<float, 1, 5, op<kSum, float>>(in, out, out_all, nblocks, 1, s);
        case 17: return run<float, 1024, 2, op<kMax, float>>(in, out, out_all, nblocks, 32, s);   /* 32 x 32 block */
        case 100:
            argmax_kernel<200><<<nblocks, 200, 0, s>>>((const double*)in, (ValIdx*)out, (ValIdx*)out_all);
            return cudaGetLastError();
        default: return cudaErrorInvalidValue;
    }
}

extern "C" __attribute__((visibility("default"))) int block_reduce_harness_warp(const int* in, int* out_sum,
                                                                               const float* fin, float* out_max,
                                                                               int nwarps, void* stream) {
    warp_kernel<<<nwarps, 32, 0, (cudaStream_t)stream>>>(in, out_sum, fin, out_max);
    return cudaGetLastError();
}
