// Test harness for include/dace_b200_fused_reduce.cuh: hand-written load functors of the kind the
// FuseMapIntoReduce rewrite generates, behind a plain C entry point (tests/test_gpu_fused_reduce.py).
#include <cuda_runtime.h>
#include <math.h>
#include "dace_b200_fused_reduce.cuh"

namespace {

struct ExpShift {   // exp(x[r, c] - m[r]): softmax denominator
    const float* x; const float* m; size_t ld;
    __device__ float operator()(size_t r, size_t c) const { return expf(x[r * ld + c] - m[r]); }
};
struct SquareD {    // x[r, c]^2: row norms in double
    const double* x; size_t ld;
    __device__ double operator()(size_t r, size_t c) const { const double v = x[r * ld + c]; return v * v; }
};
struct AbsDiffI {   // |a - b| on int32, max-reduced
    const int* a; const int* b; size_t ld;
    __device__ int operator()(size_t r, size_t c) const { const int d = a[r * ld + c] - b[r * ld + c]; return d < 0 ? -d : d; }
};
struct MulLL {      // a[r, c] * w[c] on int64, sum-reduced (dot product with a broadcast vector)
    const long long* a; const long long* w; size_t ld;
    __device__ long long operator()(size_t r, size_t c) const { return a[r * ld + c] * w[c]; }
};

// Custom WCR functors (no known neutral element): the lambda of reference tests/custom_reduce_test.py:7-9
// and an associative + commutative integer operator that is none of the built-ins (a + b + a * b).
struct MinByCompare { template <typename T> __device__ T operator()(const T& a, const T& b) const { return a < b ? a : b; } };
struct AddPlusProduct { template <typename T> __device__ T operator()(const T& a, const T& b) const { return a + b + a * b; } };
struct PlainF { const float* x; size_t ld; __device__ float operator()(size_t r, size_t c) const { return x[r * ld + c]; } };
struct PlainLL { const long long* x; size_t ld; __device__ long long operator()(size_t r, size_t c) const { return x[r * ld + c]; } };

}  // namespace

extern "C" __attribute__((visibility("default"))) int fused_reduce_harness_run(
    int case_id, const void* in0, const void* in1, void* out, long long out_stride, unsigned long long rows,
    unsigned long long n, int use_identity, void* workspace, unsigned long long workspace_bytes, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    using namespace dace_b200;
    switch (case_id) {
        case 0: { const float id = 0.f; ExpShift f{(const float*)in0, (const float*)in1, (size_t)n};
                  return fused_rows<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 1: { const double id = 0.0; SquareD f{(const double*)in0, (size_t)n};
                  return fused_rows<kSum, double, double>(f, (double*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 2: { const int id = -5; AbsDiffI f{(const int*)in0, (const int*)in1, (size_t)n};
                  return fused_rows<kMax, int, int>(f, (int*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 3: { const long long id = 11; MulLL f{(const long long*)in0, (const long long*)in1, (size_t)n};
                  return fused_rows<kSum, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        /* column folds: out[c] = fold_r load(r, c) */
        case 10: { const float id = 0.f; ExpShift f{(const float*)in0, (const float*)in1, (size_t)n};
                   return fused_cols<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 12: { const int id = -5; AbsDiffI f{(const int*)in0, (const int*)in1, (size_t)n};
                   return fused_cols<kMax, int, int>(f, (int*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 13: { const long long id = 11; MulLL f{(const long long*)in0, (const long long*)in1, (size_t)n};
                   return fused_cols<kSum, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        /* custom functors: rows (20, 21) and columns (30, 31) */
        case 20: { const float id = 9999999.f; PlainF f{(const float*)in0, (size_t)n};
                   return fused_rows_custom<MinByCompare, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 21: { const long long id = 0; PlainLL f{(const long long*)in0, (size_t)n};
                   return fused_rows_custom<AddPlusProduct, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 30: { const float id = 9999999.f; PlainF f{(const float*)in0, (size_t)n};
                   return fused_cols_custom<MinByCompare, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        case 31: { const long long id = 0; PlainLL f{(const long long*)in0, (size_t)n};
                   return fused_cols_custom<AddPlusProduct, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s); }
        default: return cudaErrorInvalidValue;
    }
}
