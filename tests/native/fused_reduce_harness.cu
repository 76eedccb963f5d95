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


// The same functors with the optional vector interface (see VECTOR LOADS in the header): what
// dace_b200/fused_expansion.py generates for operands read at the map's flat index.
template <int VMAX>
struct ExpShiftV {
    const float* x; const float* m; size_t ld;
    static constexpr int kVec = VMAX;
    using value_type = float;
    __device__ float operator()(size_t r, size_t c) const { return expf(x[r * ld + c] - m[r]); }
    template <int V> __device__ void vec(size_t r, size_t c, float (&v)[V]) const {
        float xv[V];
        dace_b200::stream_load<float, V>(x + r * ld + c, xv);
        const float mr = m[r];
#pragma unroll
        for (int k = 0; k < V; ++k) v[k] = expf(xv[k] - mr);
    }
};
struct MulLLV {     // a[r, c] * w[c]: both operands vector-loaded (w along the columns too)
    const long long* a; const long long* w; size_t ld;
    static constexpr int kVec = 4;
    using value_type = long long;
    __device__ long long operator()(size_t r, size_t c) const { return a[r * ld + c] * w[c]; }
    template <int V> __device__ void vec(size_t r, size_t c, long long (&v)[V]) const {
        long long av[V], wv[V];
        dace_b200::stream_load<long long, V>(a + r * ld + c, av);
        dace_b200::stream_load<long long, V>(w + c, wv);
#pragma unroll
        for (int k = 0; k < V; ++k) v[k] = av[k] * wv[k];
    }
};
struct PlainLLV {
    const long long* x; size_t ld;
    static constexpr int kVec = 2;
    using value_type = long long;
    __device__ long long operator()(size_t r, size_t c) const { return x[r * ld + c]; }
    template <int V> __device__ void vec(size_t r, size_t c, long long (&v)[V]) const { dace_b200::stream_load<long long, V>(x + r * ld + c, v); }
};

// Widest legal vector access for this call, the way the generated call site computes it.
inline int max_vec(int vmax, size_t elem, unsigned long long n, const void* p0, const void* p1 = nullptr) {
    for (int v = vmax; v >= 2; v /= 2)
        if (n % v == 0 && reinterpret_cast<size_t>(p0) % (v * elem) == 0 && (p1 == nullptr || reinterpret_cast<size_t>(p1) % (v * elem) == 0))
            return v;
    return 1;
}

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
        /* vector-load functors: rows (40-43) and columns (50-53) */
        case 40: { const float id = 0.f; ExpShiftV<4> f{(const float*)in0, (const float*)in1, (size_t)n};
                   return fused_rows<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                         max_vec(4, 4, n, in0)); }
        case 41: { const float id = 0.f; ExpShiftV<8> f{(const float*)in0, (const float*)in1, (size_t)n};
                   return fused_rows<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                         max_vec(8, 4, n, in0)); }
        case 42: { const long long id = 11; MulLLV f{(const long long*)in0, (const long long*)in1, (size_t)n};
                   return fused_rows<kSum, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                                 max_vec(4, 8, n, in0, in1)); }
        case 43: { const long long id = 0; PlainLLV f{(const long long*)in0, (size_t)n};
                   return fused_rows_custom<AddPlusProduct, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                                                  max_vec(2, 8, n, in0)); }
        case 50: { const float id = 0.f; ExpShiftV<4> f{(const float*)in0, (const float*)in1, (size_t)n};
                   return fused_cols<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                         max_vec(4, 4, n, in0)); }
        case 51: { const float id = 0.f; ExpShiftV<8> f{(const float*)in0, (const float*)in1, (size_t)n};
                   return fused_cols<kSum, float, float>(f, (float*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                         max_vec(8, 4, n, in0)); }
        case 52: { const long long id = 11; MulLLV f{(const long long*)in0, (const long long*)in1, (size_t)n};
                   return fused_cols<kSum, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                                 max_vec(4, 8, n, in0, in1)); }
        case 53: { const long long id = 0; PlainLLV f{(const long long*)in0, (size_t)n};
                   return fused_cols_custom<AddPlusProduct, long long, long long>(f, (long long*)out, out_stride, rows, n, use_identity ? &id : nullptr, workspace, workspace_bytes, s,
                                                                                  max_vec(2, 8, n, in0)); }
        default: return cudaErrorInvalidValue;
    }
}
