// Shared device/host helpers for the B200 (sm_100a) reduction kernel library.
//
// Everything here is written for sm_100a only: 256-bit global loads
// (LDG.E.256), redux.sync, thread-block clusters + DSMEM.  There is no other
// backend and no CPU fallback.
//
// Semantics mirrored from the reference runtime (read-only, not copied):
//   * binary operators            reference dace/runtime/include/dace/reduction.h:167-517
//   * ReductionType enum values   reference dace/runtime/include/dace/types.h:301-317
//   * xor-shuffle warp reduction  reference dace/runtime/include/dace/reduction.h:624-660
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cooperative_groups.h>
#include <cstdint>
#include <cstddef>
#include <type_traits>
#include <limits>
#include <cmath>

namespace db200 {
namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------
// Enums shared with include/dace_b200_reduce.h (kept numerically identical).
// ---------------------------------------------------------------------------
enum RedType : int {
    RT_Custom = 0, RT_Min = 1, RT_Max = 2, RT_Sum = 3, RT_Product = 4,
    RT_LAnd = 5, RT_BAnd = 6, RT_LOr = 7, RT_BOr = 8, RT_LXor = 9, RT_BXor = 10,
    RT_MinLoc = 11, RT_MaxLoc = 12, RT_Exchange = 13
};
enum DType : int {
    DT_F32 = 0, DT_F64 = 1, DT_I32 = 2, DT_I64 = 3, DT_U32 = 4, DT_U64 = 5,
    DT_BF16 = 6, DT_F16 = 7, DT_I8 = 8, DT_U8 = 9, DT_I16 = 10, DT_U16 = 11
};

__host__ __device__ inline int dtype_size(int dt) {
    switch (dt) {
        case DT_F32: case DT_I32: case DT_U32: return 4;
        case DT_F64: case DT_I64: case DT_U64: return 8;
        case DT_BF16: case DT_F16: case DT_I16: case DT_U16: return 2;
        case DT_I8: case DT_U8: return 1;
    }
    return 0;
}

constexpr int kWarp = 32;
constexpr unsigned kFullMask = 0xffffffffu;

// ---------------------------------------------------------------------------
// Conversions between storage types and accumulator types.
// ---------------------------------------------------------------------------
template <class A, class L>
__device__ __forceinline__ A to_acc(L v) { return static_cast<A>(v); }
template <>
__device__ __forceinline__ float to_acc<float, __nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <>
__device__ __forceinline__ float to_acc<float, __half>(__half v) { return __half2float(v); }

template <class S, class A>
__device__ __forceinline__ S from_acc(A v) { return static_cast<S>(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16, float>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half from_acc<__half, float>(float v) { return __float2half_rn(v); }

// ---------------------------------------------------------------------------
// Reduction operators.  `neutral()` is the operator's own neutral element used
// for padding lanes; it is NOT the user-supplied identity, which is folded in
// exactly once when the final value is written.
// ---------------------------------------------------------------------------
template <class A> struct OpSum {
    static constexpr int kRed = RT_Sum;
    static __device__ __forceinline__ A neutral() { return A(0); }
    static __device__ __forceinline__ A apply(A a, A b) { return a + b; }
};
template <class A> struct OpProduct {
    static constexpr int kRed = RT_Product;
    static __device__ __forceinline__ A neutral() { return A(1); }
    static __device__ __forceinline__ A apply(A a, A b) { return a * b; }
};
template <class A> struct OpMin {
    static constexpr int kRed = RT_Min;
    static __device__ __forceinline__ A neutral() {
        if constexpr (std::is_same<A, float>::value) return __int_as_float(0x7f800000);
        else if constexpr (std::is_same<A, double>::value) return __longlong_as_double(0x7ff0000000000000LL);
        else return std::numeric_limits<A>::max();
    }
    static __device__ __forceinline__ A apply(A a, A b) {
        if constexpr (std::is_same<A, float>::value) return fminf(a, b);
        else if constexpr (std::is_same<A, double>::value) return fmin(a, b);
        else return a < b ? a : b;
    }
};
template <class A> struct OpMax {
    static constexpr int kRed = RT_Max;
    static __device__ __forceinline__ A neutral() {
        if constexpr (std::is_same<A, float>::value) return __int_as_float(0xff800000);
        else if constexpr (std::is_same<A, double>::value) return __longlong_as_double(0xfff0000000000000LL);
        else return std::numeric_limits<A>::lowest();
    }
    static __device__ __forceinline__ A apply(A a, A b) {
        if constexpr (std::is_same<A, float>::value) return fmaxf(a, b);
        else if constexpr (std::is_same<A, double>::value) return fmax(a, b);
        else return a > b ? a : b;
    }
};
template <class A> struct OpBAnd {
    static constexpr int kRed = RT_BAnd;
    static __device__ __forceinline__ A neutral() { return static_cast<A>(~A(0)); }
    static __device__ __forceinline__ A apply(A a, A b) { return a & b; }
};
template <class A> struct OpBOr {
    static constexpr int kRed = RT_BOr;
    static __device__ __forceinline__ A neutral() { return A(0); }
    static __device__ __forceinline__ A apply(A a, A b) { return a | b; }
};
template <class A> struct OpBXor {
    static constexpr int kRed = RT_BXor;
    static __device__ __forceinline__ A neutral() { return A(0); }
    static __device__ __forceinline__ A apply(A a, A b) { return a ^ b; }
};
// Logical operators act on truthiness and yield 0/1 like the C++ `a && b`.
template <class A> struct OpLAnd {
    static constexpr int kRed = RT_LAnd;
    static __device__ __forceinline__ A neutral() { return A(1); }
    static __device__ __forceinline__ A apply(A a, A b) { return A((a != A(0)) && (b != A(0))); }
};
template <class A> struct OpLOr {
    static constexpr int kRed = RT_LOr;
    static __device__ __forceinline__ A neutral() { return A(0); }
    static __device__ __forceinline__ A apply(A a, A b) { return A((a != A(0)) || (b != A(0))); }
};
template <class A> struct OpLXor {
    static constexpr int kRed = RT_LXor;
    static __device__ __forceinline__ A neutral() { return A(0); }
    static __device__ __forceinline__ A apply(A a, A b) { return A((a != A(0)) != (b != A(0))); }
};

// ---------------------------------------------------------------------------
// Streaming global loads: read-once data, keep it out of L1 and mark it
// evict-first in L2.  32-byte form is the sm_100 LDG.E.256.
// ---------------------------------------------------------------------------
template <int BYTES> struct Words { uint32_t w[BYTES / 4]; };

__device__ __forceinline__ Words<16> ldg_stream16(const void* p) {
    Words<16> r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    return r;
}
__device__ __forceinline__ Words<8> ldg_stream8(const void* p) {
    Words<8> r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.w[0]), "=r"(r.w[1]) : "l"(p));
    return r;
}
__device__ __forceinline__ Words<32> ldg_stream32(const void* p) {
    Words<32> r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]),
                   "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]) : "l"(p));
    return r;
}
template <int BYTES> __device__ __forceinline__ Words<BYTES> ldg_stream(const void* p);
template <> __device__ __forceinline__ Words<8> ldg_stream<8>(const void* p) { return ldg_stream8(p); }
template <> __device__ __forceinline__ Words<16> ldg_stream<16>(const void* p) { return ldg_stream16(p); }
template <> __device__ __forceinline__ Words<32> ldg_stream<32>(const void* p) { return ldg_stream32(p); }

template <class L>
__device__ __forceinline__ L ldg_scalar(const L* p) {
    if constexpr (sizeof(L) == 1) {
        uint8_t v = __ldg(reinterpret_cast<const uint8_t*>(p));
        return *reinterpret_cast<L*>(&v);
    } else if constexpr (sizeof(L) == 2) {
        unsigned short v = __ldg(reinterpret_cast<const unsigned short*>(p));
        return *reinterpret_cast<L*>(&v);
    } else if constexpr (sizeof(L) == 4) {
        unsigned v = __ldg(reinterpret_cast<const unsigned*>(p));
        return *reinterpret_cast<L*>(&v);
    } else {
        unsigned long long v = __ldg(reinterpret_cast<const unsigned long long*>(p));
        return *reinterpret_cast<L*>(&v);
    }
}

// Element j of a packed vector of storage type L, converted to the accumulator type.
template <class L, class A, int BYTES>
__device__ __forceinline__ A vec_elem(const Words<BYTES>& v, int j) {
    if constexpr (std::is_same<L, __nv_bfloat16>::value) {
        // bf16 -> f32 is a 16-bit shift: exact and cheaper than a cvt.
        uint32_t w = v.w[j >> 1];
        return __uint_as_float((j & 1) ? (w & 0xffff0000u) : (w << 16));
    } else if constexpr (std::is_same<L, __half>::value) {
        uint32_t w = v.w[j >> 1];
        __half2 h = *reinterpret_cast<__half2*>(&w);
        return (j & 1) ? __high2float(h) : __low2float(h);
    } else if constexpr (sizeof(L) == 4) {
        uint32_t w = v.w[j];
        return static_cast<A>(*reinterpret_cast<L*>(&w));
    } else if constexpr (sizeof(L) == 8) {
        unsigned long long w = (static_cast<unsigned long long>(v.w[2 * j + 1]) << 32) | v.w[2 * j];
        return static_cast<A>(*reinterpret_cast<L*>(&w));
    } else if constexpr (sizeof(L) == 2) {
        uint32_t w = v.w[j >> 1];
        unsigned short s = (j & 1) ? (w >> 16) : (w & 0xffffu);
        return static_cast<A>(*reinterpret_cast<L*>(&s));
    } else {
        uint32_t w = v.w[j >> 2];
        uint8_t b = (w >> (8 * (j & 3))) & 0xffu;
        return static_cast<A>(*reinterpret_cast<L*>(&b));
    }
}

// ---------------------------------------------------------------------------
// Warp-level tree: redux.sync for 32-bit integer accumulators, xor-shuffle
// butterflies otherwise.  WIDTH lanes take part (power of two, <= 32); lanes
// are grouped as aligned WIDTH-lane segments.
// ---------------------------------------------------------------------------
template <class A>
__device__ __forceinline__ A shfl_xor(A v, int m) {
    if constexpr (sizeof(A) == 8) {
        unsigned long long u = *reinterpret_cast<unsigned long long*>(&v);
        u = __shfl_xor_sync(kFullMask, u, m);
        return *reinterpret_cast<A*>(&u);
    } else {
        return __shfl_xor_sync(kFullMask, v, m);
    }
}

template <class Op, class A, int WIDTH = 32>
__device__ __forceinline__ A warp_reduce(A v) {
    if constexpr (WIDTH == 32 && sizeof(A) == 4 && std::is_integral<A>::value) {
        // redux.sync: one instruction instead of five shuffle+op pairs.
        if constexpr (Op::kRed == RT_Sum) return static_cast<A>(__reduce_add_sync(kFullMask, v));
        else if constexpr (Op::kRed == RT_Min) return static_cast<A>(__reduce_min_sync(kFullMask, v));
        else if constexpr (Op::kRed == RT_Max) return static_cast<A>(__reduce_max_sync(kFullMask, v));
        else if constexpr (Op::kRed == RT_BAnd) return static_cast<A>(__reduce_and_sync(kFullMask, static_cast<unsigned>(v)));
        else if constexpr (Op::kRed == RT_BOr) return static_cast<A>(__reduce_or_sync(kFullMask, static_cast<unsigned>(v)));
        else if constexpr (Op::kRed == RT_BXor) return static_cast<A>(__reduce_xor_sync(kFullMask, static_cast<unsigned>(v)));
    }
#pragma unroll
    for (int m = WIDTH / 2; m >= 1; m >>= 1) v = Op::apply(v, shfl_xor(v, m));
    return v;
}

// Reduce across lanes that share `lane % stride_lanes` (stride_lanes a power of two <= 32).
template <class Op, class A>
__device__ __forceinline__ A warp_reduce_strided(A v, int stride_lanes) {
    for (int m = 16; m >= stride_lanes; m >>= 1) v = Op::apply(v, shfl_xor(v, m));
    return v;
}

// Block-wide reduce; result valid in thread 0.  `scratch` holds >= 32 elements of A.
template <class Op, class A>
__device__ __forceinline__ A block_reduce(A v, A* scratch) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    v = warp_reduce<Op, A>(v);
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    if (warp == 0) {
        A x = lane < nwarps ? scratch[lane] : Op::neutral();
        x = warp_reduce<Op, A>(x);
        v = x;
    }
    return v;
}

// ---------------------------------------------------------------------------
// Workspace layout (device memory, owned by the caller or by the library's
// lazily created per-device default): kCounters self-resetting tickets
// followed by the partials area.
// ---------------------------------------------------------------------------
constexpr size_t kCounters = 4096;
constexpr size_t kCounterBytes = kCounters * sizeof(unsigned);
constexpr size_t kWorkspaceBytes = size_t(8) << 20;
constexpr size_t kPartialBytes = kWorkspaceBytes - kCounterBytes;

struct Workspace {
    unsigned* counters;
    void* partials;
};

// How the final value is written: out = op(identity, r) when an identity is
// given, out = op(out, r) otherwise (reference dace/frontend/operations.py:251,
// reference dace/libraries/standard/nodes/reduce.py:90-105).
template <class A>
struct Finish {
    A identity;
    int has_identity;
};

template <class Op, class A, class S>
__device__ __forceinline__ void write_result(S* out, A r, const Finish<A>& fin) {
    if (fin.has_identity == 2) {  // degenerate reduction: plain copy, identity ignored
        *out = from_acc<S, A>(r);
        return;
    }
    A base = fin.has_identity ? fin.identity : to_acc<A, S>(*out);
    *out = from_acc<S, A>(Op::apply(base, r));
}

// ---------------------------------------------------------------------------
// Host-side dispatch over (storage in, accumulator, storage out) and operator.
// ---------------------------------------------------------------------------
template <class L_, class A_, class S_> struct Types { using L = L_; using A = A_; using S = S_; };

template <class F>
inline cudaError_t dispatch_types(int dtin, int dtout, F&& f) {
#define DB200_T(DI, DO, L, A, S) if (dtin == DI && dtout == DO) return f(Types<L, A, S>{});
    DB200_T(DT_F32, DT_F32, float, float, float)
    DB200_T(DT_F64, DT_F64, double, double, double)
    DB200_T(DT_I32, DT_I32, int, int, int)
    DB200_T(DT_I64, DT_I64, long long, long long, long long)
    DB200_T(DT_U32, DT_U32, unsigned, unsigned, unsigned)
    DB200_T(DT_U64, DT_U64, unsigned long long, unsigned long long, unsigned long long)
    DB200_T(DT_BF16, DT_F32, __nv_bfloat16, float, float)
    DB200_T(DT_F16, DT_F32, __half, float, float)
    DB200_T(DT_BF16, DT_BF16, __nv_bfloat16, float, __nv_bfloat16)
    DB200_T(DT_F16, DT_F16, __half, float, __half)
    DB200_T(DT_I8, DT_I8, signed char, int, signed char)
    DB200_T(DT_U8, DT_U8, unsigned char, unsigned, unsigned char)
    DB200_T(DT_I16, DT_I16, short, int, short)
    DB200_T(DT_U16, DT_U16, unsigned short, unsigned, unsigned short)
#undef DB200_T
    return cudaErrorNotSupported;
}

template <class A, class F>
inline cudaError_t dispatch_op(int red, F&& f) {
    switch (red) {
        case RT_Sum: return f(OpSum<A>{});
        case RT_Product: return f(OpProduct<A>{});
        case RT_Min: return f(OpMin<A>{});
        case RT_Max: return f(OpMax<A>{});
        default: break;
    }
    if constexpr (std::is_integral<A>::value) {
        switch (red) {
            case RT_BAnd: return f(OpBAnd<A>{});
            case RT_BOr: return f(OpBOr<A>{});
            case RT_BXor: return f(OpBXor<A>{});
            case RT_LAnd: return f(OpLAnd<A>{});
            case RT_LOr: return f(OpLOr<A>{});
            case RT_LXor: return f(OpLXor<A>{});
            default: break;
        }
    }
    return cudaErrorNotSupported;
}

// Identity arrives as a host pointer to a value of the OUTPUT storage type.
template <class A, class S>
inline Finish<A> make_finish(const void* identity) {
    Finish<A> fin;
    fin.has_identity = identity != nullptr;
    if (identity) {
        if constexpr (std::is_same<S, __nv_bfloat16>::value) {
            uint16_t b = *reinterpret_cast<const uint16_t*>(identity);
            uint32_t w = uint32_t(b) << 16;
            float f;
            memcpy(&f, &w, 4);
            fin.identity = f;
        } else if constexpr (std::is_same<S, __half>::value) {
            __half h;
            memcpy(&h, identity, 2);
            fin.identity = __half2float(h);
        } else {
            S s;
            memcpy(&s, identity, sizeof(S));
            fin.identity = static_cast<A>(s);
        }
    } else {
        fin.identity = A(0);
    }
    return fin;
}

int sm_count();

// Launch context shared by the kernels: the stream and the caller's workspace (tickets + partials,
// followed by whatever extra scratch the caller added on top of dace_b200_workspace_bytes()).
struct Launch {
    cudaStream_t stream;
    Workspace ws;
    void* extra = nullptr;       // scratch past the base workspace (two-pass `mid`, partition buffers)
    size_t extra_bytes = 0;
    // Dry run of the planner (dace_b200_reduce_nd_workspace_bytes): nothing is launched, *dry_need becomes the
    // largest scratch offset a real run of the same plan would touch.
    size_t* dry_need = nullptr;
    size_t dry_base = 0;
    void take(size_t bytes) const { if (dry_need && dry_base + bytes > *dry_need) *dry_need = dry_base + bytes; }
    // Dry runs may also record WHICH plan the planner picked (dace_b200_reduce_nd_plan): a space-separated trace.
    char* dry_plan = nullptr;
    size_t dry_plan_bytes = 0;
    bool slice_of_fold = false;  // this call plans one slice of a host-peeled REDUCED axis (api.cu, plan_dims)
    void note(const char* what) const {
        if (!dry_plan || dry_plan_bytes == 0) return;
        size_t used = 0;
        while (used < dry_plan_bytes && dry_plan[used]) ++used;
        if (used && used + 1 < dry_plan_bytes) dry_plan[used++] = ' ';
        while (*what && used + 1 < dry_plan_bytes) dry_plan[used++] = *what++;
        if (used < dry_plan_bytes) dry_plan[used] = 0;
    }
};

// Experiment knobs (environment variables), read ONCE when the library is loaded - never on a call path.
struct Tuning {
    bool no_two_pass, no_axis_sort, no_peel, no_wide_cols, no_few_rows, no_narrow_cols, no_partition, no_tiled, no_tiny_rows,
         no_outer_first;
    int wide_wc;            // DACE_B200_WIDE_WC (0 = default)
    int rows_lane_bytes;    // DACE_B200_ROWS_LANE_BYTES (0 = default)
    int part_chunk_log2;    // DACE_B200_PART_CHUNK_LOG2 (0 = default)
    int part_divert;        // DACE_B200_PART_DIVERT (-1 = default)
    int part_hint;          // DACE_B200_PART_HINT (-1 = default)
};
const Tuning& tuning();

// Per-device "done once" flags for cudaFuncSetAttribute (function attributes are per device).
// CTAs of one kernel instantiation that fit an SM (the register and shared-memory use of THIS build decide).
// Persistent grids are sized as sm_count() * this: a grid sized for 8 CTAs/SM of a kernel that fits 6 runs 1.33
// waves, and since every CTA of a persistent kernel carries the same share of the work the last third of a wave
// costs a whole one (measured, rows [8M, 33]: 0.59 -> 0.8 of the HBM peak).  Host-side arithmetic on the function's
// attributes, no synchronisation; cached per device.
template <auto Kernel>
inline int resident_ctas(int threads, size_t dynamic_smem = 0) {
    static int cache[64];                         // 0 = not asked yet
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    int v = cache[dev];
    if (v == 0) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, Kernel, threads, dynamic_smem) != cudaSuccess || v < 1) v = 1;
        cache[dev] = v;
    }
    return v;
}

struct PerDeviceOnce {
    bool done[64] = {false};
    bool* slot() {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
        return &done[dev];
    }
};

}  // namespace db200
