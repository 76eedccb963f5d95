// Test harness for include/dace_b200_shared_accum.cuh (tests/test_gpu_shared_accum.py): every block fills a
// shared tile from its slice of `src` (the thread that owns an entry folds the block's items for it - no
// shared-memory atomics needed, so every operator / type can be tested), then flushes the tile into `out`
// with dace_b200::shared_accum.  out[i * gstride] = RED(out[i * gstride], src[b][j][i] for all b, j).
#include <cuda_runtime.h>
#include "dace_b200_shared_accum.cuh"

namespace {

using namespace dace_b200;

template <int RED, typename T, int NT, int LEN>
__global__ void __launch_bounds__(NT) tile_kernel(const T* __restrict__ src, T* __restrict__ out, int items, int gstride) {
    __shared__ __align__(16) T tile[LEN];
    const int ltid = threadIdx.x;
    shared_reset<T, NT, LEN>(tile, ltid, detail::accum_neutral<RED, T>());
    const T* mine = src + size_t(blockIdx.x) * items * LEN;
    for (int i = ltid; i < LEN; i += NT) {
        T acc = tile[i];
        for (int j = 0; j < items; ++j) acc = op<RED, T>()(acc, mine[size_t(j) * LEN + i]);
        tile[i] = acc;
    }
    shared_accum<RED, T, NT, LEN>(tile, 1, out, gstride, ltid);
}

template <int RED, typename T, int NT, int LEN>
int run(const void* src, void* out, int nblocks, int items, int gstride, cudaStream_t s) {
    tile_kernel<RED, T, NT, LEN><<<nblocks, NT, 0, s>>>((const T*)src, (T*)out, items, gstride);
    return (int)cudaGetLastError();
}

}  // namespace

extern "C" __attribute__((visibility("default"))) int shared_accum_harness_len(int case_id) {
    switch (case_id) {
        case 0: return 1024; case 1: return 2048; case 2: return 1000; case 3: return 4096; case 4: return 512;
        case 5: return 300; case 6: return 256; case 7: return 128; case 8: return 777; case 9: return 64;
        default: return -1;
    }
}

extern "C" __attribute__((visibility("default"))) int shared_accum_harness_run(int case_id, const void* src, void* out,
                                                                              int nblocks, int items, int gstride,
                                                                              void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    switch (case_id) {
        case 0: return run<kSum, int, 256, 1024>(src, out, nblocks, items, gstride, s);
        case 1: return run<kSum, float, 128, 2048>(src, out, nblocks, items, gstride, s);      // gstride 1: red.global.add.v4.f32
        case 2: return run<kMax, int, 96, 1000>(src, out, nblocks, items, gstride, s);
        case 3: return run<kSum, unsigned long long, 1024, 4096>(src, out, nblocks, items, gstride, s);
        case 4: return run<kMax, float, 256, 512>(src, out, nblocks, items, gstride, s);
        case 5: return run<kMin, double, 64, 300>(src, out, nblocks, items, gstride, s);
        case 6: return run<kBitwiseXor, int, 128, 256>(src, out, nblocks, items, gstride, s);
        case 7: return run<kProduct, float, 128, 128>(src, out, nblocks, items, gstride, s);
        case 8: return run<kMin, long long, 160, 777>(src, out, nblocks, items, gstride, s);
        case 9: return run<kSum, double, 32, 64>(src, out, nblocks, items, gstride, s);
        default: return (int)cudaErrorInvalidValue;
    }
}
