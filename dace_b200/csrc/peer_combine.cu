// One-shot all-reduce of small per-GPU partials over NVLink / NVSwitch peer memory.
//
// When the sharded axis of a reduction is the reduced one (C2 full sum, C5 column sum over row
// shards), every GPU ends its local kernel with a small partial (4 B ... 64 KiB) that has to be
// combined across the GPUs of the box.  The reference's analogue is an MPI all-reduce library
// node after the local computation (reference dace/libraries/mpi/nodes/allreduce.py:9-60).
// An NCCL all-reduce of such a payload costs ~30 us of launch + protocol latency (measured: C2
// 0.590 ms kernel -> 0.620 ms step on 8 GPUs).  This kernel replaces it:
//
//   every rank owns one exchange buffer, mapped into every peer with CUDA IPC:
//       [ flags: 64 x u64 ][ slot 0: world x cap bytes ][ slot 1: world x cap bytes ]
//   step `epoch` (host counter, identical on all ranks):
//     1. PUSH: each rank stores its partial into slot (epoch & 1), region `rank`, of EVERY
//        peer's buffer (posted NVLink writes, no read round trips);
//     2. system-scope fence, then a release store of `epoch` into flags[rank] of every peer;
//     3. each rank waits (acquire loads) until its own flags[r] >= epoch for all r;
//     4. folds the `world` regions of its own slot in RANK ORDER - the same order on every
//        rank, so float results are bit-identical across ranks and from run to run - and
//        writes the result over its partial.
// Two slots are enough: a rank can only be one step ahead of the slowest peer, because step
// epoch+1 cannot complete before that peer has raised its flag for epoch+1, which its stream does
// only after its own step-`epoch` kernel (the last reader of slot epoch & 1) has finished.
#include "common.cuh"
#include "kernels.h"

namespace db200 {

constexpr int kPeerMaxWorld = 16;
constexpr size_t kPeerFlagBytes = 1024;
constexpr int kPeerThreads = 512;

struct PeerArgs {
    void* base[kPeerMaxWorld];   // this process' mapping of every rank's exchange buffer
    int rank, world;
    unsigned long long epoch;
    size_t cap_bytes;            // bytes reserved per (slot, rank) region
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <class T, class Op>
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_kernel(T* __restrict__ data, size_t nelem, PeerArgs p) {
    const unsigned tid = threadIdx.x;
    const size_t slot_off = kPeerFlagBytes + size_t(p.epoch & 1) * size_t(p.world) * p.cap_bytes;
    // Launched as a programmatic dependent of the kernel before it on the stream (the local reduction):
    // this CTA is already resident when that kernel's last CTA retires and only waits here for its
    // results to be visible - the launch gap between the two kernels (~4 us) is off the critical path.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // 1. push my partial to every rank (own buffer included)
    for (int r = 0; r < p.world; ++r) {
        const int dst_rank = (p.rank + r) % p.world;   // spread the first writes over the links
        T* dst = reinterpret_cast<T*>(static_cast<char*>(p.base[dst_rank]) + slot_off + size_t(p.rank) * p.cap_bytes);
        for (size_t i = tid; i < nelem; i += kPeerThreads) dst[i] = data[i];
    }
    // 2. publish
    __threadfence_system();
    __syncthreads();
    if (tid < unsigned(p.world))
        st_release_sys(static_cast<unsigned long long*>(p.base[tid]) + p.rank, p.epoch);
    // 3. wait for every sender
    if (tid < unsigned(p.world)) {
        const unsigned long long* f = static_cast<const unsigned long long*>(p.base[p.rank]) + tid;
        while (ld_acquire_sys(f) < p.epoch) {}
    }
    __syncthreads();
    // 4. fold in rank order (volatile loads: the regions were written by peers, not through this SM's L1)
    const char* mine = static_cast<const char*>(p.base[p.rank]) + slot_off;
    for (size_t i = tid; i < nelem; i += kPeerThreads) {
        T acc = __ldcv(reinterpret_cast<const T*>(mine) + i);
        for (int r = 1; r < p.world; ++r)
            acc = Op::apply(acc, __ldcv(reinterpret_cast<const T*>(mine + size_t(r) * p.cap_bytes) + i));
        data[i] = acc;
    }
}

template <class T>
static cudaError_t peer_launch_t(int red, void* data, size_t nelem, const PeerArgs& p, cudaStream_t stream) {
    return dispatch_op<T>(red, [&](auto op) -> cudaError_t {
        using Op = decltype(op);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(1);
        cfg.blockDim = dim3(kPeerThreads);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, peer_allreduce_kernel<T, Op>, static_cast<T*>(data), nelem, p);
    });
}

cudaError_t launch_peer_allreduce(int red, int dtype, void* data, size_t nelem, void* const* bases, int rank, int world,
                                  unsigned long long epoch, size_t cap_bytes, cudaStream_t stream) {
    if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world || epoch == 0) return cudaErrorInvalidValue;
    if (nelem == 0) return cudaSuccess;
    PeerArgs p;
    for (int r = 0; r < world; ++r) {
        if (bases[r] == nullptr) return cudaErrorInvalidValue;
        p.base[r] = bases[r];
    }
    p.rank = rank;
    p.world = world;
    p.epoch = epoch;
    p.cap_bytes = cap_bytes;
    size_t esize = 0;
    switch (dtype) {
        case DT_F32: case DT_I32: case DT_U32: esize = 4; break;
        case DT_F64: case DT_I64: case DT_U64: esize = 8; break;
        default: return cudaErrorNotSupported;
    }
    if (nelem * esize > cap_bytes) return cudaErrorInvalidValue;
    switch (dtype) {
        case DT_F32: return peer_launch_t<float>(red, data, nelem, p, stream);
        case DT_F64: return peer_launch_t<double>(red, data, nelem, p, stream);
        case DT_I32: return peer_launch_t<int>(red, data, nelem, p, stream);
        case DT_U32: return peer_launch_t<unsigned>(red, data, nelem, p, stream);
        case DT_I64: return peer_launch_t<long long>(red, data, nelem, p, stream);
        case DT_U64: return peer_launch_t<unsigned long long>(red, data, nelem, p, stream);
        default: return cudaErrorNotSupported;
    }
}

size_t peer_buffer_bytes(int world, size_t cap_bytes) { return kPeerFlagBytes + 2 * size_t(world) * cap_bytes; }

}  // namespace db200
