// Process-group communicator behind the C ABI: what a DaCe-generated program (one process per
// GPU) uses to combine the per-GPU partials of a sharded reduction.  Plays the role the MPI
// environment + Allreduce library node play in the reference (reference
// dace/libraries/mpi/environments/mpi.py:6-22, dace/libraries/mpi/nodes/allreduce.py:9-60), with
// NCCL over NVLink / NVSwitch instead of MPI, and with the one-shot peer-memory kernel
// (peer_combine.cu) for payloads <= 64 KiB.
//
// NCCL is bound at RUN time (dlopen of libnccl.so.2; an already loaded copy - e.g. the one a host
// PyTorch brought in - is reused), so the kernel library has no link-time NCCL dependency and a
// single-GPU program never touches it.  Only the handful of entry points below are used; their
// prototypes and enum values are NCCL's stable public ABI (nccl.h: ncclGetUniqueId,
// ncclCommInitRank, ncclAllReduce, ncclAllGather, ncclCommDestroy, ncclGetErrorString).
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>
#include <string>

#include "common.cuh"
#include "kernels.h"
#include "../../include/dace_b200_reduce.h"

namespace db200 {
namespace {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
enum { kNcclSum = 0, kNcclProd = 1, kNcclMax = 2, kNcclMin = 3 };
enum { kNcclInt8 = 0, kNcclUint8 = 1, kNcclInt32 = 2, kNcclUint32 = 3, kNcclInt64 = 4, kNcclUint64 = 5,
       kNcclFloat16 = 6, kNcclFloat32 = 7, kNcclFloat64 = 8, kNcclBfloat16 = 9 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* env = getenv("DACE_B200_NCCL_LIB");
        const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {   // an already loaded NCCL first: never two copies in one process
            if (n && !api.handle) api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        }
        for (const char* n : names) {
            if (n && !api.handle) api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        }
        if (!api.handle) return;
        auto sym = [&](const char* s) { return dlsym(api.handle, s); };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather;
    });
    return api;
}

constexpr size_t kPeerCap = 64 << 10;   // largest payload of the one-shot peer kernel
constexpr int kMaxWorld = 16;

struct Comm {
    int rank = 0, world = 1;
    NcclComm nccl = nullptr;
    // peer-memory exchange (optional: needs CUDA IPC between the ranks)
    bool peer = false;
    void* own = nullptr;
    void* bases[kMaxWorld] = {};
    unsigned long long epoch = 0;
    // Collectives of one communicator are chained in issue order even when the caller enqueues them on
    // different streams (independent AllReduceCombine nodes under DaCe's default stream assignment): the
    // one-shot peer kernel relies on epochs arriving in order (two slots by parity, flags overwritten), and
    // NCCL wants one order on every rank.
    std::mutex order_mutex;
    cudaEvent_t order = nullptr;
    cudaStream_t last_stream = nullptr;
    bool has_prev = false;
};

int fail(const char* what, int code) {
    if (getenv("DACE_B200_DEBUG")) {
        const char* msg = (nccl().GetErrorString && code > 0 && code < 16) ? nccl().GetErrorString(code) : "";
        fprintf(stderr, "dace_b200 comm: %s failed (%d) %s\n", what, code, msg);
    }
    return (int)cudaErrorUnknown;
}

// Maps every rank's exchange buffer into this process; all ranks agree on the outcome.
void setup_peer(Comm* c) {
    const char* off = getenv("DACE_B200_NO_PEER");
    int mine_ok = (off && off[0] == '1') ? 0 : 1;
    if (c->world > kMaxWorld) mine_ok = 0;
    const size_t bytes = peer_buffer_bytes(c->world, kPeerCap);
    unsigned char* dev_handles = nullptr;
    int* dev_flag = nullptr;
    cudaStream_t s = nullptr;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return;
    cudaIpcMemHandle_t h;
    memset(&h, 0, sizeof(h));
    if (mine_ok && dace_b200_peer_alloc(bytes, &c->own) != 0) mine_ok = 0;
    if (mine_ok && cudaIpcGetMemHandle(&h, c->own) != cudaSuccess) mine_ok = 0;
    cudaGetLastError();
    bool plumbing = cudaMalloc(&dev_handles, 64 * (size_t)c->world + 64) == cudaSuccess &&
                    cudaMalloc(&dev_flag, sizeof(int)) == cudaSuccess;
    unsigned char all[64 * kMaxWorld];
    if (plumbing && c->world <= kMaxWorld) {
        unsigned char* send = dev_handles + 64 * (size_t)c->world;
        cudaMemcpyAsync(send, &h, 64, cudaMemcpyHostToDevice, s);
        plumbing = nccl().AllGather(send, dev_handles, 64, kNcclUint8, c->nccl, s) == 0 &&
                   cudaMemcpyAsync(all, dev_handles, 64 * (size_t)c->world, cudaMemcpyDeviceToHost, s) == cudaSuccess &&
                   cudaStreamSynchronize(s) == cudaSuccess;
    } else {
        mine_ok = 0;
    }
    if (plumbing && mine_ok) {
        for (int r = 0; r < c->world && mine_ok; ++r) {
            if (r == c->rank) { c->bases[r] = c->own; continue; }
            cudaIpcMemHandle_t hr;
            memcpy(&hr, all + 64 * r, 64);
            if (cudaIpcOpenMemHandle(&c->bases[r], hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                c->bases[r] = nullptr;
                mine_ok = 0;
                cudaGetLastError();
            }
        }
    }
    // agreement (also the barrier that says "everybody has mapped everybody")
    int agreed = 0;
    if (plumbing) {
        cudaMemcpyAsync(dev_flag, &mine_ok, sizeof(int), cudaMemcpyHostToDevice, s);
        if (nccl().AllReduce(dev_flag, dev_flag, 1, kNcclInt32, kNcclMin, c->nccl, s) == 0 &&
            cudaMemcpyAsync(&agreed, dev_flag, sizeof(int), cudaMemcpyDeviceToHost, s) == cudaSuccess &&
            cudaStreamSynchronize(s) == cudaSuccess) {
            c->peer = agreed == 1;
        }
    }
    if (!c->peer) {
        for (int r = 0; r < c->world && r < kMaxWorld; ++r) {
            if (r != c->rank && c->bases[r]) cudaIpcCloseMemHandle(c->bases[r]);
            c->bases[r] = nullptr;
        }
        if (c->own) cudaFree(c->own);
        c->own = nullptr;
    }
    if (dev_handles) cudaFree(dev_handles);
    if (dev_flag) cudaFree(dev_flag);
    cudaStreamDestroy(s);
    cudaGetLastError();
}

std::mutex g_default_mutex;
Comm* g_default = nullptr;
bool g_default_owned = false;
const time_t g_loaded_at = time(nullptr);

int nccl_op(int red) {
    switch (red) {
        case RT_Sum: return kNcclSum;
        case RT_Product: return kNcclProd;
        case RT_Max: return kNcclMax;
        case RT_Min: return kNcclMin;
        default: return -1;
    }
}
int nccl_type(int dt) {
    switch (dt) {
        case DT_F32: return kNcclFloat32;
        case DT_F64: return kNcclFloat64;
        case DT_I32: return kNcclInt32;
        case DT_U32: return kNcclUint32;
        case DT_I64: return kNcclInt64;
        case DT_U64: return kNcclUint64;
        case DT_BF16: return kNcclBfloat16;
        case DT_F16: return kNcclFloat16;
        case DT_I8: return kNcclInt8;
        case DT_U8: return kNcclUint8;
        default: return -1;
    }
}
size_t type_bytes(int dt) {
    switch (dt) {
        case DT_F64: case DT_I64: case DT_U64: return 8;
        case DT_F32: case DT_I32: case DT_U32: return 4;
        case DT_BF16: case DT_F16: case DT_I16: case DT_U16: return 2;
        case DT_I8: case DT_U8: return 1;
        default: return 0;
    }
}

}  // namespace
}  // namespace db200

using namespace db200;

extern "C" {

int dace_b200_comm_unique_id(unsigned char* id128) {
    if (!id128) return (int)cudaErrorInvalidValue;
    if (!nccl().ok) return (int)cudaErrorNotSupported;
    NcclId id;
    int rc = nccl().GetUniqueId(&id);
    if (rc != 0) return fail("ncclGetUniqueId", rc);
    memcpy(id128, &id, 128);
    return 0;
}

int dace_b200_comm_create(int rank, int world, const unsigned char* id128, void** comm) {
    if (!comm || world < 1 || rank < 0 || rank >= world) return (int)cudaErrorInvalidValue;
    Comm* c = new Comm();
    c->rank = rank;
    c->world = world;
    if (world > 1) {
        if (!id128) { delete c; return (int)cudaErrorInvalidValue; }
        if (!nccl().ok) { delete c; return (int)cudaErrorNotSupported; }
        NcclId id;
        memcpy(&id, id128, 128);
        int rc = nccl().CommInitRank(&c->nccl, world, id, rank);
        if (rc != 0) { delete c; return fail("ncclCommInitRank", rc); }
        if (cudaEventCreateWithFlags(&c->order, cudaEventDisableTiming) != cudaSuccess) c->order = nullptr;
        setup_peer(c);
    }
    *comm = c;
    return 0;
}

/* RANK / WORLD_SIZE / LOCAL_RANK as set by torchrun or any launcher; the NCCL id travels through a
 * file: DACE_B200_RENDEZVOUS, default /tmp/dace_b200_nccl_<MASTER_ADDR>_<MASTER_PORT>.id. */
int dace_b200_comm_create_from_env(void** comm) {
    if (!comm) return (int)cudaErrorInvalidValue;
    const char* ws = getenv("WORLD_SIZE");
    const char* rk = getenv("RANK");
    const int world = ws ? atoi(ws) : 1;
    const int rank = rk ? atoi(rk) : 0;
    if (world <= 1) return dace_b200_comm_create(0, 1, nullptr, comm);
    std::string path;
    if (const char* p = getenv("DACE_B200_RENDEZVOUS")) {
        path = p;
    } else {
        const char* addr = getenv("MASTER_ADDR");
        const char* port = getenv("MASTER_PORT");
        path = std::string("/tmp/dace_b200_nccl_") + (addr ? addr : "local") + "_" + (port ? port : "0") + ".id";
    }
    unsigned char id[128];
    if (rank == 0) {
        int rc = dace_b200_comm_unique_id(id);
        if (rc != 0) return rc;
        std::string tmp = path + ".tmp";
        FILE* f = fopen(tmp.c_str(), "wb");
        if (!f) return (int)cudaErrorUnknown;
        const bool wrote = fwrite(id, 1, 128, f) == 128;
        fclose(f);
        if (!wrote || rename(tmp.c_str(), path.c_str()) != 0) return (int)cudaErrorUnknown;
    } else {
        bool got = false;
        for (int tries = 0; tries < 1200 && !got; ++tries) {   // up to 120 s
            struct stat st;
            // a file left behind by a crashed earlier job is older than this process: skip it
            if (stat(path.c_str(), &st) == 0 && st.st_size == 128 && st.st_mtime + 30 >= g_loaded_at) {
                FILE* f = fopen(path.c_str(), "rb");
                if (f) {
                    got = fread(id, 1, 128, f) == 128;
                    fclose(f);
                }
            }
            if (!got) usleep(100000);
        }
        if (!got) return (int)cudaErrorTimeout;
    }
    int rc = dace_b200_comm_create(rank, world, id, comm);
    if (rank == 0) unlink(path.c_str());   // CommInitRank returned: every rank has read the id
    return rc;
}

int dace_b200_comm_destroy(void* comm) {
    Comm* c = static_cast<Comm*>(comm);
    if (!c) return 0;
    {
        std::lock_guard<std::mutex> g(g_default_mutex);
        if (g_default == c) { g_default = nullptr; g_default_owned = false; }
    }
    if (c->peer) {
        cudaDeviceSynchronize();
        for (int r = 0; r < c->world; ++r)
            if (r != c->rank && c->bases[r]) cudaIpcCloseMemHandle(c->bases[r]);
    }
    if (c->nccl) nccl().CommDestroy(c->nccl);   // collective: peers have unmapped before the buffer goes away
    if (c->own) cudaFree(c->own);
    if (c->order) cudaEventDestroy(c->order);
    delete c;
    return 0;
}

int dace_b200_comm_rank(void* comm) { return comm ? static_cast<Comm*>(comm)->rank : 0; }
int dace_b200_comm_world(void* comm) { return comm ? static_cast<Comm*>(comm)->world : 1; }
int dace_b200_comm_uses_peer_memory(void* comm) { return comm && static_cast<Comm*>(comm)->peer ? 1 : 0; }

/* Process-wide communicator for generated programs: whatever the host installed with
 * dace_b200_comm_set_default(), else created from the environment on first use. */
int dace_b200_comm_set_default(void* comm) {
    std::lock_guard<std::mutex> g(g_default_mutex);
    g_default = static_cast<Comm*>(comm);
    g_default_owned = false;
    return 0;
}

int dace_b200_comm_default(void** comm) {
    if (!comm) return (int)cudaErrorInvalidValue;
    std::lock_guard<std::mutex> g(g_default_mutex);
    if (!g_default) {
        void* c = nullptr;
        int rc = dace_b200_comm_create_from_env(&c);
        if (rc != 0) return rc;
        g_default = static_cast<Comm*>(c);
        g_default_owned = true;
    }
    *comm = g_default;
    return 0;
}

int dace_b200_comm_allreduce(void* comm, int redtype, int dtype, void* data, size_t nelem, dace_b200_stream_t stream) {
    Comm* c = static_cast<Comm*>(comm);
    if (!c || (nelem > 0 && !data)) return (int)cudaErrorInvalidValue;
    const size_t esize = type_bytes(dtype);
    if (esize == 0) return (int)cudaErrorNotSupported;
    if (c->world == 1 || nelem == 0) return 0;
    int red = redtype;
    // logical partials are 0/1, so min / max / != are exact stand-ins (and NCCL has no bitwise operators)
    if (red == RT_LAnd) red = RT_Min;
    else if (red == RT_LOr) red = RT_Max;
    else if (red == RT_LXor) red = RT_BXor;
    const bool peer_type = dtype == DT_F32 || dtype == DT_F64 || dtype == DT_I32 || dtype == DT_U32 ||
                           dtype == DT_I64 || dtype == DT_U64;
    const bool peer_op = red == RT_Sum || red == RT_Product || red == RT_Min || red == RT_Max ||
                         ((red == RT_BAnd || red == RT_BOr || red == RT_BXor) && dtype != DT_F32 && dtype != DT_F64);
    const bool use_peer = c->peer && peer_type && peer_op && nelem * esize <= kPeerCap;
    const int op = nccl_op(red), ty = nccl_type(dtype);
    if (!use_peer && (op < 0 || ty < 0)) return (int)cudaErrorNotSupported;
    std::lock_guard<std::mutex> lock(c->order_mutex);
    cudaStream_t s = (cudaStream_t)stream;
    if (c->order && c->has_prev && s != c->last_stream) {
        cudaError_t e = cudaStreamWaitEvent(s, c->order, 0);   // the previous collective, on whatever stream it ran
        if (e != cudaSuccess) return (int)e;
    }
    int status;
    if (use_peer) {
        c->epoch += 1;
        status = (int)launch_peer_allreduce(red, dtype, data, nelem, c->bases, c->rank, c->world, c->epoch, kPeerCap, s);
    } else {
        int rc = nccl().AllReduce(data, data, nelem, ty, op, c->nccl, s);
        status = rc == 0 ? 0 : fail("ncclAllReduce", rc);
    }
    if (status == 0 && c->order) {
        cudaError_t e = cudaEventRecord(c->order, s);
        if (e != cudaSuccess) return (int)e;
        c->last_stream = s;
        c->has_prev = true;
    }
    return status;
}

}  // extern "C"
