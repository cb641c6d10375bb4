// crm_comm.cu — NCCL transport of the sharded step loop, called from the C++ host (crm_sim.cu
// crm_sim_steps_sharded): halo send/recv with the two slab neighbours and the all-gather of the per-step flags,
// enqueued on the sim's stream so that a batch of sharded steps never waits for the host.  The reference's
// sub-domain threads exchange boundary agents through channels inside one process (cellular_raza chili backend,
// reference src/simulation.rs:251-256,283-292); here the sub-domains are GPUs on NVLink.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy torch has already loaded when the ranks were
// started through torch.distributed, else the system one), so the library itself does not depend on it and
// single-GPU use needs no NCCL at all.  Types come from nccl.h at build time.
#include <dlfcn.h>
#include <nccl.h>

#include "crm_sim.h"

namespace {

struct Api {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

Api* api() {
    static Api a;
    static bool tried = false;
    if (tried) return &a;
    tried = true;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
        a.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (a.handle) break;
    }
    if (!a.handle) {
        a.error = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "");
        return &a;
    }
    bool ok = true;
    auto sym = [&](const char* n) {
        void* p = dlsym(a.handle, n);
        if (!p) { ok = false; a.error = std::string("missing NCCL symbol ") + n; }
        return p;
    };
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
    a.Send = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
    a.Recv = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
    if (!ok) {
        dlclose(a.handle);
        a.handle = nullptr;
    }
    return &a;
}

}  // namespace

struct crm_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    bool fused = true;  // halo send/recv and the flag all-gather in one NCCL group (CRM_B200_NCCL_FUSED=0: two launches)
};

#define NK(call)                                                                                         \
    do {                                                                                                 \
        ncclResult_t r_ = (call);                                                                        \
        if (r_ != ncclSuccess)                                                                           \
            return crm_fail(CRM_ERR_CUDA, "%s failed: %s", #call, api()->GetErrorString(r_));            \
    } while (0)

int crm_comm_rank(const crm_comm* c) { return c ? c->rank : 0; }
int crm_comm_world(const crm_comm* c) { return c ? c->world : 1; }

// both halo messages of a step: [lo] <-> rank - 1, [hi] <-> rank + 1 (byte counts; 0 = nothing on that side)
int crm_comm_sendrecv(crm_comm* c, const void* send_lo, size_t n_send_lo, void* recv_lo, size_t n_recv_lo,
                      const void* send_hi, size_t n_send_hi, void* recv_hi, size_t n_recv_hi, cudaStream_t st) {
    Api* a = api();
    NK(a->GroupStart());
    if (c->rank > 0) {
        if (n_send_lo) NK(a->Send(send_lo, n_send_lo, ncclUint8, c->rank - 1, c->comm, st));
        if (n_recv_lo) NK(a->Recv(recv_lo, n_recv_lo, ncclUint8, c->rank - 1, c->comm, st));
    }
    if (c->rank + 1 < c->world) {
        if (n_send_hi) NK(a->Send(send_hi, n_send_hi, ncclUint8, c->rank + 1, c->comm, st));
        if (n_recv_hi) NK(a->Recv(recv_hi, n_recv_hi, ncclUint8, c->rank + 1, c->comm, st));
    }
    NK(a->GroupEnd());
    return 0;
}

int crm_comm_exchange(crm_comm* c, const void* send_lo, size_t n_send_lo, void* recv_lo, size_t n_recv_lo, const void* send_hi,
                      size_t n_send_hi, void* recv_hi, size_t n_recv_hi, const void* flags, void* flags_all,
                      size_t flag_bytes, cudaStream_t st) {
    Api* a = api();
    if (c->fused) {
        NK(a->GroupStart());
        if (c->rank > 0) {
            if (n_send_lo) NK(a->Send(send_lo, n_send_lo, ncclUint8, c->rank - 1, c->comm, st));
            if (n_recv_lo) NK(a->Recv(recv_lo, n_recv_lo, ncclUint8, c->rank - 1, c->comm, st));
        }
        if (c->rank + 1 < c->world) {
            if (n_send_hi) NK(a->Send(send_hi, n_send_hi, ncclUint8, c->rank + 1, c->comm, st));
            if (n_recv_hi) NK(a->Recv(recv_hi, n_recv_hi, ncclUint8, c->rank + 1, c->comm, st));
        }
        NK(a->AllGather(flags, flags_all, flag_bytes, ncclUint8, c->comm, st));
        NK(a->GroupEnd());
        return 0;
    }
    int rc = crm_comm_sendrecv(c, send_lo, n_send_lo, recv_lo, n_recv_lo, send_hi, n_send_hi, recv_hi, n_recv_hi, st);
    return rc ? rc : crm_comm_allgather(c, flags, flags_all, flag_bytes, st);
}

int crm_comm_allgather(crm_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st) {
    NK(api()->AllGather(send, recv, bytes_per_rank, ncclUint8, c->comm, st));
    return 0;
}

void crm_comm_free(crm_comm* c) {
    if (!c) return;
    if (c->comm && api()->handle) api()->CommDestroy(c->comm);
    delete c;
}

extern "C" {

// 128-byte NCCL unique id, created by one rank and handed to all ranks by the caller (any side channel:
// torch.distributed broadcast, MPI, a file).
int crm_comm_unique_id(void* out128) {
    if (!out128) return crm_fail(CRM_ERR_INVALID, "null argument");
    Api* a = api();
    if (!a->handle) return crm_fail(CRM_ERR_CUDA, "NCCL unavailable: %s", a->error.c_str());
    static_assert(sizeof(ncclUniqueId) == 128, "unique id size");
    ncclUniqueId id;
    NK(a->GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return 0;
}

// Communicator of all slabs on `device`; rank r owns slab r.  Collective.  A communicator outlives the sims that
// use it (its connections are set up lazily by the first transfers, hundreds of milliseconds: made once per job).
int crm_comm_create(const void* id128, int rank, int world, int device, crm_comm** out) {
    if (!id128 || !out || rank < 0 || rank >= world) return crm_fail(CRM_ERR_INVALID, "bad communicator arguments");
    Api* a = api();
    if (!a->handle) return crm_fail(CRM_ERR_CUDA, "NCCL unavailable: %s", a->error.c_str());
    CK(cudaSetDevice(device));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    crm_comm* c = new crm_comm();
    c->rank = rank;
    c->world = world;
    if (const char* f = getenv("CRM_B200_NCCL_FUSED")) c->fused = f[0] != '0';
    ncclResult_t r = a->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        delete c;
        return crm_fail(CRM_ERR_CUDA, "ncclCommInitRank failed: %s", a->GetErrorString(r));
    }
    *out = c;
    return 0;
}
void crm_comm_destroy(crm_comm* c) { crm_comm_free(c); }

// The slab `s` (crm_sim_set_slab) steps with this communicator (borrowed, not owned).
int crm_sim_set_comm(crm_sim* s, crm_comm* c) {
    if (!s) return crm_fail(CRM_ERR_INVALID, "null sim");
    s->comm = c;
    return 0;
}

}  // extern "C"
