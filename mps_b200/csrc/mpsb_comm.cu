// mpsb_comm.cu — multi-GPU attachment of a handle (one process per GPU).
//
// NCCL (over NVLink 5 / NVSwitch) carries the bulk, once-per-step transfers: the bootstrap all-gather of
// the CUDA IPC handles and the all-gather of the solved pressure row blocks. The Krylov iteration itself
// never calls NCCL: its scalar reductions, barriers and halo rows move through peer-mapped HBM inside the
// persistent PCG kernel (k_pcg.cu, pcg_kernel<true>), which is what the buffers exchanged here are for.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy torch already loaded when the host is
// Python, the system one for the C++ host), so a single-GPU user needs no NCCL at all.
// The reference has no counterpart: it is a single process.
#include "mpsb_internal.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {

struct NcclApi {
    void* dl = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*)                                                        = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                                 = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t)                                                           = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t)   = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t)                                                       = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi api;
    if (api.dl) return api;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        api.dl = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.dl) break;
    }
    if (!api.dl) return api;
    api.GetUniqueId    = (decltype(api.GetUniqueId)) dlsym(api.dl, "ncclGetUniqueId");
    api.CommInitRank   = (decltype(api.CommInitRank)) dlsym(api.dl, "ncclCommInitRank");
    api.CommDestroy    = (decltype(api.CommDestroy)) dlsym(api.dl, "ncclCommDestroy");
    api.AllGather      = (decltype(api.AllGather)) dlsym(api.dl, "ncclAllGather");
    api.AllReduce      = (decltype(api.AllReduce)) dlsym(api.dl, "ncclAllReduce");
    api.GetErrorString = (decltype(api.GetErrorString)) dlsym(api.dl, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.GetErrorString;
    return api;
}

struct MpsbComm {
    ncclComm_t comm = nullptr;
    void* inbox     = nullptr;                 // local; peers map it
    void* staging   = nullptr;                 // device buffer for the handle all-gather
    void* peer_inbox[MPSB_MAX_RANKS] = {};     // cudaIpcOpenMemHandle mappings (null for the own rank)
    void* peer_p[MPSB_MAX_RANKS]     = {};
    void* peer_arena[MPSB_MAX_RANKS] = {};
};

constexpr size_t INBOX_BYTES = 4096;
constexpr size_t INBOX_RED   = 0;                                          // double[2][MAX][4]
constexpr size_t INBOX_FLAG  = sizeof(double) * 2 * MPSB_MAX_RANKS * 4;    // u64[MAX]
constexpr size_t INBOX_ARRIVE = INBOX_FLAG + sizeof(unsigned long long) * MPSB_MAX_RANKS;
constexpr size_t INBOX_EPOCH  = INBOX_ARRIVE + 64;

struct HandlePair {
    cudaIpcMemHandle_t inbox;
    cudaIpcMemHandle_t p;
    cudaIpcMemHandle_t arena;
};

int fail(mpsb_handle* h, const char* what, const char* detail) {
    snprintf(h->err, sizeof h->err, "mpsb_attach_comm: %s (%s)", what, detail ? detail : "");
    h->sticky = MPSB_ERR_COMM;
    return MPSB_ERR_COMM;
}

} // namespace

extern "C" int mpsb_comm_unique_id(void* out128) {
    if (!out128) return MPSB_ERR_ARG;
    NcclApi& N = nccl();
    if (!N.ok) return MPSB_ERR_COMM;
    ncclUniqueId id;
    if (N.GetUniqueId(&id) != ncclSuccess) return MPSB_ERR_COMM;
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
    return MPSB_OK;
}

extern "C" int mpsb_attach_comm(mpsb_handle* h, int rank, int world_size, const void* id128) {
    if (!h) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    if (world_size == 1) return MPSB_OK;
    Params& P = h->P;
    // Cases too small to shard profitably stay REPLICATED: every rank keeps running the whole single-GPU step (cluster
    // solve, graph replay, no exchange of any kind) and holds the whole, bit-identical state. dambreak_fine (6 146
    // particles) costs 1.8 ms per step on one GPU and 6.4 ms sharded over two (profiles/r1_v19_*): 236 CG iterations x
    // a few cross-GPU latencies each buy nothing at that size. MPSB_MIN_SLAB_PARTICLES=0 forces sharding (tests do).
    {
        const char* ms = getenv("MPSB_MIN_SLAB_PARTICLES");
        const int min_slab = ms ? atoi(ms) : 32768;
        if (P.N < min_slab) {
            h->replicated = world_size;
            return MPSB_OK;
        }
    }
    if (!id128 || rank < 0 || rank >= world_size || world_size > MPSB_MAX_RANKS || (P.Np / 32) % world_size != 0 ||
        h->comm) {
        snprintf(h->err, sizeof h->err, "mpsb_attach_comm: bad rank/world (%d/%d; supported world sizes: 1, 2, 4, 8) "
                 "or already attached", rank, world_size);
        return MPSB_ERR_ARG;
    }
    NcclApi& N = nccl();
    if (!N.ok) return fail(h, "libnccl.so.2 not found", dlerror());
#define CKA(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(h, #call, cudaGetErrorString(e_));                      \
    } while (0)
#define CKN(call)                                                                                  \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) return fail(h, #call, N.GetErrorString(r_));                        \
    } while (0)
    CKA(cudaSetDevice(h->device));
    MpsbComm* C = new MpsbComm();
    h->comm     = C;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    CKN(N.CommInitRank(&C->comm, world_size, id, rank));
    CKA(cudaMalloc(&C->inbox, INBOX_BYTES));
    CKA(cudaMemset(C->inbox, 0, INBOX_BYTES));
    CKA(cudaMalloc(&C->staging, sizeof(HandlePair) * world_size));
    HandlePair mine;
    CKA(cudaIpcGetMemHandle(&mine.inbox, C->inbox));
    CKA(cudaIpcGetMemHandle(&mine.p, P.cg_p));
    CKA(cudaIpcGetMemHandle(&mine.arena, h->arena));
    CKA(cudaMemcpy((char*) C->staging + sizeof(HandlePair) * rank, &mine, sizeof mine, cudaMemcpyHostToDevice));
    CKA(cudaStreamSynchronize(h->stream));
    // the all-gather also orders every rank's inbox memset before any peer's first remote store
    CKN(N.AllGather((char*) C->staging + sizeof(HandlePair) * rank, C->staging, sizeof(HandlePair), ncclChar, C->comm,
                    h->stream));
    CKA(cudaStreamSynchronize(h->stream));
    HandlePair all[MPSB_MAX_RANKS];
    CKA(cudaMemcpy(all, C->staging, sizeof(HandlePair) * world_size, cudaMemcpyDeviceToHost));
    PeerLinks& L = P.L;
    memset(&L, 0, sizeof L);
    L.arena = h->arena;
    for (int q = 0; q < world_size; q++) {
        void *ib = C->inbox, *pp = P.cg_p, *ar = h->arena;
        if (q != rank) {
            CKA(cudaIpcOpenMemHandle(&C->peer_inbox[q], all[q].inbox, cudaIpcMemLazyEnablePeerAccess));
            CKA(cudaIpcOpenMemHandle(&C->peer_p[q], all[q].p, cudaIpcMemLazyEnablePeerAccess));
            CKA(cudaIpcOpenMemHandle(&C->peer_arena[q], all[q].arena, cudaIpcMemLazyEnablePeerAccess));
            ib = C->peer_inbox[q];
            pp = C->peer_p[q];
            ar = C->peer_arena[q];
        }
        L.peer_arena[q] = (char*) ar;
        L.peer_red[q]  = (double*) ((char*) ib + INBOX_RED);
        L.peer_flag[q] = (unsigned long long*) ((char*) ib + INBOX_FLAG);
        L.peer_p[q]    = (double*) pp;
    }
    L.red    = (double*) ((char*) C->inbox + INBOX_RED);
    L.flag   = (unsigned long long*) ((char*) C->inbox + INBOX_FLAG);
    L.arrive = (unsigned long long*) ((char*) C->inbox + INBOX_ARRIVE);
    L.epoch  = (unsigned long long*) ((char*) C->inbox + INBOX_EPOCH);
    L.rank   = rank;
    L.world  = world_size;
    h->rank  = rank;
    h->world = world_size;
    // slab decomposition: this rank's fixed block of SORTED rows. The state uploaded so far is complete on every
    // rank (mpsb_create took the whole arrays and built the lists for all rows), so HOME is current.
    P.row_lo = rank * (P.Np / world_size);
    P.row_hi = P.row_lo + P.Np / world_size;
    h->home_current = true;
#undef CKA
#undef CKN
    return MPSB_OK;
}

namespace {
int nccl_fail(mpsb_handle* h, const char* what, ncclResult_t r) {
    snprintf(h->err, sizeof h->err, "%s: %s", what, nccl().GetErrorString(r));
    h->sticky = MPSB_ERR_COMM;
    return MPSB_ERR_COMM;
}
template <class T>
int gather_rows(mpsb_handle* h, T* field, size_t per_row, const char* what) {
    MpsbComm* C = (MpsbComm*) h->comm;
    size_t bytes = (size_t) (h->P.Np / h->world) * per_row * sizeof(T);
    ncclResult_t r = nccl().AllGather((char*) field + bytes * h->rank, field, bytes, ncclChar, C->comm, h->stream);
    return r == ncclSuccess ? MPSB_OK : nccl_fail(h, what, r);
}
} // namespace

// The cell keys every rank computed for its own rows of the OLD order -> the whole array on every rank (4 B per
// particle: the sort of (key, id) pairs is replicated, everything else of the build is not).
int mpsb_comm_allgather_keyrow(mpsb_handle* h) { return gather_rows(h, h->P.keyrow, 1, "ncclAllGather(keyrow)"); }

// Make HOME current on every rank (before a download or a partial upload): all-gather the row blocks of every SORTED
// field in place, then SORTED -> HOME for all rows. Collective: every rank calls it at the same point.
int mpsb_comm_sync_home(mpsb_handle* h) {
    if (h->world <= 1 || h->home_current) return MPSB_OK;
    Sorted& S = h->P.S;
    double* d[] = {S.x, S.y, S.z, S.u, S.v, S.w, S.ax, S.ay, S.az, S.p, S.minp, S.ratio};
    for (double* f : d)
        if (int rc = gather_rows(h, f, 1, "ncclAllGather(state)")) return rc;
    if (int rc = gather_rows(h, S.type, 1, "ncclAllGather(type)")) return rc;
    if (int rc = gather_rows(h, S.bc, 1, "ncclAllGather(bc)")) return rc;
    mpsb_launch_scatter_all_rows(h);
    h->home_current = true;
    return MPSB_OK;
}

// Sharded upload (mpsb_upload_shard): every rank copied its block of particle ids into its HOME arrays; all-gather
// the blocks of the uploaded fields over NVLink so that HOME is whole on every rank again.
int mpsb_comm_allgather_home(mpsb_handle* h, uint32_t mask) {
    if (h->world <= 1) return MPSB_OK;
    Home& H = h->P.H;
#define GA(bit, ptr, per)                                                                          \
    if (mask & (bit))                                                                              \
        if (int rc = gather_rows(h, ptr, per, "ncclAllGather(home)")) return rc
    GA(MPSB_F_TYPE, H.type, 1);
    GA(MPSB_F_POSITION, H.pos, 3);
    GA(MPSB_F_VELOCITY, H.vel, 3);
    GA(MPSB_F_ACCEL, H.acc, 3);
    GA(MPSB_F_PRESSURE, H.p, 1);
    GA(MPSB_F_MIN_PRESSURE, H.minp, 1);
    GA(MPSB_F_NUMBER_DENSITY_RATIO, H.ratio, 1);
    GA(MPSB_F_BOUNDARY_CONDITION, H.bc, 1);
#undef GA
    return MPSB_OK;
}

// Counters of the step: every rank counted its own rows; all-gather the structs and combine (k_halo.cu).
void mpsb_launch_combine_stats(mpsb_handle* h);
int mpsb_comm_combine_stats(mpsb_handle* h) {
    if (h->world <= 1) return MPSB_OK;
    MpsbComm* C = (MpsbComm*) h->comm;
    Params& P   = h->P;
    cudaMemcpyAsync(P.stats_all + h->rank, P.stats, sizeof(DevStats), cudaMemcpyDeviceToDevice, h->stream);
    ncclResult_t r = nccl().AllGather(P.stats_all + h->rank, P.stats_all, sizeof(DevStats), ncclChar, C->comm, h->stream);
    if (r != ncclSuccess) return nccl_fail(h, "ncclAllGather(stats)", r);
    mpsb_launch_combine_stats(h);
    return MPSB_OK;
}

// Dirichlet fixpoint (k_pressure.cu): OR of the two flags over the ranks, read back by the host that drives the loop.
int mpsb_comm_allreduce_dflags(mpsb_handle* h, int* host2) {
    MpsbComm* C = (MpsbComm*) h->comm;
    ncclResult_t r = nccl().AllReduce(h->P.dflags, h->P.dflags, 2, ncclInt, ncclMax, C->comm, h->stream);
    if (r != ncclSuccess) return nccl_fail(h, "ncclAllReduce(dirichlet flags)", r);
    if (cudaMemcpyAsync(host2, h->P.dflags, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
        cudaStreamSynchronize(h->stream) != cudaSuccess) {
        snprintf(h->err, sizeof h->err, "Dirichlet fixpoint: flag read-back failed");
        h->sticky = MPSB_ERR_CUDA;
        return MPSB_ERR_CUDA;
    }
    return MPSB_OK;
}

void mpsb_comm_release(mpsb_handle* h) {
    if (!h || !h->comm) return;
    MpsbComm* C = (MpsbComm*) h->comm;
    for (int q = 0; q < MPSB_MAX_RANKS; q++) {
        if (C->peer_inbox[q]) cudaIpcCloseMemHandle(C->peer_inbox[q]);
        if (C->peer_p[q]) cudaIpcCloseMemHandle(C->peer_p[q]);
        if (C->peer_arena[q]) cudaIpcCloseMemHandle(C->peer_arena[q]);
    }
    if (C->comm && nccl().ok) nccl().CommDestroy(C->comm);
    if (C->staging) cudaFree(C->staging);
    if (C->inbox) cudaFree(C->inbox);
    delete C;
    h->comm = nullptr;
}
