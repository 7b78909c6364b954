// mpsb_api.cu — the C ABI of libmps_b200.so (include/mps_b200.h): handle lifetime, state
// transfer, the step orchestration (reference MPS::stepForward, src/mps.cpp:111-130) and the
// inspection calls the parity tests use. No CPU fallback anywhere: every compute entry point
// needs a CUDA device and fails with MPSB_ERR_CUDA otherwise.
#include "mpsb_internal.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

namespace {

thread_local char g_create_error[512] = "";

#define CK(h, call)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            snprintf((h)->err, sizeof (h)->err, "%s:%d %s -> %s", __FILE__, __LINE__, #call,          \
                     cudaGetErrorString(e_));                                                         \
            (h)->sticky = MPSB_ERR_CUDA;                                                              \
            return MPSB_ERR_CUDA;                                                                     \
        }                                                                                             \
    } while (0)

template <class T>
cudaError_t dalloc(T** p, size_t count) {
    cudaError_t e = cudaMalloc((void**) p, sizeof(T) * std::max<size_t>(count, 1));
    if (e == cudaSuccess) e = cudaMemset(*p, 0, sizeof(T) * std::max<size_t>(count, 1));
    return e;
}

// Peer-visible arrays come out of ONE allocation per rank (PeerLinks::arena): every rank carves it in the same
// order with the same sizes, so a peer's copy of an array sits at the same offset of the peer's arena.
template <class T>
cudaError_t aalloc(mpsb_handle* h, T** p, size_t count) {
    size_t bytes = (sizeof(T) * std::max<size_t>(count, 1) + 255) & ~(size_t) 255;
    if (h->arena_used + bytes > h->arena_bytes) return cudaErrorMemoryAllocation;
    *p = reinterpret_cast<T*>(h->arena + h->arena_used);
    h->arena_used += bytes;
    return cudaSuccess;
}

void derive(Params& P) {
    const mpsb_settings& s = P.s;
    double D  = (double) s.dim;
    double len = s.re_max;
    // Bucket::Bucket (src/bucket.cpp:10-22)
    P.numX = (int) ((s.domain_max[0] - s.domain_min[0]) / len) + 3;
    P.numY = (int) ((s.domain_max[1] - s.domain_min[1]) / len) + 3;
    P.numZ = (int) ((s.domain_max[2] - s.domain_min[2]) / len) + 3;
    P.numCells = P.numX * P.numY * P.numZ;
    P.visc_coef = (s.kinematic_viscosity) * (2.0 * D) / (s.n0_viscosity * s.lambda_viscosity);
    P.beta_n0   = s.surface_threshold * s.n0_surface;
    P.mat_coef  = 2.0 * D / (s.lambda_pressure * s.n0_pressure);
    P.qc_term   = s.compressibility / (s.dt * s.dt);
    P.inv_dt2   = 1.0 / (s.dt * s.dt);
    P.src_std   = s.relaxation_coefficient * (1.0 / (s.dt * s.dt));
    P.src_ho0   = s.higher_order_gamma * (1.0 / (s.dt * s.dt));
    P.src_ho1   = (1.0 / (s.n0_pressure * s.dt)) * s.re_pressure;
    P.grad_coef = D / s.n0_pressure;
    P.inv_density = 1 / s.density;
}

int check_flags(mpsb_handle* h) {
    int f = h->host_stats.error_flags;
    if (f & 1) {
        snprintf(h->err, sizeof h->err, "neighbour-list or collision-candidate capacity exceeded (K=%d); "
                 "set MPSB_MAX_NEIGHBORS", h->P.K);
        return MPSB_ERR_CAPACITY;
    }
    if (f & 2) {
        snprintf(h->err, sizeof h->err, "a non-ghost particle left the padded cell grid (the reference would "
                 "index out of bounds here)");
        return MPSB_ERR_OUT_OF_GRID;
    }
    return MPSB_OK;
}

int fetch_stats(mpsb_handle* h) {
    const DevStats* src = h->P.stats;
    if (h->world > 1) { // every rank counted its own rows: combine (collective), leaving the per-rank struct untouched
        if (int rc = mpsb_comm_combine_stats(h)) return rc;
        src = h->P.stats_all + MPSB_MAX_RANKS;
    }
    CK(h, cudaMemcpyAsync(&h->host_stats, src, sizeof(DevStats), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    CK(h, cudaGetLastError());
    return MPSB_OK;
}

void fill_stats(mpsb_handle* h, mpsb_step_stats* st) {
    if (!st) return;
    const DevStats& d = h->host_stats;
    st->courant         = d.courant_bits_as_double;
    st->cg_rel_residual = d.cg_rel_residual;
    st->cg_iterations   = d.cg_iterations;
    st->n_active        = d.n_active;
    st->n_inner         = d.n_inner;
    st->n_surface       = d.n_surface;
    st->nnz             = d.nnz;
    st->n_ghosted       = d.n_ghosted;
    st->n_unconnected   = d.n_unconnected;
    st->n_collisions    = d.n_collisions;
    st->max_neighbors   = d.max_neighbors;
    st->ppe_columns_16bit = d.pad[0];
    st->solve_in_cluster  = d.pad[1] & 1;
    st->ppe_matrix_free   = (d.pad[1] >> 1) & 1;
    st->cg_single_reduction = (d.pad[1] >> 2) & 1;
    st->courant_exceeds_cfl = d.courant_bits_as_double > h->P.s.cfl_condition;
}

void build_lists(mpsb_handle* h, bool scatter_first = false) {
    mpsb_launch_cell_build(h, scatter_first);
    mpsb_launch_neighbors(h);
    h->have_lists = true;
}

} // namespace

extern "C" {

const char* mpsb_version(void) { return "mps_b200 0.1 sm_100a"; }

const char* mpsb_last_error(const mpsb_handle* h) { return h ? h->err : g_create_error; }

int64_t mpsb_num_particles(const mpsb_handle* h) { return h ? h->P.N : 0; }
int64_t mpsb_launch_count(const mpsb_handle* h) { return h ? h->launches : 0; }

int mpsb_create(const mpsb_settings* settings, int64_t n, const int32_t* type, const double* position,
                const double* velocity, int device, mpsb_handle** out) {
    if (!settings || !out || n < 0 || n > 0x7fffffff || (n > 0 && (!type || !position || !velocity))) {
        snprintf(g_create_error, sizeof g_create_error, "mpsb_create: bad argument");
        return MPSB_ERR_ARG;
    }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) {
        snprintf(g_create_error, sizeof g_create_error,
                 "mpsb_create: no usable CUDA device %d (%s); this library has no CPU path", device,
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count 0 or index out of range");
        return MPSB_ERR_CUDA;
    }
    mpsb_handle* h = new mpsb_handle();
    memset(h, 0, sizeof *h);
    h->device = device;
    h->world  = 1;
    Params& P = h->P;
    P.s = *settings;
    P.N = (int) n;
    P.Np = (int) ((n + 255) / 256 * 256); // whole 32-row slices, and a whole number of them per rank for 1/2/4/8 ranks
    derive(P);
    const char* envk = getenv("MPSB_MAX_NEIGHBORS");
    P.K = envk ? atoi(envk) : (settings->dim == 2 ? 64 : 384);
    P.K = (P.K + 3) & ~3; // slices are padded in groups of 4 slots
    if (P.K < 4 || P.numCells <= 0) {
        snprintf(g_create_error, sizeof g_create_error, "mpsb_create: bad neighbour capacity or cell grid");
        delete h;
        return MPSB_ERR_ARG;
    }
    h->sort_bits = 8;
    while (h->sort_bits < 32 && (1ull << h->sort_bits) <= (unsigned long long) P.numCells) h->sort_bits += 8;

#define CKC(call)                                                                                     \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            snprintf(g_create_error, sizeof g_create_error, "mpsb_create: %s -> %s", #call,           \
                     cudaGetErrorString(e_));                                                         \
            mpsb_destroy(h);                                                                          \
            return MPSB_ERR_CUDA;                                                                     \
        }                                                                                             \
    } while (0)

    CKC(cudaSetDevice(device));
    cudaDeviceProp prop;
    CKC(cudaGetDeviceProperties(&prop, device));
    h->num_sms = prop.multiProcessorCount;
    if (prop.major < 10) {
        snprintf(g_create_error, sizeof g_create_error, "mpsb_create: device %d is sm_%d%d; this library is built for sm_100a only",
                 device, prop.major, prop.minor);
        delete h;
        return MPSB_ERR_CUDA;
    }
    CKC(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    for (auto& ev : h->ev) CKC(cudaEventCreate(&ev));
    size_t N = (size_t) P.Np, KN = (size_t) P.K * N; // row arrays padded to whole slices
    Sorted& S = P.S;
    double** sd[] = {&S.x, &S.y, &S.z, &S.u, &S.v, &S.w, &S.u2, &S.v2, &S.w2, &S.ax, &S.ay, &S.az, &S.minp, &S.ratio};
    for (double** p : sd) CKC(dalloc(p, N));
    CKC(dalloc(&S.type, N)); CKC(dalloc(&S.id, N));
    Home& H = P.H;
    P.cand_cap = 1024;
    while ((size_t) P.cand_cap < N / 4) P.cand_cap <<= 1;
    P.crec_seg = std::min(P.cand_cap / MPSB_MAX_RANKS, 16384); // per-rank segment of the collision-record inbox (multi-GPU)
    // the arena: HOME, the halo fields S.bc / S.p, Dirichlet reach flags, collision records (see PeerLinks)
    h->arena_bytes = (size_t) 256 * 20 + N * (3 * 3 * sizeof(double) + 3 * sizeof(double) + 2 * sizeof(int) + sizeof(int) +
                                              sizeof(double) + sizeof(int) + sizeof(double)) +
                     (size_t) P.crec_seg * MPSB_MAX_RANKS * sizeof(CollisionRecord) + 4096;
    CKC(cudaMalloc((void**) &h->arena, h->arena_bytes));
    CKC(cudaMemset(h->arena, 0, h->arena_bytes));
    CKC(aalloc(h, &H.pos, 3 * N)); CKC(aalloc(h, &H.vel, 3 * N)); CKC(aalloc(h, &H.acc, 3 * N));
    CKC(aalloc(h, &H.p, N)); CKC(aalloc(h, &H.minp, N)); CKC(aalloc(h, &H.ratio, N));
    CKC(aalloc(h, &H.type, N)); CKC(aalloc(h, &H.bc, N));
    CKC(aalloc(h, &S.bc, N)); CKC(aalloc(h, &S.p, N)); CKC(aalloc(h, &P.reach, N)); CKC(aalloc(h, &P.ghost_ratio, N));
    CKC(aalloc(h, &P.crec, (size_t) P.crec_seg * MPSB_MAX_RANKS)); CKC(aalloc(h, &P.crec_count, MPSB_MAX_RANKS));
    P.L.arena = h->arena;
    CKC(dalloc(&H.inv, N)); CKC(dalloc(&H.inv_alt, N));
    CKC(dalloc(&P.keyrow, N)); CKC(dalloc(&P.ext, 1)); CKC(dalloc(&P.dflags, 2));
    CKC(dalloc(&P.cstate, (size_t) 6 * 2 * P.crec_seg * MPSB_MAX_RANKS)); CKC(dalloc(&P.stats_all, MPSB_MAX_RANKS + 1)); // [MAX] = the combined struct
    P.row_lo = 0;
    P.row_hi = P.Np;
    h->home_current = true;
    CKC(dalloc(&P.key, N)); CKC(dalloc(&P.key_alt, N)); CKC(dalloc(&P.val, N)); CKC(dalloc(&P.val_alt, N));
    CKC(dalloc(&P.cell_start, (size_t) P.numCells)); CKC(dalloc(&P.cell_end, (size_t) P.numCells));
    CKC(dalloc(&P.nbr_count, N)); CKC(dalloc(&P.nbr_idx, KN)); CKC(dalloc(&P.nbr_dist, KN));
    CKC(dalloc(&P.ell_count, N)); CKC(dalloc(&P.ell_width, N / 32 + 1)); CKC(dalloc(&P.ell_cmin, N / 32 + 1)); CKC(dalloc(&P.ell_cmax, N / 32 + 1)); CKC(dalloc(&P.ell_col, KN)); CKC(dalloc(&P.ell_val, KN));
    CKC(dalloc(&P.ell_col16, KN)); CKC(dalloc(&P.ell_wide, 1));
    CKC(dalloc(&P.diag, N)); CKC(dalloc(&P.rhs, N));
    CKC(dalloc(&P.cg_r, N)); CKC(dalloc(&P.cg_p, N)); CKC(dalloc(&P.cg_ap, N)); CKC(dalloc(&P.cg_invdiag, N));
    CKC(dalloc(&P.parent, N)); CKC(dalloc(&P.mark, N));
    CKC(dalloc(&P.cand, (size_t) P.cand_cap)); CKC(dalloc(&P.cand_count, 1));
    P.ev_cap = 4096;
    CKC(dalloc(&P.ev_ghost_id, (size_t) P.ev_cap)); CKC(dalloc(&P.ev_ghost_xyz, 3 * (size_t) P.ev_cap));
    CKC(dalloc(&P.ev_uncon_id, (size_t) P.ev_cap));
    CKC(dalloc(&P.stats, 1));
    P.n_active = &P.stats->n_active;
    h->block_hist_cap = 256 * (int) ((N + 2047) / 2048 + 1);
    CKC(dalloc(&h->block_hist, (size_t) h->block_hist_cap));
    h->scan_status_cap = h->block_hist_cap / 4096 + 2;
    CKC(dalloc(&h->scan_status, (size_t) h->scan_status_cap + 1)); // [cap] = the ticket counter
    if (mpsb_neighbors_configure(h) != 0 || mpsb_pcg_configure(h) != 0 || mpsb_pcg_configure_v1(h) != 0) {
        snprintf(g_create_error, sizeof g_create_error, "mpsb_create: PCG kernel cannot be made resident");
        mpsb_destroy(h);
        return MPSB_ERR_CUDA;
    }
    CKC(dalloc(&P.cg_partials, 8 * (size_t) h->cg_max_blocks + 8));
    CKC(dalloc(&P.cg_slice_pap, N / 32 + 1));
    if (h->cg_single_reduction) { CKC(dalloc(&P.cg_dir, N)); CKC(dalloc(&P.cg_s, N)); }
    CKC(dalloc(&P.cluster_done, 1));
    CKC(dalloc(&P.row_xy, N));
    mpsb_pcg_cluster_configure(h); // optional fast path for small systems; 0 = this device cannot co-schedule a cluster
    {
        const char* nc = getenv("MPSB_NO_CLUSTER_SOLVE");
        h->cl_disable  = nc && atoi(nc) ? 1 : 0;
        const char* ng = getenv("MPSB_NO_STEP_GRAPH");
        h->graph_enable = ng && atoi(ng) ? 0 : 1;
        h->n_graphs     = 0;
        const char* nt = getenv("MPSB_NEIGHBOR_TILES");
        h->neighbor_tiles = nt ? (atoi(nt) ? 1 : 0) : -1;
        const char* nb = getenv("MPSB_SMALL_BUILD_MAX_N");
        h->small_build_max_n = nb ? atoi(nb) : 2048;
        if (h->small_build_max_n > 32768) h->small_build_max_n = 32768;
    }
    // initial state: type/position/velocity from the caller, everything else 0 / Ignored (SURVEY B-1)
    CKC(cudaMemcpy(H.type, type, sizeof(int) * (size_t) n, cudaMemcpyHostToDevice));
    CKC(cudaMemcpy(H.pos, position, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice));
    CKC(cudaMemcpy(H.vel, velocity, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice));
    if (n > 0) build_lists(h);
    CKC(cudaStreamSynchronize(h->stream));
    CKC(cudaGetLastError());
#undef CKC
    *out = h;
    return MPSB_OK;
}

int mpsb_destroy(mpsb_handle* h) {
    if (!h) return MPSB_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    mpsb_comm_release(h);
    for (int k = 0; k < h->n_graphs; k++)
        if (h->graphs[k].exec) cudaGraphExecDestroy(h->graphs[k].exec);
    h->n_graphs = 0;
    Params& P = h->P;
    Sorted& S = P.S;
    Home& H   = P.H;
    void* ptrs[] = {S.x, S.y, S.z, S.u, S.v, S.w, S.u2, S.v2, S.w2, S.ax, S.ay, S.az, S.minp, S.ratio, S.type,
                    S.id, h->arena, H.inv, H.inv_alt, P.keyrow, P.ext, P.dflags, P.cstate, P.stats_all, P.key, P.key_alt,
                    P.val, P.val_alt, P.cell_start, P.cell_end, P.nbr_count, P.nbr_idx, P.nbr_dist, P.ell_count,
                    P.ell_width, P.ell_cmin, P.ell_cmax, P.ell_col, P.ell_col16, P.ell_wide, P.ell_val, P.diag, P.rhs, P.cg_r, P.cg_p, P.cg_ap, P.cg_invdiag, P.cg_partials, P.cg_slice_pap, P.cg_dir, P.cg_s, P.cluster_done, P.row_xy,
                    P.parent, P.mark, P.cand, P.cand_count, P.ev_ghost_id, P.ev_ghost_xyz, P.ev_uncon_id, P.stats,
                    h->block_hist, h->scan_status};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    for (auto& ev : h->ev)
        if (ev) cudaEventDestroy(ev);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return MPSB_OK;
}

int mpsb_upload(mpsb_handle* h, uint32_t mask, const mpsb_fields* f) {
    if (!h || !f) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    if (int rc = mpsb_comm_sync_home(h)) return rc; // several ranks: collective; fields not uploaded keep their latest values
    size_t N = (size_t) h->P.N;
    Home& H  = h->P.H;
    cudaStream_t st = h->stream;
    const bool types_only = (mask & MPSB_F_TYPE) && f->type && !((mask & MPSB_F_NUMBER_DENSITY_RATIO) && f->number_density_ratio);
    if (types_only && N > 0) mpsb_launch_ghost_mark(h); // see mpsb_launch_ghost_ratio_from_home
#define UP(bit, dst, src, bytes)                                                                  \
    if ((mask & (bit)) && (src)) CK(h, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st))
    UP(MPSB_F_TYPE, H.type, f->type, sizeof(int) * N);
    UP(MPSB_F_POSITION, H.pos, f->position, sizeof(double) * 3 * N);
    UP(MPSB_F_VELOCITY, H.vel, f->velocity, sizeof(double) * 3 * N);
    UP(MPSB_F_ACCEL, H.acc, f->acceleration, sizeof(double) * 3 * N);
    UP(MPSB_F_PRESSURE, H.p, f->pressure, sizeof(double) * N);
    UP(MPSB_F_MIN_PRESSURE, H.minp, f->minimum_pressure, sizeof(double) * N);
    UP(MPSB_F_NUMBER_DENSITY_RATIO, H.ratio, f->number_density_ratio, sizeof(double) * N);
    UP(MPSB_F_BOUNDARY_CONDITION, H.bc, f->boundary_condition, sizeof(int) * N);
#undef UP
    if ((mask & (MPSB_F_NUMBER_DENSITY_RATIO | MPSB_F_TYPE)) && N > 0) mpsb_launch_ghost_ratio_from_home(h, types_only);
    if (h->have_lists && N > 0) mpsb_launch_gather_all(h); // keep the sorted working set in step (every row, on every rank)
    h->home_current = true;
    CK(h, cudaStreamSynchronize(st));
    return MPSB_OK;
}

int mpsb_upload_device(mpsb_handle* h, const int32_t* d_type, const double* d_position, const double* d_velocity) {
    if (!h) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    if (int rc = mpsb_comm_sync_home(h)) return rc;
    size_t N = (size_t) h->P.N;
    Home& H  = h->P.H;
    if (d_type) CK(h, cudaMemcpyAsync(H.type, d_type, sizeof(int) * N, cudaMemcpyDeviceToDevice, h->stream));
    if (d_position) CK(h, cudaMemcpyAsync(H.pos, d_position, sizeof(double) * 3 * N, cudaMemcpyDeviceToDevice, h->stream));
    if (d_velocity) CK(h, cudaMemcpyAsync(H.vel, d_velocity, sizeof(double) * 3 * N, cudaMemcpyDeviceToDevice, h->stream));
    if (h->have_lists && N > 0) mpsb_launch_gather_all(h);
    h->home_current = true;
    CK(h, cudaStreamSynchronize(h->stream));
    return MPSB_OK;
}

int mpsb_download(mpsb_handle* h, uint32_t mask, const mpsb_fields* f) {
    if (!h || !f) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    if (int rc = mpsb_comm_sync_home(h)) return rc; // several ranks: collective (every rank calls download at the same point)
    size_t N = (size_t) h->P.N;
    Home& H  = h->P.H;
    cudaStream_t st = h->stream;
#define DN(bit, dst, src, bytes)                                                                  \
    if ((mask & (bit)) && (dst)) CK(h, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st))
    DN(MPSB_F_TYPE, f->type, H.type, sizeof(int) * N);
    DN(MPSB_F_POSITION, f->position, H.pos, sizeof(double) * 3 * N);
    DN(MPSB_F_VELOCITY, f->velocity, H.vel, sizeof(double) * 3 * N);
    DN(MPSB_F_ACCEL, f->acceleration, H.acc, sizeof(double) * 3 * N);
    DN(MPSB_F_PRESSURE, f->pressure, H.p, sizeof(double) * N);
    DN(MPSB_F_MIN_PRESSURE, f->minimum_pressure, H.minp, sizeof(double) * N);
    DN(MPSB_F_NUMBER_DENSITY_RATIO, f->number_density_ratio, H.ratio, sizeof(double) * N);
    DN(MPSB_F_BOUNDARY_CONDITION, f->boundary_condition, H.bc, sizeof(int) * N);
#undef DN
    CK(h, cudaStreamSynchronize(st));
    return MPSB_OK;
}

// ---- sharded state transfer (several ranks): each rank moves only ITS block of particle ids over PCIe ----------
// Block of rank g: ids [g, g + 1) * Np / W clipped to N (Np = N rounded up to whole 32-row slices x 8, the same split as
// the row blocks). The array pointers are whole-array pointers indexed by particle id, as in mpsb_upload / mpsb_download;
// only the block is read / written.
static void shard_block(const mpsb_handle* h, size_t* lo, size_t* hi) {
    const size_t N = (size_t) h->P.N, blk = (size_t) (h->P.Np / (h->world > 1 ? h->world : 1));
    *lo = std::min(N, blk * (size_t) (h->world > 1 ? h->rank : 0));
    *hi = std::min(N, *lo + blk);
}

int mpsb_shard_range(const mpsb_handle* h, int64_t* first, int64_t* last) {
    if (!h || !first || !last) return MPSB_ERR_ARG;
    size_t lo, hi;
    shard_block(h, &lo, &hi);
    *first = (int64_t) lo;
    *last  = (int64_t) hi;
    return MPSB_OK;
}

int mpsb_upload_shard(mpsb_handle* h, uint32_t mask, const mpsb_fields* f) {
    if (!h || !f) return MPSB_ERR_ARG;
    if (h->world <= 1) return mpsb_upload(h, mask, f);
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    if (int rc = mpsb_comm_sync_home(h)) return rc; // fields not uploaded keep their latest values
    size_t lo, hi;
    shard_block(h, &lo, &hi);
    const size_t n = hi - lo;
    Home& H = h->P.H;
    cudaStream_t st = h->stream;
    uint32_t present = 0;
    const bool types_only = (mask & MPSB_F_TYPE) && f->type && !((mask & MPSB_F_NUMBER_DENSITY_RATIO) && f->number_density_ratio);
    if (types_only && h->P.N > 0) mpsb_launch_ghost_mark(h); // see mpsb_launch_ghost_ratio_from_home
#define UP(bit, dst, src, per, T)                                                                                   \
    if ((mask & (bit)) && (src)) {                                                                                  \
        present |= (bit);                                                                                           \
        if (n) CK(h, cudaMemcpyAsync(dst + lo * (per), src + lo * (per), sizeof(T) * (per) * n, cudaMemcpyHostToDevice, st)); \
    }
    UP(MPSB_F_TYPE, H.type, f->type, 1, int);
    UP(MPSB_F_POSITION, H.pos, f->position, 3, double);
    UP(MPSB_F_VELOCITY, H.vel, f->velocity, 3, double);
    UP(MPSB_F_ACCEL, H.acc, f->acceleration, 3, double);
    UP(MPSB_F_PRESSURE, H.p, f->pressure, 1, double);
    UP(MPSB_F_MIN_PRESSURE, H.minp, f->minimum_pressure, 1, double);
    UP(MPSB_F_NUMBER_DENSITY_RATIO, H.ratio, f->number_density_ratio, 1, double);
    UP(MPSB_F_BOUNDARY_CONDITION, H.bc, f->boundary_condition, 1, int);
#undef UP
    if (int rc = mpsb_comm_allgather_home(h, present)) return rc; // the other ranks' blocks arrive over NVLink
    const size_t N = (size_t) h->P.N;
    if ((present & (MPSB_F_NUMBER_DENSITY_RATIO | MPSB_F_TYPE)) && N > 0) mpsb_launch_ghost_ratio_from_home(h, types_only);
    if (h->have_lists && N > 0) mpsb_launch_gather_all(h);
    h->home_current = true;
    CK(h, cudaStreamSynchronize(st));
    return MPSB_OK;
}

int mpsb_download_shard(mpsb_handle* h, uint32_t mask, const mpsb_fields* f) {
    if (!h || !f) return MPSB_ERR_ARG;
    if (h->world <= 1) return mpsb_download(h, mask, f);
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    if (int rc = mpsb_comm_sync_home(h)) return rc;
    size_t lo, hi;
    shard_block(h, &lo, &hi);
    const size_t n = hi - lo;
    Home& H = h->P.H;
    cudaStream_t st = h->stream;
#define DN(bit, dst, src, per, T)                                                                                   \
    if ((mask & (bit)) && (dst) && n) CK(h, cudaMemcpyAsync(dst + lo * (per), src + lo * (per), sizeof(T) * (per) * n, cudaMemcpyDeviceToHost, st))
    DN(MPSB_F_TYPE, f->type, H.type, 1, int);
    DN(MPSB_F_POSITION, f->position, H.pos, 3, double);
    DN(MPSB_F_VELOCITY, f->velocity, H.vel, 3, double);
    DN(MPSB_F_ACCEL, f->acceleration, H.acc, 3, double);
    DN(MPSB_F_PRESSURE, f->pressure, H.p, 1, double);
    DN(MPSB_F_MIN_PRESSURE, f->minimum_pressure, H.minp, 1, double);
    DN(MPSB_F_NUMBER_DENSITY_RATIO, f->number_density_ratio, H.ratio, 1, double);
    DN(MPSB_F_BOUNDARY_CONDITION, f->boundary_condition, H.bc, 1, int);
#undef DN
    CK(h, cudaStreamSynchronize(st));
    return MPSB_OK;
}

// The launch sequence of one step on h->stream. `timed`: bracket the phases with CUDA events (direct issue);
// inside a stream capture only the kernels and memsets are recorded.
static int issue_step(mpsb_handle* h, int want_number_density, bool timed) {
    cudaStream_t st = h->stream;
    cudaEvent_t* E  = h->ev;
#define EV(k)                                                                                     \
    do {                                                                                          \
        if (timed) CK(h, cudaEventRecord(E[k], st));                                              \
    } while (0)
    CK(h, cudaMemsetAsync(h->P.stats, 0, sizeof(DevStats), st));
    h->in_step = 1;
    EV(0);
    build_lists(h);                                  // setNeighbors
    EV(1);
    mpsb_launch_explicit(h);                         // calcGravity, calcViscosity, moveParticles
    EV(2);
    build_lists(h, true);                                  // setNeighbors
    EV(3);
    mpsb_launch_collision(h);                        // collision
    EV(4);
    build_lists(h, true);                                  // setNeighbors
    EV(5);
    mpsb_launch_flags_source(h);                     // calcPressure ...
    mpsb_launch_matrix(h);
    mpsb_launch_dirichlet(h);
    EV(6);
    mpsb_launch_pcg(h);                              // several ranks: the halo pressures travel in pressure_post
    EV(7);
    mpsb_launch_pressure_post(h);
    mpsb_launch_gradient_move(h);                    // calcPressureGradient, moveParticlesWithPressureGradient
    EV(8);
    if (want_number_density) {                       // setNumberDensity (:125-127)
        build_lists(h, true);
        mpsb_launch_number_density(h);
    }
    EV(9);
    mpsb_launch_courant(h);                          // getCourantNumber
    mpsb_launch_scatter_all(h);
    EV(10);
#undef EV
    h->in_step = 0;
    return h->sticky ? h->sticky : MPSB_OK; // a failed exchange (NCCL) inside a launcher is sticky
}

// Small cases are launch-bound (about 57 launches of a few microseconds each): the whole step is captured
// once per distinct pointer state (the velocity double buffer and the sort buffers swap inside a step) and
// replayed as ONE CUDA graph launch. Returns 1 if the step was issued through a graph, 0 if the caller must
// issue it directly, < 0 on a sticky error.
static int step_through_graph(mpsb_handle* h, int want) {
    if (!h->graph_enable || h->P.N > 32767 || h->world > 1) return 0;
    StepGraph* g = nullptr;
    for (int k = 0; k < h->n_graphs; k++)
        if (h->graphs[k].u == h->P.S.u && h->graphs[k].key == h->P.key && h->graphs[k].want == want) g = &h->graphs[k];
    if (!g) {
        if (h->n_graphs == MPSB_MAX_STEP_GRAPHS) return 0;
        const Params before = h->P;
        const int64_t launches0 = h->launches;
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            h->graph_enable = 0;
            return 0;
        }
        int rc = issue_step(h, want, false);
        cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
        cudaGraphExec_t exec = nullptr;
        if (rc == MPSB_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (rc != MPSB_OK || e != cudaSuccess || !exec) { // not capturable here: issue directly from now on
            cudaGetLastError();
            h->sticky       = 0;
            h->P            = before;
            h->launches     = launches0;
            h->graph_enable = 0;
            return 0;
        }
        g           = &h->graphs[h->n_graphs++];
        g->u        = before.S.u;
        g->key      = before.key;
        g->want     = want;
        g->exec     = exec;
        g->after    = h->P;
        g->launches = h->launches - launches0;
        h->P        = before;
        h->launches = launches0;
    }
    CK(h, cudaEventRecord(h->ev[0], h->stream));
    CK(h, cudaGraphLaunch(g->exec, h->stream));
    CK(h, cudaEventRecord(h->ev[10], h->stream));
    h->P = g->after;
    h->launches += g->launches;
    h->have_lists = true;
    return 1;
}

// MPS::stepForward (src/mps.cpp:111-130). HOME is current on entry and on exit.
int mpsb_step(mpsb_handle* h, int want_number_density, mpsb_step_stats* stats) {
    if (!h) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    if (stats) memset(stats, 0, sizeof *stats);
    if (h->P.N == 0) return MPSB_OK;
    CK(h, cudaSetDevice(h->device));
    cudaEvent_t* E = h->ev;
    const int graphed = step_through_graph(h, want_number_density ? 1 : 0);
    if (graphed < 0) return h->sticky;
    if (!graphed) {
        int rc0 = issue_step(h, want_number_density, true);
        if (rc0) return rc0;
    }
    int rc = fetch_stats(h);
    if (rc) return rc;
    fill_stats(h, stats);
    if (stats) {
        auto ms = [&](int a, int b) {
            float t = 0.f;
            cudaEventElapsedTime(&t, E[a], E[b]);
            return (double) t;
        };
        if (!graphed) {
            stats->ms_neighbors = ms(0, 1) + ms(2, 3) + ms(4, 5) + ms(8, 9);
            stats->ms_explicit  = ms(1, 2);
            stats->ms_collision = ms(3, 4);
            stats->ms_assemble  = ms(5, 6);
            stats->ms_solve     = ms(6, 7);
            stats->ms_correct   = ms(7, 8) + ms(9, 10);
        } else { // one graph launch: only the solve is stamped (device clock inside the PCG kernels)
            const DevStats& d = h->host_stats;
            stats->ms_solve = d.t_solve_end > d.t_solve_begin ? 1e-6 * (double) (d.t_solve_end - d.t_solve_begin) : 0.0;
        }
        stats->ms_total = ms(0, 10);
    }
    return check_flags(h);
}

int mpsb_run_phase(mpsb_handle* h, int phase, mpsb_step_stats* stats) {
    if (!h) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    if (stats) memset(stats, 0, sizeof *stats);
    if (h->P.N == 0) return MPSB_OK;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemsetAsync(&h->P.stats->error_flags, 0, sizeof(int), h->stream));
    switch (phase) {
    case MPSB_PH_NEIGHBORS:
        CK(h, cudaMemsetAsync(&h->P.stats->max_neighbors, 0, sizeof(int), h->stream));
        build_lists(h);
        break;
    case MPSB_PH_EXPLICIT:
        CK(h, cudaMemsetAsync(&h->P.stats->n_ghosted, 0, sizeof(int), h->stream));
        mpsb_launch_explicit(h);
        break;
    case MPSB_PH_COLLISION: mpsb_launch_collision(h); break;
    case MPSB_PH_FLAGS_SOURCE: mpsb_launch_flags_source(h); break;
    case MPSB_PH_MATRIX: mpsb_launch_matrix(h); break;
    case MPSB_PH_DIRICHLET: mpsb_launch_dirichlet(h); break;
    case MPSB_PH_SOLVE:
        mpsb_launch_pcg(h);
        break;
    case MPSB_PH_PRESSURE_POST: mpsb_launch_pressure_post(h); break;
    case MPSB_PH_GRADIENT_MOVE:
        CK(h, cudaMemsetAsync(&h->P.stats->n_ghosted, 0, sizeof(int), h->stream));
        mpsb_launch_gradient_move(h);
        break;
    case MPSB_PH_NUMBER_DENSITY:
        build_lists(h);
        mpsb_launch_number_density(h);
        break;
    case MPSB_PH_COURANT: mpsb_launch_courant(h); break;
    default: snprintf(h->err, sizeof h->err, "unknown phase %d", phase); return MPSB_ERR_ARG;
    }
    mpsb_launch_scatter_all(h);
    int rc = fetch_stats(h);
    if (rc) return rc;
    fill_stats(h, stats);
    return check_flags(h);
}

int mpsb_get_neighbors(mpsb_handle* h, int64_t* row_start, int32_t* ids, double* dist, int64_t capacity,
                       int64_t* total) {
    if (!h || !row_start || !total) return MPSB_ERR_ARG;
    if (h->world > 1) {
        snprintf(h->err, sizeof h->err, "mpsb_get_neighbors: inspection calls read one GPU's rows; not available on an attached handle");
        return MPSB_ERR_ARG;
    }
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    const Params& P = h->P;
    size_t N = (size_t) P.N;
    std::vector<int> cnt(N), inv(N), sid(N), type(N);
    CK(h, cudaMemcpy(cnt.data(), P.nbr_count, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(inv.data(), P.H.inv, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(sid.data(), P.S.id, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(type.data(), P.H.type, sizeof(int) * N, cudaMemcpyDeviceToHost));
    int kmax = 0;
    for (size_t s = 0; s < N; s++) kmax = std::max(kmax, cnt[s]);
    int64_t k = 0;
    for (size_t i = 0; i < N; i++) {
        row_start[i] = k;
        if (type[i] != MPSB_GHOST) k += cnt[inv[i]];
    }
    row_start[N] = k;
    *total       = k;
    if (!ids || !dist) return MPSB_OK;
    if (capacity < k) return MPSB_ERR_CAPACITY;
    std::vector<int> idx((size_t) P.K * P.Np);
    std::vector<double> dd((size_t) P.K * P.Np);
    if (kmax > 0) {
        CK(h, cudaMemcpy(idx.data(), P.nbr_idx, sizeof(int) * idx.size(), cudaMemcpyDeviceToHost));
        CK(h, cudaMemcpy(dd.data(), P.nbr_dist, sizeof(double) * dd.size(), cudaMemcpyDeviceToHost));
    }
    for (size_t i = 0; i < N; i++) {
        if (type[i] == MPSB_GHOST) continue;
        size_t s = (size_t) inv[i];
        for (int m = 0; m < cnt[s]; m++) {
            ids[row_start[i] + m]  = sid[idx[mpsb_sell(m, (int) s, P.K)]];
            dist[row_start[i] + m] = dd[mpsb_sell(m, (int) s, P.K)];
        }
    }
    return MPSB_OK;
}

int mpsb_get_ppe(mpsb_handle* h, int64_t* row_start, int32_t* col, double* val, double* rhs, int64_t capacity,
                 int64_t* nnz) {
    if (!h || !row_start || !nnz) return MPSB_ERR_ARG;
    if (h->world > 1) {
        snprintf(h->err, sizeof h->err, "mpsb_get_ppe: inspection calls read one GPU's rows; not available on an attached handle");
        return MPSB_ERR_ARG;
    }
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    const Params& P = h->P;
    size_t N = (size_t) P.N;
    std::vector<int> cnt(N), inv(N), sid(N), bc(N);
    std::vector<double> diag(N), b(N);
    CK(h, cudaMemcpy(cnt.data(), P.ell_count, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(inv.data(), P.H.inv, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(sid.data(), P.S.id, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(bc.data(), P.S.bc, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(diag.data(), P.diag, sizeof(double) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(b.data(), P.rhs, sizeof(double) * N, cudaMemcpyDeviceToHost));
    int kmax = 0;
    for (size_t s = 0; s < N; s++) kmax = std::max(kmax, cnt[s]);
    int64_t k = 0;
    for (size_t i = 0; i < N; i++) {
        row_start[i] = k;
        size_t s = (size_t) inv[i];
        if (bc[s] == MPSB_BC_INNER) k += cnt[s] + 1;
        if (rhs) rhs[i] = b[s];
    }
    row_start[N] = k;
    *nnz         = k;
    if (!col || !val) return MPSB_OK;
    if (capacity < k) return MPSB_ERR_CAPACITY;
    std::vector<int> ec((size_t) P.K * P.Np);
    std::vector<double> ev((size_t) P.K * P.Np);
    if (kmax > 0) {
        CK(h, cudaMemcpy(ec.data(), P.ell_col, sizeof(int) * ec.size(), cudaMemcpyDeviceToHost));
        CK(h, cudaMemcpy(ev.data(), P.ell_val, sizeof(double) * ev.size(), cudaMemcpyDeviceToHost));
    }
    std::vector<std::pair<int, double>> row;
    for (size_t i = 0; i < N; i++) {
        size_t s = (size_t) inv[i];
        if (bc[s] != MPSB_BC_INNER) continue;
        row.clear();
        for (int m = 0; m < cnt[s]; m++) row.push_back({sid[ec[mpsb_sell(m, (int) s, P.K)]], ev[mpsb_sell(m, (int) s, P.K)]});
        row.push_back({(int) i, diag[s]});
        std::sort(row.begin(), row.end(), [](const auto& a, const auto& c) { return a.first < c.first; });
        for (size_t m = 0; m < row.size(); m++) {
            col[row_start[i] + (int64_t) m] = row[m].first;
            val[row_start[i] + (int64_t) m] = row[m].second;
        }
    }
    return MPSB_OK;
}

int mpsb_debug_cg_cycles(const mpsb_handle* h, int64_t* out4) {
    if (!h || !out4) return MPSB_ERR_ARG;
    for (int k = 0; k < 4; k++) out4[k] = h->host_stats.cg_cycles[k];
    return MPSB_OK;
}

int mpsb_debug_cg_block_cycles(mpsb_handle* h, int64_t* out, int capacity_blocks, int* nblocks) {
    if (!h || !out || !nblocks) return MPSB_ERR_ARG;
    int mb = h->cg_max_blocks;
    *nblocks = mb;
    if (capacity_blocks < mb) return MPSB_ERR_CAPACITY;
    CK(h, cudaMemcpy(out, h->P.cg_partials + 3 * (size_t) mb, sizeof(int64_t) * 5 * (size_t) mb, cudaMemcpyDeviceToHost));
    return MPSB_OK;
}

// Diagnostic / parity: y = A x with the PPE of the last MATRIX/DIRICHLET phase, x and y indexed by particle id
// (entries of non-Inner particles are ignored / returned as 0). matrix_free = 0: the stored sliced-ELL stream;
// 1: the recomputation from positions that the solve uses in matrix-free mode (k_pcg.cu spmv_sweep_mf).
int mpsb_debug_spmv(mpsb_handle* h, const double* x, double* y, int matrix_free) {
    if (!h || !x || !y) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    if (h->world > 1) return MPSB_ERR_ARG;
    CK(h, cudaSetDevice(h->device));
    const Params& P = h->P;
    size_t N = (size_t) P.N;
    std::vector<int> inv(N), bc(N);
    CK(h, cudaStreamSynchronize(h->stream));
    CK(h, cudaMemcpy(inv.data(), P.H.inv, sizeof(int) * N, cudaMemcpyDeviceToHost));
    CK(h, cudaMemcpy(bc.data(), P.S.bc, sizeof(int) * N, cudaMemcpyDeviceToHost));
    std::vector<double> xs((size_t) P.Np, 0.0), ys((size_t) P.Np, 0.0);
    for (size_t i = 0; i < N; i++)
        if (bc[(size_t) inv[i]] == MPSB_BC_INNER) xs[(size_t) inv[i]] = x[i];
    CK(h, cudaMemcpy(P.cg_p, xs.data(), sizeof(double) * (size_t) P.Np, cudaMemcpyHostToDevice));
    CK(h, cudaMemset(P.cg_ap, 0, sizeof(double) * (size_t) P.Np));
    mpsb_launch_spmv_probe(h, matrix_free ? 1 : 0);
    CK(h, cudaStreamSynchronize(h->stream));
    CK(h, cudaGetLastError());
    CK(h, cudaMemcpy(ys.data(), P.cg_ap, sizeof(double) * (size_t) P.Np, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < N; i++) y[i] = bc[(size_t) inv[i]] == MPSB_BC_INNER ? ys[(size_t) inv[i]] : 0.0;
    return MPSB_OK;
}

void* mpsb_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, (size_t) (bytes ? bytes : 1), cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void mpsb_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int mpsb_set_stream(mpsb_handle* h, void* cuda_stream) {
    if (!h) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    h->stream = cuda_stream ? (cudaStream_t) cuda_stream : h->own_stream;
    return MPSB_OK;
}

int mpsb_events(mpsb_handle* h, int kind, int32_t* ids, double* xyz, int capacity, int* count) {
    if (!h || !count) return MPSB_ERR_ARG;
    if (h->sticky) return h->sticky;
    CK(h, cudaSetDevice(h->device));
    int n = kind == 0 ? h->host_stats.n_ghosted : h->host_stats.n_unconnected;
    if (h->world > 1) {
        // host_stats holds the SUM over the ranks (combine_stats_kernel); the event buffers hold this rank's events only,
        // indexed by this rank's own counter, which the combine left untouched in *P.stats
        DevStats own;
        CK(h, cudaStreamSynchronize(h->stream));
        CK(h, cudaMemcpy(&own, h->P.stats, sizeof own, cudaMemcpyDeviceToHost));
        n = kind == 0 ? own.n_ghosted : own.n_unconnected;
    }
    n     = std::min(n, h->P.ev_cap);
    *count = n;
    int m = std::min(n, capacity);
    if (m > 0 && ids)
        CK(h, cudaMemcpy(ids, kind == 0 ? h->P.ev_ghost_id : h->P.ev_uncon_id, sizeof(int) * m, cudaMemcpyDeviceToHost));
    if (m > 0 && xyz && kind == 0)
        CK(h, cudaMemcpy(xyz, h->P.ev_ghost_xyz, sizeof(double) * 3 * m, cudaMemcpyDeviceToHost));
    return MPSB_OK;
}

} // extern "C"

