// mpsb_internal.cuh — device-state layout and kernel entry points shared by the .cu files of
// libmps_b200.so (sm_100a only). See DESIGN.md §3 for the layout rationale.
//
// Two views of the particle state live in HBM:
//   * HOME  (id order, packed xyz)  — what upload/download move; authoritative between API calls.
//   * SORTED (cell-major, then id; SoA) — the working set of every kernel of the step. Built by
//     the cell-list build (key -> radix sort -> gather); all neighbour indices are SORTED indices.
// Neighbour lists and the PPE matrix are SLICED ELL (slice = the 32 rows of one warp, fixed slice
// pitch K slots): entry k of row s lives at [((s / 32) * K + k) * 32 + s % 32]. A warp reads 32
// consecutive values per slot (coalesced) AND the slots of a slice are contiguous, so a whole slice
// of the matrix is one contiguous block that the PCG kernel streams with 1-D TMA bulk copies.
#pragma once
#include "../../include/mps_b200.h"

#include <cuda_runtime.h>
#include <stdint.h>

#define MPSB_GHOST_KEY 0xFFFFFFFFu

// Device-side per-step counters, copied to the host once per step.
struct DevStats {
    double courant_bits_as_double; // written through atomicMax on the bit pattern
    double cg_rel_residual;
    double rhs_norm2;
    int cg_iterations;
    int n_active;
    int n_inner;
    int n_surface;
    int nnz;
    int n_ghosted;
    int n_unconnected;
    int n_collisions;
    int max_neighbors;
    int error_flags; // bit0 capacity, bit1 out-of-grid, bit2 event overflow
    int pad[2]; // pad[0]: the last solve streamed 16-bit column offsets
    long long cg_cycles[4]; // CTA 0's SM cycles in sweep A / B / C / reductions+grid barriers (diagnostic)
    unsigned long long t_solve_begin, t_solve_end; // %globaltimer (ns) at the start / end of the solve kernel that did the work
};

// Sorted-order SoA working set (all arrays have N entries).
struct Sorted {
    double *x, *y, *z;       // position
    double *u, *v, *w;       // velocity
    double *u2, *v2, *w2;    // velocity double buffer (explicit step reads neighbours' old velocity)
    double *ax, *ay, *az;    // acceleration
    double *p, *minp, *ratio;
    int *type;               // ParticleType values
    int *bc;                 // BoundaryCondition values
    int *id;                 // perm: sorted index -> particle id
};

struct Home {
    double *pos, *vel, *acc; // packed xyz, 3N
    double *p, *minp, *ratio;
    int *type, *bc;
    int *inv;                // particle id -> sorted index (valid after a build)
    int *inv_alt;            // multi-GPU: the perm inverse being built while `inv` (the old one) is still read
};

// Multi-GPU view of the PPE solve (mpsb_comm.cu). One process per GPU; rank g owns the
// contiguous block of 32-row slices [g, g+1) * Np/32/world of the SORTED order, i.e. a y-slab of the cell
// grid. Peers write straight into each other's HBM over NVLink/NVSwitch (buffers mapped with CUDA IPC):
// reduction partials and epoch flags into the INBOX, and the halo rows of the search direction p into
// the peer's own p vector at their natural row index. world == 1: unused.
#define MPSB_MAX_RANKS 8
struct PeerLinks {
    int rank, world;
    // my inbox (one allocation; peers hold a mapping of it)
    double* red;                 // [2][MPSB_MAX_RANKS][4]: per-rank partials, double-buffered by epoch parity
    unsigned long long* flag;    // [MPSB_MAX_RANKS]: exchange epoch each rank has reached
    // local only (same allocation, never written by peers)
    unsigned long long* arrive;  // CTA arrival counter of the running solve (zeroed by the host per launch)
    unsigned long long* epoch;   // exchanges completed so far, monotonic across solves
    // peers' inboxes and p vectors (entry [rank] is the local one)
    double* peer_red[MPSB_MAX_RANKS];
    unsigned long long* peer_flag[MPSB_MAX_RANKS];
    double* peer_p[MPSB_MAX_RANKS];
    // the peer-visible ARENA (slab decomposition of the particle phases): one allocation per rank
    // holding every array a peer stores into — HOME, the halo fields S.bc / S.p, the Dirichlet reach flags and the
    // collision-record inbox. All ranks carve it identically, so peer q's copy of a local arena pointer x is
    // peer_arena[q] + (x - arena).
    char* arena;
    char* peer_arena[MPSB_MAX_RANKS];
};

// Per-rank row extents, identical on every rank (computed from the replicated sorted keys by ext_kernel):
// rows [own_lo, own_hi) are the rank's fixed block of the SORTED order; rows [ext_lo, ext_hi) additionally cover
// its HALO: every row whose cell is within the 3x3(x3) stencil of a cell holding one of its active rows.
struct RankExtents {
    int own_lo[MPSB_MAX_RANKS], own_hi[MPSB_MAX_RANKS];
    int ext_lo[MPSB_MAX_RANKS], ext_hi[MPSB_MAX_RANKS];
};

// Everything a kernel needs, passed by value.
struct Params {
    mpsb_settings s;
    int N;        // particle slots (ghosts included)
    int Np;       // N rounded up to a whole number of 32-row slices (allocation size of row arrays)
    int K;        // neighbour-list capacity per particle (slice pitch in slots)
    int numX, numY, numZ, numCells;
    // derived coefficients, computed on the host with the reference's operation order
    double visc_coef;    // nu * (2*dim) / (n0_v * lambda_v)          src/mps.cpp:182
    double beta_n0;      // beta * n0_s                               src/mps.cpp:271
    double mat_coef;     // 2*dim / (lambda_p * n0_p)                 src/mps.cpp:343
    double qc_term;      // alpha / (dt*dt)                           src/mps.cpp:354
    double inv_dt2;      // 1 / (dt*dt)
    double src_std;      // gamma_relax * (1/(dt*dt))                 src/mps.cpp:313
    double src_ho0;      // gamma_ho * (1/(dt*dt))                    src/mps.cpp:290
    double src_ho1;      // (1/(n0_p*dt)) * re_p                      src/mps.cpp:304
    double grad_coef;    // dim / n0_p                                src/mps.cpp:530
    double inv_density;  // 1 / density                               src/particle.cpp:22
    Sorted S;
    Home H;
    // cell list
    unsigned *key, *key_alt;   // N each (sorted keys end up in `key`)
    int *val, *val_alt;        // N each (ids)
    int *cell_start, *cell_end; // numCells each
    // neighbour ELL
    int *nbr_count;            // N
    int *nbr_idx;              // K*N
    double *nbr_dist;          // K*N
    // PPE: ELL off-diagonals + diagonal + rhs; CG vectors
    int *ell_count;            // Np  true off-diagonal count of the row
    int *ell_width;            // Np/32  padded slot count of the slice (multiple of 4; padding = (col row, val 0))
    int *ell_cmin, *ell_cmax;  // Np/32  smallest / largest column of the slice (halo extent of a row block)
    int *ell_col;              // K*Np
    short *ell_col16;          // K*Np  column - first row of the slice, when every such offset fits 16 bits
    int *ell_wide;             // 1     set by the assembly when some offset does not fit: the solve then streams ell_col
    double *ell_val;           // K*Np
    double *diag, *rhs;        // N
    double *cg_r, *cg_p, *cg_ap, *cg_invdiag; // N
    double *cg_partials;       // 8 * max_blocks + 1: 3 reduction rows, 5 diagnostic rows, 2 slice-queue counters
    double *cg_dir, *cg_s;     // N each: search direction and s = A d of the single-reduction recurrence (MPSB_CG_SINGLE_REDUCTION)
    double *cg_slice_pap;      // Np/32 + 1: p.Ap of each 32-row slice (deterministic two-level sum)
    double2 *row_xy;           // Np: (x, y) of every SORTED row as one 16-byte record, written by the assembly for the
                               // matrix-free sweep A (k_pcg.cu): off-diagonals are recomputed from positions
    int *cluster_done;         // 1     set by the cluster solve (k_pcg_cluster.cu) when it did the solve
    // Dirichlet connectivity
    int *parent;               // N
    int *mark;                 // N
    // collision candidates
    unsigned long long *cand;  // cand_cap
    int cand_cap;
    int *cand_count;
    // events
    int *ev_ghost_id; double *ev_ghost_xyz; int *ev_uncon_id; int ev_cap;
    DevStats* stats;
    int* n_active; // == &stats->n_active
    PeerLinks L;
    // ---- slab decomposition (world > 1; world == 1: row_lo = 0, row_hi = Np and none of the rest is touched) ----
    int row_lo, row_hi;        // the SORTED rows this rank owns: every per-particle kernel covers exactly these
    unsigned* keyrow;          // Np: cell key of every row of the OLD order (own block computed, the rest all-gathered)
    RankExtents* ext;          // device copy of the extents of the current order
    int* reach;                // Np (arena): Dirichlet marks of boundary rows, stored by their owners
    double* ghost_ratio;       // N (arena), by particle id: n(re_p) / n0 of the list a particle had when it turned Ghost. The
                               // reference never clears a Ghost's list (setNeighbors skips it, src/mps.cpp:590-593), so its
                               // setNumberDensity keeps reporting that sum (:630-640) — and Exporter::toVtu writes it
    int* dflags;               // 2: Dirichlet fixpoint: [0] a mark changed this round, [1] an own Inner row is still unmarked
    struct CollisionRecord* crec;      // arena: W segments of crec_seg records, segment q written by rank q
    int* crec_count;                   // arena: W counts
    int crec_seg;
    double* cstate;            // 6 * 2 * W * crec_seg: compact (x, y, z, u, v, w) of the particles the records touch
    DevStats* stats_all;       // W: every rank's counters of the step (all-gathered, then combined into *stats)
};

// One collision candidate as it travels between ranks: the key fixes the reference's visiting order
// ((id_i << 32) | list slot), the rest is the pre-collision snapshot the replay starts from.
struct CollisionRecord {
    unsigned long long key;
    int row_i, row_j;
    int type_j, pad;
    double dist;
    double xi[3], ui[3], xj[3], uj[3];
};

#ifdef __CUDACC__
template <class T>
__device__ __forceinline__ T* mpsb_peer(const Params& P, T* local, int q) {
    return reinterpret_cast<T*>(P.L.peer_arena[q] + (reinterpret_cast<char*>(local) - P.L.arena));
}
#endif

// One captured step (mpsb_api.cu): valid for the pointer state it was captured in.
#define MPSB_MAX_STEP_GRAPHS 8
struct StepGraph {
    const void* u;      // P.S.u at capture (velocity double buffer parity)
    const void* key;    // P.key at capture (sort buffer parity)
    int want;           // want_number_density
    cudaGraphExec_t exec;
    Params after;       // P after the step's pointer swaps
    int64_t launches;   // kernel launches inside the graph
};

struct mpsb_handle {
    Params P;
    int device;
    cudaStream_t stream;
    cudaStream_t own_stream;
    cudaEvent_t ev[16];
    int sort_bits;
    int cg_max_blocks;      // cooperative grid limit for the PCG kernel
    int cg_threads;
    int cg_smem;            // dynamic shared memory of the PCG kernel
    int cg1_threads, cg1_smem; // the same for the 16-warp x 16-slot build (k_pcg_g16.cu), used for row blocks <= cg_small_rows
    int cg_small_rows;      // MPSB_CG_SMALL_ROWS (default 1 300 000)
    int cg_iteration_cap;   // diagnostic: stop the solve after this many iterations (0 = Eigen's 2N)
    int cg_l2_resident_kb;  // L2 policy of the matrix stream (k_pcg.cu L2Plan): KiB kept with evict_last, the rest evict_first; 0 = off
    int cg_l2_max_rows;     // ... applied to row blocks of at most this many rows
    int cg_single_reduction; // MPSB_CG_SINGLE_REDUCTION=1: Chronopoulos-Gear recurrence in the grid solve (one reduction per iteration)
    int cg_matrix_free;     // sweep A recomputes the off-diagonals from positions instead of streaming them (2-D; MPSB_MATRIX_FREE)
    int num_sms;
    int cl_size, cl_smem;   // cluster solve for small systems (k_pcg_cluster.cu): cluster size (0 = unavailable), dynamic smem
    int cl_launched;        // the cluster kernel has run on this handle (*cluster_done may be set)
    int cl_disable;         // MPSB_NO_CLUSTER_SOLVE=1: always use the grid kernel (tests)
    int neighbor_tiles;     // MPSB_NEIGHBOR_TILES: 1 / 0 force the tiled (TMA-staged) / per-thread neighbour build, -1 = default (per-thread)
    int small_build_max_n;  // single-launch cell build up to this many particles (MPSB_SMALL_BUILD_MAX_N; 0 = never)
    int in_step;            // inside mpsb_step: DevStats was cleared as a whole, launchers skip their own clears
    int graph_enable;       // small cases: replay the step as one CUDA graph (MPSB_NO_STEP_GRAPH=1 disables)
    int n_graphs;
    StepGraph graphs[MPSB_MAX_STEP_GRAPHS];
    int64_t launches;
    bool have_lists;        // a neighbour build has happened since the last position upload
    DevStats host_stats;
    char err[512];
    int sticky;
    // scratch for the radix sort
    int* block_hist; int block_hist_cap;
    unsigned long long* scan_status; int scan_status_cap; // look-back scan: one status word per scan tile, then the ticket counter
    int replicated;         // > 0: mpsb_attach_comm was called with this world size but the case stays replicated (too small to shard)
    void* comm; int rank, world; // multi-GPU (mpsb_comm.cu): MpsbComm*, mirrors P.L.rank / P.L.world
    char* arena; size_t arena_bytes, arena_used; // peer-visible arrays (PeerLinks::arena)
    bool home_current;      // multi-GPU: HOME holds the latest state of EVERY particle (after an upload / a collective sync)
    // host staging for inspection calls
};

// ---- launchers (each file owns its kernels) ----------------------------------------------------
void mpsb_launch_cell_build(mpsb_handle* h, bool scatter_first); // k_cells.cu: [scatter->]key->sort->ranges->gather
void mpsb_launch_gather_all(mpsb_handle* h);            // HOME -> SORTED with the current perm
void mpsb_launch_scatter_all(mpsb_handle* h);           // SORTED -> HOME
void mpsb_launch_neighbors(mpsb_handle* h);             // k_neighbors.cu
int mpsb_neighbors_configure(mpsb_handle* h);           // shared-memory opt-in of the tiled build
void mpsb_launch_number_density(mpsb_handle* h);
void mpsb_launch_ghost_ratio_from_home(mpsb_handle* h, bool only_new); // after an upload: a Ghost's recorded sum is its uploaded ratio (only_new: only particles the upload turned Ghost)
void mpsb_launch_ghost_mark(mpsb_handle* h); // before a types-only upload: remember who is Ghost already (P.mark)
void mpsb_launch_explicit(mpsb_handle* h);              // k_explicit.cu
void mpsb_launch_collision(mpsb_handle* h);
void mpsb_launch_gradient_move(mpsb_handle* h);
void mpsb_launch_courant(mpsb_handle* h);
void mpsb_launch_flags_source(mpsb_handle* h);          // k_pressure.cu
void mpsb_launch_matrix(mpsb_handle* h);
void mpsb_launch_dirichlet(mpsb_handle* h);
void mpsb_launch_pressure_post(mpsb_handle* h);
void mpsb_launch_pcg(mpsb_handle* h);                   // k_pcg.cu
int mpsb_pcg_configure(mpsb_handle* h);
int mpsb_pcg_configure_v1(mpsb_handle* h);              // k_pcg_g16.cu: the second ring geometry of the same kernel
void mpsb_launch_pcg_v1(mpsb_handle* h);
void mpsb_launch_spmv_probe(mpsb_handle* h, int matrix_free); // diagnostic: one sweep A, cg_p -> cg_ap
int mpsb_pcg_cluster_configure(mpsb_handle* h);         // k_pcg_cluster.cu
int mpsb_launch_pcg_cluster(mpsb_handle* h);
void mpsb_comm_release(mpsb_handle* h);
// ---- slab decomposition (world > 1) ----
int mpsb_comm_allgather_keyrow(mpsb_handle* h);         // mpsb_comm.cu: NCCL all-gather of the own-block cell keys
int mpsb_comm_sync_home(mpsb_handle* h);                // all-gather every SORTED field, then SORTED -> HOME for all rows
int mpsb_comm_allgather_home(mpsb_handle* h, uint32_t field_mask); // id blocks of the HOME fields in the mask -> whole arrays on every rank
int mpsb_comm_combine_stats(mpsb_handle* h);            // all-gather DevStats, combine into *P.stats
int mpsb_comm_allreduce_dflags(mpsb_handle* h, int* host2); // Dirichlet fixpoint: OR of the two flags over the ranks
void mpsb_launch_xbarrier(mpsb_handle* h);              // k_halo.cu: cross-rank barrier on the stream (epoch flags in the inboxes)
void mpsb_launch_halo_push_bc(mpsb_handle* h);          // own rows inside a peer's extent -> that peer's S.bc
void mpsb_launch_halo_push_p(mpsb_handle* h);           // ... -> that peer's S.p
void mpsb_launch_scatter_all_rows(mpsb_handle* h);      // SORTED -> HOME for every row (after mpsb_comm_sync_home's gathers)
static inline int mpsb_own_rows(const mpsb_handle* h) { // rows a per-particle kernel of this rank covers (< N)
    int hi = h->P.row_hi < h->P.N ? h->P.row_hi : h->P.N;
    return hi > h->P.row_lo ? hi - h->P.row_lo : 0;
}
static inline int mpsb_own_grid(const mpsb_handle* h, int block) {
    int r = mpsb_own_rows(h);
    return r > 0 ? (r + block - 1) / block : 1;
}

#define MPSB_LAUNCH(h, kernel, grid, block, smem, ...)                                   \
    do {                                                                                 \
        kernel<<<(grid), (block), (smem), (h)->stream>>>(__VA_ARGS__);                   \
        (h)->launches++;                                                                 \
    } while (0)

static inline int mpsb_div_up(int a, int b) { return (a + b - 1) / b; }

// sliced-ELL index of slot k of row s for slice pitch K
__host__ __device__ __forceinline__ size_t mpsb_sell(int k, int s, int K) {
    return ((size_t) (s >> 5) * (size_t) K + (size_t) k) * 32u + (size_t) (s & 31);
}
#define SELL(k, s) mpsb_sell((k), (s), P.K)

#ifdef __CUDACC__
// weight (reference src/mps.cpp:9-16); every operation individually rounded (no FMA: the
// library is compiled with -fmad=false)
__device__ __forceinline__ double mpsb_weight(double dist, double re) {
    return dist < re ? (re / dist) - 1.0 : 0.0;
}

// Walk the neighbour list of SORTED row s in list order, NB entries at a time: the (distance, index) loads
// of a batch are issued together, then everything `fetch(j)` gathers for them, and only then `use(k, d, j,
// rec)` runs for k ascending — so sums keep the reference's order, but a thread waits for two memory round
// trips per batch instead of three per neighbour (load d -> branch -> load j -> gather). The per-particle
// kernels are latency chains, not bandwidth: this is what sets their time at every size. Loads of the
// surplus slots of the last batch are clamped to the last entry (always inside the list) and ignored.
template <int NB, class Rec, class Fetch, class Use>
__device__ __forceinline__ void mpsb_walk_list(const Params& P, int s, int cnt, Fetch fetch, Use use) {
    for (int k0 = 0; k0 < cnt; k0 += NB) {
        double d[NB];
        int j[NB];
        Rec r[NB];
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int k = min(k0 + b, cnt - 1);
            d[b] = P.nbr_dist[SELL(k, s)];
            j[b] = P.nbr_idx[SELL(k, s)];
        }
#pragma unroll
        for (int b = 0; b < NB; b++) r[b] = fetch(j[b]);
#pragma unroll
        for (int b = 0; b < NB; b++)
            if (k0 + b < cnt) use(k0 + b, d[b], j[b], r[b]);
    }
}
struct MpsbNoRec {};

// A particle turns Ghost: record n(re_p) / n0 of the list it has NOW — the list the reference keeps for it from
// then on (see Params::ghost_ratio) — on every rank (a Ghost's row ends up in the tail block, owned by another rank).
__device__ __forceinline__ void mpsb_record_ghost_ratio(const Params& P, int s) {
    const double re = P.s.re_pressure;
    const int cnt = P.nbr_count[s];
    double n = 0.0;
    for (int k = 0; k < cnt; k++) {
        const double d = P.nbr_dist[SELL(k, s)];
        if (d < re) n += mpsb_weight(d, re);
    }
    const double v = n / P.s.n0_pressure;
    const int id = P.S.id[s];
    if (P.L.world <= 1) {
        P.ghost_ratio[id] = v;
    } else {
        for (int q = 0; q < P.L.world; q++) mpsb_peer(P, P.ghost_ratio, q)[id] = v;
    }
}
#endif
