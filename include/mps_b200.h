/* mps_b200.h — C ABI of the B200-native MPS time step (libmps_b200.so).
 *
 * This is the drop-in boundary: it sits INSIDE the reference's `MPS` class, underneath its
 * unchanged public members (reference src/mps.hpp:61-70), and replaces the private phase
 * functions of reference src/mps.cpp. Plain pointers and sizes only; no C++/torch types.
 * All arrays are host pointers unless a function says "device". Every function returns 0 on
 * success and a negative mpsb_status otherwise; it never throws. There is NO CPU fallback:
 * without a CUDA device every compute entry point fails with MPSB_ERR_CUDA.
 *
 * Environment switches read by the library (diagnostics and tests; none is needed in normal use):
 *   MPSB_MAX_NEIGHBORS=<K>       neighbour-list / PPE row capacity (default 64 in 2-D, 384 in 3-D)
 *   MPSB_NO_CLUSTER_SOLVE=1      never use the small-system cluster solve (csrc/k_pcg_cluster.cu)
 *   MPSB_NO_STEP_GRAPH=1         issue the step's launches directly instead of replaying a CUDA graph
 *   MPSB_SMALL_BUILD_MAX_N=<n>   build cell lists in one single-CTA launch up to n particles (default 2048, at most
 *                                32 768; 0 = always the multi-launch radix sort)
 *   MPSB_NEIGHBOR_TILES=0|1      neighbour build: per-thread walk through L1/L2 (0) or one CTA per 16 cells of a cell row with the
 *                                3 / 9 neighbouring cell rows staged in shared memory by TMA bulk copies (1); default 0
 *                                (the tiled build is bit-identical but measured 1.8x slower at 4 M particles)
 *   MPSB_FORCE_WIDE_COLUMNS=1    stream int32 PPE columns even when 16-bit offsets fit
 *   MPSB_CG_ITERATION_CAP=<n>    stop the solve after n iterations (profiling a bounded solve)
 *   MPSB_MATRIX_FREE=0|1         sweep A of the grid solve streams the stored off-diagonals (0) or recomputes them from
 *                                positions (1; 2-D cases on ONE GPU only: with an attached multi-GPU group the switch is
 *                                ignored and ppe_matrix_free reports 0, because positions of halo columns are not exchanged)
 *   MPSB_CG_SINGLE_REDUCTION=1   grid solve with the Chronopoulos-Gear recurrence: one global reduction and two grid-wide
 *                                synchronisations per iteration instead of two and four; same Krylov space and stopping rule,
 *                                different rounding (held to 1e-10 / +-2 % iterations, not bit-identical); never the default
 *   MPSB_CG_SMALL_ROWS=<rows>    row blocks of up to <rows> rows (default 1 300 000; twice that per rank on several GPUs) run the
 *                                16-warp x 16-slot build of the grid solve instead of the 20 x 12 one (0 = never)
 *   MPSB_L2_RESIDENT_KB=<kb>     L2 policy of the solve's matrix stream: <kb> KiB travel as evict_last, the rest as evict_first
 *                                (default 16384; 0 = no policy), for row blocks of at most MPSB_L2_MAX_ROWS rows (2 500 000)
 * and by the host mirror (mps_b200/host): MPSB_DEVICE=<ordinal>, MPSB_CSV_FLOAT=gcc5.
 *
 * Particle arrays are SoA, indexed by particle id (the reference assumes id == index,
 * src/mps.cpp:599-617, src/bucket.cpp:36-50). Vectors are packed xyz (3 doubles per particle).
 */
#ifndef MPS_B200_H
#define MPS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Values of reference `enum class ParticleType` (src/particle.hpp:5) */
enum { MPSB_GHOST = -1, MPSB_FLUID = 0, MPSB_WALL = 1, MPSB_DUMMY_WALL = 2 };
/* Values of reference `enum class BoundaryCondition` (src/particle.hpp:19) */
enum { MPSB_BC_IGNORED = 0, MPSB_BC_SURFACE = 1, MPSB_BC_INNER = 2 };

typedef enum mpsb_status {
    MPSB_OK               = 0,
    MPSB_ERR_ARG          = -1, /* bad argument */
    MPSB_ERR_CUDA         = -2, /* CUDA runtime failure or no device (sticky on the handle) */
    MPSB_ERR_CAPACITY     = -3, /* neighbour-list / candidate capacity exceeded */
    MPSB_ERR_OUT_OF_GRID  = -4, /* a non-ghost particle left the padded cell grid (reference would index OOB, SURVEY B-8) */
    MPSB_ERR_COMM         = -5  /* multi-GPU exchange failure */
} mpsb_status;

/* Flattened reference `Settings` (src/settings.hpp:35-86) plus the three `RefValues`
 * (src/mps.hpp:10-20, src/mps.cpp:132-163), which the host computes with the reference's
 * loop order so n0/lambda are bit-identical. */
typedef struct mpsb_settings {
    int32_t dim;                      /* settings.dim */
    int32_t quasi_compressibility_on; /* settings.quasiCompressibility.on */
    int32_t higher_order_source_on;   /* settings.higherOrderSourceTerm.on */
    int32_t reserved0;
    double particle_distance;         /* l0 */
    double dt;
    double cfl_condition;
    double domain_min[3];
    double domain_max[3];
    double density;
    double kinematic_viscosity;
    double gravity[3];
    double re_pressure;               /* effectiveRadius.pressure (already multiplied by l0) */
    double re_viscosity;
    double re_surface;
    double re_max;
    double surface_threshold;         /* thresholdForSurfaceDetection (beta) */
    double compressibility;           /* quasiCompressibility.compressibility (alpha) */
    double higher_order_gamma;
    double relaxation_coefficient;    /* relaxationCoefficientForPressure */
    double collision_distance;        /* already multiplied by l0 */
    double restitution;
    double n0_pressure, lambda_pressure;
    double n0_viscosity, lambda_viscosity;
    double n0_surface, lambda_surface;
} mpsb_settings;

/* Per-step report. Times are CUDA-event milliseconds on the handle's stream. */
typedef struct mpsb_step_stats {
    double courant;          /* return value of reference getCourantNumber (src/mps.cpp:547-571) */
    double cg_rel_residual;  /* sqrt(|r|^2/|b|^2) at exit, Eigen's error() */
    double ms_total;
    double ms_neighbors;     /* all neighbour-list builds of the step */
    double ms_explicit;      /* gravity + viscosity + move */
    double ms_collision;
    double ms_assemble;      /* flags, source term, matrix, Dirichlet connectivity */
    double ms_solve;         /* the PCG solve (reference solvePoissonEquation) */
    double ms_correct;       /* clamp, minimum pressure, gradient, correction, Courant */
    int32_t cg_iterations;
    int32_t n_active;        /* non-ghost particles at the end of the step */
    int32_t n_inner;         /* PPE unknowns (BoundaryCondition::Inner) */
    int32_t n_surface;
    int32_t nnz;             /* stored PPE entries incl. diagonal */
    int32_t n_ghosted;       /* particles turned Ghost in this step (debugLogCount += ) */
    int32_t n_unconnected;   /* Inner particles with no Dirichlet connection (debugLogCount += ) */
    int32_t n_collisions;    /* collision candidates replayed (dist < collisionDistance) */
    int32_t max_neighbors;   /* longest neighbour list seen in the step */
    int32_t courant_exceeds_cfl; /* 1 if courant > cfl_condition (debugLogCount += ) */
    int32_t ppe_columns_16bit;   /* 1: the solve streamed 16-bit column offsets (10 B per entry), 0: int32 columns (12 B) */
    int32_t solve_in_cluster;    /* 1: the small-system cluster kernel did the solve (k_pcg_cluster.cu), 0: the grid kernel */
    int32_t ppe_matrix_free;     /* 1: sweep A of the solve recomputed the off-diagonals from positions (2 B per entry streamed) */
    int32_t cg_single_reduction; /* 1: the solve ran the single-reduction (Chronopoulos-Gear) recurrence (MPSB_CG_SINGLE_REDUCTION=1) */
} mpsb_step_stats;

typedef struct mpsb_handle mpsb_handle;

/* Phases for mpsb_run_phase (parity tests drive the step piecewise, teacher-forced).
 * Each names the reference function(s) it replaces. */
typedef enum mpsb_phase {
    MPSB_PH_NEIGHBORS       = 0, /* setNeighbors            src/mps.cpp:585-619 + bucket.cpp:24-51 */
    MPSB_PH_EXPLICIT        = 1, /* calcGravity+calcViscosity+moveParticles   :165-215 */
    MPSB_PH_COLLISION       = 2, /* collision               :217-250 */
    MPSB_PH_FLAGS_SOURCE    = 3, /* setBoundaryCondition+setSourceTerm        :262-320 */
    MPSB_PH_MATRIX          = 4, /* setMatrix               :322-360 */
    MPSB_PH_DIRICHLET       = 5, /* ensureDirichletBoundaryConnection         :362-415 */
    MPSB_PH_SOLVE           = 6, /* solvePoissonEquation    :417-470 */
    MPSB_PH_PRESSURE_POST   = 7, /* removeNegativePressure+setMinimumPressure :472-503 */
    MPSB_PH_GRADIENT_MOVE   = 8, /* calcPressureGradient+moveParticlesWithPressureGradient :505-545 */
    MPSB_PH_NUMBER_DENSITY  = 9, /* setNumberDensity (includes its own setNeighbors) :627-641 */
    MPSB_PH_COURANT         = 10 /* getCourantNumber        :547-571 */
} mpsb_phase;

/* Field mask bits for upload/download. */
enum {
    MPSB_F_TYPE     = 1 << 0, /* int32[n]   */
    MPSB_F_POSITION = 1 << 1, /* double[3n] */
    MPSB_F_VELOCITY = 1 << 2, /* double[3n] */
    MPSB_F_ACCEL    = 1 << 3, /* double[3n] */
    MPSB_F_PRESSURE = 1 << 4, /* double[n]  */
    MPSB_F_MIN_PRESSURE = 1 << 5, /* double[n] */
    MPSB_F_NUMBER_DENSITY_RATIO = 1 << 6, /* double[n] */
    MPSB_F_BOUNDARY_CONDITION   = 1 << 7, /* int32[n] */
    MPSB_F_ALL = 0xFF
};

typedef struct mpsb_fields {
    int32_t* type;
    double* position;
    double* velocity;
    double* acceleration;
    double* pressure;
    double* minimum_pressure;
    double* number_density_ratio;
    int32_t* boundary_condition;
} mpsb_fields;

/* --- lifetime ------------------------------------------------------------------------- */

/* Replaces MPS::loadInitialState's resize/Bucket construction (src/mps.cpp:96-104): creates the
 * device state for n particles (id == index) on CUDA device `device` and uploads type, position,
 * velocity. Pressure, min pressure, ratio start at 0 and boundary condition at Ignored
 * (the reference leaves them uninitialised, SURVEY B-1). Does NOT run setNumberDensity —
 * call mpsb_run_phase(MPSB_PH_NUMBER_DENSITY) as loadInitialState does (:106). */
int mpsb_create(const mpsb_settings* settings, int64_t n, const int32_t* type,
                const double* position, const double* velocity, int device, mpsb_handle** out);
int mpsb_destroy(mpsb_handle* h);
/* Last error text for the handle (or for a failed mpsb_create when h == NULL). */
const char* mpsb_last_error(const mpsb_handle* h);
/* Library identification: "mps_b200 <version> sm_100a". */
const char* mpsb_version(void);

/* --- the hot path --------------------------------------------------------------------- */

/* Replaces MPS::stepForward (src/mps.cpp:111-130). `want_number_density` = isTimeToExport ||
 * higherOrderSourceTerm.on (:125-127). Synchronises the stream before returning. */
int mpsb_step(mpsb_handle* h, int want_number_density, mpsb_step_stats* stats);
/* Run one phase of the step on the current device state (tests / teacher forcing). */
int mpsb_run_phase(mpsb_handle* h, int phase, mpsb_step_stats* stats);

/* --- state transfer (host mirror of `std::vector<Particle>`, src/particle.hpp:21-46) ----- */
/* With an attached multi-GPU group both calls are COLLECTIVE: every rank calls them at the same point of its program
 * (the ranks first bring HOME up to date from each other's row blocks); each rank passes / receives whole arrays. */
int mpsb_upload(mpsb_handle* h, uint32_t field_mask, const mpsb_fields* src);
int mpsb_download(mpsb_handle* h, uint32_t field_mask, const mpsb_fields* dst);
int64_t mpsb_num_particles(const mpsb_handle* h);

/* Sharded variants for several ranks (both COLLECTIVE): the pointers are still whole arrays indexed by particle id, but rank g
 * reads / writes only ITS block of ids [first, last) (mpsb_shard_range: [g, g+1) * Np / W clipped to n) — per-rank PCIe
 * traffic is 1/W of the state. After mpsb_upload_shard the blocks are all-gathered between the GPUs over NVLink, so every
 * rank's device state is whole, exactly as after mpsb_upload. With one rank they are mpsb_upload / mpsb_download. */
int mpsb_upload_shard(mpsb_handle* h, uint32_t field_mask, const mpsb_fields* src);
int mpsb_download_shard(mpsb_handle* h, uint32_t field_mask, const mpsb_fields* dst);
int mpsb_shard_range(const mpsb_handle* h, int64_t* first, int64_t* last);

/* Page-locked host memory for the buffers passed to mpsb_upload / mpsb_download (cudaHostAlloc: the copies then run
 * as DMA at full PCIe rate with no staging copy inside the driver). Returns NULL on failure; free with mpsb_host_free. */
void* mpsb_host_alloc(uint64_t bytes);
void mpsb_host_free(void* p);

/* --- inspection (parity tests) -------------------------------------------------------------
 * Neighbour lists of the last build as CSR by particle id, entries in the reference's order
 * (src/mps.cpp:599-617). Call with ids == NULL to get only row_start (n+1 entries) and size. */
int mpsb_get_neighbors(mpsb_handle* h, int64_t* row_start, int32_t* ids, double* dist,
                       int64_t capacity, int64_t* total);
/* The PPE of the last MATRIX/DIRICHLET phase as CSR over particle ids (n rows, empty rows for
 * non-Inner particles like the reference's N x N matrix, src/mps.cpp:99-100), columns ascending
 * within a row as after Eigen setFromTriplets, plus the source term. */
int mpsb_get_ppe(mpsb_handle* h, int64_t* row_start, int32_t* col, double* val, double* rhs,
                 int64_t capacity, int64_t* nnz);
/* Ids of particles behind the last step's debug-log events (src/mps.cpp:405-410, 655-663).
 * kind 0 = ghosted (xyz = position at the time, 3 doubles each), 1 = Dirichlet-unconnected.
 * With an attached multi-GPU group *count and the ids are those of the CALLING rank's rows only (the counters in
 * mpsb_step_stats are sums over the ranks); the union over the ranks is the single-GPU list. */
int mpsb_events(mpsb_handle* h, int kind, int32_t* ids, double* xyz, int capacity, int* count);

/* --- multi-GPU (one process per GPU; slab decomposition by rows of the cell-major order) ---------------------
 * Attach an exchange group before the first step. From then on every rank runs every phase of the step on its own
 * block of rows only (halo particles, particle migration, halo flags / pressures, collision records and the Krylov
 * exchanges travel over NVLink peer memory; NCCL carries the 4-byte cell keys, the counters and state syncs).
 * mpsb_step / mpsb_run_phase / mpsb_upload / mpsb_download are then collective; the inspection calls
 * (mpsb_get_neighbors, mpsb_get_ppe, mpsb_debug_spmv) are single-GPU only; mpsb_events lists the calling rank's events. `nccl_unique_id` is the 128-byte ncclUniqueId
 * produced on rank 0 and distributed by the host (torch.distributed / a launcher).
 * Cases below MPSB_MIN_SLAB_PARTICLES (default 32 768) are not sharded: attach succeeds, every rank keeps running the whole
 * single-GPU step on the whole state (replicas, bit-identical, no communication) — a few thousand particles are a latency
 * problem that cross-GPU exchanges only make worse. MPSB_MIN_SLAB_PARTICLES=0 shards every size.
 * Collision capacity of a sharded handle: each rank may record at most min(C / 8, 16 384) colliding pairs of its own rows per
 * step, C = the single-GPU candidate capacity (the power of two >= max(1024, n / 4)); the records of all ranks are then
 * replayed in the reference's serial order by one thread on every rank. More pairs than that on one rank end the step with
 * MPSB_ERR_CAPACITY on every rank (the same state may still fit the single-GPU path, whose limit is C). */
int mpsb_comm_unique_id(void* out128);
int mpsb_attach_comm(mpsb_handle* h, int rank, int world_size, const void* nccl_unique_id128);

/* --- device-resident entry points (bench: inputs already in HBM) ------------------------------
 * Same as mpsb_upload for position/velocity/type but from DEVICE pointers on the handle's device. */
int mpsb_upload_device(mpsb_handle* h, const int32_t* d_type, const double* d_position,
                       const double* d_velocity);
/* Run all subsequent work of this handle on the caller's CUDA stream (a cudaStream_t passed as
 * void*; NULL restores the handle's own stream), so the host can bracket steps with its own events. */
int mpsb_set_stream(mpsb_handle* h, void* cuda_stream);
/* Diagnostic: SM cycles CTA 0 spent in the last solve's sweep A (SpMV), sweep B, sweep C, and in the
 * reductions + grid barriers (4 values). */
int mpsb_debug_cg_cycles(const mpsb_handle* h, int64_t* out4);
/* Diagnostic: the same four counters for every CTA of the last solve plus its SM id, as 5 rows of
 * *nblocks int64 (row-major: A, B, C, sync, smid). */
int mpsb_debug_cg_block_cycles(mpsb_handle* h, int64_t* out, int capacity_blocks, int* nblocks);
/* Diagnostic / parity: y = A x with the PPE of the last MATRIX/DIRICHLET phase, x and y host vectors indexed by
 * particle id (entries of non-Inner particles ignored / returned 0). matrix_free = 0 applies the stored entries,
 * 1 the recomputation from positions the solve uses in matrix-free mode. Single GPU only. */
int mpsb_debug_spmv(mpsb_handle* h, const double* x, double* y, int matrix_free);
/* Counters since create: kernel launches issued by this library on this handle. */
int64_t mpsb_launch_count(const mpsb_handle* h);

/* --- for callers that must not allocate ----------------------------------------------------------------------
 * No counterpart in the reference (it has no device code and no allocator to avoid); the value computed is the one
 * of getCourantNumber (src/mps.cpp:547-571), bit-identical to mpsb_run_phase with MPSB_PH_COURANT.
 *
 * Everything above works on a handle, and a handle is made of cudaMalloc'ed arrays. This one call is for programs that
 * hold a particle state in device memory of their own and cannot let a library allocate when they call it: a host that
 * gave all device memory to its own pool, a caller inside a section where cudaMalloc's implicit device-wide
 * synchronisation would stall its other streams, a watchdog that only wants to know whether a live state still obeys the
 * CFL limit. It performs no cudaMalloc / cudaFree and calls no allocating routine, and it needs no handle.
 *
 * Inputs and the output are DEVICE pointers on CUDA device `device`: d_type int32[n], d_velocity packed xyz double[3n],
 * d_courant one double. The kernel is queued on `cuda_stream` (a cudaStream_t passed as void*; NULL = the default
 * stream) and the call returns WITHOUT synchronising: *d_courant is valid once that stream has reached this point.
 * Only particles of type MPSB_FLUID count; n = 0 gives 0.0.
 *
 * The kernel's scratch memory (one partial maximum per thread block + a ticket counter) is a fixed array inside the
 * library image, MPSB_STATIC_SCRATCH_BYTES on every GPU the library is loaded on, held until the process ends. Hence:
 *   - n may not exceed MPSB_STATIC_MAX_PARTICLES; a larger n returns MPSB_ERR_ARG, launches nothing and says so in error_text;
 *   - all calls on one GPU share the array, and it keeps the previous call's contents. The call itself clears what has to
 *     start at zero, so calls that follow one another on the GPU are independent; calls that may OVERLAP on the GPU
 *     (two streams with no ordering between them, two host threads) race on it — serialise them;
 *   - the first call on each device looks up the array's device address (once per device); make that first call before,
 *     not during, any stream capture.
 * error_text (may be NULL) receives a message of at most error_capacity bytes when the return value is not MPSB_OK. */
#define MPSB_STATIC_MAX_PARTICLES 67108864LL
#define MPSB_STATIC_SCRATCH_BYTES 262400
int mpsb_courant_static(const int32_t* d_type, const double* d_velocity, int64_t n, double dt, double particle_distance,
                        int device, void* cuda_stream, double* d_courant, char* error_text, int error_capacity);

#ifdef __cplusplus
}
#endif
#endif /* MPS_B200_H */
