// k_pcg.cu — the pressure Poisson solve: Jacobi-preconditioned conjugate gradient as ONE persistent
// cooperative kernel (one launch per solve, one CTA per SM; grid-wide barriers between the three
// sweeps of an iteration; no host round trips).
//
// Replaces reference MPS::solvePoissonEquation (src/mps.cpp:417-470), i.e. Eigen 3.4.0
// ConjugateGradient<SparseMatrix<double>, Lower, DiagonalPreconditioner> with all defaults:
//   conjugate_gradient()  Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:28-91
//   x0 = 0, tol = DBL_EPSILON on |r|/|b| with the recursively updated r, maxIters = 2 * cols
//                         IterativeSolverBase.h:281-284, 400-413
//   invdiag = 1/A_jj, or 1 where the diagonal is missing or zero   BasicPreconditioners.h:65-79
// The reference solves twice with identical input (:457,:460); once is enough (SURVEY B-4).
// Iteration numbering follows Eigen: `iterations()` is the loop counter at exit.
//
// Per iteration, over n rows with sliced-ELL off-diagonals (fp64 value + int32 column):
//   sweep A  ap = A p, partial p.ap     the matrix streams HBM -> shared memory through 1-D TMA bulk
//                                       copies (cp.async.bulk + mbarrier), one ring of stages per warp,
//                                       so the bytes in flight do not depend on registers/occupancy;
//                                       p[col] gathers are served by L1/L2 (p is 8n bytes, L2-resident)
//   sweep B  r -= a ap; |r|^2, r.z              read r,ap,invdiag, write r = 32n        (16-byte vectors)
//   sweep C  x += a p; p = z + b p (z = invdiag r)   read x,r,invdiag,p, write x,p = 48n
//            (x += a p rides in sweep C, which holds p in registers anyway: 8n bytes fewer per iteration than the
//            textbook placement in sweep B — 80n instead of the 88n the algorithmic figure counts; same operands, same rounding)
// = 12 nnz + 108 n algorithmic bytes, the figure bench.py's roofline uses.
//
// Multi-GPU (template MULTI): rank g solves the rows of its y-slab (a contiguous block of
// slices) in the SAME persistent kernel; nothing returns to the host between iterations and no NCCL call
// sits on the Krylov path. The three per-iteration global steps are fused device-side exchanges over
// NVLink peer memory: (1) p.ap, (2) |r|^2 and r.z — each rank's partial is stored into every peer's inbox
// and every rank adds the partials in rank order, so all ranks hold bit-identical scalars and take the
// same branches; (3) the halo rows of the new p, which sweep C stores directly into the neighbouring
// ranks' p vectors, published by the same flag exchange that acts as the cross-GPU barrier.
#include "mpsb_internal.cuh"

// This file is compiled twice: as itself (ring geometry 20 warps x 2 stages x 12 slots, tuned at 4 M rows) and through
// k_pcg_g16.cu (PCG_VARIANT 1: 16 warps x 2 x 16 slots, a whole ~28-slot slice in flight at once — faster for row blocks
// of up to ~1.3 M rows, i.e. a rank's block of the 4 M case on 4 or 8 GPUs). Variant 0's launcher dispatches.
#ifndef PCG_VARIANT
#define PCG_VARIANT 0
#endif
#if PCG_VARIANT == 1
#define mpsb_pcg_configure mpsb_pcg_configure_v1
#define mpsb_launch_pcg mpsb_launch_pcg_v1
#define mpsb_launch_spmv_probe mpsb_launch_spmv_probe_v1
#define cg_threads cg1_threads
#define cg_smem cg1_smem
#endif

#include <cfloat>
#include <cstdlib>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

#ifndef PCG_NSTAGES
#define PCG_NSTAGES 2
#endif
#ifndef PCG_NCH
#define PCG_NCH 12
#endif
#ifndef PCG_NWARPS
#define PCG_NWARPS 20
#endif
#ifndef PCG_NBATCH
#define PCG_NBATCH 2
#endif
// L2 reuse between the sweeps of one iteration (the vectors of 4 M rows are 5 x 32 MB, L2 is 126 MB): a sweep
// that walks the rows in the direction opposite to its predecessor meets the lines that one touched last.
// Measured at 4.04 M rows (profiles/r1_v23_pcg_l2_reuse_variants.txt): sweep C walking DOWN after sweep B
// walked up takes 12 % fewer cycles (shipped); sweep B walking down after sweep A gains nothing (the matrix
// stream has flushed A's p / ap lines) and costs 5 %; L2 eviction-priority hints (evict_first on the matrix
// stream, evict_last on the vectors) change nothing either way. Results do not depend on any of these.
#ifndef PCG_B_DESC
#define PCG_B_DESC 0
#endif
#ifndef PCG_C_DESC
#define PCG_C_DESC 1
#endif
#ifndef PCG_STREAM_HINT
#define PCG_STREAM_HINT 0 // 1: L2 evict_first policy on the matrix stream
#endif
#ifndef PCG_VEC_HINT
#define PCG_VEC_HINT 0 // 1: L2 evict_last policy on the p / ap / r accesses of sweeps B and C
#endif
#ifndef MPSB_MATRIX_FREE_DEFAULT
#define MPSB_MATRIX_FREE_DEFAULT 0 // sweep A recomputes off-diagonals from positions (see spmv_sweep_mf); env MPSB_MATRIX_FREE overrides
#endif
// L2 policy of the matrix stream (L2Plan): on for row blocks of up to MPSB_L2_MAX_ROWS rows — the sizes whose five CG
// vectors (40 B per row) fit the L2 next to the stream; measured neutral-to-negative above (4 M rows on one GPU)
#ifndef MPSB_L2_RESIDENT_KB_DEFAULT
#define MPSB_L2_RESIDENT_KB_DEFAULT 16384
#endif
#ifndef MPSB_L2_MAX_ROWS_DEFAULT
#define MPSB_L2_MAX_ROWS_DEFAULT 2500000
#endif
#ifndef PCG_SLICE_SUM_MAX
#define PCG_SLICE_SUM_MAX 40960 // several GPUs: blocks of up to this many slices sum p.Ap in the exchange's last CTA (no grid barrier)
#endif
#ifndef PCG_SPLIT_SYNC
#define PCG_SPLIT_SYNC 0 // diagnostic: book the barrier/exchange after sweep B under B and after sweep C under C
#endif

namespace {

constexpr int PCG_WARPS   = PCG_NWARPS;    // warps per CTA (one CTA per SM)
constexpr int PCG_THREADS = PCG_WARPS * 32;
constexpr int PCG_STAGES  = PCG_NSTAGES;   // ring depth per warp
constexpr int PCG_CH      = PCG_NCH;       // slots per stage (multiple of 4): 8 * 384 B = 3 KB + diag
// one stage: CH slots of 32 values, the slice's 32 diagonals, CH slots of 32 columns
struct __align__(128) Stage {
    double val[PCG_CH * 32];
    double diag[32];
    int col[PCG_CH * 32]; // 16-bit column mode: the first half holds short offsets from the slice's first row
};
struct __align__(16) StageMeta {
    int slice;
    int nslots;
    int first; // bit 0: first chunk of the slice (start the row sum from diag * p); bit 1: the slice may read a peer's rows
    int last;  // last chunk of the slice (store ap, accumulate p.ap)
};
struct Smem {
    Stage stage[PCG_WARPS][PCG_STAGES];
    StageMeta meta[PCG_WARPS][PCG_STAGES];
    unsigned long long full[PCG_WARPS][PCG_STAGES];
    double red[66];
    int need[MPSB_MAX_RANKS][2]; // MULTI: my rows each peer reads, [lo, hi)
    double xr[MPSB_MAX_RANKS][2]; // MULTI: every rank's pair of the last exchange
    double xr3[MPSB_MAX_RANKS][3]; // MULTI, single-reduction recurrence: every rank's triple
    int flag;                    // MULTI: this CTA arrived last at the running exchange
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    unsigned ok = 0, addr = smem_u32(bar), spins = 0;
    while (!ok) {
        if (++spins > (1u << 26)) __trap(); // a lost TMA completion must fail loudly, never hang the GPU
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(addr), "r"(parity)
            : "memory");
    }
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
#if PCG_STREAM_HINT
    // the matrix stream is read once per iteration: mark its lines first to go, so the vectors stay in L2
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
                 : "memory");
#else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
#endif
}
// the same with an L2 eviction-priority policy chosen at run time (resident part of the matrix: evict_last, the rest:
// evict_first — see L2Plan)
__device__ __forceinline__ void bulk_g2s_pol(void* dst, const void* src, unsigned bytes, unsigned long long* bar, unsigned long long pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
                 : "memory");
}
// 16-byte accesses of the vector sweeps, optionally tagged "last to go" in L2 (PCG_VEC_HINT)
__device__ __forceinline__ double2 ld2_keep(const double* base, int i) {
#if PCG_VEC_HINT
    double2 v;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(reinterpret_cast<const double2*>(base) + i), "l"(pol) : "memory");
    return v;
#else
    return reinterpret_cast<const double2*>(base)[i];
#endif
}
__device__ __forceinline__ void st2_keep(double* base, int i, double2 v) {
#if PCG_VEC_HINT
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(reinterpret_cast<double2*>(base) + i), "d"(v.x), "d"(v.y), "l"(pol) : "memory");
#else
    reinterpret_cast<double2*>(base)[i] = v;
#endif
}

// Deterministic block-wide reduction of two values; result broadcast to every thread.
// MINMAX: a is reduced with min, b with max (halo extents); otherwise both are summed.
template <bool MINMAX>
__device__ __forceinline__ void combine2(double& a, double& b, double oa, double ob) {
    if (MINMAX) {
        a = fmin(a, oa);
        b = fmax(b, ob);
    } else {
        a += oa;
        b += ob;
    }
}
template <bool MINMAX = false>
__device__ __forceinline__ void block_sum2(double& a, double& b, double* sh /* 66 */) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        combine2<MINMAX>(a, b, __shfl_xor_sync(0xffffffffu, a, o), __shfl_xor_sync(0xffffffffu, b, o));
    __syncthreads(); // protect sh from the previous use
    if (lane == 0) {
        sh[warp]      = a;
        sh[32 + warp] = b;
    }
    __syncthreads();
    if (warp == 0) {
        double va = sh[lane < PCG_WARPS ? lane : 0], vb = sh[32 + (lane < PCG_WARPS ? lane : 0)];
        if (!MINMAX && lane >= PCG_WARPS) va = vb = 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            combine2<MINMAX>(va, vb, __shfl_xor_sync(0xffffffffu, va, o), __shfl_xor_sync(0xffffffffu, vb, o));
        if (lane == 0) {
            sh[64] = va;
            sh[65] = vb;
        }
    }
    __syncthreads();
    a = sh[64];
    b = sh[65];
}

// Every CTA sums the same per-CTA partials in the same order -> identical scalars grid-wide.
__device__ __forceinline__ void grid_sum2(const double* pa, const double* pb, int nblocks, double& a, double& b,
                                          double* sh) {
    double la = 0.0, lb = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += PCG_THREADS) {
        la += __ldcg(pa + i);
        lb += __ldcg(pb + i);
    }
    block_sum2(la, lb, sh);
    a = la;
    b = lb;
}

// ---- cross-GPU exchange (MULTI) ------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct Xchg {
    unsigned long long epoch; // exchanges completed by this rank since attach (same on every rank)
    unsigned long long count; // exchanges completed in this launch
};

// One fused all-to-all + barrier over all CTAs of all ranks. Every CTA contributes (a, b); on return
// xr[q][0..1] (shared memory) holds rank q's combined pair on every CTA of every rank. Steps:
//   1. CTA partials -> local array; fence; arrive on the local counter.
//   2. The CTA that arrives last combines the partials in index order (deterministic, whoever is last) and
//      stores the rank's pair into the inbox of every rank.
//   3. Every CTA polls its OWN inbox (local HBM, no polling over NVLink) until all W ranks have delivered.
// Two wire formats:
//   XCHG_SUM / XCHG_MINMAX  carry data and nothing else has to be visible with them -> flag-in-data lines
//       (each double travels as one 16-byte store {lo32, flag, hi32, flag}, flag = low 32 bits of the epoch;
//       only 8-byte store atomicity is assumed). One NVLink one-way latency, no fence, no acknowledgement.
//   XCHG_PUBLISH            carries no data but publishes the halo rows of p that sweep C stored into the
//       peers' p vectors: every CTA fences at system scope before arriving, the last CTA stores the epoch flag
//       with release.sys, readers poll it with acquire.sys.
// It subsumes a grid-wide barrier. A rank that never arrives (crashed peer) trips the timeout -> trap,
// which fails the launch loudly instead of hanging the box.
//   XCHG_SLICES             XCHG_SUM whose local part is not a sum of per-CTA partials: the CTA that arrives last adds
//       the `ns` per-slice values of `slices` (p.Ap of every 32-row slice, stored by whichever warp took the slice) in
//       index order. Saves the grid barrier that would otherwise have to separate sweep A from a fixed per-CTA split
//       of that sum; deterministic whoever is last. Used when a rank's block is short (several GPUs).
enum { XCHG_SUM = 0, XCHG_MINMAX = 1, XCHG_PUBLISH = 2, XCHG_SLICES = 3 };

__device__ __forceinline__ void xchg_timeout(const Params& P, unsigned long long t0) {
    if (global_timer_ns() - t0 > 20000000000ull) { // 20 s: a peer is gone
        atomicOr(&P.stats->error_flags, 8);
        __threadfence_system();
        __trap();
    }
}

template <int MODE>
__device__ __forceinline__ void xchg_all(const Params& P, Xchg& xs, double a, double b, double* part0, double* part1,
                                         double* sh, int* sflag, double (*xr)[2], const double* slices = nullptr, int ns = 0) {
    constexpr bool MINMAX = MODE == XCHG_MINMAX;
    constexpr bool SLICES = MODE == XCHG_SLICES;
    const PeerLinks& L = P.L;
    const int W = L.world, nblocks = gridDim.x;
    xs.epoch++;
    xs.count++;
    const unsigned fl = (unsigned) xs.epoch;
    if (MODE != XCHG_PUBLISH && !SLICES) block_sum2<MINMAX>(a, b, sh);
    else __syncthreads(); // all halo stores (per-slice stores) of the CTA precede thread 0's fence
    if (threadIdx.x == 0) {
        if (SLICES) {
            __threadfence();
        } else if (MODE != XCHG_PUBLISH) {
            part0[blockIdx.x] = a;
            part1[blockIdx.x] = b;
            __threadfence();
        } else {
            __threadfence_system();
        }
        unsigned long long t = atomicAdd(L.arrive, 1ull);
        const int last = (t + 1ull == xs.count * (unsigned long long) nblocks) ? 1 : 0;
        if (last) {
            if (MODE != XCHG_PUBLISH) __threadfence();
            else __threadfence_system();
        }
        *sflag = last;
    }
    __syncthreads();
    if (*sflag) {
        if (MODE != XCHG_PUBLISH) {
            double la = MINMAX ? DBL_MAX : 0.0, lb = MINMAX ? -DBL_MAX : 0.0;
            if (SLICES) {
                // sixteen loads in flight per thread (a block of 31 k slices is 49 per thread: four L2 round trips, not
                // twelve), added in index order
                for (int i = threadIdx.x; i < ns; i += 16 * PCG_THREADS) {
                    double v[16];
#pragma unroll
                    for (int k = 0; k < 16; k++) v[k] = (i + k * PCG_THREADS < ns) ? __ldcg(slices + i + k * PCG_THREADS) : 0.0;
#pragma unroll
                    for (int k = 0; k < 16; k++) la += v[k];
                }
            } else {
                for (int i = threadIdx.x; i < nblocks; i += PCG_THREADS) combine2<MINMAX>(la, lb, __ldcg(part0 + i), __ldcg(part1 + i));
            }
            block_sum2<MINMAX>(la, lb, sh);
            if ((int) threadIdx.x < 2 * W) {
                const int q = threadIdx.x >> 1, ln = threadIdx.x & 1;
                const double v = ln ? lb : la;
                double* dst = L.peer_red[q] + ((size_t) (xs.epoch & 1ull) * MPSB_MAX_RANKS + L.rank) * 4 + ln * 2;
                asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(dst), "r"((unsigned) __double2loint(v)), "r"(fl),
                             "r"((unsigned) __double2hiint(v))
                             : "memory");
            }
        } else if ((int) threadIdx.x < W) {
            st_release_sys(L.peer_flag[threadIdx.x] + L.rank, xs.epoch);
        }
    }
    if (MODE != XCHG_PUBLISH) {
        if ((int) threadIdx.x < 2 * W) {
            const int q = threadIdx.x >> 1, ln = threadIdx.x & 1;
            const double* src = L.red + ((size_t) (xs.epoch & 1ull) * MPSB_MAX_RANKS + q) * 4 + ln * 2;
            const unsigned long long t0 = global_timer_ns();
            unsigned x, f0, y, f1, spins = 0;
            for (;;) {
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(f0), "=r"(y), "=r"(f1) : "l"(src) : "memory");
                if (f0 == fl && f1 == fl) break;
                if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
            }
            xr[q][ln] = __hiloint2double((int) y, (int) x);
            __threadfence(); // acquire side: what this rank's CTAs wrote before arriving is visible after the barrier below
        }
    } else if ((int) threadIdx.x < W) {
        const unsigned long long t0 = global_timer_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(L.flag + threadIdx.x) < xs.epoch)
            if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
    }
    __syncthreads();
}
// Three sums in ONE exchange (single-reduction recurrence, MPSB_CG_SINGLE_REDUCTION): (a, b) from per-CTA partials and c
// from the `ns` per-slice values of `slices`, both added by the CTA that arrives last, in index order. Wire format as
// XCHG_SUM, three 16-byte lines per rank, in a second region of the inbox (1 KiB behind the pairs). Result: xr3[q][0..2].
__device__ __forceinline__ void xchg_sum3(const Params& P, Xchg& xs, double a, double b, double* part0, double* part1, double* sh, int* sflag,
                                          double (*xr3)[3], const double* slices, int ns) {
    const PeerLinks& L = P.L;
    const int W = L.world, nblocks = gridDim.x;
    xs.epoch++;
    xs.count++;
    const unsigned fl = (unsigned) xs.epoch;
    block_sum2(a, b, sh);
    if (threadIdx.x == 0) {
        part0[blockIdx.x] = a;
        part1[blockIdx.x] = b;
        __threadfence();
        unsigned long long t = atomicAdd(L.arrive, 1ull);
        const int last = (t + 1ull == xs.count * (unsigned long long) nblocks) ? 1 : 0;
        if (last) __threadfence();
        *sflag = last;
    }
    __syncthreads();
    if (*sflag) {
        double la = 0.0, lb = 0.0, lc = 0.0, zero = 0.0;
        for (int i = threadIdx.x; i < nblocks; i += PCG_THREADS) {
            la += __ldcg(part0 + i);
            lb += __ldcg(part1 + i);
        }
        for (int i = threadIdx.x; i < ns; i += 16 * PCG_THREADS) {
            double v[16];
#pragma unroll
            for (int k = 0; k < 16; k++) v[k] = (i + k * PCG_THREADS < ns) ? __ldcg(slices + i + k * PCG_THREADS) : 0.0;
#pragma unroll
            for (int k = 0; k < 16; k++) lc += v[k];
        }
        block_sum2(la, lb, sh);
        block_sum2(lc, zero, sh);
        if ((int) threadIdx.x < 3 * W) {
            const int q = threadIdx.x / 3, ln = threadIdx.x - 3 * q;
            const double v = ln == 0 ? la : (ln == 1 ? lb : lc);
            double* dst = L.peer_red[q] + 128 + ((size_t) (xs.epoch & 1ull) * MPSB_MAX_RANKS + L.rank) * 6 + ln * 2;
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(dst), "r"((unsigned) __double2loint(v)), "r"(fl),
                         "r"((unsigned) __double2hiint(v))
                         : "memory");
        }
    }
    if ((int) threadIdx.x < 3 * W) {
        const int q = threadIdx.x / 3, ln = threadIdx.x - 3 * q;
        const double* src = L.red + 128 + ((size_t) (xs.epoch & 1ull) * MPSB_MAX_RANKS + q) * 6 + ln * 2;
        const unsigned long long t0 = global_timer_ns();
        unsigned x, f0, y, f1, spins = 0;
        for (;;) {
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(f0), "=r"(y), "=r"(f1) : "l"(src) : "memory");
            if (f0 == fl && f1 == fl) break;
            if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
        }
        xr3[q][ln] = __hiloint2double((int) y, (int) x);
        __threadfence();
    }
    __syncthreads();
}

// sum over ranks in rank order: bit-identical on every rank
__device__ __forceinline__ void xchg_sum2(const Params& P, double (*xr)[2], double& a, double& b) {
    a = 0.0;
    b = 0.0;
    for (int q = 0; q < P.L.world; q++) {
        a += xr[q][0];
        b += xr[q][1];
    }
}

// The exchange that ends sweep C, split in two so that nothing on this rank waits for NVLink (see HaloWait):
//   local barrier   every CTA fences at GPU scope and arrives; the last one releases this rank's own flag;
//                   every CTA waits for that flag only and goes on to sweep A.
//   publication     off the critical path: a CTA that stored halo rows into a peer's p vector fences at system
//                   scope, then every CTA arrives on a second counter; the last one fences and stores the epoch
//                   into the peers' inboxes. Peers look at it only when a warp reaches a boundary slice.
__device__ __forceinline__ void publish_split(const Params& P, Xchg& xs, unsigned long long& publishes, bool pushed) {
    const PeerLinks& L = P.L;
    const unsigned long long nblocks = gridDim.x;
    xs.epoch++;
    xs.count++;
    publishes++;
    const int any_pushed = __syncthreads_or(pushed ? 1 : 0); // also: the CTA's stores precede thread 0's fences
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(L.arrive, 1ull) + 1ull == xs.count * nblocks) {
            __threadfence();
            asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(L.flag + L.rank), "l"(xs.epoch) : "memory");
        }
        if (any_pushed) __threadfence_system();
        if (atomicAdd(L.arrive + 1, 1ull) + 1ull == publishes * nblocks) {
            __threadfence_system();
            for (int q = 0; q < L.world; q++)
                if (q != L.rank) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(L.peer_flag[q] + L.rank), "l"(xs.epoch) : "memory");
        }
        const unsigned long long t0 = global_timer_ns();
        unsigned spins = 0;
        unsigned long long seen;
        do {
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(seen) : "l"(L.flag + L.rank) : "memory");
            if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
        } while (seen < xs.epoch);
    }
    __syncthreads();
}

// Ring bookkeeping of one warp; serial numbers keep running across sweeps so the mbarrier phase
// parity of stage (serial % STAGES) is (serial / STAGES) & 1.
struct Pipe {
    unsigned issued;
    unsigned consumed;
};

constexpr int PCG_BATCH = PCG_NBATCH; // slices taken from the global queue per atomic

// Sweep A for the calling warp: ap = A p. Slices are handed out DYNAMICALLY (a global ticket counter,
// PCG_BATCH slices per ticket, tickets fetched two ahead so neither the atomic nor the width loads sit
// on the critical path): a static split leaves fast warps idle while HBM serves the slow ones, which
// cost 25 % of the sweep. Determinism is kept by storing p.ap PER SLICE (warp-reduced in a fixed order)
// into slice_pap[]; the caller sums that array in a fixed order. Rows of a slice beyond n carry diag 0 /
// padding only and p = 0 there.
// The sweep covers slices [sbeg, nslices) (a rank's row block; sbeg = 0 on one GPU).
// MULTI only: the order in which a rank walks its slices and what the slices at the ends of its block wait for.
// The exchange that ends sweep C is split: every CTA waits only for its OWN rank's flag (the local barrier),
// sweep A starts at once in the MIDDLE of the block (tickets are rotated by `shift`), and a warp waits for the
// peers' flags — the halo rows of p they stored into this rank's vector — only when it reaches a slice outside
// [gap_lo, gap_hi), i.e. one whose columns can leave the block. By then the flags have long arrived, so the
// NVLink latency of the halo publication is hidden behind the interior of the sweep.
struct HaloWait {
    int shift;                       // ticket v walks slice sbeg + (v - sbeg + shift) mod (nslices - sbeg)
    int gap_lo, gap_hi;              // slices [gap_lo, gap_hi) read no row of another rank
    const unsigned long long* flags; // this rank's inbox flags
    unsigned long long epoch;        // the publishing exchange the halo rows belong to
    int world;
};

// L2 policy of the matrix stream. The stream of a row block is read once per iteration in full; with the default
// policy it displaces the five CG vectors (40 B per row) from the L2 on every pass, although they are re-read within
// the iteration. With `on`, the stream travels as evict_first — the vectors stay L2-resident whenever they fit (sweeps B
// and C then run from L2, sweep A's p gathers and ap stores hit) — except the first slices [keep_lo, keep_hi), a small
// share sized by the host (MPSB_L2_RESIDENT_KB, 16 MB) that travels as evict_last and is served from L2 from the second
// iteration on. Measured on one B200 (profiles/r2_v4_l2_policy_experiment.txt): 48.9 -> 44.5 us/iteration at 0.5 M rows
// (a rank's block on 8 GPUs), 79.9 -> 73.4 at 1.0 M, 145.0 -> 140.8 at 2.0 M, 269.7 -> 272 at 4.0 M (vectors 160 MB: off
// there, MPSB_L2_MAX_ROWS). Larger kept shares gain nothing: what the L2 keeps usefully is about 60 MB, not 126 (data
// read from both dies appears to be held twice); the driver's persisting carve-out changes nothing. Results do not
// depend on any of this.
struct L2Plan {
    int on;
    int keep_lo, keep_hi;
    unsigned long long pol_keep, pol_stream;
};

template <bool C16>
__device__ __forceinline__ void spmv_sweep(const Params& P, Smem& sm, Pipe& pipe, int sbeg, int nslices, const double* p,
                                           double* ap, unsigned* ticket, double* slice_pap, const HaloWait& hw, const L2Plan& l2) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t K = (size_t) P.K;
    const int nsl_total = nslices - sbeg;
    auto rot = [&](int v) -> int { // ticket -> slice
        int x = v - sbeg + hw.shift;
        if (x >= nsl_total) x -= nsl_total;
        return sbeg + x;
    };
    bool halo_ready = false;
    auto take = [&]() -> int { // lane 0 holds the result; shuffled when needed
        int v = 0;
        if (lane == 0) v = sbeg + (int) atomicAdd(ticket, (unsigned) PCG_BATCH);
        return v;
    };
    auto widths = [&](int s0) -> int { return (lane < PCG_BATCH && s0 + lane < nslices) ? P.ell_width[rot(s0 + lane)] : 0; };
    // batch pipeline: cur (being issued), nxt (id known, widths loading), t2 (ticket in flight)
    int cur = __shfl_sync(0xffffffffu, take(), 0);
    int cur_w = widths(cur);
    int nxt = __shfl_sync(0xffffffffu, take(), 0);
    int nxt_w = widths(nxt);
    int t2 = take();
    int pi = 0, pk = 0; // producer cursor: slice index inside cur, slot offset
    int pw = __shfl_sync(0xffffffffu, cur_w, 0);
    auto produce = [&]() -> bool {
        if (cur + pi >= nslices || pipe.issued - pipe.consumed >= (unsigned) PCG_STAGES) return false;
        const int slice = rot(cur + pi);
        const int nsl   = min(PCG_CH, pw - pk);
        const int halo  = (slice < hw.gap_lo || slice >= hw.gap_hi) ? 2 : 0;
        const bool first = pk == 0, last = pk + nsl >= pw;
        const unsigned st = pipe.issued % PCG_STAGES;
        if (lane == 0) {
            Stage& S = sm.stage[warp][st];
            sm.meta[warp][st] = StageMeta{slice, nsl, (first ? 1 : 0) | halo, last ? 1 : 0};
            unsigned long long* bar = &sm.full[warp][st];
            mbar_arrive_expect_tx(bar, (first ? 256u : 0u) + (unsigned) nsl * (C16 ? 320u : 384u));
            const size_t off = ((size_t) slice * K + (size_t) pk) * 32;
            if (l2.on) {
                const unsigned long long pol = (slice >= l2.keep_lo && slice < l2.keep_hi) ? l2.pol_keep : l2.pol_stream;
                if (first) bulk_g2s_pol(S.diag, P.diag + (size_t) slice * 32, 256u, bar, pol);
                if (nsl > 0) {
                    bulk_g2s_pol(S.val, P.ell_val + off, (unsigned) nsl * 256u, bar, pol);
                    if (C16) bulk_g2s_pol(S.col, P.ell_col16 + off, (unsigned) nsl * 64u, bar, pol);
                    else bulk_g2s_pol(S.col, P.ell_col + off, (unsigned) nsl * 128u, bar, pol);
                }
            } else {
                if (first) bulk_g2s(S.diag, P.diag + (size_t) slice * 32, 256u, bar);
                if (nsl > 0) {
                    bulk_g2s(S.val, P.ell_val + off, (unsigned) nsl * 256u, bar);
                    if (C16) bulk_g2s(S.col, P.ell_col16 + off, (unsigned) nsl * 64u, bar);
                    else bulk_g2s(S.col, P.ell_col + off, (unsigned) nsl * 128u, bar);
                }
            }
        }
        pipe.issued++;
        pk += nsl;
        if (last) {
            pk = 0;
            if (++pi == PCG_BATCH) { // next batch; keep two tickets ahead
                pi = 0;
                cur = nxt; cur_w = nxt_w;
                nxt = __shfl_sync(0xffffffffu, t2, 0);
                nxt_w = widths(nxt);
                t2 = nxt < nslices ? take() : nslices;
            }
            pw = __shfl_sync(0xffffffffu, cur_w, pi);
        }
        return true;
    };
    double sum = 0.0, pr = 0.0;
    int row = 0;
    for (;;) {
        while (produce()) {}
        if (pipe.consumed == pipe.issued) break;
        __syncwarp(); // lane 0's meta of freshly issued stages is visible to the warp
        const unsigned st = pipe.consumed % PCG_STAGES, parity = (pipe.consumed / PCG_STAGES) & 1u;
        const StageMeta m = sm.meta[warp][st];
        const Stage& S    = sm.stage[warp][st];
        if (m.first & 1) { // own-row p: issued before the wait so its latency hides behind the TMA
            row = m.slice * 32 + lane;
            pr  = p[row];
        }
        mbar_wait(&sm.full[warp][st], parity);
        if ((m.first & 2) && !halo_ready) { // first slice of this warp that may read a peer's rows
            if (lane == 0) {
                const unsigned long long t0 = global_timer_ns();
                unsigned spins = 0;
                for (int q = 0; q < hw.world; q++)
                    while (ld_acquire_sys(hw.flags + q) < hw.epoch)
                        if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
            }
            __syncwarp();
            halo_ready = true;
        }
        if (m.first & 1) sum = S.diag[lane] * pr;
        double g[PCG_CH]; // all gathers of the stage in flight together
#pragma unroll
        for (int k = 0; k < PCG_CH; k++) {
            int c = 0;
            if (k < m.nslots) {
                if (C16) c = m.slice * 32 + (int) reinterpret_cast<const short*>(S.col)[k * 32 + lane];
                else c = S.col[k * 32 + lane];
            }
            g[k] = p[c];
        }
#pragma unroll
        for (int k = 0; k < PCG_CH; k += 4) {
            if (k < m.nslots) {
                double v0 = S.val[(k + 0) * 32 + lane], v1 = S.val[(k + 1) * 32 + lane];
                double v2 = S.val[(k + 2) * 32 + lane], v3 = S.val[(k + 3) * 32 + lane];
                sum += v0 * g[k + 0]; sum += v1 * g[k + 1]; sum += v2 * g[k + 2]; sum += v3 * g[k + 3];
            }
        }
        if (m.last) {
            ap[row] = sum;
            double v = pr * sum;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) slice_pap[m.slice] = v;
        }
        __syncwarp(); // every lane is done with the stage before lane 0 refills it
        pipe.consumed++;
    }
}


// ---- matrix-free sweep A (2-D) -------------------------------------------------------------------------
// The stored sweep streams 10 B per off-diagonal (fp64 value + 16-bit column) and sits at the HBM roof for
// that format (frac 0.99). Every off-diagonal is a function of one distance,
//     a_ij = -(mat_coef / density) * (re_p / d_ij - 1)             (src/mps.cpp:343-349)
// so this sweep streams only the 2-byte column and RECOMPUTES the value from the two positions: one 16-byte
// gather of (x_j, y_j) next to the 8-byte gather of p_j (both L1/L2-resident: SORTED order keeps a row's
// neighbours within +-1 cell row), 11 fp64 operations per entry on the otherwise idle FP64 pipe:
//     q   = (x_j - x_i)^2 + (y_j - y_i)^2                          differences of the stored doubles
//     z0  = rsqrt.approx(q)                                         MUFU.RSQ64H, ~22 bits of 1/d
//     z   = z0 + z0 e (1/2 + 3/8 e),  e = 1 - q z0^2               one third-order step: 22 -> ~64 bits
//     acc += (re z - 1) p_j
//     ap_i = diag_i p_i - (mat_coef / density) * acc
// The diagonal (which also carries the quasi-compressibility term and the Dirichlet doubling) stays stored.
// Entries differ from the assembled ones (which mpsb_get_ppe still returns bit-exactly) by a few ulp: the
// contract on matrix entries is 1e-10 relative (north_star), on the pressure 1e-10; both are tested. The
// operator stays exactly symmetric: (x_j - x_i)^2 == (x_i - x_j)^2 bit for bit.
// One ring stage holds the columns of up to MF_CH slots (the `val` area of a Stage reinterpreted), so a
// slice is normally ONE bulk copy of width * 64 B + its 256 B of diagonals.
constexpr int MF_CH = (PCG_CH * 256) / 64; // 16-bit column slots that fit the value area of a stage (48)
constexpr int MF_G  = 8;                    // entries whose gathers are in flight together

__device__ __forceinline__ double mf_weight(double dx, double dy, double re) {
    const double q = fma(dy, dy, dx * dx); // d^2
    double z;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(z) : "d"(q)); // MUFU.RSQ64H: ~22 bits of 1/d
    const double t  = z * z;
    const double e  = fma(-q, t, 1.0);
    const double g  = fma(0.375, e, 0.5);
    const double ge = g * e;
    z = fma(z, ge, z);      // third-order step: 1/d to ~1 ulp
    return fma(re, z, -1.0); // re / d - 1
}

__device__ __forceinline__ void spmv_sweep_mf(const Params& P, Smem& sm, Pipe& pipe, int sbeg, int nslices, const double* p,
                                              double* ap, unsigned* ticket, double* slice_pap, const HaloWait& hw) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t K = (size_t) P.K;
    const int nsl_total = nslices - sbeg;
    const double re = P.s.re_pressure;
    const double neg_coef = -(P.mat_coef / P.s.density);
    const double2* xy = P.row_xy;
    auto rot = [&](int v) -> int {
        int x = v - sbeg + hw.shift;
        if (x >= nsl_total) x -= nsl_total;
        return sbeg + x;
    };
    bool halo_ready = false;
    auto take = [&]() -> int {
        int v = 0;
        if (lane == 0) v = sbeg + (int) atomicAdd(ticket, (unsigned) PCG_BATCH);
        return v;
    };
    auto widths = [&](int s0) -> int { return (lane < PCG_BATCH && s0 + lane < nslices) ? P.ell_width[rot(s0 + lane)] : 0; };
    int cur = __shfl_sync(0xffffffffu, take(), 0);
    int cur_w = widths(cur);
    int nxt = __shfl_sync(0xffffffffu, take(), 0);
    int nxt_w = widths(nxt);
    int t2 = take();
    int pi = 0, pk = 0;
    int pw = __shfl_sync(0xffffffffu, cur_w, 0);
    auto produce = [&]() -> bool {
        if (cur + pi >= nslices || pipe.issued - pipe.consumed >= (unsigned) PCG_STAGES) return false;
        const int slice = rot(cur + pi);
        const int nsl   = min(MF_CH, pw - pk);
        const int halo  = (slice < hw.gap_lo || slice >= hw.gap_hi) ? 2 : 0;
        const bool first = pk == 0, last = pk + nsl >= pw;
        const unsigned st = pipe.issued % PCG_STAGES;
        if (lane == 0) {
            Stage& S = sm.stage[warp][st];
            sm.meta[warp][st] = StageMeta{slice, nsl, (first ? 1 : 0) | halo | (pk << 8), last ? 1 : 0};
            unsigned long long* bar = &sm.full[warp][st];
            mbar_arrive_expect_tx(bar, (first ? 256u : 0u) + (unsigned) nsl * 64u);
            if (first) bulk_g2s(S.diag, P.diag + (size_t) slice * 32, 256u, bar);
            if (nsl > 0) bulk_g2s(S.val, P.ell_col16 + ((size_t) slice * K + (size_t) pk) * 32, (unsigned) nsl * 64u, bar);
        }
        pipe.issued++;
        pk += nsl;
        if (last) {
            pk = 0;
            if (++pi == PCG_BATCH) {
                pi = 0;
                cur = nxt; cur_w = nxt_w;
                nxt = __shfl_sync(0xffffffffu, t2, 0);
                nxt_w = widths(nxt);
                t2 = nxt < nslices ? take() : nslices;
            }
            pw = __shfl_sync(0xffffffffu, cur_w, pi);
        }
        return true;
    };
    double acc = 0.0, pr = 0.0, dpr = 0.0;
    double2 me = make_double2(0.0, 0.0);
    int row = 0, cnt = 0;
    for (;;) {
        while (produce()) {}
        if (pipe.consumed == pipe.issued) break;
        __syncwarp();
        const unsigned st = pipe.consumed % PCG_STAGES, parity = (pipe.consumed / PCG_STAGES) & 1u;
        const StageMeta m = sm.meta[warp][st];
        const Stage& S    = sm.stage[warp][st];
        if (m.first & 1) { // own-row loads: issued before the wait so their latency hides behind the TMA
            row = m.slice * 32 + lane;
            pr  = p[row];
            me  = xy[row];
            cnt = P.ell_count[row];
        }
        mbar_wait(&sm.full[warp][st], parity);
        if ((m.first & 2) && !halo_ready) {
            if (lane == 0) {
                const unsigned long long t0 = global_timer_ns();
                unsigned spins = 0;
                for (int q = 0; q < hw.world; q++)
                    while (ld_acquire_sys(hw.flags + q) < hw.epoch)
                        if ((++spins & 1023u) == 0) xchg_timeout(P, t0);
            }
            __syncwarp();
            halo_ready = true;
        }
        if (m.first & 1) {
            dpr = S.diag[lane] * pr;
            acc = 0.0;
        }
        const short* c16 = reinterpret_cast<const short*>(S.val);
        const int base = m.slice * 32, k0 = m.first >> 8;
#pragma unroll 1
        for (int k = 0; k < m.nslots; k += MF_G) { // nslots is a multiple of 4: the second half of a group may be absent
            double2 q[MF_G];
            double g[MF_G];
#pragma unroll
            for (int b = 0; b < MF_G; b++) {
                const int c = (k + b < m.nslots) ? base + (int) c16[(k + b) * 32 + lane] : row;
                q[b] = xy[c];
                g[b] = p[c];
            }
#pragma unroll
            for (int b = 0; b < MF_G; b++) {
                const double w = mf_weight(q[b].x - me.x, q[b].y - me.y, re);
                // padding slots (column = own row, d = 0 -> w = inf/NaN) and absent ones are selected away, not multiplied
                acc = fma((k0 + k + b < cnt) ? w : 0.0, g[b], acc);
            }
        }
        if (m.last) {
            const double sum = fma(neg_coef, acc, dpr);
            ap[row] = sum;
            double v = pr * sum;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) slice_pap[m.slice] = v;
        }
        __syncwarp();
        pipe.consumed++;
    }
}

// MULTI = false: one GPU, rows [0, n). MULTI = true: this rank's row block; see the file header.
template <bool MULTI, bool SR>
__global__ void __launch_bounds__(PCG_THREADS, 1) pcg_kernel(Params P, int max_blocks, int iteration_cap, int matrix_free, int resident_kb) {
    if (*P.cluster_done) { // the cluster solve (k_pcg_cluster.cu) already did it; uniform over the grid
        if (blockIdx.x == 0 && threadIdx.x == 0) *P.ell_wide = 0; // consumed: the next assembly starts from "fits"
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) P.stats->t_solve_begin = global_timer_ns();
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    double* sh = sm.red;
    const int n       = *P.n_active;
    const int nblocks = gridDim.x;
    const int tid     = blockIdx.x * PCG_THREADS + threadIdx.x;
    const int stride  = nblocks * PCG_THREADS;
    double* x    = P.S.p;
    double* r    = P.cg_r;
    double* p    = P.cg_p; // deliberately not __restrict__/const: gathers must be coherent ld.global
    double* ap   = P.cg_ap;
    double* invd = P.cg_invdiag;
    double* part0 = P.cg_partials;
    double* part1 = part0 + max_blocks;
    double* part2 = part1 + max_blocks;
    unsigned* tickets = reinterpret_cast<unsigned*>(P.cg_partials + 8 * (size_t) max_blocks); // 2 slice queues

    // this rank's rows: [row_lo, row_hi) of the padded Np rows, of which [row_lo, act_hi) are active (< n)
    const int W = MULTI ? P.L.world : 1, rank = MULTI ? P.L.rank : 0;
    const int rows_per_rank = P.Np / W;               // Np is a multiple of 32 * 8
    const int row_lo = rank * rows_per_rank, row_hi = row_lo + rows_per_rank;
    const int act_hi = max(row_lo, min(row_hi, n));
    const int sl_lo = row_lo >> 5, sl_hi = (act_hi + 31) >> 5; // slices with at least one active row
    Xchg xs{0ull, 0ull};
    if (MULTI) xs.epoch = *P.L.epoch;
    // columns stream as 16-bit offsets from the slice's first row (10 B per entry instead of 12) unless the
    // assembly met an offset that does not fit (matrix_kernel sets ell_wide); uniform over the grid
    const bool col16 = *P.ell_wide == 0;

    if (threadIdx.x < PCG_WARPS * PCG_STAGES) mbar_init(&sm.full[0][0] + threadIdx.x, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    Pipe pipe{0u, 0u};

    // L2 residency plan (see L2Plan): keep the first slices of the block whose stream adds up to the budget. Every CTA
    // adds the block's slice widths itself (the array is 4 B per 32 rows and L2-resident; once per solve).
    L2Plan l2{0, 0, 0, 0ull, 0ull};
    if (resident_kb > 0 && sl_hi > sl_lo) {
        double w = 0.0, unused = 0.0;
        for (int sl = sl_lo + (int) threadIdx.x; sl < sl_hi; sl += PCG_THREADS) w += (double) P.ell_width[sl];
        block_sum2(w, unused, sh);
        const double bytes = w * (col16 ? 320.0 : 384.0) + 256.0 * (double) (sl_hi - sl_lo);
        const double share = fmin(1.0, (double) resident_kb * 1024.0 / fmax(bytes, 1.0));
        l2.on      = 1;
        l2.keep_lo = sl_lo;
        l2.keep_hi = sl_lo + (int) (share * (double) (sl_hi - sl_lo));
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(l2.pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(l2.pol_stream));
    }

    // setup: x = 0, r = b, invdiag, p = invdiag r (0 beyond n, up to the slice padding); |b|^2 and r.p
    double l0 = 0.0, l1 = 0.0;
    for (int i = row_lo + tid; i < row_hi; i += stride) {
        if (i < P.N) x[i] = 0.0;
        double pi = 0.0;
        if (i < n) {
            double bi = P.rhs[i], d = P.diag[i];
            double id = d != 0.0 ? 1.0 / d : 1.0;
            pi = id * bi;
            r[i] = bi; invd[i] = id;
            l0 += bi * bi;
            l1 += bi * pi;
        }
        p[i] = pi;
    }
    // MULTI: which of my rows each peer reads in its sweep A. Every rank publishes the column extent of
    // its row block [need_lo, need_hi]; the rows of mine inside a peer's extent are pushed to that peer
    // whenever p changes (here and in sweep C). push_gap_* bracket the rows nobody else needs.
    int push_gap_lo = row_lo, push_gap_hi = row_hi;
    unsigned long long halo_epoch = 0ull; // epoch of the last exchange that published halo rows
    if (MULTI) {
        double cmn = DBL_MAX, cmx = -DBL_MAX;
        for (int sl = sl_lo + tid; sl < sl_hi; sl += stride) {
            cmn = fmin(cmn, (double) P.ell_cmin[sl]);
            cmx = fmax(cmx, (double) P.ell_cmax[sl]);
        }
        xchg_all<XCHG_MINMAX>(P, xs, cmn, cmx, part0, part1, sh, &sm.flag, sm.xr);
        if ((int) threadIdx.x < W) {
            const int q = threadIdx.x;
            double a = sm.xr[q][0], b = sm.xr[q][1];
            int lo = 0, hi = 0; // empty: a rank without active rows needs nothing
            if (a <= b && q != rank) {
                lo = max((int) a, row_lo);
                hi = min((int) b + 1, row_hi);
                if (hi < lo) hi = lo;
            }
            sm.need[q][0] = lo;
            sm.need[q][1] = hi;
        }
        __syncthreads();
        for (int q = 0; q < W; q++) {
            if (sm.need[q][1] <= sm.need[q][0]) continue;
            if (q < rank) push_gap_lo = max(push_gap_lo, sm.need[q][1]);
            else push_gap_hi = min(push_gap_hi, sm.need[q][0]);
        }
        // first halo push: rows of the initial p that peers read (written above by other threads; the
        // exchange just passed is a barrier over this rank's CTAs)
        for (int q = 0; q < W; q++) {
            double* dst = P.L.peer_p[q];
            for (int i = sm.need[q][0] + tid; i < sm.need[q][1]; i += stride) dst[i] = p[i];
        }
        xchg_all<XCHG_PUBLISH>(P, xs, 0.0, 0.0, part0, part1, sh, &sm.flag, sm.xr); // the halo rows are visible before any sweep A
        halo_epoch = xs.epoch; // flags carry the epochs of PUBLISHING exchanges only
    }
    double rhsNorm2, absNew;
    if (MULTI) {
        xchg_all<XCHG_SUM>(P, xs, l0, l1, part0, part1, sh, &sm.flag, sm.xr);
        xchg_sum2(P, sm.xr, rhsNorm2, absNew);
    } else {
        block_sum2(l0, l1, sh);
        if (threadIdx.x == 0) {
            part0[blockIdx.x] = l0;
            part1[blockIdx.x] = l1;
        }
        grid.sync();
        grid_sum2(part0, part1, nblocks, rhsNorm2, absNew, sh);
    }
    int it = 0;
    double residualNorm2 = rhsNorm2;
    bool done = false;
    if (rhsNorm2 == 0.0) { // x = 0, iters = 0, error = 0 (ConjugateGradient.h:44-50)
        residualNorm2 = 0.0;
        done = true;
    }
    const double threshold = fmax(DBL_EPSILON * DBL_EPSILON * rhsNorm2, DBL_MIN);
    if (!done && residualNorm2 < threshold) done = true;
    if (!MULTI) grid.sync(); // part0/part1 are reused below
    const int maxIters = iteration_cap > 0 ? iteration_cap : 2 * P.N;
    // sweeps B and C walk the active rows of the block two at a time (row_lo is even)
    const int pair_lo = row_lo >> 1, pair_hi = pair_lo + ((act_hi - row_lo) >> 1);
    const bool odd_tail = ((act_hi - row_lo) & 1) != 0; // row act_hi - 1 is left over

    // MULTI: store row i of the new p into the p vector of every peer that reads it
    bool did_push = false;              // this thread stored into a peer's p vector in the running sweep C
    unsigned long long publishes = 0ull; // split publications so far in this launch (second arrival counter)
    auto push_row = [&](int i, double v) {
        if (i >= push_gap_lo && i < push_gap_hi) return;
        for (int q = 0; q < W; q++)
            if (i >= sm.need[q][0] && i < sm.need[q][1]) {
                P.L.peer_p[q][i] = v;
                did_push = true;
            }
    };

    // sweep A's walk (see HaloWait): single GPU = natural order, nothing to wait for
    HaloWait hw{0, sl_lo, sl_hi, nullptr, 0ull, 1};
    if (MULTI) {
        hw.shift  = (sl_hi - sl_lo) / 2;
        hw.gap_lo = (max(push_gap_lo, row_lo) + 31) >> 5;
        hw.gap_hi = min(push_gap_hi, row_hi) >> 5;
        hw.flags  = P.L.flag;
        hw.world  = W;
    }
    // pairs of active rows at the high end of the block that some peer reads (sweep C handles them first)
    const int high_pairs = MULTI ? max(0, min(pair_hi - pair_lo, pair_hi - (max(push_gap_hi, row_lo) >> 1))) : 0;
    long long cyc[4] = {0, 0, 0, 0}, t0 = clock64(), t1;
#define PCG_TICK(slot) do { t1 = clock64(); cyc[slot] += t1 - t0; t0 = t1; } while (0)
    if (SR) {
        // ---- single-reduction recurrence (Chronopoulos & Gear 1989), MPSB_CG_SINGLE_REDUCTION=1, never the default ----
        // Same Krylov space, same stopping rule on the recursively updated |r|^2, but ONE global reduction per
        // iteration instead of two (and two grid-wide synchronisations instead of four): with u = D^-1 r, w = A u,
        //     gamma = r.u, delta = u.w, |r|^2            one reduction, after the sweep that forms w
        //     beta = gamma / gamma_old, alpha = gamma / (delta - beta gamma / alpha_old)
        //     d = u + beta d; s = w + beta s; x += alpha d; r -= alpha s; u = D^-1 r        one fused vector sweep
        // (d is the search direction p of the textbook loop, s = A d by recurrence instead of by product). The
        // scalars differ from Eigen's in rounding (delta - beta gamma / alpha against p.Ap), so this mode is held to
        // the 1e-10 / +-2 % bars, not to bit-identity with the default loop. The convergence test of update k arrives
        // with the reduction of sweep k + 1: one product is spent after convergence. u lives in cg_p (the vector sweep A
        // gathers and the one whose halo rows travel), w in cg_ap.
        double* u = p;
        double* w = ap;
        double* d = P.cg_dir;
        double* s = P.cg_s;
        double gamma = absNew, alpha = 0.0, lg = 0.0, lrr = 0.0;
        bool first = true;
        unsigned sweeps = 0; // sweep A count: the slice queue alternates per SWEEP (`it` lags one sweep behind here)
        while (!done) {
            unsigned* ticket = tickets + (sweeps & 1u);
            if (MULTI) hw.epoch = halo_epoch;
            if (col16) spmv_sweep<true>(P, sm, pipe, sl_lo, sl_hi, u, w, ticket, P.cg_slice_pap, hw, l2);
            else spmv_sweep<false>(P, sm, pipe, sl_lo, sl_hi, u, w, ticket, P.cg_slice_pap, hw, l2);
            __syncthreads();
            PCG_TICK(0);
            double G, RR, D;
            if (MULTI) {
                xchg_sum3(P, xs, lg, lrr, part0, part1, sh, &sm.flag, sm.xr3, P.cg_slice_pap + sl_lo, sl_hi - sl_lo);
                G = RR = D = 0.0;
                for (int q = 0; q < W; q++) {
                    G += sm.xr3[q][0];
                    RR += sm.xr3[q][1];
                    D += sm.xr3[q][2];
                }
                if (tid == 0) tickets[(sweeps + 1u) & 1u] = 0u;
            } else {
                grid.sync(); // every slice of the sweep has its u.w stored
                if (tid == 0) tickets[(sweeps + 1u) & 1u] = 0u;
                double ld = 0.0, zero = 0.0;
                {
                    const int nsl = sl_hi - sl_lo, per = (nsl + nblocks - 1) / nblocks;
                    const int lo = sl_lo + blockIdx.x * per, hi = min(lo + per, sl_hi);
                    for (int i = lo + (int) threadIdx.x; i < hi; i += PCG_THREADS) ld += __ldcg(P.cg_slice_pap + i);
                }
                block_sum2(lg, lrr, sh);
                block_sum2(ld, zero, sh);
                if (threadIdx.x == 0) {
                    part0[blockIdx.x] = lg;
                    part1[blockIdx.x] = lrr;
                    part2[blockIdx.x] = ld;
                }
                grid.sync();
                grid_sum2(part0, part1, nblocks, G, RR, sh);
                grid_sum2(part2, part2, nblocks, D, zero, sh); // (the partials are rewritten two grid barriers from here)
            }
            sweeps++;
            double beta = 0.0;
            if (first) {
                alpha = gamma / D;
            } else {
                residualNorm2 = RR;
                if (RR < threshold) break;
                it++;
                if (it >= maxIters) break;
                beta  = G / gamma;
                alpha = G / (D - beta * G / alpha);
                gamma = G;
            }
            PCG_TICK(3);
            // ---- fused vector sweep: d, s, x, r, u; partials of r.u and |r|^2 ----
            lg = 0.0;
            lrr = 0.0;
            for (int j = tid; j < pair_hi - pair_lo; j += stride) {
                const int i = MULTI ? (j < high_pairs ? pair_hi - high_pairs + j : pair_lo + j - high_pairs) : pair_lo + j;
                double2 uu = reinterpret_cast<const double2*>(u)[i], ww = reinterpret_cast<const double2*>(w)[i];
                double2 dd = uu, ss = ww;
                if (!first) {
                    const double2 d0 = reinterpret_cast<const double2*>(d)[i], s0 = reinterpret_cast<const double2*>(s)[i];
                    dd.x = uu.x + beta * d0.x; dd.y = uu.y + beta * d0.y;
                    ss.x = ww.x + beta * s0.x; ss.y = ww.y + beta * s0.y;
                }
                double2 xx = reinterpret_cast<const double2*>(x)[i], rr = reinterpret_cast<const double2*>(r)[i];
                const double2 id = reinterpret_cast<const double2*>(invd)[i];
                xx.x += alpha * dd.x; xx.y += alpha * dd.y;
                rr.x -= alpha * ss.x; rr.y -= alpha * ss.y;
                uu.x = id.x * rr.x; uu.y = id.y * rr.y;
                reinterpret_cast<double2*>(d)[i] = dd;
                reinterpret_cast<double2*>(s)[i] = ss;
                reinterpret_cast<double2*>(x)[i] = xx;
                reinterpret_cast<double2*>(r)[i] = rr;
                reinterpret_cast<double2*>(u)[i] = uu;
                lg += rr.x * uu.x; lg += rr.y * uu.y;
                lrr += rr.x * rr.x; lrr += rr.y * rr.y;
                if (MULTI) {
                    push_row(2 * i, uu.x);
                    push_row(2 * i + 1, uu.y);
                }
            }
            if (odd_tail && tid == 0) {
                const int i = act_hi - 1;
                const double di = first ? u[i] : u[i] + beta * d[i], si = first ? w[i] : w[i] + beta * s[i];
                const double ri = r[i] - alpha * si, ui = invd[i] * ri;
                x[i] += alpha * di;
                d[i] = di; s[i] = si; r[i] = ri; u[i] = ui;
                lg += ri * ui;
                lrr += ri * ri;
                if (MULTI) push_row(i, ui);
            }
            first = false;
            PCG_TICK(1);
            if (MULTI) {
                publish_split(P, xs, publishes, did_push);
                did_push   = false;
                halo_epoch = xs.epoch;
            } else grid.sync();
            PCG_TICK(3);
        }
    } else
    while (!done && it < maxIters) {
        // ---- sweep A: ap = A p, p.ap ----
        unsigned* ticket = tickets + (it & 1);
        if (MULTI) hw.epoch = halo_epoch;
        if (col16 && matrix_free) spmv_sweep_mf(P, sm, pipe, sl_lo, sl_hi, p, ap, ticket, P.cg_slice_pap, hw);
        else if (col16) spmv_sweep<true>(P, sm, pipe, sl_lo, sl_hi, p, ap, ticket, P.cg_slice_pap, hw, l2);
        else spmv_sweep<false>(P, sm, pipe, sl_lo, sl_hi, p, ap, ticket, P.cg_slice_pap, hw, l2);
        __syncthreads();
        PCG_TICK(0);
        double pap, unused;
        if (MULTI && sl_hi - sl_lo <= PCG_SLICE_SUM_MAX) {
            // short block: the exchange's last CTA adds the per-slice values itself, no grid barrier before it
            xchg_all<XCHG_SLICES>(P, xs, 0.0, 0.0, part0, part1, sh, &sm.flag, sm.xr, P.cg_slice_pap + sl_lo, sl_hi - sl_lo);
            xchg_sum2(P, sm.xr, pap, unused);
            if (tid == 0) tickets[(it + 1) & 1] = 0u; // every CTA of the rank is past sweep A; idle until the next one
        } else {
            grid.sync();
            // p.ap = sum of the per-slice partials in a fixed order: this CTA's fixed range, then all CTAs
            if (tid == 0) tickets[(it + 1) & 1] = 0u; // the other counter is idle until the next sweep A
            double lpap = 0.0, dummy = 0.0;
            {
                const int nsl = sl_hi - sl_lo, per = (nsl + nblocks - 1) / nblocks;
                const int lo = sl_lo + blockIdx.x * per, hi = min(lo + per, sl_hi);
                for (int i = lo + (int) threadIdx.x; i < hi; i += PCG_THREADS) lpap += __ldcg(P.cg_slice_pap + i);
            }
            if (MULTI) {
                xchg_all<XCHG_SUM>(P, xs, lpap, dummy, part0, part1, sh, &sm.flag, sm.xr);
                xchg_sum2(P, sm.xr, pap, unused);
            } else {
                block_sum2(lpap, dummy, sh);
                if (threadIdx.x == 0) part0[blockIdx.x] = lpap;
                grid.sync();
                grid_sum2(part0, part0, nblocks, pap, unused, sh);
            }
        }
        const double alpha = absNew / pap;
        PCG_TICK(3);
        // ---- sweep B: x += alpha p; r -= alpha ap; |r|^2; r.z  (two rows per thread, 16-byte accesses) ----
        double lrr = 0.0, lrz = 0.0;
#pragma unroll 2
        for (int j = tid; j < pair_hi - pair_lo; j += stride) {
            const int i = PCG_B_DESC ? pair_hi - 1 - j : pair_lo + j; // see PCG_B_DESC
            double2 rr = ld2_keep(r, i);
            double2 aa = ld2_keep(ap, i), dd = reinterpret_cast<const double2*>(invd)[i];
            rr.x -= alpha * aa.x; rr.y -= alpha * aa.y;
            st2_keep(r, i, rr);
            lrr += rr.x * rr.x; lrr += rr.y * rr.y;
            lrz += rr.x * (dd.x * rr.x); lrz += rr.y * (dd.y * rr.y);
        }
        if (odd_tail && tid == 0) {
            int i = act_hi - 1;
            double ri = r[i] - alpha * ap[i];
            r[i] = ri;
            lrr += ri * ri;
            lrz += ri * (invd[i] * ri);
        }
        PCG_TICK(1);
        double rz;
        if (MULTI) {
            xchg_all<XCHG_SUM>(P, xs, lrr, lrz, part1, part2, sh, &sm.flag, sm.xr);
            xchg_sum2(P, sm.xr, residualNorm2, rz);
        } else {
            block_sum2(lrr, lrz, sh);
            if (threadIdx.x == 0) {
                part1[blockIdx.x] = lrr;
                part2[blockIdx.x] = lrz;
            }
            grid.sync();
            grid_sum2(part1, part2, nblocks, residualNorm2, rz, sh);
        }
        if (residualNorm2 < threshold) { // converged: the x update sweep C would have carried (same operands, same rounding)
            for (int j = tid; j < pair_hi - pair_lo; j += stride) {
                const int i = pair_lo + j;
                double2 pp = reinterpret_cast<const double2*>(p)[i], xx = reinterpret_cast<const double2*>(x)[i];
                xx.x += alpha * pp.x; xx.y += alpha * pp.y;
                reinterpret_cast<double2*>(x)[i] = xx;
            }
            if (odd_tail && tid == 0) x[act_hi - 1] += alpha * p[act_hi - 1];
            break;
        }
        const double beta = rz / absNew;
        absNew = rz;
        PCG_TICK(PCG_SPLIT_SYNC ? 1 : 3);
        // ---- sweep C: p = z + beta p (MULTI: halo rows also go to the peers that read them) ----
        // MULTI: the rows peers read above this block (the high end) go first, then the block from its low end,
        // so every remote store is issued early and its acknowledgement is back before the publishing fence
#pragma unroll 2
        for (int j = tid; j < pair_hi - pair_lo; j += stride) {
            const int i = MULTI ? (j < high_pairs ? pair_hi - high_pairs + j : pair_lo + j - high_pairs)
                                : (PCG_C_DESC ? pair_hi - 1 - j : pair_lo + j);
            double2 pp = ld2_keep(p, i), rr = ld2_keep(r, i);
            double2 dd = reinterpret_cast<const double2*>(invd)[i];
            double2 xx = reinterpret_cast<const double2*>(x)[i];
            xx.x += alpha * pp.x; xx.y += alpha * pp.y; // x += alpha p of THIS iteration, carried here so sweep B need not read p
            reinterpret_cast<double2*>(x)[i] = xx;
            pp.x = dd.x * rr.x + beta * pp.x;
            pp.y = dd.y * rr.y + beta * pp.y;
            st2_keep(p, i, pp);
            if (MULTI) {
                push_row(2 * i, pp.x);
                push_row(2 * i + 1, pp.y);
            }
        }
        if (odd_tail && tid == 0) {
            int i = act_hi - 1;
            x[i] += alpha * p[i];
            double v = invd[i] * r[i] + beta * p[i];
            p[i] = v;
            if (MULTI) push_row(i, v);
        }
        it++;
        PCG_TICK(2);
        if (MULTI) { // local barrier + publication of the halo rows; the peers' flags are awaited inside sweep A
            publish_split(P, xs, publishes, did_push);
            did_push   = false;
            halo_epoch = xs.epoch;
        } else grid.sync();
        PCG_TICK(PCG_SPLIT_SYNC ? 2 : 3);
    }
    if (threadIdx.x == 0) { // per-CTA diagnostic rows behind the three reduction rows
        long long* dbg = reinterpret_cast<long long*>(P.cg_partials + 3 * (size_t) max_blocks);
        for (int k = 0; k < 4; k++) dbg[(size_t) k * max_blocks + blockIdx.x] = cyc[k];
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        dbg[(size_t) 4 * max_blocks + blockIdx.x] = smid;
    }
    if (tid == 0) {
        for (int k = 0; k < 4; k++) P.stats->cg_cycles[k] = cyc[k];
        P.stats->cg_iterations   = it;
        P.stats->pad[0]          = col16 ? 1 : 0;
        P.stats->pad[1]          = ((col16 && matrix_free) ? 2 : 0) | (SR ? 4 : 0); // bit 1: sweep A recomputed the off-diagonals; bit 2: single-reduction recurrence
        // leave the launch-scoped scratch cleared for the next solve (saves memset nodes): the two slice-queue
        // counters (no sweep A is running any more) and the column-width flag of the assembly just consumed
        tickets[0] = 0u;
        tickets[1] = 0u;
        *P.ell_wide = 0;
        P.stats->t_solve_end     = global_timer_ns();
        P.stats->rhs_norm2       = rhsNorm2;
        P.stats->cg_rel_residual = rhsNorm2 > 0.0 ? sqrt(residualNorm2 / rhsNorm2) : 0.0;
        if (MULTI) *P.L.epoch = xs.epoch; // every CTA read the old value before its first exchange
    }
}


// Diagnostic / parity: ONE sweep A (y = A x over the SORTED rows of the assembled system) outside the solve, on
// the stored stream or on the matrix-free recomputation, so tests can compare the two operators entry by entry
// (unit vectors) and on random vectors. x lives in cg_p, y in cg_ap; single GPU only.
__global__ void __launch_bounds__(PCG_THREADS, 1) spmv_probe_kernel(Params P, int max_blocks, int matrix_free) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    if (threadIdx.x < PCG_WARPS * PCG_STAGES) mbar_init(&sm.full[0][0] + threadIdx.x, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    Pipe pipe{0u, 0u};
    const int n = *P.n_active, sl_hi = (n + 31) >> 5;
    unsigned* tickets = reinterpret_cast<unsigned*>(P.cg_partials + 8 * (size_t) max_blocks);
    HaloWait hw{0, 0, sl_hi, nullptr, 0ull, 1};
    const bool col16 = *P.ell_wide == 0;
    const L2Plan l2{0, 0, 0, 0ull, 0ull};
    if (col16 && matrix_free) spmv_sweep_mf(P, sm, pipe, 0, sl_hi, P.cg_p, P.cg_ap, tickets, P.cg_slice_pap, hw);
    else if (col16) spmv_sweep<true>(P, sm, pipe, 0, sl_hi, P.cg_p, P.cg_ap, tickets, P.cg_slice_pap, hw, l2);
    else spmv_sweep<false>(P, sm, pipe, 0, sl_hi, P.cg_p, P.cg_ap, tickets, P.cg_slice_pap, hw, l2);
}

} // namespace

int mpsb_pcg_configure(mpsb_handle* h) {
    int smem = (int) sizeof(Smem);
    if (cudaFuncSetAttribute(pcg_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(pcg_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(pcg_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(pcg_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    // opt-in: one global reduction per iteration (Chronopoulos-Gear recurrence); different rounding, never the default
    const char* sr = getenv("MPSB_CG_SINGLE_REDUCTION");
    h->cg_single_reduction = sr && atoi(sr) ? 1 : 0;
    if (cudaFuncSetAttribute(spmv_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pcg_kernel<true, true>, PCG_THREADS, smem);
    if (e != cudaSuccess || per_sm < 1) return -1;
    h->cg_threads    = PCG_THREADS;
    h->cg_smem       = smem;
    h->cg_max_blocks = h->num_sms; // one CTA per SM
    const char* cap  = getenv("MPSB_CG_ITERATION_CAP"); // diagnostic (profiling a bounded solve); 0 = Eigen's 2N
    h->cg_iteration_cap = cap ? atoi(cap) : 0;
    // matrix-free sweep A: 2-D cases with a flat z (all particles in one plane) only; MPSB_MATRIX_FREE=0/1 overrides
    const mpsb_settings& s = h->P.s;
    const char* mf = getenv("MPSB_MATRIX_FREE");
    // part of the matrix stream kept L2-resident across iterations (L2Plan): MPSB_L2_RESIDENT_MB, 0 = off
    h->cg_small_rows = getenv("MPSB_CG_SMALL_ROWS") ? atoi(getenv("MPSB_CG_SMALL_ROWS")) : 1300000;
    const char* l2kb = getenv("MPSB_L2_RESIDENT_KB");
    h->cg_l2_resident_kb = l2kb ? atoi(l2kb) : MPSB_L2_RESIDENT_KB_DEFAULT;
    h->cg_l2_max_rows    = getenv("MPSB_L2_MAX_ROWS") ? atoi(getenv("MPSB_L2_MAX_ROWS")) : MPSB_L2_MAX_ROWS_DEFAULT;
    h->cg_matrix_free = (s.dim == 2 && s.domain_max[2] == s.domain_min[2]) ? (mf ? (atoi(mf) ? 1 : 0) : MPSB_MATRIX_FREE_DEFAULT) : 0;
    return 0;
}

void mpsb_launch_pcg(mpsb_handle* h) {
    Params& P = h->P;
    const bool multi = P.L.world > 1;
    // The matrix-free sweep gathers positions (row_xy) of halo columns, and row_xy holds only this rank's own rows (matrix_kernel
    // writes [row_lo, row_hi); it is not in the peer arena and no halo push fills it): one GPU only, several GPUs stream the stored entries.
    const int mfree = multi ? 0 : h->cg_matrix_free;
#if PCG_VARIANT == 0
    mpsb_launch_pcg_cluster(h); // small systems: one cluster, problem resident in shared memory; sets *cluster_done
    // short row blocks: the 16-warp x 16-slot build of this file (k_pcg_g16.cu). Measured (profiles/r2_v13_ring_geometry.txt), us per
    // iteration, 16 x 16 against 20 x 12: one GPU 32.1 / 35.3 at 0.25 M rows, 43.7 / 44.7 at 0.5 M, 70.0 / 73.3 at 1.0 M, 141.6 / 142.6
    // at 2.0 M, 288.1 / 285.2 at 4.0 M (its 512-thread vector sweeps are slower); two GPUs 49.3 / 53.3 at 0.5 M rows per rank,
    // 76.8 / 85.8 at 1.0 M, 142.1 / 152.0 at 2.0 M. Hence: up to cg_small_rows on one GPU, up to twice that per rank on several.
    if (P.Np / (multi ? P.L.world : 1) <= (multi ? 2 : 1) * h->cg_small_rows && !mfree) {
        mpsb_launch_pcg_v1(h);
        return;
    }
#endif
    // one 32-row slice per warp at least; never more CTAs than can be co-resident (cooperative launch)
    int want = mpsb_div_up(P.Np / 32 / (multi ? P.L.world : 1), PCG_WARPS);
    int grid = want < h->cg_max_blocks ? want : h->cg_max_blocks;
    if (grid < 1) grid = 1;
    int max_blocks = h->cg_max_blocks;
    if (multi) cudaMemsetAsync(P.L.arrive, 0, 2 * sizeof(unsigned long long), h->stream); // both arrival counters
    int cap = h->cg_iteration_cap;
    int resident_kb = (P.Np / (multi ? P.L.world : 1) <= h->cg_l2_max_rows) ? h->cg_l2_resident_kb : 0;
    void* args[] = {(void*) &P, (void*) &max_blocks, (void*) &cap, (void*) &mfree, (void*) &resident_kb};
    const bool sr = h->cg_single_reduction && !mfree;
    const void* fn = multi ? (sr ? (const void*) pcg_kernel<true, true> : (const void*) pcg_kernel<true, false>)
                           : (sr ? (const void*) pcg_kernel<false, true> : (const void*) pcg_kernel<false, false>);
    cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(PCG_THREADS), args, (size_t) h->cg_smem, h->stream);
    h->launches++;
}

// y = A x for the last assembled PPE; x and y are DEVICE vectors over SORTED rows already placed in cg_p / read from
// cg_ap by the caller (mpsb_debug_spmv, mpsb_api.cu).
void mpsb_launch_spmv_probe(mpsb_handle* h, int matrix_free) {
    Params& P = h->P;
    int grid = mpsb_div_up(P.Np / 32, PCG_WARPS);
    if (grid > h->cg_max_blocks) grid = h->cg_max_blocks;
    if (grid < 1) grid = 1;
    cudaMemsetAsync(P.cg_partials + 8 * (size_t) h->cg_max_blocks, 0, 2 * sizeof(double), h->stream); // slice-queue counters
    MPSB_LAUNCH(h, spmv_probe_kernel, grid, PCG_THREADS, (size_t) h->cg_smem, P, h->cg_max_blocks, matrix_free);
    cudaMemsetAsync(P.cg_partials + 8 * (size_t) h->cg_max_blocks, 0, 2 * sizeof(double), h->stream);
}
