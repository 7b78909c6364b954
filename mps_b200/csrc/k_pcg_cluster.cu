// k_pcg_cluster.cu — the pressure Poisson solve for SMALL systems (every case the reference ships: 458 to
// 6 146 particles) as ONE thread-block cluster with the whole problem resident in shared memory.
//
// Same method and stopping rule as k_pcg.cu (Eigen's Jacobi-preconditioned CG, reference
// src/mps.cpp:417-470; ConjugateGradient.h:28-91), different machine mapping: at these sizes the
// solve is pure latency — 236 iterations x 3 dependent global reductions — and a grid-wide barrier through
// L2 (about 4 us) costs more than an iteration's arithmetic. Here
//   * the cluster's CTAs (8 or 16, one per SM of a GPC) each own a block of 32-row slices and copy their
//     part of the sliced-ELL matrix (fp64 values, 16-bit column offsets, diagonal) into shared memory ONCE;
//   * every CTA keeps a WINDOW of the search direction p — its own rows plus the rows its block's columns
//     reach (one cell row either side in the cell-major order) — and a window of z = D^-1 r; sweep B stores
//     each new z row into every window that holds it through distributed shared memory (DSMEM), and
//     p = z + beta p is then formed locally on the whole window, so no barrier is needed for p;
//   * reductions are per-warp partials stored into every CTA's shared memory and summed in one fixed order
//     (every thread of the cluster gets bit-identical scalars); the only synchronisation is the hardware
//     cluster barrier, TWO per iteration (the two dependent reductions of PCG).
// If the problem does not fit (shared-memory capacity, 16-bit offsets) the kernel leaves *cluster_done == 0
// and the persistent grid kernel of k_pcg.cu, launched right after it, does the solve; when it fits, that
// kernel sees *cluster_done == 1 and returns at once. No host round trip decides anything.
#include "mpsb_internal.cuh"

#include <cfloat>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

constexpr int CL_THREADS = 512;
constexpr int CL_WARPS   = CL_THREADS / 32;
constexpr int CL_MAX     = 16; // largest cluster tried

struct ClusterLayout { // byte offsets into the dynamic shared memory of one CTA
    int pwin, zwin, val, col, diag, invd, x, r, ap, slice_pap, off, total;
};

__device__ __forceinline__ int align16(int b) { return (b + 15) & ~15; }

// window: rows of p this CTA reads (its own rows and the columns of its block); own_slices: its slices;
// entries: its stored slots * 32
__device__ __forceinline__ ClusterLayout make_layout(int window, int own_slices, int entries) {
    ClusterLayout L;
    int b = 0;
    L.pwin = b; b = align16(b + window * 8);
    L.zwin = b; b = align16(b + window * 8);
    L.val = b; b = align16(b + entries * 8);
    L.col = b; b = align16(b + entries * 2);
    const int own_rows = own_slices * 32;
    L.diag = b; b = align16(b + own_rows * 8);
    L.invd = b; b = align16(b + own_rows * 8);
    L.x = b; b = align16(b + own_rows * 8);
    L.r = b; b = align16(b + own_rows * 8);
    L.ap = b; b = align16(b + own_rows * 8);
    L.slice_pap = b; b = align16(b + own_slices * 8);
    L.off = b; b = align16(b + (own_slices + 1) * 4);
    L.total = b;
    return L;
}

// CTA-wide (sum, min, max) of three ints, result in every thread.
__device__ __forceinline__ void cta_sum_min_max(int& sum, int& mn, int& mx, int* shi /* 3 * CL_WARPS */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __syncthreads();
    if (lane == 0) {
        shi[warp]                = sum;
        shi[CL_WARPS + warp]     = mn;
        shi[2 * CL_WARPS + warp] = mx;
    }
    __syncthreads();
    sum = 0;
    for (int w = 0; w < CL_WARPS; w++) {
        sum += shi[w];
        mn = min(mn, shi[CL_WARPS + w]);
        mx = max(mx, shi[2 * CL_WARPS + w]);
    }
}

__global__ void __launch_bounds__(CL_THREADS, 1) pcg_cluster_kernel(Params P, int smem_bytes, int iteration_cap) {
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ int shi[3 * CL_WARPS];
    // [exchange parity][source rank][source warp][value]: every warp of every CTA stores its partial here, in
    // every CTA, through DSMEM; summed by every warp in the same fixed order after the cluster barrier
    __shared__ double red[2][CL_MAX][CL_WARPS][2];
    __shared__ int meta[CL_MAX][3]; // per rank: does-not-fit flag, window [lo, hi); written through DSMEM

    const int C = (int) cluster.num_blocks(), rank = (int) cluster.block_rank();
    if (rank == 0 && threadIdx.x == 0) *P.cluster_done = 0; // set again at the end if this kernel does the solve
    unsigned long long t_begin;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = *P.n_active;
    const int slices = (n + 31) >> 5;
    const int per = (slices + C - 1) / C;
    const int s0 = min(rank * per, slices), s1 = min(s0 + per, slices), own = s1 - s0;
    const int row0 = s0 * 32, own_rows = own * 32;

    // ---- size of the block and the window of p it reads; does it fit? all CTAs agree through DSMEM
    int ent = 0, wlo = row0, whi = row0 + own_rows - 1;
    for (int s = s0 + (int) threadIdx.x; s < s1; s += CL_THREADS) {
        ent += P.ell_width[s];
        wlo = min(wlo, P.ell_cmin[s]);
        whi = max(whi, P.ell_cmax[s]);
    }
    cta_sum_min_max(ent, wlo, whi, shi);
    wlo &= ~31;                             // whole slices: rows of a slice beyond n are padding with p = 0
    whi = own > 0 ? ((whi + 32) & ~31) : wlo; // exclusive
    const int window = whi - wlo, entries = ent * 32;
    const ClusterLayout L = make_layout(window, own, entries);
    const int bad = (L.total > smem_bytes || *P.ell_wide != 0 || n > 32767) ? 1 : 0;
    if (threadIdx.x < (unsigned) C) {
        int* m = cluster.map_shared_rank(&meta[rank][0], threadIdx.x);
        m[0] = bad;
        m[1] = wlo;
        m[2] = whi;
    }
    cluster.sync();
    int any_bad = 0;
    for (int q = 0; q < C; q++) any_bad |= meta[q][0];
    if (any_bad) return; // uniform over the cluster; *cluster_done stays 0 -> the grid kernel solves

    double* pwin = reinterpret_cast<double*>(dyn + L.pwin) - wlo; // indexed by global row
    double* zwin = reinterpret_cast<double*>(dyn + L.zwin) - wlo;
    double* val  = reinterpret_cast<double*>(dyn + L.val);
    short* col   = reinterpret_cast<short*>(dyn + L.col);
    double* diag = reinterpret_cast<double*>(dyn + L.diag);
    double* invd = reinterpret_cast<double*>(dyn + L.invd);
    double* x    = reinterpret_cast<double*>(dyn + L.x);
    double* r    = reinterpret_cast<double*>(dyn + L.r);
    double* ap   = reinterpret_cast<double*>(dyn + L.ap);
    double* spap = reinterpret_cast<double*>(dyn + L.slice_pap);
    int* off     = reinterpret_cast<int*>(dyn + L.off);
    // the p / z windows of CTA q, indexed by global row (valid inside q's window [meta[q][1], meta[q][2]) only);
    // every CTA's layout starts with pwin at offset 0 and zwin right after a window of ITS OWN size
    auto peer_p = [&](int q) -> double* { return reinterpret_cast<double*>(cluster.map_shared_rank(dyn, q)) - meta[q][1]; };
    auto peer_z = [&](int q) -> double* {
        return reinterpret_cast<double*>(cluster.map_shared_rank(dyn, q) + align16((meta[q][2] - meta[q][1]) * 8)) - meta[q][1];
    };

    // ---- load the block: slot offsets (serial prefix over a few slices), then the matrix
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int s = 0; s < own; s++) {
            off[s] = acc;
            acc += P.ell_width[s0 + s] * 32;
        }
        off[own] = acc;
    }
    for (int j = wlo + (int) threadIdx.x; j < whi; j += CL_THREADS) { // rows nobody owns (beyond n) stay 0
        pwin[j] = 0.0;
        zwin[j] = 0.0;
    }
    cluster.sync(); // windows are cleared everywhere before any CTA stores into them
    for (int sl = warp; sl < own; sl += CL_WARPS) {
        const int s = s0 + sl, w = (off[sl + 1] - off[sl]) >> 5;
        const size_t g = (size_t) s * (size_t) P.K * 32u;
        for (int k = 0; k < w; k++) {
            val[off[sl] + k * 32 + lane] = P.ell_val[g + (size_t) k * 32 + lane];
            col[off[sl] + k * 32 + lane] = P.ell_col16[g + (size_t) k * 32 + lane];
        }
    }
    // store v as row `row` of p (which = 0) or z (which = 1) in every CTA whose window holds that row
    // (windows grow with the rank, so the destinations of a row are a contiguous range of ranks; the range of
    // the thread's first row is kept in registers, further rows of large blocks recompute it)
    auto dest_range = [&](int row, int& qa, int& qb) {
        qa = C;
        qb = -1;
        for (int q = 0; q < C; q++)
            if (row >= meta[q][1] && row < meta[q][2]) {
                qa = min(qa, q);
                qb = max(qb, q);
            }
    };
    int my_qa, my_qb;
    dest_range(row0 + (int) threadIdx.x, my_qa, my_qb);
    auto broadcast = [&](int row, double v, int which) {
        int qa = my_qa, qb = my_qb;
        if (row != row0 + (int) threadIdx.x) dest_range(row, qa, qb);
        for (int q = qa; q <= qb; q++)
            if (row >= meta[q][1] && row < meta[q][2]) (which ? peer_z(q) : peer_p(q))[row] = v;
    };
    // setup: x = 0, r = b, invdiag, p = invdiag r (0 beyond n); |b|^2 and r.p
    double l0 = 0.0, l1 = 0.0;
    for (int i = threadIdx.x; i < own_rows; i += CL_THREADS) {
        const int row = row0 + i;
        double d = 0.0, id = 1.0, bi = 0.0;
        if (row < n) {
            bi = P.rhs[row];
            d  = P.diag[row];
            id = d != 0.0 ? 1.0 / d : 1.0;
            const double pi = id * bi;
            l0 += bi * bi;
            l1 += bi * pi;
            broadcast(row, pi, 0);
        }
        diag[i] = d; invd[i] = id; x[i] = 0.0; r[i] = bi; ap[i] = 0.0;
    }
    // particle slots that are not rows of the system keep pressure 0 (as the grid kernel's setup does)
    for (int i = n + rank * CL_THREADS + (int) threadIdx.x; i < P.N; i += C * CL_THREADS) P.S.p[i] = 0.0;

    int parity = 0;
    // All-CTA sum of (a, b): every warp reduces its lanes and stores the pair into red[parity][rank][warp] of
    // EVERY CTA (lane q stores to rank q); after the cluster barrier every warp adds the C x 16 pairs in the
    // same fixed order (lane-strided, then the shuffle tree) -> bit-identical scalars in every thread of the
    // cluster, and a single barrier per reduction. The barrier also publishes the window stores made before it.
    // (Measured alternative: per-warp remote mbarrier arrivals instead of barrier.cluster — 25 % slower.)
    auto exchange = [&](double& a, double& b) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane < C) {
            double* dst = cluster.map_shared_rank(&red[parity][rank][warp][0], lane);
            dst[0] = a;
            dst[1] = b;
        }
        cluster.sync();
        double sa = 0.0, sb = 0.0;
        const double* src = &red[parity][0][0][0];
        for (int e = lane; e < C * CL_WARPS; e += 32) {
            sa += src[2 * e];
            sb += src[2 * e + 1];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sa += __shfl_xor_sync(0xffffffffu, sa, o);
            sb += __shfl_xor_sync(0xffffffffu, sb, o);
        }
        a = sa;
        b = sb;
        parity ^= 1;
    };
    exchange(l0, l1); // also publishes the p windows and orders the matrix load before the first sweep
    const double rhsNorm2 = l0;
    double absNew = l1, residualNorm2 = rhsNorm2;
    bool done = false;
    if (rhsNorm2 == 0.0) { // x = 0, iters = 0, error = 0 (ConjugateGradient.h:44-50)
        residualNorm2 = 0.0;
        done = true;
    }
    const double threshold = fmax(DBL_EPSILON * DBL_EPSILON * rhsNorm2, DBL_MIN);
    if (!done && residualNorm2 < threshold) done = true;
    const int maxIters = iteration_cap > 0 ? iteration_cap : 2 * P.N;
    int it = 0;
    long long cyc[4] = {0, 0, 0, 0}, c0 = clock64(), c1;
#define CL_TICK(slot) do { c1 = clock64(); cyc[slot] += c1 - c0; c0 = c1; } while (0)
    while (!done && it < maxIters) {
        // ---- sweep A: ap = A p; p.ap per slice
        for (int sl = warp; sl < own; sl += CL_WARPS) {
            const int base = (s0 + sl) * 32, w = (off[sl + 1] - off[sl]) >> 5, o = off[sl];
            const double* pb = pwin + base;
            const double pr  = pb[lane];
            double sum = diag[sl * 32 + lane] * pr;
            for (int k = 0; k < w; k += 4) { // widths are multiples of 4: four gathers in flight, adds in slot order
                const int k0 = col[o + (k + 0) * 32 + lane], k1 = col[o + (k + 1) * 32 + lane];
                const int k2 = col[o + (k + 2) * 32 + lane], k3 = col[o + (k + 3) * 32 + lane];
                const double p0 = pb[k0], p1 = pb[k1], p2 = pb[k2], p3 = pb[k3];
                const double v0 = val[o + (k + 0) * 32 + lane], v1 = val[o + (k + 1) * 32 + lane];
                const double v2 = val[o + (k + 2) * 32 + lane], v3 = val[o + (k + 3) * 32 + lane];
                sum += v0 * p0; sum += v1 * p1; sum += v2 * p2; sum += v3 * p3;
            }
            ap[sl * 32 + lane] = sum;
            double v = pr * sum;
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
            if (lane == 0) spap[sl] = v;
        }
        __syncthreads();
        CL_TICK(0);
        double lpap = 0.0, unused = 0.0;
        for (int sl = threadIdx.x; sl < own; sl += CL_THREADS) lpap += spap[sl];
        exchange(lpap, unused);
        const double alpha = absNew / lpap;
        CL_TICK(3);
        // ---- sweep B: x += alpha p; r -= alpha ap; |r|^2; r.z; z = invdiag r goes to every window holding the row
        double lrr = 0.0, lrz = 0.0;
        for (int i = threadIdx.x; i < own_rows; i += CL_THREADS) {
            const int row = row0 + i;
            if (row < n) {
                const double ri = r[i] - alpha * ap[i];
                x[i] += alpha * pwin[row];
                r[i] = ri;
                const double zi = invd[i] * ri;
                lrr += ri * ri;
                lrz += ri * zi;
                broadcast(row, zi, 1);
            }
        }
        CL_TICK(1);
        exchange(lrr, lrz); // its barrier publishes the z windows
        residualNorm2 = lrr;
        if (residualNorm2 < threshold) break;
        const double beta = lrz / absNew;
        absNew = lrz;
        CL_TICK(3);
        // ---- sweep C: p = z + beta p on the whole window, locally (no third barrier)
        for (int j = wlo + (int) threadIdx.x; j < whi; j += CL_THREADS) pwin[j] = zwin[j] + beta * pwin[j];
        it++;
        __syncthreads();
        CL_TICK(2);
    }
    for (int i = threadIdx.x; i < own_rows; i += CL_THREADS)
        if (row0 + i < n) P.S.p[row0 + i] = x[i];
    if (rank == 0 && threadIdx.x == 0) {
        for (int k = 0; k < 4; k++) P.stats->cg_cycles[k] = cyc[k];
        P.stats->cg_iterations   = it;
        P.stats->pad[0]          = 1;
        P.stats->pad[1]          = 1;
        P.stats->rhs_norm2       = rhsNorm2;
        P.stats->cg_rel_residual = rhsNorm2 > 0.0 ? sqrt(residualNorm2 / rhsNorm2) : 0.0;
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        P.stats->t_solve_begin   = t_begin;
        P.stats->t_solve_end     = t_end;
        *P.cluster_done          = 1;
    }
    cluster.sync(); // no CTA exits while its shared memory may still be addressed by a peer
}

} // namespace

// Picks the cluster size this device can co-schedule with the full shared-memory carve-out. 0 = unavailable.
int mpsb_pcg_cluster_configure(mpsb_handle* h) {
    h->cl_size = 0;
    h->cl_smem = 0;
    int max_optin = 0;
    if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device) != cudaSuccess) return 0;
    const int smem = max_optin - 12 * 1024; // the kernel's static shared memory (8.4 KB) + a margin
    if (cudaFuncSetAttribute(pcg_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return 0;
    cudaFuncSetAttribute(pcg_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int c : {16, 8, 4, 2}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim          = dim3(c);
        cfg.blockDim         = dim3(CL_THREADS);
        cfg.dynamicSmemBytes = (size_t) smem;
        cudaLaunchAttribute at[1];
        at[0].id               = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = c;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs    = at;
        cfg.numAttrs = 1;
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, pcg_cluster_kernel, &cfg) == cudaSuccess && nclusters >= 1) {
            h->cl_size = c;
            h->cl_smem = smem;
            break;
        }
    }
    cudaGetLastError(); // a rejected size is not an error of the handle
    return h->cl_size;
}

// Launches the cluster solve when the case is small enough to have a chance of fitting; returns 1 if launched.
int mpsb_launch_pcg_cluster(mpsb_handle* h) {
    Params& P = h->P;
    // *cluster_done is 0 unless this kernel solved: it clears the flag itself when it starts (no memset node)
    if (h->cl_size <= 0 || P.L.world > 1 || P.N > 32767 || h->cl_disable) {
        if (h->cl_launched) { // an earlier solve may have left the flag set (e.g. a communicator attached since)
            cudaMemsetAsync(P.cluster_done, 0, sizeof(int), h->stream);
            h->cl_launched = 0;
        }
        return 0;
    }
    h->cl_launched = 1;
    // cluster size: the smallest power of two whose CTAs are expected to hold their block (rows of the widest
    // lists seen in practice), so tiny systems do not pay a 16-CTA barrier; the kernel itself checks the fit
    const int slices = (P.N + 31) / 32, width = P.s.dim == 2 ? 36 : 128;
    int csize = 1;
    while (csize < h->cl_size) {
        const int per = (slices + csize - 1) / csize;
        const int win    = per + 64 < slices ? per + 64 : slices; // own slices + the cell rows either side
        const long bytes = (long) win * 32 * 16 + (long) per * 32 * ((long) width * 10 + 48) + 1024;
        if (bytes <= h->cl_smem) break;
        csize *= 2;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3(csize);
    cfg.blockDim         = dim3(CL_THREADS);
    cfg.dynamicSmemBytes = (size_t) h->cl_smem;
    cfg.stream           = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id               = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = csize;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs    = at;
    cfg.numAttrs = 1;
    int smem = h->cl_smem, cap = h->cg_iteration_cap;
    cudaLaunchKernelEx(&cfg, pcg_cluster_kernel, P, smem, cap);
    h->launches++;
    return 1;
}
