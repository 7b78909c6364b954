// k_pcg.cu — the pressure Poisson solve: Jacobi-preconditioned conjugate gradient as ONE persistent
// cooperative kernel (one launch per solve; grid-wide barriers between the three sweeps of an
// iteration; no host round trips, so the small shipped cases are not launch-bound and the 4 M case
// streams the matrix at HBM rate).
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
// Per iteration, over n rows with ELL off-diagonals (fp64 value + int32 column):
//   sweep A  ap = A p, partial p.ap            12 B per stored entry + 4n (count) + 8n diag + 8n p + 8n ap
//   sweep B  x += a p; r -= a ap; |r|^2, r.z   read x,p,r,ap,invdiag, write x,r = 56n
//   sweep C  p = z + b p (z = invdiag r)       read r,invdiag,p, write p = 32n
// = 12 nnz + 108 n algorithmic bytes, the figure bench.py's roofline uses.
#include "mpsb_internal.cuh"

#include <cfloat>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

constexpr int PCG_THREADS = 256;

// Deterministic block-wide sum of up to two values; result broadcast to every thread.
__device__ __forceinline__ void block_sum2(double& a, double& b, double* sh /* 2*32 + 2 */) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    __syncthreads(); // protect sh from the previous use
    if (lane == 0) {
        sh[warp]      = a;
        sh[32 + warp] = b;
    }
    __syncthreads();
    if (warp == 0) {
        constexpr int W = PCG_THREADS / 32;
        double va = lane < W ? sh[lane] : 0.0, vb = lane < W ? sh[32 + lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            va += __shfl_xor_sync(0xffffffffu, va, o);
            vb += __shfl_xor_sync(0xffffffffu, vb, o);
        }
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

__global__ void __launch_bounds__(PCG_THREADS) pcg_kernel(Params P, int max_blocks, int iteration_cap) {
    cg::grid_group grid = cg::this_grid();
    __shared__ double sh[66];
    const int n       = *P.n_active;
    const int nblocks = gridDim.x;
    const int tid     = blockIdx.x * PCG_THREADS + threadIdx.x;
    const int stride  = nblocks * PCG_THREADS;
    const size_t N    = (size_t) P.N;
    double* __restrict__ x  = P.S.p;
    double* __restrict__ r  = P.cg_r;
    double* __restrict__ p  = P.cg_p;
    double* __restrict__ ap = P.cg_ap;
    double* __restrict__ invd = P.cg_invdiag;
    const double* __restrict__ diag = P.diag;
    const double* __restrict__ ell_val = P.ell_val;
    const int* __restrict__ ell_col = P.ell_col;
    const int* __restrict__ ell_count = P.ell_count;
    double* part0 = P.cg_partials;
    double* part1 = part0 + max_blocks;
    double* part2 = part1 + max_blocks;

    // setup: x = 0, r = b, invdiag, p = invdiag r; |b|^2 and r.p
    double l0 = 0.0, l1 = 0.0;
    for (int i = tid; i < P.N; i += stride) {
        x[i] = 0.0;
        if (i < n) {
            double bi = P.rhs[i], d = diag[i];
            double id = d != 0.0 ? 1.0 / d : 1.0;
            double pi = id * bi;
            r[i] = bi; invd[i] = id; p[i] = pi;
            l0 += bi * bi;
            l1 += bi * pi;
        }
    }
    block_sum2(l0, l1, sh);
    if (threadIdx.x == 0) {
        part0[blockIdx.x] = l0;
        part1[blockIdx.x] = l1;
    }
    grid.sync();
    double rhsNorm2, absNew;
    grid_sum2(part0, part1, nblocks, rhsNorm2, absNew, sh);
    int it = 0;
    double residualNorm2 = rhsNorm2;
    bool done = false;
    if (rhsNorm2 == 0.0) { // x = 0, iters = 0, error = 0 (ConjugateGradient.h:44-50)
        residualNorm2 = 0.0;
        done = true;
    }
    const double threshold = fmax(DBL_EPSILON * DBL_EPSILON * rhsNorm2, DBL_MIN);
    if (!done && residualNorm2 < threshold) done = true;
    grid.sync(); // part0/part1 are reused below
    const int maxIters = iteration_cap > 0 ? iteration_cap : 2 * P.N;

    while (!done && it < maxIters) {
        // ---- sweep A: ap = A p, p.ap ----
        double lpap = 0.0, dummy = 0.0;
        for (int row = tid; row < n; row += stride) {
            int cnt    = ell_count[row];
            double pr  = p[row];
            double sum = diag[row] * pr;
            int k = 0;
            for (; k + 4 <= cnt; k += 4) {
                int c0 = ell_col[(k + 0) * N + row], c1 = ell_col[(k + 1) * N + row];
                int c2 = ell_col[(k + 2) * N + row], c3 = ell_col[(k + 3) * N + row];
                double v0 = ell_val[(k + 0) * N + row], v1 = ell_val[(k + 1) * N + row];
                double v2 = ell_val[(k + 2) * N + row], v3 = ell_val[(k + 3) * N + row];
                double p0 = p[c0], p1 = p[c1], p2 = p[c2], p3 = p[c3];
                sum += v0 * p0; sum += v1 * p1; sum += v2 * p2; sum += v3 * p3;
            }
            for (; k < cnt; k++) sum += ell_val[k * N + row] * p[ell_col[k * N + row]];
            ap[row] = sum;
            lpap += pr * sum;
        }
        block_sum2(lpap, dummy, sh);
        if (threadIdx.x == 0) part0[blockIdx.x] = lpap;
        grid.sync();
        double pap, unused;
        grid_sum2(part0, part0, nblocks, pap, unused, sh);
        const double alpha = absNew / pap;
        // ---- sweep B: x += alpha p; r -= alpha ap; |r|^2; r.z ----
        double lrr = 0.0, lrz = 0.0;
        for (int i = tid; i < n; i += stride) {
            double pi = p[i], ri = r[i], api = ap[i], idg = invd[i];
            x[i] += alpha * pi;
            ri -= alpha * api;
            r[i] = ri;
            lrr += ri * ri;
            lrz += ri * (idg * ri);
        }
        block_sum2(lrr, lrz, sh);
        if (threadIdx.x == 0) {
            part1[blockIdx.x] = lrr;
            part2[blockIdx.x] = lrz;
        }
        grid.sync();
        double rz;
        grid_sum2(part1, part2, nblocks, residualNorm2, rz, sh);
        if (residualNorm2 < threshold) break;
        const double beta = rz / absNew;
        absNew = rz;
        // ---- sweep C: p = z + beta p ----
        for (int i = tid; i < n; i += stride) p[i] = invd[i] * r[i] + beta * p[i];
        it++;
        grid.sync();
    }
    if (tid == 0) {
        P.stats->cg_iterations   = it;
        P.stats->rhs_norm2       = rhsNorm2;
        P.stats->cg_rel_residual = rhsNorm2 > 0.0 ? sqrt(residualNorm2 / rhsNorm2) : 0.0;
    }
}

} // namespace

int mpsb_pcg_configure(mpsb_handle* h) {
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pcg_kernel, PCG_THREADS, 0);
    if (e != cudaSuccess || per_sm < 1) return -1;
    h->cg_threads    = PCG_THREADS;
    h->cg_max_blocks = per_sm * h->num_sms;
    return 0;
}

void mpsb_launch_pcg(mpsb_handle* h) {
    Params& P = h->P;
    // enough CTAs to cover the rows, never more than can be co-resident (cooperative launch)
    int want = mpsb_div_up(P.N, PCG_THREADS);
    int grid = want < h->cg_max_blocks ? want : h->cg_max_blocks;
    if (grid < 1) grid = 1;
    int max_blocks = h->cg_max_blocks;
    int cap = 0;
    void* args[] = {(void*) &P, (void*) &max_blocks, (void*) &cap};
    cudaLaunchCooperativeKernel((const void*) pcg_kernel, dim3(grid), dim3(PCG_THREADS), args, 0, h->stream);
    h->launches++;
}
