// k_pressure.cu — everything of calcPressure (src/mps.cpp:252-260) except the Krylov solve:
//   flags_kernel / source_ho_kernel : setBoundaryCondition + setSourceTerm      (:262-320)
//   matrix_kernel                   : setMatrix, straight from the neighbour lists into ELL (:322-360)
//   dirichlet_*                     : ensureDirichletBoundaryConnection as connected components (:362-415)
//   clamp_kernel / min_pressure_kernel : removeNegativePressure + setMinimumPressure (:472-503)
#include "mpsb_internal.cuh"

#include <cstdlib>

namespace {

// Boundary-condition flag for every slot and, for the standard source term, b_i in the same pass
// over the distances (re_s and re_p sums are kept separate so differing radii stay exact).
__global__ void __launch_bounds__(256) flags_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= P.N) return;
    int bc = MPSB_BC_IGNORED;
    double b = 0.0;
    if (s < *P.n_active) {
        int type = P.S.type[s];
        if (type != MPSB_GHOST && type != MPSB_DUMMY_WALL) {
            const double re_s = P.s.re_surface, re_p = P.s.re_pressure;
            int cnt = P.nbr_count[s];
            double ns = 0.0, np = 0.0;
            for (int k0 = 0; k0 < cnt; k0 += 8) { // distances only: eight loads in flight, sums in list order
                double d[8];
#pragma unroll
                for (int b = 0; b < 8; b++) d[b] = P.nbr_dist[SELL(min(k0 + b, cnt - 1), s)];
#pragma unroll
                for (int b = 0; b < 8; b++)
                    if (k0 + b < cnt) {
                        if (d[b] < re_s) ns += (re_s / d[b]) - 1.0;
                        if (d[b] < re_p) np += (re_p / d[b]) - 1.0;
                    }
            }
            if (ns < P.beta_n0) {
                bc = MPSB_BC_SURFACE;
            } else {
                bc = MPSB_BC_INNER;
                // gamma * (1/(dt*dt)) * (n - n0) / n0      (src/mps.cpp:313)
                b = P.src_std * (np - P.s.n0_pressure) / P.s.n0_pressure;
            }
        }
    }
    P.S.bc[s] = bc;
    P.rhs[s]  = b;
    if (bc == MPSB_BC_INNER) atomicAdd(&P.stats->n_inner, 1);
    if (bc == MPSB_BC_SURFACE) atomicAdd(&P.stats->n_surface, 1);
}

// Higher-order source term (src/mps.cpp:288-307); needs the neighbours' flags, hence a second pass.
__global__ void __launch_bounds__(256) source_ho_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.bc[s] != MPSB_BC_INNER) return;
    const double re = P.s.re_pressure;
    const double xi = P.S.x[s], yi = P.S.y[s], zi = P.S.z[s];
    const double ui = P.S.u[s], vi = P.S.v[s], wi = P.S.w[s];
    double b = P.src_ho0 * (P.S.ratio[s] - 1.0);
    int cnt = P.nbr_count[s];
    struct Nb { double x, y, z, u, v, w; int bc; };
    mpsb_walk_list<4, Nb>(
        P, s, cnt, [&](int j) { return Nb{P.S.x[j], P.S.y[j], P.S.z[j], P.S.u[j], P.S.v[j], P.S.w[j], P.S.bc[j]}; },
        [&](int, double d, int, const Nb& q) {
            if (q.bc == MPSB_BC_IGNORED) return;
            if (d < re) {
                double rx = q.x - xi, ry = q.y - yi, rz = q.z - zi;
                double ux = q.u - ui, uy = q.v - vi, uz = q.w - wi;
                double ru = (rx * ux + ry * uy) + rz * uz;
                b -= P.src_ho1 * ru / (d * d * d);
            }
        });
    P.rhs[s] = b;
}

// Row s of the PPE: diagonal kept apart, off-diagonals (Inner-Inner only) compacted into the slots of
// the row's 32-row slice. a = (2D/(lambda n0)) * w / rho (src/mps.cpp:343); Surface neighbours add to the
// diagonal only. Every slice is padded to a common width (the longest row of the slice, rounded up to a
// multiple of 4) with entries (column = own row, value = +0.0) so the PCG kernel streams whole slices
// without per-row trip counts; adding +0.0 * p never changes a finite sum. One warp = one slice, so
// the grid covers Np rows and no thread exits before the warp-wide maximum.
__global__ void __launch_bounds__(256) matrix_kernel(Params P, int force_wide) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s == P.row_lo && force_wide) *P.ell_wide = 1;
    if (s >= P.Np) return; // Np is a multiple of 32: whole warps leave together
    int m = 0, cmin = s, cmax = s, wide = 0;
    double diag = 0.0;
    if (s < *P.n_active && P.S.bc[s] == MPSB_BC_INNER) {
        const double re = P.s.re_pressure;
        int cnt = P.nbr_count[s];
        mpsb_walk_list<8, int>(
            P, s, cnt, [&](int j) { return P.S.bc[j]; },
            [&](int, double d, int j, const int& bcj) {
                if (d < re) {
                    if (bcj == MPSB_BC_IGNORED) return;
                    double a = P.mat_coef * ((re / d) - 1.0) / P.s.density;
                    diag += a;
                    if (bcj == MPSB_BC_INNER) {
                        P.ell_col[SELL(m, s)] = j;
                        P.ell_val[SELL(m, s)] = -a;
                        const int off = j - (s & ~31); // 16-bit form: offset from the slice's first row
                        P.ell_col16[SELL(m, s)] = (short) off;
                        if (off < -32768 || off > 32767) wide = 1;
                        cmin = min(cmin, j);
                        cmax = max(cmax, j);
                        m++;
                    }
                }
            });
        if (P.s.quasi_compressibility_on) diag += P.qc_term;
        atomicAdd(&P.stats->nnz, m + 1);
    }
    int width = m;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        width = max(width, __shfl_xor_sync(0xffffffffu, width, o));
        cmin  = min(cmin, __shfl_xor_sync(0xffffffffu, cmin, o));
        cmax  = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
    }
    width = min((width + 3) & ~3, P.K & ~3);
    for (int k = m; k < width; k++) {
        P.ell_col[SELL(k, s)]   = s;
        P.ell_col16[SELL(k, s)] = (short) (s & 31);
        P.ell_val[SELL(k, s)]   = 0.0;
    }
    if (wide) *P.ell_wide = 1;
    if ((threadIdx.x & 31) == 0) {
        P.ell_width[s >> 5] = width;
        P.ell_cmin[s >> 5]  = cmin; // column extent of the slice: the halo a row block needs (k_pcg.cu)
        P.ell_cmax[s >> 5]  = cmax;
    }
    P.ell_count[s] = m;
    P.diag[s]      = diag;
    P.row_xy[s]    = make_double2(P.S.x[s], P.S.y[s]); // rows >= N: zero-filled allocation, never a column
}

// ---- Dirichlet connectivity ------------------------------------------------------------------------
// The reference runs a BFS from every Surface particle over the r < re_p graph through ALL
// non-ghost particles (DummyWall included) and doubles the diagonal of Inner particles it never
// reaches. The neighbour relation is symmetric, so "reached" == "in a connected component that
// holds a Surface particle": lock-free union-find (hook larger root under smaller, CAS) + marking.
__device__ __forceinline__ bool in_extent(const Params& P, int s) {
    if (P.L.world <= 1) return true;
    const RankExtents& E = *P.ext;
    const int q = P.L.rank;
    return (s >= E.own_lo[q] && s < E.own_hi[q]) || (s >= E.ext_lo[q] && s < E.ext_hi[q]);
}

__device__ __forceinline__ int uf_find(int* parent, int x) {
    int p = parent[x];
    while (p != x) {
        int gp = parent[p];
        if (gp != p) parent[x] = gp; // path halving (benign race: only ever points closer to a root)
        x = p;
        p = gp;
    }
    return x;
}

__global__ void dirichlet_init_kernel(Params P) {
    int s = blockIdx.x * blockDim.x + threadIdx.x; // rows of this rank's extent: own + halo (one GPU: all rows)
    if (s < P.N && !in_extent(P, s)) return;
    if (s >= P.N) return;
    P.parent[s] = s;
    P.mark[s]   = 0;
}

__global__ void __launch_bounds__(256) dirichlet_union_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active) return;
    const double re = P.s.re_pressure;
    int cnt = P.nbr_count[s];
    mpsb_walk_list<8, MpsbNoRec>(
        P, s, cnt, [](int) { return MpsbNoRec{}; },
        [&](int, double d, int j, const MpsbNoRec&) {
            // each undirected edge once; an edge to a halo row above this rank's block belongs to nobody else here
            if (!(d < re) || (j > s && j < P.row_hi)) return;
            int a = uf_find(P.parent, s), b = uf_find(P.parent, j);
            while (a != b) {
                if (a < b) { int t = a; a = b; b = t; }
                int old = atomicCAS(&P.parent[a], a, b);
                if (old == a) break;
                a = uf_find(P.parent, old);
                b = uf_find(P.parent, b);
            }
        });
}

__global__ void dirichlet_mark_kernel(Params P) {
    int s = blockIdx.x * blockDim.x + threadIdx.x; // rows of this rank's extent: own + halo (one GPU: all rows)
    if (s < P.N && !in_extent(P, s)) return;
    if (s >= *P.n_active) return;
    if (P.S.bc[s] == MPSB_BC_SURFACE) P.mark[uf_find(P.parent, s)] = 1;
}

__global__ void dirichlet_apply_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.bc[s] != MPSB_BC_INNER) return;
    if (!P.mark[uf_find(P.parent, s)]) {
        P.diag[s] *= 2.0; // coefficientMatrix.coeffRef(i, i) *= 2.0 (src/mps.cpp:412)
        int e = atomicAdd(&P.stats->n_unconnected, 1);
        if (e < P.ev_cap) P.ev_uncon_id[e] = P.S.id[s];
    }
}

// Several ranks (slab decomposition): a rank's union-find covers the edges incident to its OWN rows, with its halo
// rows as extra nodes, so a component that reaches a Surface particle only through another rank's rows is not
// marked yet. Marks are propagated to a fixpoint: every owner stores, for each own row that is some peer's halo
// row, whether the row's local component is marked into that peer's `reach` array; after the cross-rank barrier
// every rank folds the reach flags of its halo rows into its local components. Each edge of the global graph
// lies in the local graph of the owner(s) of its end points and neighbouring local graphs share those end points,
// so the fixpoint is the reference's answer (reached from a Surface particle <=> component holds one). Rounds are
// driven by the host (one 2-int all-reduce each): typically one per slab between a row and the free surface.
__global__ void __launch_bounds__(256) dirichlet_reach_push_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.row_hi || s >= *P.n_active) return;
    const RankExtents& E = *P.ext;
    bool boundary = false;
    for (int q = 0; q < P.L.world; q++) boundary |= q != P.L.rank && s >= E.ext_lo[q] && s < E.ext_hi[q];
    const bool inner = P.S.bc[s] == MPSB_BC_INNER;
    if (!boundary && !inner) return;
    const int v = P.mark[uf_find(P.parent, s)];
    if (inner && !v) P.dflags[1] = 1; // still unresolved somewhere: another round may be needed
    if (!boundary) return;
    for (int q = 0; q < P.L.world; q++)
        if (q != P.L.rank && s >= E.ext_lo[q] && s < E.ext_hi[q]) mpsb_peer(P, P.reach, q)[s] = v;
}

__global__ void __launch_bounds__(256) dirichlet_absorb_kernel(Params P) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= *P.n_active || (s >= P.row_lo && s < P.row_hi) || !in_extent(P, s)) return;
    if (P.reach[s]) {
        const int r = uf_find(P.parent, s);
        if (!P.mark[r]) {
            P.mark[r]   = 1;
            P.dflags[0] = 1;
        }
    }
}

// ---- after the solve -----------------------------------------------------------------------------------
__global__ void clamp_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= P.N) return;
    if (P.S.p[s] < 0.0) P.S.p[s] = 0.0; // NaN passes through, like the reference (:475)
}

__global__ void __launch_bounds__(256) min_pressure_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.bc[s] == MPSB_BC_IGNORED) return;
    const double re = P.s.re_pressure;
    double m = P.S.p[s];
    int cnt  = P.nbr_count[s];
    struct Nb { double p; int type; };
    mpsb_walk_list<8, Nb>(
        P, s, cnt, [&](int j) { return Nb{P.S.p[j], P.S.type[j]}; },
        [&](int, double d, int, const Nb& q) {
            if (d < re) {
                if (q.type == MPSB_DUMMY_WALL) return;
                if (m > q.p) m = q.p;
            }
        });
    P.S.minp[s] = m;
}

} // namespace

void mpsb_launch_flags_source(mpsb_handle* h) {
    Params& P = h->P;
    int grid  = mpsb_own_grid(h, 256);
    if (!h->in_step) cudaMemsetAsync(&P.stats->n_inner, 0, 2 * sizeof(int), h->stream); // n_inner, n_surface (a step clears all stats at once)
    MPSB_LAUNCH(h, flags_kernel, grid, 256, 0, P);
    if (h->world > 1) { // the assembly, the higher-order source term and the Dirichlet check read the neighbours' flags
        mpsb_launch_halo_push_bc(h);
        mpsb_launch_xbarrier(h);
    }
    if (P.s.higher_order_source_on) MPSB_LAUNCH(h, source_ho_kernel, grid, 256, 0, P);
}

void mpsb_launch_matrix(mpsb_handle* h) {
    Params& P = h->P;
    if (!h->in_step) cudaMemsetAsync(&P.stats->nnz, 0, sizeof(int), h->stream);
    const char* force = getenv("MPSB_FORCE_WIDE_COLUMNS"); // tests: exercise the 32-bit column stream
    // *ell_wide is cleared by the solve that consumed it (pcg_kernel), not by a memset node here
    MPSB_LAUNCH(h, matrix_kernel, mpsb_div_up(P.row_hi - P.row_lo, 256), 256, 0, P, force && atoi(force) ? 1 : 0);
}

void mpsb_launch_dirichlet(mpsb_handle* h) {
    Params& P = h->P;
    int grid = mpsb_own_grid(h, 256), grid_all = mpsb_div_up(P.N, 256);
    if (!h->in_step) cudaMemsetAsync(&P.stats->n_unconnected, 0, sizeof(int), h->stream);
    MPSB_LAUNCH(h, dirichlet_init_kernel, grid_all, 256, 0, P);
    MPSB_LAUNCH(h, dirichlet_union_kernel, grid, 256, 0, P);
    MPSB_LAUNCH(h, dirichlet_mark_kernel, grid_all, 256, 0, P);
    if (h->world > 1) {
        for (int round = 0; round < 4096; round++) {
            cudaMemsetAsync(P.dflags, 0, 2 * sizeof(int), h->stream);
            MPSB_LAUNCH(h, dirichlet_reach_push_kernel, grid, 256, 0, P);
            mpsb_launch_xbarrier(h);
            MPSB_LAUNCH(h, dirichlet_absorb_kernel, grid_all, 256, 0, P);
            int flags[2] = {0, 0};
            if (mpsb_comm_allreduce_dflags(h, flags)) return; // sticky error set
            if (!flags[0] || !flags[1]) break; // no mark changed anywhere, or every Inner row everywhere is connected
        }
    }
    MPSB_LAUNCH(h, dirichlet_apply_kernel, grid, 256, 0, P);
}

void mpsb_launch_pressure_post(mpsb_handle* h) {
    Params& P = h->P;
    int grid  = mpsb_own_grid(h, 256);
    MPSB_LAUNCH(h, clamp_kernel, grid, 256, 0, P);
    if (h->world > 1) { // minimum pressure and the gradient read the neighbours' (clamped) pressures
        mpsb_launch_halo_push_p(h);
        mpsb_launch_xbarrier(h);
    }
    MPSB_LAUNCH(h, min_pressure_kernel, grid, 256, 0, P);
}
