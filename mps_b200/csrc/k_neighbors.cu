// k_neighbors.cu — neighbour-list build and particle number density.
//
// Replaces reference MPS::setNeighbors (src/mps.cpp:585-619) and MPS::setNumberDensity /
// getNumberDensity (:573-583, :627-641). One thread owns one sorted particle and walks the 27
// cells in the reference's order (jx outer, jy, jz inner; ids ascending inside a cell), so the
// list ORDER and every distance are bit-identical and all later per-particle sums can follow
// list order. Lists are slot-major ELL: coalesced stores here and coalesced
// loads in every consumer.
#include "mpsb_internal.cuh"

namespace {

__global__ void __launch_bounds__(128) neighbor_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    int n_active = *P.n_active;
    if (s >= P.N) return;
    if (s >= n_active) {
        P.nbr_count[s] = 0;
        return;
    }
    const double xi = P.S.x[s], yi = P.S.y[s], zi = P.S.z[s];
    const double re = P.s.re_max;
    const double re2_out = (re * re) * (1.0 + 0x1p-50); // >= re^2 in exact arithmetic (the product errs by < 2^-53 relative)
    unsigned key = P.key[s];
    int nxy = P.numX * P.numY;
    int iz = key / nxy, rem = key - iz * nxy;
    int iy = rem / P.numX, ix = rem - iy * P.numX;
    int count = 0;
    // A domain of zero thickness in z (every 2-D case) keeps all non-ghost particles at z == zmin, i.e. in the
    // plane iz == 1 (anything else is out of the domain and a ghost): the planes iz +- 1 are empty, skip them.
    const int zspan = P.s.domain_max[2] == P.s.domain_min[2] ? 0 : 1;
    // The cells are walked in the reference's order as one rolled loop q = (jx, jy, jz) (27 unrolled copies of
    // the body would overflow the instruction cache). Two things keep a thread from waiting one memory round
    // trip per candidate: the range of cell q + 1 is loaded before cell q is processed, and the candidates of
    // a cell are loaded eight at a time (surplus slots clamped to the last particle of the cell and ignored).
    const int nz_cells = 2 * zspan + 1, ncell = 9 * nz_cells;
    auto cell_range = [&](int q, int& b, int& e) {
        const int dx = q / (3 * nz_cells), r = q - dx * (3 * nz_cells), dy = r / nz_cells, dz = r - dy * nz_cells;
        const int jx = ix - 1 + dx, jy = iy - 1 + dy, jz = iz - zspan + dz;
        b = 0;
        e = 0;
        if (jx >= 0 && jx < P.numX && jy >= 0 && jy < P.numY && jz >= 0 && jz < P.numZ) {
            const int c = jx + jy * P.numX + jz * nxy;
            b = P.cell_start[c];
            e = P.cell_end[c];
        }
    };
    int nb, ne;
    cell_range(0, nb, ne);
#pragma unroll 1
    for (int q = 0; q < ncell; q++) {
        const int b = nb, e = ne;
        if (q + 1 < ncell) cell_range(q + 1, nb, ne);
#pragma unroll 1
        for (int j0 = b; j0 < e; j0 += 8) {
            double X[8], Y[8], Z[8];
#pragma unroll
            for (int t = 0; t < 8; t++) {
                const int j = min(j0 + t, e - 1);
                X[t] = P.S.x[j];
                Y[t] = P.S.y[j];
                Z[t] = P.S.z[j];
            }
#pragma unroll
            for (int t = 0; t < 8; t++) {
                const int j = j0 + t;
                if (j >= e) break;
                double dx = X[t] - xi, dy = Y[t] - yi, dz = Z[t] - zi;
                double d2 = (dx * dx + dy * dy) + dz * dz;
                // 70 % of the candidates of a 3x3(x3) cell block are out of range: skip their square root.
                // d2 >= re2_out implies sqrt(d2) >= re exactly, hence RN(sqrt(d2)) >= re: the reference's
                // `dist < re` is false for them whatever the rounding, so the kept set is unchanged.
                if (d2 >= re2_out) continue;
                double dist = sqrt(d2); // Eigen (pj - pi).norm()
                if (j != s && dist < re) {
                    if (count < P.K) {
                        P.nbr_idx[SELL(count, s)]  = j;
                        P.nbr_dist[SELL(count, s)] = dist;
                    }
                    count++;
                }
            }
        }
    }
    if (count > P.K) {
        atomicOr(&P.stats->error_flags, 1);
        count = P.K;
    }
    P.nbr_count[s] = count;
    // longest list of the step (warp max first to keep atomics rare)
    int m = count;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(&P.stats->max_neighbors, m);
}

// ratio = Surface ? 0 : n(re_p) / n0_p   (src/mps.cpp:630-640), Ghosts included (their list is frozen).
__global__ void number_density_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= P.N) return;
    if (s >= *P.n_active) { // a Ghost: the sum over the list it had when it left (Params::ghost_ratio), 0 while flagged Surface
        P.S.ratio[s] = P.S.bc[s] == MPSB_BC_SURFACE ? 0.0 : P.ghost_ratio[P.S.id[s]];
        return;
    }
    double ratio = 0.0;
    if (P.S.bc[s] != MPSB_BC_SURFACE) {
        const double re = P.s.re_pressure;
        int cnt = P.nbr_count[s];
        double n = 0.0;
        for (int k0 = 0; k0 < cnt; k0 += 8) { // eight distance loads in flight, sum in list order
            double d[8];
#pragma unroll
            for (int b = 0; b < 8; b++) d[b] = P.nbr_dist[SELL(min(k0 + b, cnt - 1), s)];
#pragma unroll
            for (int b = 0; b < 8; b++)
                if (k0 + b < cnt && d[b] < re) n += mpsb_weight(d[b], re);
        }
        ratio = n / P.s.n0_pressure;
    }
    P.S.ratio[s] = ratio;
}

} // namespace

void mpsb_launch_neighbors(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, neighbor_kernel, mpsb_own_grid(h, 128), 128, 0, P);
}

namespace {
__global__ void ghost_ratio_from_home_kernel(Params P) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.N && P.H.type[i] == MPSB_GHOST) P.ghost_ratio[i] = P.H.ratio[i];
}
} // namespace

void mpsb_launch_ghost_ratio_from_home(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, ghost_ratio_from_home_kernel, mpsb_div_up(P.N, 256), 256, 0, P);
}

void mpsb_launch_number_density(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, number_density_kernel, mpsb_own_grid(h, 256), 256, 0, P);
}
