// k_neighbors.cu — neighbour-list build and particle number density.
//
// Replaces reference MPS::setNeighbors (src/mps.cpp:585-619) and MPS::setNumberDensity /
// getNumberDensity (:573-583, :627-641). One thread owns one sorted particle and walks the 27
// cells in the reference's order (jx outer, jy, jz inner; ids ascending inside a cell), so the
// list ORDER and every distance are bit-identical and all later per-particle sums can follow
// list order. Lists are slot-major ELL: coalesced stores here and coalesced
// loads in every consumer.
#include "mpsb_internal.cuh"

#include <cstdlib>

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

// ---- tiled build: neighbour-cell rows staged in shared memory by TMA bulk copies -------------------------------
// One CTA owns NT_CX consecutive cells of one cell row (iy, iz). SORTED order is cell-major with x fastest, so
//   * the particles of the tile are ONE contiguous range of rows, and
//   * for each of the 3 (2-D) or 9 (3-D) neighbouring cell rows (jy, jz), the candidates of EVERY particle of the
//     tile — cells c0 - 1 .. c0 + NT_CX of that row — are one contiguous range as well.
// Lane 0 issues one 1-D bulk copy (cp.async.bulk ... mbarrier::complete_tx::bytes) per coordinate array and
// neighbouring row (9 or 27 per tile) into shared memory, the CTA waits on the mbarrier once, and each thread then
// walks its particle's 9 / 27 cells in the reference's order (jx outer, jy, jz inner; rows ascending inside a cell
// = ids ascending) with every candidate coordinate coming from shared memory: the per-thread kernel read each
// candidate ~9 times per cell through L1/L2 (88 % L1 hits, but a 1.4 ms latency chain at 4 M particles); here HBM/L2
// is touched once per tile and row. Arithmetic, skip test, order and stores are those of neighbor_kernel, so the lists
// stay bit-identical (tests/test_gpu_parity.py compares ids, order and distances with the reference).
// A neighbouring row range that does not fit its shared-memory slot (NT_CAP particles: denser than ~32 per cell) is
// read from global memory instead — same loop, pointer switched — so there is no capacity limit.
constexpr int NT_CX      = 16;  // cells per tile along x
constexpr int NT_THREADS = 128;

__device__ __forceinline__ unsigned nt_smem_u32(const void* p) { return (unsigned) __cvta_generic_to_shared(p); }

template <int NROWS, int NT_CAP> // NROWS: 3 (flat z) or 9; NT_CAP: staged particles per neighbouring row (even)
__global__ void __launch_bounds__(NT_THREADS) neighbor_tile_kernel(Params P, int tiles_x) {
    constexpr int NZ = NROWS / 3;       // cell planes walked in z
    constexpr int SLOT = NT_CAP + 2;    // doubles per staged array (room for the 16-byte alignment of both ends)
    extern __shared__ __align__(16) unsigned char nt_raw[];
    double* sxyz = reinterpret_cast<double*>(nt_raw);                       // [NROWS][3][SLOT]
    int* cs = reinterpret_cast<int*>(sxyz + (size_t) NROWS * 3 * SLOT);     // [NROWS][NT_CX + 2] cell starts
    int* ce = cs + NROWS * (NT_CX + 2);                                     // [NROWS][NT_CX + 2] cell ends
    int* rng = ce + NROWS * (NT_CX + 2);                                    // [NROWS][2] staged range (aligned begin, staged flag); [NROWS*2..] own range
    __shared__ __align__(8) unsigned long long bar;

    const int n_active = *P.n_active;
    // rows beyond the active ones (Ghosts) keep an empty list, as in neighbor_kernel
    {
        const int lo = max(P.row_lo, n_active), hi = min(P.row_hi, P.N);
        for (int s = lo + blockIdx.x * NT_THREADS + threadIdx.x; s < hi; s += gridDim.x * NT_THREADS) P.nbr_count[s] = 0;
    }
    const int tile = blockIdx.x, tx = tile % tiles_x, rowid = tile / tiles_x;
    const int iy = rowid % P.numY, izr = rowid / P.numY;   // izr: 0 .. (NZ == 1 ? 0 : numZ - 1)
    const int iz = NZ == 1 ? 1 : izr;                      // flat z: every active particle lives in the plane iz == 1
    const int c0 = tx * NT_CX, nxy = P.numX * P.numY;
    // cell ranges of the 3 / 9 neighbouring rows, cells c0 - 1 .. c0 + NT_CX
    for (int t = threadIdx.x; t < NROWS * (NT_CX + 2); t += NT_THREADS) {
        const int r = t / (NT_CX + 2), k = t - r * (NT_CX + 2);
        const int jy = iy - 1 + r / NZ, jz = NZ == 1 ? iz : iz - 1 + r % NZ, jx = c0 - 1 + k;
        int b = 0, e = 0;
        if (jx >= 0 && jx < P.numX && jy >= 0 && jy < P.numY && jz >= 0 && jz < P.numZ) {
            const int c = jx + jy * P.numX + jz * nxy;
            b = P.cell_start[c];
            e = P.cell_end[c];
        }
        cs[t] = b;
        ce[t] = e;
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(nt_smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // own rows: the non-empty cells k = 1 .. NT_CX of the centre row r = NZ (jy = iy, jz = iz)
        const int rc = 1 * NZ + (NZ == 1 ? 0 : 1);
        int ob = 0, oe = 0;
        for (int k = 1; k <= NT_CX; k++) {
            const int b = cs[rc * (NT_CX + 2) + k], e = ce[rc * (NT_CX + 2) + k];
            if (e > b) {
                if (oe == 0) ob = b;
                oe = e;
            }
        }
        ob = max(ob, P.row_lo);
        oe = min(oe, min(P.row_hi, n_active));
        rng[NROWS * 2]     = ob;
        rng[NROWS * 2 + 1] = oe;
        unsigned bytes = 0;
        if (oe > ob) {
            for (int r = 0; r < NROWS; r++) {
                int rb = 0, re = 0;
                for (int k = 0; k < NT_CX + 2; k++) {
                    const int b = cs[r * (NT_CX + 2) + k], e = ce[r * (NT_CX + 2) + k];
                    if (e > b) {
                        if (re == 0) rb = b;
                        re = e;
                    }
                }
                const int al = rb & ~1, cnt = ((re - al) + 1) & ~1; // 16-byte aligned begin, whole 16-byte units
                const int staged = (re > rb && cnt <= SLOT) ? 1 : 0;
                rng[2 * r]     = al;
                rng[2 * r + 1] = staged;
                if (staged) bytes += 3u * (unsigned) cnt * 8u;
            }
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(nt_smem_u32(&bar)), "r"(bytes) : "memory");
            for (int r = 0; r < NROWS; r++) {
                if (!rng[2 * r + 1]) continue;
                int rb = 0, re = 0;
                for (int k = 0; k < NT_CX + 2; k++) {
                    const int b = cs[r * (NT_CX + 2) + k], e = ce[r * (NT_CX + 2) + k];
                    if (e > b) {
                        if (re == 0) rb = b;
                        re = e;
                    }
                }
                const int al = rb & ~1, cnt = ((re - al) + 1) & ~1;
                const double* src[3] = {P.S.x + al, P.S.y + al, P.S.z + al};
                for (int a = 0; a < 3; a++)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                     nt_smem_u32(sxyz + ((size_t) r * 3 + a) * SLOT)),
                                 "l"(src[a]), "r"((unsigned) cnt * 8u), "r"(nt_smem_u32(&bar))
                                 : "memory");
            }
        }
    }
    __syncthreads();
    const int ob = rng[NROWS * 2], oe = rng[NROWS * 2 + 1];
    if (oe <= ob) return; // empty tile, or none of its rows is this rank's
    { // wait for the staged rows (phase 0 of the barrier)
        unsigned ok = 0, spins = 0;
        const unsigned addr = nt_smem_u32(&bar);
        while (!ok) {
            if (++spins > (1u << 26)) __trap(); // a lost TMA completion must fail loudly, never hang the GPU
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok)
                         : "r"(addr), "r"(0u)
                         : "memory");
        }
    }
    const double re = P.s.re_max;
    const double re2_out = (re * re) * (1.0 + 0x1p-50); // see neighbor_kernel
    int wmax = 0;
    for (int s = ob + (int) threadIdx.x; s < oe; s += NT_THREADS) {
        const double xi = P.S.x[s], yi = P.S.y[s], zi = P.S.z[s];
        const unsigned key = P.key[s];
        const int ix = (int) (key % (unsigned) P.numX); // the row (iy, iz) is the tile's
        const int kx = ix - c0;                          // 0 .. NT_CX - 1; neighbour cells k = kx .. kx + 2
        int count = 0;
#pragma unroll 1
        for (int q = 0; q < 3 * NROWS; q++) { // q = (dx, dy, dz): the reference's order
            const int dx = q / NROWS, r = q - dx * NROWS;
            const int b = cs[r * (NT_CX + 2) + kx + dx], e = ce[r * (NT_CX + 2) + kx + dx];
            if (e <= b) continue;
            const bool staged = rng[2 * r + 1] != 0;
            const int al = rng[2 * r];
            const double* X = staged ? sxyz + ((size_t) r * 3 + 0) * SLOT - al : P.S.x;
            const double* Y = staged ? sxyz + ((size_t) r * 3 + 1) * SLOT - al : P.S.y;
            const double* Z = staged ? sxyz + ((size_t) r * 3 + 2) * SLOT - al : P.S.z;
#pragma unroll 1
            for (int j0 = b; j0 < e; j0 += 4) {
                double cx[4], cy[4], cz[4];
#pragma unroll
                for (int t = 0; t < 4; t++) {
                    const int j = min(j0 + t, e - 1);
                    cx[t] = X[j];
                    cy[t] = Y[j];
                    cz[t] = Z[j];
                }
#pragma unroll
                for (int t = 0; t < 4; t++) {
                    const int j = j0 + t;
                    if (j >= e) break;
                    const double ddx = cx[t] - xi, ddy = cy[t] - yi, ddz = cz[t] - zi;
                    const double d2 = (ddx * ddx + ddy * ddy) + ddz * ddz;
                    if (d2 >= re2_out) continue;
                    const double dist = sqrt(d2);
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
        wmax = max(wmax, count);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if ((threadIdx.x & 31) == 0 && wmax > 0) atomicMax(&P.stats->max_neighbors, wmax);
}

template <int NROWS, int NT_CAP>
constexpr size_t nt_smem_bytes() {
    return (size_t) NROWS * 3 * (NT_CAP + 2) * sizeof(double) + (size_t) (2 * NROWS * (NT_CX + 2) + 2 * NROWS + 2) * sizeof(int);
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

int mpsb_neighbors_configure(mpsb_handle*) { // once per handle, outside any stream capture
    if (cudaFuncSetAttribute(neighbor_tile_kernel<3, 640>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) nt_smem_bytes<3, 640>()) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(neighbor_tile_kernel<9, 448>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) nt_smem_bytes<9, 448>()) != cudaSuccess) return -1;
    return 0;
}

// Default: the per-thread walk. The tiled kernel (MPSB_NEIGHBOR_TILES=1) is parity-green on every case but MEASURED SLOWER
// at 4.04 M particles on one B200 (profiles/r2_v6_neighbor_tiles_experiment.txt): 2.57 ms per build against 1.43 ms.
// The per-thread walk already takes 88 % of its candidate loads from L1, so staging saves little traffic, while a
// tile pays a serial range set-up and one exposed TMA round trip with nothing to overlap it (one tile per CTA), and
// 154 particles per 128-thread tile leave the second pass of the CTA mostly idle. Kept as an opt-in with its tests.
void mpsb_launch_neighbors(mpsb_handle* h) {
    Params& P = h->P;
    const bool tiles = h->neighbor_tiles > 0;
    if (!tiles) {
        MPSB_LAUNCH(h, neighbor_kernel, mpsb_own_grid(h, 128), 128, 0, P);
        return;
    }
    const bool flat = P.s.domain_max[2] == P.s.domain_min[2];
    const int tiles_x = mpsb_div_up(P.numX, NT_CX);
    if (flat) {
        constexpr size_t smem = nt_smem_bytes<3, 640>();
        MPSB_LAUNCH(h, (neighbor_tile_kernel<3, 640>), tiles_x * P.numY, NT_THREADS, smem, P, tiles_x);
    } else {
        constexpr size_t smem = nt_smem_bytes<9, 448>();
        MPSB_LAUNCH(h, (neighbor_tile_kernel<9, 448>), tiles_x * P.numY * P.numZ, NT_THREADS, smem, P, tiles_x);
    }
}

namespace {
// only_new: touch only particles that are Ghost now and were not before the upload (P.mark, ghost_mark_kernel)
__global__ void ghost_ratio_from_home_kernel(Params P, int only_new) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.N && P.H.type[i] == MPSB_GHOST && !(only_new && P.mark[i])) P.ghost_ratio[i] = P.H.ratio[i];
}
__global__ void ghost_mark_kernel(Params P) { // P.mark is per-phase scratch (collision replay, Dirichlet): free between calls
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.N) P.mark[i] = P.H.type[i] == MPSB_GHOST;
}
} // namespace

// An upload that carries the ratio sets every Ghost's recorded sum to its uploaded ratio. An upload of the types alone must
// not: a particle that turned Ghost in a step without a number-density pass still has its pre-ghost ratio in HOME, and the
// sum mpsb_record_ghost_ratio froze for it is the value the reference reports. Only a particle the upload itself turns
// Ghost takes its HOME ratio (mark before the copy with mpsb_launch_ghost_mark, refresh after it with only_new).
void mpsb_launch_ghost_ratio_from_home(mpsb_handle* h, bool only_new) {
    Params& P = h->P;
    MPSB_LAUNCH(h, ghost_ratio_from_home_kernel, mpsb_div_up(P.N, 256), 256, 0, P, only_new ? 1 : 0);
}

void mpsb_launch_ghost_mark(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, ghost_mark_kernel, mpsb_div_up(P.N, 256), 256, 0, P);
}

void mpsb_launch_number_density(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, number_density_kernel, mpsb_own_grid(h, 256), 256, 0, P);
}
