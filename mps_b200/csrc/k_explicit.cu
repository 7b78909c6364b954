// k_explicit.cu — explicit half of the step and the final correction:
//   explicit_kernel   : calcGravity + calcViscosity + moveParticles        (src/mps.cpp:165-215)
//   collision kernels : collision                                          (src/mps.cpp:217-250)
//   gradient_kernel   : calcPressureGradient                               (src/mps.cpp:505-533)
//   move2_kernel      : moveParticlesWithPressureGradient                  (src/mps.cpp:535-545)
//   courant_kernel    : getCourantNumber                                   (src/mps.cpp:547-571)
// All sums follow neighbour-list order and every fp64 operation is individually rounded
// (-fmad=false) so results are bit-identical to the reference.
#include "mpsb_internal.cuh"

namespace {

// checkBoundaryViolation (src/mps.cpp:643-667): strict '<' on both sides, per axis.
__device__ __forceinline__ bool out_of_domain(const mpsb_settings& s, double x, double y, double z) {
    bool in = true;
    if (x < s.domain_min[0] || s.domain_max[0] < x) in = false;
    if (y < s.domain_min[1] || s.domain_max[1] < y) in = false;
    if (z < s.domain_min[2] || s.domain_max[2] < z) in = false;
    return !in;
}

__device__ __forceinline__ void ghost_event(const Params& P, int s, double x, double y, double z) {
    mpsb_record_ghost_ratio(P, s);
    int e = atomicAdd(&P.stats->n_ghosted, 1);
    if (e < P.ev_cap) {
        P.ev_ghost_id[e]          = P.S.id[s];
        P.ev_ghost_xyz[3 * e]     = x;
        P.ev_ghost_xyz[3 * e + 1] = y;
        P.ev_ghost_xyz[3 * e + 2] = z;
    } else {
        atomicOr(&P.stats->error_flags, 4);
    }
}

// Reads neighbours' OLD velocity (u,v,w), writes the new one to (u2,v2,w2); the host swaps.
__global__ void __launch_bounds__(256) explicit_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= P.N) return;
    double u = P.S.u[s], v = P.S.v[s], w = P.S.w[s];
    if (s < *P.n_active && P.S.type[s] == MPSB_FLUID) {
        const double re = P.s.re_viscosity;
        int cnt = P.nbr_count[s];
        double tx = 0.0, ty = 0.0, tz = 0.0;
        struct Vel { double u, v, w; };
        mpsb_walk_list<8, Vel>(
            P, s, cnt, [&](int j) { return Vel{P.S.u[j], P.S.v[j], P.S.w[j]}; },
            [&](int, double d, int, const Vel& q) {
                if (d < re) {
                    double wgt = (re / d) - 1.0;
                    tx += (q.u - u) * wgt;
                    ty += (q.v - v) * wgt;
                    tz += (q.w - w) * wgt;
                }
            });
        tx *= P.visc_coef; ty *= P.visc_coef; tz *= P.visc_coef;
        double ax = P.s.gravity[0] + tx, ay = P.s.gravity[1] + ty, az = P.s.gravity[2] + tz;
        const double dt = P.s.dt;
        u += ax * dt; v += ay * dt; w += az * dt;
        double x = P.S.x[s] + u * dt, y = P.S.y[s] + v * dt, z = P.S.z[s] + w * dt;
        P.S.x[s] = x; P.S.y[s] = y; P.S.z[s] = z;
        if (out_of_domain(P.s, x, y, z)) {
            ghost_event(P, s, x, y, z);
            P.S.type[s] = MPSB_GHOST;
        }
    }
    P.S.u2[s] = u; P.S.v2[s] = v; P.S.w2[s] = w;
    P.S.ax[s] = 0.0; P.S.ay[s] = 0.0; P.S.az[s] = 0.0; // acceleration.setZero() for every type (:213)
}

// ---- collision -----------------------------------------------------------------------------------
// The reference loop is sequential and in place (Gauss-Seidel): later pairs see earlier updates.
// Pass 1 (parallel) collects the candidate pairs — list distance < collisionDistance, Fluid i,
// skipping Fluid j with id >= i — keyed by (id_i, list slot), which is exactly the reference's
// visiting order. Pass 2 sorts the (few) keys and replays them serially with live positions and
// velocities. Candidate membership depends only on the stale list distances and on types, neither
// of which changes during the loop, so the candidate set equals the reference's.
__global__ void __launch_bounds__(256) collision_candidates_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.type[s] != MPSB_FLUID) return;
    int cnt = P.nbr_count[s], id = P.S.id[s];
    // candidates are rare: only the distances are batched, the neighbour's type and id are read on a hit
    mpsb_walk_list<8, MpsbNoRec>(
        P, s, cnt, [](int) { return MpsbNoRec{}; },
        [&](int k, double d, int j, const MpsbNoRec&) {
            if (d < P.s.collision_distance) {
                if (P.S.type[j] == MPSB_FLUID && P.S.id[j] >= id) return;
                int c = atomicAdd(P.cand_count, 1);
                if (c < P.cand_cap)
                    P.cand[c] = ((unsigned long long) (unsigned) id << 32) | (unsigned) k;
                else
                    atomicOr(&P.stats->error_flags, 1);
            }
        });
}

__global__ void __launch_bounds__(1024) collision_replay_kernel(Params P) {
    int M = min(*P.cand_count, P.cand_cap);
    __syncthreads(); // every thread has read the count: leave it cleared for the next candidate pass
    if (threadIdx.x == 0) {
        *P.cand_count         = 0;
        P.stats->n_collisions = M;
    }
    if (M == 0) return;
    // bitonic sort of the keys in global memory (padded with ~0 to a power of two)
    int M2 = 1;
    while (M2 < M) M2 <<= 1;
    for (int i = M + threadIdx.x; i < M2; i += blockDim.x) P.cand[i] = ~0ull;
    __syncthreads();
    for (int size = 2; size <= M2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (M2 >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                bool up = (lo & size) == 0;
                unsigned long long a = P.cand[lo], b = P.cand[hi];
                if ((a > b) == up) {
                    P.cand[lo] = b;
                    P.cand[hi] = a;
                }
            }
            __syncthreads();
        }
    if (threadIdx.x != 0) return;
    const double col = P.s.collision_distance, e = P.s.restitution;
    for (int c = 0; c < M; c++) {
        unsigned long long key = P.cand[c];
        int s = P.H.inv[(int) (key >> 32)], k = (int) (key & 0xffffffffu);
        int j = P.nbr_idx[SELL(k, s)];
        double dist = P.nbr_dist[SELL(k, s)];
        double invMi = P.inv_density;                                       // i is Fluid
        double invMj = P.S.type[j] == MPSB_FLUID ? P.inv_density : 0.0;    // src/particle.cpp:19-29
        double reducedMass = 1.0 / (invMi + invMj);
        double dx = P.S.x[j] - P.S.x[s], dy = P.S.y[j] - P.S.y[s], dz = P.S.z[j] - P.S.z[s];
        double z2 = (dx * dx + dy * dy) + dz * dz;
        double nx = dx, ny = dy, nz = dz; // Eigen normalized(): divide by sqrt when > 0
        if (z2 > 0.0) {
            double sz = sqrt(z2);
            nx = dx / sz; ny = dy / sz; nz = dz / sz;
        }
        double du = P.S.u[j] - P.S.u[s], dv = P.S.v[j] - P.S.v[s], dw = P.S.w[j] - P.S.w[s];
        double rel = (du * nx + dv * ny) + dw * nz;
        if (rel < 0.0) {
            double impulse = -(1.0 + e) * rel * reducedMass;
            double ci = impulse * invMi, cj = impulse * invMj;
            P.S.u[s] -= ci * nx; P.S.v[s] -= ci * ny; P.S.w[s] -= ci * nz;
            P.S.u[j] += cj * nx; P.S.v[j] += cj * ny; P.S.w[j] += cj * nz;
            double depth = col - dist;
            double posImpulse = depth * reducedMass;
            double pi = posImpulse * invMi, pj = posImpulse * invMj;
            P.S.x[s] -= pi * nx; P.S.y[s] -= pi * ny; P.S.z[s] -= pi * nz;
            P.S.x[j] += pj * nx; P.S.y[j] += pj * ny; P.S.z[j] += pj * nz;
        }
    }
}


// ---- collision across ranks (slab decomposition) -------------------------------------------------------------
// The reference loop is ONE serial sweep over all particles (src/mps.cpp:217-250): a pair in one slab can feed a
// chain that ends in another. So the candidates are found per rank (own rows), but every rank replays ALL of them,
// identically: each owner stores its candidates as self-contained records — visiting-order key plus the
// pre-collision position / velocity of both particles (halo copies are current: the lists were just built) — into
// its segment of EVERY rank's record inbox over NVLink, a cross-rank barrier publishes them, and one CTA per rank
// sorts the keys and replays the pairs serially on a compact copy of the particles the records touch (P.mark maps
// row -> compact slot), exactly the arithmetic of collision_replay_kernel. Own rows then take their new state from
// the compact copy; halo copies are refreshed by the list build that follows.
__global__ void __launch_bounds__(256) collision_records_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.row_hi || s >= *P.n_active || P.S.type[s] != MPSB_FLUID) return;
    int cnt = P.nbr_count[s], id = P.S.id[s];
    mpsb_walk_list<8, MpsbNoRec>(
        P, s, cnt, [](int) { return MpsbNoRec{}; },
        [&](int k, double d, int j, const MpsbNoRec&) {
            if (d < P.s.collision_distance) {
                const int tj = P.S.type[j];
                if (tj == MPSB_FLUID && P.S.id[j] >= id) return;
                int c = atomicAdd(P.cand_count, 1);
                if (c >= P.crec_seg) {
                    atomicOr(&P.stats->error_flags, 1);
                    return;
                }
                CollisionRecord r;
                r.key    = ((unsigned long long) (unsigned) id << 32) | ((unsigned long long) (unsigned) k << 20);
                r.row_i  = s;
                r.row_j  = j;
                r.type_j = tj;
                r.pad    = 0;
                r.dist   = d;
                r.xi[0] = P.S.x[s]; r.xi[1] = P.S.y[s]; r.xi[2] = P.S.z[s];
                r.ui[0] = P.S.u[s]; r.ui[1] = P.S.v[s]; r.ui[2] = P.S.w[s];
                r.xj[0] = P.S.x[j]; r.xj[1] = P.S.y[j]; r.xj[2] = P.S.z[j];
                r.uj[0] = P.S.u[j]; r.uj[1] = P.S.v[j]; r.uj[2] = P.S.w[j];
                const size_t at = (size_t) P.L.rank * P.crec_seg + c;
                for (int q = 0; q < P.L.world; q++) mpsb_peer(P, P.crec, q)[at] = r;
            }
        });
}

__global__ void collision_count_push_kernel(Params P) {
    if (threadIdx.x != 0) return;
    const int n = min(*P.cand_count, P.crec_seg);
    *P.cand_count = 0;
    for (int q = 0; q < P.L.world; q++) mpsb_peer(P, P.crec_count, q)[P.L.rank] = n;
}

__global__ void __launch_bounds__(1024) collision_replay_records_kernel(Params P) {
    __shared__ int offset[MPSB_MAX_RANKS + 1];
    const int W = P.L.world;
    if (threadIdx.x == 0) {
        offset[0] = 0;
        for (int q = 0; q < W; q++) offset[q + 1] = offset[q] + P.crec_count[q];
        P.stats->n_collisions = offset[W];
    }
    __syncthreads();
    const int M = offset[W];
    if (M == 0) return;
    // keys with the record index in the low 20 bits; forget stale row -> slot entries
    for (int q = 0; q < W; q++)
        for (int c = threadIdx.x; c < P.crec_count[q]; c += blockDim.x) {
            const int at = q * P.crec_seg + c;
            const CollisionRecord& r = P.crec[at];
            P.cand[offset[q] + c] = r.key | (unsigned long long) at;
            P.mark[r.row_i] = 0;
            P.mark[r.row_j] = 0;
        }
    int M2 = 1;
    while (M2 < M) M2 <<= 1;
    for (int i = M + threadIdx.x; i < M2; i += blockDim.x) P.cand[i] = ~0ull;
    __syncthreads();
    for (int size = 2; size <= M2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (M2 >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                bool up = (lo & size) == 0;
                unsigned long long a = P.cand[lo], b = P.cand[hi];
                if ((a > b) == up) {
                    P.cand[lo] = b;
                    P.cand[hi] = a;
                }
            }
            __syncthreads();
        }
    const int cap = 2 * W * P.crec_seg;
    double *cx = P.cstate, *cy = cx + cap, *cz = cy + cap, *cu = cz + cap, *cv = cu + cap, *cw = cv + cap;
    if (threadIdx.x == 0) {
        const double col = P.s.collision_distance, e = P.s.restitution;
        int used = 0;
        auto slot_of = [&](int row, const double* x, const double* u) -> int {
            int m = P.mark[row];
            if (m == 0) {
                m = ++used;
                P.mark[row] = m;
                cx[m - 1] = x[0]; cy[m - 1] = x[1]; cz[m - 1] = x[2];
                cu[m - 1] = u[0]; cv[m - 1] = u[1]; cw[m - 1] = u[2];
            }
            return m - 1;
        };
        for (int c = 0; c < M; c++) {
            const CollisionRecord& r = P.crec[(int) (P.cand[c] & 0xFFFFFull)];
            const int a = slot_of(r.row_i, r.xi, r.ui), b = slot_of(r.row_j, r.xj, r.uj);
            const double dist = r.dist;
            double invMi = P.inv_density;
            double invMj = r.type_j == MPSB_FLUID ? P.inv_density : 0.0;
            double reducedMass = 1.0 / (invMi + invMj);
            double dx = cx[b] - cx[a], dy = cy[b] - cy[a], dz = cz[b] - cz[a];
            double z2 = (dx * dx + dy * dy) + dz * dz;
            double nx = dx, ny = dy, nz = dz;
            if (z2 > 0.0) {
                double sz = sqrt(z2);
                nx = dx / sz; ny = dy / sz; nz = dz / sz;
            }
            double du = cu[b] - cu[a], dv = cv[b] - cv[a], dw = cw[b] - cw[a];
            double rel = (du * nx + dv * ny) + dw * nz;
            if (rel < 0.0) {
                double impulse = -(1.0 + e) * rel * reducedMass;
                double ci = impulse * invMi, cj = impulse * invMj;
                cu[a] -= ci * nx; cv[a] -= ci * ny; cw[a] -= ci * nz;
                cu[b] += cj * nx; cv[b] += cj * ny; cw[b] += cj * nz;
                double depth = col - dist;
                double posImpulse = depth * reducedMass;
                double pi = posImpulse * invMi, pj = posImpulse * invMj;
                cx[a] -= pi * nx; cy[a] -= pi * ny; cz[a] -= pi * nz;
                cx[b] += pj * nx; cy[b] += pj * ny; cz[b] += pj * nz;
            }
        }
    }
    __syncthreads();
    // own rows take their state from the compact copy (several records may name the same row: same value)
    for (int q = 0; q < W; q++)
        for (int c = threadIdx.x; c < P.crec_count[q]; c += blockDim.x) {
            const CollisionRecord& r = P.crec[q * P.crec_seg + c];
            const int rows[2] = {r.row_i, r.row_j};
            for (int t = 0; t < 2; t++) {
                const int row = rows[t];
                if (row < P.row_lo || row >= P.row_hi) continue;
                const int m = P.mark[row] - 1;
                P.S.x[row] = cx[m]; P.S.y[row] = cy[m]; P.S.z[row] = cz[m];
                P.S.u[row] = cu[m]; P.S.v[row] = cv[m]; P.S.w[row] = cw[m];
            }
        }
}

// ---- pressure gradient and correction ----------------------------------------------------------------
__global__ void __launch_bounds__(256) gradient_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.type[s] != MPSB_FLUID) return;
    const double re = P.s.re_pressure;
    const double xi = P.S.x[s], yi = P.S.y[s], zi = P.S.z[s], pmin = P.S.minp[s];
    int cnt = P.nbr_count[s];
    double gx = 0.0, gy = 0.0, gz = 0.0;
    struct Nb { double x, y, z, p; int type; };
    mpsb_walk_list<8, Nb>(
        P, s, cnt, [&](int j) { return Nb{P.S.x[j], P.S.y[j], P.S.z[j], P.S.p[j], P.S.type[j]}; },
        [&](int, double d, int, const Nb& q) {
            if (d < re) {
                if (q.type == MPSB_DUMMY_WALL) return;
                double dp = q.p - pmin, wgt = (re / d) - 1.0, d2 = d * d;
                gx += (q.x - xi) * dp * wgt / d2;
                gy += (q.y - yi) * dp * wgt / d2;
                gz += (q.z - zi) * dp * wgt / d2;
            }
        });
    gx *= P.grad_coef; gy *= P.grad_coef; gz *= P.grad_coef;
    P.S.ax[s] = -1.0 * gx / P.s.density;
    P.S.ay[s] = -1.0 * gy / P.s.density;
    P.S.az[s] = -1.0 * gz / P.s.density;
}

__global__ void __launch_bounds__(256) move2_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    if (s >= *P.n_active || P.S.type[s] != MPSB_FLUID) return;
    const double dt = P.s.dt;
    double ax = P.S.ax[s], ay = P.S.ay[s], az = P.S.az[s];
    P.S.u[s] += ax * dt; P.S.v[s] += ay * dt; P.S.w[s] += az * dt;
    double x = P.S.x[s] + ax * dt * dt, y = P.S.y[s] + ay * dt * dt, z = P.S.z[s] + az * dt * dt;
    P.S.x[s] = x; P.S.y[s] = y; P.S.z[s] = z;
    if (out_of_domain(P.s, x, y, z)) {
        ghost_event(P, s, x, y, z);
        P.S.type[s] = MPSB_GHOST;
    }
    // acceleration is deliberately left as is: `pi.acceleration.Zero()` is a no-op (SURVEY B-3)
}

__global__ void __launch_bounds__(256) courant_kernel(Params P) {
    int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x; // own rows (one GPU: all of them)
    if (s >= P.row_hi) return;
    double c = 0.0;
    if (s < *P.n_active && P.S.type[s] == MPSB_FLUID) {
        double u = P.S.u[s], v = P.S.v[s], w = P.S.w[s];
        double cc = (sqrt((u * u + v * v) + w * w) * P.s.dt) / P.s.particle_distance;
        if (cc > c) c = cc; // std::max(max, c): a NaN never replaces the running maximum
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c = fmax(c, __shfl_xor_sync(0xffffffffu, c, o));
    // non-negative doubles order like their bit patterns
    if ((threadIdx.x & 31) == 0 && c > 0.0)
        atomicMax((unsigned long long*) &P.stats->courant_bits_as_double, (unsigned long long) __double_as_longlong(c));
}

} // namespace

void mpsb_launch_explicit(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, explicit_kernel, mpsb_own_grid(h, 256), 256, 0, P);
    double* t;
    t = P.S.u; P.S.u = P.S.u2; P.S.u2 = t;
    t = P.S.v; P.S.v = P.S.v2; P.S.v2 = t;
    t = P.S.w; P.S.w = P.S.w2; P.S.w2 = t;
}

void mpsb_launch_collision(mpsb_handle* h) {
    Params& P = h->P;
    if (h->world > 1) {
        MPSB_LAUNCH(h, collision_records_kernel, mpsb_own_grid(h, 256), 256, 0, P);
        MPSB_LAUNCH(h, collision_count_push_kernel, 1, 32, 0, P);
        mpsb_launch_xbarrier(h);
        MPSB_LAUNCH(h, collision_replay_records_kernel, 1, 1024, 0, P);
        return;
    }
    MPSB_LAUNCH(h, collision_candidates_kernel, mpsb_own_grid(h, 256), 256, 0, P);
    MPSB_LAUNCH(h, collision_replay_kernel, 1, 1024, 0, P);
}

void mpsb_launch_gradient_move(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, gradient_kernel, mpsb_own_grid(h, 256), 256, 0, P);
    MPSB_LAUNCH(h, move2_kernel, mpsb_own_grid(h, 256), 256, 0, P);
}

void mpsb_launch_courant(mpsb_handle* h) {
    Params& P = h->P;
    if (!h->in_step) cudaMemsetAsync(&P.stats->courant_bits_as_double, 0, sizeof(double), h->stream); // a step clears all stats at once
    MPSB_LAUNCH(h, courant_kernel, mpsb_own_grid(h, 256), 256, 0, P);
}
