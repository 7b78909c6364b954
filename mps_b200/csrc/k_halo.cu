// k_halo.cu — the exchanges of the slab-decomposed step (world > 1).
//
// Every rank owns a fixed block of the SORTED rows (a y-slab of the cell grid: SORTED order is cell-major with x
// fastest) and runs every per-particle kernel on its own rows only. What the reference does in one address space
// (src/mps.cpp:585-619 setNeighbors, :262-360 flags / matrix, :472-533 minimum pressure / gradient) needs three
// kinds of cross-rank traffic here, all of it PUSHED by the owner into the peer's HBM over NVLink peer memory and
// published by one cross-rank barrier — nothing polls or reads over NVLink:
//   * halo particles + migration: folded into the cell build (k_cells.cu): after the (replicated) sort of the
//     (key, id) pairs an owner stores each of its particles into the HOME arrays of every rank whose new extent
//     holds the particle's new row — its own slab (migration) or its halo;
//   * halo fields: boundary-condition flags before the assembly, pressures before minimum pressure / gradient
//     (halo_push kernels below);
//   * collision records and Dirichlet reach flags (k_explicit.cu, k_pressure.cu).
// The Krylov loop's own exchanges live in k_pcg.cu.
#include "mpsb_internal.cuh"

namespace {

__device__ __forceinline__ unsigned long long halo_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Cross-rank barrier on the stream: everything this rank's earlier kernels stored (locally or into peers) is
// visible to every rank's later kernels. One CTA: bump the shared epoch counter (the same monotonic sequence the
// PCG kernel's exchanges use, so flags never go backwards), fence at system scope, store the epoch into every
// peer's inbox flag, wait until every peer's epoch has arrived in mine. A peer that never arrives trips a 20 s
// timeout -> trap -> MPSB_ERR_CUDA on the host, never a hang.
__global__ void xbarrier_kernel(Params P) {
    __shared__ unsigned long long epoch;
    const PeerLinks& L = P.L;
    if (threadIdx.x == 0) {
        epoch    = *L.epoch + 1ull;
        *L.epoch = epoch;
        __threadfence_system();
    }
    __syncthreads();
    if ((int) threadIdx.x < L.world) {
        const int q = threadIdx.x;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(L.peer_flag[q] + L.rank), "l"(epoch) : "memory");
        const unsigned long long t0 = halo_timer_ns();
        unsigned long long seen;
        unsigned spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(L.flag + q) : "memory");
            if ((++spins & 1023u) == 0 && halo_timer_ns() - t0 > 20000000000ull) {
                atomicOr(&P.stats->error_flags, 8);
                __threadfence_system();
                __trap();
            }
        } while (seen < epoch);
    }
}

// Row s of this rank lies inside peer q's extent (its halo) -> store the field value into q's copy of the array.
template <class T>
__global__ void halo_push_kernel(Params P, T* field) {
    const int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.row_hi || s >= P.N) return;
    const RankExtents& E = *P.ext;
    T v;
    bool loaded = false;
    for (int q = 0; q < P.L.world; q++) {
        if (q == P.L.rank || s < E.ext_lo[q] || s >= E.ext_hi[q]) continue;
        if (!loaded) {
            v      = field[s];
            loaded = true;
        }
        mpsb_peer(P, field, q)[s] = v;
    }
}

// Counters: sums, maxima and flag ORs over the ranks; the solve's figures are identical on every rank already.
__global__ void combine_stats_kernel(Params P) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    DevStats out = P.stats_all[P.L.rank];
    out.n_inner = out.n_surface = out.nnz = out.n_ghosted = out.n_unconnected = out.n_collisions = 0;
    out.max_neighbors = 0;
    out.error_flags   = 0;
    unsigned long long courant = 0ull;
    for (int q = 0; q < P.L.world; q++) {
        const DevStats& d = P.stats_all[q];
        out.n_inner += d.n_inner;
        out.n_surface += d.n_surface;
        out.nnz += d.nnz;
        out.n_ghosted += d.n_ghosted;
        out.n_unconnected += d.n_unconnected;
        out.max_neighbors = max(out.max_neighbors, d.max_neighbors);
        out.error_flags |= d.error_flags;
        const unsigned long long c = (unsigned long long) __double_as_longlong(d.courant_bits_as_double);
        courant = c > courant ? c : courant; // non-negative doubles order like their bit patterns
    }
    out.n_collisions           = P.stats_all[P.L.rank].n_collisions; // the replay is replicated: every rank counted them all
    out.courant_bits_as_double = __longlong_as_double((long long) courant);
    P.stats_all[MPSB_MAX_RANKS] = out; // fetch_stats reads this; the rank's own counters stay as they are
}

} // namespace

void mpsb_launch_xbarrier(mpsb_handle* h) {
    MPSB_LAUNCH(h, xbarrier_kernel, 1, 32, 0, h->P);
}

void mpsb_launch_halo_push_bc(mpsb_handle* h) {
    MPSB_LAUNCH(h, halo_push_kernel<int>, mpsb_own_grid(h, 256), 256, 0, h->P, h->P.S.bc);
}

void mpsb_launch_halo_push_p(mpsb_handle* h) {
    MPSB_LAUNCH(h, halo_push_kernel<double>, mpsb_own_grid(h, 256), 256, 0, h->P, h->P.S.p);
}

void mpsb_launch_combine_stats(mpsb_handle* h) {
    MPSB_LAUNCH(h, combine_stats_kernel, 1, 32, 0, h->P);
}
