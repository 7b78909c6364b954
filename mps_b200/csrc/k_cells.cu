// k_cells.cu — cell-list (bucket) construction: particle -> cell key, stable LSD radix sort of
// (key, id), cell start/end ranges, and the HOME <-> SORTED permutation of the particle state.
//
// Replaces reference Bucket::storeParticles (src/bucket.cpp:24-51): the reference threads every
// cell's particles on a linked list in id order; here a STABLE sort by cell key starting from id
// order gives every cell as one contiguous range of the sorted arrays with ids ascending inside
// it, i.e. the same traversal order. Ghosts get the maximum key and sink to the tail
// (the reference skips them, bucket.cpp:37-38).
#include "mpsb_internal.cuh"
#include <cstdio>

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS   = 8;
constexpr int SORT_TILE    = SORT_THREADS * SORT_ITEMS;
constexpr int RADIX        = 256;

// Cell key from the HOME position, by particle id. ix = (int)((x - xmin)/len) + 1 with a true
// division and C truncation (src/bucket.cpp:40-43). A non-ghost particle whose cell would make the
// reference's 27-cell scan index outside its arrays raises the out-of-grid flag (SURVEY B-8).
// It also clears the cell ranges and the active count of the previous build (they are rewritten by
// cell_range_kernel, several kernels later), which saves three memset nodes per build.
__device__ __forceinline__ unsigned cell_key_of(const Params& P, int i) {
    unsigned key = MPSB_GHOST_KEY;
    if (P.H.type[i] != MPSB_GHOST) {
        double len = P.s.re_max;
        int ix = (int) ((P.H.pos[3 * i + 0] - P.s.domain_min[0]) / len) + 1;
        int iy = (int) ((P.H.pos[3 * i + 1] - P.s.domain_min[1]) / len) + 1;
        int iz = (int) ((P.H.pos[3 * i + 2] - P.s.domain_min[2]) / len) + 1;
        bool bad = ix < 1 || ix > P.numX - 2 || iy < 1 || iy > P.numY - 2 || iz < 1 || iz > P.numZ - 2;
        if (bad) {
            atomicOr(&P.stats->error_flags, 2);
            ix = min(max(ix, 0), P.numX - 1);
            iy = min(max(iy, 0), P.numY - 1);
            iz = min(max(iz, 0), P.numZ - 1);
        }
        key = (unsigned) (ix + iy * P.numX + iz * P.numX * P.numY);
    }
    return key;
}

__device__ __forceinline__ void scatter_row(const Params& P, int s);

// scatter_first: thread i first writes SORTED row i back to HOME (the sorted copy is newer) and then keys the
// particle that row belongs to — one launch instead of scatter_all_kernel + key_kernel.
__global__ void key_kernel(Params P, int scatter_first) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int c = i; c < P.numCells; c += gridDim.x * blockDim.x) {
        P.cell_start[c] = 0;
        P.cell_end[c]   = 0;
    }
    if (i == 0) *P.n_active = 0;
    if (i >= P.N) return;
    int id = i;
    if (scatter_first) {
        scatter_row(P, i);
        id = P.S.id[i];
    }
    P.key[id] = cell_key_of(P, id);
    P.val[id] = id;
}

// Ghost keys map to digit 255 in every pass so they stay at the tail for any number of passes.
__device__ __forceinline__ unsigned digit_of(unsigned key, int shift) { return (key >> shift) & (RADIX - 1); }

// Pass 1 of a radix pass: per-tile digit histogram, stored digit-major: hist[d * numTiles + tile].
__global__ void radix_hist_kernel(const unsigned* __restrict__ key, int n, int shift, int numTiles,
                                  int* __restrict__ hist) {
    __shared__ int sh[RADIX];
    sh[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; it++) {
        int i = base + it * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&sh[digit_of(key[i], shift)], 1);
    }
    __syncthreads();
    hist[threadIdx.x * numTiles + blockIdx.x] = sh[threadIdx.x];
}

// Exclusive scan of the digit-major tile histograms (256 * numTiles ints, 505 k at 4 M particles), in place, as ONE
// multi-CTA launch with decoupled look-back (Merrill & Garland): a tile of 4 096 elements per CTA; every CTA
// publishes its tile aggregate at once, then one warp walks back over its predecessors' status words (32 at a
// time) adding aggregates until it meets an inclusive prefix, and publishes its own inclusive prefix. Round 1
// scanned this array with ONE CTA: 124 us per call, 9 calls per step — more than the histogram and scatter passes
// together.
// No memset and no host-side epoch, so the launch is CUDA-graph safe: tiles are handed out by a 64-bit ticket
// counter that runs on across launches (every launch of a handle has the same tile count T, launches are
// stream-ordered), ticket t -> launch epoch t / T, tile t % T; a status word packs {epoch:30, flag:2, value:32},
// and a word from another epoch reads as "not ready". Taking tiles by ticket also gives the forward-progress
// guarantee look-back needs: a tile's predecessors have all been started by resident CTAs.
constexpr int SCAN_THREADS = 512, SCAN_ITEMS = 8, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
constexpr unsigned long long SCAN_AGG = 1ull, SCAN_PREFIX = 2ull;

__device__ __forceinline__ unsigned long long scan_pack(unsigned long long epoch, unsigned long long flag, int v) {
    return ((epoch & 0x3FFFFFFFull) << 34) | (flag << 32) | (unsigned long long) (unsigned) v;
}

__device__ __forceinline__ unsigned long long scan_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_lookback_kernel(int* __restrict__ data, int n, int ntiles,
                                                                    unsigned long long* __restrict__ status,
                                                                    unsigned long long* __restrict__ ticket) {
    __shared__ int warp_sums[SCAN_THREADS / 32];
    __shared__ unsigned long long ticket_sh;
    __shared__ int excl_sh;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) ticket_sh = atomicAdd(ticket, 1ull);
    __syncthreads();
    const unsigned long long epoch = ticket_sh / (unsigned long long) ntiles;
    const int tile = (int) (ticket_sh % (unsigned long long) ntiles);
    const int idx  = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS];
    if (idx + SCAN_ITEMS <= n) {
        const int4 a = reinterpret_cast<const int4*>(data + idx)[0], b = reinterpret_cast<const int4*>(data + idx)[1];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) v[k] = idx + k < n ? data[idx + k] : 0;
    }
    int sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) sum += v[k];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const int ws = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
        int wi = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = wi - ws; // exclusive warp offsets
        const int aggregate = __shfl_sync(0xffffffffu, wi, 31);
        volatile unsigned long long* st = status;
        if (lane == 0) st[tile] = scan_pack(epoch, tile == 0 ? SCAN_PREFIX : SCAN_AGG, aggregate);
        int exclusive = 0;
        for (int look = tile - 1; look >= 0; look -= 32) { // warp-uniform
            const int j = look - lane;
            unsigned long long w = scan_pack(epoch, SCAN_PREFIX, 0); // lanes before tile 0: an empty inclusive prefix
            if (j >= 0) {
                // A predecessor publishes within microseconds (tiles are handed out in ticket order, so it is already
                // running). A status word that never turns up means a launch with another tile count or a ticket out
                // of step with the epochs: fail loudly (trap -> MPSB_ERR_CUDA on the host) instead of hanging the GPU.
                unsigned spins = 0;
                unsigned long long t0 = 0ull;
                do {
                    w = st[j];
                    if ((++spins & 4095u) == 0u) {
                        const unsigned long long now = scan_timer_ns();
                        if (t0 == 0ull) t0 = now;
                        else if (now - t0 > 5000000000ull) __trap();
                    }
                } while ((w >> 34) != (epoch & 0x3FFFFFFFull) || ((w >> 32) & 3ull) == 0ull);
            }
            const unsigned has_prefix = __ballot_sync(0xffffffffu, ((w >> 32) & 3ull) == SCAN_PREFIX);
            const int stop = has_prefix ? __ffs(has_prefix) - 1 : 31; // nearest predecessor holding an inclusive prefix
            int part = lane <= stop ? (int) (unsigned) w : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            exclusive += part;
            if (has_prefix) break;
        }
        if (lane == 0) {
            if (tile > 0) st[tile] = scan_pack(epoch, SCAN_PREFIX, exclusive + aggregate);
            excl_sh = exclusive;
        }
    }
    __syncthreads();
    int excl = excl_sh + warp_sums[warp] + (incl - sum);
    if (idx + SCAN_ITEMS <= n) {
        int4 a, b;
        a.x = excl; excl += v[0]; a.y = excl; excl += v[1]; a.z = excl; excl += v[2]; a.w = excl; excl += v[3];
        b.x = excl; excl += v[4]; b.y = excl; excl += v[5]; b.z = excl; excl += v[6]; b.w = excl;
        reinterpret_cast<int4*>(data + idx)[0] = a;
        reinterpret_cast<int4*>(data + idx)[1] = b;
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            if (idx + k < n) data[idx + k] = excl;
            excl += v[k];
        }
    }
}

// Pass 2 of a radix pass: stable scatter. The tile is consumed 256 elements at a time in index
// order; inside each group a warp ranks equal digits with __match_any_sync (lane order = index
// order), warps are prefix-summed per digit, and a per-digit running offset carries between groups.
__global__ void radix_scatter_kernel(const unsigned* __restrict__ key_in, const int* __restrict__ val_in,
                                     unsigned* __restrict__ key_out, int* __restrict__ val_out, int n, int shift,
                                     int numTiles, const int* __restrict__ hist) {
    constexpr int WARPS = SORT_THREADS / 32;
    __shared__ int warp_cnt[WARPS][RADIX];
    __shared__ int running[RADIX];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    running[threadIdx.x] = hist[threadIdx.x * numTiles + blockIdx.x];
    int base = blockIdx.x * SORT_TILE;
    for (int it = 0; it < SORT_ITEMS; it++) {
#pragma unroll
        for (int w = 0; w < WARPS; w++) warp_cnt[w][threadIdx.x] = 0;
        __syncthreads();
        int i      = base + it * SORT_THREADS + threadIdx.x;
        bool valid = i < n;
        unsigned k = valid ? key_in[i] : 0u;
        int v      = valid ? val_in[i] : 0;
        unsigned d = valid ? digit_of(k, shift) : RADIX; // invalid lanes get a digit nobody shares
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank_in_warp == 0) warp_cnt[warp][d] = __popc(peers);
        __syncthreads();
        {
            int dgt = threadIdx.x, acc = running[dgt];
#pragma unroll
            for (int w = 0; w < WARPS; w++) {
                int c = warp_cnt[w][dgt];
                warp_cnt[w][dgt] = acc;
                acc += c;
            }
            running[dgt] = acc;
        }
        __syncthreads();
        if (valid) {
            int dst      = warp_cnt[warp][d] + rank_in_warp;
            key_out[dst] = k;
            val_out[dst] = v;
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void gather_row(const Params& P, int s, int from_val);

// cell_start / cell_end from the sorted keys, the number of non-ghost particles, and (same thread, same row)
// perm, inverse perm and HOME -> SORTED for every field.
__global__ void cell_range_gather_kernel(Params P) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.N) return;
    unsigned k = P.key[s];
    if (k != MPSB_GHOST_KEY) {
        if (s == 0 || P.key[s - 1] != k) P.cell_start[k] = s;
        unsigned next = (s == P.N - 1) ? MPSB_GHOST_KEY : P.key[s + 1];
        if (next != k) P.cell_end[k] = s + 1;
        if (next == MPSB_GHOST_KEY) *P.n_active = s + 1;
    }
    gather_row(P, s, 1);
}

// HOME (id order) -> SORTED (cell order) for every field; also perm and its inverse.
__device__ __forceinline__ void gather_row(const Params& P, int s, int from_val) {
    int id = from_val ? P.val[s] : P.S.id[s];
    if (from_val) {
        P.S.id[s]   = id;
        P.H.inv[id] = s;
    }
    const Home& H = P.H;
    P.S.x[s] = H.pos[3 * id]; P.S.y[s] = H.pos[3 * id + 1]; P.S.z[s] = H.pos[3 * id + 2];
    P.S.u[s] = H.vel[3 * id]; P.S.v[s] = H.vel[3 * id + 1]; P.S.w[s] = H.vel[3 * id + 2];
    P.S.ax[s] = H.acc[3 * id]; P.S.ay[s] = H.acc[3 * id + 1]; P.S.az[s] = H.acc[3 * id + 2];
    P.S.p[s] = H.p[id]; P.S.minp[s] = H.minp[id]; P.S.ratio[s] = H.ratio[id];
    P.S.type[s] = H.type[id]; P.S.bc[s] = H.bc[id];
}
__global__ void gather_all_kernel(Params P, int from_val) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < P.N) gather_row(P, s, from_val);
}

// SORTED -> HOME.
__device__ __forceinline__ void scatter_row(const Params& P, int s) {
    int id = P.S.id[s];
    const Home& H = P.H;
    H.pos[3 * id] = P.S.x[s]; H.pos[3 * id + 1] = P.S.y[s]; H.pos[3 * id + 2] = P.S.z[s];
    H.vel[3 * id] = P.S.u[s]; H.vel[3 * id + 1] = P.S.v[s]; H.vel[3 * id + 2] = P.S.w[s];
    H.acc[3 * id] = P.S.ax[s]; H.acc[3 * id + 1] = P.S.ay[s]; H.acc[3 * id + 2] = P.S.az[s];
    H.p[id] = P.S.p[s]; H.minp[id] = P.S.minp[s]; H.ratio[id] = P.S.ratio[s];
    H.type[id] = P.S.type[s]; H.bc[id] = P.S.bc[s];
}
__global__ void scatter_all_kernel(Params P) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < P.N) scatter_row(P, s);
}


// ---- slab decomposition (world > 1): the build with halo particles and migration folded in -------------------
// Ownership is by ROW: rank g owns rows [g, g+1) * Np / W of the SORTED order for good; a particle changes owner
// when the sort gives it a row in another block (migration), and is a halo particle of every rank whose extent
// (RankExtents) holds its row. Sequence (mpsb_launch_cell_build):
//   key_rows_kernel      own rows of the OLD order: cell key from the SORTED position      (1/W of the particles)
//   NCCL all-gather      4 B per particle: the only replicated input
//   key_by_id_kernel     keys into id order, so that the stable sort leaves ids ascending inside a cell   (replicated)
//   radix passes         (key, id) pairs                                                             (replicated)
//   cell_range_kernel    cell ranges, active count, the NEW inverse permutation                     (replicated)
//   ext_kernel           every rank's extent in the new order, from the sorted keys (binary searches)
//   scatter_push_kernel  own rows of the OLD order: the particle's state goes into the HOME arrays of every rank
//                        whose new extent holds its new row — local stores for the bulk, NVLink peer stores for the
//                        particles that migrate or sit in a neighbour's halo
//   cross-rank barrier   (k_halo.cu)
//   gather_ext_kernel    rows of the new extent: HOME -> SORTED, all from local memory
__device__ __forceinline__ unsigned cell_key_of_position(const Params& P, int type, double x, double y, double z) {
    if (type == MPSB_GHOST) return MPSB_GHOST_KEY;
    const double len = P.s.re_max;
    int ix = (int) ((x - P.s.domain_min[0]) / len) + 1;
    int iy = (int) ((y - P.s.domain_min[1]) / len) + 1;
    int iz = (int) ((z - P.s.domain_min[2]) / len) + 1;
    const bool bad = ix < 1 || ix > P.numX - 2 || iy < 1 || iy > P.numY - 2 || iz < 1 || iz > P.numZ - 2;
    if (bad) {
        atomicOr(&P.stats->error_flags, 2);
        ix = min(max(ix, 0), P.numX - 1);
        iy = min(max(iy, 0), P.numY - 1);
        iz = min(max(iz, 0), P.numZ - 1);
    }
    return (unsigned) (ix + iy * P.numX + iz * P.numX * P.numY);
}

__global__ void key_rows_kernel(Params P) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    for (int c = i; c < P.numCells; c += gridDim.x * blockDim.x) {
        P.cell_start[c] = 0;
        P.cell_end[c]   = 0;
    }
    if (i == 0) *P.n_active = 0;
    const int s = P.row_lo + i;
    if (s >= P.row_hi) return;
    P.keyrow[s] = s < P.N ? cell_key_of_position(P, P.S.type[s], P.S.x[s], P.S.y[s], P.S.z[s]) : MPSB_GHOST_KEY;
}

__global__ void key_by_id_kernel(Params P) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.N) return;
    const int id = P.S.id[s];
    P.key[id] = P.keyrow[s];
    P.val[id] = id;
}

__global__ void cell_range_kernel(Params P) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.N) return;
    const unsigned k = P.key[s];
    if (k != MPSB_GHOST_KEY) {
        if (s == 0 || P.key[s - 1] != k) P.cell_start[k] = s;
        const unsigned next = (s == P.N - 1) ? MPSB_GHOST_KEY : P.key[s + 1];
        if (next != k) P.cell_end[k] = s + 1;
        if (next == MPSB_GHOST_KEY) *P.n_active = s + 1;
    }
    P.H.inv_alt[P.val[s]] = s;
}

// Thread q: rank q's extent in the new order. Neighbour cells of cell c have keys within c +- D, D = 1 + numX
// (+ numX numY when the domain has thickness in z), so the halo of a block whose active rows carry keys
// [kmin, kmax] is every active row with a key in [kmin - D, kmax + D].
__global__ void ext_kernel(Params P) {
    const int q = threadIdx.x, W = P.L.world;
    if (q >= W) return;
    RankExtents& E = *P.ext;
    const int per = P.Np / W, n = *P.n_active;
    const int lo = q * per, hi = lo + per;
    E.own_lo[q] = lo;
    E.own_hi[q] = hi;
    if (lo >= n) { // no active row in the block: nothing to search neighbours for
        E.ext_lo[q] = lo;
        E.ext_hi[q] = lo;
        return;
    }
    const int last = min(hi, n) - 1;
    const unsigned flat = P.s.domain_max[2] == P.s.domain_min[2] ? 0u : 1u;
    const unsigned D = 1u + (unsigned) P.numX + flat * (unsigned) (P.numX * P.numY);
    const unsigned kmin = P.key[lo], kmax = P.key[last];
    const unsigned klo = kmin > D ? kmin - D : 0u, khi = kmax + D;
    int a = 0, b = n; // first row with key >= klo
    while (a < b) {
        const int m = (a + b) >> 1;
        if (P.key[m] < klo) a = m + 1; else b = m;
    }
    E.ext_lo[q] = min(a, lo);
    a = 0; b = n; // first row with key > khi
    while (a < b) {
        const int m = (a + b) >> 1;
        if (P.key[m] <= khi) a = m + 1; else b = m;
    }
    E.ext_hi[q] = max(a, last + 1);
}

__device__ __forceinline__ bool rank_holds_row(const RankExtents& E, int q, int row) {
    return (row >= E.own_lo[q] && row < E.own_hi[q]) || (row >= E.ext_lo[q] && row < E.ext_hi[q]);
}

__global__ void __launch_bounds__(256) scatter_push_kernel(Params P) {
    const int s = P.row_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.row_hi || s >= P.N) return;
    const int id = P.S.id[s], nr = P.H.inv_alt[id];
    const RankExtents& E = *P.ext;
    const double x = P.S.x[s], y = P.S.y[s], z = P.S.z[s], u = P.S.u[s], v = P.S.v[s], w = P.S.w[s];
    const double ax = P.S.ax[s], ay = P.S.ay[s], az = P.S.az[s], pr = P.S.p[s], mp = P.S.minp[s], ra = P.S.ratio[s];
    const int ty = P.S.type[s], bc = P.S.bc[s];
    const Home& H = P.H;
    for (int q = 0; q < P.L.world; q++) {
        if (!rank_holds_row(E, q, nr)) continue;
        double *pos = H.pos, *vel = H.vel, *acc = H.acc, *hp = H.p, *hm = H.minp, *hr = H.ratio;
        int *ht = H.type, *hb = H.bc;
        if (q != P.L.rank) {
            pos = mpsb_peer(P, pos, q); vel = mpsb_peer(P, vel, q); acc = mpsb_peer(P, acc, q);
            hp = mpsb_peer(P, hp, q); hm = mpsb_peer(P, hm, q); hr = mpsb_peer(P, hr, q);
            ht = mpsb_peer(P, ht, q); hb = mpsb_peer(P, hb, q);
        }
        pos[3 * id] = x; pos[3 * id + 1] = y; pos[3 * id + 2] = z;
        vel[3 * id] = u; vel[3 * id + 1] = v; vel[3 * id + 2] = w;
        acc[3 * id] = ax; acc[3 * id + 1] = ay; acc[3 * id + 2] = az;
        hp[id] = pr; hm[id] = mp; hr[id] = ra;
        ht[id] = ty; hb[id] = bc;
    }
}

__global__ void __launch_bounds__(256) gather_ext_kernel(Params P) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P.N) return;
    const int id = P.val[s];
    P.S.id[s] = id; // the permutation is replicated: every row, whoever owns it
    if (!rank_holds_row(*P.ext, P.L.rank, s)) return;
    const Home& H = P.H;
    P.S.x[s] = H.pos[3 * id]; P.S.y[s] = H.pos[3 * id + 1]; P.S.z[s] = H.pos[3 * id + 2];
    P.S.u[s] = H.vel[3 * id]; P.S.v[s] = H.vel[3 * id + 1]; P.S.w[s] = H.vel[3 * id + 2];
    P.S.ax[s] = H.acc[3 * id]; P.S.ay[s] = H.acc[3 * id + 1]; P.S.az[s] = H.acc[3 * id + 2];
    P.S.minp[s] = H.minp[id]; P.S.ratio[s] = H.ratio[id];
    P.S.type[s] = H.type[id];
    // The halo copies of the flags and the pressure have ONE writer: the owner's halo push (k_halo.cu), which may
    // already be under way on a faster peer while this kernel still runs. Only own rows take them from HOME.
    if (s >= P.row_lo && s < P.row_hi) {
        P.S.p[s]  = H.p[id];
        P.S.bc[s] = H.bc[id];
    }
}

// ---- small systems: the whole build in ONE launch ----------------------------------------------------
// The shipped cases (458 - 6 146 particles) are bound by the length of the launch chain, and the build above
// is 10 launches, three times per step. Up to small_build_max_n particles one CTA does all of it, phases
// separated by __syncthreads: (0) SORTED -> HOME when the sorted copy is newer, (1) keys and per-cell
// counts, (2) exclusive scan of the counts, (3) placement with per-cell cursors (any order), (4) ids put in
// ascending order inside every cell (rank = smaller ids of the same cell) and the ghost tail (a scan), (5) sorted keys,
// perm, inverse perm and HOME -> SORTED. A stable sort by key has exactly one result, so every array the
// rest of the step reads is identical to the radix path's (empty cells: start == end instead of 0, 0).
constexpr int SMALL_BUILD_THREADS = 1024;
constexpr int SMALL_BUILD_MAX_CELLS = 1 << 18; // the particle limit is h->small_build_max_n

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    __syncthreads(); // warp_sums may still be read from an earlier call
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int ws = warp_sums[lane], wi = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        warp_sums[lane] = wi - ws;
        if (lane == 31) warp_sums[32] = wi;
    }
    __syncthreads();
    total = warp_sums[32];
    return warp_sums[warp] + incl - v;
}

__global__ void __launch_bounds__(SMALL_BUILD_THREADS) small_cell_build_kernel(Params P, int scatter_first) {
    __shared__ int warp_sums[33];
    const int T = SMALL_BUILD_THREADS, tid = threadIdx.x, N = P.N, C = P.numCells;
    unsigned* key_by_id = P.key_alt; // scratch: key of particle id
    int* cursor         = P.cell_end;
    // (0) + clear the counts
    if (scatter_first)
        for (int s = tid; s < N; s += T) scatter_row(P, s);
    for (int c = tid; c < C; c += T) cursor[c] = 0;
    __syncthreads();
    // (1) keys, counts
    for (int i = tid; i < N; i += T) {
        unsigned k   = cell_key_of(P, i);
        key_by_id[i] = k;
        if (k != MPSB_GHOST_KEY) atomicAdd(&cursor[k], 1);
    }
    __syncthreads();
    // (2) scan: a contiguous chunk of cells per thread
    const int per = (C + T - 1) / T, c0 = min(tid * per, C), c1 = min(c0 + per, C);
    int sum = 0;
    for (int c = c0; c < c1; c++) sum += cursor[c];
    int n_active;
    int off = block_exclusive_scan(sum, warp_sums, n_active);
    for (int c = c0; c < c1; c++) {
        int cnt         = cursor[c];
        P.cell_start[c] = off;
        cursor[c]       = off; // advanced back to start + count by the placement
        off += cnt;
    }
    // ghosts keep id order at the tail: rank = number of ghosts with a smaller id (chunked scan over ids)
    const int iper = (N + T - 1) / T, i0 = min(tid * iper, N), i1 = min(i0 + iper, N);
    int g = 0;
    for (int i = i0; i < i1; i++) g += key_by_id[i] == MPSB_GHOST_KEY;
    int n_ghost;
    int goff = n_active + block_exclusive_scan(g, warp_sums, n_ghost);
    for (int i = i0; i < i1; i++)
        if (key_by_id[i] == MPSB_GHOST_KEY) P.val_alt[goff++] = i;
    __syncthreads();
    // (3) placement
    for (int i = tid; i < N; i += T) {
        unsigned k = key_by_id[i];
        if (k != MPSB_GHOST_KEY) P.val_alt[atomicAdd(&cursor[k], 1)] = i;
    }
    __syncthreads();
    // (4) ids ascending inside every cell: the final place of an id is the cell's start plus the number of
    // smaller ids in the cell (independent loads, no sorting loop); ghosts are in order already
    for (int s = tid; s < N; s += T) {
        const int v      = P.val_alt[s];
        const unsigned k = key_by_id[v];
        int dst          = s;
        if (k != MPSB_GHOST_KEY) {
            const int b = P.cell_start[k], e = cursor[k];
            dst = b;
            for (int q = b; q < e; q++) dst += P.val_alt[q] < v;
        }
        P.val[dst] = v;
    }
    if (tid == 0) *P.n_active = n_active;
    __syncthreads();
    // (5) sorted keys, perm, state
    for (int s = tid; s < N; s += T) {
        P.key[s] = key_by_id[P.val[s]];
        gather_row(P, s, 1);
    }
}

} // namespace

void mpsb_launch_gather_all(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, gather_all_kernel, mpsb_div_up(P.N, 256), 256, 0, P, 0);
}

// End of a step / phase: one GPU writes every row back to HOME; with several ranks HOME is brought up to date
// lazily and collectively (mpsb_comm_sync_home) only when the host asks for the state.
void mpsb_launch_scatter_all(mpsb_handle* h) {
    Params& P = h->P;
    if (h->world > 1) {
        h->home_current = false;
        return;
    }
    MPSB_LAUNCH(h, scatter_all_kernel, mpsb_div_up(P.N, 256), 256, 0, P);
}

void mpsb_launch_scatter_all_rows(mpsb_handle* h) {
    Params& P = h->P;
    MPSB_LAUNCH(h, scatter_all_kernel, mpsb_div_up(P.N, 256), 256, 0, P);
}

// scatter_first: the SORTED copy is newer than HOME (it was modified since the last scatter).
void mpsb_launch_cell_build(mpsb_handle* h, bool scatter_first) {
    Params& P   = h->P;
    int N       = P.N;
    if (mpsb_div_up(RADIX * mpsb_div_up(N, SORT_TILE), SCAN_TILE) > h->scan_status_cap) { // one status word per scan tile
        snprintf(h->err, sizeof h->err, "cell build: %d scan tiles exceed the %d status words of this handle",
                 mpsb_div_up(RADIX * mpsb_div_up(N, SORT_TILE), SCAN_TILE), h->scan_status_cap);
        h->sticky = MPSB_ERR_ARG;
        return;
    }
    if (h->world > 1) {
        const int grid = mpsb_div_up(N, 256), tiles = mpsb_div_up(N, SORT_TILE);
        const int scan_tiles = mpsb_div_up(RADIX * tiles, SCAN_TILE);
        MPSB_LAUNCH(h, key_rows_kernel, mpsb_div_up(P.Np / h->world, 256), 256, 0, P);
        if (mpsb_comm_allgather_keyrow(h)) return;
        MPSB_LAUNCH(h, key_by_id_kernel, grid, 256, 0, P);
        for (int shift = 0; shift < h->sort_bits; shift += 8) {
            MPSB_LAUNCH(h, radix_hist_kernel, tiles, SORT_THREADS, 0, P.key, N, shift, tiles, h->block_hist);
            MPSB_LAUNCH(h, scan_lookback_kernel, scan_tiles, SCAN_THREADS, 0, h->block_hist, RADIX * tiles, scan_tiles,
                        h->scan_status, h->scan_status + h->scan_status_cap);
            MPSB_LAUNCH(h, radix_scatter_kernel, tiles, SORT_THREADS, 0, P.key, P.val, P.key_alt, P.val_alt, N, shift,
                        tiles, h->block_hist);
            unsigned* tk = P.key; P.key = P.key_alt; P.key_alt = tk;
            int* tv = P.val; P.val = P.val_alt; P.val_alt = tv;
        }
        MPSB_LAUNCH(h, cell_range_kernel, grid, 256, 0, P);
        MPSB_LAUNCH(h, ext_kernel, 1, 32, 0, P);
        MPSB_LAUNCH(h, scatter_push_kernel, mpsb_own_grid(h, 256), 256, 0, P);
        mpsb_launch_xbarrier(h);
        MPSB_LAUNCH(h, gather_ext_kernel, grid, 256, 0, P);
        int* ti = P.H.inv; P.H.inv = P.H.inv_alt; P.H.inv_alt = ti;
        h->home_current = false;
        return;
    }
    if (N <= h->small_build_max_n && P.numCells <= 8 * h->small_build_max_n && P.numCells <= SMALL_BUILD_MAX_CELLS) {
        MPSB_LAUNCH(h, small_cell_build_kernel, 1, SMALL_BUILD_THREADS, 0, P, scatter_first ? 1 : 0);
        return;
    }
    int grid    = mpsb_div_up(N, 256);
    int tiles   = mpsb_div_up(N, SORT_TILE);
    const int scan_tiles = mpsb_div_up(RADIX * tiles, SCAN_TILE); // fixed per handle (N is): see scan_lookback_kernel
    MPSB_LAUNCH(h, key_kernel, grid, 256, 0, P, scatter_first ? 1 : 0);
    for (int shift = 0; shift < h->sort_bits; shift += 8) {
        MPSB_LAUNCH(h, radix_hist_kernel, tiles, SORT_THREADS, 0, P.key, N, shift, tiles, h->block_hist);
        MPSB_LAUNCH(h, scan_lookback_kernel, scan_tiles, SCAN_THREADS, 0, h->block_hist, RADIX * tiles, scan_tiles,
                    h->scan_status, h->scan_status + h->scan_status_cap);
        MPSB_LAUNCH(h, radix_scatter_kernel, tiles, SORT_THREADS, 0, P.key, P.val, P.key_alt, P.val_alt, N, shift,
                    tiles, h->block_hist);
        unsigned* tk = P.key; P.key = P.key_alt; P.key_alt = tk;
        int* tv = P.val; P.val = P.val_alt; P.val_alt = tv;
    }
    MPSB_LAUNCH(h, cell_range_gather_kernel, grid, 256, 0, P);
}
