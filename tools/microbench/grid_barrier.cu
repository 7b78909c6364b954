// Microbenchmark: cost of one grid-wide barrier (148 CTAs x 640 threads, one CTA per SM) for several designs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o grid_barrier grid_barrier.cu && ./grid_barrier
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

#define THREADS 640
#define STRIDE 1056 // u32 between per-CTA slots (spread over L2 slices)

struct Bufs {
    unsigned long long* ctr;  // counters
    unsigned* slots;          // [2][nb + 8][STRIDE]
    unsigned* packed;         // [2][256] contiguous flags
    double* junk;             // per-thread store target (simulates sweep stores before the fence)
};

__device__ __forceinline__ unsigned ld_vol(const unsigned* p) { unsigned v; asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_rlx(const unsigned* p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_acq(const unsigned* p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_vol(unsigned* p, unsigned v) { asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_rlx(unsigned* p, unsigned v) { asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_rel(unsigned* p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// V1: one counter, thread 0 polls it
__device__ void bar_counter(Bufs B, unsigned ep) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(B.ctr, 1ull);
        const unsigned long long want = (unsigned long long) ep * gridDim.x;
        while (*(volatile unsigned long long*) B.ctr < want) {}
        __threadfence();
    }
    __syncthreads();
}
// V2: 16 group counters + top counter
__device__ void bar_hier(Bufs B, unsigned ep) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const int G = 16, g = blockIdx.x % G;
        const int members = (gridDim.x - g + G - 1) / G;
        unsigned long long* gc = B.ctr + 32 * (1 + g); // 256 B apart
        if (atomicAdd(gc, 1ull) + 1ull == (unsigned long long) ep * members) {
            __threadfence();
            atomicAdd(B.ctr, 1ull);
        }
        const unsigned long long want = (unsigned long long) ep * G;
        while (*(volatile unsigned long long*) B.ctr < want) {}
        __threadfence();
    }
    __syncthreads();
}
// V3: leader gathers 4-byte flags (one 32-byte-aligned slot per CTA, spread), broadcasts one result word (8 replicas)
__device__ __forceinline__ void st_atom(unsigned* p, unsigned v) { atomicExch(p, v); }
template <int MODE> // 0: volatile, 1: relaxed.gpu + fences, 2: release/acquire (no fences), 3: atomicExch flags + fences, 4: relaxed store + trailing fence
__device__ void bar_leader(Bufs B, unsigned ep) {
    const int nb = gridDim.x;
    unsigned* base = B.slots + (size_t) (ep & 1u) * (nb + 8) * STRIDE;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        if (lane == 0) {
            if (MODE == 2) st_rel(base + (size_t) blockIdx.x * STRIDE, ep);
            else { __threadfence(); if (MODE == 0) st_vol(base + (size_t) blockIdx.x * STRIDE, ep); else if (MODE == 3) st_atom(base + (size_t) blockIdx.x * STRIDE, ep); else st_rlx(base + (size_t) blockIdx.x * STRIDE, ep);
                   if (MODE == 4) __threadfence(); }
        }
        if (blockIdx.x == 0) {
            for (int c = lane; c < nb; c += 32) {
                const unsigned* src = base + (size_t) c * STRIDE;
                if (MODE == 0) while (ld_vol(src) != ep) {}
                else if (MODE == 2) while (ld_acq(src) != ep) {}
                else while (ld_rlx(src) != ep) {}
            }
            __syncwarp();
            if (lane == 0) {
                if (MODE != 2) __threadfence();
                for (int r = 0; r < 8; r++) {
                    unsigned* dst = base + (size_t) (nb + r) * STRIDE;
                    if (MODE == 0) st_vol(dst, ep); else if (MODE == 2) st_rel(dst, ep); else if (MODE == 3) st_atom(dst, ep); else st_rlx(dst, ep);
                }
                if (MODE == 4) __threadfence();
            }
        } else if (lane == 0) {
            const unsigned* src = base + (size_t) (nb + blockIdx.x % 8) * STRIDE;
            if (MODE == 0) while (ld_vol(src) != ep) {}
            else if (MODE == 2) while (ld_acq(src) != ep) {}
            else while (ld_rlx(src) != ep) {}
            if (MODE != 2) __threadfence();
        }
    }
    __syncthreads();
}
// leader gather with flag-in-data 16-byte lines: data+flag delivered by two 64-bit atomic exchanges, result likewise; returns the sum
__device__ double bar_leader_data(Bufs B, unsigned ep, double v) {
    const int nb = gridDim.x;
    unsigned* base = B.slots + (size_t) (ep & 1u) * (nb + 8) * STRIDE;
    __shared__ double res;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        auto put = [&](unsigned* dst, double x) {
            unsigned long long w0 = ((unsigned long long) ep << 32) | (unsigned) __double2loint(x);
            unsigned long long w1 = ((unsigned long long) ep << 32) | (unsigned) __double2hiint(x);
            atomicExch((unsigned long long*) dst, w0);
            atomicExch((unsigned long long*) dst + 1, w1);
        };
        auto get = [&](const unsigned* src) -> double {
            unsigned x, f0, y, f1;
            for (;;) {
                asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(f0), "=r"(y), "=r"(f1) : "l"(src) : "memory");
                if (f0 == ep && f1 == ep) break;
            }
            return __hiloint2double((int) y, (int) x);
        };
        if (lane == 0) { __threadfence(); put(base + (size_t) blockIdx.x * STRIDE, v); }
        if (blockIdx.x == 0) {
            double s = 0.0;
            for (int c = lane; c < nb; c += 32) s += get(base + (size_t) c * STRIDE);
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) {
                __threadfence();
                for (int r = 0; r < 8; r++) put(base + (size_t) (nb + r) * STRIDE, s);
                res = s;
            }
        } else if (lane == 0) {
            res = get(base + (size_t) (nb + blockIdx.x % 8) * STRIDE);
            __threadfence();
        }
    }
    __syncthreads();
    return res;
}
// V5: all-to-all over packed 4-byte flags
__device__ void bar_all2all(Bufs B, unsigned ep) {
    const int nb = gridDim.x;
    unsigned* base = B.packed + (size_t) (ep & 1u) * 256;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        if (lane == 0) { __threadfence(); st_rlx(base + blockIdx.x, ep); }
        for (int c = lane; c < nb; c += 32) while (ld_rlx(base + c) != ep) {}
        __syncwarp();
        if (lane == 0) __threadfence();
    }
    __syncthreads();
}

template <int V>
__global__ void __launch_bounds__(THREADS, 1) bench(Bufs B, int iters, int work, unsigned ep0, long long* cycles) {
    cg::grid_group grid = cg::this_grid();
    unsigned ep = ep0;
    const int tid = blockIdx.x * THREADS + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        for (int w = 0; w < work; w++) B.junk[(size_t) w * gridDim.x * THREADS + tid] = (double) (i + w); // stores the fence must drain
        ep++;
        if (V == 0) grid.sync();
        else if (V == 1) bar_counter(B, ep);
        else if (V == 2) bar_hier(B, ep);
        else if (V == 3) bar_leader<0>(B, ep);
        else if (V == 4) bar_leader<1>(B, ep);
        else if (V == 5) bar_leader<2>(B, ep);
        else if (V == 6) bar_all2all(B, ep);
        else if (V == 7) bar_leader<3>(B, ep);
        else if (V == 8) bar_leader<4>(B, ep);
        else if (V == 9) { double r = bar_leader_data(B, ep, 1.0); if (r != (double) gridDim.x) __trap(); }
    }
    if (tid == 0) *cycles = clock64() - t0;
}

template <int V>
void run(const char* name, Bufs B, int nb, int work) {
    const int iters = 2000;
    long long* cyc;
    cudaMalloc(&cyc, 8);
    cudaMemset(B.ctr, 0, 8 * 1024);
    cudaMemset(B.slots, 0, sizeof(unsigned) * 2 * (nb + 8) * STRIDE);
    cudaMemset(B.packed, 0, sizeof(unsigned) * 512);
    unsigned ep0 = 0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int it = iters;
    void* args[] = {&B, &it, &work, &ep0, &cyc};
    cudaEventRecord(e0);
    cudaError_t err = cudaLaunchCooperativeKernel((const void*) bench<V>, dim3(nb), dim3(THREADS), args, 0, 0);
    cudaEventRecord(e1);
    cudaError_t e2 = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long hc = 0;
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s work=%d  %7.3f us/barrier  %8.0f cycles  (%s %s)\n", name, work, 1e3 * ms / iters, (double) hc / iters, cudaGetErrorString(err),
           cudaGetErrorString(e2));
    cudaFree(cyc);
}

int main() {
    int dev = 0, nsm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    Bufs B;
    cudaMalloc(&B.ctr, 8 * 1024);
    cudaMalloc(&B.slots, sizeof(unsigned) * 2 * (nsm + 8) * STRIDE);
    cudaMalloc(&B.packed, sizeof(unsigned) * 512);
    cudaMalloc(&B.junk, sizeof(double) * 8 * (size_t) nsm * THREADS);
    printf("SMs %d, %d threads per CTA\n", nsm, THREADS);
    for (int work : {0, 4}) {
        run<0>("cg grid.sync", B, nsm, work);
        run<1>("one counter, atomicAdd + volatile poll", B, nsm, work);
        run<2>("16 group counters + top counter", B, nsm, work);
        run<3>("leader gather, volatile flags + fences", B, nsm, work);
        run<4>("leader gather, relaxed.gpu flags + fences", B, nsm, work);
        run<5>("leader gather, release/acquire flags", B, nsm, work);
        run<6>("all-to-all packed relaxed flags + fences", B, nsm, work);
        run<7>("leader gather, atomicExch flags + fences", B, nsm, work);
        run<8>("leader gather, relaxed store + trailing fence", B, nsm, work);
        run<9>("leader gather+sum, 16 B lines by atomicExch64", B, nsm, work);
    }
    return 0;
}
