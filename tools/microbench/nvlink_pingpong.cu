// Microbenchmark: one-way latency of a flag pushed into a peer GPU's memory over NVLink (store on GPU A -> visible to a
// polling thread on GPU B), for the store / poll flavours the PCG exchanges could use. Two GPUs, one process, peer access.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nvlink_pingpong nvlink_pingpong.cu && ./nvlink_pingpong
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE> // 0 st.volatile / ld.volatile, 1 st.relaxed.sys / ld.relaxed.sys, 2 st.release.sys / ld.acquire.sys, 3 atomicExch_system / ld.volatile,
                    // 4 st.volatile + trailing __threadfence_system, 5 16-byte flag-in-data line (st.volatile.v4 / ld.volatile.v4)
__global__ void pingpong(unsigned* mine, unsigned* peer, int iters, int first, long long* cycles) {
    long long t0 = clock64();
    for (int i = 1; i <= iters; i++) {
        const unsigned v = (unsigned) i;
        if (!first) { // wait for the ping, then answer
            if (MODE == 5) { unsigned a, b, c, d; do { asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(mine) : "memory"); } while (b != v || d != v); }
            else if (MODE == 1) { unsigned x; do { asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
            else if (MODE == 2) { unsigned x; do { asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
            else { unsigned x; do { asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
        }
        if (MODE == 5) asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%1,%2};" ::"l"(peer), "r"(v + 7u), "r"(v) : "memory");
        else if (MODE == 1) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(v) : "memory");
        else if (MODE == 2) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(v) : "memory");
        else if (MODE == 3) atomicExch_system(peer, v);
        else { asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(peer), "r"(v) : "memory"); if (MODE == 4) __threadfence_system(); }
        if (first) { // wait for the answer
            if (MODE == 5) { unsigned a, b, c, d; do { asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(mine) : "memory"); } while (b != v || d != v); }
            else if (MODE == 1) { unsigned x; do { asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
            else if (MODE == 2) { unsigned x; do { asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
            else { unsigned x; do { asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(x) : "l"(mine) : "memory"); } while (x != v); }
        }
    }
    *cycles = clock64() - t0;
}

template <int MODE>
void run(const char* name, unsigned* b0, unsigned* b1, cudaStream_t s0, cudaStream_t s1, long long* c0) {
    const int iters = 20000;
    cudaSetDevice(0); cudaMemset(b0, 0, 64);
    cudaSetDevice(1); cudaMemset(b1, 0, 64); cudaDeviceSynchronize();
    cudaSetDevice(0); cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaSetDevice(1);
    long long* c1; cudaMalloc(&c1, 8);
    pingpong<MODE><<<1, 1, 0, s1>>>(b1, b0, iters, 0, c1);
    cudaSetDevice(0);
    cudaEventRecord(e0, s0);
    pingpong<MODE><<<1, 1, 0, s0>>>(b0, b1, iters, 1, c0);
    cudaEventRecord(e1, s0);
    cudaError_t err = cudaStreamSynchronize(s0);
    cudaSetDevice(1); cudaError_t err1 = cudaStreamSynchronize(s1);
    cudaSetDevice(0);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-52s one-way %.3f us  (%s %s)\n", name, 1e3 * ms / iters / 2, cudaGetErrorString(err), cudaGetErrorString(err1));
}

int main() {
    int n = 0; cudaGetDeviceCount(&n);
    if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
    int can = 0; cudaDeviceCanAccessPeer(&can, 0, 1);
    printf("peer access 0->1: %d\n", can);
    unsigned *b0, *b1; long long* c0; cudaStream_t s0, s1;
    cudaSetDevice(0); cudaDeviceEnablePeerAccess(1, 0); cudaMalloc(&b0, 64); cudaMalloc(&c0, 8); cudaStreamCreate(&s0);
    cudaSetDevice(1); cudaDeviceEnablePeerAccess(0, 0); cudaMalloc(&b1, 64); cudaStreamCreate(&s1);
    run<0>("st.volatile -> ld.volatile poll", b0, b1, s0, s1, c0);
    run<1>("st.relaxed.sys -> ld.relaxed.sys poll", b0, b1, s0, s1, c0);
    run<2>("st.release.sys -> ld.acquire.sys poll", b0, b1, s0, s1, c0);
    run<3>("atomicExch_system -> ld.volatile poll", b0, b1, s0, s1, c0);
    run<4>("st.volatile + __threadfence_system -> ld.volatile", b0, b1, s0, s1, c0);
    run<5>("16-byte flag-in-data line (v4 volatile)", b0, b1, s0, s1, c0);
    return 0;
}
