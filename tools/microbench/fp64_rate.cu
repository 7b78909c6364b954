// Measures per-SM issue rates that bound the matrix-free sweep A (k_pcg.cu): dependent-free DFMA chains, MUFU.RSQ64H,
// and 16-byte L1-resident gathers. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_rate tools/microbench/fp64_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters) {
    double a[8];
    for (int k = 0; k < 8; k++) a[k] = threadIdx.x * 1e-3 + k;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = fma(a[k], m, c);
    }
    double s = 0;
    for (int k = 0; k < 8; k++) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void rsq_kernel(double* out, int iters) {
    double a[8];
    for (int k = 0; k < 8; k++) a[k] = 1.0 + threadIdx.x * 1e-3 + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) asm volatile("rsqrt.approx.ftz.f64 %0, %0;" : "+d"(a[k]));
    }
    double s = 0;
    for (int k = 0; k < 8; k++) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 32 lanes gather 16-byte records at a stride pattern typical of the sweep (neighbouring rows, 1-3 lines per request)
__global__ void gather_kernel(const double2* __restrict__ tab, double* out, int iters, int spread) {
    const int lane = threadIdx.x & 31;
    int idx = (threadIdx.x >> 5) * 64 + (lane * spread) % 97;
    double s = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            double2 v = tab[(idx + k * 5) & 4095];
            s += v.x;
            idx += (int) v.y; // v.y == 0: keeps the loads dependent-free but not hoistable
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
float time_ms(F f) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f();
    cudaEventRecord(a); f(); cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount, threads = 640, iters = 20000;
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double* out; cudaMalloc(&out, sizeof(double) * sms * threads);
    double2* tab; cudaMalloc(&tab, sizeof(double2) * 4096); cudaMemset(tab, 0, sizeof(double2) * 4096);
    float ms = time_ms([&] { dfma_kernel<<<sms, threads>>>(out, iters); });
    double ops = (double) sms * threads * iters * 8;
    printf("DFMA: %.1f G lane-ops/s/SM = %.1f lanes/clk/SM at the nominal %.0f MHz (%.2f TFLOP/s chip)\n", ops / ms / 1e6 / sms,
           ops / (ms * 1e-3) / sms / (khz * 1e3), khz / 1e3, 2 * ops / ms / 1e9);
    ms = time_ms([&] { rsq_kernel<<<sms, threads>>>(out, iters); });
    printf("MUFU.RSQ64H: %.1f lanes/clk/SM\n", ops / (ms * 1e-3) / sms / (khz * 1e3));
    for (int spread : {1, 2, 4}) {
        ms = time_ms([&] { gather_kernel<<<sms, threads>>>(tab, out, iters / 4, spread); });
        double reqs = (double) sms * (threads / 32) * (iters / 4) * 8;
        printf("LDG.128 gather (lane stride %d records): %.2f clk per warp request per SM\n", spread,
               (ms * 1e-3) * (khz * 1e3) / (reqs / sms));
    }
    return 0;
}
