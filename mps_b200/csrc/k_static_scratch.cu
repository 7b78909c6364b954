// mpsb_courant_static: the Courant number of a particle set WITHOUT any allocation by the library.
//
// Who needs this. mpsb_create builds a handle with cudaMalloc; some programs that want to call into this library
// are not allowed to do that at the moment they call: a host that has handed all of its device memory to its own
// pool allocator and must not let a second allocator fragment it; code running inside a region where the CUDA
// runtime refuses to allocate or where an implicit device synchronisation (cudaMalloc / cudaFree are synchronising)
// would stall other streams; a monitoring or watchdog thread that wants the CFL number of a live simulation state
// it only holds device pointers to. For those callers this file offers one reduction of the step — the Courant
// number, reference MPS::getCourantNumber (src/mps.cpp:547-571) — as a free function over caller-owned device
// arrays. Its scratch memory (one partial maximum per thread block and a ticket counter) does not come from an
// allocator: it is a file-scope __device__ array that is part of the loaded module on every GPU.
//
// What that costs and what the caller must know:
//   * every GPU on which this library is loaded holds MPSB_STATIC_SCRATCH_BYTES (256 KiB + 256 B) until the process exits;
//   * at most MPSB_STATIC_MAX_PARTICLES (67 108 864) particles per call — larger calls are refused, nothing is launched;
//   * all calls on one GPU share that one array. It keeps whatever the previous call left in it; the entry point clears
//     the ticket counter on the caller's stream before each launch, and every block overwrites its partial before it is
//     read, so calls that run one after the other are independent. Two calls that may overlap on the GPU (different
//     streams without an ordering between them, or two host threads) race on the array and return garbage: serialise them;
//   * the first call on a device looks up the device address of the array (cudaGetSymbolAddress, once per device, with
//     that device current). Make that first call outside any stream capture.
//
// The arithmetic is the one of courant_kernel (k_explicit.cu): (sqrt((u*u + v*v) + w*w) * dt) / l0 with every operation
// rounded on its own (-fmad=false), and a maximum does not depend on the order it is taken in, so the result is
// bit-identical to mpsb_run_phase(MPSB_PH_COURANT) and to the reference's serial loop.
#include "../../include/mps_b200.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

namespace {

constexpr int THREADS = 256;
constexpr int PER_THREAD = 8;
constexpr int TILE = THREADS * PER_THREAD;      // particles per thread block
constexpr int MAX_BLOCKS = 32768;               // one partial maximum each
constexpr int HEADER_BYTES = 256;               // ticket counter, padded so the partials stay 256-byte aligned
constexpr long long MAX_PARTICLES = (long long) MAX_BLOCKS * TILE;
constexpr size_t SCRATCH_BYTES = HEADER_BYTES + sizeof(double) * (size_t) MAX_BLOCKS;
constexpr int MAX_DEVICES = 64;

static_assert(MAX_PARTICLES == MPSB_STATIC_MAX_PARTICLES, "header and kernel disagree on the particle limit");
static_assert(SCRATCH_BYTES == MPSB_STATIC_SCRATCH_BYTES, "header and kernel disagree on the scratch size");

// The scratch memory itself. In host code this name is only a handle for the runtime's symbol table, NOT an address
// a kernel could use; device_scratch() below asks the runtime for the real device address once per device.
static __device__ __align__(256) unsigned char g_static_scratch[SCRATCH_BYTES];

__global__ void __launch_bounds__(THREADS) courant_static_kernel(const int* __restrict__ type,
                                                                 const double* __restrict__ vel, long long n, double dt,
                                                                 double l0, unsigned char* scratch, double* out) {
    unsigned int* ticket = (unsigned int*) scratch;
    double* partial = (double*) (scratch + HEADER_BYTES);
    __shared__ double warp_max[THREADS / 32];
    __shared__ bool last;

    double c = 0.0;
    const long long base = (long long) blockIdx.x * TILE;
#pragma unroll
    for (int k = 0; k < PER_THREAD; k++) {
        const long long i = base + (long long) k * THREADS + threadIdx.x;
        if (i < n && type[i] == MPSB_FLUID) {
            const double u = vel[3 * i], v = vel[3 * i + 1], w = vel[3 * i + 2];
            const double cc = (sqrt((u * u + v * v) + w * w) * dt) / l0;
            if (cc > c) c = cc; // std::max(max, c): a NaN never replaces the running maximum
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c = fmax(c, __shfl_xor_sync(0xffffffffu, c, o));
    if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < THREADS / 32; w++) c = fmax(c, warp_max[w]);
        partial[blockIdx.x] = c;
        __threadfence(); // the partial is visible before the ticket says so
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the block that drew the last ticket folds this launch's partials (and only those: slots past gridDim.x may
    // still hold an earlier, larger call's values)
    c = 0.0;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += THREADS) c = fmax(c, __ldcg(partial + b));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c = fmax(c, __shfl_xor_sync(0xffffffffu, c, o));
    if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < THREADS / 32; w++) c = fmax(c, warp_max[w]);
        *out = c;
    }
}

std::mutex g_mu;
unsigned char* g_scratch_addr[MAX_DEVICES]; // device address of g_static_scratch on each device, nullptr until looked up

void say(char* text, int cap, const char* what, const char* detail) {
    if (text && cap > 0) snprintf(text, (size_t) cap, "mpsb_courant_static: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
}

// Device address of the scratch array on the CURRENT device (which must be `device`); looked up once per device.
cudaError_t device_scratch(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = (unsigned char*) p;
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

} // namespace

extern "C" int mpsb_courant_static(const int32_t* d_type, const double* d_velocity, int64_t n, double dt,
                                   double particle_distance, int device, void* cuda_stream, double* d_courant,
                                   char* error_text, int error_capacity) {
    if (error_text && error_capacity > 0) error_text[0] = 0;
    if (n < 0 || !d_courant || (n > 0 && (!d_type || !d_velocity))) {
        say(error_text, error_capacity, "bad argument (null device pointer or negative particle count)", nullptr);
        return MPSB_ERR_ARG;
    }
    if (n > MAX_PARTICLES) {
        char detail[160];
        snprintf(detail, sizeof detail, "%lld particles asked for, the static scratch array serves at most %lld per call",
                 (long long) n, MAX_PARTICLES);
        say(error_text, error_capacity, "too many particles", detail);
        return MPSB_ERR_ARG;
    }
    if (device < 0 || device >= MAX_DEVICES) {
        say(error_text, error_capacity, "bad device ordinal", nullptr);
        return MPSB_ERR_ARG;
    }
    int previous = -1;
    cudaError_t e = cudaGetDevice(&previous);
    if (e == cudaSuccess && previous != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        say(error_text, error_capacity, "no CUDA device (there is no CPU path)", cudaGetErrorString(e));
        return MPSB_ERR_CUDA;
    }
    cudaStream_t stream = (cudaStream_t) cuda_stream;
    unsigned char* scratch = nullptr;
    const char* step = "cudaGetSymbolAddress(static scratch)";
    e = device_scratch(device, &scratch);
    if (e == cudaSuccess) {
        step = "clearing the ticket counter";
        e = cudaMemsetAsync(scratch, 0, HEADER_BYTES, stream);
    }
    if (e == cudaSuccess) {
        step = "launching courant_static_kernel";
        const unsigned int blocks = n > 0 ? (unsigned int) ((n + TILE - 1) / TILE) : 1u;
        courant_static_kernel<<<blocks, THREADS, 0, stream>>>(d_type, d_velocity, (long long) n, dt, particle_distance,
                                                              scratch, d_courant);
        e = cudaGetLastError();
    }
    if (previous != device) cudaSetDevice(previous);
    if (e != cudaSuccess) {
        char detail[200];
        snprintf(detail, sizeof detail, "%s -> %s", step, cudaGetErrorString(e));
        say(error_text, error_capacity, "CUDA failure", detail);
        return MPSB_ERR_CUDA;
    }
    return MPSB_OK;
}
