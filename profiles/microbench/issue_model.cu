// Does an FP64 warp instruction on B200 occupy the SMSP dispatch port for 2 cycles (so that integer /
// LDS instructions are NOT free next to a DFMA stream), or do the pipes overlap?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a issue_model.cu -o issue_model && ./issue_model
// Per loop iteration and thread: 8 independent DFMA plus 8*K independent integer LOP3/IADD (K = 0..3)
// or K LDS.64.  Prints cycles per warp-iteration per SMSP.
#include <cstdio>
#include <cuda_runtime.h>

template <int K, int MODE>
__global__ void __launch_bounds__(256) kern(double* out, int iters, unsigned seed) {
    __shared__ double tab[1024];
    for (int i = threadIdx.x; i < 1024; i += 256) tab[i] = 1.0 + i * 1e-9;
    __syncthreads();
    double a[8];
    unsigned x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 1e-9 + i; x[i] = seed + threadIdx.x * 8 + i; }
    const double b = 1.0000001, c = 1e-9;
    double s = 0.0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(b), "d"(c));
            if (MODE == 0) {
#pragma unroll
                for (int k = 0; k < K; ++k) asm volatile("xor.b32 %0, %0, %1;" : "+r"(x[i]) : "r"(seed + k));
            } else if (MODE == 1) {
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    double t;
                    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t) : "r"((unsigned)__cvta_generic_to_shared(tab) + ((x[i] + k * 16) & 127) * 8));
                    s += 0.0 * 0 + (t > 3.0 ? 1.0 : 0.0) * 0;   // keep the load alive without FP64 work
                    x[i] ^= (unsigned)__double2loint(t) & 0;
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; ++k) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(seed | 1), "r"(k));
            }
        }
    }
    double r = s;
    unsigned y = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { r += a[i]; y ^= x[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + y;
}

template <int K, int MODE>
void run(const char* name, double* out, int sms, double mhz) {
    const int iters = 1 << 14, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kern<K, MODE><<<blocks, 256>>>(out, iters, 12345u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    // warps per SMSP = 8 blocks * 8 warps / 4 = 16; per warp-iteration: 8 DFMA + 8K other
    const double cyc = best * 1e-3 * mhz * 1e6 / ((double)iters * 16);
    printf("%-8s K=%d  %8.3f ms  %6.2f cycles per (8 DFMA + %d other) per SMSP  -> %5.2f cycles per DFMA-group\n", name, K, best,
           cyc, 8 * K, cyc / 8);
}

// DFMA dependent-issue latency / throughput versus independent chains per warp and warps per SMSP
template <int CH>
__global__ void lat_kern(double* out, int iters) {
    double a[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) a[i] = threadIdx.x * 1e-9 + i;
    const double b = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 8; ++rep)
#pragma unroll
            for (int i = 0; i < CH; ++i) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(b), "d"(c));
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int CH>
void run_lat(double* out, int sms, double mhz, int warps_per_smsp) {
    const int iters = 1 << 13;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        lat_kern<CH><<<sms, 128 * warps_per_smsp>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double cyc = best * 1e-3 * mhz * 1e6 / ((double)iters * 8);   // cycles per round of CH DFMAs per warp
    printf("chains %d warps/SMSP %d: %7.2f cycles per dependent step, %5.2f cycles per DFMA per SMSP\n", CH, warps_per_smsp, cyc,
           cyc / (CH * warps_per_smsp));
}

int main() {
    int sms, khz;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    printf("SMs %d, clock %.0f MHz (nominal max; cycles assume the GPU runs at it)\n", sms, mhz);
    double* out; cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    run<0, 0>("xor", out, sms, mhz); run<1, 0>("xor", out, sms, mhz); run<2, 0>("xor", out, sms, mhz); run<3, 0>("xor", out, sms, mhz);
    run<1, 2>("imad", out, sms, mhz); run<2, 2>("imad", out, sms, mhz);
    run<1, 1>("lds64", out, sms, mhz); run<2, 1>("lds64", out, sms, mhz);
    for (int w : {1, 2, 4, 6, 8}) { run_lat<1>(out, sms, mhz, w); run_lat<2>(out, sms, mhz, w); run_lat<4>(out, sms, mhz, w); }
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
