// Throughput of the FP64 instruction forms kernel (c') uses: DFMA/DMUL/DADD with register,
// immediate and uniform-register operands.  8 independent chains x 16 warps per SMSP.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a fp64_forms.cu -o fp64_forms && ./fp64_forms
#include <cstdio>
#include <cuda_runtime.h>

template <int FORM>
__global__ void __launch_bounds__(256) kern(double* out, int iters, double ub, double uc) {
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i * 1e-3;
    const double b = 1.0 + threadIdx.x * 1e-12, c = threadIdx.x * 1e-13;   // per-thread registers
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (FORM == 0) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(b), "d"(c));
            if (FORM == 1) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(a[i]) : "d"(b));
            if (FORM == 2) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(a[i]) : "d"(c));
            if (FORM == 3) asm volatile("fma.rn.f64 %0, %0, %1, 0d4338000000000000;" : "+d"(a[i]) : "d"(b));
            if (FORM == 4) asm volatile("add.rn.f64 %0, %0, 0dC338000000000000;" : "+d"(a[i]));
            if (FORM == 5) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(ub), "d"(uc));   // kernel params -> UR/const
            if (FORM == 6) asm volatile("fma.rn.f64 %0, %0, %1, 0d3FE0000000000000;" : "+d"(a[i]) : "d"(b));   // + 0.5
            if (FORM == 7) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(a[i]) : "d"(-b));
            if (FORM == 8) asm volatile("fma.rn.f64 %0, %1, %0, %1;" : "+d"(a[i]) : "d"(b));             // fma(T, q, T)
        }
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int FORM>
void run(const char* name, double* out, int sms, double mhz) {
    const int iters = 1 << 14, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kern<FORM><<<blocks, 256>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double cyc = best * 1e-3 * mhz * 1e6 / ((double)iters * 16 * 8);
    printf("%-34s %8.3f ms  %5.2f cycles per warp instruction per SMSP\n", name, best, cyc);
}

int main() {
    int sms, khz;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    double* out; cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    run<0>("DFMA r,r,r", out, sms, mhz);
    run<1>("DMUL r,r", out, sms, mhz);
    run<2>("DADD r,r", out, sms, mhz);
    run<3>("DFMA r,r,imm(1.5*2^52)", out, sms, mhz);
    run<4>("DADD r,imm(-1.5*2^52)", out, sms, mhz);
    run<5>("DFMA r,param,param (UR/const)", out, sms, mhz);
    run<6>("DFMA r,r,imm(0.5)", out, sms, mhz);
    run<7>("DMUL r,-r", out, sms, mhz);
    run<8>("DFMA T,q,T", out, sms, mhz);
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
