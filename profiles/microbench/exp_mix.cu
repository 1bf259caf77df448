// Which non-FP64 instruction of the table-driven exp keeps the FP64 pipe from saturating?
// Variants drop one instruction class at a time (results then wrong on purpose) and report
// cycles per warp-exp per SMSP; 10 FP64 instructions per exp (incl. the accumulate) -> 20 is the floor.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr double kMagic = 6755399441055744.0, kScale = 92.332482616893656877, kLn2_64 = 1.0830424696249145e-02;

// MASK bits: 1 clamp (VIMNMX), 2 table address (AND+IMAD), 4 LDS, 8 exponent patch (IMAD), 16 argument update (IADD)
template <int MASK>
__global__ void __launch_bounds__(256) kern(double* out, int iters, const double* __restrict__ tab) {
    __shared__ __align__(16) double etab[64 * 16];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) etab[i] = tab[i / 16];
    __syncthreads();
    const uint32_t tab_lane = (uint32_t)__cvta_generic_to_shared(etab) + (threadIdx.x & 15) * 8u;
    double t[8], s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { t[i] = -1e-3 * (threadIdx.x + 1) * (i + 1); s[i] = 0.0; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            double u = t[i];
            if (MASK & 1) {
                const unsigned hi = min((unsigned)__double2hiint(u), 0xC085E000u);
                u = __hiloint2double((int)hi, __double2loint(u));
            }
            double kd = fma(u, kScale, kMagic);
            const int k = __double2loint(kd);
            kd -= kMagic;
            const double r = fma(kd, -kLn2_64, u);
            uint32_t addr = tab_lane;
            if (MASK & 2) addr = tab_lane + ((k & 63) << 7);
            int t_lo = 0, t_hi = 0x3ff00000;
            if (MASK & 4) asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(t_lo), "=r"(t_hi) : "r"(addr));
            else { t_lo = (int)addr; }
            if (MASK & 8) t_hi += k << 14;
            const double T = __hiloint2double(t_hi, t_lo);
            double p = fma(r, 4.1666666666666664e-02, 1.6666666666666666e-01);
            p = fma(p, r, 0.5);
            p = fma(p, r, 1.0);
            const double q = r * p;
            s[i] = fma(0.5, fma(T, q, T), s[i]);
            if (MASK & 16) t[i] = __hiloint2double(__double2hiint(t[i]), __double2loint(t[i]) + 64);
        }
    }
    double r = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += s[i] + t[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MASK>
void run(const char* name, double* out, const double* tab, int sms, double mhz) {
    const int iters = 1 << 12, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kern<MASK><<<blocks, 256>>>(out, iters, tab);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double cyc = best * 1e-3 * mhz * 1e6 / ((double)iters * 8 * 16);
    printf("%-44s %7.3f ms  %5.2f cycles per warp-exp per SMSP  (FP64 floor 20)\n", name, best, cyc);
}

int main() {
    int sms, khz;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    double *out, *tab;
    cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    cudaMalloc(&tab, 64 * 8);
    double h[64];
    for (int j = 0; j < 64; ++j) h[j] = 1.0 + j / 64.0;
    cudaMemcpy(tab, h, sizeof h, cudaMemcpyHostToDevice);
    run<31>("full exp", out, tab, sms, mhz);
    run<30>("no clamp (VIMNMX)", out, tab, sms, mhz);
    run<29>("no table address (AND+IMAD)", out, tab, sms, mhz);
    run<27>("no LDS", out, tab, sms, mhz);
    run<23>("no exponent patch (IMAD)", out, tab, sms, mhz);
    run<15>("no argument update (IADD)", out, tab, sms, mhz);
    run<16>("FP64 only + argument update", out, tab, sms, mhz);
    run<0>("FP64 only", out, tab, sms, mhz);
    run<4>("FP64 + LDS", out, tab, sms, mhz);
    run<1>("FP64 + clamp", out, tab, sms, mhz);
    run<8>("FP64 + patch", out, tab, sms, mhz);
    run<2>("FP64 + address", out, tab, sms, mhz);
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
