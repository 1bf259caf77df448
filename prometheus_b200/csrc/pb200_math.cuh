// pb200_math.cuh -- fp64 device math of the transit hot path (sm_100a).
//
//   pb_exp_neg      exp(-t), t >= 0        inner loop of kernel (c'): Beer-Lambert, gasProperties.py:1241
//   pb_interp_log   np.interp + 10**v - offset   == n_interp_log, gasProperties.py:34-51
//   pb_faddeeva_re  Re w(x+iy)             == scipy.special.wofz(...).real inside voigt_profile
//   pb_voigt        scipy.special.voigt_profile(x, sigma, gamma), call site gasProperties.py:687-688
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace pb200 {

// constants of the reference (pythonScripts/constants.py:13-17); mirrored in constants.py
constexpr double kElemCharge = 4.803e-10;
constexpr double kElectronMass = 9.109e-28;
constexpr double kLightSpeed = 2.998e10;
constexpr double kBoltzmann = 1.381e-16;
constexpr double kPi = 3.141592653589793;

// ---------------------------------------------------------------------------------------------
// exp(-t) for t >= 0.  Cody-Waite reduction t = -k ln2 + r, |r| <= ln2/2, degree-11 polynomial
// (max relative error 2e-17 before rounding, measured against long double), exponent patched
// in with integer adds.  t is clamped at 700 (exp(-700) = 1e-304 stands in for "0"), which
// keeps the result normal and the code branch-free.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double pb_exp_neg(double t) {
    t = fmin(t, 700.0);
    const double kMagic = 6755399441055744.0;  // 1.5 * 2^52: rint() through the adder
    double kd = fma(t, -1.4426950408889634, kMagic);
    const int k = __double2loint(kd);
    kd -= kMagic;
    double r = fma(kd, -6.93147180369123816490e-01, -t);
    r = fma(kd, -1.90821492927058770002e-10, r);
    double p = 0x1.af354eec9216ep-26;
    p = fma(p, r, 0x1.28917965df995p-22);
    p = fma(p, r, 0x1.71de0fd1838abp-19);
    p = fma(p, r, 0x1.a019b9189c259p-16);
    p = fma(p, r, 0x1.a01a01a6cffcap-13);
    p = fma(p, r, 0x1.6c16c178884cep-10);
    p = fma(p, r, 0x1.1111111110c74p-7);
    p = fma(p, r, 0x1.5555555553d6fp-5);
    p = fma(p, r, 0x1.5555555555555p-3);
    p = fma(p, r, 0x1.0000000000001p-1);
    p = fma(p, r * r, r);
    p += 1.0;
    return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

// ---------------------------------------------------------------------------------------------
// exp(u) for u <= 0 -- the Beer-Lambert inner loop of kernel (c') -- with a 256-entry table:
//   k = rint(u * 256/ln2) = 256 e + j,  r = u - k ln2/256  (|r| <= ln2/512 = 1.35e-3)
//   exp(u) = 2^e * T[j] * (1 + r + r^2/2 + r^3/6),   T[j] = 2^(j/256)
// 8 FP64-pipe instructions (+1 for u = N*sigma, +1 for the accumulate) instead of 18 for a
// polynomial-only exp: the kernel is bound by the FP64 pipe (64 lanes/SM, 2 cycles per warp
// instruction) and, to a lesser degree, by the integer instructions issued next to it
// (profiles/microbench/exp_mix.cu), so both are kept minimal:
//   * table entry j is stored with (j << 12) subtracted from its high word, so adding (k << 12)
//     gives the high word of 2^e * 2^(j/256) in one integer op;
//   * the table is laid out [j][8] in shared memory: a half-warp's 16 lanes hit 8 bank pairs,
//     at most 2-way conflicts, which the otherwise idle LSU absorbs;
//   * the clamp u >= -700 (keeps 2^e normal) is only compiled into the CLAMP variant, used by warps
//     whose largest possible tau can exceed 690.
// Accuracy: truncation r^4/24 <= 1.4e-13 relative, reduction error <= 1.1e-16 |k| ln2/256.
// ---------------------------------------------------------------------------------------------
constexpr int kExpTabEntries = 256;
constexpr int kExpTabCopies = 8;
constexpr int kExpTabShift = 12;                           // 20 - log2(kExpTabEntries)
constexpr double kExpMagic = 6755399441055744.0;           // 1.5 * 2^52: rint() through the adder
constexpr double kExpScale = 369.32993046757462751;        // 256 / ln 2
constexpr double kExpStep = 2.7076061740622862e-03;        // ln 2 / 256
constexpr double kExpClampTau = 690.0;

__host__ __device__ inline double pb_exp_tab_entry(double two_pow_frac, int j) {
    union { double d; uint64_t u; } v;
    v.d = two_pow_frac;
    v.u -= (uint64_t)j << (kExpTabShift + 32);
    return v.d;
}

// T = 2^e * 2^(j/256) for k = 256 e + j.  tab_addr: shared-window byte address of this lane's column.
// AND + IMAD + LDS.64 + IMAD, all off the FP64 pipe.
template <int COPIES = kExpTabCopies>
__device__ __forceinline__ double pb_exp_tab_load(uint32_t tab_addr, int k) {
    int t_lo, t_hi;
    asm volatile("{\n"
        ".reg .b32 j, a;\n"
        "and.b32 j, %2, 255;\n"
        "mad.lo.u32 a, j, %4, %3;\n"
        "ld.shared.v2.b32 {%0, %1}, [a];\n"
        "}\n"
        : "=r"(t_lo), "=r"(t_hi)
        : "r"(k), "r"(tab_addr), "n"(COPIES * 8));
    return __hiloint2double(t_hi + (k << kExpTabShift), t_lo);
}

// The three constants that do not fit an FP64 immediate (only the high word can be encoded) are
// taken from kernel parameters by the hot kernels, so that they are constant-bank operands of the
// DFMAs instead of registers re-materialised with MOVs inside the chord loop.
struct ExpConsts {
    double scale;   // 256 / ln 2
    double step;    // ln 2 / 256
    double sixth;   // 1 / 6
};
__host__ __device__ inline ExpConsts pb_exp_consts() { return ExpConsts{kExpScale, kExpStep, 1.6666666666666666e-01}; }

template <bool CLAMP, int COPIES = kExpTabCopies>
__device__ __forceinline__ double pb_exp_neg_tab(double u, uint32_t tab_addr, const ExpConsts& ec) {
    if (CLAMP) {
        const unsigned hi = min((unsigned)__double2hiint(u), 0xC085E000u);   // u >= -700 (-0.0/+tiny pass)
        u = __hiloint2double((int)hi, __double2loint(u));
    }
    double kd = fma(u, ec.scale, kExpMagic);
    const int k = __double2loint(kd);
    kd -= kExpMagic;
    const double r = fma(-kd, ec.step, u);
    const double T = pb_exp_tab_load<COPIES>(tab_addr, k);
    double p = fma(r, ec.sixth, 0.5);
    p = fma(p, r, 1.0);
    const double q = r * p;
    return fma(T, q, T);
}

template <bool CLAMP, int COPIES = kExpTabCopies>
__device__ __forceinline__ double pb_exp_neg_tab(double u, uint32_t tab_addr) {
    return pb_exp_neg_tab<CLAMP, COPIES>(u, tab_addr, pb_exp_consts());
}

// ---------------------------------------------------------------------------------------------
// Big-table form used by kernel (c') (pb200_transit.cu).  The kernel works in SCALED units
// v = u * 4096/ln2 (the scale is folded into the per-thread -sigma), so
//     exp(u) = 2^(k/4096) * exp(r ln2/4096),   k = rint(v),  r = v - k,  |r| <= 1/2,
// with 2^(k/4096) from a 4096-entry shared-memory table (32 KB, one copy: neighbouring lanes hold
// neighbouring wavelengths, so their k are equal or close and bank conflicts are rare -- ncu) and
// a QUADRATIC in r: truncation |r^3| (ln2/4096)^3 / 6 <= 1.0e-13, alternating in sign.
// The chord weight is folded into the polynomial: the record carries {w, w c1, w c2}, so
//     w exp(u) = T * (w + r (w c1 + r w c2))            2 DFMA + the accumulate DFMA.
// For one absorber k and r come straight from the product N * nsig:
//     t = fma(N, nsig, magic); kf = t - magic; r = fma(N, nsig, -kf)    (r exact to 2^-54)
// => 6 FP64 instructions per chord*lambda sample (S + 6 for S > 1, + 1 integer op when clamped).
// ---------------------------------------------------------------------------------------------
constexpr int kExpBigEntries = 4096;
constexpr int kExpBigShift = 8;                              // 20 - log2(4096)
constexpr double kExpBigScale = 5909.278887481194;           // 4096 / ln 2
constexpr double kExpBigC1 = 1.6922538587889289e-04;         // ln 2 / 4096
constexpr double kExpBigC2 = 1.43186156129301e-08;           // (ln 2 / 4096)^2 / 2
constexpr unsigned kExpBigClampHi = 0xC14F47D0u;             // high word of -4.1e6 scaled units (tau = 693.8)
constexpr double kExpBigClampDecision = 690.0 * 5909.278887481194;   // warps that can exceed it take the clamped loop

__host__ __device__ inline double pb_exp_big_entry(double two_pow_frac, int j) {
    union { double d; uint64_t u; } v;
    v.d = two_pow_frac;
    v.u -= (uint64_t)j << (kExpBigShift + 32);
    return v.d;
}

// T = 2^(k/4096) = 2^e 2^(j/4096) for k = 4096 e + j >= -4.2e6.  LOP3 + IMAD + LDS.64 + IMAD.
__device__ __forceinline__ double pb_exp_big_load(uint32_t tab_base, int k) {
    int t_lo, t_hi;
    asm volatile("{\n"
        ".reg .b32 j, a;\n"
        "and.b32 j, %2, 4095;\n"
        "mad.lo.u32 a, j, 8, %3;\n"
        "ld.shared.v2.b32 {%0, %1}, [a];\n"
        "}\n"
        : "=r"(t_lo), "=r"(t_hi)
        : "r"(k), "r"(tab_base));
    return __hiloint2double(t_hi + (k << kExpBigShift), t_lo);
}

// ---------------------------------------------------------------------------------------------
// Table-driven log2 / exp2 from GLOBAL memory (L1-resident: 4 KB + 32 KB), for the x-sums of the column kernel,
// where (R/r)^q = exp2(q log2 R - (q/2) log2 r^2) per line-of-sight sample is the cost (libdevice: sqrt + log + exp
// ~ 60 FP64 instructions; this form: 10 + 9).
//   log2: r2 = 2^e m, m in [1, 2); idx = top 8 mantissa bits; {inv, l2} = table[idx] with inv ~ 1/m_centre and
//         l2 = -log2(inv) (long double on the host); u = fma(m, inv, -1), |u| <= 2^-9; log2(1+u) by a degree-7
//         polynomial (truncation u^8/8 < 1e-22).  Absolute error ~ ulp(e + l2)/2 <= 7e-15 for r2 < 2^128 -- the
//         same as libdevice's log at that magnitude.
//   exp2: v = 4096 y, k = rint(v), r = v - k; 2^(k/4096) from the big table of kernel (c') (entries pre-biased, see
//         pb_exp_big_entry) times a CUBIC in r (truncation (ln2/4096)^4/24/16 = 2e-18).  y < -1020: exact exp2()
//         (denormal tails of a Gaussian stay what they are).
// ---------------------------------------------------------------------------------------------
constexpr int kLogTabEntries = 256;
constexpr double kExpBigC3 = 8.0768402961349e-13;            // (ln 2 / 4096)^3 / 6

// Coefficients whose low word is not zero cannot be FP64 immediates; as literals the compiler re-materialises each
// with two moves INSIDE the x loops (23 of 120 instructions of the power-law loop).  From __constant__ memory they are
// constant-bank operands of the DFMAs themselves.
static __constant__ double pb_kc[12] = {
    2.0609929155555190e-01,    // [0]  1/(7 ln2)
    -2.4044917348149410e-01,   // [1] -1/(6 ln2)
    2.8853900817779268e-01,    // [2]  1/(5 ln2)
    -3.6067376022224085e-01,   // [3] -1/(4 ln2)
    4.8089834696298783e-01,    // [4]  1/(3 ln2)
    -7.2134752044448170e-01,   // [5] -1/(2 ln2)
    1.4426950408889634e+00,    // [6]  1/ln2 = log2(e)
    8.0768402961349e-13,       // [7]  kExpBigC3
    1.43186156129301e-08,      // [8]  kExpBigC2
    1.6922538587889289e-04,    // [9]  kExpBigC1
    0.0, 0.0};

__device__ __forceinline__ double pb_log2_tab(const double2* __restrict__ ltab, double x) {
    const int hi = __double2hiint(x);
    const int e = (hi >> 20) - 1023;
    const double2 t = __ldg(ltab + ((hi >> 12) & 255));
    const double m = __hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, __double2loint(x));
    const double u = fma(m, t.x, -1.0);
    double p = fma(pb_kc[0], u, pb_kc[1]);
    p = fma(p, u, pb_kc[2]);
    p = fma(p, u, pb_kc[3]);
    p = fma(p, u, pb_kc[4]);
    p = fma(p, u, pb_kc[5]);
    p = fma(p, u, pb_kc[6]);
    return fma(u, p, (double)e + t.y);
}

__device__ __forceinline__ double pb_exp2_tab(const double* __restrict__ big, double y) {
    if (y < -1020.0) return (y < -1076.0) ? 0.0 : exp2(y);   // below 2^-1075 the result rounds to 0; denormals: exact exp2
    const double t = fma(y, 4096.0, kExpMagic);
    const int k = __double2loint(t);
    const double kf = t - kExpMagic;
    const double r = fma(y, 4096.0, -kf);
    const double T0 = __ldg(big + (k & (kExpBigEntries - 1)));
    const double T = __hiloint2double(__double2hiint(T0) + (k << kExpBigShift), __double2loint(T0));
    double p = fma(r, kExpBigC3, kExpBigC2);
    p = fma(p, r, kExpBigC1);
    return fma(T, r * p, T);
}

// Branch-free forms for the x-sums of kernels (a) / (a'): a loop body without branches, calls or select chains lets
// the compiler interleave the independent sample chains and keeps the instructions issued next to the FP64 pipe few.
//   pb_rsqrt_nb(s):   1/sqrt(s) for normal s > 0 from the MUFU seed (rel. error < 2^-20) and ONE third-order step
//                     y (1 + e/2 + 3 e^2/8), e = 1 - s y^2: error (5/16) e^3 < 2^-62 before rounding, <= 1.5 ulp after.
//   pb_exp2_core(y):  the table path of pb_exp2_tab without its range checks: exact for y >= -1020.0005; below, the
//                     result is garbage (the table index stays in range) and must be discarded by the caller.
//   pb_exp2_class(y): the range checks as INTEGER compares on the high word (for negative y the unsigned high word
//                     grows with |y|): `tiny` -- y < -1020: the caller returns 0; `rare` is raised when y lies in the band
//                     [-1076, -1020) whose true results are denormal: the caller redoes that chord with exp2().
constexpr unsigned kHiM1020 = 0xC08FE000u;   // high word of -1020.0
constexpr unsigned kHiM1076 = 0xC090D000u;   // high word of -1076.0
__device__ __forceinline__ double pb_rsqrt_nb(double s) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(s));
    const double e = fma(-(s * y), y, 1.0);
    return fma(y * e, fma(e, 0.375, 0.5), y);
}

// SH: `big` is a copy of the table in shared memory (a persistent CTA loads it once) instead of the one in global memory
template <bool SH = false>
__device__ __forceinline__ double pb_exp2_core(const double* __restrict__ big, double y) {
    const double t = fma(y, 4096.0, kExpMagic);
    const int k = __double2loint(t);
    const double kf = t - kExpMagic;
    const double r = fma(y, 4096.0, -kf);
    const double T0 = SH ? big[k & (kExpBigEntries - 1)] : __ldg(big + (k & (kExpBigEntries - 1)));
    const double T = __hiloint2double(__double2hiint(T0) + (k << kExpBigShift), __double2loint(T0));
    double p = fma(r, pb_kc[7], pb_kc[8]);
    p = fma(p, r, pb_kc[9]);
    return fma(T, r * p, T);
}

__device__ __forceinline__ bool pb_exp2_class(double y, bool& rare) {
    const unsigned hy = (unsigned)__double2hiint(y);
    rare |= (hy - (kHiM1020 + 1u)) <= (kHiM1076 - kHiM1020);   // one subtract + one compare for the band
    return hy > kHiM1020;
}

// exp(a) for |a| <= 700 (same kernel as pb_exp_neg, two-sided); used for 10**v = exp(v ln10) of the
// molecular tables, where |v| <= 50 keeps the extra rounding of v*ln10 below 1e-14 relative.
__device__ __forceinline__ double pb_exp(double a) {
    a = fmax(fmin(a, 700.0), -700.0);
    const double kMagic = 6755399441055744.0;
    double kd = fma(a, 1.4426950408889634, kMagic);
    const int k = __double2loint(kd);
    kd -= kMagic;
    double r = fma(kd, -6.93147180369123816490e-01, a);
    r = fma(kd, -1.90821492927058770002e-10, r);
    double p = 0x1.af354eec9216ep-26;
    p = fma(p, r, 0x1.28917965df995p-22);
    p = fma(p, r, 0x1.71de0fd1838abp-19);
    p = fma(p, r, 0x1.a019b9189c259p-16);
    p = fma(p, r, 0x1.a01a01a6cffcap-13);
    p = fma(p, r, 0x1.6c16c178884cep-10);
    p = fma(p, r, 0x1.1111111110c74p-7);
    p = fma(p, r, 0x1.5555555553d6fp-5);
    p = fma(p, r, 0x1.5555555555555p-3);
    p = fma(p, r, 0x1.0000000000001p-1);
    p = fma(p, r * r, r);
    p += 1.0;
    return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}

// ---------------------------------------------------------------------------------------------
// np.interp semantics on a sorted table (numba's implementation follows numpy's
// compiled_base.c): clamp to the end values outside [x0, x_{n-1}], otherwise
// slope*(x - x_j) + y_j with j = largest index with x_j <= x.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t pb_search_le(const double* __restrict__ xs, int64_t n, double x) {
    // largest j in [0, n-2] with xs[j] <= x  (x is inside (xs[0], xs[n-1]))
    int64_t lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(xs + mid) <= x) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ double pb_interp(const double* __restrict__ xs,
                                            const double* __restrict__ ys, int64_t n, double x) {
    if (!(x == x)) return x;  // NaN passes through
    if (n == 1) return __ldg(ys);
    if (x <= __ldg(xs)) return __ldg(ys);
    if (x >= __ldg(xs + n - 1)) return __ldg(ys + n - 1);
    const int64_t j = pb_search_le(xs, n, x);
    const double x0 = __ldg(xs + j), x1 = __ldg(xs + j + 1);
    const double y0 = __ldg(ys + j), y1 = __ldg(ys + j + 1);
    const double slope = (y1 - y0) / (x1 - x0);
    return slope * (x - x0) + y0;
}

// bucket of x in the bucket index of a table (DevConstituent::bucket); the index is built with this same expression
__device__ __forceinline__ int pb_bucket_of(double x, double x0, double inv, int n_bucket) {
    const double fb = (x - x0) * inv;
    return (fb > 0.0) ? ((fb < (double)n_bucket) ? (int)fb : n_bucket - 1) : 0;   // NaN -> 0
}

// The same value as pb_interp, with the binary search started from a bracket [lo, hi] that is known to contain the
// answer (bucket index of the table, DevConstituent::bucket): 3-4 dependent loads instead of ~19 on a 3e5-entry table.
__device__ __forceinline__ double pb_interp_bracket(const double* __restrict__ xs, const double* __restrict__ ys, int64_t n,
                                                    double x, int64_t lo, int64_t hi) {
    if (!(x == x)) return x;
    if (n == 1) return __ldg(ys);
    if (x <= __ldg(xs)) return __ldg(ys);
    if (x >= __ldg(xs + n - 1)) return __ldg(ys + n - 1);
    // invariant of pb_search_le: xs[lo] <= x < xs[hi]
    if (hi > n - 1) hi = n - 1;
    if (lo < 0) lo = 0;
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(xs + mid) <= x) lo = mid; else hi = mid;
    }
    const double x0 = __ldg(xs + lo), x1 = __ldg(xs + lo + 1);
    const double y0 = __ldg(ys + lo), y1 = __ldg(ys + lo + 1);
    const double slope = (y1 - y0) / (x1 - x0);
    return slope * (x - x0) + y0;
}

__device__ __forceinline__ double pb_interp_log(const double* __restrict__ xs,
                                                const double* __restrict__ ys, int64_t n, double x,
                                                double offset) {
    return exp10(pb_interp(xs, ys, n, x)) - offset;
}

// ---------------------------------------------------------------------------------------------
// Re w(x + iy), y >= 0 (Faddeeva function).  Own two-region algorithm, validated against
// scipy.special.wofz to <= 3e-14 relative over |x| <= 1e5, 1e-8 <= y <= 10 (tests/test_voigt.py):
//   |z|^2 >= 100 : asymptotic series  w ~ i/(sqrt(pi) z) * sum_k (2k-1)!!/2^k z^-2k, 2..12 terms
//   otherwise    : trapezoid rule for (i/pi) int e^{-t^2}/(z-t) dt with step h = 1/2 on nodes
//                  t_n = x + (n+1/2)h that straddle x symmetrically; every term is positive and
//                  the pole correction 2 e^{-z^2}/(1 + e^{2 pi y/h}) is free of cancellation
//                  (residual error e^{-pi^2/h^2} = 7e-18).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double pb_faddeeva_re_far(double x, double y, double r2) {
    // u = 1/z^2 = conj(z)^2 / |z|^4
    const double inv = 1.0 / r2;
    const double ur = (x * x - y * y) * inv * inv;
    const double ui = -2.0 * x * y * inv * inv;
    int nt;
    if (r2 >= 1e8) nt = 2; else if (r2 >= 1e6) nt = 3; else if (r2 >= 1e5) nt = 4;
    else if (r2 >= 1e4) nt = 5; else if (r2 >= 1e3) nt = 6; else if (r2 >= 400.0) nt = 8;
    else if (r2 >= 256.0) nt = 9; else if (r2 >= 144.0) nt = 10; else nt = 12;
    // c_k = (2k-1)!!/2^k
    const double c[12] = {1.0, 0.5, 0.75, 1.875, 6.5625, 29.53125, 162.421875, 1055.7421875,
                          7918.06640625, 67303.564453125, 639383.8623046875, 6713530.554199219};
    double sr = 0.0, si = 0.0;
#pragma unroll 1
    for (int k = nt - 1; k >= 0; --k) {
        const double tr = fma(sr, ur, -si * ui) + c[k];
        si = fma(sr, ui, si * ur);
        sr = tr;
    }
    // w = i/(sqrt(pi) z) * s ;  i/z = (y + i x)/|z|^2  ->  Re = (y*sr - x*si)/(sqrt(pi)|z|^2)
    return 0.5641895835477563 * (y * sr - x * si) * inv;
}

static __device__ __noinline__ double pb_faddeeva_re_near(double x, double y) {
    const double h = 0.5, T = 6.3;
    const double ax = fabs(x);
    // nodes t = ax + (n + 1/2) h with |t| < T
    const int n_lo = (int)ceil((-T - ax) / h - 0.5);
    const int n_hi = (int)floor((T - ax) / h - 0.5);
    const double y2 = y * y;
    double s = 0.0;
    for (int n = n_lo; n <= n_hi; ++n) {
        const double u = (n + 0.5) * h;
        const double t = ax + u;
        s += exp(-t * t) / fma(u, u, y2);
    }
    s *= h * y / kPi;
    if (y < 5.0) {
        const double g = 2.0 * exp(y2 - ax * ax) * cos(2.0 * ax * y) / (1.0 + exp(2.0 * kPi * y / h));
        s += g;
    }
    return s;
}

__device__ __forceinline__ double pb_faddeeva_re(double x, double y) {
    const double r2 = fma(x, x, y * y);
    if (r2 >= 100.0) return pb_faddeeva_re_far(x, y, r2);
    return pb_faddeeva_re_near(x, y);
}

// scipy.special.voigt_profile(x, sigma, gamma) incl. its degenerate branches
__device__ __forceinline__ double pb_voigt(double x, double sigma, double gamma) {
    const double kInvSqrt2 = 0.7071067811865476;
    const double kSqrt2Pi = 2.5066282746310002;
    if (sigma == 0.0) {
        if (gamma == 0.0) return (x == 0.0) ? INFINITY : 0.0;
        return gamma / kPi / (x * x + gamma * gamma);
    }
    if (gamma == 0.0) {
        const double q = x / sigma;
        return 1.0 / kSqrt2Pi / sigma * exp(-0.5 * q * q);
    }
    const double zr = x / sigma * kInvSqrt2;
    const double zi = gamma / sigma * kInvSqrt2;
    return pb_faddeeva_re(zr, zi) / sigma / kSqrt2Pi;
}

// warp-level fp64 sum, result in every lane
__device__ __forceinline__ double pb_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace pb200
