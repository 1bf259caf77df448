// pb200_density.cu -- kernels (a)/(a'): chord weights, blocking mask, number density and
// column density on the chord grid.  SURVEY 8(a) rows a2, a4, a6, a7, a8 (column part).
//
// Work decomposition: atoms-only scenes (the run path) -- column_cta_kernel, one CTA per 256 chords of a phase, 8 lanes
// per chord in the x-sum; scenes with molecules -- column_kernel, one warp per chord, which also stores the density
// rows kernel (m) needs.  Chord axes are a few hundred doubles and live in L1 / shared memory; the only HBM traffic is
// the coalesced store of the result.
#include <cooperative_groups.h>

#include "pb200_internal.cuh"

namespace pb200 {

// ---- a2 + a4: per sky cell weights, per (phase, cell) blocking ---------------------------------
// w_out[m] = rho*clv (gasProperties.py:1179-1180, :1232 with F* = 1); star_shift[m] (:1182-1183)
__global__ void sky_cell_kernel(const DevScene* __restrict__ scp) {
    const DevScene& sc = *scp;
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= sc.Mp) return;
    if (m >= sc.M) { sc.w_out[m] = 0.0; sc.star_shift[m] = 1.0; return; }
    const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
    const double rho = sc.rho[irho];
    // mu = sqrt(clip(1 - rho^2/R^2, 0, 1)); clv = 1 - u1 (1-mu) - u2 (1-mu)^2, no FMA contraction
    double q = __dsub_rn(1.0, __ddiv_rn(__dmul_rn(rho, rho), __dmul_rn(sc.star_R, sc.star_R)));
    q = fmin(fmax(q, 0.0), 1.0);
    const double om = __dsub_rn(1.0, __dsqrt_rn(q));
    const double clv = __dsub_rn(__dsub_rn(1.0, __dmul_rn(sc.clv_u1, om)),
                                 __dmul_rn(sc.clv_u2, __dmul_rn(om, om)));
    sc.w_out[m] = __dmul_rn(rho, clv);
    const double v = __dmul_rn(__ddiv_rn(__dmul_rn(sc.vsini, rho), sc.star_R), sc.cos_phi_rot[iphi]);
    const double beta = __ddiv_rn(v, kLightSpeed);
    sc.star_shift[m] = __dsqrt_rn(__ddiv_rn(__dsub_rn(1.0, beta), __dadd_rn(1.0, beta)));
}

// blocked(p, m): planet disk sqrt((y-y_p)^2 + z^2) < R_p, moon disks (y-y_m)^2 + z^2 < R_m^2
// (gasProperties.py:1224-1230).  Same operation order as numpy, round-to-nearest, no FMA.
__device__ __forceinline__ bool pb_blocked(const DevScene& sc, int p, double y, double z) {
    const double z2 = __dmul_rn(z, z);
    const double dyp = __dsub_rn(y, sc.planet_y[p]);
    bool blocked = __dsqrt_rn(__dadd_rn(__dmul_rn(dyp, dyp), z2)) < (sc.planet_R_phase ? sc.planet_R_phase[p] : sc.planet_R);
    for (int k = 0; k < sc.n_moons; ++k) {
        const double dym = __dsub_rn(y, sc.moon_y[k * sc.n_orb + p]);
        const double Rm = sc.moon_R_per_phase ? sc.moon_R[k * sc.n_orb + p] : sc.moon_R[k];
        blocked |= __dadd_rn(__dmul_rn(dym, dym), z2) < __dmul_rn(Rm, Rm);
    }
    return blocked;
}

// ---- a6 + a7 + a8: density and column density, one warp per chord -----------------------------
// WRITE_DENSITY: also store n(c, x) of scenarios that carry molecules (needed by kernel (m)).
__global__ void __launch_bounds__(256)
column_kernel(const DevScene* __restrict__ scp, int phase_lo, int n_phase, int* __restrict__ oob_flag) {
    const DevScene& sc = *scp;
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    const int64_t n_chords = (int64_t)n_phase * sc.Mp;
    for (int64_t cid = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); cid < n_chords;
         cid += (int64_t)gridDim.x * warps_per_block) {
        const int p = phase_lo + (int)(cid / sc.Mp);
        const int m = (int)(cid % sc.Mp);
        const int64_t row = (int64_t)p * sc.Mp + m;
        if (m >= sc.M) {   // padding of the phase-major rows: zero weight, zero column
            if (lane == 0) {
                sc.w_in[row] = 0.0;
                for (int a = 0; a < sc.n_atoms; ++a) sc.ncol[(int64_t)a * sc.n_orb * sc.Mp + row] = 0.0;
            }
            continue;
        }
        const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
        const double rho = sc.rho[irho];
        const double y = __dmul_rn(rho, sc.sin_phi[iphi]);
        const double z = __dmul_rn(rho, sc.cos_phi[iphi]);
        const bool blocked = pb_blocked(sc, p, y, z);
        if (lane == 0) {
            const double wv = blocked ? 0.0 : sc.w_out[m];
            sc.w_in[row] = wv;
        }
        for (int s = 0; s < sc.n_scen; ++s) {
            const DevScenario& S = sc.scen[s];
            double sum = 0.0;
            // the reference evaluates only unblocked chords (gasProperties.py:1235-1240)
            if (!blocked) {
                const double amp = S.amplitude ? S.amplitude[p] : 1.0;
                double* dens = (S.density_slot >= 0)
                                   ? sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x
                                   : nullptr;
                if (S.type == PB200_SCN_SERPENS) {
                    int oob = 0;
                    for (int ix = lane; ix < sc.n_x; ix += 32) {
                        const double n = pb_trilinear(S, sc.x[ix], y, z, &oob) * amp;
                        if (dens) dens[ix] = n;
                        sum += n;
                    }
                    if (oob) *oob_flag = 1;
                } else {
                    const double dy = y - S.centre_y[p];
                    const double cx = S.centre_x[p];
                    const double* sp = pb_params(S, p);
                    for (int ix = lane; ix < sc.n_x; ix += 32) {
                        const double n = pb_density(S, sp, amp, sc.x[ix] - cx, dy, z);
                        if (dens) dens[ix] = n;
                        sum += n;
                    }
                }
                sum = pb_warp_sum(sum);
            } else if (S.density_slot >= 0) {
                double* dens = sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x;
                for (int ix = lane; ix < sc.n_x; ix += 32) dens[ix] = 0.0;
            }
            if (lane == 0) {
                for (int k = S.first_cons; k < S.first_cons + S.n_cons; ++k) {
                    const DevConstituent& C = sc.cons[k];
                    if (C.kind == 0) {   // col = sum_x(n*chi) * delta_x, gasProperties.py:940
                        const double col = (sum * C.chi) * sc.delta_x;
                        sc.ncol[(int64_t)C.atom_index * sc.n_orb * sc.Mp + row] = col;
                        // densities are >= 0, so the bit patterns order like the values
                        if (col > 0.0) atomicMax(sc.nmax + (int64_t)C.atom_index * sc.n_orb + p, (unsigned long long)__double_as_longlong(col));
                    }
                }
            }
        }
    }
}

// ---- (a') fast form for scenes without molecules ------------------------------------------------
// The analytic scenarios along one chord.  What does not depend on x is hoisted twice:
//   * per (scenario, phase) -- ScnConst: radii, scale heights, exponents, q log2 R, the density centre; uniform over
//     a CTA whose chords share the phase;
//   * per chord -- (pre, cc): n(x) = pre * term(dx^2, cc) with cc = dy^2 + z^2 (spherical scenarios) or dy^2 (torus,
//     whose exp(-(z/H)^2) goes into pre).
// term() is branch-free (pb_rsqrt_nb, pb_exp2_tab_nb, pb_log2_tab: no sqrt / division / libm slow paths), so the
// independent samples of a lane interleave; the rare denormal band of the exp is redone exactly by pb_term_exact.
// Accuracy against the libm forms of pb_density: <= 3e-14 relative (Ncol is tested to 1e-12).
struct ScnConst {
    int type;
    double n0;        // P[0] * amplitude(p)
    double a, b, c;   // barometric {R, log2e/H, -}; hydrostatic {R, G mu M/(kB T), jc/R}; power law {R^2, -q/2, q log2 R};
                      // torus {a_t, 1/(4 H), 1/H}
    double cx, cy;    // density centre at this phase
};

__device__ __forceinline__ ScnConst pb_scn_const(const DevScenario& S, int p) {
    ScnConst k;
    const double* __restrict__ P = pb_params(S, p);
    k.type = S.type;
    k.n0 = P[0] * (S.amplitude ? S.amplitude[p] : 1.0);
    k.cx = S.centre_x ? S.centre_x[p] : 0.0;
    k.cy = S.centre_y ? S.centre_y[p] : 0.0;
    k.a = k.b = k.c = 0.0;
    const double kLog2e = 1.4426950408889634;
    switch (S.type) {
        case PB200_SCN_BAROMETRIC: k.a = P[1]; k.b = kLog2e / P[2]; break;                                  // gasProperties.py:155-160
        case PB200_SCN_HYDROSTATIC: k.a = P[1]; k.b = P[2] / P[3]; k.c = k.b / P[1]; break;                   // :197-203
        case PB200_SCN_POWERLAW: k.a = P[1] * P[1]; k.b = -0.5 * P[2]; k.c = P[2] * log2(P[1]); break;       // :240-243, :326-329, :371-373, :453-460
        case PB200_SCN_TORUS: k.a = P[1]; k.b = 1.0 / (4.0 * P[2]); k.c = 1.0 / P[2]; break;                 // :506-515
        default: break;
    }
    return k;
}

// per chord: n(x) = pre * term(dx^2, cc)
// exp2(y) for any y <= 0: the table path, 0 below 2^-1076, libm in the denormal band (per-chord use only)
__device__ __forceinline__ double pb_exp2_any(const double* __restrict__ big, double y) {
    bool rare = false;
    const bool tiny = pb_exp2_class(y, rare);
    if (rare) return exp2(y);
    return tiny ? 0.0 : pb_exp2_core(big, y);
}

__device__ __forceinline__ void pb_chord_pre(const ScnConst& k, const double* __restrict__ big, double y, double z, double& pre, double& cc) {
    const double dy = y - k.cy;
    if (k.type == PB200_SCN_TORUS) {
        const double qz = z * k.c;
        pre = k.n0 * pb_exp2_any(big, -(qz * qz) * pb_kc[6]);
        cc = dy * dy;
    } else {
        pre = k.n0;
        cc = fma(dy, dy, z * z);
    }
}

// n(x)/pre at one sample.  `rare`: see pb_exp2_class.  No branch, no call; the only selects are the final one and the
// Jeans term of the hydrostatic law.  The heaviside of the reference (H(r - R), = 1 at r = R) shares the predicate
// of the exp's underflow.
template <int TYPE, bool SH = false>
__device__ __forceinline__ double pb_term(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                          double dx2, double cc, bool& rare) {
    const double kLog2e = pb_kc[6];
    const double s = dx2 + cc;                                  // r^2 (torus: a^2); > 0 on a midpoint grid
    double yv;
    bool dead = false;
    if constexpr (TYPE == PB200_SCN_POWERLAW) {
        // (R/r)^q = exp2(q log2 R - (q/2) log2 r^2); r < R <=> r^2 < R^2 up to a rounding of the boundary sample
        dead = s < k.a;
        yv = fma(k.b, pb_log2_tab(ltab, s), k.c);
    } else {
        const double ri = pb_rsqrt_nb(s), r = s * ri;
        if constexpr (TYPE == PB200_SCN_BAROMETRIC) {
            dead = r < k.a;
            yv = (k.a - r) * k.b;
        } else if constexpr (TYPE == PB200_SCN_HYDROSTATIC) {
            const double J = (r < k.a) ? 0.0 : k.b * ri;        // non-zero inside the planet, as the reference
            yv = (J - k.c) * kLog2e;
        } else {                                                // torus
            const double qa = (r - k.a) * k.b;
            yv = -(qa * qa) * kLog2e;
        }
    }
    const double v = pb_exp2_core<SH>(big, yv);
    const bool zero = pb_exp2_class(yv, rare) | dead;   // one predicate, one select
    return zero ? 0.0 : v;
}

// the same value through libm (exp2 keeps the denormal band exact): for the chords flagged `rare`
template <int TYPE>
__device__ __forceinline__ double pb_term_exact(const ScnConst& k, const double2* __restrict__ ltab, double dx2, double cc) {
    const double kLog2e = 1.4426950408889634;
    const double s = dx2 + cc;
    if (TYPE == PB200_SCN_POWERLAW) return (s < k.a) ? 0.0 : exp2(fma(k.b, pb_log2_tab(ltab, s), k.c));
    const double ri = pb_rsqrt_nb(s), r = s * ri;
    if (TYPE == PB200_SCN_BAROMETRIC) return (r < k.a) ? 0.0 : exp2((k.a - r) * k.b);
    if (TYPE == PB200_SCN_HYDROSTATIC) return exp2((((r < k.a) ? 0.0 : k.b * ri) - k.c) * kLog2e);
    const double qa = (r - k.a) * k.b;
    return exp2(-(qa * qa) * kLog2e);
}

// x-sum of one chord over the samples ix = l8, l8 + 8, ... of one of the 8 lanes that share the chord: two samples per
// iteration, two accumulators.  XS: (x - cx)^2 comes from shared memory (s_dx2), else it is formed from the x axis.
// *rare_out is raised when a sample fell into the denormal band of the exp: the caller then takes pb_lane_sum_exact.
constexpr int kColLanes = 8;
template <int TYPE, bool XS>
__device__ __forceinline__ double pb_lane_sum(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                              const double* __restrict__ s_dx2, const double* __restrict__ xs, int n_x, int l8,
                                              double cc, bool& rare) {
    double a0 = 0.0, a1 = 0.0;
    int ix = l8;
    for (; ix + kColLanes < n_x; ix += 2 * kColLanes) {
        double d0, d1;
        if constexpr (XS) { d0 = s_dx2[ix]; d1 = s_dx2[ix + kColLanes]; }
        else { const double u = xs[ix] - k.cx, w = xs[ix + kColLanes] - k.cx; d0 = u * u; d1 = w * w; }
        a0 += pb_term<TYPE>(k, big, ltab, d0, cc, rare);
        a1 += pb_term<TYPE>(k, big, ltab, d1, cc, rare);
    }
    if (ix < n_x) {
        double d0;
        if constexpr (XS) d0 = s_dx2[ix]; else { const double u = xs[ix] - k.cx; d0 = u * u; }
        a0 += pb_term<TYPE>(k, big, ltab, d0, cc, rare);
    }
    return a0 + a1;
}

template <int TYPE>
static __device__ __noinline__ double pb_lane_sum_exact(const ScnConst k, const double2* __restrict__ ltab,
                                                        const double* __restrict__ xs, int n_x, int l8, double cc) {
    double a0 = 0.0, a1 = 0.0;   // same samples, same order of additions as pb_lane_sum
    int ix = l8;
    for (; ix + kColLanes < n_x; ix += 2 * kColLanes) {
        const double u = xs[ix] - k.cx, w = xs[ix + kColLanes] - k.cx;
        a0 += pb_term_exact<TYPE>(k, ltab, u * u, cc);
        a1 += pb_term_exact<TYPE>(k, ltab, w * w, cc);
    }
    if (ix < n_x) { const double u = xs[ix] - k.cx; a0 += pb_term_exact<TYPE>(k, ltab, u * u, cc); }
    return a0 + a1;
}

template <int TYPE, bool XS>
__device__ __forceinline__ double pb_lane_sum_checked(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                                      const double* __restrict__ s_dx2, const double* __restrict__ xs, int n_x,
                                                      int l8, double cc) {
    bool rare = false;
    const double sum = pb_lane_sum<TYPE, XS>(k, big, ltab, s_dx2, xs, n_x, l8, cc, rare);
    // a dropped term is < 2^-1020: it changes the sum (relative 2^-53) only if the sum itself is below 2^-960
    return (rare && sum < 1.0261342003245941e-289) ? pb_lane_sum_exact<TYPE>(k, ltab, xs, n_x, l8, cc) : sum;
}

template <bool XS>
__device__ __forceinline__ double pb_lane_sum_any(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                                  const double* __restrict__ s_dx2, const double* __restrict__ xs, int n_x, int l8,
                                                  double cc) {
    switch (k.type) {   // uniform over the CTA
        case PB200_SCN_BAROMETRIC: return pb_lane_sum_checked<PB200_SCN_BAROMETRIC, XS>(k, big, ltab, s_dx2, xs, n_x, l8, cc);
        case PB200_SCN_HYDROSTATIC: return pb_lane_sum_checked<PB200_SCN_HYDROSTATIC, XS>(k, big, ltab, s_dx2, xs, n_x, l8, cc);
        case PB200_SCN_POWERLAW: return pb_lane_sum_checked<PB200_SCN_POWERLAW, XS>(k, big, ltab, s_dx2, xs, n_x, l8, cc);
        case PB200_SCN_TORUS: return pb_lane_sum_checked<PB200_SCN_TORUS, XS>(k, big, ltab, s_dx2, xs, n_x, l8, cc);
        default: return 0.0;
    }
}

// Upper bound of the x-sum of one scenario on a chord (every term <= its value at dx = 0, the closest approach
// r_min^2 = dy^2 + z^2), from the per-phase constants with the table math of the sample loop (a 1e-9 margin covers its
// rounding); +inf where no cheap bound exists.  Used by the run path only, to skip the sum on chords that
// compact_kernel will classify as transparent anyway; the bound is what such a chord gets stored as its Ncol.
__device__ __forceinline__ double pb_column_bound(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                                  int n_x, double y, double z) {
    const double kInf = __longlong_as_double(0x7ff0000000000000ll);
    const double n0 = fabs(k.n0) * (double)n_x * (1.0 + 1e-9);
    const double dy = y - k.cy;
    const double s = fma(dy, dy, z * z);
    double yv;
    switch (k.type) {
        case PB200_SCN_BAROMETRIC: { const double r = s * pb_rsqrt_nb(s); if (r < k.a) return n0; yv = (k.a - r) * k.b; break; }
        case PB200_SCN_HYDROSTATIC: { const double ri = pb_rsqrt_nb(s); if (s * ri < k.a) return n0; yv = (k.b * ri - k.c) * pb_kc[6]; break; }
        case PB200_SCN_POWERLAW:
            // (R/r)^q falls with r only for q > 0; for q <= 0 (the reference accepts any q) the closest approach is not
            // the maximum: no cheap bound
            if (!(k.b < 0.0)) return kInf;
            if (s < k.a) return n0;
            yv = fma(k.b, pb_log2_tab(ltab, s), k.c);
            break;
        case PB200_SCN_TORUS: { const double qz = z * k.c; yv = -(qz * qz) * pb_kc[6]; break; }
        default: return kInf;   // tabulated density
    }
    if (!(yv == yv)) return kInf;
    bool rare = false;
    const bool tiny = pb_exp2_class(yv, rare);
    return n0 * (tiny ? 1.1125369292536007e-307 : pb_exp2_core(big, yv));   // 2^-1020 bounds everything below
}

// One CTA = 256 consecutive chords of ONE phase (rows are padded to 256 chords).
//   Phase A, one thread per chord: y, z, blocking mask and weight (bit-exact operation order), then -- run path -- the
//     decisions which sums compact_kernel will not need (see run_mode below); the chords that need an exact sum are
//     compacted into a work list, so that no lane idles in phase B for a blocked, mirrored or culled neighbour.
//   Phase B, per scenario: (x - cx(p))^2 of the phase goes to shared memory once; 8 lanes share a listed chord and
//     stride x (pb_lane_sum), their partial sums are added in a fixed order (xor tree) -- deterministic.
//   Then one thread per chord again writes Ncol (coalesced) and feeds the running maxima (one atomicMax per warp).
// run_mode = 0: every unblocked chord gets its exact column densities, one per atomic constituent (inner seams).
// run_mode = 2 (pb200_run with PB200_OPT_CULL = 0): exact sums for every unblocked chord, absorbers as in run_mode 1.
// run_mode = 1 (pb200_run): the sums that compact_kernel does not need are skipped --
//   * the mirror chord of a pair whose primary chord is not blocked (sc.mirror);
//   * chords whose upper bound sum_a bound_a * max_w sigma_a(p, w) is already below 2^-54: they are
//     transparent whatever the exact sum is, and get the bound stored as Ncol (so that compact_kernel's own
//     test classifies them the same way).
// Running maximum of the column densities of a phase (DevScene::nmax: clamp decision, thin test and window order of
// kernel (c') -- an UPPER bound is what they need).  Columns are >= 0, so their bit patterns order like the values: one
// warp-wide integer max of the high words (redux.sync), rounded up by one unit of the high word (2^-20 relative), one
// atomic per warp.
__device__ __forceinline__ void pb_feed_max(unsigned long long* __restrict__ dst, double col, int lane) {
    const unsigned hi = (col > 0.0) ? (unsigned)__double2hiint(col) + 1u : 0u;   // NaN compares false: not fed (as before)
    const unsigned wmax = __reduce_max_sync(0xffffffffu, hi);
    if (lane == 0 && wmax) atomicMax(dst, (unsigned long long)wmax << 32);
}

constexpr int kColCta = 256;
constexpr int kColMaxXs = 2048;   // x samples whose (x - cx)^2 are kept in shared memory (16 KB); beyond: formed per sample
__global__ void __launch_bounds__(kColCta, 4)
column_cta_kernel(const DevScene* __restrict__ scp, int phase_lo, int n_phase, int* __restrict__ oob_flag, int run_mode) {
    const DevScene& sc = *scp;
    __shared__ double s_y[kColCta], s_z[kColCta], s_val[kColCta], s_pre[kColCta], s_cc[kColCta], s_dx2[kColMaxXs];
    __shared__ unsigned char s_list[kColCta];
    __shared__ int s_cnt[kColCta / 32];
    __shared__ ScnConst s_k[kMaxScenarios];
    __shared__ double s_sigmax[PB200_MAX_ABSORBERS];
    __shared__ int s_abs_scn[PB200_MAX_ABSORBERS];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int p = phase_lo + blockIdx.y;            // grid: (Mp / 256 tiles of a phase, phases)
    const int m = blockIdx.x * kColCta + tid;
    const int64_t row = (int64_t)p * sc.Mp + m;
    const int n_x = sc.n_x;
    const double* __restrict__ xs = sc.x;
    const double* __restrict__ big = sc.exp_tab + kExpTabEntries;
    const double2* __restrict__ ltab = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries + kExpBigEntries);

    // ---- phase A -------------------------------------------------------------------------------------
    bool blocked = true, need_sum = false;
    double y = 0.0, z = 0.0, wv = 0.0;
    if (m < sc.M) {
        const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
        const double rho = sc.rho[irho];
        y = __dmul_rn(rho, sc.sin_phi[iphi]);
        z = __dmul_rn(rho, sc.cos_phi[iphi]);
        blocked = pb_blocked(sc, p, y, z);
        wv = blocked ? 0.0 : sc.w_out[m];
        need_sum = !blocked;      // the reference evaluates only unblocked chords (gasProperties.py:1235-1240)
        if (run_mode == 1 && sc.mirror && need_sum) {
            const int n2 = sc.n_phi / 2, h = iphi / n2, i = iphi - h * n2, ip = n2 - 1 - i;
            if (i > ip) {   // mirror chord of the pair: needed only if the primary chord is blocked
                const int jphi = h * n2 + ip;
                need_sum = pb_blocked(sc, p, __dmul_rn(rho, sc.sin_phi[jphi]), __dmul_rn(rho, sc.cos_phi[jphi]));
            }
        }
    }
    sc.w_in[row] = wv;
    // run statistics (also reset by sigma_phase_kernel, which the sigma-cell route does not launch): before compaction adds to them
    if (run_mode && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) { sc.stat[1] = 0ull; sc.stat[2] = 0ull; sc.stat[4] = 0ull; }
    if (tid < sc.n_scen) s_k[tid] = pb_scn_const(sc.scen[tid], p);   // log2(R), 1/H, ...: once per CTA, not once per thread
    if (run_mode && tid >= 64 && tid < 64 + sc.n_abs) s_abs_scn[tid - 64] = sc.cons[sc.abs_first[tid - 64]].scenario;
    if (run_mode == 1 && tid >= 32 && tid < 32 + sc.n_abs) {
        // upper bound of the absorber's cross section at ANY wavelength: sum_k chi_k (10^max(tab_y_k) - offset_k) (the
        // table lookup is a convex combination of table values).  Independent of sigma_phase_kernel, which therefore
        // runs beside this kernel; compact_kernel's own test uses the run's exact maxima, which are not larger.
        s_sigmax[tid - 32] = pb_abs_sigma_bound(sc, tid - 32);   // negative or NaN: +inf, never cull
    }
    __syncthreads();
    bool skip = false;
    if (need_sum && run_mode == 1) {   // transparent whatever the exact sums are?
        double bound = 0.0;
        for (int g = 0; g < sc.n_abs; ++g) {
            const double ub = pb_column_bound(s_k[s_abs_scn[g]], big, ltab, n_x, y, z);
            bound += ub * sc.delta_x * s_sigmax[g];
        }
        skip = bound < 5.5511151231257827e-17;   // 2^-54; NaN (0 * inf) -> not skipped
    }
    s_y[tid] = y;
    s_z[tid] = z;
    const bool listed = need_sum && !skip;
    const unsigned bal = __ballot_sync(0xffffffffu, listed);
    if (lane == 0) s_cnt[wid] = __popc(bal);
    __syncthreads();
    int base = 0, n_list = 0;
#pragma unroll
    for (int w = 0; w < kColCta / 32; ++w) { base += (w < wid) ? s_cnt[w] : 0; n_list += s_cnt[w]; }
    if (listed) s_list[base + __popc(bal & ((1u << lane) - 1u))] = (unsigned char)tid;

    // ---- per scenario: phase B + the writes ---------------------------------------------------------------
    const int grp = tid / kColLanes, l8 = tid % kColLanes;
    for (int s = 0; s < sc.n_scen; ++s) {
        const DevScenario& S = sc.scen[s];
        int g_lo = 0, g_hi = -1;     // run path: the absorbers that share this scenario's sum (consecutive)
        if (run_mode) {
            g_lo = sc.n_abs;
            for (int g = 0; g < sc.n_abs; ++g)
                if (s_abs_scn[g] == s) { g_lo = g < g_lo ? g : g_lo; g_hi = g; }
            if (g_hi < 0) continue;   // uniform
        }
        const bool serp = S.type == PB200_SCN_SERPENS;
        const double amp = S.amplitude ? S.amplitude[p] : 1.0;
        const bool xs_smem = !serp && n_x <= kColMaxXs;
        __syncthreads();   // work list complete; the previous scenario's s_val / s_dx2 are no longer read
        const ScnConst k = s_k[s];
        if (xs_smem)
            for (int ix = tid; ix < n_x; ix += kColCta) { const double dx = xs[ix] - k.cx; s_dx2[ix] = dx * dx; }
        s_val[tid] = (need_sum && skip) ? pb_column_bound(k, big, ltab, n_x, y, z) : 0.0;
        if (listed && !serp) {
            double pre, cc;
            pb_chord_pre(k, big, y, z, pre, cc);
            // the sample loop decides underflow on bit patterns: a NaN (NaN scene parameters) would read as "tiny".
            // It is caught here, once per chord: pre = NaN makes the sum NaN, as the reference's arithmetic would
            if (!(cc == cc) || !((k.a + k.b) + (k.c + k.cx) == (k.a + k.b) + (k.c + k.cx))) pre = __longlong_as_double(0x7ff8000000000000ll);
            s_pre[tid] = pre;
            s_cc[tid] = cc;
        }
        __syncthreads();
        for (int i0 = 0; i0 < n_list; i0 += kColCta / kColLanes) {   // every lane of a warp makes the same number of rounds
            const int i = i0 + grp;
            const bool act = i < n_list;
            const int t = act ? s_list[i] : 0;
            double sum = 0.0;
            if (act) {
                if (serp) {
                    int oob = 0;
                    for (int ix = l8; ix < n_x; ix += kColLanes) sum += pb_trilinear(S, xs[ix], s_y[t], s_z[t], &oob);
                    if (oob) *oob_flag = 1;
                    sum *= amp;
                } else {
                    const double pre = s_pre[t];   // every term is finite and <= 1: pre == 0 gives an exact 0 (torus far off the plane)
                    if (pre != 0.0)
                        sum = pre * (xs_smem ? pb_lane_sum_any<true>(k, big, ltab, s_dx2, xs, n_x, l8, s_cc[t])
                                             : pb_lane_sum_any<false>(k, big, ltab, s_dx2, xs, n_x, l8, s_cc[t]));
                }
            }
            sum += __shfl_xor_sync(0xffffffffu, sum, 4);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            if (act && l8 == 0) s_val[t] = sum;
        }
        __syncthreads();
        const double val = s_val[tid];
        if (run_mode) {
            // one absorber per group of atomic constituents that share a scenario (DevScene::n_abs):
            // Ncol_g = sum_x n dx, the mixing ratios are in the group's cross section
            for (int g = g_lo; g <= g_hi; ++g) {
                if (s_abs_scn[g] != s) continue;
                const double col = val * sc.delta_x;
                sc.ncol[(int64_t)g * sc.n_orb * sc.Mp + row] = col;
                pb_feed_max(sc.nmax + (int64_t)g * sc.n_orb + p, col, lane);
            }
        } else {
            for (int c = S.first_cons; c < S.first_cons + S.n_cons; ++c) {
                const DevConstituent& C = sc.cons[c];
                if (C.kind != 0) continue;
                const double col = (val * C.chi) * sc.delta_x;                      // gasProperties.py:940
                sc.ncol[(int64_t)C.atom_index * sc.n_orb * sc.Mp + row] = col;
                pb_feed_max(sc.nmax + (int64_t)C.atom_index * sc.n_orb + p, col, lane);
            }
        }
    }
}

// ---- chord records of kernel (c'), transparent chords culled ------------------------------------
// One CTA per phase walks the phase's chords in order.  A chord is
//   * dead        if w_in == 0 (blocked by a body, or row padding): it contributes nothing;
//   * transparent if sum_a Ncol_a * max_w sigma_a(p, w) < 2^-54: at every wavelength of the run
//                 exp(-tau) rounds to exactly 1.0 (as it does in the reference), so the chord adds exactly
//                 w_in to F_in -- these weights are summed once into w0[p];
//   * active      otherwise: it gets a record {w, w c1, w c2, Ncol_0..} (the quadratic of the table exp
//                 folded into the weight, pb200_math.cuh "Big-table form": w exp(r c1) ~= w + r (w c1 + r w c2)).
// Active records are packed in chord order at the front of the phase's rec row (block-wide exclusive
// scan, deterministic), the last tile is padded with all-zero records, n_act[p] = number of tiles.
// CTA size: 1024 threads for large sky grids; 256 for small ones (a dense phase sweep like C3 has 401 phases of 1 200
// chord pairs: four small CTAs per SM run concurrently where one large one spent its time in barriers).
template <int kCompactThreads>
__global__ void __launch_bounds__(kCompactThreads) compact_kernel(const DevScene* __restrict__ scp, const TransitParams prm, int cull,
                                                                  int n_pow2) {
    const DevScene& sc = *scp;
    const int phase_lo = prm.phase_lo;
    const int p = phase_lo + blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int S = sc.n_abs, R = pb_rec_stride(S);
    __shared__ double s_smax[kMaxConstituents];
    __shared__ int s_warp[32];
    __shared__ double s_part[32];
    __shared__ int s_base;
    const bool cells = prm.cell_range != nullptr;   // sigma cells: sigma_phase_kernel runs BESIDE this kernel, the bound comes from the tables
    if (tid < S) {
        const double v = cells ? pb_abs_sigma_bound(sc, tid) : __longlong_as_double((long long)sc.smax[(int64_t)tid * sc.n_orb + p]);
        s_smax[tid] = (cull && v >= 0.0) ? v : __longlong_as_double(0x7ff0000000000000ll);   // negative or NaN: never cull
    }
    if (tid == 0) s_base = 0;
    __syncthreads();
    const int64_t row0 = (int64_t)p * sc.Mp;
    double* rec_row = sc.rec + row0 * R;
    double w0 = 0.0;
    double tt[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};   // this thread's share of thin_tot
    // mirror pairing (sc.mirror, set by pb200_set_scene when it holds): every body and density centre lies in
    // the plane z = 0 and every density depends on z through z^2 only, so the chords (phi, rho) and
    // (pi - phi, rho) have the same column densities; with an even, mirror-symmetric phi axis they are grid
    // chords i and n_phi/2 - 1 - i of the same half turn.  One record then stands for both: weight =
    // w_in(i) + w_in(i') (each is zero if that chord is blocked), Ncol from whichever is not blocked.
    const int n2 = sc.n_phi / 2, P = (n2 + 1) / 2;
    const int n_items = sc.mirror ? 2 * P * sc.n_rho : sc.Mp;
    for (int m0 = 0; m0 < n_items; m0 += kCompactThreads) {
        const int q = m0 + tid;
        double w = 0.0, bound = 0.0;
        int m = q;
        if (q < n_items) {
            if (sc.mirror) {
                const int h = q / (P * sc.n_rho), rem = q - h * (P * sc.n_rho);
                const int i = rem / sc.n_rho, irho = rem - i * sc.n_rho;
                m = (h * n2 + i) * sc.n_rho + irho;
                const int ip = n2 - 1 - i;
                const double wa = sc.w_in[row0 + m];
                if (ip != i) {
                    const int mb = (h * n2 + ip) * sc.n_rho + irho;
                    const double wb = sc.w_in[row0 + mb];
                    if (wa == 0.0) m = mb;               // Ncol of a blocked chord is not evaluated: take the mirror's
                    w = wa + wb;
                } else {
                    w = wa;
                }
            } else {
                w = sc.w_in[row0 + m];
            }
            for (int a = 0; a < S; ++a) bound = fma(sc.ncol[(int64_t)a * sc.n_orb * sc.Mp + row0 + m], s_smax[a], bound);
        }
        const bool alive = (w != 0.0);
        const bool active = alive && !(bound < 5.5511151231257827e-17);          // 2^-54; NaN -> active
        if (alive && !active) w0 += w;
        const unsigned bal = __ballot_sync(0xffffffffu, active);
        if (lane == 0) s_warp[wid] = __popc(bal);
        __syncthreads();
        if (wid == 0) {   // exclusive scan of the 32 warp counts
            const int c = (lane < kCompactThreads / 32) ? s_warp[lane] : 0;
            int x = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, x, d);
                if (lane >= d) x += y;
            }
            s_warp[lane] = x - c;
            if (lane == 31) s_part[0] = (double)x;   // chunk total (exact: <= 1024)
        }
        __syncthreads();
        const int base = s_base;
        if (active) {
            double* r = rec_row + (int64_t)(base + s_warp[wid] + __popc(bal & ((1u << lane) - 1u))) * R;
            const double w1 = w * kExpBigC1, w2 = w * kExpBigC2;
            r[0] = w; r[1] = w1; r[2] = w2;
            for (int a = 0; a < S; ++a) r[3 + a] = sc.ncol[(int64_t)a * sc.n_orb * sc.Mp + row0 + m];
            if (3 + S < R) r[3 + S] = 0.0;
            if (S <= 2) {   // same operations as ThinSums::add of kernel (c')
                const double N0 = r[3], N1 = (S == 2) ? r[4] : 0.0;
                tt[0] += w;
                tt[1] = fma(N0, w1, tt[1]);
                tt[2] = fma(N1, w1, tt[2]);
                const double t0 = N0 * w2, t1 = N1 * w2;
                tt[3] = fma(t0, N0, tt[3]);
                tt[4] = fma(t0, N1, tt[4]);
                tt[5] = fma(t1, N1, tt[5]);
            }
        }
        __syncthreads();
        if (tid == 0) s_base = base + (int)s_part[0];
        __syncthreads();
    }
    const int n_rec = s_base;
    const int n_tiles = (n_rec + kChordTile - 1) / kChordTile;
    for (int i = n_rec * R + tid; i < n_tiles * kChordTile * R; i += kCompactThreads) rec_row[i] = 0.0;   // padding
    // sum of the transparent weights and the thin totals, fixed order: warp sums -> s_red[k][warp], then warp k
    // adds row k (one barrier pair for all seven sums)
    __shared__ double s_red[7][32];
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const double v = pb_warp_sum(k == 0 ? w0 : tt[k == 0 ? 0 : k - 1]);
        if (lane == 0) s_red[k][wid] = v;
    }
    __syncthreads();
    if (wid < 7) {
        const double t = pb_warp_sum((lane < kCompactThreads / 32) ? s_red[wid][lane] : 0.0);
        if (lane == 0) {
            if (wid == 0) sc.w0[p] = t; else sc.thin_tot[(int64_t)p * 8 + wid - 1] = t;
        }
    }
    if (tid == 0) {
        sc.n_act[p] = n_tiles;
        atomicAdd(sc.stat + 1, (unsigned long long)n_tiles * kChordTile);
    }
    if (tid < S && !cells) sc.smax[(int64_t)tid * sc.n_orb + p] = 0ull;   // clean for the atomicMax of the next run

    // ---- order of the wavelength windows of this phase for kernel (c') -------------------------------
    // A CTA of kernel (c') owns one window; windows in a line core take the table exp on most chords, windows
    // in the continuum none.  They are launched in decreasing order of optical thickness,
    // key = sum_a max_m Ncol_a(p, m) * max_{w in window} |nsig_a(p, w)| (the largest scaled tau of the window),
    // so that the long CTAs start first (bitonic network in shared memory; the order only matters for the
    // load balance, never for the results).  More than kMaxWindows windows: identity order.
    // ---- sigma cells (one absorber): the cells this phase needs, from "thin for every chord" (N_max a < thin_v: those
    // wavelengths take the order-2 totals above) up to the phase's largest |nsig| -------------------------------------------
    if (prm.cell_range != nullptr) {
        if (tid < kCellCap / 32) prm.cell_count[blockIdx.x * (kCellCap / 32) + tid] = 0u;
        if (tid == 0) {
            const double mx = pb_abs_sigma_bound(sc, 0) * kExpBigScale;   // largest |nsig| any wavelength can have
            const double nmax = __longlong_as_double((long long)sc.nmax[p]);
            int2 r = make_int2(0, 0);
            const double a_thin = prm.thin_v / nmax;
            if (nmax > 0.0 && n_tiles > 0 && mx >= a_thin && a_thin > 0.0) {
                r.x = pb_cell_index(a_thin);
                // (a window with a NaN cross section reports inf: every cell then; wavelengths beyond the arrays: exact sum in cell_eval_kernel)
                const int n = (mx < __longlong_as_double(0x7ff0000000000000ll)) ? pb_cell_index(mx) - r.x + 1 : kCellCap;
                r.y = n < 1 ? 1 : (n > kCellCap ? kCellCap : n);
            }
            prm.cell_range[blockIdx.x] = r;
        }
    }
    if (prm.order == nullptr) return;   // sigma cells / sigma-sorted windows: the launch order is not taken from here
    extern __shared__ __align__(16) unsigned char wsm[];
    double* key = reinterpret_cast<double*>(wsm);
    int* idx = reinterpret_cast<int*>(key + n_pow2);
    int* out = prm.order + (int64_t)blockIdx.x * prm.n_win;
    if (n_pow2 == 0) {
        for (int i = tid; i < prm.n_win; i += kCompactThreads) out[i] = i;
        return;
    }
    for (int i = tid; i < n_pow2; i += kCompactThreads) {
        double k = -1.0;                              // padding sorts last
        if (i < prm.n_win) {
            k = 0.0;
            for (int a = 0; a < S; ++a) {
                const double nm = __longlong_as_double((long long)sc.nmax[(int64_t)a * sc.n_orb + p]);
                const double wm = prm.winmax[((int64_t)a * prm.n_phase + blockIdx.x) * prm.n_win + i];
                k += (nm > 0.0) ? nm * wm : 0.0;
            }
            if (!(k == k)) k = __longlong_as_double(0x7ff0000000000000ll);
        }
        key[i] = k;
        idx[i] = i;
    }
    __syncthreads();
    for (int size = 2; size <= n_pow2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = tid; i < n_pow2; i += kCompactThreads) {
                const int j = i ^ stride;
                if (j > i) {
                    const bool desc = (i & size) == 0;
                    const double ki = key[i], kj = key[j];
                    const int ii = idx[i], ij = idx[j];
                    // total order (key, then index): the result does not depend on the network
                    const bool i_first = (ki > kj) || (ki == kj && ii < ij);
                    if (i_first != desc) { key[i] = kj; key[j] = ki; idx[i] = ij; idx[j] = ii; }
                }
            }
            __syncthreads();
        }
    for (int i = tid; i < prm.n_win; i += kCompactThreads) out[i] = idx[i];
}

// ---- the same for large sky grids whose kernel needs no window order (sigma cells, sorted windows): a CLUSTER of
// kCompactCluster CTAs per phase.  One CTA per phase (above) walks 36 000 chords of C5 in 18 rounds of four barriers --
// 83 us on 21 of 148 SMs.  Here CTA r of the cluster takes the r-th contiguous share of the phase's chords: pass 1 counts
// its active chords, the counts are exchanged through distributed shared memory (cluster.sync), pass 2 packs the records
// at base = sum of the counts of the ranks before it -- still chord order, still deterministic.  w0 and the thin totals
// are added per CTA and then across the ranks in rank order by CTA 0, which also leaves n_act, the padding of the last
// tile, the statistics and the cell range of the phase.
constexpr int kCompactCluster = 8;
constexpr int kCcThreads = 512;
struct CompactItem { double w; int m; bool alive, active; };
__device__ __forceinline__ CompactItem pb_compact_item(const DevScene& sc, int S, const double* s_smax, int64_t row0, int q, int n_items,
                                                       int P, int n2) {
    CompactItem it{0.0, q, false, false};
    if (q >= n_items) return it;
    double bound = 0.0;
    if (sc.mirror) {
        const int h = q / (P * sc.n_rho), rem = q - h * (P * sc.n_rho);
        const int i = rem / sc.n_rho, irho = rem - i * sc.n_rho;
        it.m = (h * n2 + i) * sc.n_rho + irho;
        const int ip = n2 - 1 - i;
        const double wa = sc.w_in[row0 + it.m];
        if (ip != i) {
            const int mb = (h * n2 + ip) * sc.n_rho + irho;
            const double wb = sc.w_in[row0 + mb];
            if (wa == 0.0) it.m = mb;               // Ncol of a blocked chord is not evaluated: take the mirror's
            it.w = wa + wb;
        } else {
            it.w = wa;
        }
    } else {
        it.w = sc.w_in[row0 + it.m];
    }
    for (int a = 0; a < S; ++a) bound = fma(sc.ncol[(int64_t)a * sc.n_orb * sc.Mp + row0 + it.m], s_smax[a], bound);
    it.alive = (it.w != 0.0);
    it.active = it.alive && !(bound < 5.5511151231257827e-17);          // 2^-54; NaN -> active
    return it;
}

__global__ void __cluster_dims__(kCompactCluster, 1, 1) __launch_bounds__(kCcThreads)
compact_cluster_kernel(const DevScene* __restrict__ scp, const TransitParams prm, int cull) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const DevScene& sc = *scp;
    const int rank = (int)cluster.block_rank();
    const int pi = blockIdx.x / kCompactCluster, p = prm.phase_lo + pi;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    constexpr int NW = kCcThreads / 32;
    const int S = sc.n_abs, R = pb_rec_stride(S);
    __shared__ double s_smax[kMaxConstituents];
    __shared__ int s_warp[32];
    __shared__ int s_cnt, s_base, s_chunk;
    __shared__ double s_red[7][NW];
    __shared__ double s_sums[7];
    const bool cells = prm.cell_range != nullptr;   // as in compact_kernel
    if (tid < S) {
        const double v = cells ? pb_abs_sigma_bound(sc, tid) : __longlong_as_double((long long)sc.smax[(int64_t)tid * sc.n_orb + p]);
        s_smax[tid] = (cull && v >= 0.0) ? v : __longlong_as_double(0x7ff0000000000000ll);   // negative or NaN: never cull
    }
    __syncthreads();
    const int64_t row0 = (int64_t)p * sc.Mp;
    double* rec_row = sc.rec + row0 * R;
    const int n2 = sc.n_phi / 2, P = (n2 + 1) / 2;
    const int n_items = sc.mirror ? 2 * P * sc.n_rho : sc.Mp;
    const int per_cta = ((n_items + kCompactCluster * kCcThreads - 1) / (kCompactCluster * kCcThreads)) * kCcThreads;
    const int i0 = rank * per_cta;
    int i1 = i0 + per_cta;
    if (i1 > n_items) i1 = n_items;

    // ---- pass 1: this CTA's number of active chords ------------------------------------------------------------
    int cnt = 0;
    for (int q = i0 + tid; q < i1; q += kCcThreads) cnt += pb_compact_item(sc, S, s_smax, row0, q, n_items, P, n2).active ? 1 : 0;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if (lane == 0) s_warp[wid] = cnt;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int i = 0; i < NW; ++i) t += s_warp[i];
        s_cnt = t;
    }
    cluster.sync();
    int base0 = 0, n_rec = 0;
    for (int r = 0; r < kCompactCluster; ++r) {
        const int c = *cluster.map_shared_rank(&s_cnt, r);
        if (r < rank) base0 += c;
        n_rec += c;
    }
    if (tid == 0) s_base = base0;
    __syncthreads();

    // ---- pass 2: pack the records of this share in chord order ---------------------------------------------------
    double w0 = 0.0;
    double tt[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int m0 = i0; m0 < i1; m0 += kCcThreads) {
        const CompactItem it = pb_compact_item(sc, S, s_smax, row0, (m0 + tid < i1) ? m0 + tid : n_items, n_items, P, n2);
        if (it.alive && !it.active) w0 += it.w;
        const unsigned bal = __ballot_sync(0xffffffffu, it.active);
        if (lane == 0) s_warp[wid] = __popc(bal);
        __syncthreads();
        if (wid == 0) {   // exclusive scan of the warp counts
            const int c = (lane < NW) ? s_warp[lane] : 0;
            int x = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, x, d);
                if (lane >= d) x += y;
            }
            s_warp[lane] = x - c;
            if (lane == 31) s_chunk = x;
        }
        __syncthreads();
        const int base = s_base;
        if (it.active) {
            double* r = rec_row + (int64_t)(base + s_warp[wid] + __popc(bal & ((1u << lane) - 1u))) * R;
            const double w = it.w, w1 = w * kExpBigC1, w2 = w * kExpBigC2;
            r[0] = w; r[1] = w1; r[2] = w2;
            for (int a = 0; a < S; ++a) r[3 + a] = sc.ncol[(int64_t)a * sc.n_orb * sc.Mp + row0 + it.m];
            if (3 + S < R) r[3 + S] = 0.0;
            if (S <= 2) {   // same operations as ThinSums::add of kernel (c')
                const double N0 = r[3], N1 = (S == 2) ? r[4] : 0.0;
                tt[0] += w;
                tt[1] = fma(N0, w1, tt[1]);
                tt[2] = fma(N1, w1, tt[2]);
                const double t0 = N0 * w2, t1 = N1 * w2;
                tt[3] = fma(t0, N0, tt[3]);
                tt[4] = fma(t0, N1, tt[4]);
                tt[5] = fma(t1, N1, tt[5]);
            }
        }
        __syncthreads();
        if (tid == 0) s_base = base + s_chunk;
        __syncthreads();
    }
    // this CTA's sums (fixed order), then CTA 0 adds the ranks in rank order
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const double v = pb_warp_sum(k == 0 ? w0 : tt[k == 0 ? 0 : k - 1]);
        if (lane == 0) s_red[k][wid] = v;
    }
    __syncthreads();
    if (tid < 7) {
        double t = 0.0;
        for (int i = 0; i < NW; ++i) t += s_red[tid][i];
        s_sums[tid] = t;
    }
    cluster.sync();
    if (rank == 0) {
        const int n_tiles = (n_rec + kChordTile - 1) / kChordTile;
        for (int i = n_rec * R + tid; i < n_tiles * kChordTile * R; i += kCcThreads) rec_row[i] = 0.0;   // padding
        if (tid < 7) {
            double t = 0.0;
            for (int r = 0; r < kCompactCluster; ++r) t += cluster.map_shared_rank(s_sums, r)[tid];
            if (tid == 0) sc.w0[p] = t; else sc.thin_tot[(int64_t)p * 8 + tid - 1] = t;
        }
        if (tid == 0) {
            sc.n_act[p] = n_tiles;
            atomicAdd(sc.stat + 1, (unsigned long long)n_tiles * kChordTile);
        }
        if (tid < S && !cells) sc.smax[(int64_t)tid * sc.n_orb + p] = 0ull;   // clean for the atomicMax of the next run (every rank has read it)
        if (prm.cell_range != nullptr) {   // as in compact_kernel
            if (tid < kCellCap / 32) prm.cell_count[pi * (kCellCap / 32) + tid] = 0u;
            if (tid == 0) {
                const double mx = pb_abs_sigma_bound(sc, 0) * kExpBigScale;
                const double nmax = __longlong_as_double((long long)sc.nmax[p]);
                int2 rg = make_int2(0, 0);
                const double a_thin = prm.thin_v / nmax;
                if (nmax > 0.0 && n_tiles > 0 && mx >= a_thin && a_thin > 0.0) {
                    rg.x = pb_cell_index(a_thin);
                    const int n = (mx < __longlong_as_double(0x7ff0000000000000ll)) ? pb_cell_index(mx) - rg.x + 1 : kCellCap;
                    rg.y = n < 1 ? 1 : (n > kCellCap ? kCellCap : n);
                }
                prm.cell_range[pi] = rg;
            }
        }
    }
    cluster.sync();   // the other ranks' shared memory stays alive until CTA 0 has read it
}

// ---- calculateNumberDensity materialised (chord-major, reference chord order) ------------------
// out[c*n_x + ix], c = m*n_orb + p.  One warp per chord: rows are written as coalesced 8-byte stores.
__global__ void __launch_bounds__(256)
number_density_grid_kernel(const DevScene* __restrict__ scp, int scenario, double* __restrict__ out,
                           int* __restrict__ oob_flag) {
    const DevScene& sc = *scp;
    const DevScenario& S = sc.scen[scenario];
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    const int64_t n_chords = (int64_t)sc.M * sc.n_orb;
    for (int64_t c = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); c < n_chords;
         c += (int64_t)gridDim.x * warps_per_block) {
        const int m = (int)(c / sc.n_orb), p = (int)(c % sc.n_orb);
        const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
        const double rho = sc.rho[irho];
        const double y = __dmul_rn(rho, sc.sin_phi[iphi]);
        const double z = __dmul_rn(rho, sc.cos_phi[iphi]);
        const double amp = S.amplitude ? S.amplitude[p] : 1.0;
        double* row = out + c * sc.n_x;
        if (S.type == PB200_SCN_SERPENS) {
            int oob = 0;
            for (int ix = lane; ix < sc.n_x; ix += 32) row[ix] = pb_trilinear(S, sc.x[ix], y, z, &oob) * amp;
            if (oob) *oob_flag = 1;
        } else {
            const double dy = y - S.centre_y[p], cx = S.centre_x[p];
            const double* sp = pb_params(S, p);
            for (int ix = lane; ix < sc.n_x; ix += 32) row[ix] = pb_density(S, sp, amp, sc.x[ix] - cx, dy, z);
        }
    }
}

// (a) fast form for the analytic scenarios, materialised n(chord, x) in the reference's chord order c = m*n_orb + p.
// A tile is 256 consecutive chords: one thread per chord hoists its constants (ScnConst of its phase, pre, cc) into
// shared memory, then 8 lanes share a chord row and stride x two samples at a time -- each lane stores 16 bytes, the 8
// lanes of a chord one 128-byte line, a warp four of them: no staging through shared memory, no partial sectors.
// CTAs are persistent over the tiles and keep the 4096-entry exp table (32 KB) and the x axis (up to kRowsMaxXs samples)
// in shared memory: with the table in global memory the kernel waited on those loads (ncu: long_scoreboard the top
// stall, issue slots 54 % busy).  8 B of HBM traffic per sample against ~24 FP64 and ~51 warp
// instructions per 32 samples: the bound is instruction issue, see DESIGN.md.
constexpr int kRowsCta = 256;
constexpr int kRowsMaxXs = 512;
template <int TYPE, bool XSH>
__device__ __forceinline__ void pb_lane_rows(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                             const double* __restrict__ xs, int n_x, int l8, double pre, double cc,
                                             double* __restrict__ row, bool vec) {
    // the values themselves are the output here, so a sample in the denormal band of the exp is redone through exp2()
    // on the spot (a rarely taken branch after the pair, not a call inside the arithmetic)
    bool rare = false;
    if (vec) {
        int ix = 2 * l8;
        if constexpr (XSH) {
            // four samples in flight per lane (two 16-byte stores): the chain of one sample is ~24 dependent FP64 instructions
            for (; ix + 2 * kColLanes < n_x; ix += 4 * kColLanes) {
                const double2 xa = *reinterpret_cast<const double2*>(xs + ix), xb = *reinterpret_cast<const double2*>(xs + ix + 2 * kColLanes);
                const double u0 = xa.x - k.cx, u1 = xa.y - k.cx, u2 = xb.x - k.cx, u3 = xb.y - k.cx;
                const double d0 = u0 * u0, d1 = u1 * u1, d2 = u2 * u2, d3 = u3 * u3;
                double t0 = pb_term<TYPE, true>(k, big, ltab, d0, cc, rare);
                double t1 = pb_term<TYPE, true>(k, big, ltab, d1, cc, rare);
                double t2 = pb_term<TYPE, true>(k, big, ltab, d2, cc, rare);
                double t3 = pb_term<TYPE, true>(k, big, ltab, d3, cc, rare);
                if (rare) {
                    t0 = pb_term_exact<TYPE>(k, ltab, d0, cc); t1 = pb_term_exact<TYPE>(k, ltab, d1, cc);
                    t2 = pb_term_exact<TYPE>(k, ltab, d2, cc); t3 = pb_term_exact<TYPE>(k, ltab, d3, cc);
                    rare = false;
                }
                *reinterpret_cast<double2*>(row + ix) = make_double2(pre * t0, pre * t1);
                *reinterpret_cast<double2*>(row + ix + 2 * kColLanes) = make_double2(pre * t2, pre * t3);
            }
        }
        for (; ix < n_x; ix += 2 * kColLanes) {     // n_x is even here
            double x0, x1;
            if constexpr (XSH) { const double2 xx = *reinterpret_cast<const double2*>(xs + ix); x0 = xx.x; x1 = xx.y; }
            else { x0 = xs[ix]; x1 = xs[ix + 1]; }
            const double u = x0 - k.cx, w = x1 - k.cx;
            const double d0 = u * u, d1 = w * w;
            double t0 = pb_term<TYPE, true>(k, big, ltab, d0, cc, rare);
            double t1 = pb_term<TYPE, true>(k, big, ltab, d1, cc, rare);
            if (rare) { t0 = pb_term_exact<TYPE>(k, ltab, d0, cc); t1 = pb_term_exact<TYPE>(k, ltab, d1, cc); rare = false; }
            *reinterpret_cast<double2*>(row + ix) = make_double2(pre * t0, pre * t1);
        }
    } else {
        for (int ix = l8; ix < n_x; ix += kColLanes) {
            const double u = xs[ix] - k.cx, d0 = u * u;
            double t0 = pb_term<TYPE, true>(k, big, ltab, d0, cc, rare);
            if (rare) { t0 = pb_term_exact<TYPE>(k, ltab, d0, cc); rare = false; }
            row[ix] = pre * t0;
        }
    }
}

template <bool XSH>
__device__ __forceinline__ void pb_lane_rows_any(const ScnConst& k, const double* __restrict__ big, const double2* __restrict__ ltab,
                                                 const double* __restrict__ xs, int n_x, int l8, double pre, double cc,
                                                 double* __restrict__ row, bool vec) {
    switch (k.type) {
        case PB200_SCN_BAROMETRIC: pb_lane_rows<PB200_SCN_BAROMETRIC, XSH>(k, big, ltab, xs, n_x, l8, pre, cc, row, vec); break;
        case PB200_SCN_HYDROSTATIC: pb_lane_rows<PB200_SCN_HYDROSTATIC, XSH>(k, big, ltab, xs, n_x, l8, pre, cc, row, vec); break;
        case PB200_SCN_POWERLAW: pb_lane_rows<PB200_SCN_POWERLAW, XSH>(k, big, ltab, xs, n_x, l8, pre, cc, row, vec); break;
        case PB200_SCN_TORUS: pb_lane_rows<PB200_SCN_TORUS, XSH>(k, big, ltab, xs, n_x, l8, pre, cc, row, vec); break;
        default: break;
    }
}

__global__ void __launch_bounds__(kRowsCta, 3)
number_density_rows_kernel(const DevScene* __restrict__ scp, int scenario, double* __restrict__ out, int64_t n_tiles,
                           unsigned long long* __restrict__ next_tile) {
    const DevScene& sc = *scp;
    const DevScenario& S = sc.scen[scenario];
    extern __shared__ __align__(16) double s_dyn[];   // exp table [kExpBigEntries] | x axis [n_x] (n_x <= kRowsMaxXs)
    __shared__ ScnConst s_k[kRowsCta];
    __shared__ double s_pre[kRowsCta], s_cc[kRowsCta];
    __shared__ long long s_tile;
    const int tid = threadIdx.x, grp = tid / kColLanes, l8 = tid % kColLanes;
    const int64_t n_chords = (int64_t)sc.M * sc.n_orb;
    const int n_x = sc.n_x;
    const double* __restrict__ gbig = sc.exp_tab + kExpTabEntries;
    const double2* __restrict__ ltab = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries + kExpBigEntries);
    const bool vec = (n_x % 2 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    const bool xsh = n_x <= kRowsMaxXs;
    double* s_big = s_dyn;
    double* s_xs = s_dyn + kExpBigEntries;
    {
        const double2* src = reinterpret_cast<const double2*>(gbig);
        double2* dst = reinterpret_cast<double2*>(s_big);
        for (int i = tid; i < kExpBigEntries / 2; i += kRowsCta) dst[i] = __ldg(src + i);
        if (xsh) for (int i = tid; i < n_x; i += kRowsCta) s_xs[i] = sc.x[i];
    }
    // persistent CTAs; the first tile is the CTA's own, further ones come from a counter (rows differ in cost: a torus row
    // off the plane is all zeros), so the CTAs finish together
    for (int64_t tile = blockIdx.x; tile < n_tiles;) {
        const int64_t c0 = tile * kRowsCta, c = c0 + tid;
        __syncthreads();   // the previous tile's constants are no longer read (first pass: the tables are complete)
        if (tid == 0) s_tile = (long long)gridDim.x + (long long)atomicAdd(next_tile, 1ull);
        if (c < n_chords) {
            const int m = (int)(c / sc.n_orb), p = (int)(c % sc.n_orb);
            const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
            const double rho = sc.rho[irho];
            const ScnConst k = pb_scn_const(S, p);
            double pre, cc;
            pb_chord_pre(k, gbig, __dmul_rn(rho, sc.sin_phi[iphi]), __dmul_rn(rho, sc.cos_phi[iphi]), pre, cc);
            s_k[tid] = k;
            s_pre[tid] = pre;
            s_cc[tid] = cc;
        }
        __syncthreads();
        const int rows = (int)((n_chords - c0 < kRowsCta) ? n_chords - c0 : kRowsCta);
        for (int i = grp; i < rows; i += kRowsCta / kColLanes) {
            const ScnConst k = s_k[i];
            const double pre = s_pre[i], cc = s_cc[i];
            double* __restrict__ row = out + (c0 + i) * n_x;
            if (pre == 0.0) {   // every term is finite and <= 1: the whole row is exactly 0 (torus far off the plane)
                if (vec) for (int ix = 2 * l8; ix < n_x; ix += 2 * kColLanes) *reinterpret_cast<double2*>(row + ix) = make_double2(0.0, 0.0);
                else for (int ix = l8; ix < n_x; ix += kColLanes) row[ix] = 0.0;
                continue;
            }
            if (xsh) pb_lane_rows_any<true>(k, s_big, ltab, s_xs, n_x, l8, pre, cc, row, vec);
            else pb_lane_rows_any<false>(k, s_big, ltab, sc.x, n_x, l8, pre, cc, row, vec);
        }
        tile = s_tile;   // written before the barrier in the middle of this iteration, read after the row loop
    }
}

// same for caller-supplied chords (API parity with <Scenario>.calculateNumberDensity(x, phi, rho, orb))
__global__ void __launch_bounds__(256)
number_density_chords_kernel(const DevScene* __restrict__ scp, int scenario, int64_t n,
                             const double* __restrict__ y, const double* __restrict__ z,
                             const double* __restrict__ cx, const double* __restrict__ cy,
                             const double* __restrict__ amp, double* __restrict__ out,
                             int* __restrict__ oob_flag) {
    const DevScene& sc = *scp;
    const DevScenario& S = sc.scen[scenario];
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    for (int64_t c = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); c < n;
         c += (int64_t)gridDim.x * warps_per_block) {
        const double a = amp ? amp[c] : 1.0;
        double* row = out + c * sc.n_x;
        if (S.type == PB200_SCN_SERPENS) {
            int oob = 0;
            for (int ix = lane; ix < sc.n_x; ix += 32) row[ix] = pb_trilinear(S, sc.x[ix], y[c], z[c], &oob) * a;
            if (oob) *oob_flag = 1;
        } else {
            const double dy = y[c] - cy[c];
            for (int ix = lane; ix < sc.n_x; ix += 32) row[ix] = pb_density(S, S.p, a, sc.x[ix] - cx[c], dy, z[c]);
        }
    }
}

// ---- F_out for F* == 1: sum_m rho*clv, one block, fixed order (deterministic) -------------------
__global__ void __launch_bounds__(1024) fout_scalar_kernel(const DevScene* __restrict__ scp) {
    const DevScene& sc = *scp;
    __shared__ double part[32];
    double s = 0.0;
    for (int m = threadIdx.x; m < sc.M; m += blockDim.x) s += sc.w_out[m];
    s = pb_warp_sum(s);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        double t = (threadIdx.x < (blockDim.x >> 5)) ? part[threadIdx.x] : 0.0;
        t = pb_warp_sum(t);
        if (threadIdx.x == 0) sc.fout[0] = t;
    }
}

// ---- host-callable launchers ---------------------------------------------------------------------
static inline int grid_for_warps(int64_t n_warp_items, int sm_count) {
    const int64_t blocks = (n_warp_items + 7) / 8;   // 8 warps per 256-thread block
    const int64_t cap = (int64_t)sm_count * 8;       // 8 resident blocks per SM, grid-stride beyond
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

void launch_sky_cells(const DevScene* d_scene, int Mp, cudaStream_t st) {
    sky_cell_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(d_scene);
}
void launch_fout_scalar(const DevScene* d_scene, cudaStream_t st) {
    fout_scalar_kernel<<<1, 1024, 0, st>>>(d_scene);
}
void launch_column(const DevScene* d_scene, int Mp, int phase_lo, int n_phase, int* oob, int sm_count,
                   cudaStream_t st, bool has_molecules, int run_mode) {
    if (has_molecules) {   // density rows n(c, x) are needed by kernel (m): warp per chord, coalesced row stores
        column_kernel<<<grid_for_warps((int64_t)n_phase * Mp, sm_count), 256, 0, st>>>(d_scene, phase_lo, n_phase, oob);
    } else {
        const int64_t n = (int64_t)n_phase * Mp;   // Mp is a multiple of kColCta: a CTA never straddles two phases
        (void)n;
        for (int p0 = 0; p0 < n_phase; p0 += 65535) {   // gridDim.y limit; a retrieval batch can stack more phases than that
            const int np = n_phase - p0 < 65535 ? n_phase - p0 : 65535;
            column_cta_kernel<<<dim3((unsigned)(Mp / kColCta), (unsigned)np), kColCta, 0, st>>>(d_scene, phase_lo + p0, np, oob, run_mode);
        }
    }
}
constexpr int kMaxWindows = 8192;
cudaError_t launch_compact(const DevScene* d_scene, const TransitParams& prm, bool cull, int chords_per_phase, cudaStream_t st) {
    const bool small = chords_per_phase <= 4096;
    int n_pow2 = 1;
    while (n_pow2 < prm.n_win) n_pow2 <<= 1;
    if (n_pow2 > kMaxWindows) n_pow2 = 0;
    const size_t smem = (size_t)n_pow2 * (sizeof(double) + sizeof(int));
    if (smem > 40 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(compact_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(compact_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    // no window order wanted (sigma cells, sorted windows) and a large sky grid: a cluster of CTAs per phase (C5, 36 000 chords
    // per phase; at C4's 9 000 the single CTA is as fast and the cluster launch costs more than it saves)
    if (chords_per_phase >= (1 << 14) && prm.order == nullptr) {
        compact_cluster_kernel<<<prm.n_phase * kCompactCluster, kCcThreads, 0, st>>>(d_scene, prm, cull ? 1 : 0);
        return cudaGetLastError();
    }
    if (small) compact_kernel<256><<<prm.n_phase, 256, smem, st>>>(d_scene, prm, cull ? 1 : 0, n_pow2);
    else compact_kernel<1024><<<prm.n_phase, 1024, smem, st>>>(d_scene, prm, cull ? 1 : 0, n_pow2);
    return cudaGetLastError();
}
void launch_number_density_grid(const DevScene* d_scene, int64_t n_chords, int scenario, double* out,
                                int* oob, int sm_count, cudaStream_t st) {
    number_density_grid_kernel<<<grid_for_warps(n_chords, sm_count), 256, 0, st>>>(d_scene, scenario, out, oob);
}
void launch_number_density_thread(const DevScene* d_scene, int64_t n_chords, int n_x, int scenario, double* out, int sm_count,
                                  unsigned long long* next_tile, cudaStream_t st) {
    const int64_t n_tiles = (n_chords + kRowsCta - 1) / kRowsCta;
    const int64_t slots = (int64_t)sm_count * 3;     // resident CTAs (<= 85 registers x 256 threads)
    const int64_t ctas = n_tiles < slots ? n_tiles : slots;
    cudaMemsetAsync(next_tile, 0, sizeof(unsigned long long), st);
    const size_t smem = (kExpBigEntries + (n_x <= kRowsMaxXs ? ((n_x + 1) & ~1) : 0)) * sizeof(double);
    static bool attr_set = false;      // 32 KB of table + the static arrays exceed the 48 KB default
    if (!attr_set) {
        cudaFuncSetAttribute(number_density_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)((kExpBigEntries + kRowsMaxXs) * sizeof(double)));
        attr_set = true;
    }
    number_density_rows_kernel<<<(unsigned)(ctas > 0 ? ctas : 1), kRowsCta, smem, st>>>(d_scene, scenario, out, n_tiles, next_tile);
}
void launch_number_density_chords(const DevScene* d_scene, int scenario, int64_t n, const double* y,
                                  const double* z, const double* cx, const double* cy, const double* amp,
                                  double* out, int* oob, int sm_count, cudaStream_t st) {
    number_density_chords_kernel<<<grid_for_warps(n, sm_count), 256, 0, st>>>(d_scene, scenario, n, y, z, cx,
                                                                              cy, amp, out, oob);
}

}  // namespace pb200
