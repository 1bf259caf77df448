// pb200_density.cu -- kernels (a)/(a'): chord weights, blocking mask, number density and
// column density on the chord grid.  SURVEY 8(a) rows a2, a4, a6, a7, a8 (column part).
//
// Work decomposition: one warp per chord (phase p, sky cell m = i_phi*n_rho + i_rho); the 32 lanes
// stride over the line-of-sight axis x and are reduced with warp shuffles in fp64.  Chord axes are a
// few hundred doubles and live in L1; the only HBM traffic is the coalesced store of the result.
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

// ---- (a') fast form for scenes without molecules: one THREAD per chord -------------------------
// All 32 lanes stay busy whatever n_x is, there are no shuffles, x[ix] is a warp-uniform load, and the
// per-chord pieces of the formulas are hoisted out of the x loop (torus: exp(-(z/H)^2); power law:
// (R/r)^q = exp(q (ln R - ln(r^2)/2)) instead of pow: <= 3e-14 relative, Ncol is tested to 1e-12).
// Four partial sums per thread give ILP and halve the rounding growth of the long sum.
template <class F>
__device__ __forceinline__ double pb_sum4(int n, F term) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int ix = 0;
    for (; ix + 4 <= n; ix += 4) {
        a0 += term(ix); a1 += term(ix + 1); a2 += term(ix + 2); a3 += term(ix + 3);
    }
    for (; ix < n; ++ix) a0 += term(ix);
    return (a0 + a1) + (a2 + a3);
}

// The analytic scenarios along one chord: use(prefactor, term) is called once with the per-chord prefactor and a
// functor term(ix) such that n(x_ix) = prefactor * term(ix) -- everything that does not depend on x is hoisted
// (torus: exp(-(z/H)^2); power law: q log2 R), exp / pow go through the L1-resident tables of pb200_math.cuh.
// Shared by the fused x-sum (kernel (a'), use = pb_sum4) and the materialising kernel (a) (use = store).
template <class Use>
__device__ __forceinline__ double pb_chord_terms(const DevScene& sc, const DevScenario& S, int p, double amp, double y,
                                                 double z, Use&& use) {
    const double* __restrict__ xs = sc.x;
    const double cx = S.centre_x[p];
    const double dy = y - S.centre_y[p];
    const double* __restrict__ P = pb_params(S, p);
    const double n0 = P[0] * amp;
    const double* __restrict__ big = sc.exp_tab + kExpTabEntries;
    const double2* __restrict__ ltab = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries + kExpBigEntries);
    const double kLog2e = 1.4426950408889634;
    switch (S.type) {
        case PB200_SCN_BAROMETRIC: {   // n0 exp((R - r)/H) H(r - R), gasProperties.py:155-160
            const double R = P[1], invH = kLog2e / P[2], c2 = fma(dy, dy, z * z);
            return use(n0, [&](int ix) {
                const double dx = xs[ix] - cx;
                const double r = sqrt(fma(dx, dx, c2));
                return (r < R) ? 0.0 : pb_exp2_tab(big, (R - r) * invH);
            });
        }
        case PB200_SCN_HYDROSTATIC: {  // n0 exp(J(r) H(r - R) - J(R)), gasProperties.py:197-203
            const double R = P[1], jc = P[2] / P[3], j0 = jc / R, c2 = fma(dy, dy, z * z);
            return use(n0, [&](int ix) {
                const double dx = xs[ix] - cx;
                const double r = sqrt(fma(dx, dx, c2));
                return pb_exp2_tab(big, (((r < R) ? 0.0 : jc / r) - j0) * kLog2e);
            });
        }
        case PB200_SCN_POWERLAW: {     // n0 (R/r)^q H(r - R), gasProperties.py:240-243, :326-329, :371-373, :453-460
            // (R/r)^q = exp2(q log2 R - (q/2) log2 r^2); r < R <=> r^2 < R^2 up to a rounding of the boundary sample,
            // which the reference's own sqrt does not define better
            const double R = P[1], q = P[2], qlog2R = q * log2(R), hq = -0.5 * q, c2 = fma(dy, dy, z * z), R2 = R * R;
            return use(n0, [&](int ix) {
                const double dx = xs[ix] - cx;
                const double r2 = fma(dx, dx, c2);
                return (r2 < R2) ? 0.0 : pb_exp2_tab(big, fma(hq, pb_log2_tab(ltab, r2), qlog2R));
            });
        }
        case PB200_SCN_TORUS: {        // n0 exp(-((a - a_t)/4H)^2) exp(-(z/H)^2), gasProperties.py:506-515
            const double at = P[1], inv4H = 1.0 / (4.0 * P[2]), qz = z / P[2], dy2 = dy * dy;
            const double ez = exp(-(qz * qz));
            return use(n0 * ez, [&](int ix) {
                const double dx = xs[ix] - cx;
                const double qa = (sqrt(fma(dx, dx, dy2)) - at) * inv4H;
                return pb_exp2_tab(big, -(qa * qa) * kLog2e);
            });
        }
        default:
            return 0.0;
    }
}

__device__ __forceinline__ double pb_column_sum(const DevScene& sc, const DevScenario& S, int p, double amp, double y,
                                                double z, int* oob_flag) {
    const int n_x = sc.n_x;
    if (S.type == PB200_SCN_SERPENS) {
        const double* __restrict__ xs = sc.x;
        int oob = 0;
        const double sum = pb_sum4(n_x, [&](int ix) { return pb_trilinear(S, xs[ix], y, z, &oob); });
        if (oob) *oob_flag = 1;
        return sum * amp;
    }
    return pb_chord_terms(sc, S, p, amp, y, z, [&](double pre, auto term) {
        return (pre == 0.0) ? 0.0 : pre * pb_sum4(n_x, term);   // terms are finite and <= 1
    });
}

// Upper bound of the x-sum of one scenario on a chord (every term <= its value at dx = 0, the closest
// approach r_min = sqrt(dy^2 + z^2)); +inf where no cheap bound exists.  Used by the run path only, to skip
// the sum on chords that compact_kernel will classify as transparent anyway.
__device__ __forceinline__ double pb_column_bound(const DevScene& sc, const DevScenario& S, int p, double amp, double y,
                                                  double z) {
    if (S.type == PB200_SCN_SERPENS) return __longlong_as_double(0x7ff0000000000000ll);
    const double dy = y - S.centre_y[p];
    const double* P = pb_params(S, p);
    const double n0 = fabs(P[0] * amp) * sc.n_x;
    const double rmin = sqrt(fma(dy, dy, z * z));
    switch (S.type) {
        case PB200_SCN_BAROMETRIC: return n0 * ((rmin < P[1]) ? 1.0 : exp((P[1] - rmin) / P[2]));
        case PB200_SCN_HYDROSTATIC: {
            const double jc = P[2] / P[3];
            return n0 * ((rmin < P[1]) ? 1.0 : exp(jc / rmin - jc / P[1]));
        }
        case PB200_SCN_POWERLAW:
            // (R/r)^q falls with r only for q > 0; for q <= 0 (the reference accepts any q) the closest approach is not
            // the maximum: no cheap bound
            if (!(P[2] > 0.0)) return __longlong_as_double(0x7ff0000000000000ll);
            return n0 * ((rmin < P[1]) ? 1.0 : exp(P[2] * (log(P[1]) - log(rmin))));
        case PB200_SCN_TORUS: { const double qz = z / P[2]; return n0 * exp(-(qz * qz)); }
        default: return __longlong_as_double(0x7ff0000000000000ll);
    }
}

// run_mode = 0: every unblocked chord gets its exact column densities (inner seams of the API).
// run_mode = 2 (pb200_run with PB200_NO_CULL=1): exact sums for every unblocked chord, absorbers as in run_mode 1.
// run_mode = 1 (pb200_run): the sums that compact_kernel does not need are skipped --
//   * the mirror chord of a pair whose primary chord is not blocked (sc.mirror);
//   * chords whose upper bound sum_a bound_a * max_w sigma_a(p, w) is already below 2^-54: they are
//     transparent whatever the exact sum is, and get the bound stored as Ncol (so that compact_kernel's own
//     test classifies them the same way).
__global__ void __launch_bounds__(128, 6)
column_thread_kernel(const DevScene* __restrict__ scp, int phase_lo, int n_phase, int* __restrict__ oob_flag,
                     int run_mode) {
    const DevScene& sc = *scp;
    const int64_t cid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cid >= (int64_t)n_phase * sc.Mp) return;
    const int p = phase_lo + (int)(cid / sc.Mp);
    const int m = (int)(cid % sc.Mp);
    const int64_t row = (int64_t)p * sc.Mp + m;
    bool blocked = true, need_sum = false;
    double y = 0.0, z = 0.0, wv = 0.0;
    if (m < sc.M) {
        const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
        const double rho = sc.rho[irho];
        y = __dmul_rn(rho, sc.sin_phi[iphi]);
        z = __dmul_rn(rho, sc.cos_phi[iphi]);
        blocked = pb_blocked(sc, p, y, z);
        wv = blocked ? 0.0 : sc.w_out[m];
        need_sum = !blocked;
        if (run_mode == 1 && sc.mirror && need_sum) {
            const int n2 = sc.n_phi / 2, h = iphi / n2, i = iphi - h * n2, ip = n2 - 1 - i;
            if (i > ip) {   // mirror chord of the pair: needed only if the primary chord is blocked
                const int jphi = h * n2 + ip;
                need_sum = pb_blocked(sc, p, __dmul_rn(rho, sc.sin_phi[jphi]), __dmul_rn(rho, sc.cos_phi[jphi]));
            }
        }
    }
    sc.w_in[row] = wv;
    bool skip = false;
    if (run_mode) {
        // The run path works with one absorber per group of atomic constituents that share a scenario
        // (DevScene::n_abs): Ncol_g = sum_x n dx, the mixing ratios are in the group's cross section.
        if (need_sum && run_mode == 1) {   // transparent whatever the exact sums are?
            double bound = 0.0;
            for (int g = 0; g < sc.n_abs; ++g) {
                const DevScenario& S = sc.scen[sc.cons[sc.abs_first[g]].scenario];
                const double ub = pb_column_bound(sc, S, p, S.amplitude ? S.amplitude[p] : 1.0, y, z);
                const double sm = __longlong_as_double((long long)sc.smax[(int64_t)g * sc.n_orb + p]);
                bound += ub * sc.delta_x * ((sm >= 0.0) ? sm : __longlong_as_double(0x7ff0000000000000ll));
            }
            skip = bound < 5.5511151231257827e-17;   // 2^-54; NaN (0 * inf) -> not skipped
        }
        int prev_scen = -1;
        double sum = 0.0;
        for (int g = 0; g < sc.n_abs; ++g) {
            const int s = sc.cons[sc.abs_first[g]].scenario;
            const DevScenario& S = sc.scen[s];
            if (s != prev_scen) {   // (PB200_NO_GROUP: several absorbers of one scenario share the sum)
                const double amp = S.amplitude ? S.amplitude[p] : 1.0;
                sum = 0.0;
                if (need_sum) sum = skip ? pb_column_bound(sc, S, p, amp, y, z) : pb_column_sum(sc, S, p, amp, y, z, oob_flag);
                prev_scen = s;
            }
            const double col = sum * sc.delta_x;
            sc.ncol[(int64_t)g * sc.n_orb * sc.Mp + row] = col;
            double mx = (col > 0.0) ? col : 0.0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
            if ((threadIdx.x & 31) == 0 && mx > 0.0)
                atomicMax(sc.nmax + (int64_t)g * sc.n_orb + p, (unsigned long long)__double_as_longlong(mx));
        }
        return;
    }
    for (int s = 0; s < sc.n_scen; ++s) {
        const DevScenario& S = sc.scen[s];
        const double amp = S.amplitude ? S.amplitude[p] : 1.0;
        // the reference evaluates only unblocked chords (gasProperties.py:1235-1240)
        double sum = 0.0;
        if (need_sum) sum = skip ? pb_column_bound(sc, S, p, amp, y, z) : pb_column_sum(sc, S, p, amp, y, z, oob_flag);
        for (int k = S.first_cons; k < S.first_cons + S.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            if (C.kind != 0) continue;
            const double col = (sum * C.chi) * sc.delta_x;                      // gasProperties.py:940
            sc.ncol[(int64_t)C.atom_index * sc.n_orb * sc.Mp + row] = col;
            // max over the warp first (a CTA never straddles two phases: Mp is a multiple of 256): one atomic
            // per warp instead of 32 on the same address
            double mx = (col > 0.0) ? col : 0.0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
            if ((threadIdx.x & 31) == 0 && mx > 0.0)
                atomicMax(sc.nmax + (int64_t)C.atom_index * sc.n_orb + p, (unsigned long long)__double_as_longlong(mx));
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
constexpr int kCompactThreads = 1024;
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
    if (tid < S) {
        const double v = __longlong_as_double((long long)sc.smax[(int64_t)tid * sc.n_orb + p]);
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
            const int c = s_warp[lane];
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
        const double t = pb_warp_sum(s_red[wid][lane]);
        if (lane == 0) {
            if (wid == 0) sc.w0[p] = t; else sc.thin_tot[(int64_t)p * 8 + wid - 1] = t;
        }
    }
    if (tid == 0) {
        sc.n_act[p] = n_tiles;
        atomicAdd(sc.stat + 1, (unsigned long long)n_tiles * kChordTile);
    }
    if (tid < S) sc.smax[(int64_t)tid * sc.n_orb + p] = 0ull;   // clean for the atomicMax of the next run

    // ---- order of the wavelength windows of this phase for kernel (c') -------------------------------
    // A CTA of kernel (c') owns one window; windows in a line core take the table exp on most chords, windows
    // in the continuum none.  They are launched in decreasing order of optical thickness,
    // key = sum_a max_m Ncol_a(p, m) * max_{w in window} |nsig_a(p, w)| (the largest scaled tau of the window),
    // so that the long CTAs start first (bitonic network in shared memory; the order only matters for the
    // load balance, never for the results).  More than kMaxWindows windows: identity order.
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

// (a) fast form for the analytic scenarios: one THREAD per chord (per-chord factors hoisted once, all lanes busy
// whatever n_x is), 16 samples at a time staged through shared memory so that the warp writes whole 128-byte
// segments: a store instruction covers two rows x 16 doubles.  HBM-bound by construction (8 B per sample).
constexpr int kDensChunk = 16;
__global__ void __launch_bounds__(128)
number_density_thread_kernel(const DevScene* __restrict__ scp, int scenario, double* __restrict__ out) {
    const DevScene& sc = *scp;
    const DevScenario& S = sc.scen[scenario];
    __shared__ double tile[4][32][kDensChunk + 1];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t n_chords = (int64_t)sc.M * sc.n_orb;
    const int64_t c0 = ((int64_t)blockIdx.x * 4 + wid) * 32;
    if (c0 >= n_chords) return;
    const int64_t c = c0 + lane;
    const bool have = c < n_chords;
    const int64_t cc = have ? c : n_chords - 1;
    const int m = (int)(cc / sc.n_orb), p = (int)(cc % sc.n_orb);
    const int iphi = m / sc.n_rho, irho = m - iphi * sc.n_rho;
    const double rho = sc.rho[irho];
    const double y = __dmul_rn(rho, sc.sin_phi[iphi]);
    const double z = __dmul_rn(rho, sc.cos_phi[iphi]);
    const double amp = S.amplitude ? S.amplitude[p] : 1.0;
    const int n_x = sc.n_x;
    double (*my)[kDensChunk + 1] = tile[wid];
    const int rows = (int)((n_chords - c0 < 32) ? n_chords - c0 : 32);
    // blockIdx.y splits the x axis in whole chunks: enough tasks that the last wave is short
    const int n_chunks = (n_x + kDensChunk - 1) / kDensChunk;
    const int x_lo = (int)((int64_t)blockIdx.y * n_chunks / gridDim.y) * kDensChunk;
    int x_hi = (int)((int64_t)(blockIdx.y + 1) * n_chunks / gridDim.y) * kDensChunk;
    if (x_hi > n_x) x_hi = n_x;
    pb_chord_terms(sc, S, p, amp, y, z, [&](double pre, auto term) {
        const bool dead = (pre == 0.0);   // every term is finite and <= 1: the whole row is exactly 0 (torus far off the plane)
        for (int x0 = x_lo; x0 < x_hi; x0 += kDensChunk) {
            const int nk = (x_hi - x0 < kDensChunk) ? x_hi - x0 : kDensChunk;
#pragma unroll 4
            for (int k = 0; k < nk; ++k) my[lane][k] = dead ? 0.0 : pre * term(x0 + k);
            __syncwarp();
            const int col = lane & 15, half = lane >> 4;
            if (col < nk) {
#pragma unroll 4
                for (int r = half; r < rows; r += 2) out[(c0 + r) * n_x + x0 + col] = my[r][col];
            }
            __syncwarp();
        }
        return 0.0;
    });
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
        const int64_t n = (int64_t)n_phase * Mp;
        column_thread_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_scene, phase_lo, n_phase, oob, run_mode);
    }
}
constexpr int kMaxWindows = 8192;
cudaError_t launch_compact(const DevScene* d_scene, const TransitParams& prm, bool cull, cudaStream_t st) {
    int n_pow2 = 1;
    while (n_pow2 < prm.n_win) n_pow2 <<= 1;
    if (n_pow2 > kMaxWindows) n_pow2 = 0;
    const size_t smem = (size_t)n_pow2 * (sizeof(double) + sizeof(int));
    if (smem > 40 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    compact_kernel<<<prm.n_phase, kCompactThreads, smem, st>>>(d_scene, prm, cull ? 1 : 0, n_pow2);
    return cudaGetLastError();
}
void launch_number_density_grid(const DevScene* d_scene, int64_t n_chords, int scenario, double* out,
                                int* oob, int sm_count, cudaStream_t st) {
    number_density_grid_kernel<<<grid_for_warps(n_chords, sm_count), 256, 0, st>>>(d_scene, scenario, out, oob);
}
void launch_number_density_thread(const DevScene* d_scene, int64_t n_chords, int n_x, int scenario, double* out, int sm_count,
                                  cudaStream_t st) {
    const int64_t blocks = (n_chords + 127) / 128;
    const int n_chunks = (n_x + kDensChunk - 1) / kDensChunk;
    int64_t xs = ((int64_t)sm_count * 9 * 8 + blocks - 1) / blocks;   // ~8 tasks per resident block slot
    if (xs > n_chunks) xs = n_chunks;
    if (xs < 1) xs = 1;
    number_density_thread_kernel<<<dim3((unsigned)blocks, (unsigned)xs), 128, 0, st>>>(d_scene, scenario, out);
}
void launch_number_density_chords(const DevScene* d_scene, int scenario, int64_t n, const double* y,
                                  const double* z, const double* cx, const double* cy, const double* amp,
                                  double* out, int* oob, int sm_count, cudaStream_t st) {
    number_density_chords_kernel<<<grid_for_warps(n, sm_count), 256, 0, st>>>(d_scene, scenario, n, y, z, cx,
                                                                              cy, amp, out, oob);
}

}  // namespace pb200
