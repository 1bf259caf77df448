// pb200_transit.cu -- kernel (c'): fused optical depth + Beer-Lambert + stellar-disk sum.
//
//   R(p, w) = sum_m w_in(p,m) exp(-sum_a Ncol_a(p,m) sigma_a(p,w)) / sum_m w_out(m)
//
// restates Transit.sumOverChords (gasProperties.py:1205-1258) together with the atomic branch of
// Atmosphere.getLOSopticalDepth_Batch (:937-954) without ever materialising tau(chord, lambda):
//   * sigma_phase_kernel looks up -sigma_a(p, w) * 4096/ln2 once per (phase, wavelength)
//     (np.interp + 10**v - offset, i.e. n_interp_log :34-51) -- the chord-axis splits of one
//     (phase, wavelength) tile would otherwise each repeat the binary search;
//   * one CTA of transit_kernel = one window of 32 WPT wavelengths of one orbital phase; a thread keeps
//     its WPT values of every absorber in registers, the 8 warps hold the same wavelengths and split the
//     chords;
//   * the ACTIVE chord records {w, w c1, w c2, Ncol_0..Ncol_{S-1}} of the phase (compact_kernel:
//     transparent chords removed, mirror chords paired) stream through shared memory in 256-chord tiles,
//     one 1-D TMA bulk copy per tile (cp.async.bulk + mbarrier, double buffered); every thread of a warp
//     reads the same record, so the LDS is a broadcast;
//   * optically thin chords (warp-uniform test) are summed as a second-order expansion per chord
//     (ThinSums), the rest take the fp64 table exp (see "Big-table form" in pb200_math.cuh: the kernel
//     is bound by the FP64 pipe and by the instructions issued next to it, so the exp is built to need
//     as few of both as possible: 6 FP64 + 3 integer + 1 LDS per sample for one absorber); fp64
//     accumulators; the warps' sums are added in a fixed order (deterministic).  Normally the kernel
//     writes R itself; only when a run has fewer (window, phase) pairs than CTA slots is the chord axis
//     split over blockIdx.z and the partial sums reduced by finalize_kernel.
#include "pb200_internal.cuh"

namespace pb200 {


// ---- mbarrier / TMA bulk-copy wrappers (PTX ISA 8.x, sm_90+) -----------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}


__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds_v2f64(uint32_t addr, double& a, double& b) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "r"(addr));
}

// (b1) for the whole output tile: nsig[a][p][w] = -sigma_a(p, lambda_w) * 4096/ln2, and the largest
// sigma_a of the phase over the run's wavelengths (smax, for the transparent-chord test of compact_kernel)
__global__ void __launch_bounds__(256) sigma_phase_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    __shared__ double s_wmax[8];
    const int64_t wi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = wi < prm.n_wav_sel;
    const int pi = blockIdx.y, p = prm.phase_lo + pi;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) { sc.stat[1] = 0ull; sc.stat[2] = 0ull; sc.stat[4] = 0ull; }   // run statistics
    const double lam = live ? __ldg(sc.wavelength + pb_wav_index(prm, wi)) : 0.0;
    const int wpw = prm.win / 32;   // warps per window: this 256-wavelength CTA covers 8 / wpw whole windows
    for (int a = 0; a < sc.n_abs; ++a) {   // absorber = the atomic constituents of one scenario (DevScene::n_abs)
        unsigned long long bits = 0ull;
        double ns_abs = 0.0;
        if (live) {
            const int k0 = sc.abs_first[a];
            const double sl = __ldg(sc.scen[sc.cons[k0].scenario].shift + p) * lam;             // :917
            double v = 0.0;
            for (int k = k0; k < k0 + sc.abs_count[a]; ++k) {
                const DevConstituent& C = sc.cons[k];
                v = fma(C.chi, exp10(pb_cons_log_sigma(C, sl)) - C.offset, v);                                      // :940 chi, :950-951 sigma
            }
            const double ns = -v * kExpBigScale;
            prm.nsig[((int64_t)a * prm.n_phase + pi) * prm.n_wav_sel + wi] = ns;
            bits = (v >= 0.0) ? (unsigned long long)__double_as_longlong(v) : ~0ull;   // negative or NaN: never cull
            ns_abs = (ns == ns) ? fabs(ns) : __longlong_as_double(0x7ff0000000000000ll);      // NaN: never thin
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, bits, d);
            bits = o > bits ? o : bits;
            ns_abs = fmax(ns_abs, __shfl_xor_sync(0xffffffffu, ns_abs, d));
        }
        if (lane == 0 && bits) atomicMax(sc.smax + (int64_t)a * sc.n_orb + p, bits);
        // window maxima of |nsig_a| for the thin test / the thickness order of kernel (c')
        __syncthreads();
        if (lane == 0) s_wmax[wid] = ns_abs;
        __syncthreads();
        if (threadIdx.x < 8 / wpw) {
            const int window = blockIdx.x * (8 / wpw) + threadIdx.x;
            if (window < prm.n_win) {
                double mx = 0.0;
                for (int i = 0; i < wpw; ++i) mx = fmax(mx, s_wmax[threadIdx.x * wpw + i]);
                prm.winmax[((int64_t)a * prm.n_phase + pi) * prm.n_win + window] = mx;
            }
        }
    }
}

// Optically thin chords.  If sum_a N_a max|nsig_a| over the warp's wavelengths is below kThinV (tau < 2.1e-5),
// then k = 0 and T = 1 for all of them and w exp(-tau) = w + v (w c1 + v w c2), v = sum_a N_a nsig_a, with a
// truncation error tau^3/6 < 1.6e-15.  For one or two absorbers that polynomial separates into chord sums that
// do not depend on the wavelength,
//     sum_m w_m exp(-tau_mj) = SW + nsig_j (SA + nsig_j SB),   SW = sum w, SA = sum N w c1, SB = sum N^2 w c2,
// so a thin chord costs 4 FP64 instructions per THREAD (8 for two absorbers) instead of 6 FP64 + 4 integer
// per SAMPLE; the sums are applied to the accumulators once, after the last tile (ThinSums::finish).  The
// test is warp-uniform (N is a broadcast, the maxima are warp-wide), so the loop never diverges.
constexpr double kThinV = 0.125;
template <int S>
struct ThinSums {
    double sw = 0.0, sa[S <= 2 ? S : 1] = {}, sb[S == 1 ? 1 : 3] = {};
    int thick = 0;   // chords of this warp that took the full path (statistics)
    __device__ __forceinline__ void add(double w0, double w1, double w2, const double (&N)[S]) {
        sw += w0;
        if constexpr (S == 1) {
            sa[0] = fma(N[0], w1, sa[0]);
            sb[0] = fma(N[0] * N[0], w2, sb[0]);
        } else if constexpr (S == 2) {
            sa[0] = fma(N[0], w1, sa[0]);
            sa[1] = fma(N[1], w1, sa[1]);
            const double t0 = N[0] * w2, t1 = N[1] * w2;
            sb[0] = fma(t0, N[0], sb[0]);
            sb[1] = fma(t0, N[1], sb[1]);
            sb[2] = fma(t1, N[1], sb[2]);
        }
    }
    template <int WPT>
    __device__ __forceinline__ void finish(const double (&nsig)[S][WPT], double (&acc)[WPT]) const {
        if constexpr (S == 1) {
#pragma unroll
            for (int j = 0; j < WPT; ++j) acc[j] += fma(nsig[0][j], fma(nsig[0][j], sb[0], sa[0]), sw);
        } else if constexpr (S == 2) {
#pragma unroll
            for (int j = 0; j < WPT; ++j) {
                const double a = nsig[0][j], b = nsig[1][j];
                double q = fma(a, sa[0], sw);
                q = fma(b, sa[1], q);
                q = fma(a * a, sb[0], q);
                q = fma(2.0 * a * b, sb[1], q);
                acc[j] += fma(b * b, sb[2], q);
            }
        }
    }
};

// Sigma-sorted windows (one absorber, pb200_sort.cu): a window's wavelengths have nearly the same sigma, so two more
// warp-uniform decisions pay.  (1) tau < 0.1 on the whole window: w exp(-tau) = w sum_k (v c1)^k / k!, v = N nsig_j,
// to order 8 (truncation 0.1^9/9! = 2.8e-15): per chord and THREAD the nine sums A_k = sum w (N c1)^k / k!
// (25 FP64 instructions) instead of 8 samples through the table exp; applied by Horner at the end.  (2) tau > 50 on
// the whole window: w exp(-tau) < 2e-22 w, the chord is skipped (the reference's own term is below 1e-21 of the sum).
constexpr double kThinHiV = 0.1 * kExpBigScale;
constexpr double kOpaqueV = 50.0 * kExpBigScale;
struct ThinHi {
    double a[9] = {};
    __device__ __forceinline__ void add(double w0, double N) { add_scaled(w0, N * kExpBigC1); }
    // a[k] += w0 t^k / k!   (two FP64 instructions per order: the running power and one FMA with 1/k!)
    __device__ __forceinline__ void add_scaled(double w0, double t) {
        double p = w0;
        a[0] += p;
        p *= t;  a[1] += p;
        p *= t;  a[2] = fma(p, 0.5, a[2]);
        p *= t;  a[3] = fma(p, 1.0 / 6.0, a[3]);
        p *= t;  a[4] = fma(p, 1.0 / 24.0, a[4]);
        p *= t;  a[5] = fma(p, 1.0 / 120.0, a[5]);
        p *= t;  a[6] = fma(p, 1.0 / 720.0, a[6]);
        p *= t;  a[7] = fma(p, 1.0 / 5040.0, a[7]);
        p *= t;  a[8] = fma(p, 1.0 / 40320.0, a[8]);
    }
    __device__ __forceinline__ double value(double n) const {
        double q = fma(a[8], n, a[7]);
        q = fma(q, n, a[6]); q = fma(q, n, a[5]); q = fma(q, n, a[4]); q = fma(q, n, a[3]);
        q = fma(q, n, a[2]); q = fma(q, n, a[1]); q = fma(q, n, a[0]);
        return q;
    }
    template <int WPT>
    __device__ __forceinline__ void finish(const double (&nsig)[WPT], double (&acc)[WPT]) const {
#pragma unroll
        for (int j = 0; j < WPT; ++j) acc[j] += value(nsig[j]);
    }
};

template <int S, int WPT, bool CLAMP, int NWARPS, bool HI>
__device__ __forceinline__ void transit_tile(uint32_t stage, const double (&nsig)[S][WPT], double (&acc)[WPT],
                                             uint32_t tab_base, const double (&smw)[S], ThinSums<S>& thin, int wid, double thin_v,
                                             double hi_v, double smin0, ThinHi& hi) {
    constexpr int R = pb_rec_stride(S);
    constexpr uint32_t kRecBytes = R * sizeof(double);
    // the warps of the CTA share the window (same wavelengths) and split the chords: warp w takes chords
    // w, w + NWARPS, ... of the tile
#pragma unroll 2
    for (int i = 0; i < kChordTile / NWARPS; ++i) {
        const uint32_t rec = stage + (uint32_t)(i * NWARPS + wid) * kRecBytes;
        double d[R], N[S];
#pragma unroll
        for (int i = 0; i < R / 2; ++i) lds_v2f64(rec + 16 * i, d[2 * i], d[2 * i + 1]);   // broadcast LDS.128
        const double w0 = d[0], w1 = d[1], w2 = d[2];
#pragma unroll
        for (int a = 0; a < S; ++a) N[a] = d[3 + a];
        double bound = N[0] * smw[0];
#pragma unroll
        for (int a = 1; a < S; ++a) bound = fma(N[a], smw[a], bound);
        if (bound < thin_v) {   // warp-uniform
            if constexpr (S <= 2) {
                thin.add(w0, w1, w2, N);
            } else {
#pragma unroll
                for (int j = 0; j < WPT; ++j) {
                    double v = N[0] * nsig[0][j];
#pragma unroll
                    for (int a = 1; a < S; ++a) v = fma(N[a], nsig[a][j], v);
                    acc[j] += fma(fma(v, w2, w1), v, w0);
                }
            }
            continue;
        }
        if constexpr (HI) {
            if (bound < hi_v) { hi.add(w0, N[0]); continue; }   // warp-uniform
            if (N[0] * smin0 > kOpaqueV) continue;              // opaque over the whole window
        }
        ++thin.thick;
#pragma unroll
        for (int j = 0; j < WPT; ++j) {
            double t, r;
            if constexpr (S == 1 && !CLAMP) {
                t = fma(N[0], nsig[0][j], kExpMagic);                              // -tau in scaled units, :954
                r = fma(N[0], nsig[0][j], -(t - kExpMagic));
            } else {
                double v = N[0] * nsig[0][j];
#pragma unroll
                for (int a = 1; a < S; ++a) v = fma(N[a], nsig[a][j], v);
                if (CLAMP) {   // tau <= 693.8 keeps 2^e normal; exp(-693.8) w is < 1e-301
                    const unsigned hi = min((unsigned)__double2hiint(v), kExpBigClampHi);
                    v = __hiloint2double((int)hi, __double2loint(v));
                }
                t = v + kExpMagic;
                r = v - (t - kExpMagic);
            }
            const double T = pb_exp_big_load(tab_base, __double2loint(t));
            double q = fma(r, w2, w1);
            q = fma(q, r, w0);                                                     // w exp(r ln2/4096)
            acc[j] = fma(T, q, acc[j]);                                            // :1241, :1244/:1255
        }
    }
}

// Tile step of the sigma-sorted variant (one absorber).  The per-chord decisions do not depend on the wavelength,
// so they are made LANE-PARALLEL: lane l classifies chord l * NWARPS + wid of the tile and, if the chord is thin,
// adds it to its own copies of the chord sums (ThinSums order 2 below 2.1e-5, ThinHi order 8 below 0.1; the lanes'
// copies are added once, after the last tile) -- 1/32 of the work of letting every thread repeat it.  Only the
// chords that need the table exp (ballot) are then walked by the whole warp, their records re-read as broadcasts.
template <int S, int WPT, bool CLAMP, int NWARPS, bool HI>
__device__ __forceinline__ void transit_tile_sorted(uint32_t stage, const double (&nsig)[S][WPT], double (&acc)[WPT],
                                                    uint32_t tab_base, const double (&smw)[S], ThinSums<S>& thin, int wid, int lane,
                                                    double thin_v, double hi_v, double smin0, ThinHi& hi) {
    static_assert(kChordTile / NWARPS == 32, "one chord per lane");
    static_assert(S <= 2 && (!HI || S == 1), "chord sums separate for one or two absorbers");
    constexpr int R = pb_rec_stride(S);
    constexpr uint32_t kRecBytes = R * sizeof(double);   // {w, w c1, w c2, N_0[, N_1, 0]}
    double d[R], N[S];
    {
        const uint32_t rec = stage + (uint32_t)(lane * NWARPS + wid) * kRecBytes;
#pragma unroll
        for (int i = 0; i < R / 2; ++i) lds_v2f64(rec + 16 * i, d[2 * i], d[2 * i + 1]);
    }
#pragma unroll
    for (int a = 0; a < S; ++a) N[a] = d[3 + a];
    double bound = N[0] * smw[0];
#pragma unroll
    for (int a = 1; a < S; ++a) bound = fma(N[a], smw[a], bound);
    bool need_exp = false;
    if (bound < thin_v) {
        thin.add(d[0], d[1], d[2], N);
    } else if (HI && bound < hi_v) {
        hi.add(d[0], N[0]);
    } else if (!(HI && N[0] * smin0 > kOpaqueV)) {
        need_exp = true;
    }
    unsigned mask = __ballot_sync(0xffffffffu, need_exp);
    thin.thick += (lane == 0) ? __popc(mask) : 0;
    while (mask) {
        const int l = __ffs(mask) - 1;
        mask &= mask - 1;
        const uint32_t rec = stage + (uint32_t)(l * NWARPS + wid) * kRecBytes;
        double b[R];
#pragma unroll
        for (int i = 0; i < R / 2; ++i) lds_v2f64(rec + 16 * i, b[2 * i], b[2 * i + 1]);   // broadcast
#pragma unroll
        for (int j = 0; j < WPT; ++j) {
            double t, r;
            if constexpr (S == 1 && !CLAMP) {
                t = fma(b[3], nsig[0][j], kExpMagic);                              // -tau in scaled units, :954
                r = fma(b[3], nsig[0][j], -(t - kExpMagic));
            } else {
                double v = b[3] * nsig[0][j];
#pragma unroll
                for (int a = 1; a < S; ++a) v = fma(b[3 + a], nsig[a][j], v);
                if (CLAMP) {   // tau <= 693.8 keeps 2^e normal
                    const unsigned hi32 = min((unsigned)__double2hiint(v), kExpBigClampHi);
                    v = __hiloint2double((int)hi32, __double2loint(v));
                }
                t = v + kExpMagic;
                r = v - (t - kExpMagic);
            }
            const double T = pb_exp_big_load(tab_base, __double2loint(t));
            double q = fma(r, b[2], b[1]);
            q = fma(q, r, b[0]);                                                   // w exp(r ln2/4096)
            acc[j] = fma(T, q, acc[j]);                                            // :1241, :1244/:1255
        }
    }
}

// NARROW sigma-sorted windows.  For one absorber R(p, w) depends on the wavelength only through sigma(p, w):
//     sum_m w_m exp(-N_m sigma) =: f_p(sigma),   a smooth function of ONE variable.
// In a sorted window the cross sections lie in [s_lo, s_hi]; with the centre s_c, the half width h and
// sigma_j = s_c + h x_j, |x_j| <= 1:
//     w exp(-N sigma_j) = [w exp(-N s_c)] * sum_k (-N h x_j)^k / k!  =  sum_k x_j^k * [E u^k / k!],  E = w exp(-N s_c), u = -N h,
// so the window needs per chord ONE table exp and the nine running sums A_k = sum_m E_m u_m^k / k! (lane-parallel, 26 FP64
// instructions per chord and window) instead of 256 samples through the exp (1536), and per wavelength a Horner
// evaluation at the end.  Truncation after k = 8 with tau_c = N s_c and eps = h / s_c:  the remainder of a chord is below
// w e^{-tau_c (1 - eps)} (tau_c eps)^9 / 9!  <=  w (9 eps / (e (1 - eps)))^9 / 9!   (maximum over tau_c, at tau_c = 9 / (1 - eps));
// eps <= 1/40 (kCentreEps) gives 6.3e-16 w if EVERY chord sat at that optical depth -- a few units in the last place of
// R for any distribution of column densities, so the test needs nothing but the window's own sigma range.  Opaque chords (tau > 50 on the whole window) are skipped as in
// the other sorted windows, which also keeps the table exp inside its unclamped range (tau_c <= 50 (1 + 2 eps)).
// The sums live in the registers of ThinHi (the two expansions never occur in the same window).
constexpr double kCentreEps = 1.0 / 40.0;
// GROUPS = false: the window as a whole is narrow; the warps split the chords of the tile (one chord per lane).
// GROUPS = true: the window is not, but each of its 32-wavelength groups is (line wings: sigma changes by a few per cent
// across 256 sorted wavelengths, by a few per mille across 32): warp g owns group g -- centre, half width and opaque
// limit of ITS wavelengths -- and walks all 256 chords of the tile, 8 per lane: 8 x 26 FP64 instructions per chord
// and window, still 7 times fewer than the table exp per sample.
template <int NWARPS, bool GROUPS>
__device__ __forceinline__ void transit_tile_centre(uint32_t stage, uint32_t tab_base, int wid, int lane, double nc, double hu,
                                                    double smin0, ThinHi& hi, int& n_centre) {
    static_assert(kChordTile / NWARPS == 32, "one chord per lane");
    constexpr uint32_t kRecBytes = pb_rec_stride(1) * sizeof(double);   // {w, w c1, w c2, N}
#pragma unroll 2
    for (int i = 0; i < (GROUPS ? NWARPS : 1); ++i) {
        const uint32_t rec = stage + (uint32_t)(lane * NWARPS + (GROUPS ? i : wid)) * kRecBytes;
        double w0, w1, w2, N;
        lds_v2f64(rec, w0, w1);
        lds_v2f64(rec + 16, w2, N);
        const bool live = !(N * smin0 > kOpaqueV);     // padding records have w = 0 and N = 0: they add exact zeros
        if (live) {
            const double t = fma(N, nc, kExpMagic);                                    // -tau_c in scaled units
            const double r = fma(N, nc, -(t - kExpMagic));
            const double T = pb_exp_big_load(tab_base, __double2loint(t));
            double q = fma(r, w2, w1);
            q = fma(q, r, w0);
            hi.add_scaled(T * q, N * hu);                                              // E = w exp(-tau_c), u = N h ln2/4096
        }
        const unsigned took = __ballot_sync(0xffffffffu, live);   // every lane votes (not inside the conditional below)
        n_centre += (lane == 0) ? __popc(took) : 0;
    }
}

// S atomic constituents (compile time), WPT wavelengths per thread, THREADS threads per CTA.
// One CTA = one wavelength window (32 * WPT consecutive wavelengths) x one phase x one chord-axis split.
// All warps of the CTA hold the SAME wavelengths (lane = fastest index, j = 32-wavelength group) and split
// the chords of every tile between them, so they see the same mix of thin and thick chords and stay in
// step; their partial sums are added in a fixed order at the end.  Windows are taken in decreasing order of
// optical thickness (compact_kernel: longest CTAs first).  A window that is thin for every chord of
// the phase (S <= 2) needs no chord loop at all: its sums are the per-phase totals of compact_kernel.
template <int S, int WPT, int THREADS, bool HI>
__global__ void __launch_bounds__(THREADS, (S == 1 && !HI) ? 3 : 2)
transit_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    static_assert(!HI || S == 1, "sigma-sorted windows are for one absorber");
    const DevScene& sc = *scp;
    // shared memory: exp table [4096] | stage 0 records | stage 1 records | per-warp sums [NWARPS][WIN]
    extern __shared__ __align__(128) double smem[];
    __shared__ uint64_t full_bar[2];
    constexpr int NWARPS = THREADS / 32, WIN = 32 * WPT;
    constexpr int R = pb_rec_stride(S);
    constexpr int kStageDoubles = R * kChordTile;
    constexpr uint32_t kStageBytes = kStageDoubles * sizeof(double);
    double* etab = smem;
    double* stages = smem + kExpBigEntries;
    double* red = stages + 2 * kStageDoubles;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int pi = blockIdx.y, p = prm.phase_lo + pi;
    const int split = blockIdx.z;
    // sorted wavelengths: window 0 holds the largest cross sections, the launch order is the thickness order already
    const int window = HI ? (int)blockIdx.x : __ldg(prm.order + (int64_t)pi * prm.n_win + blockIdx.x);
    const int64_t win_lo = (int64_t)window * WIN;                         // relative to wav_lo
    // this CTA's share of the phase's ACTIVE tiles (compact_kernel): even shares over the splits, at least
    // prm.min_tiles per working CTA; splits beyond the need have nothing to do
    const int n_act = sc.n_act[p];
    const int tps = pb_tiles_per_split(n_act, prm.n_split, prm.min_tiles);
    const int tile0 = split * tps;
    int n_tiles = n_act - tile0;
    if (n_tiles > tps) n_tiles = tps;
    if (n_tiles <= 0 && prm.n_split > 1) return;   // finalize_kernel skips this split (pb_splits_used)

    // ---- prologue: -sigma_a(p, w) 4096/ln2 of the window (every warp loads the same values) ---------
    double nsig[S][WPT];
    double acc[WPT];
    bool bad = false;   // NaN cross-section (NaN wavelength): the reference propagates NaN into R
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        acc[j] = 0.0;
        const int64_t w = win_lo + j * 32 + lane;
        const bool live = w < prm.n_wav_sel;
#pragma unroll
        for (int a = 0; a < S; ++a) {
            const double v = live ? __ldg(prm.nsig + ((int64_t)a * prm.n_phase + pi) * prm.n_wav_sel + w) : 0.0;
            bad |= (v != v);
            nsig[a][j] = v;
        }
    }
    // window-wide max |nsig_a| for the thin test (NaN -> +inf: never thin), the largest scaled tau of the
    // window on the densest chord of the phase, and from it: does any chord need the clamped exp
    // (tau > 690), does any chord need the exp at all
    double smw[S];
    double key = 0.0;
    const bool any_bad = __any_sync(0xffffffffu, bad);
#pragma unroll
    for (int a = 0; a < S; ++a) {
        double mx = 0.0;
#pragma unroll
        for (int j = 0; j < WPT; ++j) mx = fmax(mx, fabs(nsig[a][j]));
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
        smw[a] = any_bad ? __longlong_as_double(0x7ff0000000000000ll) : mx;
        key = fma(__longlong_as_double((long long)__ldg(sc.nmax + (int64_t)a * sc.n_orb + p)), smw[a], key);
    }
    const bool clamp = !(key < kExpBigClampDecision);   // NaN -> clamped variant (harmless)
    const bool all_thin = (S <= 2) && (key < prm.thin_v);
    ThinSums<S> thin;
    ThinHi hi;
    double smin0 = 0.0;   // smallest |nsig| of the window's live wavelengths (0 with a NaN: never opaque)
    if constexpr (HI) {
        double mn = __longlong_as_double(0x7ff0000000000000ll);
#pragma unroll
        for (int j = 0; j < WPT; ++j)
            if (win_lo + j * 32 + lane < prm.n_wav_sel) mn = fmin(mn, fabs(nsig[0][j]));
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, d));
        smin0 = any_bad ? 0.0 : mn;
    }
    const double hi_v = (prm.thin_v > 0.0) ? kThinHiV : 0.0;
    // narrow window (transit_tile_centre): centre and half width of the window's nsig (all <= 0 up to rounding of sigma = 0)
    bool narrow = false;
    double nc = 0.0, hw = 0.0;
    int n_centre = 0;
    bool groups = false;   // every 32-wavelength group is narrow about its own centre: warp g takes group g
    double smin_c = smin0;
    if constexpr (HI) {
        const double span = smw[0] - smin0, mid2 = smw[0] + smin0;
        narrow = prm.centre >= 1 && !any_bad && span <= kCentreEps * mid2 && !all_thin;   // inf / NaN compare false
        nc = -0.5 * mid2;
        hw = 0.5 * span;
        if (!narrow && prm.centre >= 2 && !any_bad && !all_thin) {
            static_assert(WPT == NWARPS, "one 32-wavelength group per warp");
            bool ok = true;
#pragma unroll
            for (int j = 0; j < WPT; ++j) {
                const double a = fabs(nsig[0][j]);
                double mx = a, mn = (win_lo + j * 32 + lane < prm.n_wav_sel) ? a : __longlong_as_double(0x7ff0000000000000ll);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, d));
                    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, d));
                }
                if (mn > mx) mn = mx;   // a group of the last window without live wavelengths (mx = 0): results never stored
                ok = ok && (mx - mn <= kCentreEps * (mx + mn));
                if (j == wid) { nc = -0.5 * (mx + mn); hw = 0.5 * (mx - mn); smin_c = mn; }
            }
            groups = narrow = ok;    // the same for every warp of the CTA (all of them looked at all eight groups)
            if (!ok) { nc = -0.5 * mid2; hw = 0.5 * span; smin_c = smin0; }
        }
    }

    if (n_tiles > 0 && !all_thin) {
        if (tid == 0) {
            mbar_init(&full_bar[0], 1);
            mbar_init(&full_bar[1], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        const double* rec_row = sc.rec + ((int64_t)p * sc.Mp + (int64_t)tile0 * kChordTile) * R;
        auto issue = [&](int tile, int stage) {
            mbar_expect_tx(&full_bar[stage], kStageBytes);
            tma_bulk_g2s(stages + stage * kStageDoubles, rec_row + (int64_t)tile * kStageDoubles, kStageBytes,
                         &full_bar[stage]);
        };
        if (tid == 0) issue(0, 0);
        {   // the exp table: 32 KB from L2, 16-byte loads
            const double2* src = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries);
            double2* dst = reinterpret_cast<double2*>(etab);
            for (int i = tid; i < kExpBigEntries / 2; i += THREADS) dst[i] = __ldg(src + i);
        }
        __syncthreads();   // exp table complete
        // laundered through an opaque mov so that the address lives in a register instead of being
        // re-derived (S2UR CgaCtaId + ULEA) for every chord
        uint32_t tab_base;
        asm volatile("mov.u32 %0, %1;" : "=r"(tab_base) : "r"(smem_u32(etab)));
        const uint32_t stages_addr = smem_u32(stages);
        for (int t = 0; t < n_tiles; ++t) {
            const int stage = t & 1;
            if (tid == 0 && t + 1 < n_tiles) issue(t + 1, stage ^ 1);
            mbar_wait(&full_bar[stage], (t >> 1) & 1);
            const uint32_t sw = stages_addr + stage * kStageBytes;
            if (HI && narrow) {
                if constexpr (HI) {
                    if (groups) transit_tile_centre<NWARPS, true>(sw, tab_base, wid, lane, nc, hw * kExpBigC1, smin_c, hi, n_centre);
                    else transit_tile_centre<NWARPS, false>(sw, tab_base, wid, lane, nc, hw * kExpBigC1, smin_c, hi, n_centre);
                }
            } else if constexpr (S <= 2) {   // one or two absorbers: lane-parallel per-chord decisions (sorted or not)
                if (clamp) transit_tile_sorted<S, WPT, true, NWARPS, HI>(sw, nsig, acc, tab_base, smw, thin, wid, lane, prm.thin_v, hi_v, smin0, hi);
                else transit_tile_sorted<S, WPT, false, NWARPS, HI>(sw, nsig, acc, tab_base, smw, thin, wid, lane, prm.thin_v, hi_v, smin0, hi);
            } else {
                if (clamp) transit_tile<S, WPT, true, NWARPS, HI>(sw, nsig, acc, tab_base, smw, thin, wid, prm.thin_v, hi_v, smin0, hi);
                else transit_tile<S, WPT, false, NWARPS, HI>(sw, nsig, acc, tab_base, smw, thin, wid, prm.thin_v, hi_v, smin0, hi);
            }
            __syncthreads();   // stage may be refilled by the producer in the next iteration
        }
        if constexpr (S <= 2) {   // the lanes' copies of the chord sums -> warp totals (fixed order)
            thin.sw = pb_warp_sum(thin.sw);
#pragma unroll
            for (int a = 0; a < S; ++a) thin.sa[a] = pb_warp_sum(thin.sa[a]);
#pragma unroll
            for (int a = 0; a < (S == 1 ? 1 : 3); ++a) thin.sb[a] = pb_warp_sum(thin.sb[a]);
            if constexpr (HI) {
#pragma unroll
                for (int k = 0; k < 9; ++k) hi.a[k] = pb_warp_sum(hi.a[k]);
            }
        }
        thin.finish(nsig, acc);
        if constexpr (HI) {
            if (narrow) {   // Horner in x_j = (nsig_j - centre) / half width
                const double inv_h = hw > 0.0 ? 1.0 / hw : 0.0;
                double xs[WPT];
#pragma unroll
                for (int j = 0; j < WPT; ++j) xs[j] = (nsig[0][j] - nc) * inv_h;
                if (groups) {   // this warp's sums belong to its own group of wavelengths only
                    double x = 0.0;
#pragma unroll
                    for (int j = 0; j < WPT; ++j)
                        if (j == wid) x = xs[j];
                    const double q = hi.value(x);
#pragma unroll
                    for (int j = 0; j < WPT; ++j)
                        if (j == wid) acc[j] += q;
                } else {
                    hi.finish(xs, acc);
                }
                if (lane == 0 && n_centre) atomicAdd(sc.stat + 4, (unsigned long long)n_centre);
            } else {
                hi.finish(nsig[0], acc);
            }
        }
        if (lane == 0 && thin.thick) atomicAdd(sc.stat + 2, (unsigned long long)thin.thick);
    } else if (n_tiles > 0 && wid == 0 && split == 0) {
        // every chord of the phase is thin in this window: the chord sums are the phase totals
        if constexpr (S <= 2) {
            const double* tt = sc.thin_tot + (int64_t)p * 8;
            thin.sw = tt[0];
            thin.sa[0] = tt[1];
            if constexpr (S == 2) { thin.sa[1] = tt[2]; thin.sb[0] = tt[3]; thin.sb[1] = tt[4]; thin.sb[2] = tt[5]; }
            else thin.sb[0] = tt[3];
            thin.finish(nsig, acc);
        }
    }

    // ---- epilogue: add the warps' sums in a fixed order, one thread per wavelength --------------------
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        double v = acc[j];
        if (any_bad) {
#pragma unroll
            for (int a = 0; a < S; ++a)
                if (nsig[a][j] != nsig[a][j]) v = nsig[a][j];
        }
        red[wid * WIN + j * 32 + lane] = v;
    }
    __syncthreads();
    for (int t = tid; t < WIN; t += THREADS) {
        const int64_t w = win_lo + t;
        if (w >= prm.n_wav_sel) continue;
        double v = red[t];
#pragma unroll
        for (int k = 1; k < NWARPS; ++k) v += red[k * WIN + t];
        const int64_t o = (int64_t)pi * prm.n_wav_sel + w;
        const int64_t wl = HI ? (int64_t)__ldg(prm.perm + o) : w;   // sorted position -> local wavelength index
        if (prm.n_split == 1) prm.R[pb_out_index(prm, pi, wl)] = (v + sc.w0[p]) / sc.fout[0];           // :1258
        else prm.partial[(int64_t)split * prm.n_phase * prm.n_wav_sel + o] = v;
    }
}

// R = (sum over splits, fixed order) / F_out
__global__ void finalize_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    const int64_t n = (int64_t)prm.n_phase * prm.n_wav_sel;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int p = prm.phase_lo + (int)(i / prm.n_wav_sel);
        double s = prm.culled ? sc.w0[p] : 0.0;                                 // transparent chords
        const int used = prm.culled ? pb_splits_used(sc.n_act[p], prm.n_split, prm.min_tiles) : prm.n_split;
        for (int k = 0; k < used; ++k) s += prm.partial[(int64_t)k * n + i];
        const double fout = sc.fout[sc.fout_per_wav ? prm.wav_lo + i % prm.n_wav_sel : 0];
        const int64_t wl = prm.perm ? (int64_t)__ldg(prm.perm + i) : i % prm.n_wav_sel;   // partial sums are in sorted order
        prm.R[pb_out_index(prm, (int)(i / prm.n_wav_sel), wl)] = s / fout;
    }
}

// =====================================================================================================================
// Sigma cells (one absorber): R(p, w) = f_p(sigma(p, w)) with f_p(s) = sum_m w_m exp(-N_m s), a smooth function of the
// cross section ALONE -- so it is tabulated once per phase on a geometric grid of sigma cells and every wavelength only
// evaluates a polynomial.  Cell k holds the scaled cross sections a = |nsig| in [g^k, g^(k+1)), g = (1 + eps)/(1 - eps),
// eps = 3/40 (pb200_internal.cuh): about its centre a_c = (g^k + g^(k+1))/2 with half width h = eps a_c,
//     w exp(-N a) = sum_k x^k [E u^k / k!],   E = w exp(-N a_c), u = -N h, x = (a - a_c)/h in [-1, 1],
// truncated after k = 12: with tau_c = N a_c a chord's remainder is below w e^{-tau_c (1 - eps)} (tau_c eps)^13 / 13!
// <= w (13 eps / (e (1 - eps)))^13 / 13! = 7.2e-16 w WHATEVER its optical depth (the maximum sits at tau_c = 13 / (1 - eps)).
// (Order 8 needs eps = 1/40 for the same bound: three times the cells at 25 instead of 33 FP64 instructions per chord
// and cell -- 2.3 times the work.)  cell_moment_kernel forms the 13 sums A_k = sum_m E_m u_m^k / k! of every cell (ONE table
// exp per chord and cell; lane = cell, the warps of a CTA and the chord-axis splits share the chords) and the last split
// to arrive adds the splits in a fixed order; cell_eval_kernel applies Horner per wavelength.  The work no longer depends
// on the number of wavelengths: C5 has 300 306 per phase but ~215 cells between "every chord thin" (N_max a < 2.1e-5
// scaled: the order-2 totals of compact_kernel, which also leaves the cell range of the phase) and its largest cross
// section.  No sort, no permutation.
// =====================================================================================================================
// 1/k!: in the constant bank, so that they are operands of the FMAs instead of immediates rebuilt (UMOV pairs) in every iteration
static __constant__ double pb_cell_inv_fact[kCellSums] = {1.0, 1.0, 0.5, 0.16666666666666666, 0.041666666666666664, 0.008333333333333333,
                                                          0.001388888888888889, 0.0001984126984126984, 2.48015873015873e-05,
                                                          2.7557319223985893e-06, 2.755731922398589e-07, 2.505210838544172e-08,
                                                          2.08767569878681e-09};
struct CellSums {
    double a[kCellSums] = {};
    // a[k] += E u^k / k!: the running power and one FMA with 1/k! per order
    __device__ __forceinline__ void add(double E, double u) {
        const double* __restrict__ inv_fact = pb_cell_inv_fact;
        double p = E;
        a[0] += p;
        p *= u;
        a[1] += p;
#pragma unroll
        for (int k = 2; k < kCellSums; ++k) {
            p *= u;
            a[k] = fma(p, inv_fact[k], a[k]);
        }
    }
};

// partial[((pi * n_cs + split) * kCellSums + k) * kCellCap + cell];  sums[(pi * (kCellSums + 2) + k) * kCellCap + cell], rows
// kCellSums, kCellSums + 1 of sums: the cell's centre (of nsig, <= 0) and 1 / half width for cell_eval_kernel
__global__ void __launch_bounds__(256, 3)
cell_moment_kernel(const DevScene* __restrict__ scp, const TransitParams prm, double* __restrict__ partial, double* __restrict__ sums, int n_cs) {
    const DevScene& sc = *scp;
    extern __shared__ __align__(128) double smem[];
    __shared__ uint64_t full_bar[2];
    __shared__ int s_last;
    constexpr int NWARPS = 8;
    constexpr int R = pb_rec_stride(1);
    constexpr int kStageDoubles = R * kChordTile;
    constexpr uint32_t kStageBytes = kStageDoubles * sizeof(double);
    constexpr uint32_t kRecBytes = R * sizeof(double);
    double* etab = smem;
    double* stages = smem + kExpBigEntries;
    double* red = etab;                            // [NWARPS][kCellSums][32]: after the chord loop the table's 32 KB are free
    static_assert(NWARPS * kCellSums * 32 <= kExpBigEntries, "the warps' sums fit the exp table's memory");
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int pi = blockIdx.y, p = prm.phase_lo + pi, split = blockIdx.z;
    const int2 rg = prm.cell_range[pi];
    if ((int)blockIdx.x * 32 >= rg.y) return;      // no such cells in this phase (the same for the whole CTA)
    const int cell = blockIdx.x * 32 + lane;
    const int n_act = sc.n_act[p];
    const int tps = pb_tiles_per_split(n_act, n_cs, 1);
    const int tile0 = split * tps;
    int n_tiles = n_act - tile0;
    if (n_tiles > tps) n_tiles = tps;
    const double e_lo = pb_cell_edge(rg.x + cell), e_hi = pb_cell_edge(rg.x + cell + 1);
    const double nc = -0.5 * (e_lo + e_hi);                           // centre of the cell's nsig (<= 0)
    const double hw = 0.5 * (e_hi - e_lo);

    CellSums cs;
    int n_live = 0;
    if (n_tiles > 0) {
        const double hu = hw * kExpBigC1;                             // half width as an exponent
        if (tid == 0) {
            mbar_init(&full_bar[0], 1);
            mbar_init(&full_bar[1], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        const double* rec_row = sc.rec + ((int64_t)p * sc.Mp + (int64_t)tile0 * kChordTile) * R;
        auto issue = [&](int tile, int stage) {
            mbar_expect_tx(&full_bar[stage], kStageBytes);
            tma_bulk_g2s(stages + stage * kStageDoubles, rec_row + (int64_t)tile * kStageDoubles, kStageBytes, &full_bar[stage]);
        };
        if (tid == 0) issue(0, 0);
        {
            const double2* src = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries);
            double2* dst = reinterpret_cast<double2*>(etab);
            for (int i = tid; i < kExpBigEntries / 2; i += 256) dst[i] = __ldg(src + i);
        }
        __syncthreads();
        uint32_t tab_base;
        asm volatile("mov.u32 %0, %1;" : "=r"(tab_base) : "r"(smem_u32(etab)));
        const uint32_t stages_addr = smem_u32(stages);
        for (int t = 0; t < n_tiles; ++t) {
            const int stage = t & 1;
            if (tid == 0 && t + 1 < n_tiles) issue(t + 1, stage ^ 1);
            mbar_wait(&full_bar[stage], (t >> 1) & 1);
            const uint32_t sw = stages_addr + stage * kStageBytes;
#pragma unroll 2
            for (int i = 0; i < kChordTile / NWARPS; ++i) {            // warp w takes records w, w + 8, ...: every lane the same record
                const uint32_t rec = sw + (uint32_t)(i * NWARPS + wid) * kRecBytes;
                double w0, w1, w2, N;
                lds_v2f64(rec, w0, w1);                                 // broadcast
                lds_v2f64(rec + 16, w2, N);
                const bool live = !(N * e_lo > kOpaqueV) && w0 != 0.0; // opaque over the whole cell / padding record
                if (live) {
                    const double tt = fma(N, nc, kExpMagic);            // -tau at the centre, scaled
                    const double r = fma(N, nc, -(tt - kExpMagic));
                    const double T = pb_exp_big_load(tab_base, __double2loint(tt));
                    double q = fma(r, w2, w1);
                    q = fma(q, r, w0);
                    cs.add(T * q, N * hu);
                    ++n_live;
                }
            }
            __syncthreads();   // stage may be refilled by the producer in the next iteration
        }
    }
    // the warps' sums in a fixed order; a split without tiles writes zeros (the reduction below adds every split)
#pragma unroll
    for (int k = 0; k < kCellSums; ++k) red[(wid * kCellSums + k) * 32 + lane] = cs.a[k];
    __syncthreads();
    for (int i = tid; i < kCellSums * 32; i += 256) {
        double v = red[i];
#pragma unroll
        for (int w = 1; w < NWARPS; ++w) v += red[w * kCellSums * 32 + i];
        const int k = i >> 5, l = i & 31;
        partial[(((int64_t)pi * n_cs + split) * kCellSums + k) * kCellCap + blockIdx.x * 32 + l] = v;
    }
    n_live = (int)pb_warp_sum((double)n_live);
    if (lane == 0 && n_live) atomicAdd(sc.stat + 4, (unsigned long long)n_live);
    // the last split of this (phase, cell block) to arrive adds all of them, always in split order: deterministic
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(prm.cell_count + pi * (kCellCap / 32) + blockIdx.x, 1u) == (unsigned)(n_cs - 1));
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double* out = sums + (int64_t)pi * (kCellSums + 2) * kCellCap + blockIdx.x * 32;
    for (int i = tid; i < kCellSums * 32; i += 256) {
        const int k = i >> 5, l = i & 31;
        double v = 0.0;
        for (int s = 0; s < n_cs; ++s) v += __ldcg(partial + (((int64_t)pi * n_cs + s) * kCellSums + k) * kCellCap + blockIdx.x * 32 + l);
        out[k * kCellCap + l] = v;
    }
    if (wid == 0) {
        out[kCellSums * kCellCap + lane] = nc;
        out[(kCellSums + 1) * kCellCap + lane] = 1.0 / hw;
    }
}

__global__ void __launch_bounds__(256) cell_eval_kernel(const DevScene* __restrict__ scp, const TransitParams prm, const double* __restrict__ sums) {
    const DevScene& sc = *scp;
    const int64_t wi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (wi >= prm.n_wav_sel) return;
    const int pi = blockIdx.y, p = prm.phase_lo + pi;
    // (b1) -sigma(p, lambda) 4096/ln2 of the one absorber, as sigma_phase_kernel forms it (np.interp + 10**v - offset, :34-51,
    // chi :940): looked up here, where it is used -- no nsig array, no kernel beside the column stage
    double ns;
    {
        const double lam = __ldg(sc.wavelength + pb_wav_index(prm, wi));
        const int k0 = sc.abs_first[0];
        const double sl = __ldg(sc.scen[sc.cons[k0].scenario].shift + p) * lam;                 // :917
        double v = 0.0;
        for (int k = k0; k < k0 + sc.abs_count[0]; ++k) {
            const DevConstituent& C = sc.cons[k];
            v = fma(C.chi, exp10(pb_cons_log_sigma(C, sl)) - C.offset, v);
        }
        ns = -v * kExpBigScale;
    }
    const double a = fabs(ns);
    const int2 rg = prm.cell_range[pi];
    double v;
    int k = 0;
    if (!(a < __longlong_as_double(0x7ff0000000000000ll))) {
        v = ns * 0.0;                                                   // NaN (or infinite) cross section: NaN in R, as the reference
    } else if (rg.y == 0 || (k = pb_cell_index(a) - rg.x) < 0 || !(a * __longlong_as_double((long long)sc.nmax[p]) >= prm.thin_v)) {
        const double* tt = sc.thin_tot + (int64_t)p * 8;                // thin for every chord of the phase: order-2 totals
        v = fma(ns, fma(ns, tt[3], tt[1]), tt[0]);
    } else if (k < kCellCap) {
        const double* A = sums + (int64_t)pi * (kCellSums + 2) * kCellCap + k;
        const double x = (ns - A[kCellSums * kCellCap]) * A[(kCellSums + 1) * kCellCap];
        double q = A[kCellOrder * kCellCap];
#pragma unroll
        for (int j = kCellOrder - 1; j >= 0; --j) q = fma(q, x, A[j * kCellCap]);
        v = q;
    } else {                                                            // beyond the cell arrays (tau spans > 33 decades): exact sum
        v = 0.0;
        const double* rec = sc.rec + (int64_t)p * sc.Mp * pb_rec_stride(1);
        const int64_t n_rec = (int64_t)sc.n_act[p] * kChordTile;
        for (int64_t m = 0; m < n_rec; ++m) v = fma(rec[m * 4], exp(rec[m * 4 + 3] * ns * kExpBigC1), v);
    }
    prm.R[pb_out_index(prm, pi, wi)] = (v + sc.w0[p]) / sc.fout[0];    // :1258
}

int cell_chord_splits(int sm_count, int n_phase, int n_tiles) {
    int n_cs = (sm_count * 16 + 7 * n_phase - 1) / (7 * n_phase);       // ~7 cell blocks per phase are in use (~215 cells)
    if (n_cs > n_tiles / 2) n_cs = n_tiles / 2;
    return n_cs < 1 ? 1 : n_cs;
}
static size_t cell_round(size_t b) { return (b + 255) & ~size_t(255); }
size_t cell_scratch_bytes(int n_phase, int n_cs) {
    return cell_round((size_t)n_phase * sizeof(int2)) + cell_round((size_t)n_phase * (kCellCap / 32) * sizeof(unsigned)) +
           (size_t)n_phase * ((size_t)n_cs * kCellSums + kCellSums + 2) * kCellCap * sizeof(double);
}
// where compact_kernel leaves the cell ranges / finds the counters to clear: set before launch_compact
void cell_bind(TransitParams& prm, void* scratch) {
    prm.cell_range = reinterpret_cast<int2*>(scratch);
    prm.cell_count = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(scratch) + cell_round((size_t)prm.n_phase * sizeof(int2)));
}
static double* cell_partial(const TransitParams& prm, void* scratch) {
    return reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + cell_round((size_t)prm.n_phase * sizeof(int2)) +
                                     cell_round((size_t)prm.n_phase * (kCellCap / 32) * sizeof(unsigned)));
}
// the records, thin totals, w0 and cell ranges of the run (compaction) -> the cells' Taylor sums; does not need nsig
cudaError_t launch_cell_moments(const DevScene* d_scene, const TransitParams& prm, int n_cs, void* scratch, cudaStream_t st) {
    double* partial = cell_partial(prm, scratch);
    double* sums = partial + (size_t)prm.n_phase * n_cs * kCellSums * kCellCap;
    const size_t smem = (kExpBigEntries + 2 * pb_rec_stride(1) * kChordTile) * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(cell_moment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cell_moment_kernel<<<dim3(kCellCap / 32, (unsigned)prm.n_phase, (unsigned)n_cs), 256, smem, st>>>(d_scene, prm, partial, sums, n_cs);
    return cudaGetLastError();
}
// sigma(p, lambda) looked up per wavelength + the sums -> R
cudaError_t launch_cell_eval(const DevScene* d_scene, const TransitParams& prm, int n_cs, void* scratch, cudaStream_t st) {
    const double* sums = cell_partial(prm, scratch) + (size_t)prm.n_phase * n_cs * kCellSums * kCellCap;
    cell_eval_kernel<<<dim3((unsigned)((prm.n_wav_sel + 255) / 256), (unsigned)prm.n_phase), 256, 0, st>>>(d_scene, prm, sums);
    return cudaGetLastError();
}

template <int S, int WPT, int THREADS = 256, bool HI = false>
static cudaError_t launch_one(const DevScene* d_scene, const TransitParams& prm, cudaStream_t st) {
    const size_t smem = (kExpBigEntries + 2 * pb_rec_stride(S) * kChordTile + (THREADS / 32) * 32 * WPT) * sizeof(double);
    auto kern = transit_kernel<S, WPT, THREADS, HI>;
    {   // dynamic + static shared memory is always above the 48 KB default limit
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    if (prm.win != 32 * WPT) return cudaErrorInvalidValue;
    dim3 grid((unsigned)prm.n_win, (unsigned)prm.n_phase, (unsigned)prm.n_split);
    kern<<<grid, THREADS, smem, st>>>(d_scene, prm);
    return cudaGetLastError();
}

// bucket index of every atomic table: with bfun(x) = clamp(int((x - x0) * inv), 0, n_bucket - 1) -- the SAME expression the
// lookup evaluates -- bucket[b] = smallest j with bfun(tab_x[j]) >= b (n_tab if none), b = 0 .. n_bucket.  One thread per
// table entry fills the buckets between its predecessor's and its own: no search, no dependent loads.
__global__ void __launch_bounds__(256) bucket_build_kernel(const DevScene* __restrict__ scp) {
    const DevScene& sc = *scp;
    const int lane = threadIdx.x & 31;
    for (int k = 0; k < sc.n_cons; ++k) {   // largest log10(sigma + offset) of every atomic table (DevScene::tab_ymax)
        const DevConstituent& C = sc.cons[k];
        if (C.kind != 0) continue;
        unsigned long long mx = 0ull;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < C.n_tab; i += (int64_t)gridDim.x * blockDim.x) {
            const double y = __ldg(C.tab_y + i);
            const unsigned long long b = (y == y) ? pb_ordered_bits(y) : ~0ull;   // NaN in a table: no bound
            mx = b > mx ? b : mx;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, mx, d); mx = o > mx ? o : mx; }
        if (lane == 0 && mx) atomicMax(sc.tab_ymax + k, mx);
    }
    for (int k = 0; k < sc.n_cons; ++k) {
        const DevConstituent& C = sc.cons[k];
        if (C.kind != 0 || !C.bucket) continue;
        int* out = const_cast<int*>(C.bucket);
        const int64_t n = C.n_tab, n_round = (n + 31) & ~int64_t(31);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += (int64_t)gridDim.x * blockDim.x) {
            int b0 = 1, b1 = 0;   // buckets (b0 .. b1) get the value i: those above the predecessor's bucket up to this entry's
            if (i < n) {
                b1 = pb_bucket_of(__ldg(C.tab_x + i), C.bucket_x0, C.bucket_inv, C.n_bucket);
                b0 = (i > 0 ? pb_bucket_of(__ldg(C.tab_x + i - 1), C.bucket_x0, C.bucket_inv, C.n_bucket) : -1) + 1;
            }
            const bool big = (b1 - b0) >= 4;   // a gap of the coarse part of the grid: filled by the whole warp
            if (!big)
                for (int b = b0; b <= b1; ++b) out[b] = (int)i;
            unsigned mask = __ballot_sync(0xffffffffu, big);
            while (mask) {
                const int l = __ffs(mask) - 1;
                mask &= mask - 1;
                const int g0 = __shfl_sync(0xffffffffu, b0, l), g1 = __shfl_sync(0xffffffffu, b1, l);
                const int gi = (int)__shfl_sync(0xffffffffu, (int)i, l);
                for (int b = g0 + lane; b <= g1; b += 32) out[b] = gi;
            }
            if (i == n - 1)   // beyond the last entry's bucket
                for (int b = b1 + 1; b <= C.n_bucket; ++b) out[b] = (int)n;
        }
    }
}
void launch_bucket_build(const DevScene* d_scene, int sm_count, cudaStream_t st) {
    bucket_build_kernel<<<sm_count * 8, 256, 0, st>>>(d_scene);
}

void launch_sigma_phase(const DevScene* d_scene, const TransitParams& prm, cudaStream_t st) {
    dim3 grid((unsigned)((prm.n_wav_sel + 255) / 256), (unsigned)prm.n_phase);
    sigma_phase_kernel<<<grid, 256, 0, st>>>(d_scene, prm);
}

// CTA shape: 256 threads x WPT wavelengths, WPT = 8 for 1-2 absorbers, 4 for 3-4, 2 above (register
// budget: 2 S WPT + temporaries must leave >= 2 CTAs per SM).
// returns cudaSuccess, or cudaErrorInvalidValue when n_atoms is outside the compiled range
cudaError_t launch_transit(const DevScene* d_scene, int n_atoms, const TransitParams& prm, cudaStream_t st) {
    switch (n_atoms) {
        case 1: return prm.perm ? launch_one<1, 8, 256, true>(d_scene, prm, st) : launch_one<1, 8>(d_scene, prm, st);
        case 2: return launch_one<2, 8>(d_scene, prm, st);
        case 3: return launch_one<3, 4>(d_scene, prm, st);
        case 4: return launch_one<4, 4>(d_scene, prm, st);
        case 5: return launch_one<5, 2>(d_scene, prm, st);
        case 6: return launch_one<6, 2>(d_scene, prm, st);
        case 7: return launch_one<7, 2>(d_scene, prm, st);
        case 8: return launch_one<8, 2>(d_scene, prm, st);
        case 9: return launch_one<9, 2>(d_scene, prm, st);
        case 10: return launch_one<10, 2>(d_scene, prm, st);
        case 11: return launch_one<11, 2>(d_scene, prm, st);
        case 12: return launch_one<12, 2>(d_scene, prm, st);
        default: return cudaErrorInvalidValue;
    }
}

void launch_finalize(const DevScene* d_scene, const TransitParams& prm, int sm_count, cudaStream_t st) {
    const int64_t n = (int64_t)prm.n_phase * prm.n_wav_sel;
    int64_t blocks = (n + 255) / 256;
    if (blocks > (int64_t)sm_count * 8) blocks = (int64_t)sm_count * 8;
    finalize_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_scene, prm);
}

}  // namespace pb200
