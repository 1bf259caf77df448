// pb200_sigma.cu -- kernels (b1)/(b2): cross-section lookup and Voigt line synthesis, the
// materialised optical depth, and the FP64 machine probes.  SURVEY 8(a) rows a8-a11.
#include "pb200_internal.cuh"
#include "pb200_molecule.cuh"

namespace pb200 {

// ---- (b1) n_interp_log (gasProperties.py:34-51): out = 10**interp(wl) - offset ------------------
__global__ void sigma_lookup_kernel(const double* __restrict__ tab_x, const double* __restrict__ tab_y,
                                    int64_t n_tab, double offset, const double* __restrict__ wl, int64_t n,
                                    double* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = pb_interp_log(tab_x, tab_y, n_tab, wl[i], offset);
}

// sigma_u[a][p][w] = sigma_a(shift_a(p) * lambda_w): the reference's cached per-unique-shift table
// (gasProperties.py:941-953), materialised only for pb200_optical_depth.
__global__ void sigma_u_kernel(const DevScene* __restrict__ scp, int64_t wav_lo, int64_t n_wav_sel,
                               double* __restrict__ out) {
    const DevScene& sc = *scp;
    const int a = blockIdx.z, p = blockIdx.y;
    int k = 0;
    for (; k < sc.n_cons; ++k)
        if (sc.cons[k].kind == 0 && sc.cons[k].atom_index == a) break;
    const DevConstituent& C = sc.cons[k];
    const double shift = sc.scen[C.scenario].shift[p];
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_wav_sel; w += (int64_t)gridDim.x * blockDim.x)
        out[((int64_t)a * sc.n_orb + p) * n_wav_sel + w] =
            pb_interp_log(C.tab_x, C.tab_y, C.n_tab, shift * sc.wavelength[wav_lo + w], C.offset);
}

// ---- (b2) Voigt synthesis: sigma(wl) = sum_lines K f V(c/wl - c/wl0; sigma_v/wl0, gamma) -------
// A thread owns kVoigtWpt refined wavelengths and keeps their line sums in registers; line parameters
// are staged through shared memory in chunks as three 16-byte records per line (broadcast LDS.128),
// one coalesced store per wavelength.  (gasProperties.py:672-692)
//
// Almost every (wavelength, line) pair is far out in the Lorentzian wing (|z|^2 >= 1e8: the lines
// span thousands of Angstrom, a Doppler width is ~0.01 A), where the asymptotic series of w(z) needs
// two terms:  Re w = y/(sqrt(pi) |z|^2) * (1 + (1.5 x^2 - 0.5 y^2)/|z|^4)  (next term 0.75/|z|^4 < 1e-16).
// That branch is 13 FP64 instructions (reciprocal: MUFU.RCP64H seed + two Newton steps); everything
// nearer goes through the general pb_faddeeva_re.  FP64-pipe-bound.
constexpr int kLineChunk = 512;
constexpr int kVoigtWpt = 2;
constexpr int kVoigtThreads = 128;

__device__ __forceinline__ double pb_rcp(double a) {   // 1/a for normal a > 0, <= 1 ulp
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    double e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    return fma(r, e, r);
}

__global__ void __launch_bounds__(kVoigtThreads)
voigt_sigma_kernel(const double* __restrict__ line_wl, const double* __restrict__ line_gamma,
                   const double* __restrict__ line_f, int64_t n_lines, double sigma_v,
                   const double* __restrict__ wl, int64_t n, double offset, double* __restrict__ out_sigma,
                   double* __restrict__ out_log10) {
    // per line: {nu0, 1/(sigma_nu sqrt2)} {-y^2/2, amp*y/sqrt(pi)} {y, amp},  amp = K f /(sigma_nu sqrt(2 pi))
    __shared__ double2 s_a[kLineChunk], s_b[kLineChunk], s_c[kLineChunk];
    const double kInvSqrt2 = 0.7071067811865476, kSqrt2Pi = 2.5066282746310002, kInvSqrtPi = 0.5641895835477563;
    // np.pi * e**2 / (m_e * c), evaluated left to right as the reference does (:689-690)
    const double K = (kPi * (kElemCharge * kElemCharge)) / (kElectronMass * kLightSpeed);
    const int64_t i0 = (int64_t)blockIdx.x * (kVoigtThreads * kVoigtWpt) + threadIdx.x;
    double nu[kVoigtWpt], acc[kVoigtWpt];
#pragma unroll
    for (int j = 0; j < kVoigtWpt; ++j) {
        const int64_t i = i0 + (int64_t)j * kVoigtThreads;
        nu[j] = (i < n) ? kLightSpeed / wl[i] : 1.0;
        acc[j] = 0.0;
    }
    for (int64_t l0 = 0; l0 < n_lines; l0 += kLineChunk) {
        const int cnt = (int)((n_lines - l0 < kLineChunk) ? n_lines - l0 : kLineChunk);
        __syncthreads();
        for (int l = threadIdx.x; l < cnt; l += kVoigtThreads) {
            const double w0 = line_wl[l0 + l];
            const double sig_nu = sigma_v / w0;           // :688
            const double rs = 1.0 / sig_nu;
            const double y = line_gamma[l0 + l] * rs * kInvSqrt2;
            const double amp = K * line_f[l0 + l] * (rs / kSqrt2Pi);
            s_a[l] = make_double2(kLightSpeed / w0, rs * kInvSqrt2);
            s_b[l] = make_double2(-0.5 * y * y, amp * y * kInvSqrtPi);
            s_c[l] = make_double2(y, amp);
        }
        __syncthreads();
        for (int l = 0; l < cnt; ++l) {
            const double2 a = s_a[l], b = s_b[l];
#pragma unroll
            for (int j = 0; j < kVoigtWpt; ++j) {
                const double x = (nu[j] - a.x) * a.y;
                const double x2 = x * x;
                const double r2 = fma(-2.0, b.x, x2);
                if (r2 >= 1e8) {
                    const double inv = pb_rcp(r2);
                    const double corr = fma(inv * inv, fma(1.5, x2, b.x), 1.0);
                    acc[j] = fma(b.y * inv, corr, acc[j]);
                } else {
                    const double2 c = s_c[l];
                    acc[j] = fma(c.y, pb_faddeeva_re(x, c.x), acc[j]);
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kVoigtWpt; ++j) {
        const int64_t i = i0 + (int64_t)j * kVoigtThreads;
        if (i < n) {
            if (out_sigma) out_sigma[i] = acc[j];
            if (out_log10) out_log10[i] = log10(acc[j] + offset);   // :713-714
        }
    }
}

__global__ void voigt_profile_kernel(const double* __restrict__ x, const double* __restrict__ sigma,
                                     const double* __restrict__ gamma, int64_t n, double* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = pb_voigt(x[i], sigma[i], gamma[i]);
}

// ---- MolecularConstituent.getSigmaAbs materialised (gasProperties.py:789-818) -------------------
__global__ void molecular_sigma_kernel(MolTable tab, double offset, const double* __restrict__ P, double T,
                                       const double* __restrict__ wl, int64_t n_c, int64_t n_x, int64_t n_wav,
                                       double* __restrict__ out) {
    const int64_t total = n_c * n_x * n_wav;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t w = i % n_wav, cx = i / n_wav, c = cx / n_x;
        const double Pc = fmax(P[cx], 1e-4);    // np.clip(P, a_min=1e-4) :805
        out[i] = exp10(pb_mol_log_sigma(tab, Pc, T, wl[c * n_wav + w])) - offset;   // :814
    }
}

// ---- getLOSopticalDepth_Batch materialised (gasProperties.py:885-956) ---------------------------
// tau[c][w], c = m*n_orb + p (reference chord order).  One CTA per (chord, wavelength tile).
__global__ void __launch_bounds__(256)
optical_depth_kernel(const DevScene* __restrict__ scp, const double* __restrict__ sigma_u, int64_t wav_lo,
                     int64_t n_wav_sel, int64_t chord0, double* __restrict__ tau) {
    const DevScene& sc = *scp;
    const int64_t c = chord0 + blockIdx.y;
    const int m = (int)(c / sc.n_orb), p = (int)(c % sc.n_orb);
    const int64_t row = (int64_t)p * sc.Mp + m;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_wav_sel; w += (int64_t)gridDim.x * blockDim.x) {
        double t = 0.0;
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            const DevScenario& S = sc.scen[C.scenario];
            if (C.kind == 0) {
                const double N = sc.ncol[(int64_t)C.atom_index * sc.n_orb * sc.Mp + row];
                t += N * sigma_u[((int64_t)C.atom_index * sc.n_orb + p) * n_wav_sel + w];     // :954
            } else {
                MolTable tab = pb_mol_table(C);
                const double lam = S.shift[p] * sc.wavelength[wav_lo + w];
                const double* dens = sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x;
                double s = 0.0;
                const double Tk = pb_scn_T(S, p);
                for (int ix = 0; ix < sc.n_x; ++ix) {
                    const double n = dens[ix];
                    const double Pc = fmax(n * kBoltzmann * Tk, 1e-4);                        // :927, :805
                    const double sg = exp10(pb_mol_log_sigma(tab, Pc, Tk, lam)) - C.offset;
                    s = fma(n * C.chi, sg, s);                                                // :932-936
                }
                t += s * sc.delta_x;
            }
        }
        tau[c * n_wav_sel + w] = t;
    }
}

// ---- Atmosphere.getLOSopticalDepth_Batch for CALLER-SUPPLIED chords (gasProperties.py:885-956) ------------------
// One CTA per (chord, wavelength tile).  The chord's geometry arrives as arrays the host evaluated with the objects' own
// methods: y, z and, per scenario, the density centre (cx, cy), the Doppler factor and the amplitude -- [n_scen][n].
// The CTA first forms the x-sum of every scenario (libm forms of pb_density, fixed-order block reduction; the density
// row of a scenario with molecules stays in shared memory), then every thread walks its wavelengths.
__global__ void __launch_bounds__(256)
optical_depth_chords_kernel(const DevScene* __restrict__ scp, int64_t n, int64_t chord0, const double* __restrict__ y,
                            const double* __restrict__ z, const double* __restrict__ cx, const double* __restrict__ cy,
                            const double* __restrict__ shift, const double* __restrict__ amp, int64_t wav_lo, int64_t n_wav_sel,
                            double* __restrict__ tau, int* __restrict__ oob_flag) {
    const DevScene& sc = *scp;
    extern __shared__ double od_sm[];                 // [n_scen] sums | [8] warp partials | [n_density_slots][n_x] density rows
    double* s_sum = od_sm;
    double* s_part = od_sm + kMaxScenarios;
    double* s_dens = s_part + 8;
    const int64_t c = chord0 + blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const double yc = y[c], zc = z[c];
    for (int s = 0; s < sc.n_scen; ++s) {
        const DevScenario& S = sc.scen[s];
        const double a = amp ? amp[(int64_t)s * n + c] : 1.0;
        double part = 0.0;
        int oob = 0;
        for (int ix = tid; ix < sc.n_x; ix += blockDim.x) {
            double d;
            if (S.type == PB200_SCN_SERPENS) d = pb_trilinear(S, sc.x[ix], yc, zc, &oob) * a;
            else d = pb_density(S, S.p, a, sc.x[ix] - cx[(int64_t)s * n + c], yc - cy[(int64_t)s * n + c], zc);
            if (S.density_slot >= 0) s_dens[(size_t)S.density_slot * sc.n_x + ix] = d;
            part += d;
        }
        if (oob) *oob_flag = 1;
        part = pb_warp_sum(part);
        if (lane == 0) s_part[wid] = part;
        __syncthreads();
        if (tid == 0) { double t = 0.0; for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_part[w]; s_sum[s] = t; }
        __syncthreads();
    }
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + tid; w < n_wav_sel; w += (int64_t)gridDim.x * blockDim.x) {
        const double lam0 = sc.wavelength[wav_lo + w];
        double t = 0.0;
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            const DevScenario& S = sc.scen[C.scenario];
            const double lam = shift[(int64_t)C.scenario * n + c] * lam0;                        // :917
            if (C.kind == 0) {
                const double N = (s_sum[C.scenario] * C.chi) * sc.delta_x;                       // :940
                t += N * (exp10(pb_cons_log_sigma(C, lam)) - C.offset);                          // :950-954
            } else {
                MolTable tab = pb_mol_table(C);
                const double* dens = s_dens + (size_t)S.density_slot * sc.n_x;
                double acc = 0.0;
                for (int ix = 0; ix < sc.n_x; ++ix) {
                    const double nd = dens[ix];
                    const double Pc = fmax(nd * kBoltzmann * S.T, 1e-4);                          // :927, :805
                    acc = fma(nd * C.chi, exp10(pb_mol_log_sigma(tab, Pc, S.T, lam)) - C.offset, acc);   // :932-936
                }
                t += acc * sc.delta_x;
            }
        }
        tau[c * n_wav_sel + w] = t;
    }
}

// ---- table producers: log10(values + offset), optionally with the last axis reversed ---------------------------
// MolecularConstituent.constructLookupFunction (gasProperties.py:773-781: xsecarr[:, :, ::-1], log10(sigma + offset)) and
// Star.addFstarFunction (celestialBodies.py:263: log10(F)).  HBM-bound: 16 B per element.
__global__ void __launch_bounds__(256)
log_table_kernel(const double* __restrict__ in, int64_t n_rows, int64_t n_cols, int reverse, double offset, double* __restrict__ out) {
    const int64_t n = n_rows * n_cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / n_cols, col = i - r * n_cols;
        const double v = in[reverse ? r * n_cols + (n_cols - 1 - col) : i];
        out[i] = log10(__dadd_rn(v, offset));
    }
}
void launch_log_table(const double* in, int64_t n_rows, int64_t n_cols, int reverse, double offset, double* out, int sm_count,
                      cudaStream_t st) {
    int64_t blocks = (n_rows * n_cols + 255) / 256;
    if (blocks > (int64_t)sm_count * 16) blocks = (int64_t)sm_count * 16;
    log_table_kernel<<<(unsigned)(blocks > 0 ? blocks : 1), 256, 0, st>>>(in, n_rows, n_cols, reverse, offset, out);
}

// ---- FP64 probes: DFMA chain and exp throughput --------------------------------------------------
__global__ void __launch_bounds__(256) probe_dfma_kernel(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
// throughput of kernel (c')'s sample (one absorber: 6 FP64 + LOP3 + LEA + LDS.64 + IMAD, see "Big-table
// form" in pb200_math.cuh) with nothing else around it: 8 independent samples per thread and iteration
__global__ void __launch_bounds__(256) probe_exp_kernel(double* out, int iters, const double* __restrict__ exp_tab) {
    extern __shared__ __align__(16) double etab[];
    for (int i = threadIdx.x; i < kExpBigEntries; i += blockDim.x) etab[i] = exp_tab[kExpTabEntries + i];
    __syncthreads();
    const uint32_t tab_base = (uint32_t)__cvta_generic_to_shared(etab);
    double ns[8], s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { ns[i] = -1e-3 * (threadIdx.x + 1) * (i + 1); s[i] = 0.0; }
    double N = 1.0;
    const double w0 = 0.5, w1 = 0.5 * kExpBigC1, w2 = 0.5 * kExpBigC2;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const double t = fma(N, ns[i], kExpMagic);
            const double r = fma(N, ns[i], -(t - kExpMagic));
            const double T = pb_exp_big_load(tab_base, __double2loint(t));
            double q = fma(r, w2, w1);
            q = fma(q, r, w0);
            s[i] = fma(T, q, s[i]);
        }
        N = __hiloint2double(__double2hiint(N), __double2loint(N) + 4096);   // next chord: integer pipe
    }
    double r = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// ---- launchers ------------------------------------------------------------------------------------
static inline unsigned blocks_for(int64_t n, int threads, int sm_count) {
    int64_t b = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)sm_count * 16;
    if (b > cap) b = cap;
    return (unsigned)(b > 0 ? b : 1);
}
void launch_sigma_lookup(const double* tx, const double* ty, int64_t n_tab, double offset, const double* wl,
                         int64_t n, double* out, int sm_count, cudaStream_t st) {
    sigma_lookup_kernel<<<blocks_for(n, 256, sm_count), 256, 0, st>>>(tx, ty, n_tab, offset, wl, n, out);
}
void launch_sigma_u(const DevScene* d_scene, int n_atoms, int n_orb, int64_t wav_lo, int64_t n_wav_sel,
                    double* out, int sm_count, cudaStream_t st) {
    dim3 grid(blocks_for(n_wav_sel, 256, sm_count), n_orb, n_atoms);
    sigma_u_kernel<<<grid, 256, 0, st>>>(d_scene, wav_lo, n_wav_sel, out);
}
void launch_voigt_sigma(const double* lw, const double* lg, const double* lf, int64_t n_lines, double sigma_v,
                        const double* wl, int64_t n, double offset, double* out_sigma, double* out_log10,
                        cudaStream_t st) {
    const int64_t per = kVoigtThreads * kVoigtWpt;
    voigt_sigma_kernel<<<(unsigned)((n + per - 1) / per), kVoigtThreads, 0, st>>>(lw, lg, lf, n_lines, sigma_v, wl, n, offset,
                                                                                 out_sigma, out_log10);
}
void launch_voigt_profile(const double* x, const double* s, const double* g, int64_t n, double* out, int sm_count,
                          cudaStream_t st) {
    voigt_profile_kernel<<<blocks_for(n, 256, sm_count), 256, 0, st>>>(x, s, g, n, out);
}
void launch_molecular_sigma(const MolTable& tab, double offset, const double* P, double T, const double* wl,
                            int64_t n_c, int64_t n_x, int64_t n_wav, double* out, int sm_count, cudaStream_t st) {
    molecular_sigma_kernel<<<blocks_for(n_c * n_x * n_wav, 256, sm_count), 256, 0, st>>>(tab, offset, P, T, wl, n_c,
                                                                                         n_x, n_wav, out);
}
void launch_optical_depth(const DevScene* d_scene, const double* sigma_u, int64_t n_chords, int64_t wav_lo,
                          int64_t n_wav_sel, double* tau, cudaStream_t st) {
    unsigned bx = (unsigned)((n_wav_sel + 255) / 256);
    if (bx > 64) bx = 64;
    // gridDim.y is limited to 65535: chunk the chord axis
    for (int64_t c0 = 0; c0 < n_chords; c0 += 65535) {
        const unsigned by = (unsigned)((n_chords - c0 < 65535) ? n_chords - c0 : 65535);
        optical_depth_kernel<<<dim3(bx, by), 256, 0, st>>>(d_scene, sigma_u, wav_lo, n_wav_sel, c0, tau);
    }
}
cudaError_t launch_optical_depth_chords(const DevScene* d_scene, int n_x, int n_density_slots, int64_t n, const double* y,
                                        const double* z, const double* cx, const double* cy, const double* shift,
                                        const double* amp, int64_t wav_lo, int64_t n_wav_sel, double* tau, int* oob,
                                        cudaStream_t st) {
    unsigned bx = (unsigned)((n_wav_sel + 255) / 256);
    if (bx > 32) bx = 32;
    const size_t smem = (size_t)(kMaxScenarios + 8 + (size_t)n_density_slots * n_x) * sizeof(double);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    if (smem > 40 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(optical_depth_chords_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    for (int64_t c0 = 0; c0 < n; c0 += 65535) {   // gridDim.y limit
        const unsigned by = (unsigned)((n - c0 < 65535) ? n - c0 : 65535);
        optical_depth_chords_kernel<<<dim3(bx, by), 256, smem, st>>>(d_scene, n, c0, y, z, cx, cy, shift, amp, wav_lo, n_wav_sel, tau, oob);
    }
    return cudaGetLastError();
}
void launch_probe_dfma(double* out, int blocks, int iters, cudaStream_t st) {
    probe_dfma_kernel<<<blocks, 256, 0, st>>>(out, iters);
}
void launch_probe_exp(double* out, int blocks, int iters, const double* exp_tab, cudaStream_t st) {
    probe_exp_kernel<<<blocks, 256, kExpBigEntries * sizeof(double), st>>>(out, iters, exp_tab);
}

}  // namespace pb200
