// pb200_transit.cu -- kernel (c'): fused optical depth + Beer-Lambert + stellar-disk sum.
//
//   R(p, w) = sum_m w_in(p,m) exp(-sum_a Ncol_a(p,m) sigma_a(p,w)) / sum_m w_out(m)
//
// restates Transit.sumOverChords (gasProperties.py:1205-1258) together with the atomic branch of
// Atmosphere.getLOSopticalDepth_Batch (:937-954) without ever materialising tau(chord, lambda):
//   * sigma_phase_kernel looks up -sigma_a(p, w) * 4096/ln2 once per (phase, wavelength)
//     (np.interp + 10**v - offset, i.e. n_interp_log :34-51) -- the chord-axis splits of one
//     (phase, wavelength) tile would otherwise each repeat the binary search;
//   * a thread of transit_kernel owns WPT wavelengths of one orbital phase and keeps those values
//     in registers;
//   * the chord records {w, w c1, w c2, Ncol_0..Ncol_{S-1}} of the phase stream through shared
//     memory in 256-chord tiles, one 1-D TMA bulk copy per tile (cp.async.bulk + mbarrier, double
//     buffered); every thread of the CTA reads the same record, so the LDS is a broadcast;
//   * fp64 exp (table-driven, see "Big-table form" in pb200_math.cuh: the kernel is bound by the
//     FP64 pipe and by the instructions issued next to it, so the exp is built to need as few of
//     both as possible: 6 FP64 + 3 integer + 1 LDS per sample for one absorber), fp64
//     accumulators; the chord axis is split over blockIdx.z to get many more CTAs than resident
//     slots (tail effect), partial sums reduced in a fixed order by finalize_kernel (deterministic).
#include <cstdlib>

#include "pb200_internal.cuh"

namespace pb200 {

constexpr int kRecExtra = 3;           // {w, w c1, w c2} in front of the S column densities

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
    const int64_t wi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = wi < prm.n_wav_sel;
    const int pi = blockIdx.y, p = prm.phase_lo + pi;
    const double lam = live ? __ldg(sc.wavelength + prm.wav_lo + wi) : 0.0;
    int a = 0;
    for (int k = 0; k < sc.n_cons; ++k) {
        const DevConstituent& C = sc.cons[k];
        if (C.kind != 0) continue;
        unsigned long long bits = 0ull;
        if (live) {
            const double shift = __ldg(sc.scen[C.scenario].shift + p);
            const double v = pb_interp_log(C.tab_x, C.tab_y, C.n_tab, shift * lam, C.offset);   // :917, :950-951
            prm.nsig[((int64_t)a * prm.n_phase + pi) * prm.n_wav_sel + wi] = -v * kExpBigScale;
            bits = (v >= 0.0) ? (unsigned long long)__double_as_longlong(v) : ~0ull;   // negative or NaN: never cull
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, bits, d);
            bits = o > bits ? o : bits;
        }
        if ((threadIdx.x & 31) == 0 && bits) atomicMax(sc.smax + (int64_t)a * sc.n_orb + p, bits);
        ++a;
    }
}

// One 256-chord tile: acc[j] += w_m * exp(-sum_a N_a(m) sigma_a(j)) for the thread's WPT wavelengths.
// `stage` is the shared-window address of the tile's records (explicit ld.shared: no generic->shared
// address conversion inside the loop).  There is no branch on blocked chords or padding: their record
// is all zeros (w = w c1 = w c2 = 0, N = 0), so they add T * 0 = +0 exactly (F_in = 0,
// gasProperties.py:1233) -- a few per cent of wasted samples buy a branch-free loop body that ptxas
// can software-pipeline across chords.
template <int S, int WPT, bool CLAMP, int UNROLL = 2, int DROP = 0>
__device__ __forceinline__ void transit_tile(uint32_t stage, const double (&nsig)[S][WPT], double (&acc)[WPT],
                                             uint32_t tab_base) {
    constexpr uint32_t kRecBytes = (S + kRecExtra) * sizeof(double);
#pragma unroll UNROLL
    for (int m = 0; m < kChordTile; ++m) {
        const uint32_t rec = stage + m * kRecBytes;
        double w0, w1, w2, N[S];
        if constexpr (S == 1) {
            lds_v2f64(rec, w0, w1);                                                // broadcast LDS.128 x 2
            lds_v2f64(rec + 16, w2, N[0]);
        } else {
            w0 = lds_f64(rec); w1 = lds_f64(rec + 8); w2 = lds_f64(rec + 16);
#pragma unroll
            for (int a = 0; a < S; ++a) N[a] = lds_f64(rec + 24 + 8 * a);
        }
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
            double T;
            if constexpr (DROP == 0) T = pb_exp_big_load(tab_base, __double2loint(t));
            else if constexpr (DROP == 1) T = __hiloint2double(0x3ff00000 + (__double2loint(t) << 8), 0);           // no LOP3/LEA/LDS
            else if constexpr (DROP == 2) { int lo_, hi_; asm volatile("{.reg .b32 j, a; and.b32 j, %2, 4095; mad.lo.u32 a, j, 8, %3; ld.shared.v2.b32 {%0, %1}, [a];}" : "=r"(lo_), "=r"(hi_) : "r"(__double2loint(t)), "r"(tab_base)); T = __hiloint2double(hi_, lo_); }   // no IMAD
            else T = __hiloint2double(0x3ff00000, __double2loint(t));                                                  // FP64 only
            double q = fma(r, w2, w1);
            q = fma(q, r, w0);                                                     // w exp(r ln2/4096)
            acc[j] = fma(T, q, acc[j]);                                            // :1241, :1244/:1255
        }
    }
}

// S atomic constituents (compile time), WPT wavelengths per thread, THREADS threads per CTA
template <int S, int WPT, int THREADS, int VAR = 0>
__global__ void __launch_bounds__(THREADS, (VAR & 1) ? 4 : (THREADS == 256 ? 3 : 6))
transit_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    // shared memory: exp table [4096] | stage 0 records | stage 1 records; a stage holds kChordTile records
    extern __shared__ __align__(128) double smem[];
    __shared__ uint64_t full_bar[2];
    constexpr int R = S + kRecExtra;
    constexpr int kStageDoubles = R * kChordTile;
    constexpr uint32_t kStageBytes = kStageDoubles * sizeof(double);
    double* etab = smem;
    double* stages = smem + kExpBigEntries;

    const int tid = threadIdx.x;
    const int p = prm.phase_lo + blockIdx.y;
    const int split = blockIdx.z;
    // this CTA's share of the phase's ACTIVE tiles (compact_kernel): even shares, at least 4 tiles per
    // working CTA so that the table load and the prologue stay amortised; CTAs beyond the need store zeros
    const int n_act = sc.n_act[p];
    const int tps = pb_tiles_per_split(n_act, prm.n_split, prm.min_tiles);
    const int tile0 = split * tps;
    int n_tiles = n_act - tile0;
    if (n_tiles > tps) n_tiles = tps;
    const int64_t w_base = (int64_t)blockIdx.x * (THREADS * WPT) + tid;     // relative to wav_lo
    if (n_tiles <= 0) {   // nothing to do (CTA-uniform): finalize_kernel skips this split; R = w0 / F_out for the only split
        if (prm.n_split > 1) return;
#pragma unroll
        for (int j = 0; j < WPT; ++j) {
            const int64_t w = w_base + (int64_t)j * THREADS;
            if (w >= prm.n_wav_sel) continue;
            const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + w;
            {
                double v = sc.w0[p] / sc.fout[0];
                for (int a = 0; a < S; ++a) {   // NaN cross-section: the reference propagates NaN into R
                    const double ns = __ldg(prm.nsig + ((int64_t)a * prm.n_phase + blockIdx.y) * prm.n_wav_sel + w);
                    if (ns != ns) v = ns;
                }
                prm.R[o] = v;
            }
        }
        return;
    }

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
    if (tid == 0 && n_tiles > 0) issue(0, 0);
    {   // the exp table: 32 KB from L2, 16-byte loads
        const double2* src = reinterpret_cast<const double2*>(sc.exp_tab + kExpTabEntries);
        double2* dst = reinterpret_cast<double2*>(etab);
        for (int i = tid; i < kExpBigEntries / 2; i += THREADS) dst[i] = __ldg(src + i);
    }

    // ---- prologue: -sigma_a(p, w) 4096/ln2 for this thread's wavelengths ---------------------------
    double nsig[S][WPT];
    double acc[WPT];
    bool bad = false;   // NaN cross-section (NaN wavelength): the reference propagates NaN into R
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        acc[j] = 0.0;
        const int64_t w = w_base + (int64_t)j * THREADS;
        const bool live = w < prm.n_wav_sel;
#pragma unroll
        for (int a = 0; a < S; ++a) {
            const double v = live ? __ldg(prm.nsig + ((int64_t)a * prm.n_phase + blockIdx.y) * prm.n_wav_sel + w) : 0.0;
            bad |= (v != v);
            nsig[a][j] = v;
        }
    }

    // A warp needs the clamped exp only if one of its wavelengths can reach tau > 690 on the densest
    // chord of this phase (warp-uniform choice, so the two loop variants never diverge).
    bool clamp = false;
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        double tmax = 0.0;
#pragma unroll
        for (int a = 0; a < S; ++a)
            tmax = fma(__longlong_as_double((long long)__ldg(sc.nmax + (int64_t)a * sc.n_orb + p)), -nsig[a][j], tmax);
        clamp |= !(tmax < kExpBigClampDecision);   // NaN -> clamped variant (harmless)
    }
    clamp = __any_sync(0xffffffffu, clamp);
    __syncthreads();   // exp table complete

    // ---- main loop over chord tiles -----------------------------------------------------------
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
        if (clamp) transit_tile<S, WPT, true>(sw, nsig, acc, tab_base);
        else transit_tile<S, WPT, false, (VAR & 2) ? 4 : 2, (VAR >> 2)>(sw, nsig, acc, tab_base);
        __syncthreads();   // stage may be refilled by the producer in the next iteration
    }

    // ---- epilogue -------------------------------------------------------------------------------
    const double fout = (prm.n_split == 1) ? sc.fout[0] : 1.0;
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        const int64_t w = w_base + (int64_t)j * THREADS;
        if (w >= prm.n_wav_sel) continue;
        const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + w;
        double v = acc[j];
        if (bad) {
#pragma unroll
            for (int a = 0; a < S; ++a)
                if (nsig[a][j] != nsig[a][j]) v = nsig[a][j];
        }
        if (prm.n_split == 1) prm.R[o] = (v + sc.w0[p]) / fout;                // :1258
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
        prm.R[i] = s / fout;
    }
}

template <int S, int WPT, int THREADS, int VAR = 0>
static cudaError_t launch_one(const DevScene* d_scene, const TransitParams& prm, cudaStream_t st) {
    const size_t smem = (kExpBigEntries + 2 * (S + kRecExtra) * kChordTile) * sizeof(double);
    auto kern = transit_kernel<S, WPT, THREADS, VAR>;
    {   // dynamic + static shared memory is always above the 48 KB default limit
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const int64_t per_block = (int64_t)THREADS * WPT;
    dim3 grid((unsigned)((prm.n_wav_sel + per_block - 1) / per_block), (unsigned)prm.n_phase, (unsigned)prm.n_split);
    kern<<<grid, THREADS, smem, st>>>(d_scene, prm);
    return cudaGetLastError();
}

static int env_int(const char* name) {
    const char* e = getenv(name);
    return e ? atoi(e) : 0;
}

// CTA shape: 256 threads x WPT wavelengths, WPT = 8 for 1-2 absorbers, 4 for 3-4, 2 above (register
// budget: 2 S WPT + temporaries must leave >= 2 CTAs per SM).  PB200_THREADS=128 selects 128-thread
// CTAs for 1-2 absorbers (tuning experiments only).
static int transit_threads() {
    static const int v = env_int("PB200_THREADS");
    return v == 128 ? 128 : 256;
}

int transit_cta_wavelengths(int n_atoms) {
    return n_atoms <= 2 ? transit_threads() * 8 : n_atoms <= 4 ? 256 * 4 : 256 * 2;
}

void launch_sigma_phase(const DevScene* d_scene, const TransitParams& prm, cudaStream_t st) {
    dim3 grid((unsigned)((prm.n_wav_sel + 255) / 256), (unsigned)prm.n_phase);
    sigma_phase_kernel<<<grid, 256, 0, st>>>(d_scene, prm);
}

// returns cudaSuccess, or cudaErrorInvalidValue when n_atoms is outside the compiled range
cudaError_t launch_transit(const DevScene* d_scene, int n_atoms, const TransitParams& prm, cudaStream_t st) {
    const bool t256 = transit_threads() == 256;
    static const int var = env_int("PB200_TV");
    if (n_atoms == 1 && var == 1) return launch_one<1, 8, 256, 1>(d_scene, prm, st);
    if (n_atoms == 1 && var == 2) return launch_one<1, 8, 256, 2>(d_scene, prm, st);
    if (n_atoms == 1 && var == 3) return launch_one<1, 8, 256, 3>(d_scene, prm, st);
    if (n_atoms == 1 && var == 6) return launch_one<1, 8, 256, 6>(d_scene, prm, st);
    if (n_atoms == 1 && var == 10) return launch_one<1, 8, 256, 10>(d_scene, prm, st);
    if (n_atoms == 1 && var == 14) return launch_one<1, 8, 256, 14>(d_scene, prm, st);
    switch (n_atoms) {
        case 1: return t256 ? launch_one<1, 8, 256>(d_scene, prm, st) : launch_one<1, 8, 128>(d_scene, prm, st);
        case 2: return t256 ? launch_one<2, 8, 256>(d_scene, prm, st) : launch_one<2, 8, 128>(d_scene, prm, st);
        case 3: return launch_one<3, 4, 256>(d_scene, prm, st);
        case 4: return launch_one<4, 4, 256>(d_scene, prm, st);
        case 5: return launch_one<5, 2, 256>(d_scene, prm, st);
        case 6: return launch_one<6, 2, 256>(d_scene, prm, st);
        case 7: return launch_one<7, 2, 256>(d_scene, prm, st);
        case 8: return launch_one<8, 2, 256>(d_scene, prm, st);
        case 9: return launch_one<9, 2, 256>(d_scene, prm, st);
        case 10: return launch_one<10, 2, 256>(d_scene, prm, st);
        case 11: return launch_one<11, 2, 256>(d_scene, prm, st);
        case 12: return launch_one<12, 2, 256>(d_scene, prm, st);
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
