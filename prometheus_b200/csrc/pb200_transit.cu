// pb200_transit.cu -- kernel (c'): fused optical depth + Beer-Lambert + stellar-disk sum.
//
//   R(p, w) = sum_m w_in(p,m) exp(-sum_a Ncol_a(p,m) sigma_a(p,w)) / sum_m w_out(m)
//
// restates Transit.sumOverChords (gasProperties.py:1205-1258) together with the atomic branch of
// Atmosphere.getLOSopticalDepth_Batch (:937-954) without ever materialising tau(chord, lambda):
//   * a thread owns WPT wavelengths of one orbital phase and keeps -sigma_a(p, w) in registers
//     (looked up once in the prologue: np.interp + 10**v - offset, i.e. n_interp_log :34-51);
//   * the chord records {w_in, Ncol_0..Ncol_{S-1}} of the phase stream through shared memory in
//     256-chord tiles, one 1-D TMA bulk copy per tile (cp.async.bulk + mbarrier, double buffered);
//     every thread of the CTA reads the same record, so the LDS is a broadcast;
//   * fp64 exp (table-driven, pb_exp_neg_tab: the kernel is bound by the FP64 pipe, so the exp is
//     built to need as few FP64 instructions as possible), fp64 accumulators; the chord axis is
//     split over blockIdx.z to get many more CTAs than resident slots (tail effect), partial sums
//     reduced in a fixed order by finalize_kernel (deterministic).
// The chord-level early-out (w_in == 0: blocked/padding) is CTA-uniform.
#include <cstdlib>

#include "pb200_internal.cuh"

namespace pb200 {

constexpr int kTransitThreads = 128;

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

// One 256-chord tile: acc[j] += w_m * exp(-sum_a N_a(m) sigma_a(j)) for the thread's WPT wavelengths.
// `stage` is the shared-window address of the tile's records (explicit ld.shared: no generic->shared
// address conversion inside the loop).
template <int S, int WPT, bool CLAMP>
__device__ __forceinline__ void transit_tile(uint32_t stage, const double (&nsig)[S][WPT], double (&acc)[WPT],
                                             uint32_t tab_lane, const ExpConsts& ec) {
    constexpr uint32_t kRecBytes = (S + 1) * sizeof(double);
#pragma unroll 2
    for (int m = 0; m < kChordTile; ++m) {
        const uint32_t rec = stage + m * kRecBytes;
        double wv, N[S];
        if constexpr (S == 1) {
            lds_v2f64(rec, wv, N[0]);                                              // broadcast LDS.128
        } else {
            wv = lds_f64(rec);
#pragma unroll
            for (int a = 0; a < S; ++a) N[a] = lds_f64(rec + 8 + 8 * a);
        }
        // blocked chord or padding: F_in = 0 (gasProperties.py:1233); CTA-uniform, integer test
        if ((__double2hiint(wv) | __double2loint(wv)) == 0) continue;
#pragma unroll
        for (int j = 0; j < WPT; ++j) {
            double u = N[0] * nsig[0][j];                                          // -tau, :954
#pragma unroll
            for (int a = 1; a < S; ++a) u = fma(N[a], nsig[a][j], u);
            acc[j] = fma(wv, pb_exp_neg_tab<CLAMP>(u, tab_lane, ec), acc[j]);     // :1241, :1244/:1255
        }
    }
}

// S atomic constituents (compile time), WPT wavelengths per thread
template <int S, int WPT>
__global__ void __launch_bounds__(kTransitThreads)
transit_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    // shared memory: exp table [64][16] | stage 0 records | stage 1 records; a record is
    // {w_in, Ncol_0 .. Ncol_{S-1}}, a stage holds kChordTile records
    extern __shared__ __align__(128) double smem[];
    __shared__ uint64_t full_bar[2];
    constexpr int R = S + 1;
    constexpr int kTabDoubles = kExpTabEntries * kExpTabCopies;
    constexpr int kStageDoubles = R * kChordTile;
    constexpr uint32_t kStageBytes = kStageDoubles * sizeof(double);
    double* etab = smem;
    double* stages = smem + kTabDoubles;

    const int tid = threadIdx.x;
    const int p = prm.phase_lo + blockIdx.y;
    const int split = blockIdx.z;
    const int tile0 = split * prm.tiles_per_split;
    const int n_tiles_total = sc.Mp / kChordTile;
    int n_tiles = n_tiles_total - tile0;
    if (n_tiles > prm.tiles_per_split) n_tiles = prm.tiles_per_split;

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < kTabDoubles; i += kTransitThreads) etab[i] = __ldg(sc.exp_tab + (i / kExpTabCopies));
    __syncthreads();

    const double* rec_row = sc.rec + ((int64_t)p * sc.Mp + (int64_t)tile0 * kChordTile) * R;
    auto issue = [&](int tile, int stage) {
        mbar_expect_tx(&full_bar[stage], kStageBytes);
        tma_bulk_g2s(stages + stage * kStageDoubles, rec_row + (int64_t)tile * kStageDoubles, kStageBytes,
                     &full_bar[stage]);
    };
    if (tid == 0 && n_tiles > 0) issue(0, 0);

    // ---- prologue: -sigma_a(p, w) for this thread's wavelengths (kernel (b1) inlined) -----------
    const int64_t w_base = prm.wav_lo + (int64_t)blockIdx.x * (kTransitThreads * WPT) + tid;
    double nsig[S][WPT];
    double acc[WPT];
    bool bad = false;   // NaN cross-section (NaN wavelength): the reference propagates NaN into R
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        acc[j] = 0.0;
        const int64_t w = w_base + (int64_t)j * kTransitThreads;
        const bool live = w < prm.wav_lo + prm.n_wav_sel;
        const double lam = live ? __ldg(sc.wavelength + w) : 0.0;
        int a = 0;
        for (int k = 0; k < sc.n_cons && a < S; ++k) {
            const DevConstituent& C = sc.cons[k];
            if (C.kind != 0) continue;
            double v = 0.0;
            if (live) {
                const double shift = __ldg(sc.scen[C.scenario].shift + p);
                v = pb_interp_log(C.tab_x, C.tab_y, C.n_tab, shift * lam, C.offset);   // :917, :950-951
                bad |= (v != v);
            }
#pragma unroll
            for (int aa = 0; aa < S; ++aa)
                if (aa == a) nsig[aa][j] = -v;
            ++a;
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
        clamp |= !(tmax < kExpClampTau);   // NaN -> clamped variant (harmless)
    }
    clamp = __any_sync(0xffffffffu, clamp);

    // ---- main loop over chord tiles -----------------------------------------------------------
    const uint32_t tab_lane = smem_u32(etab) + (tid & (kExpTabCopies - 1)) * (uint32_t)sizeof(double);
    const uint32_t stages_addr = smem_u32(stages);
    for (int t = 0; t < n_tiles; ++t) {
        const int stage = t & 1;
        if (tid == 0 && t + 1 < n_tiles) issue(t + 1, stage ^ 1);
        mbar_wait(&full_bar[stage], (t >> 1) & 1);
        const uint32_t sw = stages_addr + stage * kStageBytes;
        if (clamp) transit_tile<S, WPT, true>(sw, nsig, acc, tab_lane, pb_exp_consts());
        else transit_tile<S, WPT, false>(sw, nsig, acc, tab_lane, pb_exp_consts());
        __syncthreads();   // stage may be refilled by the producer in the next iteration
    }

    // ---- epilogue -------------------------------------------------------------------------------
    const double fout = (prm.n_split == 1) ? sc.fout[0] : 1.0;
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        const int64_t w = w_base + (int64_t)j * kTransitThreads;
        if (w >= prm.wav_lo + prm.n_wav_sel) continue;
        const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + (w - prm.wav_lo);
        double v = acc[j];
        if (bad) {
#pragma unroll
            for (int a = 0; a < S; ++a)
                if (nsig[a][j] != nsig[a][j]) v = nsig[a][j];
        }
        if (prm.n_split == 1) prm.R[o] = v / fout;                             // :1258
        else prm.partial[(int64_t)split * prm.n_phase * prm.n_wav_sel + o] = v;
    }
}

// R = (sum over splits, fixed order) / F_out
__global__ void finalize_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    const int64_t n = (int64_t)prm.n_phase * prm.n_wav_sel;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < prm.n_split; ++k) s += prm.partial[(int64_t)k * n + i];
        const double fout = sc.fout[sc.fout_per_wav ? prm.wav_lo + i % prm.n_wav_sel : 0];
        prm.R[i] = s / fout;
    }
}

template <int S, int WPT>
static cudaError_t launch_one(const DevScene* d_scene, const TransitParams& prm, cudaStream_t st) {
    const size_t smem = (kExpTabEntries * kExpTabCopies + 2 * (S + 1) * kChordTile) * sizeof(double);
    auto kern = transit_kernel<S, WPT>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const int64_t per_block = (int64_t)kTransitThreads * WPT;
    dim3 grid((unsigned)((prm.n_wav_sel + per_block - 1) / per_block), (unsigned)prm.n_phase, (unsigned)prm.n_split);
    kern<<<grid, kTransitThreads, smem, st>>>(d_scene, prm);
    return cudaGetLastError();
}

static int env_wpt() {
    static const int v = [] { const char* e = getenv("PB200_WPT"); return e ? atoi(e) : 0; }();
    return v;
}

int transit_wpt_for(int n_atoms) {
    const int forced = env_wpt();
    if (n_atoms <= 4) return (forced == 4 || forced == 8) ? forced : 8;
    return 2;
}

// returns cudaSuccess, or cudaErrorInvalidValue when n_atoms is outside the compiled range
cudaError_t launch_transit(const DevScene* d_scene, int n_atoms, const TransitParams& prm, cudaStream_t st) {
    const bool w8 = transit_wpt_for(n_atoms) == 8;
    switch (n_atoms) {
        case 1: return w8 ? launch_one<1, 8>(d_scene, prm, st) : launch_one<1, 4>(d_scene, prm, st);
        case 2: return w8 ? launch_one<2, 8>(d_scene, prm, st) : launch_one<2, 4>(d_scene, prm, st);
        case 3: return w8 ? launch_one<3, 8>(d_scene, prm, st) : launch_one<3, 4>(d_scene, prm, st);
        case 4: return w8 ? launch_one<4, 8>(d_scene, prm, st) : launch_one<4, 4>(d_scene, prm, st);
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
