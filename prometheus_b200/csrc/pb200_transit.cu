// pb200_transit.cu -- kernel (c'): fused optical depth + Beer-Lambert + stellar-disk sum.
//
//   R(p, w) = sum_m w_in(p,m) exp(-sum_a Ncol_a(p,m) sigma_a(p,w)) / sum_m w_out(m)
//
// restates Transit.sumOverChords (gasProperties.py:1205-1258) together with the atomic branch of
// Atmosphere.getLOSopticalDepth_Batch (:937-954) without ever materialising tau(chord, lambda):
//   * a thread owns WPT wavelengths of one orbital phase and keeps sigma_a(p, w) in registers
//     (looked up once in the prologue: np.interp + 10**v - offset, i.e. n_interp_log :34-51);
//   * the chord records {w_in, Ncol_0..Ncol_{S-1}} of the phase stream through shared memory in
//     256-chord tiles, fetched by 1-D TMA bulk copies (cp.async.bulk + mbarrier, double buffered);
//     every thread of the CTA reads the same record, so the LDS is a broadcast;
//   * fp64 exp, fp64 accumulators; optional split of the chord axis over blockIdx.z for small
//     (phase x lambda) grids, reduced in a fixed order by finalize_kernel (deterministic).
// Chord-level early-outs (w_in == 0: blocked/padding; all Ncol == 0: exp(0) = 1) are CTA-uniform.
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


// S atomic constituents (compile time), WPT wavelengths per thread
template <int S, int WPT>
__global__ void __launch_bounds__(kTransitThreads)
transit_kernel(const DevScene* __restrict__ scp, const TransitParams prm) {
    const DevScene& sc = *scp;
    // stage layout: [stage][row][kChordTile], row 0 = w_in, rows 1..S = Ncol_a
    extern __shared__ __align__(128) double smem[];
    __shared__ uint64_t full_bar[2];
    const int tid = threadIdx.x;
    const int p = prm.phase_lo + blockIdx.y;
    const int split = blockIdx.z;
    const int tile0 = split * prm.tiles_per_split;
    const int n_tiles_total = sc.Mp / kChordTile;
    int n_tiles = n_tiles_total - tile0;
    if (n_tiles > prm.tiles_per_split) n_tiles = prm.tiles_per_split;

    constexpr uint32_t kRowBytes = kChordTile * sizeof(double);
    constexpr int kStageDoubles = (S + 1) * kChordTile;

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const double* w_in_row = sc.w_in + (int64_t)p * sc.Mp;
    const double* ncol_row = sc.ncol + (int64_t)p * sc.Mp;
    const int64_t atom_stride = (int64_t)sc.n_orb * sc.Mp;
    auto issue = [&](int tile, int stage) {
        double* dst = smem + stage * kStageDoubles;
        const int64_t off = (int64_t)(tile0 + tile) * kChordTile;
        mbar_expect_tx(&full_bar[stage], (S + 1) * kRowBytes);
        tma_bulk_g2s(dst, w_in_row + off, kRowBytes, &full_bar[stage]);
#pragma unroll
        for (int a = 0; a < S; ++a)
            tma_bulk_g2s(dst + (a + 1) * kChordTile, ncol_row + a * atom_stride + off, kRowBytes, &full_bar[stage]);
    };
    if (tid == 0 && n_tiles > 0) issue(0, 0);

    // ---- prologue: sigma_a(p, w) for this thread's wavelengths (kernel (b1) inlined) ------------
    const int64_t w_base = prm.wav_lo + (int64_t)blockIdx.x * (kTransitThreads * WPT) + tid;
    double sig[S][WPT];
    double acc[WPT];
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
            }
#pragma unroll
            for (int aa = 0; aa < S; ++aa)
                if (aa == a) sig[aa][j] = v;
            ++a;
        }
    }

    // ---- main loop over chord tiles -----------------------------------------------------------
    for (int t = 0; t < n_tiles; ++t) {
        const int stage = t & 1;
        if (tid == 0 && t + 1 < n_tiles) issue(t + 1, stage ^ 1);
        mbar_wait(&full_bar[stage], (t >> 1) & 1);
        const double* sw = smem + stage * kStageDoubles;
#pragma unroll 2
        for (int m = 0; m < kChordTile; ++m) {
            const double wv = sw[m];
            if (wv == 0.0) continue;   // blocked chord or padding: F_in = 0 (gasProperties.py:1233)
            double N[S];
            bool any = false;
#pragma unroll
            for (int a = 0; a < S; ++a) {
                N[a] = sw[(a + 1) * kChordTile + m];
                any |= (N[a] != 0.0);
            }
            if (!any) {   // tau == 0 exactly
#pragma unroll
                for (int j = 0; j < WPT; ++j) acc[j] += wv;
                continue;
            }
#pragma unroll
            for (int j = 0; j < WPT; ++j) {
                double tau = N[0] * sig[0][j];
#pragma unroll
                for (int a = 1; a < S; ++a) tau = fma(N[a], sig[a][j], tau);   // :954
                acc[j] = fma(wv, pb_exp_neg(tau), acc[j]);                    // :1241, :1244/:1255
            }
        }
        __syncthreads();   // stage may be refilled by the producer in the next iteration
    }

    // ---- epilogue -------------------------------------------------------------------------------
    const double fout = (prm.n_split == 1) ? sc.fout[0] : 1.0;
#pragma unroll
    for (int j = 0; j < WPT; ++j) {
        const int64_t w = w_base + (int64_t)j * kTransitThreads;
        if (w >= prm.wav_lo + prm.n_wav_sel) continue;
        const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + (w - prm.wav_lo);
        if (prm.n_split == 1) prm.R[o] = acc[j] / fout;                        // :1258
        else prm.partial[(int64_t)split * prm.n_phase * prm.n_wav_sel + o] = acc[j];
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
    const size_t smem = 2 * (S + 1) * kChordTile * sizeof(double);
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

int transit_wpt_for(int n_atoms) { return n_atoms <= 4 ? 4 : 2; }

// returns cudaSuccess, or cudaErrorInvalidValue when n_atoms is outside the compiled range
cudaError_t launch_transit(const DevScene* d_scene, int n_atoms, const TransitParams& prm, cudaStream_t st) {
    switch (n_atoms) {
        case 1: return launch_one<1, 4>(d_scene, prm, st);
        case 2: return launch_one<2, 4>(d_scene, prm, st);
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
