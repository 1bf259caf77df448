// pb200_general.cu -- kernel (m): the general form of the fused transit kernel, used when the scene
// has molecular constituents (gasProperties.py:925-936, :789-818) and/or a tabulated stellar
// spectrum seen through stellar rotation (per-chord F*, gasProperties.py:1213-1222).
//
// One thread owns one wavelength of one phase.  Per molecule it first reduces the (P, T, lambda)
// table to U_j = log10 sigma(P_j, T, shift*lambda) for every pressure node j (T and lambda
// interpolation hoisted out of the chord loop; the column U lives in shared memory, [j][thread],
// conflict-free).  The CTA then walks the chords of its split; for each chord its threads
// cooperatively turn the density row n(x) into records {j, t_P, n*chi*dx} (pressure interval
// search and the division hoisted out of the wavelength dimension), after which every thread sums
//     tau_mol = sum_x n chi dx * (10**((1-t_P) U_j + t_P U_{j+1}) - offset)
// over the records (broadcast LDS) with its own U column.  sigma(chord, x, lambda) is never stored.
#include "pb200_internal.cuh"
#include "pb200_molecule.cuh"

namespace pb200 {

constexpr int kGenThreads = 128;
constexpr int kMaxMolNodes = 64;   // sum over molecules of pressure nodes held in shared memory

struct MolRecord {
    double t;     // fractional position inside the pressure interval, or -1: out of grid -> fill
    double coef;  // n * chi * delta_x
    int j;        // pressure interval
    int pad;
};

// F_out[w] = sum_m rho*clv*F*(lambda_w / shift*(m))   (gasProperties.py:1213-1222, :1232, :1249/:1256)
__global__ void __launch_bounds__(kGenThreads)
fout_table_kernel(const DevScene* __restrict__ scp) {
    const DevScene& sc = *scp;
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= sc.n_wav) return;
    const double lam = sc.wavelength[w];
    double s = 0.0;
    for (int m = 0; m < sc.M; ++m) {
        const double fs = pb_interp_log(sc.fstar_x, sc.fstar_y, sc.n_fstar, lam / sc.star_shift[m], 0.0);
        s += sc.w_out[m] * fs;
    }
    sc.fout[w] = s;
}

__global__ void __launch_bounds__(kGenThreads)
transit_general_kernel(const DevScene* __restrict__ scp, const TransitParams prm, int rec_cap) {
    const DevScene& sc = *scp;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // layout: exp table [256][8] | U[node_total][kGenThreads] | sig[n_atoms][kGenThreads] | rec[2][rec_cap]
    int node_total = 0;
    for (int k = 0; k < sc.n_cons; ++k)
        if (sc.cons[k].kind == 1) node_total += sc.cons[k].dims[0];
    double* etab = reinterpret_cast<double*>(smem_raw);
    for (int i = threadIdx.x; i < kExpTabEntries * kExpTabCopies; i += kGenThreads) etab[i] = __ldg(sc.exp_tab + i / kExpTabCopies);
    const uint32_t tab_lane = (uint32_t)__cvta_generic_to_shared(etab) + (threadIdx.x & (kExpTabCopies - 1)) * 8u;
    double* U = etab + kExpTabEntries * kExpTabCopies;
    double* sig = U + (size_t)node_total * kGenThreads;
    MolRecord* rec = reinterpret_cast<MolRecord*>(sig + (size_t)sc.n_atoms * kGenThreads);

    const int tid = threadIdx.x;
    const int p = prm.phase_lo + blockIdx.y;
    const int64_t w = prm.wav_lo + (int64_t)blockIdx.x * kGenThreads + tid;
    const bool live = w < prm.wav_lo + prm.n_wav_sel;
    const double lam = live ? sc.wavelength[w] : sc.wavelength[prm.wav_lo];
    const bool per_chord_fstar = sc.fout_per_wav != 0;

    // ---- prologue: atomic sigma and molecular U columns -----------------------------------------
    {
        int node0 = 0;
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            const DevScenario& S = sc.scen[C.scenario];
            const double sl = S.shift[p] * lam;                                               // :917
            if (C.kind == 0) {
                sig[C.atom_index * kGenThreads + tid] = pb_interp_log(C.tab_x, C.tab_y, C.n_tab, sl, C.offset);
            } else {
                const MolTable t = pb_mol_table(C);
                const bool inside = (S.T >= t.gT[0] && S.T <= t.gT[t.nT - 1] && sl >= t.gW[0] && sl <= t.gW[t.nW - 1]);
                int iT = 0, iW = 0;
                double fT = 0.0, fW = 0.0;
                if (inside) {
                    iT = pb_interval(t.gT, t.nT, S.T); iW = pb_interval(t.gW, t.nW, sl);
                    fT = (S.T - t.gT[iT]) / (t.gT[iT + 1] - t.gT[iT]);
                    fW = (sl - t.gW[iW]) / (t.gW[iW + 1] - t.gW[iW]);
                }
                for (int j = 0; j < t.nP; ++j) {
                    double u = t.fill;
                    if (inside) {
                        const double* v = t.values + ((int64_t)j * t.nT + iT) * t.nW + iW;
                        u = (v[0] * (1.0 - fW) + v[1] * fW) * (1.0 - fT) + (v[t.nW] * (1.0 - fW) + v[t.nW + 1] * fW) * fT;
                    } else {
                        u = nan("");   // marks "outside in T or lambda": every sample takes the fill value
                    }
                    U[(node0 + j) * kGenThreads + tid] = u;
                }
                node0 += t.nP;
            }
        }
    }

    const int m_lo = blockIdx.z * prm.tiles_per_split * kChordTile;
    int m_hi = m_lo + prm.tiles_per_split * kChordTile;
    if (m_hi > sc.M) m_hi = sc.M;
    const int64_t atom_stride = (int64_t)sc.n_orb * sc.Mp;
    const double kLn10 = 2.302585092994046;
    double acc = 0.0;
    int buf = 0;

    for (int m = m_lo; m < m_hi; ++m) {
        const int64_t row = (int64_t)p * sc.Mp + m;
        const double wv = sc.w_in[row];
        if (wv == 0.0) continue;   // blocked (CTA-uniform)
        // ---- cooperative records for this chord ---------------------------------------------------
        MolRecord* R = rec + (size_t)buf * rec_cap;
        int n_rec = 0;
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            if (C.kind != 1) continue;
            const DevScenario& S = sc.scen[C.scenario];
            const double* dens = sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x;
            for (int ix = tid; ix < sc.n_x; ix += kGenThreads) {
                const double n = dens[ix];
                const double P = fmax(n * kBoltzmann * S.T, 1e-4);                            // :927, :805
                MolRecord r;
                r.coef = n * C.chi * sc.delta_x;                                              // :932, :936
                r.pad = 0;
                if (P >= C.grid_P[0] && P <= C.grid_P[C.dims[0] - 1]) {
                    r.j = pb_interval(C.grid_P, C.dims[0], P);
                    r.t = (P - C.grid_P[r.j]) / (C.grid_P[r.j + 1] - C.grid_P[r.j]);
                } else {
                    r.j = 0; r.t = -1.0;
                }
                R[n_rec + ix] = r;
            }
            n_rec += sc.n_x;
        }
        __syncthreads();
        // ---- per-thread optical depth ----------------------------------------------------------------
        double tau = 0.0;
        for (int a = 0; a < sc.n_atoms; ++a)
            tau = fma(sc.ncol[a * atom_stride + row], sig[a * kGenThreads + tid], tau);        // :954
        {
            int node0 = 0, r0 = 0;
            for (int k = 0; k < sc.n_cons; ++k) {
                const DevConstituent& C = sc.cons[k];
                if (C.kind != 1) continue;
                const double* Uc = U + (size_t)node0 * kGenThreads + tid;
                const double fill_sigma = pb_exp_neg_tab<true>(C.fill * kLn10, tab_lane) - C.offset;
                double s = 0.0;
                for (int i = 0; i < sc.n_x; ++i) {
                    const MolRecord r = R[r0 + i];
                    const double u0 = Uc[r.j * kGenThreads], u1 = Uc[(r.j + 1) * kGenThreads];
                    double sg;
                    if (r.t < 0.0 || u0 != u0) sg = fill_sigma;
                    else sg = pb_exp_neg_tab<true>(((1.0 - r.t) * u0 + r.t * u1) * kLn10, tab_lane) - C.offset;   // :814
                    s = fma(r.coef, sg, s);
                }
                tau += s;
                node0 += C.dims[0];
                r0 += sc.n_x;
            }
        }
        double f = wv;
        if (per_chord_fstar) f *= pb_interp_log(sc.fstar_x, sc.fstar_y, sc.n_fstar, lam / sc.star_shift[m], 0.0);
        acc = fma(f, pb_exp_neg_tab<true>(-tau, tab_lane), acc);                                    // :1241
        buf ^= 1;   // the other record buffer is free: its readers passed the barrier above
    }

    if (!live) return;
    const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + (w - prm.wav_lo);
    if (prm.n_split == 1) prm.R[o] = acc / sc.fout[per_chord_fstar ? w : 0];                  // :1258
    else prm.partial[(int64_t)blockIdx.z * prm.n_phase * prm.n_wav_sel + o] = acc;
}

cudaError_t launch_transit_general(const DevScene* d_scene, const DevScene& h, const TransitParams& prm, int sm_count,
                                   cudaStream_t st, int* n_launch) {
    (void)sm_count;
    *n_launch = 0;
    int node_total = 0, n_mol = 0;
    for (int k = 0; k < h.n_cons; ++k)
        if (h.cons[k].kind == 1) { node_total += h.cons[k].dims[0]; ++n_mol; }
    if (node_total > kMaxMolNodes) return cudaErrorInvalidValue;
    if (h.fout_per_wav) {
        fout_table_kernel<<<(unsigned)((h.n_wav + kGenThreads - 1) / kGenThreads), kGenThreads, 0, st>>>(d_scene);
        ++*n_launch;
    }
    const int rec_cap = n_mol * h.n_x > 0 ? n_mol * h.n_x : 1;
    const size_t smem = ((size_t)kExpTabEntries * kExpTabCopies + ((size_t)node_total + h.n_atoms) * kGenThreads) * sizeof(double) +
                        2 * (size_t)rec_cap * sizeof(MolRecord);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(transit_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)((prm.n_wav_sel + kGenThreads - 1) / kGenThreads), (unsigned)prm.n_phase, (unsigned)prm.n_split);
    transit_general_kernel<<<grid, kGenThreads, smem, st>>>(d_scene, prm, rec_cap);
    ++*n_launch;
    return cudaGetLastError();
}

}  // namespace pb200
