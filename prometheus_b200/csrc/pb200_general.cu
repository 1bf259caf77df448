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
#include <cstdlib>

namespace pb200 {

constexpr int kGenThreads = 256;
constexpr int kGenTabCopies = 4;   // exp table [256][4]: 8 KB, so that four CTAs fit one SM
constexpr int kGenWarps = kGenThreads / 32;
constexpr int kMaxMolNodes = 64;   // sum over molecules of (pressure nodes + 1) held in shared memory

// one (x sample, molecule) of the current chord; records with the same pressure node form runs
struct __align__(16) MolRecord {
    double t;      // fractional position inside the pressure interval (0 for the fill node)
    double coef;   // n * chi * delta_x
};
struct MolRun {
    int joff;      // node index (global over molecules) * 8 B * kGenThreads: byte offset of V_j in shared memory
    int start;     // first record of the run; the run ends where the next one starts
};

// coef * sigma of one record for this thread's wavelength: ln sigma = V_j + t (V_{j+1} - V_j), then the
// table exp (no clamp: ln sigma >= ln(1e-50) = -115).  1 + 8 + 1 FP64 instructions, one broadcast LDS.128.
__device__ __forceinline__ double mol_term(uint32_t rec_addr, double v, double d, uint32_t tab_lane) {
    double t, coef;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(t), "=d"(coef) : "r"(rec_addr));
    return coef * pb_exp_neg_tab<false, kGenTabCopies>(fma(t, d, v), tab_lane);
}

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

// Prologue shared by the two forms of kernel (m): consumers fill sig[a][thread] (atomic sigma at this thread's
// wavelength) and the molecular columns V[node][thread] = ln sigma(P_node, T, shift*lambda) (T and lambda already
// interpolated; one extra "fill" node per molecule, one zero pad row at the end); the producer warp copies the
// pressure grids to pgrid (may be NULL).
__device__ __forceinline__ void pb_general_prologue(const DevScene& sc, int p, double lam, bool producer, int tid, int lane,
                                                    double* V, double* sig, double* pgrid, const MolLayout& lay) {
    const double kLn10 = 2.302585092994046;
    if (!producer) {
        int imol = 0;
        if (lay.grouped) {   // one effective cross section per absorber group: tau_atoms = (sum_x n dx) * sum_k chi_k sigma_k
            for (int g = 0; g < sc.n_abs; ++g) {
                double v = 0.0;
                for (int k = sc.abs_first[g]; k < sc.abs_first[g] + sc.abs_count[g]; ++k) {
                    const DevConstituent& C = sc.cons[k];
                    const double sl = sc.scen[C.scenario].shift[p] * lam;                     // :917
                    v = fma(C.chi, exp10(pb_cons_log_sigma(C, sl)) - C.offset, v);
                }
                sig[g * kGenThreads + tid] = v;
            }
        }
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            const DevScenario& S = sc.scen[C.scenario];
            const double sl = S.shift[p] * lam;                                               // :917
            if (C.kind == 0) {
                if (!lay.grouped) sig[C.atom_index * kGenThreads + tid] = exp10(pb_cons_log_sigma(C, sl)) - C.offset;
            } else {
                const MolTable t = pb_mol_table(C);
                const double Tk = pb_scn_T(S, p);
                // outside the table in T or lambda: every sample takes the fill value (:779 bounds_error=False)
                const bool inside = (Tk >= t.gT[0] && Tk <= t.gT[t.nT - 1] && sl >= t.gW[0] && sl <= t.gW[t.nW - 1]);
                int iT = 0, iW = 0;
                double fT = 0.0, fW = 0.0;
                if (inside) {
                    iT = pb_interval(t.gT, t.nT, Tk); iW = pb_interval(t.gW, t.nW, sl);
                    fT = (Tk - t.gT[iT]) / (t.gT[iT + 1] - t.gT[iT]);
                    fW = (sl - t.gW[iW]) / (t.gW[iW + 1] - t.gW[iW]);
                }
                const int row0 = lay.row0[imol], jmin = lay.jmin[imol], cnt = lay.cnt[imol];
                for (int r = 0; r < cnt; ++r) {
                    const int j = jmin + r;
                    double u = t.fill;
                    if (inside) {
                        const double* v = t.values + ((int64_t)j * t.nT + iT) * t.nW + iW;
                        u = (v[0] * (1.0 - fW) + v[1] * fW) * (1.0 - fT) + (v[t.nW] * (1.0 - fW) + v[t.nW + 1] * fW) * fT;
                    }
                    u *= kLn10;
                    V[(row0 + r) * kGenThreads + tid] = u;
                }
                V[(row0 + cnt) * kGenThreads + tid] = t.fill * kLn10;       // fill node (records point at it with t = 0)
                ++imol;
            }
        }
        V[(size_t)lay.rows * kGenThreads + tid] = 0.0;     // pad row: "V_{j+1}" of the last fill node
    } else if (pgrid) {
        int node0 = 0;
        for (int k = 0; k < sc.n_cons; ++k) {
            const DevConstituent& C = sc.cons[k];
            if (C.kind != 1) continue;
            for (int j = lane; j < C.dims[0]; j += 32) pgrid[node0 + j] = C.grid_P[j];
            node0 += C.dims[0] + 1;
        }
    }
}

// named barriers (id 0 is __syncthreads): producer warp <-> 4 consumer warps, two chord buffers
__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
constexpr int kBarFull = 1, kBarEmpty = 3;
constexpr int kGenBlock = kGenThreads + 32;   // + one producer warp

// Warp-specialised: warp 4 (producer) walks the chords of this CTA's split, skips blocked ones, and
// turns each chord's density rows into a shared-memory item
//     { w_in, star_shift, valid flag | Ncol_a | sum of coef per molecule | records {t_P, n chi dx, node} }
// (global loads, pressure-interval search and the division happen here, off the consumers' path);
// warps 0-3 (consumers, one wavelength per thread) wait on the item and sum
//     tau = sum_a Ncol_a sigma_a + sum_records coef * exp(V_j + t (V_{j+1} - V_j)) - offset * sum coef.
// Shared memory: exp table [256][4] | V[node_total + 1][kGenThreads], V = ln sigma at pressure node j
// (T and lambda already interpolated for this thread's wavelength; every molecule has one extra
// "fill" node ln(fill sigma); one zero pad row at the end) | sig[n_atoms][kGenThreads] |
// pgrid[node_total] | item[2]
__global__ void __launch_bounds__(kGenBlock)
transit_general_kernel(const DevScene* __restrict__ scp, const TransitParams prm, int item_doubles, int node_total, int n_mol,
                       const MolLayout lay) {
    const DevScene& sc = *scp;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* etab = reinterpret_cast<double*>(smem_raw);
    double* V = etab + kExpTabEntries * kGenTabCopies;
    double* sig = V + (size_t)(node_total + 1) * kGenThreads;
    double* pgrid = sig + (size_t)sc.n_atoms * kGenThreads;
    double* items = pgrid + ((node_total + 1) & ~1);
    // item layout (doubles): [0] w_in  [1] star_shift  [2] valid (int), n_runs (int)  [4..) Ncol_a  [..) offset * sum coef
    //                        [run_off..) runs {joff, start} (n_rec + 1 slots)  [rec_off..) records {t, coef}
    const int n_rec_all = n_mol * sc.n_x;
    const int hdr = 4 + sc.n_atoms + 1;
    const int run_off = (hdr + 1) & ~1;
    const int rec_off = run_off + ((n_rec_all + 1 + 1) & ~1);

    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= kGenThreads;
    for (int i = tid; i < kExpTabEntries * kGenTabCopies; i += kGenBlock) etab[i] = __ldg(sc.exp_tab + i / kGenTabCopies);
    const uint32_t tab_lane = (uint32_t)__cvta_generic_to_shared(etab) + (tid & (kGenTabCopies - 1)) * 8u;
    const int p = prm.phase_lo + blockIdx.y;
    const int64_t w = prm.wav_lo + (int64_t)blockIdx.x * kGenThreads + tid;
    const bool live = !producer && w < prm.wav_lo + prm.n_wav_sel;
    const double lam = live ? sc.wavelength[w] : sc.wavelength[prm.wav_lo];
    const bool per_chord_fstar = sc.fout_per_wav != 0;

    pb_general_prologue(sc, p, lam, producer, tid, lane, V, sig, pgrid, lay);
    __syncthreads();

    // Mirror pairing (sc.mirror, see compact_kernel): the chords (phi, rho) and (pi - phi, rho) have equal densities
    // along x, so one item with weight w_in(i) + w_in(i') stands for both; the walk is over the pair index q.
    const int n2 = sc.n_phi / 2, Pm = (n2 + 1) / 2;
    const int n_items = sc.mirror ? 2 * Pm * sc.n_rho : sc.M;
    const int m_lo = blockIdx.z * prm.tiles_per_split * kChordTile;
    int m_hi = m_lo + prm.tiles_per_split * kChordTile;
    if (m_hi > n_items) m_hi = n_items;
    const int64_t atom_stride = (int64_t)sc.n_orb * sc.Mp;

    if (producer) {
        // ======================= producer warp ==================================================
        int it = 0;
        for (int q = m_lo; q <= m_hi; ++q) {
            const bool end = (q >= m_hi);
            int64_t row = 0;
            double wv = 0.0;
            int m = q;
            if (!end) {
                if (sc.mirror) {
                    const int h = q / (Pm * sc.n_rho), rem = q - h * (Pm * sc.n_rho);
                    const int i = rem / sc.n_rho, irho = rem - i * sc.n_rho, ip = n2 - 1 - i;
                    m = (h * n2 + i) * sc.n_rho + irho;
                    wv = sc.w_in[(int64_t)p * sc.Mp + m];
                    if (ip != i) {
                        const int mb = (h * n2 + ip) * sc.n_rho + irho;
                        const double wb = sc.w_in[(int64_t)p * sc.Mp + mb];
                        if (wv == 0.0) m = mb;   // the density row of a blocked chord is zero: take the mirror's
                        wv += wb;
                    }
                } else {
                    wv = sc.w_in[(int64_t)p * sc.Mp + m];
                }
                row = (int64_t)p * sc.Mp + m;
                if (wv == 0.0) continue;   // blocked: F_in = 0 (gasProperties.py:1233)
            }
            const int b = it & 1;
            if (it >= 2) bar_sync(kBarEmpty + b, kGenBlock);
            double* item = items + (size_t)b * item_doubles;
            if (lane == 0) {
                item[0] = wv;
                item[1] = (!end && per_chord_fstar) ? sc.star_shift[m] : 1.0;
                reinterpret_cast<int*>(item + 2)[0] = end ? 0 : 1;
            }
            if (!end) {
                for (int a = lane; a < sc.n_atoms; a += 32) item[4 + a] = sc.ncol[a * atom_stride + row];
                MolRecord* R = reinterpret_cast<MolRecord*>(item + rec_off);
                MolRun* runs = reinterpret_cast<MolRun*>(item + run_off);
                int n_rec = 0, node0 = 0, n_runs = 0, prev_joff = -1;
                double off_sum = 0.0;
                for (int k = 0; k < sc.n_cons; ++k) {
                    const DevConstituent& C = sc.cons[k];
                    if (C.kind != 1) continue;
                    const DevScenario& S = sc.scen[C.scenario];
                    const double* dens = sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x;
                    const int nP = C.dims[0];
                    const double* g = pgrid + node0;
                    const double kT = kBoltzmann * pb_scn_T(S, p), g0 = g[0], g1 = g[nP - 1];
                    double part = 0.0;
                    for (int ix0 = 0; ix0 < sc.n_x; ix0 += 32) {
                        const int ix = ix0 + lane;
                        const bool have = ix < sc.n_x;
                        int joff = -2;
                        if (have) {
                            const double n = dens[ix];
                            const double P = fmax(n * kT, 1e-4);                              // :927, :805
                            MolRecord r;
                            r.coef = n * C.chi * sc.delta_x;                                  // :932, :936
                            if (P >= g0 && P <= g1) {
                                int lo = 0, hi = nP - 1;
                                while (hi - lo > 1) {
                                    const int mid = (lo + hi) >> 1;
                                    if (g[mid] <= P) lo = mid; else hi = mid;
                                }
                                r.t = (P - g[lo]) / (g[lo + 1] - g[lo]);
                                joff = (node0 + lo) * kGenThreads * 8;
                            } else {
                                r.t = 0.0;
                                joff = (node0 + nP) * kGenThreads * 8;
                            }
                            part += r.coef;
                            R[n_rec + ix] = r;
                        }
                        // run boundaries of this 32-record chunk (records are in x order)
                        int before = __shfl_up_sync(0xffffffffu, joff, 1);
                        if (lane == 0) before = prev_joff;
                        const bool head = have && joff != before;
                        const unsigned heads = __ballot_sync(0xffffffffu, head);
                        if (head) {
                            const int idx = n_runs + __popc(heads & ((1u << lane) - 1u));
                            runs[idx].joff = joff;
                            runs[idx].start = n_rec + ix;
                        }
                        n_runs += __popc(heads);
                        const int last = min(31, sc.n_x - 1 - ix0);
                        prev_joff = __shfl_sync(0xffffffffu, joff, last);
                    }
                    part = pb_warp_sum(part);                  // fixed order: deterministic
                    off_sum = fma(C.offset, part, off_sum);
                    n_rec += sc.n_x;
                    node0 += nP + 1;
                    prev_joff = -1;                            // a new molecule always starts a new run
                }
                if (lane == 0) {
                    runs[n_runs].joff = 0;
                    runs[n_runs].start = n_rec;                // sentinel: end of the last run
                    item[4 + sc.n_atoms] = off_sum;
                    reinterpret_cast<int*>(item + 2)[1] = n_runs;
                }
            }
            __threadfence_block();
            bar_arrive(kBarFull + b, kGenBlock);
            ++it;
        }
        // the consumers' last "empty" arrival has no matching wait yet: drain it so the barrier is idle at exit
        if (it >= 2) bar_sync(kBarEmpty + ((it - 2) & 1), kGenBlock);
        return;
    }

    // ======================= consumer warps =====================================================
    const uint32_t vd_lane = (uint32_t)__cvta_generic_to_shared(V) + tid * 8u;
    double acc = 0.0;
    for (int it = 0;; ++it) {
        const int b = it & 1;
        bar_sync(kBarFull + b, kGenBlock);
        const double* item = items + (size_t)b * item_doubles;
        if (reinterpret_cast<const int*>(item + 2)[0] == 0) break;
        const double wv = item[0];
        double tau = 0.0;
        for (int a = 0; a < sc.n_atoms; ++a) tau = fma(item[4 + a], sig[a * kGenThreads + tid], tau);   // :954
        {
            const int n_runs = reinterpret_cast<const int*>(item + 2)[1];
            const MolRun* runs = reinterpret_cast<const MolRun*>(item + run_off);
            const uint32_t ra = (uint32_t)__cvta_generic_to_shared(item + rec_off);
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int start = runs[0].start;
            for (int q = 0; q < n_runs; ++q) {
                const int joff = runs[q].joff, stop = runs[q + 1].start;
                double v, d;   // this thread's V_j and V_{j+1} - V_j: loaded once per run instead of per record
                asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(vd_lane + (uint32_t)joff));
                asm volatile("ld.shared.f64 %0, [%1];" : "=d"(d) : "r"(vd_lane + (uint32_t)joff + kGenThreads * 8u));
                d -= v;
                int i = start;
                for (; i + 4 <= stop; i += 4) {
                    s0 += mol_term(ra + i * 16, v, d, tab_lane);
                    s1 += mol_term(ra + i * 16 + 16, v, d, tab_lane);
                    s2 += mol_term(ra + i * 16 + 32, v, d, tab_lane);
                    s3 += mol_term(ra + i * 16 + 48, v, d, tab_lane);
                }
                for (; i < stop; ++i) s0 += mol_term(ra + i * 16, v, d, tab_lane);
                start = stop;
            }
            tau += ((s0 + s1) + (s2 + s3)) - item[4 + sc.n_atoms];   // sum coef (sigma - offset), :814, :936
        }
        double f = wv;
        if (per_chord_fstar) f *= pb_interp_log(sc.fstar_x, sc.fstar_y, sc.n_fstar, lam / item[1], 0.0);
        bar_arrive(kBarEmpty + b, kGenBlock);   // every read of the item is done
        acc = fma(f, pb_exp_neg_tab<true, kGenTabCopies>(-tau, tab_lane), acc);                              // :1241
    }

    if (!live) return;
    const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + (w - prm.wav_lo);
    if (prm.n_split == 1) prm.R[pb_out_index(prm, blockIdx.y, w - prm.wav_lo)] = acc / sc.fout[per_chord_fstar ? w : 0];   // :1258
    else prm.partial[(int64_t)blockIdx.z * prm.n_phase * prm.n_wav_sel + o] = acc;
}

// =================================================================================================
// Moment form of kernel (m).
// Along a chord, the records of one molecule whose pressure falls into the same interval j (and the same of
// `bins` equal sub-intervals of t) share V_j(lambda), d_j(lambda) = V_{j+1} - V_j:
//     sum_rec coef exp(V_j + t d_j) = exp(V_j + tc d_j) * sum_k d_j^k [ sum_rec coef (t - tc)^k / k! ]
// The bracket -- moments of the records about the bin centre tc -- does not depend on the wavelength.  They are
// computed ONCE per chord (mol_moment_kernel, one thread per chord walking its density row) instead of once per
// CTA of 128 wavelengths, and a consumer thread then pays one table exp and kMolK FMAs per ITEM (a handful per
// chord and molecule: the pressure along a chord crosses a few decades) instead of per x sample.
// Truncation: |(t - tc) d| <= d_max / (2 bins) <= 0.19  =>  0.19^9 / 9! = 9e-13 relative on a sigma term;
// bins = ceil(d_max / 0.38) with d_max = max |V_{j+1} - V_j| over the table nodes (mol_dmax_kernel; interpolation in
// T and lambda is a convex combination of node values, so the node maximum bounds every thread's d_j).
// =================================================================================================
constexpr int kMolK = 8;
constexpr int kMolItemD = 12;   // doubles per item: {V byte offset (int), tc, M_0/0! .. M_8/8!, pad}
// Per (phase, chord pair) one BLOCK in global memory, blocks of consecutive pairs contiguous (stride_d doubles):
//   [0] weight w_in(i) + w_in(i') (0: blocked or padding -- skipped)   [1] star_shift   [2] n_items (int)
//   [3] offset * sum coef   [4 ..) Ncol_a   [hdr ..) items
// so that the producer warp of transit_moment_kernel moves kMolTile whole blocks with ONE sweep of independent
// coalesced loads: no dependent global load is left on its path (w_in -> chord -> header -> items took ~2400 cycles
// per chord when it walked the chords one by one).
__host__ __device__ inline int mol_hdr_doubles(int n_atoms) { return (4 + n_atoms + 1) & ~1; }

// d_max over the nodes of every molecular table: stat[0] = bit pattern of max |v[iP+1] - v[iP]| * ln 10
__global__ void __launch_bounds__(256) mol_dmax_kernel(const DevScene* __restrict__ scp, unsigned long long* __restrict__ stat) {
    const DevScene& sc = *scp;
    double mx = 0.0;
    for (int k = 0; k < sc.n_cons; ++k) {
        const DevConstituent& C = sc.cons[k];
        if (C.kind != 1) continue;
        const int64_t plane = (int64_t)C.dims[1] * C.dims[2], n = (int64_t)(C.dims[0] - 1) * plane;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
            const double d = fabs(__ldg(C.values + i + plane) - __ldg(C.values + i)) * 2.302585092994046;
            mx = (d == d) ? fmax(mx, d) : __longlong_as_double(0x7ff0000000000000ll);   // NaN in a table: no bound
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0.0) atomicMax(stat, (unsigned long long)__double_as_longlong(mx));
}

// One thread per (phase, pair slot).  COUNT: only the number of items (atomicMax into stat[1]); else the blocks.
// n_slots = pairs per phase rounded up to the tile; slots beyond the pairs get weight 0.
template <bool COUNT>
__global__ void __launch_bounds__(128)
mol_moment_kernel(const DevScene* __restrict__ scp, int phase_lo, int n_phase, int n_slots, int bins, double* __restrict__ blocks,
                  int stride_d, unsigned long long* __restrict__ stat, const MolLayout lay) {
    const DevScene& sc = *scp;
    const int64_t cid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int n_items = 0;
    bool live_block = false;
    if (cid < (int64_t)n_phase * n_slots) {
        const int pi = (int)(cid / n_slots), q = (int)(cid % n_slots), p = phase_lo + pi;
        const int n2 = sc.n_phi / 2, Pm = (n2 + 1) / 2;
        const int n_pairs = sc.mirror ? 2 * Pm * sc.n_rho : sc.M;
        double* out = COUNT ? nullptr : blocks + cid * stride_d;
        const int hdr = mol_hdr_doubles(sc.n_abs);
        double off_sum = 0.0, wv = 0.0;
        int m = q;
        if (q < n_pairs) {
            // mirror pairing (sc.mirror, see compact_kernel): chords (phi, rho) and (pi - phi, rho) have equal densities
            // along x; one block with weight w_in(i) + w_in(i') stands for both
            if (sc.mirror) {
                const int h = q / (Pm * sc.n_rho), rem = q - h * (Pm * sc.n_rho);
                const int i = rem / sc.n_rho, irho = rem - i * sc.n_rho, ip = n2 - 1 - i;
                m = (h * n2 + i) * sc.n_rho + irho;
                wv = sc.w_in[(int64_t)p * sc.Mp + m];
                if (ip != i) {
                    const int mb = (h * n2 + ip) * sc.n_rho + irho;
                    const double wb = sc.w_in[(int64_t)p * sc.Mp + mb];
                    if (wv == 0.0) m = mb;   // the density row of a blocked chord is zero: take the mirror's
                    wv += wb;
                }
            } else {
                wv = sc.w_in[(int64_t)p * sc.Mp + m];
            }
        }
        if (wv != 0.0) {   // blocked: F_in = 0 (gasProperties.py:1233)
            int imol = -1;
            live_block = true;
            for (int k = 0; k < sc.n_cons; ++k) {
                const DevConstituent& C = sc.cons[k];
                if (C.kind != 1) continue;
                const DevScenario& S = sc.scen[C.scenario];
                const double* __restrict__ dens = sc.density[S.density_slot] + ((int64_t)p * sc.M + m) * sc.n_x;
                const double* __restrict__ g = C.grid_P;
                const int nP = C.dims[0];
                const double kT = kBoltzmann * pb_scn_T(S, p), g0 = __ldg(g), g1 = __ldg(g + nP - 1);
                double part = 0.0, M[kMolK + 1], tc = 0.0;
                int cur = -1, cur_joff = 0;
                ++imol;
                int lo_min = 1 << 30, lo_max = -1;               // COUNT: pressure intervals this chord reaches
                const int row0 = lay.row0[imol], jmin = lay.jmin[imol], fill_row = lay.row0[imol] + lay.cnt[imol];
                auto flush = [&]() {
                    if (cur < 0) return;
                    if (!COUNT) {
                        double* it = out + hdr + (int64_t)n_items * kMolItemD;
                        it[0] = __hiloint2double(0, cur_joff);
                        it[1] = tc;
                        double fact = 1.0;
#pragma unroll
                        for (int j = 0; j <= kMolK; ++j) { it[2 + j] = M[j] / fact; fact *= (double)(j + 1); }
                        it[2 + kMolK + 1] = 0.0;
                    }
                    ++n_items;
                };
                for (int ix = 0; ix < sc.n_x; ++ix) {
                    const double n = dens[ix];
                    const double coef = n * C.chi * sc.delta_x;                                   // :932, :936
                    if (coef == 0.0) continue;
                    part += coef;
                    const double P = fmax(n * kT, 1e-4);                                          // :927, :805
                    int key, joff;
                    double t = 0.0, centre = 0.0;
                    if (P >= g0 && P <= g1) {
                        const int lo = pb_interval(g, nP, P);
                        const double ga = __ldg(g + lo);
                        t = (P - ga) / (__ldg(g + lo + 1) - ga);
                        int b = (int)(t * bins);
                        b = b < bins ? b : bins - 1;
                        centre = (b + 0.5) / bins;
                        key = lo * bins + b;
                        joff = (row0 + lo - jmin) * kGenThreads * 8;
                        lo_min = lo < lo_min ? lo : lo_min;
                        lo_max = lo > lo_max ? lo : lo_max;
                    } else {          // outside the pressure grid: the fill node, sigma does not depend on t
                        key = nP * bins;
                        joff = fill_row * kGenThreads * 8;
                    }
                    if (key != cur) {
                        flush();
                        cur = key; cur_joff = joff; tc = centre;
#pragma unroll
                        for (int j = 0; j <= kMolK; ++j) M[j] = 0.0;
                    }
                    const double u = t - tc;
                    double pw = coef;
#pragma unroll
                    for (int j = 0; j <= kMolK; ++j) { M[j] += pw; pw *= u; }
                }
                flush();
                off_sum = fma(C.offset, part, off_sum);
                if (COUNT && lo_max >= 0 && imol < kMaxMols) {   // (one atomic pair per chord and molecule, count pass only)
                    atomicMin(stat + 4 + 2 * imol, (unsigned long long)lo_min);
                    atomicMax(stat + 5 + 2 * imol, (unsigned long long)lo_max + 1ull);
                }
            }
        }
        if (!COUNT) {
            out[0] = wv;
            out[1] = (wv != 0.0 && sc.fout_per_wav) ? sc.star_shift[m] : 1.0;
            out[2] = __hiloint2double(0, n_items);
            out[3] = off_sum;
            const int64_t row = (int64_t)p * sc.Mp + m, atom_stride = (int64_t)sc.n_orb * sc.Mp;
            // one column per absorber group: sum_x n dx = Ncol_k / chi_k of a constituent with chi_k != 0
            for (int g = 0; g < sc.n_abs; ++g) {
                double N = 0.0;
                if (wv != 0.0)
                    for (int k = sc.abs_first[g]; k < sc.abs_first[g] + sc.abs_count[g]; ++k)
                        if (sc.cons[k].chi != 0.0) { N = sc.ncol[sc.cons[k].atom_index * atom_stride + row] / sc.cons[k].chi; break; }
                out[4 + g] = N;
            }
        }
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n_items = max(n_items, __shfl_xor_sync(0xffffffffu, n_items, o));
        if ((threadIdx.x & 31) == 0 && n_items > 0) atomicMax(stat + 1, (unsigned long long)n_items);
    } else {   // run statistics: items and live blocks of the run (the work kernel (m) does per wavelength)
        int live = n_items > 0 || live_block ? 1 : 0, tot = n_items;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { tot += __shfl_xor_sync(0xffffffffu, tot, o); live += __shfl_xor_sync(0xffffffffu, live, o); }
        if ((threadIdx.x & 31) == 0 && live > 0) { atomicAdd(stat + 2, (unsigned long long)tot); atomicAdd(stat + 3, (unsigned long long)live); }
    }
}

// one item of the current chord for this thread's wavelength: exp(V_j + tc d_j) * sum_k d_j^k M_k / k!
// (the item is read with six broadcast 16-byte loads)
__device__ __forceinline__ double mol_item_term(const double* __restrict__ it, uint32_t vd_lane, uint32_t tab_lane) {
    const uint32_t ia = (uint32_t)__cvta_generic_to_shared(it);
    double f0, tc, m0, m1, m2, m3, m4, m5, m6, m7, m8, pad;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(f0), "=d"(tc) : "r"(ia));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(m0), "=d"(m1) : "r"(ia + 16u));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(m2), "=d"(m3) : "r"(ia + 32u));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(m4), "=d"(m5) : "r"(ia + 48u));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(m6), "=d"(m7) : "r"(ia + 64u));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(m8), "=d"(pad) : "r"(ia + 80u));
    const int joff = __double2loint(f0);
    double v, d;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(vd_lane + (uint32_t)joff));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(d) : "r"(vd_lane + (uint32_t)joff + kGenThreads * 8u));
    d -= v;
    const double e = pb_exp_neg_tab<false, kGenTabCopies>(fma(tc, d, v), tab_lane);
    double p = fma(m8, d, m7);
    p = fma(p, d, m6); p = fma(p, d, m5); p = fma(p, d, m4); p = fma(p, d, m3);
    p = fma(p, d, m2); p = fma(p, d, m1); p = fma(p, d, m0);
    return e * p;
}

// Same decomposition as transit_general_kernel (one consumer thread per wavelength, one producer warp, two
// buffers, named barriers), but the unit that moves is a TILE of `tile` consecutive pair blocks: the producer warp
// copies tile * stride_d doubles with one sweep of independent coalesced 16-byte loads.
// Shared memory: exp table | V[node_total + 1][kGenThreads] | sig[n_atoms][kGenThreads] | buf[2][tile * stride_d]
__global__ void __launch_bounds__(kGenBlock)
transit_moment_kernel(const DevScene* __restrict__ scp, const TransitParams prm, const double* __restrict__ blocks,
                      int stride_d, int n_slots, int tile, const MolLayout lay) {
    const DevScene& sc = *scp;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* etab = reinterpret_cast<double*>(smem_raw);
    double* V = etab + kExpTabEntries * kGenTabCopies;
    double* sig = V + (size_t)(lay.rows + 1) * kGenThreads;
    double* bufs = sig + (size_t)sc.n_abs * kGenThreads;
    const int hdr = mol_hdr_doubles(sc.n_abs);
    const int n_grp = sc.n_abs;
    const int buf_doubles = tile * stride_d;

    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= kGenThreads;
    for (int i = tid; i < kExpTabEntries * kGenTabCopies; i += kGenBlock) etab[i] = __ldg(sc.exp_tab + i / kGenTabCopies);
    const uint32_t tab_lane = (uint32_t)__cvta_generic_to_shared(etab) + (tid & (kGenTabCopies - 1)) * 8u;
    const int p = prm.phase_lo + blockIdx.y;
    const int64_t w = prm.wav_lo + (int64_t)blockIdx.x * kGenThreads + tid;
    const bool live = !producer && w < prm.wav_lo + prm.n_wav_sel;
    const double lam = live ? sc.wavelength[w] : sc.wavelength[prm.wav_lo];
    const bool per_chord_fstar = sc.fout_per_wav != 0;
    pb_general_prologue(sc, p, lam, producer, tid, lane, V, sig, nullptr, lay);
    __syncthreads();

    // this split's slots [q_lo, q_hi): whole tiles (n_slots and the split size are multiples of the tile)
    const int q_lo = blockIdx.z * prm.tiles_per_split * kChordTile;
    int q_hi = q_lo + prm.tiles_per_split * kChordTile;
    if (q_hi > n_slots) q_hi = n_slots;
    const int n_tiles = (q_hi > q_lo) ? (q_hi - q_lo) / tile : 0;
    const double* __restrict__ src0 = blocks + ((int64_t)blockIdx.y * n_slots + q_lo) * stride_d;

    if (producer) {
        for (int it = 0; it < n_tiles; ++it) {
            const int b = it & 1;
            if (it >= 2) bar_sync(kBarEmpty + b, kGenBlock);
            const double2* __restrict__ s2 = reinterpret_cast<const double2*>(src0 + (int64_t)it * buf_doubles);
            double2* d2 = reinterpret_cast<double2*>(bufs + (size_t)b * buf_doubles);
#pragma unroll 8
            for (int i = lane; i < buf_doubles / 2; i += 32) d2[i] = __ldg(s2 + i);
            __threadfence_block();
            bar_arrive(kBarFull + b, kGenBlock);
        }
        // the consumers' last "empty" arrivals have no matching wait yet: drain them so the barriers are idle at exit
        if (n_tiles >= 2) bar_sync(kBarEmpty + ((n_tiles - 2) & 1), kGenBlock);
        if (n_tiles >= 1) bar_sync(kBarEmpty + ((n_tiles - 1) & 1), kGenBlock);
        return;
    }

    const uint32_t vd_lane = (uint32_t)__cvta_generic_to_shared(V) + tid * 8u;
    double acc = 0.0;
    for (int it = 0; it < n_tiles; ++it) {
        const int b = it & 1;
        bar_sync(kBarFull + b, kGenBlock);
        const double* buf = bufs + (size_t)b * buf_doubles;
        for (int c = 0; c < tile; ++c) {
            const double* blk = buf + (size_t)c * stride_d;
            const double wv = blk[0];
            if (wv == 0.0) continue;   // blocked or padding (block-uniform branch)
            double tau = 0.0;
            for (int g = 0; g < n_grp; ++g) tau = fma(blk[4 + g], sig[g * kGenThreads + tid], tau);   // :954, atoms of a scenario as one absorber
            const int n_items = __double2loint(blk[2]);
            const double* items = blk + hdr;
            double s0 = 0.0, s1 = 0.0;   // (four interleaved items: 72 registers, 20.4 against 18.7 ms at C2)
            int i = 0;
            for (; i + 2 <= n_items; i += 2) {
                s0 += mol_item_term(items + (size_t)i * kMolItemD, vd_lane, tab_lane);
                s1 += mol_item_term(items + (size_t)(i + 1) * kMolItemD, vd_lane, tab_lane);
            }
            if (i < n_items) s0 += mol_item_term(items + (size_t)i * kMolItemD, vd_lane, tab_lane);
            tau += (s0 + s1) - blk[3];   // sum coef (sigma - offset), :814, :936
            double f = wv;
            if (per_chord_fstar) f *= pb_interp_log(sc.fstar_x, sc.fstar_y, sc.n_fstar, lam / blk[1], 0.0);
            acc = fma(f, pb_exp_neg_tab<true, kGenTabCopies>(-tau, tab_lane), acc);                         // :1241
        }
        bar_arrive(kBarEmpty + b, kGenBlock);   // every read of the buffer is done
    }

    if (!live) return;
    const int64_t o = (int64_t)blockIdx.y * prm.n_wav_sel + (w - prm.wav_lo);
    if (prm.n_split == 1) prm.R[pb_out_index(prm, blockIdx.y, w - prm.wav_lo)] = acc / sc.fout[per_chord_fstar ? w : 0];   // :1258
    else prm.partial[(int64_t)blockIdx.z * prm.n_phase * prm.n_wav_sel + o] = acc;
}

void launch_mol_dmax(const DevScene* d_scene, unsigned long long* stat, int sm_count, cudaStream_t st) {
    mol_dmax_kernel<<<sm_count * 4, 256, 0, st>>>(d_scene, stat);
}
void launch_mol_moments(const DevScene* d_scene, int n_slots, int phase_lo, int n_phase, int bins, double* blocks, int stride_d,
                        unsigned long long* stat, bool count, const MolLayout& lay, cudaStream_t st) {
    const int64_t n = (int64_t)n_phase * n_slots;
    const unsigned nb = (unsigned)((n + 127) / 128);
    if (count) mol_moment_kernel<true><<<nb, 128, 0, st>>>(d_scene, phase_lo, n_phase, n_slots, bins, nullptr, 0, stat, lay);
    else mol_moment_kernel<false><<<nb, 128, 0, st>>>(d_scene, phase_lo, n_phase, n_slots, bins, blocks, stride_d, stat, lay);
}

// Row layout of the V columns.  ranges == NULL: every pressure node of every molecule (per-record form); else
// ranges[2 i], ranges[2 i + 1] = first interval and last interval + 1 the chords of the scene reach in molecule i
// (0, 0: none) -- interval lo needs the nodes lo and lo + 1.
MolLayout mol_layout(const DevScene& h, const unsigned long long* ranges, int grouped) {
    MolLayout lay{};
    int row = 0, i = 0;
    for (int k = 0; k < h.n_cons && i < kMaxMols; ++k) {
        if (h.cons[k].kind != 1) continue;
        const int nP = h.cons[k].dims[0];
        int jmin = 0, cnt = nP;
        if (ranges) {
            const long long lo = (long long)ranges[2 * i], hi1 = (long long)ranges[2 * i + 1];
            if (hi1 <= 0 || lo >= nP) { jmin = 0; cnt = 0; }
            else { jmin = (int)lo; cnt = (int)(hi1 - lo) + 1; if (jmin + cnt > nP) cnt = nP - jmin; }
        }
        lay.row0[i] = row; lay.jmin[i] = jmin; lay.cnt[i] = cnt;
        row += cnt + 1;
        ++i;
    }
    lay.n_mol = i;
    lay.rows = row;
    lay.grouped = grouped;
    return lay;
}
int mol_block_stride(int n_groups, int max_items) { return mol_hdr_doubles(n_groups) + max_items * kMolItemD; }
// blocks per tile: a power of two <= 8 (divides kChordTile) within `budget` bytes for the two buffers, at least one
// block.  (Tiles of 2 blocks to fit four CTAs per SM instead of three were slower: 22.8 against 18.7 ms at C2.)
int mol_tile(int stride_d, size_t fixed_bytes, size_t budget) {
    int t = 8;   // (16: 25.7 ms, 32: 43.9 ms, 8: 18.7 ms at C2 -- occupancy matters more than barrier rounds)
    while (t > 1 && fixed_bytes + (size_t)2 * t * stride_d * sizeof(double) > budget) t >>= 1;
    return t;
}

cudaError_t launch_transit_moments(const DevScene* d_scene, const DevScene& h, const TransitParams& prm, const double* blocks,
                                   int stride_d, int n_slots, const MolLayout& lay, cudaStream_t st, int* n_launch) {
    *n_launch = 0;
    if (lay.rows > kMaxMolNodes || h.n_mols > kMaxMols) return cudaErrorInvalidValue;
    if (h.fout_per_wav) {
        fout_table_kernel<<<(unsigned)((h.n_wav + kGenThreads - 1) / kGenThreads), kGenThreads, 0, st>>>(d_scene);
        ++*n_launch;
    }
    const size_t fixed = (size_t)kExpTabEntries * kGenTabCopies * sizeof(double) +
                         (size_t)(lay.rows + 1) * kGenThreads * sizeof(double) + (size_t)h.n_abs * kGenThreads * sizeof(double);
    const int tile = mol_tile(stride_d, 0, 48 * 1024);
    const size_t smem = fixed + 2 * (size_t)tile * stride_d * sizeof(double);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(transit_moment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)((prm.n_wav_sel + kGenThreads - 1) / kGenThreads), (unsigned)prm.n_phase, (unsigned)prm.n_split);
    transit_moment_kernel<<<grid, kGenBlock, smem, st>>>(d_scene, prm, blocks, stride_d, n_slots, tile, lay);
    ++*n_launch;
    return cudaGetLastError();
}

cudaError_t launch_transit_general(const DevScene* d_scene, const DevScene& h, const TransitParams& prm, int sm_count,
                                   cudaStream_t st, int* n_launch) {
    (void)sm_count;
    *n_launch = 0;
    int node_total = 0, n_mol = 0;
    for (int k = 0; k < h.n_cons; ++k)
        if (h.cons[k].kind == 1) { node_total += h.cons[k].dims[0] + 1; ++n_mol; }
    if (node_total > kMaxMolNodes || n_mol > kMaxMols) return cudaErrorInvalidValue;
    if (h.fout_per_wav) {
        fout_table_kernel<<<(unsigned)((h.n_wav + kGenThreads - 1) / kGenThreads), kGenThreads, 0, st>>>(d_scene);
        ++*n_launch;
    }
    const MolLayout lay = mol_layout(h, nullptr, 0);   // every node; one sigma column per atomic constituent
    const int n_rec_all = n_mol * h.n_x;
    const int run_off = ((4 + h.n_atoms + 1) + 1) & ~1;
    const int item_doubles = run_off + ((n_rec_all + 1 + 1) & ~1) + n_rec_all * (int)(sizeof(MolRecord) / sizeof(double));
    const size_t smem = (size_t)kExpTabEntries * kGenTabCopies * sizeof(double) +
                        (size_t)(node_total + 1) * kGenThreads * sizeof(double) + (size_t)h.n_atoms * kGenThreads * sizeof(double) +
                        (size_t)((node_total + 1) & ~1) * sizeof(double) + 2 * (size_t)item_doubles * sizeof(double);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(transit_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)((prm.n_wav_sel + kGenThreads - 1) / kGenThreads), (unsigned)prm.n_phase, (unsigned)prm.n_split);
    transit_general_kernel<<<grid, kGenBlock, smem, st>>>(d_scene, prm, item_doubles, node_total, n_mol, lay);
    ++*n_launch;
    return cudaGetLastError();
}

}  // namespace pb200
