// pb200_internal.cuh -- device-side scene description shared by the kernels and the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/prometheus_b200.h"
#include "pb200_math.cuh"

namespace pb200 {

constexpr int kMaxScenarios = PB200_MAX_SCENARIOS;
constexpr int kMaxConstituents = PB200_MAX_CONSTITUENTS;
constexpr int kMaxMoons = PB200_MAX_MOONS;
constexpr int kChordTile = 256;   // chords per shared-memory tile of kernel (c'); rows are padded to it

struct DevScenario {
    int type, n_cons, first_cons, has_molecule;
    double T;
    double p[8];
    const double* p_phase;   // NULL or [n_orb][8]: per-(virtual-)phase parameters of a retrieval batch (pb200_set_models);
                             // slot 7 carries the model's temperature T (no scenario uses more than p[0..3])
    const double* centre_x;
    const double* centre_y;
    const double* shift;
    const double* amplitude;
    const double* gx; const double* gy; const double* gz; const double* values;
    int dims[3];
    int density_slot;   // index into DevScene::density (molecular scenarios) or -1
};

struct DevConstituent {
    int kind, scenario, atom_index, mol_index;
    double chi, offset, fill;
    const double* tab_x; const double* tab_y; int64_t n_tab;
    // bucket index of tab_x (atoms, built on the device before the first run of a scene): bucket[b] = smallest j with
    // pb_bucket_of(tab_x[j]) >= b (n_tab if none), b = 0 .. n_bucket; brackets the binary search of sigma_phase_kernel
    const int* bucket; int n_bucket; int pad2;
    double bucket_x0, bucket_inv;
    const double* grid_P; const double* grid_T; const double* grid_w; const double* values;
    int dims[3];
    int pad;
};

// log10(sigma + offset) of an atomic constituent at wavelength sl (np.interp semantics), through the bucket index of
// its table when it has one: the same index, hence the same bits, as the full binary search
__device__ __forceinline__ double pb_cons_log_sigma(const DevConstituent& C, double sl) {
    if (C.bucket) {   // the answer j (largest with tab_x[j] <= sl) lies in [bucket[b] - 1, bucket[b + 1])
        const int b = pb_bucket_of(sl, C.bucket_x0, C.bucket_inv, C.n_bucket);
        return pb_interp_bracket(C.tab_x, C.tab_y, C.n_tab, sl, (int64_t)__ldg(C.bucket + b) - 1, __ldg(C.bucket + b + 1));
    }
    return pb_interp(C.tab_x, C.tab_y, C.n_tab, sl);
}

struct DevScene {
    int n_x, n_rho, n_phi, n_orb;
    int n_scen, n_cons, n_moons, n_atoms, n_mols;
    int M;            // chords per phase  = n_phi * n_rho
    int Mp;           // M padded to a multiple of kChordTile
    int fout_per_wav; // 1: fout is [n_wav] (per-chord F* table with stellar rotation), 0: fout is [1]
    int moon_R_per_phase;   // 1: moon_R is [n_moons][n_orb] (retrieval batch), 0: [n_moons]
    // Absorbers of the run path (atomic scenes): the atomic constituents of ONE scenario share its density, so
    // tau = sum_k Ncol_k sigma_k = (sum_x n dx) * sum_k chi_k sigma_k -- one absorber per scenario with the
    // effective cross section sum_k chi_k sigma_k(p, lambda) instead of one per constituent (C5: 2 -> 1, C2a: 8 -> 1).
    // abs_first/abs_count: constituent range of absorber g.  PB200_NO_GROUP=1: one absorber per constituent.
    int n_abs;
    int abs_first[kMaxConstituents], abs_count[kMaxConstituents];
    int n_orb_model;        // phases per model (== n_orb without a batch); row m*n_orb_model + p is model m, phase p
    int64_t n_wav;
    double delta_x, star_R, clv_u1, clv_u2, vsini, planet_R;
    const double* x; const double* rho; const double* sin_phi; const double* cos_phi;
    const double* cos_phi_rot; const double* wavelength;
    const double* fstar_x; const double* fstar_y; int64_t n_fstar;
    const double* planet_y; const double* moon_R; const double* moon_y;
    const double* planet_R_phase;   // NULL, or [n_orb]: planet radius of the model a (virtual) phase belongs to
    DevScenario scen[kMaxScenarios];
    DevConstituent cons[kMaxConstituents];
    // derived, device-resident, phase-major ("m" = i_phi*n_rho + i_rho):
    double* w_out;        // [Mp]            rho*clv, 0 in the padding
    double* star_shift;   // [Mp]            Doppler factor of the stellar surface element
    double* w_in;         // [n_orb][Mp]     rho*clv, 0 where blocked by planet/moon or padding
    double* ncol;         // [n_atoms][n_orb][Mp]  column densities
    double* density[kMaxScenarios];  // [n_orb][M][n_x] number density of molecular scenarios
    double* fout;         // [1] or [n_wav]  sum of unattenuated flux
    // chord records of kernel (c'): [n_orb][<= Mp][pb_rec_stride(n_atoms)] = {w_in, w_in c1, w_in c2, Ncol_0 .. Ncol_{S-1}[, 0]}, one TMA
    // bulk copy per 256-chord tile
    double* rec;
    const double* exp_tab;   // [256] 2^(j/256) then [4096] 2^(j/4096), correctly rounded, high words biased (pb200_math.cuh),
                             // then [256] {1/m_j, -log2(1/m_j)} of pb_log2_tab
    unsigned long long* nmax;   // [n_atoms][n_orb] bit pattern of max_m Ncol (>= 0), set by kernel (a')
    // transparent-chord culling (compact_kernel): a chord whose optical depth stays below 2^-54 at every
    // wavelength of the run contributes exactly w_in (exp(-tau) rounds to 1.0), so only the other
    // ("active") chords get a record; per phase they are packed at the front of the phase's rec row
    unsigned long long* smax;   // [n_atoms][n_orb] bit pattern of max_w sigma_a(p, w) over the run's wavelengths
    int mirror;                 // 1: chords (phi, rho) and (pi - phi, rho) share one record (see compact_kernel)
    int* n_act;                 // [n_orb] 256-chord tiles of active records
    double* w0;                 // [n_orb] sum of w_in over the transparent chords
    double* thin_tot;           // [n_orb][8] sums over the ACTIVE chords of {w, N0 w c1, N1 w c1, N0^2 w c2, N0 N1 w c2,
                                //   N1^2 w c2}: what kernel (c') needs for a window in which every chord is thin (S <= 2)
    unsigned long long* stat;   // [2]: [0] low word = tabulated-density out-of-grid flag, [1] active records of the run
    // [n_cons] order-preserving bit pattern of max(tab_y) of every atomic table (bucket_build_kernel, once per scene; 0:
    // not known): the column kernel bounds sigma with it and so does not wait for sigma_phase_kernel
    unsigned long long* tab_ymax;
};

// order-preserving map double -> uint64 (larger double, larger integer; 0 is below every double) and back
__host__ __device__ inline unsigned long long pb_ordered_bits(double v) {
    union { double d; unsigned long long u; } x; x.d = v;
    return (x.u >> 63) ? ~x.u : (x.u | 0x8000000000000000ull);
}
__host__ __device__ inline double pb_from_ordered_bits(unsigned long long b) {
    union { double d; unsigned long long u; } x;
    x.u = (b >> 63) ? (b & 0x7fffffffffffffffull) : ~b;
    return x.d;
}

// doubles per chord record of kernel (c'): {w, w c1, w c2, Ncol_0..Ncol_{S-1}} padded to an even count, so that a
// record is a whole number of 16-byte shared-memory loads
__host__ __device__ constexpr int pb_rec_stride(int n_atoms) { return (n_atoms + 4) & ~1; }

// kernel (c'): tiles per working CTA and number of working splits of a phase with n_act active tiles
__host__ __device__ inline int pb_tiles_per_split(int n_act, int n_split, int min_tiles) {
    const int tps = (n_act + n_split - 1) / n_split;
    return tps < min_tiles ? min_tiles : tps;
}
__host__ __device__ inline int pb_splits_used(int n_act, int n_split, int min_tiles) {
    const int tps = pb_tiles_per_split(n_act, n_split, min_tiles);
    return (n_act + tps - 1) / tps;
}

// global wavelength index of local index wi of a run
__host__ __device__ inline int64_t pb_wav_index(const struct TransitParams& prm, int64_t wi);

// launch parameters of kernel (c') and its variants
struct TransitParams {
    int phase_lo, n_phase;
    int64_t wav_lo, n_wav_sel;
    // block-cyclic wavelength selection (pb200_run_device_cyclic): local index wi stands for the global index
    // ((wi / cyc_block) * cyc_n + cyc_rank) * cyc_block + wi % cyc_block; cyc_block = 0: contiguous [wav_lo, wav_lo + n_wav_sel)
    int64_t cyc_block;
    int cyc_n, cyc_rank;
    int n_split;          // chord-axis splits (gridDim.z)
    int tiles_per_split;  // kChordTile-chord tiles per split (kernel (m); kernel (c') shares the ACTIVE tiles, see below)
    int min_tiles;        // kernel (c'): least number of active tiles a working CTA takes
    int culled;           // 1: kernel (c') (w0 / n_act valid), 0: kernel (m)
    double thin_v;        // kernel (c'): scaled-tau threshold of the thin path (kThinV; 0 switches it off)
    int centre;           // kernel (c'), sorted windows: 0 no Taylor sums about the centre, 1 whole windows, 2 also per 32-wavelength group
    int win;              // kernel (c'): wavelengths per warp window (32 * WPT)
    int n_win;            // windows per phase = ceil(n_wav_sel / win)
    int* order;           // [n_phase][n_win] window indices sorted by decreasing optical thickness (compact_kernel)
    double* winmax;       // [n_atoms][n_phase][n_win] max |nsig_a| of each window (sigma_phase_kernel)
    double* partial;      // [n_split][n_phase][n_wav_sel] sums of w_in*exp(-tau)  (n_split > 1); kernel (c') only writes
                          // the slots of splits that have tiles (pb_splits_used)
    double* R;            // [n_phase][n_wav_sel] final ratio, or (R_row_stride != 0) row 0 of the caller's FULL result:
                          // element (pi, wi) goes to R[pi * R_row_stride + global wavelength index] -- the tile is
                          // written in place into memory the device can reach (mapped host memory, a peer GPU)
    int64_t R_row_stride;
    double* nsig;         // [n_atoms][n_phase][n_wav_sel] -sigma_a(p, w) * 4096/ln2 (sigma_phase_kernel); with perm != NULL
                          // in the sorted order of each phase
    const int* perm;      // NULL, or [n_phase][n_wav_sel]: local wavelength index of sorted position i (pb200_sort.cu)
    int2* cell_range;     // NULL, or [n_phase] {absolute index of the first sigma cell, number of cells} (compact_kernel -> cell_*_kernel)
    unsigned* cell_count; // [n_phase][kCellCap / 32] arrival counters of cell_moment_kernel's chord-axis splits (zeroed by compact_kernel)
};

// Upper bound of absorber g's cross section at ANY wavelength: sum_k chi_k (10^max(tab_y_k) - offset_k) -- the table lookup is
// a convex combination of table values (tab_ymax: bucket_build_kernel).  +inf where a table holds a NaN or the sum is negative.
__device__ __forceinline__ double pb_abs_sigma_bound(const DevScene& sc, int g) {
    double sm = 0.0;
    for (int c = sc.abs_first[g]; c < sc.abs_first[g] + sc.abs_count[g]; ++c) {
        const unsigned long long b = sc.tab_ymax[c];
        const double ymax = (b == 0ull || b == ~0ull) ? __longlong_as_double(0x7ff0000000000000ll) : pb_from_ordered_bits(b);
        sm += sc.cons[c].chi * (exp10(ymax) - sc.cons[c].offset);
    }
    return (sm >= 0.0) ? sm * (1.0 + 1e-12) : __longlong_as_double(0x7ff0000000000000ll);
}

// Sigma cells (pb200_transit.cu): cell k holds the scaled cross sections a = |nsig| in [g^k, g^(k+1)), g = (1 + eps)/(1 - eps),
// eps = 3/40; Taylor sums of order kCellOrder about the centre (truncation < 7.2e-16 of a chord's weight, see cell_moment_kernel)
constexpr int kCellCap = 512;                               // cells per phase the arrays hold (g^512 = 2.6e33 in sigma)
constexpr int kCellOrder = 12;
constexpr int kCellSums = kCellOrder + 1;
constexpr double kCellPerOctave = 4.612303829032796;        // 1 / log2((1 + 3/40) / (1 - 3/40))
__device__ __forceinline__ int pb_cell_index(double a) { return (int)floor(log2(a) * kCellPerOctave); }
__device__ __forceinline__ double pb_cell_edge(int k) { return exp2((double)k * (1.0 / kCellPerOctave)); }
__host__ __device__ inline int64_t pb_wav_index(const TransitParams& prm, int64_t wi) {
    if (prm.cyc_block == 0) return prm.wav_lo + wi;
    return ((wi / prm.cyc_block) * prm.cyc_n + prm.cyc_rank) * prm.cyc_block + wi % prm.cyc_block;
}
// where element (pi, wi) of the run's tile goes in prm.R
__host__ __device__ inline int64_t pb_out_index(const TransitParams& prm, int pi, int64_t wi) {
    return prm.R_row_stride ? (int64_t)pi * prm.R_row_stride + pb_wav_index(prm, wi) : (int64_t)pi * prm.n_wav_sel + wi;
}


// ---------------------------------------------------------------------------------------------
// number density of one scenario at offsets (dx, dy, z) from its centre -- SURVEY 8(a) row a7
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double pb_heaviside1(double u) { return (u < 0.0) ? 0.0 : 1.0; }

// scalar parameters of a scenario at (virtual) phase p
__device__ __forceinline__ const double* pb_params(const DevScenario& s, int p) {
    return s.p_phase ? s.p_phase + (int64_t)p * 8 : s.p;
}

// temperature of a scenario at (virtual) phase p: P = n kB T, the T axis of molecular tables (gasProperties.py:927)
__device__ __forceinline__ double pb_scn_T(const DevScenario& s, int p) {
    return s.p_phase ? s.p_phase[(int64_t)p * 8 + 7] : s.T;
}

__device__ __forceinline__ double pb_density(const DevScenario& s, const double* __restrict__ sp, double amp, double dx,
                                             double dy, double z) {
    switch (s.type) {
        case PB200_SCN_BAROMETRIC: {   // gasProperties.py:155-160
            const double r = sqrt(dx * dx + dy * dy + z * z);
            return sp[0] * amp * exp((sp[1] - r) / sp[2]) * pb_heaviside1(r - sp[1]);
        }
        case PB200_SCN_HYDROSTATIC: {  // gasProperties.py:197-203 (non-zero inside the planet)
            const double r = sqrt(dx * dx + dy * dy + z * z);
            const double jeans0 = sp[2] / (sp[3] * sp[1]);
            const double jeans = sp[2] / (sp[3] * r) * pb_heaviside1(r - sp[1]);
            return sp[0] * amp * exp(jeans - jeans0);
        }
        case PB200_SCN_POWERLAW: {     // gasProperties.py:240-243, :326-329, :371-373, :453-460
            const double r = sqrt(dx * dx + dy * dy + z * z);
            return sp[0] * amp * pow(sp[1] / r, sp[2]) * pb_heaviside1(r - sp[1]);
        }
        case PB200_SCN_TORUS: {        // gasProperties.py:506-515, celestialBodies.py:438-470
            const double a = sqrt(dx * dx + dy * dy);
            const double qa = (a - sp[1]) / (4.0 * sp[2]);
            const double qz = z / sp[2];
            return sp[0] * amp * (exp(-(qa * qa)) * exp(-(qz * qz)));
        }
        default:
            return 0.0;
    }
}

// trilinear interpolation on a regular grid (scipy RegularGridInterpolator, method='linear'),
// used by the serpens scenario (gasProperties.py:580-600).  *oob is set when the point lies
// outside the grid (the reference raises there).
__device__ __forceinline__ int pb_grid_cell(const double* __restrict__ g, int n, double v, double* frac,
                                            int* oob) {
    if (v < g[0] || v > g[n - 1]) *oob = 1;
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (g[mid] <= v) lo = mid; else hi = mid;
    }
    // scipy: interval with g[i] <= v < g[i+1], clipped to [0, n-2]
    *frac = (v - g[lo]) / (g[lo + 1] - g[lo]);
    return lo;
}

__device__ __forceinline__ double pb_trilinear(const DevScenario& s, double px, double py, double pz,
                                               int* oob) {
    double fx, fy, fz;
    const int ix = pb_grid_cell(s.gx, s.dims[0], px, &fx, oob);
    const int iy = pb_grid_cell(s.gy, s.dims[1], py, &fy, oob);
    const int iz = pb_grid_cell(s.gz, s.dims[2], pz, &fz, oob);
    const int64_t sy = s.dims[2], sx = (int64_t)s.dims[1] * s.dims[2];
    const double* v = s.values + ix * sx + iy * sy + iz;
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int bx = (c >> 2) & 1, by = (c >> 1) & 1, bz = c & 1;
        const double wgt = (bx ? fx : 1.0 - fx) * (by ? fy : 1.0 - fy) * (bz ? fz : 1.0 - fz);
        acc += v[bx * sx + by * sy + bz] * wgt;
    }
    return acc;
}

}  // namespace pb200
