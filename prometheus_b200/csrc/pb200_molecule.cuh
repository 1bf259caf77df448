// pb200_molecule.cuh -- molecular cross-section table lookup.
//
// scipy.interpolate.RegularGridInterpolator(method='linear', bounds_error=False, fill_value=f)
// on log10(sigma + offset) over (P, T, lambda)  (reference: gasProperties.py:779-780, :814):
//   per axis  i = interval with g[i] <= v < g[i+1] clipped to [0, n-2],  t = (v - g[i])/(g[i+1]-g[i])
//   value     = sum over the 8 cell corners of values[corner] * prod(t or 1-t)
//   any coordinate outside [g[0], g[n-1]] -> fill_value
#pragma once
#include "pb200_internal.cuh"

namespace pb200 {

struct MolTable {
    const double* gP; const double* gT; const double* gW; const double* values;
    int nP, nT, nW;
    double fill;
};

__device__ __host__ __forceinline__ MolTable pb_mol_table(const DevConstituent& C) {
    MolTable t;
    t.gP = C.grid_P; t.gT = C.grid_T; t.gW = C.grid_w; t.values = C.values;
    t.nP = C.dims[0]; t.nT = C.dims[1]; t.nW = C.dims[2];
    t.fill = C.fill;
    return t;
}

// Rows of the per-thread columns V[row][thread] = ln sigma(P_node, T, shift * lambda) that kernel (m) keeps in shared
// memory.  The per-record form holds every pressure node; the moment form only the nodes the scene's chords reach
// (found once per scene by the count pass of mol_moment_kernel): C2's pressures span 9 of the 13 decades of its tables,
// and with the atoms of a scenario folded into one effective cross section that is the difference between two and three
// CTAs per SM.
constexpr int kMaxMols = 8;
struct MolLayout {
    int n_mol;
    int row0[kMaxMols];   // first row of molecule i
    int jmin[kMaxMols];   // pressure node held in row0[i]
    int cnt[kMaxMols];    // node rows held: nodes jmin .. jmin + cnt - 1; the molecule's fill row follows at row0 + cnt
    int rows;             // all rows; one zero pad row follows ("V_{j+1}" of the last fill row)
    int grouped;          // 1: sig[g] = sum_k chi_k sigma_k per absorber group (DevScene::n_abs); 0: sig[a] per atomic constituent
};

// interval index with g[i] <= v < g[i+1], clipped to [0, n-2]
__device__ __forceinline__ int pb_interval(const double* __restrict__ g, int n, double v) {
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(g + mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ double pb_mol_log_sigma(const MolTable& t, double P, double T, double lam) {
    if (!(P >= t.gP[0] && P <= t.gP[t.nP - 1] && T >= t.gT[0] && T <= t.gT[t.nT - 1] && lam >= t.gW[0] &&
          lam <= t.gW[t.nW - 1]))
        return t.fill;
    const int iP = pb_interval(t.gP, t.nP, P), iT = pb_interval(t.gT, t.nT, T), iW = pb_interval(t.gW, t.nW, lam);
    const double fP = (P - t.gP[iP]) / (t.gP[iP + 1] - t.gP[iP]);
    const double fT = (T - t.gT[iT]) / (t.gT[iT + 1] - t.gT[iT]);
    const double fW = (lam - t.gW[iW]) / (t.gW[iW + 1] - t.gW[iW]);
    const int64_t sT = t.nW, sP = (int64_t)t.nT * t.nW;
    const double* v = t.values + iP * sP + iT * sT + iW;
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int bP = (c >> 2) & 1, bT = (c >> 1) & 1, bW = c & 1;
        const double wgt = (bP ? fP : 1.0 - fP) * (bT ? fT : 1.0 - fT) * (bW ? fW : 1.0 - fW);
        acc += __ldg(v + bP * sP + bT * sT + bW) * wgt;
    }
    return acc;
}

}  // namespace pb200
