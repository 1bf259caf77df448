// pb200_retrieval.cu -- light-curve epilogue of a retrieval batch.
//
// The consumer sketched by the reference's mainRetrieval.py:69-93: after R(phase, lambda) of a forward model,
//     SEL[p, w]     = OR_l ( lambda_w >= centre_l*shift_p - bw/2  AND  lambda_w <= centre_l*shift_p + bw/2 )
//     lightcurve[p] = mean(R[p, SEL[p]]) / max(R[p, :])
// With the models of a batch stacked along the phase axis (pb200_set_models) a row of R is (model, phase) and
// one launch does every light curve of the batch; only n_models*n_orb doubles cross PCIe instead of R.
#include "pb200_internal.cuh"

namespace pb200 {

constexpr int kMaxBandLines = 16;
constexpr int kLcThreads = 256;

struct BandParams {
    int n_lines;
    int n_orb_model;   // phases per model: the band centres move with shift[row % n_orb_model]
    int phase_lo;      // first row of R_rows (row index of the stacked scene)
    double half_bw;    // filterBandwidth / 2., evaluated on the host like the reference does
    double centre[kMaxBandLines];
    const double* shift;
};

// One CTA per row.  The selection is index work and must be bit-exact: the band edges are formed with the
// reference's operation order (product, then -/+ half the bandwidth) without FMA contraction.  Sums run in a
// fixed order (thread-strided, warp shuffles, then the warps in order) => bit-reproducible.
__global__ void __launch_bounds__(kLcThreads)
lightcurve_kernel(const double* __restrict__ R_rows, const double* __restrict__ wavelength, int64_t n_wav, BandParams b,
                  double* __restrict__ out) {
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const double shift = b.shift[(b.phase_lo + row) % b.n_orb_model];
    double lo[kMaxBandLines], hi[kMaxBandLines];
#pragma unroll
    for (int l = 0; l < kMaxBandLines; ++l) {
        const double c = __dmul_rn(l < b.n_lines ? b.centre[l] : 0.0, shift);
        // unused slots select nothing (lo > hi)
        lo[l] = l < b.n_lines ? __dsub_rn(c, b.half_bw) : 1.0;
        hi[l] = l < b.n_lines ? __dadd_rn(c, b.half_bw) : 0.0;
    }
    const double* __restrict__ r = R_rows + (int64_t)row * n_wav;
    double sum = 0.0, mx = -__longlong_as_double(0x7ff0000000000000ll);
    long long cnt = 0;
    int nan_seen = 0;
    for (int64_t w = tid; w < n_wav; w += kLcThreads) {
        const double v = r[w], lam = wavelength[w];
        bool sel = false;
#pragma unroll
        for (int l = 0; l < kMaxBandLines; ++l) sel |= (lam >= lo[l]) & (lam <= hi[l]);
        if (sel) { sum += v; ++cnt; }
        nan_seen |= (v != v);
        mx = fmax(mx, v);
    }
    __shared__ double s_sum[kLcThreads / 32], s_max[kLcThreads / 32];
    __shared__ long long s_cnt[kLcThreads / 32];
    __shared__ int s_nan[kLcThreads / 32];
    sum = pb_warp_sum(sum);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        nan_seen |= __shfl_xor_sync(0xffffffffu, nan_seen, o);
    }
    if (lane == 0) { s_sum[wid] = sum; s_max[wid] = mx; s_cnt[wid] = cnt; s_nan[wid] = nan_seen; }
    __syncthreads();
    if (tid == 0) {
        double S = 0.0, M = s_max[0];
        long long C = 0;
        int N = 0;
        for (int k = 0; k < kLcThreads / 32; ++k) { S += s_sum[k]; M = fmax(M, s_max[k]); C += s_cnt[k]; N |= s_nan[k]; }
        if (N) M = __longlong_as_double(0x7ff8000000000000ll);   // np.max propagates NaN
        // np.mean of an empty selection is nan (0/0); then / max
        out[row] = __ddiv_rn(__ddiv_rn(S, (double)C), M);
    }
}

cudaError_t launch_lightcurve(const double* R_rows, const double* wavelength, int64_t n_wav, int n_rows, int phase_lo,
                              int n_orb_model, const pb200_band& band, const double* shift_dev, double* out,
                              cudaStream_t st) {
    if (band.n_lines < 0 || band.n_lines > kMaxBandLines) return cudaErrorInvalidValue;
    BandParams b;
    b.n_lines = band.n_lines; b.n_orb_model = n_orb_model; b.phase_lo = phase_lo;
    b.half_bw = band.bandwidth / 2.0;
    for (int l = 0; l < kMaxBandLines; ++l) b.centre[l] = l < band.n_lines ? band.line_centre[l] : 0.0;
    b.shift = shift_dev;
    lightcurve_kernel<<<n_rows, kLcThreads, 0, st>>>(R_rows, wavelength, n_wav, b, out);
    return cudaGetLastError();
}

}  // namespace pb200
