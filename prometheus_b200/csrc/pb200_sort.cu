// pb200_sort.cu -- wavelengths of a phase ordered by cross section, for kernel (c') with one absorber.
//
// A window of kernel (c') is 256 wavelengths that share the per-chord decisions "optically thin: expansion",
// "opaque: contributes nothing", "else: table exp".  On a grid where every 256 consecutive wavelengths hold a
// line core AND its wings (C5: 4096 lines on 300 306 wavelengths) those tests never fire, because the window's
// largest and smallest sigma differ by ten orders of magnitude.  R(p, lambda) is independent per wavelength, so the
// wavelengths of a phase may be processed in ANY order: sorted by sigma, a window is homogeneous and most
// (chord, window) pairs are decided without an exp.  One device-wide radix sort of 64-bit keys
// (phase << 44 | top 44 bits of the order-preserving image of nsig; the sort looks at 32 of them) with the local
// wavelength index as value, then a gather of nsig into the sorted order; kernel (c') scatters R back through the
// permutation.
#include <cub/device/device_radix_sort.cuh>

#include "pb200_internal.cuh"

namespace pb200 {

constexpr int kSortPhaseBits = 20;

__global__ void __launch_bounds__(256)
sort_keys_kernel(const double* __restrict__ nsig, int n_phase, int64_t n_wav_sel, unsigned long long* __restrict__ keys,
                 int* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)n_phase * n_wav_sel) return;
    const unsigned long long pi = (unsigned long long)(i / n_wav_sel);
    unsigned long long b = (unsigned long long)__double_as_longlong(nsig[i]);
    b ^= (b >> 63) ? ~0ull : 0x8000000000000000ull;          // ascending order of the doubles (NaN: somewhere, harmless)
    keys[i] = (pi << (64 - kSortPhaseBits)) | (b >> kSortPhaseBits);
    vals[i] = (int)(i % n_wav_sel);
}

__global__ void __launch_bounds__(256)
sort_gather_kernel(const double* __restrict__ nsig, const int* __restrict__ perm, int n_phase, int64_t n_wav_sel,
                   double* __restrict__ sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)n_phase * n_wav_sel) return;
    sorted[i] = nsig[(i / n_wav_sel) * n_wav_sel + perm[i]];
}

size_t sigma_sort_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (const int*)nullptr, (int*)nullptr, (int)n);
    return bytes;
}

// nsig [n_phase][n_wav_sel] -> perm (sorted position -> local wavelength index) and nsig in sorted order
cudaError_t launch_sigma_sort(const double* nsig, int n_phase, int64_t n_wav_sel, unsigned long long* keys_in,
                              unsigned long long* keys_out, int* vals_in, int* perm, void* temp, size_t temp_bytes,
                              double* nsig_sorted, cudaStream_t st) {
    const int64_t n = (int64_t)n_phase * n_wav_sel;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    sort_keys_kernel<<<blocks, 256, 0, st>>>(nsig, n_phase, n_wav_sel, keys_in, vals_in);
    // The order only has to make the windows homogeneous (kernel (c') takes every window's sigma range from the values
    // themselves), so the sort looks at the phase and the leading bits of sigma only: 32 key bits = 4 radix passes instead
    // of 8 (sign + exponent + >= 12 mantissa bits: neighbours in the order differ by < 2.5e-4 relative).  Stable, hence
    // deterministic.
    int phase_bits = 1;
    while ((1 << phase_bits) < n_phase) ++phase_bits;
    int sigma_bits = 32 - phase_bits;
    if (sigma_bits < 24) sigma_bits = 24;
    const int key_lo = (64 - kSortPhaseBits) - sigma_bits;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, perm, (int)n, key_lo,
                                                    (64 - kSortPhaseBits) + phase_bits, st);
    if (e != cudaSuccess) return e;
    sort_gather_kernel<<<blocks, 256, 0, st>>>(nsig, perm, n_phase, n_wav_sel, nsig_sorted);
    return cudaGetLastError();
}

}  // namespace pb200
