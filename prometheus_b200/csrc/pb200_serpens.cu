// pb200_serpens.cu -- the producer of the tabulated (SERPENS) density: SURVEY 8(f) rank 4, the CUDA part of
// SerpensExosphere.addInterpolatedDensity (reference: pythonScripts/gasProperties.py:548-583):
//     n_histogram = np.histogramdd(particlePos, bins=[xBins, yBins, zBins])[0] * N / (n_particles * cellVolume)
//     n_histogram = scipy.ndimage.gaussian_filter(n_histogram, sigma)            (if sigma > 0)
// Binning is index work and is bit-exact: a particle falls into bin j = searchsorted(edges, v, 'right') - 1 of each
// axis, a value equal to the last edge goes to the last bin, anything outside is dropped (numpy's histogramdd).
// The filter follows scipy's correlate1d for symmetric kernels (centre term first, then the pairs from the
// outermost inwards, 'reflect' boundary), one pass per axis in the order x, y, z; the weights come from the host,
// computed with scipy's own expression.
#include "pb200_internal.cuh"

namespace pb200 {

// j = (number of edges <= v) - 1, with v == last edge moved into the last bin; -1 if outside
__device__ __forceinline__ int pb_hist_bin(const double* __restrict__ e, int n_edges, double v) {
    if (!(v >= e[0]) || !(v <= e[n_edges - 1])) return -1;   // NaN falls out here too
    int lo = 0, hi = n_edges;                                 // first index with e[idx] > v
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (e[mid] <= v) lo = mid + 1; else hi = mid;
    }
    int j = lo - 1;
    if (j == n_edges - 1) j = n_edges - 2;                    // on the rightmost edge
    return j;
}

__global__ void __launch_bounds__(256)
serpens_histogram_kernel(const double* __restrict__ pos, int64_t n, const double* __restrict__ xb, int nxb,
                         const double* __restrict__ yb, int nyb, const double* __restrict__ zb, int nzb,
                         unsigned int* __restrict__ counts) {
    const int ny = nyb - 1, nz = nzb - 1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int jx = pb_hist_bin(xb, nxb, pos[3 * i]);
        const int jy = pb_hist_bin(yb, nyb, pos[3 * i + 1]);
        const int jz = pb_hist_bin(zb, nzb, pos[3 * i + 2]);
        if ((jx | jy | jz) < 0) continue;
        atomicAdd(counts + ((int64_t)jx * ny + jy) * nz + jz, 1u);   // integer counts: order independent
    }
}

// counts * N / (n_particles * cellVolume), the reference's operation order (gasProperties.py:565-566)
__global__ void __launch_bounds__(256)
serpens_scale_kernel(const unsigned int* __restrict__ counts, int64_t n_cells, double N, double denom,
                     double* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_cells) out[i] = __ddiv_rn(__dmul_rn((double)counts[i], N), denom);
}

// scipy 'reflect': (d c b a | a b c d | d c b a)
__device__ __forceinline__ int pb_reflect(int i, int n) {
    while (i < 0 || i >= n) i = (i < 0) ? -i - 1 : 2 * n - 1 - i;
    return i;
}

// one axis of gaussian_filter: out[..., l, ...] = w[r] in[l] + sum_{k=r..1} (in[l-k] + in[l+k]) w[r-k]
__global__ void __launch_bounds__(256)
serpens_filter_axis_kernel(const double* __restrict__ in, double* __restrict__ out, int64_t n_cells, int len, int64_t stride,
                           const double* __restrict__ w, int radius) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cells) return;
    const int l = (int)((i / stride) % len);
    const double* __restrict__ line = in + (i - (int64_t)l * stride);
    double acc = __dmul_rn(line[(int64_t)l * stride], w[radius]);
    for (int k = radius; k >= 1; --k) {
        const double a = line[(int64_t)pb_reflect(l - k, len) * stride], b = line[(int64_t)pb_reflect(l + k, len) * stride];
        acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(a, b), w[radius - k]));
    }
    out[i] = acc;
}

// the two diagnostic sums the reference prints (gasProperties.py:569-577): all cells, and the cells outside the
// planetary disk but inside the stellar disk; one block, fixed order
__global__ void __launch_bounds__(1024)
serpens_sums_kernel(const double* __restrict__ n, int nx, int ny, int nz, const double* __restrict__ yp,
                    const double* __restrict__ zp, double Rp2, double Rs2, double* __restrict__ sums) {
    __shared__ double s_a[32], s_b[32];
    double a = 0.0, b = 0.0;
    const int64_t cells = (int64_t)nx * ny * nz;
    for (int64_t i = threadIdx.x; i < cells; i += blockDim.x) {
        const int iz = (int)(i % nz), iy = (int)((i / nz) % ny);
        const double v = n[i], r2 = yp[iy] * yp[iy] + zp[iz] * zp[iz];
        a += v;
        if (r2 > Rp2 && r2 < Rs2) b += v;
    }
    a = pb_warp_sum(a); b = pb_warp_sum(b);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a; s_b[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x < 32) {
        a = pb_warp_sum(s_a[threadIdx.x]); b = pb_warp_sum(s_b[threadIdx.x]);
        if (threadIdx.x == 0) { sums[0] = a; sums[1] = b; }
    }
}

void launch_serpens_histogram(const double* pos, int64_t n, const double* xb, int nxb, const double* yb, int nyb,
                              const double* zb, int nzb, unsigned int* counts, int sm_count, cudaStream_t st) {
    int64_t blocks = (n + 255) / 256;
    if (blocks > (int64_t)sm_count * 16) blocks = (int64_t)sm_count * 16;
    if (blocks < 1) blocks = 1;
    serpens_histogram_kernel<<<(unsigned)blocks, 256, 0, st>>>(pos, n, xb, nxb, yb, nyb, zb, nzb, counts);
}
void launch_serpens_scale(const unsigned int* counts, int64_t n_cells, double N, double denom, double* out, cudaStream_t st) {
    serpens_scale_kernel<<<(unsigned)((n_cells + 255) / 256), 256, 0, st>>>(counts, n_cells, N, denom, out);
}
void launch_serpens_filter_axis(const double* in, double* out, int64_t n_cells, int len, int64_t stride, const double* w,
                                int radius, cudaStream_t st) {
    serpens_filter_axis_kernel<<<(unsigned)((n_cells + 255) / 256), 256, 0, st>>>(in, out, n_cells, len, stride, w, radius);
}
void launch_serpens_sums(const double* n, int nx, int ny, int nz, const double* yp, const double* zp, double Rp2, double Rs2,
                         double* sums, cudaStream_t st) {
    serpens_sums_kernel<<<1, 1024, 0, st>>>(n, nx, ny, nz, yp, zp, Rp2, Rs2, sums);
}

// ---- allocation-free variant of the histogram (pb200_serpens_histogram_static) -----------------------------------
// The integer count grid of serpens_histogram_kernel is scratch.  Here it lives in a file-scope device array that
// every GPU which loads the library holds until exit (8 MiB), so the call needs no cudaMalloc / cudaFree: usable
// under CUDA graph capture, in serving loops and in harnesses that forbid allocation.  In host code the array's name
// is only a stand-in; its device address is looked up once per device with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned int g_serpens_static_counts[PB200_SERPENS_STATIC_MAX_CELLS];

int ctx_fail(pb200_ctx* c, int code, const char* msg);
int ctx_info(pb200_ctx* c, int* device, cudaStream_t* stream, int* sm_count);
void ctx_add_launches(pb200_ctx* c, int n);

}  // namespace pb200

#include <cstdio>
#include <mutex>

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
unsigned int* g_static_counts_addr[kMaxDevices] = {};   // device address of g_serpens_static_counts, per device
}  // namespace

extern "C" int pb200_serpens_histogram_static(pb200_ctx* c, const double* pos_dev, int64_t n, const double* xb_dev,
                                              int32_t nxb, const double* yb_dev, int32_t nyb, const double* zb_dev,
                                              int32_t nzb, double N, double denom, double* out_dev) {
    using namespace pb200;
    int device = 0, sm_count = 0;
    cudaStream_t st = nullptr;
    if (ctx_info(c, &device, &st, &sm_count)) return PB200_ERR_INVALID;
    if (n < 0 || (n && !pos_dev) || !xb_dev || !yb_dev || !zb_dev || nxb < 2 || nyb < 2 || nzb < 2 || !out_dev)
        return ctx_fail(c, PB200_ERR_INVALID, "pb200_serpens_histogram_static: bad arguments");
    const int64_t cells = (int64_t)(nxb - 1) * (nyb - 1) * (nzb - 1);
    char msg[256];
    if (cells > (int64_t)PB200_SERPENS_STATIC_MAX_CELLS) {
        snprintf(msg, sizeof msg, "pb200_serpens_histogram_static: %lld cells exceed the static scratch of %lld cells "
                 "(PB200_SERPENS_STATIC_MAX_CELLS); use pb200_serpens_density", (long long)cells,
                 (long long)PB200_SERPENS_STATIC_MAX_CELLS);
        return ctx_fail(c, PB200_ERR_INVALID, msg);
    }
    if (device < 0 || device >= kMaxDevices) return ctx_fail(c, PB200_ERR_INVALID, "pb200_serpens_histogram_static: device index out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { snprintf(msg, sizeof msg, "cudaSetDevice: %s", cudaGetErrorString(e)); return ctx_fail(c, PB200_ERR_CUDA, msg); }
    unsigned int* counts = nullptr;
    {   // once per device, with that device current
        std::lock_guard<std::mutex> lock(g_static_mutex);
        if (!g_static_counts_addr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_serpens_static_counts);
            if (e != cudaSuccess || !p) {
                snprintf(msg, sizeof msg, "cudaGetSymbolAddress(static histogram scratch): %s", cudaGetErrorString(e));
                return ctx_fail(c, PB200_ERR_CUDA, msg);
            }
            g_static_counts_addr[device] = static_cast<unsigned int*>(p);
        }
        counts = g_static_counts_addr[device];
    }
    // the array keeps the previous call's counts: the cells of this call start at zero
    e = cudaMemsetAsync(counts, 0, (size_t)cells * sizeof(unsigned int), st);
    if (e != cudaSuccess) { snprintf(msg, sizeof msg, "cudaMemsetAsync: %s", cudaGetErrorString(e)); return ctx_fail(c, PB200_ERR_CUDA, msg); }
    if (n) launch_serpens_histogram(pos_dev, n, xb_dev, nxb, yb_dev, nyb, zb_dev, nzb, counts, sm_count, st);
    launch_serpens_scale(counts, cells, N, denom, out_dev, st);
    ctx_add_launches(c, n ? 2 : 1);
    e = cudaGetLastError();
    if (e != cudaSuccess) { snprintf(msg, sizeof msg, "static histogram launch: %s", cudaGetErrorString(e)); return ctx_fail(c, PB200_ERR_CUDA, msg); }
    return PB200_OK;
}
