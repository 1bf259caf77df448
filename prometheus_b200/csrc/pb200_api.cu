// pb200_api.cu -- the C ABI of include/prometheus_b200.h: context, scene upload, orchestration.
// All compute is in the kernels of pb200_density.cu / pb200_sigma.cu / pb200_transit.cu /
// pb200_molecular.cu; there is no host fallback for any of it.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "pb200_internal.cuh"
#include "pb200_molecule.cuh"

namespace pb200 {


// launchers implemented in the kernel translation units
void launch_sky_cells(const DevScene*, int Mp, cudaStream_t);
void launch_fout_scalar(const DevScene*, cudaStream_t);
void launch_column(const DevScene*, int Mp, int phase_lo, int n_phase, int* oob, int sm_count, cudaStream_t, bool has_molecules,
                   int run_mode);
void launch_number_density_grid(const DevScene*, int64_t n_chords, int scenario, double* out, int* oob, int sm_count,
                                cudaStream_t);
void launch_number_density_thread(const DevScene*, int64_t n_chords, int n_x, int scenario, double* out, int sm_count,
                                  unsigned long long* next_tile, cudaStream_t);
void launch_number_density_chords(const DevScene*, int scenario, int64_t n, const double* y, const double* z,
                                  const double* cx, const double* cy, const double* amp, double* out, int* oob,
                                  int sm_count, cudaStream_t);
cudaError_t launch_transit(const DevScene*, int n_atoms, const TransitParams&, cudaStream_t);
cudaError_t launch_transit_general(const DevScene*, const DevScene& host_copy, const TransitParams&, int sm_count,
                                   cudaStream_t, int* n_launch);
void launch_sigma_phase(const DevScene*, const TransitParams&, cudaStream_t);
void launch_bucket_build(const DevScene*, int sm_count, cudaStream_t);
cudaError_t launch_compact(const DevScene*, const TransitParams&, bool cull, int chords_per_phase, cudaStream_t);
void launch_finalize(const DevScene*, const TransitParams&, int sm_count, cudaStream_t);
int cell_chord_splits(int sm_count, int n_phase, int n_tiles);
size_t cell_scratch_bytes(int n_phase, int n_cs);
cudaError_t launch_cell_moments(const DevScene* d_scene, const TransitParams& prm, int n_cs, void* scratch, cudaStream_t st);
cudaError_t launch_cell_eval(const DevScene* d_scene, const TransitParams& prm, int n_cs, void* scratch, cudaStream_t st);
void cell_bind(TransitParams& prm, void* scratch);
void launch_sigma_lookup(const double*, const double*, int64_t, double, const double*, int64_t, double*, int,
                         cudaStream_t);
void launch_sigma_u(const DevScene*, int n_atoms, int n_orb, int64_t wav_lo, int64_t n_wav_sel, double* out, int,
                    cudaStream_t);
void launch_voigt_sigma(const double*, const double*, const double*, int64_t, double, const double*, int64_t, double,
                        double*, double*, cudaStream_t);
void launch_voigt_profile(const double*, const double*, const double*, int64_t, double*, int, cudaStream_t);
void launch_molecular_sigma(const MolTable&, double offset, const double* P, double T, const double* wl, int64_t n_c,
                            int64_t n_x, int64_t n_wav, double* out, int sm_count, cudaStream_t);
void launch_optical_depth(const DevScene*, const double* sigma_u, int64_t n_chords, int64_t wav_lo, int64_t n_wav_sel,
                          double* tau, cudaStream_t);
cudaError_t launch_optical_depth_chords(const DevScene*, int n_x, int n_density_slots, int64_t n, const double* y, const double* z,
                                        const double* cx, const double* cy, const double* shift, const double* amp,
                                        int64_t wav_lo, int64_t n_wav_sel, double* tau, int* oob, cudaStream_t);
void launch_log_table(const double* in, int64_t n_rows, int64_t n_cols, int reverse, double offset, double* out, int sm_count,
                      cudaStream_t);
void launch_mol_dmax(const DevScene*, unsigned long long* stat, int sm_count, cudaStream_t);
void launch_mol_moments(const DevScene*, int n_slots, int phase_lo, int n_phase, int bins, double* blocks, int stride_d,
                        unsigned long long* stat, bool count, const MolLayout& lay, cudaStream_t);
int mol_block_stride(int n_groups, int max_items);
MolLayout mol_layout(const DevScene& host_copy, const unsigned long long* ranges, int grouped);
cudaError_t launch_transit_moments(const DevScene*, const DevScene& host_copy, const TransitParams&, const double* blocks,
                                   int stride_d, int n_slots, const MolLayout& lay, cudaStream_t, int* n_launch);
size_t sigma_sort_temp_bytes(int64_t n);
cudaError_t launch_sigma_sort(const double* nsig, int n_phase, int64_t n_wav_sel, unsigned long long* keys_in,
                              unsigned long long* keys_out, int* vals_in, int* perm, void* temp, size_t temp_bytes,
                              double* nsig_sorted, cudaStream_t);
cudaError_t launch_lightcurve(const double* R_rows, const double* wavelength, int64_t n_wav, int n_rows, int phase_lo,
                              int n_orb_model, const pb200_band& band, const double* shift_dev, double* out, cudaStream_t);
void launch_serpens_histogram(const double* pos, int64_t n, const double* xb, int nxb, const double* yb, int nyb,
                              const double* zb, int nzb, unsigned int* counts, int sm_count, cudaStream_t);
void launch_serpens_scale(const unsigned int* counts, int64_t n_cells, double N, double denom, double* out, cudaStream_t);
void launch_serpens_filter_axis(const double* in, double* out, int64_t n_cells, int len, int64_t stride, const double* w,
                                int radius, cudaStream_t);
void launch_serpens_sums(const double* n, int nx, int ny, int nz, const double* yp, const double* zp, double Rp2, double Rs2,
                         double* sums, cudaStream_t);
void launch_probe_dfma(double* out, int blocks, int iters, cudaStream_t);
void launch_probe_exp(double* out, int blocks, int iters, const double* exp_tab, cudaStream_t);

}  // namespace pb200

using namespace pb200;

static thread_local std::string g_create_error;

struct pb200_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    int64_t launches = 0;
    bool has_scene = false;
    bool general_path = false;   // molecules or a per-chord F* table: transit_general kernels
    DevScene h_scene;            // host copy, pointers are device pointers
    DevScene* d_scene = nullptr;
    void* arena = nullptr;       // one allocation holding every uploaded array
    size_t arena_bytes = 0;
    void* stage = nullptr;       // pinned host staging buffer for the upload
    size_t stage_bytes = 0;
    void* darena = nullptr;      // one allocation holding every derived array (w_out, rec, ncol, ...)
    size_t darena_bytes = 0;     // both arenas are grow-only: repeated pb200_set_scene calls do not cudaMalloc
    double* scratch = nullptr;   // partial sums / R tile / temporary outputs
    size_t scratch_bytes = 0;
    int* d_oob = nullptr;        // 16 bytes: [0] out-of-grid flag (int), [8..16) active chord records of the last run
    int64_t last_n_wav_sel = 0;
    int last_wpt = 8;
    double* d_exp_tab = nullptr; // 2^(j/64), j < 64
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // 6,7: scene upload
    cudaStream_t aux = nullptr;          // sigma_phase (+ sort) of a run goes here, beside the column kernel on `stream`
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    pb200_timing last = {};
    double upload_ms = 0.0;
    // retrieval batches (pb200_set_models): the scene as uploaded, the stacked per-phase arrays of the batch
    DevScene base;               // h_scene at the end of pb200_set_scene
    void* marena = nullptr;      // device: stacked per-phase arrays + per-phase scenario parameters
    size_t marena_bytes = 0;
    void* mstage = nullptr;      // pinned staging of the same
    size_t mstage_bytes = 0;
    bool mstage_busy = false;    // a copy out of mstage may still be in flight
    double* d_band = nullptr;    // device: band shift [n_orb_model] then light curve [rows]
    size_t band_bytes = 0;
    int lastR_phase_lo = 0, lastR_n_phase = 0;   // rows of R left by the last run (in `scratch`, in the caller's device
    bool lastR_full = false;                     // buffer or in mapped host memory) over the whole wavelength grid
    const double* lastR_ptr = nullptr;
    // moment form of kernel (m): per-chord items (grow-only), bins of the current scene (-1: not yet known)
    double* mol_items = nullptr;
    size_t mol_items_bytes = 0;
    unsigned long long* d_mol_stat = nullptr;   // [0] bit pattern of d_max, [1] largest item count of a chord, [2] items and
                                                // [3] live (phase, chord pair) blocks of the last fill pass
    int mol_bins = -1;
    int mol_max_items = -1;      // largest item count of a chord over ALL phases of the resident scene (-1: not yet counted)
    MolLayout mol_lay = {};      // rows of the V columns the scene's chords reach (same count pass)
    bool mol_stats_pending = false;   // the last run was kernel (m) in moment form: check_oob reads [2], [3]
    bool buckets_ready = false;   // bucket indices of the sigma tables (built on the device before the first run of a scene)
    std::vector<std::pair<void*, int64_t>> regions;   // sharded uploads of the current scene: (device base, doubles per rank)
    // pb200_prefetch_tables: the large page-locked arrays of the last pb200_set_scene (source, bytes, arena offset), and
    // those of them whose copy a prefetch has already put on the stream for the NEXT pb200_set_scene
    struct UpItem { const void* src; size_t bytes; size_t offset; };
    std::vector<UpItem> last_big, prefetched;
    void* prefetched_arena = nullptr;
    // sigma-sorted order of the last run (pb200_last_sort_order; lives in `scratch` until the next run)
    const int* last_perm = nullptr;
    const double* last_nsig_sorted = nullptr;
    int64_t last_sort_n = 0;
    // run-time options (pb200_set_option); the library never reads the environment
    int64_t opt[PB200_OPT_COUNT] = {1, 1, 1, 1, -1, 0, 1, 1, 3};
};

static int fail(pb200_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    return code;
}

#define PB_CUDA(c, call)                                                                          \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail((c), PB200_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                                      \
    } while (0)

static int ensure_scratch(pb200_ctx* c, size_t bytes) {
    if (bytes <= c->scratch_bytes) return PB200_OK;
    if (c->scratch) { PB_CUDA(c, cudaStreamSynchronize(c->stream)); PB_CUDA(c, cudaFree(c->scratch)); c->scratch = nullptr; c->scratch_bytes = 0; }
    PB_CUDA(c, cudaMalloc((void**)&c->scratch, bytes));
    c->scratch_bytes = bytes;
    return PB200_OK;
}

// page-locked blocks handed out by pb200_host_alloc: arrays inside them are copied to the device
// straight from where they lie (no staging copy), see pb200_set_scene
static std::mutex g_pinned_mutex;
struct PinnedBlock { size_t bytes; uintptr_t dev; bool registered; };   // dev: the block's address as the device sees it
static std::map<uintptr_t, PinnedBlock> g_pinned;
static bool is_pinned(const void* p, size_t bytes) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    auto it = g_pinned.upper_bound(a);
    if (it == g_pinned.begin()) return false;
    --it;
    return a >= it->first && a + bytes <= it->first + it->second.bytes;
}
// device-side address of [p, p + bytes) if it lies in a page-locked block of this library, else NULL
static double* device_view(const void* p, size_t bytes) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    auto it = g_pinned.upper_bound(a);
    if (it == g_pinned.begin()) return nullptr;
    --it;
    if (!(a >= it->first && a + bytes <= it->first + it->second.bytes) || !it->second.dev) return nullptr;
    return reinterpret_cast<double*>(it->second.dev + (a - it->first));
}

static void free_scene(pb200_ctx* c) {
    if (c->darena) { cudaFree(c->darena); c->darena = nullptr; c->darena_bytes = 0; }
    if (c->arena) { cudaFree(c->arena); c->arena = nullptr; c->arena_bytes = 0; }
    c->has_scene = false;
}

extern "C" {

int pb200_abi_version(void) { return PB200_ABI_VERSION; }

int pb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* pb200_last_error(pb200_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int pb200_create(int device, pb200_ctx** out) {
    if (!out) return fail(nullptr, PB200_ERR_INVALID, "pb200_create: out == NULL");
    *out = nullptr;
    int n = pb200_device_count();
    if (n <= 0) return fail(nullptr, PB200_ERR_NO_DEVICE, "no CUDA device visible; libprometheus_b200 has no CPU fallback");
    if (device < 0 || device >= n) return fail(nullptr, PB200_ERR_INVALID, "device %d out of range [0, %d)", device, n);
    pb200_ctx* c = new pb200_ctx();
    c->device = device;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) { c->own_stream = true; e = cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device); }
    for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreate(&c->ev[i]);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->aux, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_scene, sizeof(DevScene));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_oob, 64);   // [0] out-of-grid flag, [8..16) active records, [16..24) thick (warp, chord) pairs of the run,
                                                                 // [24..32) tile counter of kernel (a), [32..40) (chord, narrow window) pairs of the run
    if (e == cudaSuccess) e = cudaMemset(c->d_oob, 0, 64);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_exp_tab, (kExpTabEntries + kExpBigEntries + 2 * kLogTabEntries) * sizeof(double));
    if (e == cudaSuccess) {
        double tab[kExpTabEntries];
        for (int j = 0; j < kExpTabEntries; ++j) tab[j] = pb_exp_tab_entry((double)exp2l((long double)j / kExpTabEntries), j);
        e = cudaMemcpy(c->d_exp_tab, tab, sizeof tab, cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) {
        static double big[kExpBigEntries];
        for (int j = 0; j < kExpBigEntries; ++j) big[j] = pb_exp_big_entry((double)exp2l((long double)j / kExpBigEntries), j);
        e = cudaMemcpy(c->d_exp_tab + kExpTabEntries, big, sizeof big, cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) {   // pb_log2_tab: {inv, -log2(inv)} per 1/256 of the mantissa range
        static double lt[2 * kLogTabEntries];
        for (int j = 0; j < kLogTabEntries; ++j) {
            const double inv = 1.0 / (1.0 + (j + 0.5) / kLogTabEntries);
            lt[2 * j] = inv;
            lt[2 * j + 1] = (double)(-log2l((long double)inv));
        }
        e = cudaMemcpy(c->d_exp_tab + kExpTabEntries + kExpBigEntries, lt, sizeof lt, cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        int rc = fail(nullptr, PB200_ERR_CUDA, "pb200_create: %s", cudaGetErrorString(e));
        delete c;
        return rc;
    }
    memset(&c->h_scene, 0, sizeof(DevScene));
    *out = c;
    return PB200_OK;
}

void pb200_destroy(pb200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_scene(c);
    if (c->scratch) cudaFree(c->scratch);
    if (c->stage) cudaFreeHost(c->stage);
    if (c->mstage) cudaFreeHost(c->mstage);
    if (c->marena) cudaFree(c->marena);
    if (c->mol_items) cudaFree(c->mol_items);
    if (c->d_mol_stat) cudaFree(c->d_mol_stat);
    if (c->d_band) cudaFree(c->d_band);
    if (c->d_scene) cudaFree(c->d_scene);
    if (c->d_oob) cudaFree(c->d_oob);
    if (c->d_exp_tab) cudaFree(c->d_exp_tab);
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->aux) { cudaStreamSynchronize(c->aux); cudaStreamDestroy(c->aux); }
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int pb200_set_stream(pb200_ctx* c, void* cuda_stream) {
    if (!c) return PB200_ERR_INVALID;
    if (c->stream == (cudaStream_t)cuda_stream && !c->own_stream) return PB200_OK;
    PB_CUDA(c, cudaSetDevice(c->device));
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->own_stream && c->stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
    c->stream = (cudaStream_t)cuda_stream;
    return PB200_OK;
}

int64_t pb200_launch_count(pb200_ctx* c) { return c ? c->launches : 0; }

int pb200_set_option(pb200_ctx* c, int option, int64_t value) {
    if (!c) return PB200_ERR_INVALID;
    if (option < 0 || option >= PB200_OPT_COUNT) return fail(c, PB200_ERR_INVALID, "pb200_set_option: unknown option %d", option);
    const bool ok = option == PB200_OPT_SORT ? (value >= -1 && value <= 1)
                  : option == PB200_OPT_MIN_TILES ? (value >= 1 && value <= (1 << 20))
                  : option == PB200_OPT_CENTRE ? (value >= 0 && value <= 3)
                  : (value == 0 || value == 1);
    if (!ok) return fail(c, PB200_ERR_INVALID, "pb200_set_option: value %lld out of range for option %d", (long long)value, option);
    c->opt[option] = value;
    return PB200_OK;
}

int pb200_get_option(pb200_ctx* c, int option, int64_t* value) {
    if (!c || !value) return PB200_ERR_INVALID;
    if (option < 0 || option >= PB200_OPT_COUNT) return fail(c, PB200_ERR_INVALID, "pb200_get_option: unknown option %d", option);
    *value = c->opt[option];
    return PB200_OK;
}

int pb200_host_alloc(size_t bytes, void** out) {
    if (!out) return PB200_ERR_INVALID;
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
    if (e != cudaSuccess) return fail(nullptr, PB200_ERR_CUDA, "cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e));
    void* dev = nullptr;
    if (cudaHostGetDevicePointer(&dev, *out, 0) != cudaSuccess) { cudaGetLastError(); dev = nullptr; }
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned[reinterpret_cast<uintptr_t>(*out)] = PinnedBlock{bytes ? bytes : 1, reinterpret_cast<uintptr_t>(dev), false};
    }
    return PB200_OK;
}

int pb200_host_register(void* p, size_t bytes, void** dev_ptr) {
    if (!p || !bytes) return PB200_ERR_INVALID;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(nullptr, PB200_ERR_CUDA, "cudaHostRegister(%zu): %s", bytes, cudaGetErrorString(e)); }
    void* dev = nullptr;
    e = cudaHostGetDevicePointer(&dev, p, 0);
    if (e != cudaSuccess) { cudaGetLastError(); cudaHostUnregister(p); return fail(nullptr, PB200_ERR_CUDA, "cudaHostGetDevicePointer: %s", cudaGetErrorString(e)); }
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned[reinterpret_cast<uintptr_t>(p)] = PinnedBlock{bytes, reinterpret_cast<uintptr_t>(dev), true};
    }
    if (dev_ptr) *dev_ptr = dev;
    return PB200_OK;
}

int pb200_host_unregister(void* p) {
    if (!p) return PB200_OK;
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        auto it = g_pinned.find(reinterpret_cast<uintptr_t>(p));
        if (it == g_pinned.end() || !it->second.registered) return fail(nullptr, PB200_ERR_INVALID, "pb200_host_unregister: not a registered block");
        g_pinned.erase(it);
    }
    cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(nullptr, PB200_ERR_CUDA, "cudaHostUnregister: %s", cudaGetErrorString(e)); }
    return PB200_OK;
}

int pb200_host_free(void* p) {
    if (!p) return PB200_OK;
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned.erase(reinterpret_cast<uintptr_t>(p));
    }
    cudaError_t e = cudaFreeHost(p);
    if (e != cudaSuccess) return fail(nullptr, PB200_ERR_CUDA, "cudaFreeHost: %s", cudaGetErrorString(e));
    return PB200_OK;
}

}  // extern "C"

// =================================================================================================
// scene upload
// =================================================================================================
namespace {

struct ArenaPlan {
    struct Item { const void* src; size_t bytes; size_t offset; size_t chunk; };   // chunk != 0: sharded upload, bytes per rank
    std::vector<Item> items;
    size_t total = 0;
    // a large table of which every rank uploads one n_ranks-th; the ranks complete it with an all-gather
    // (pb200_set_scene_sharded): room for n_ranks whole chunks
    size_t add_sharded(const void* src, size_t bytes, int n_ranks) {
        const size_t per = (((bytes + n_ranks - 1) / n_ranks) + 255) & ~size_t(255);
        const size_t off = total;
        items.push_back({src, bytes, off, per});
        total += per * n_ranks;
        return off;
    }
    // returns the offset (in bytes) the array will have inside the arena
    size_t add(const void* src, size_t bytes) {
        const size_t off = total;
        items.push_back({src, bytes, off, 0});
        total += (bytes + 255) & ~size_t(255);   // 256-byte alignment: TMA bulk copies need 16
        return off;
    }
};

template <typename T>
const T* at(void* base, size_t off) { return reinterpret_cast<const T*>(static_cast<char*>(base) + off); }

}  // namespace


// Lays out the derived device arrays (phase-major rows, see DevScene) for c->h_scene.n_orb phases in the grow-only
// derived arena and clears the running maxima.  Called by pb200_set_scene and by pb200_set_models (where the
// number of phases is n_models * n_orb).
static int plan_derived(pb200_ctx* c) {
    DevScene& h = c->h_scene;
    const int n_atoms = h.n_atoms;
    const bool fstar_per_chord = h.n_fstar > 0 && h.vsini != 0.0;
    struct Want { double** p; size_t n; };
    std::vector<Want> wants;
    double *nmax_d = nullptr, *smax_d = nullptr, *nact_d = nullptr, *ymax_d = nullptr;
    wants.push_back({&h.w_out, (size_t)h.Mp});
    wants.push_back({&h.star_shift, (size_t)h.Mp});
    wants.push_back({&h.w_in, (size_t)h.n_orb * h.Mp});
    wants.push_back({&h.ncol, (size_t)n_atoms * h.n_orb * h.Mp});
    wants.push_back({&h.rec, (size_t)pb_rec_stride(n_atoms) * h.n_orb * h.Mp});
    wants.push_back({&nmax_d, (size_t)(n_atoms ? n_atoms : 1) * h.n_orb});
    wants.push_back({&smax_d, (size_t)(n_atoms ? n_atoms : 1) * h.n_orb});
    wants.push_back({&nact_d, (size_t)(h.n_orb + 1) / 2});
    wants.push_back({&ymax_d, (size_t)kMaxConstituents});
    wants.push_back({&h.w0, (size_t)h.n_orb});
    wants.push_back({&h.thin_tot, (size_t)h.n_orb * 8});
    for (int i = 0; i < h.n_scen; ++i)
        if (h.scen[i].density_slot >= 0)
            wants.push_back({&h.density[h.scen[i].density_slot], (size_t)h.n_orb * h.M * h.n_x});
    wants.push_back({&h.fout, fstar_per_chord ? (size_t)h.n_wav : 1});
    double* bucket_d[kMaxConstituents] = {};
    for (int k = 0; k < h.n_cons; ++k)
        if (h.cons[k].kind == 0 && h.cons[k].n_bucket > 0) wants.push_back({&bucket_d[k], (size_t)(h.cons[k].n_bucket + 2) / 2 + 1});
    size_t need = 0;
    for (const auto& w : wants) need += ((w.n ? w.n : 1) * sizeof(double) + 255) & ~size_t(255);
    if (need > c->darena_bytes) {
        if (c->darena) PB_CUDA(c, cudaFree(c->darena));
        c->darena = nullptr; c->darena_bytes = 0;
        PB_CUDA(c, cudaMalloc(&c->darena, need));
        c->darena_bytes = need;
    }
    {
        size_t off = 0;
        for (const auto& w : wants) {
            *w.p = reinterpret_cast<double*>(static_cast<char*>(c->darena) + off);
            off += ((w.n ? w.n : 1) * sizeof(double) + 255) & ~size_t(255);
        }
    }
    for (int k = 0; k < h.n_cons; ++k)
        h.cons[k].bucket = (h.cons[k].kind == 0 && h.cons[k].n_bucket > 0) ? reinterpret_cast<const int*>(bucket_d[k]) : nullptr;
    c->buckets_ready = false;
    h.exp_tab = c->d_exp_tab;
    h.nmax = reinterpret_cast<unsigned long long*>(nmax_d);
    h.smax = reinterpret_cast<unsigned long long*>(smax_d);
    h.n_act = reinterpret_cast<int*>(nact_d);
    h.tab_ymax = reinterpret_cast<unsigned long long*>(ymax_d);
    PB_CUDA(c, cudaMemsetAsync(ymax_d, 0, kMaxConstituents * sizeof(double), c->stream));
    h.stat = reinterpret_cast<unsigned long long*>(c->d_oob);
    PB_CUDA(c, cudaMemsetAsync(nmax_d, 0, (size_t)(n_atoms ? n_atoms : 1) * h.n_orb * sizeof(double), c->stream));
    PB_CUDA(c, cudaMemsetAsync(smax_d, 0, (size_t)(n_atoms ? n_atoms : 1) * h.n_orb * sizeof(double), c->stream));
    return PB200_OK;
}

static int set_scene_impl(pb200_ctx* c, const pb200_scene* s, int n_ranks, int rank);

// Start the host-to-device copies of the large page-locked tables of the LAST scene again, at the places they had: a
// caller that is about to flatten the same objects (Transit.sumOverChords called again -- a retrieval loop, a light curve
// per night) lets the DMA run while the host prepares the small arrays.  The next pb200_set_scene skips every copy that
// is already on the stream (same source pointer, size and arena offset) and makes all the others as usual, so a
// different scene costs only the wasted transfer.  No counterpart in the reference (its tables never leave the host).
extern "C" int pb200_prefetch_tables(pb200_ctx* c) {
    if (!c) return PB200_ERR_INVALID;
    if (!c->arena || c->last_big.empty()) return PB200_OK;
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    c->prefetched.clear();
    c->prefetched_arena = c->arena;
    for (const auto& it : c->last_big) {
        if (!is_pinned(it.src, it.bytes) || it.offset + it.bytes > c->arena_bytes) continue;   // freed since
        PB_CUDA(c, cudaMemcpyAsync(static_cast<char*>(c->arena) + it.offset, it.src, it.bytes, cudaMemcpyHostToDevice, c->stream));
        c->prefetched.push_back(it);
    }
    return PB200_OK;
}

extern "C" int pb200_set_scene(pb200_ctx* c, const pb200_scene* s) { return set_scene_impl(c, s, 1, 0); }

// As pb200_set_scene, for rank `rank` of n_ranks processes that upload the SAME scene to their GPUs: of every large
// atomic sigma table only this rank's n_ranks-th goes over PCIe; the ranks then complete the tables with one
// all-gather per region over NVLink (pb200_table_regions; prometheus_b200/distributed.py) before the first run.
extern "C" int pb200_set_scene_sharded(pb200_ctx* c, const pb200_scene* s, int n_ranks, int rank) {
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(c, PB200_ERR_INVALID, "pb200_set_scene_sharded: rank %d of %d", rank, n_ranks);
    return set_scene_impl(c, s, n_ranks, rank);
}

extern "C" int pb200_table_regions(pb200_ctx* c, int max_regions, void** base, int64_t* chunk_doubles, int* n_regions) {
    if (!c || !n_regions) return PB200_ERR_INVALID;
    *n_regions = (int)c->regions.size();
    for (int i = 0; i < (int)c->regions.size() && i < max_regions; ++i) {
        if (base) base[i] = c->regions[i].first;
        if (chunk_doubles) chunk_doubles[i] = c->regions[i].second;
    }
    return PB200_OK;
}

static int set_scene_impl(pb200_ctx* c, const pb200_scene* s, int n_ranks, int rank) {
    if (!c || !s) return PB200_ERR_INVALID;
    if (s->abi_version != PB200_ABI_VERSION)
        return fail(c, PB200_ERR_INVALID, "scene abi_version %d != library %d", s->abi_version, PB200_ABI_VERSION);
    if (s->n_x <= 0 || s->n_rho <= 0 || s->n_phi <= 0 || s->n_orb <= 0 || s->n_wav <= 0)
        return fail(c, PB200_ERR_INVALID, "empty axis: n_x=%d n_rho=%d n_phi=%d n_orb=%d n_wav=%lld", s->n_x, s->n_rho,
                    s->n_phi, s->n_orb, (long long)s->n_wav);
    if (s->n_scenarios <= 0 || s->n_scenarios > kMaxScenarios)
        return fail(c, PB200_ERR_INVALID, "n_scenarios=%d outside [1, %d]", s->n_scenarios, kMaxScenarios);
    if (s->n_constituents < 0 || s->n_constituents > kMaxConstituents)
        return fail(c, PB200_ERR_INVALID, "n_constituents=%d outside [0, %d]", s->n_constituents, kMaxConstituents);
    if (s->n_moons < 0 || s->n_moons > kMaxMoons)
        return fail(c, PB200_ERR_INVALID, "n_moons=%d outside [0, %d]", s->n_moons, kMaxMoons);
    if (!s->x || !s->rho || !s->sin_phi || !s->cos_phi || !s->cos_phi_rot || !s->wavelength || !s->planet_y ||
        !s->scenarios || (s->n_constituents && !s->constituents) || (s->n_moons && (!s->moon_R || !s->moon_y)))
        return fail(c, PB200_ERR_INVALID, "scene has NULL array pointers");
    if ((int64_t)s->n_phi * s->n_rho > (int64_t)1 << 30) return fail(c, PB200_ERR_INVALID, "sky grid too large");

    PB_CUDA(c, cudaSetDevice(c->device));
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->has_scene = false;
    PB_CUDA(c, cudaEventRecord(c->ev[6], c->stream));

    DevScene& h = c->h_scene;
    memset(&h, 0, sizeof h);
    h.n_x = s->n_x; h.n_rho = s->n_rho; h.n_phi = s->n_phi; h.n_orb = s->n_orb; h.n_wav = s->n_wav;
    h.n_scen = s->n_scenarios; h.n_cons = s->n_constituents; h.n_moons = s->n_moons;
    h.M = s->n_phi * s->n_rho;
    h.Mp = ((h.M + kChordTile - 1) / kChordTile) * kChordTile;
    h.delta_x = s->delta_x; h.star_R = s->star_R; h.clv_u1 = s->clv_u1; h.clv_u2 = s->clv_u2; h.vsini = s->vsini;
    h.planet_R = s->planet_R;
    h.n_fstar = (s->fstar_x && s->fstar_y) ? s->n_fstar : 0;

    // ---- plan the arena --------------------------------------------------------------------------
    ArenaPlan plan;
    const size_t D = sizeof(double);
    const size_t o_x = plan.add(s->x, s->n_x * D), o_rho = plan.add(s->rho, s->n_rho * D);
    const size_t o_sin = plan.add(s->sin_phi, s->n_phi * D), o_cos = plan.add(s->cos_phi, s->n_phi * D);
    const size_t o_rot = plan.add(s->cos_phi_rot, s->n_phi * D), o_wav = plan.add(s->wavelength, s->n_wav * D);
    const size_t o_py = plan.add(s->planet_y, s->n_orb * D);
    size_t o_fx = 0, o_fy = 0, o_mR = 0, o_my = 0;
    if (h.n_fstar) { o_fx = plan.add(s->fstar_x, h.n_fstar * D); o_fy = plan.add(s->fstar_y, h.n_fstar * D); }
    if (s->n_moons) { o_mR = plan.add(s->moon_R, s->n_moons * D); o_my = plan.add(s->moon_y, (size_t)s->n_moons * s->n_orb * D); }

    struct ScnOff { size_t cx, cy, sh, amp, gx, gy, gz, val; };
    struct ConOff { size_t tx, ty, gP, gT, gW, val; };
    std::vector<ScnOff> so(s->n_scenarios);
    std::vector<ConOff> co(s->n_constituents);
    int n_atoms = 0, n_mols = 0, n_density = 0, covered = 0;
    for (int i = 0; i < s->n_scenarios; ++i) {
        const pb200_scenario& q = s->scenarios[i];
        if (q.type < PB200_SCN_BAROMETRIC || q.type > PB200_SCN_SERPENS)
            return fail(c, PB200_ERR_UNSUPPORTED, "scenario %d: unknown type %d", i, q.type);
        if (!q.shift) return fail(c, PB200_ERR_INVALID, "scenario %d: shift array is NULL", i);
        if (q.type != PB200_SCN_SERPENS && (!q.centre_x || !q.centre_y))
            return fail(c, PB200_ERR_INVALID, "scenario %d: centre arrays are NULL", i);
        if (q.first_constituent != covered || q.n_constituents < 0 || covered + q.n_constituents > s->n_constituents)
            return fail(c, PB200_ERR_INVALID, "scenario %d: constituent range [%d,+%d) is not contiguous", i,
                        q.first_constituent, q.n_constituents);
        covered += q.n_constituents;
        so[i].cx = q.centre_x ? plan.add(q.centre_x, s->n_orb * D) : 0;
        so[i].cy = q.centre_y ? plan.add(q.centre_y, s->n_orb * D) : 0;
        so[i].sh = plan.add(q.shift, s->n_orb * D);
        so[i].amp = q.amplitude ? plan.add(q.amplitude, s->n_orb * D) : 0;
        if (q.type == PB200_SCN_SERPENS) {
            if (!q.gx || !q.gy || !q.gz || !q.values || q.dims[0] < 2 || q.dims[1] < 2 || q.dims[2] < 2)
                return fail(c, PB200_ERR_INVALID, "scenario %d: serpens table missing", i);
            so[i].gx = plan.add(q.gx, q.dims[0] * D); so[i].gy = plan.add(q.gy, q.dims[1] * D);
            so[i].gz = plan.add(q.gz, q.dims[2] * D);
            so[i].val = plan.add(q.values, (size_t)q.dims[0] * q.dims[1] * q.dims[2] * D);
        }
    }
    if (covered != s->n_constituents) return fail(c, PB200_ERR_INVALID, "constituents not covered by the scenarios");
    for (int k = 0; k < s->n_constituents; ++k) {
        const pb200_constituent& q = s->constituents[k];
        if (q.kind == 0) {
            if (!q.tab_x || !q.tab_y || q.n_tab < 1) return fail(c, PB200_ERR_INVALID, "constituent %d: sigma table missing", k);
            if (n_ranks > 1 && (size_t)q.n_tab * D >= (size_t(1) << 20)) {
                co[k].tx = plan.add_sharded(q.tab_x, q.n_tab * D, n_ranks); co[k].ty = plan.add_sharded(q.tab_y, q.n_tab * D, n_ranks);
            } else {
                co[k].tx = plan.add(q.tab_x, q.n_tab * D); co[k].ty = plan.add(q.tab_y, q.n_tab * D);
            }
        } else if (q.kind == 1) {
            if (!q.grid_P || !q.grid_T || !q.grid_w || !q.values || q.dims[0] < 2 || q.dims[1] < 2 || q.dims[2] < 2)
                return fail(c, PB200_ERR_INVALID, "constituent %d: molecular table missing", k);
            co[k].gP = plan.add(q.grid_P, q.dims[0] * D); co[k].gT = plan.add(q.grid_T, q.dims[1] * D);
            co[k].gW = plan.add(q.grid_w, q.dims[2] * D);
            co[k].val = plan.add(q.values, (size_t)q.dims[0] * q.dims[1] * q.dims[2] * D);
        } else {
            return fail(c, PB200_ERR_UNSUPPORTED, "constituent %d: unknown kind %d", k, q.kind);
        }
    }

    // ---- stage + one H2D copy --------------------------------------------------------------------
    if (plan.total > c->stage_bytes) {
        if (c->stage) cudaFreeHost(c->stage);
        c->stage = nullptr; c->stage_bytes = 0;
        PB_CUDA(c, cudaMallocHost(&c->stage, plan.total));
        c->stage_bytes = plan.total;
    }
    if (plan.total > c->arena_bytes) {
        if (c->arena) PB_CUDA(c, cudaFree(c->arena));
        c->arena = nullptr; c->arena_bytes = 0;
        PB_CUDA(c, cudaMalloc(&c->arena, plan.total));
        c->arena_bytes = plan.total;
    }
    {   // large arrays that already lie in page-locked memory of pb200_host_alloc (the sigma tables built by
        // pb200_voigt_sigma through the Python mirror) go to the device from where they are; everything
        // else is gathered in the staging buffer and sent as runs of consecutive items
        size_t run_lo = 0, run_hi = 0;   // staged byte range not yet sent
        auto flush_run = [&]() -> cudaError_t {
            if (run_hi == run_lo) return cudaSuccess;
            cudaError_t e = cudaMemcpyAsync(static_cast<char*>(c->arena) + run_lo, static_cast<char*>(c->stage) + run_lo,
                                            run_hi - run_lo, cudaMemcpyHostToDevice, c->stream);
            run_lo = run_hi;
            return e;
        };
        c->regions.clear();
        std::vector<pb200_ctx::UpItem> big_now;
        for (const auto& it : plan.items) {
            if (it.chunk) {   // this rank's slice only
                PB_CUDA(c, flush_run());
                const size_t lo = it.chunk * rank;
                const size_t n = lo < it.bytes ? (it.bytes - lo < it.chunk ? it.bytes - lo : it.chunk) : 0;
                char* dst = static_cast<char*>(c->arena) + it.offset + lo;
                const char* src = static_cast<const char*>(it.src) + lo;
                if (n && is_pinned(src, n)) {
                    PB_CUDA(c, cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, c->stream));
                } else if (n) {
                    memcpy(static_cast<char*>(c->stage) + it.offset + lo, src, n);
                    PB_CUDA(c, cudaMemcpyAsync(dst, static_cast<char*>(c->stage) + it.offset + lo, n, cudaMemcpyHostToDevice, c->stream));
                }
                run_lo = run_hi = it.offset + it.chunk * n_ranks;
                c->regions.push_back({static_cast<char*>(c->arena) + it.offset, (int64_t)(it.chunk / sizeof(double))});
            } else if (it.bytes >= (64u << 10) && is_pinned(it.src, it.bytes)) {
                PB_CUDA(c, flush_run());
                bool done = false;   // already on the stream (pb200_prefetch_tables): same source, same size, same place
                if (c->prefetched_arena == c->arena)
                    for (const auto& q : c->prefetched) done = done || (q.src == it.src && q.bytes == it.bytes && q.offset == it.offset);
                if (!done)
                    PB_CUDA(c, cudaMemcpyAsync(static_cast<char*>(c->arena) + it.offset, it.src, it.bytes,
                                               cudaMemcpyHostToDevice, c->stream));
                big_now.push_back({it.src, it.bytes, it.offset});
                run_lo = run_hi = it.offset + ((it.bytes + 255) & ~size_t(255));
            } else {
                memcpy(static_cast<char*>(c->stage) + it.offset, it.src, it.bytes);
                run_hi = it.offset + ((it.bytes + 255) & ~size_t(255));
            }
        }
        if (run_hi > plan.total) run_hi = plan.total;
        PB_CUDA(c, flush_run());
        c->last_big.swap(big_now);
        c->prefetched.clear();
    }

    void* A = c->arena;
    h.x = at<double>(A, o_x); h.rho = at<double>(A, o_rho); h.sin_phi = at<double>(A, o_sin);
    h.cos_phi = at<double>(A, o_cos); h.cos_phi_rot = at<double>(A, o_rot); h.wavelength = at<double>(A, o_wav);
    h.planet_y = at<double>(A, o_py);
    if (h.n_fstar) { h.fstar_x = at<double>(A, o_fx); h.fstar_y = at<double>(A, o_fy); }
    if (s->n_moons) { h.moon_R = at<double>(A, o_mR); h.moon_y = at<double>(A, o_my); }
    for (int i = 0; i < s->n_scenarios; ++i) {
        const pb200_scenario& q = s->scenarios[i];
        DevScenario& d = h.scen[i];
        d.type = q.type; d.n_cons = q.n_constituents; d.first_cons = q.first_constituent; d.T = q.T;
        memcpy(d.p, q.p, sizeof d.p);
        d.centre_x = q.centre_x ? at<double>(A, so[i].cx) : nullptr;
        d.centre_y = q.centre_y ? at<double>(A, so[i].cy) : nullptr;
        d.shift = at<double>(A, so[i].sh);
        d.amplitude = q.amplitude ? at<double>(A, so[i].amp) : nullptr;
        d.density_slot = -1;
        if (q.type == PB200_SCN_SERPENS) {
            d.gx = at<double>(A, so[i].gx); d.gy = at<double>(A, so[i].gy); d.gz = at<double>(A, so[i].gz);
            d.values = at<double>(A, so[i].val);
            memcpy(d.dims, q.dims, sizeof d.dims);
        }
        for (int k = q.first_constituent; k < q.first_constituent + q.n_constituents; ++k) {
            const pb200_constituent& qc = s->constituents[k];
            DevConstituent& dc = h.cons[k];
            dc.kind = qc.kind; dc.scenario = i; dc.chi = qc.chi; dc.offset = qc.offset; dc.fill = qc.fill;
            dc.atom_index = dc.mol_index = -1;
            if (qc.kind == 0) {
                dc.atom_index = n_atoms++;
                dc.tab_x = at<double>(A, co[k].tx); dc.tab_y = at<double>(A, co[k].ty); dc.n_tab = qc.n_tab;
                dc.bucket = nullptr; dc.n_bucket = 0;
                const double span = qc.tab_x[qc.n_tab - 1] - qc.tab_x[0];
                if (qc.n_tab >= 1024 && span > 0.0 && span < 1e300) {   // worth an index (and a finite, ascending table)
                    dc.n_bucket = (int)(qc.n_tab < (1 << 22) ? qc.n_tab : (1 << 22));
                    dc.bucket_x0 = qc.tab_x[0];
                    dc.bucket_inv = (double)dc.n_bucket / span;
                }
            } else {
                dc.mol_index = n_mols++;
                d.has_molecule = 1;
                dc.grid_P = at<double>(A, co[k].gP); dc.grid_T = at<double>(A, co[k].gT);
                dc.grid_w = at<double>(A, co[k].gW); dc.values = at<double>(A, co[k].val);
                memcpy(dc.dims, qc.dims, sizeof dc.dims);
            }
        }
        if (d.has_molecule) d.density_slot = n_density++;
    }
    h.n_atoms = n_atoms; h.n_mols = n_mols;
    {   // absorbers of the run path (see DevScene::n_abs)
        const bool group = c->opt[PB200_OPT_GROUP] != 0;
        h.n_abs = 0;
        for (int i = 0; i < s->n_scenarios; ++i) {
            const pb200_scenario& q = s->scenarios[i];
            bool open = false;
            for (int k = q.first_constituent; k < q.first_constituent + q.n_constituents; ++k) {
                if (s->constituents[k].kind != 0) { open = false; continue; }
                if (group && open) { h.abs_count[h.n_abs - 1] = k + 1 - h.abs_first[h.n_abs - 1]; continue; }
                h.abs_first[h.n_abs] = k; h.abs_count[h.n_abs] = 1; ++h.n_abs;
                open = true;
            }
        }
    }

    c->general_path = (n_mols > 0) || (h.n_fstar > 0 && h.vsini != 0.0);
    {
        int rc = plan_derived(c);
        if (rc) return rc;
    }
    {   // mirror pairing of chords (compact_kernel): all bodies lie in z = 0 and every analytic density depends
        // on z^2 only; the tabulated (SERPENS) density does not.  The phi axis must be mirror symmetric about
        // pi/2 and 3 pi/2 on the grid itself: sin(phi_i') = sin(phi_i), cos(phi_i') = -cos(phi_i), i' = n/2-1-i.
        // kernel (m) pairs too, unless F* is seen through stellar rotation (star_shift differs within a pair)
        bool ok = !(h.n_fstar > 0 && h.vsini != 0.0) && (s->n_phi % 2 == 0);
        for (int i = 0; ok && i < s->n_scenarios; ++i) ok = s->scenarios[i].type != PB200_SCN_SERPENS;
        const int n2 = s->n_phi / 2;
        for (int hh = 0; ok && hh < 2; ++hh)
            for (int i = 0; ok && i < n2; ++i) {
                const int a = hh * n2 + i, b = hh * n2 + n2 - 1 - i;
                ok = fabs(s->sin_phi[a] - s->sin_phi[b]) <= 1e-14 && fabs(s->cos_phi[a] + s->cos_phi[b]) <= 1e-14;
            }
        h.mirror = (ok && c->opt[PB200_OPT_MIRROR] != 0) ? 1 : 0;
    }
    h.fout_per_wav = (h.n_fstar > 0 && h.vsini != 0.0) ? 1 : 0;
    h.n_orb_model = h.n_orb;
    h.moon_R_per_phase = 0;

    PB_CUDA(c, cudaMemcpyAsync(c->d_scene, &h, sizeof(DevScene), cudaMemcpyHostToDevice, c->stream));
    // phase-independent prologue: flux weights and stellar surface shifts
    launch_sky_cells(c->d_scene, h.Mp, c->stream);
    launch_fout_scalar(c->d_scene, c->stream);
    c->launches += 2;
    PB_CUDA(c, cudaGetLastError());
    PB_CUDA(c, cudaEventRecord(c->ev[7], c->stream));
    // no synchronisation here: the staging buffer is only reused after the synchronize at the top of
    // the next pb200_set_scene, and the upload time is read back lazily (collect_timing)
    c->upload_ms = -1.0;
    c->has_scene = true;
    c->base = h;
    c->lastR_full = false;
    c->mol_bins = -1;
    c->mol_max_items = -1;
    return PB200_OK;
}

// =================================================================================================
// Transit.sumOverChords
// =================================================================================================
namespace {

// Sigma-sorted windows (pb200_sort.cu): one absorber, a wavelength axis and a sky grid large enough that the sort (a
// device-wide radix sort of n_phase * n_wav keys) is small next to the chord loops it shortens: C5 (36 000 chords per
// phase x 300 306 wavelengths) 72 -> 24 ms; C2 without its molecules (3 000 x 95 897) would lose 0.2 ms to the sort.
// PB200_OPT_SORT = 0 / 1 forces it.
bool want_sorted(const pb200_ctx* c, int n_phase, int64_t n_wav_sel) {
    const DevScene& h = c->h_scene;
    const bool possible = !c->general_path && h.n_abs == 1 && n_phase < (1 << 20) && (int64_t)n_phase * n_wav_sel < (1ll << 31);
    if (c->opt[PB200_OPT_SORT] >= 0) return possible && c->opt[PB200_OPT_SORT] == 1;
    // (the wavelength count is this RUN's: a rank of an 8-GPU job holds an eighth of the grid and still gains -- C5 / 8:
    // 37 538 wavelengths, 9.1 ms unsorted against 2.9 sorted)
    return possible && n_wav_sel >= (1 << 13) && h.M >= (1 << 14);
}

// Sigma cells (pb200_transit.cu, cell_*_kernel): where the sorted windows would run and the expansions are allowed, R is
// tabulated per phase on a grid of sigma cells instead -- no sort, work independent of the wavelength count.
// There is no sort to pay for, so the sky grid may be small: the cells cost ~650 x 25 FP64 instructions per active chord
// record whatever the wavelength count (C4, 9 000 chords per phase x 31 969 wavelengths: transit stage 0.100 -> 0.070 ms).
bool want_cells(const pb200_ctx* c, int n_phase, int64_t n_wav_sel) {
    const DevScene& h = c->h_scene;
    if (c->opt[PB200_OPT_CENTRE] < 3 || c->opt[PB200_OPT_THIN] == 0 || c->general_path || h.n_abs != 1) return false;
    if (c->opt[PB200_OPT_SORT] >= 0) return c->opt[PB200_OPT_SORT] == 1;
    return n_wav_sel >= (1 << 13);
}

int check_scene(pb200_ctx* c) {
    if (!c) return PB200_ERR_INVALID;
    if (!c->has_scene) return fail(c, PB200_ERR_NO_SCENE, "no scene: call pb200_set_scene first");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return PB200_OK;
}

// enqueue kernels (a') and (c') for the tile; R_dev receives [n_phase][n_wav_sel]
int enqueue_run(pb200_ctx* c, int phase_lo, int phase_hi, int64_t wav_lo, int64_t wav_hi, double* R_dev,
                bool R_in_scratch, int64_t cyc_block = 0, int cyc_n = 1, int cyc_rank = 0, int64_t R_row_stride = 0) {
    const DevScene& h = c->h_scene;
    if (phase_lo < 0 || phase_hi > h.n_orb || phase_lo >= phase_hi || wav_lo < 0 || wav_hi > h.n_wav || wav_lo >= wav_hi)
        return fail(c, PB200_ERR_INVALID, "tile [%d,%d) x [%lld,%lld) outside the (%d, %lld) grid", phase_lo, phase_hi,
                    (long long)wav_lo, (long long)wav_hi, h.n_orb, (long long)h.n_wav);
    const int n_phase = phase_hi - phase_lo;
    const int64_t n_wav_sel = wav_hi - wav_lo;
    const int64_t n_out = (int64_t)n_phase * n_wav_sel;
    const int64_t launches0 = c->launches;
    c->lastR_full = false;
    c->last_perm = nullptr; c->last_nsig_sorted = nullptr; c->last_sort_n = 0;
    c->mol_stats_pending = false;

    TransitParams prm;
    prm.phase_lo = phase_lo; prm.n_phase = n_phase; prm.wav_lo = wav_lo; prm.n_wav_sel = n_wav_sel;
    prm.cyc_block = cyc_block; prm.cyc_n = cyc_n; prm.cyc_rank = cyc_rank;
    prm.R_row_stride = R_in_scratch ? 0 : R_row_stride;
    if (cyc_block && c->general_path)
        return fail(c, PB200_ERR_UNSUPPORTED, "block-cyclic wavelength tiles are implemented for atomic scenes (kernel (c')) only");
    // kernel (m) walks mirror pairs when the scene allows it: about half the chords
    const int n_pairs = 2 * ((h.n_phi / 2 + 1) / 2) * h.n_rho;
    const int n_tiles = (c->general_path && h.mirror) ? (n_pairs + kChordTile - 1) / kChordTile : h.Mp / kChordTile;
    // wavelengths per CTA: kernel (m) 128; kernel (c') one window of 32 * WPT (WPT = 8, 4, 2 for <= 2, <= 4, more absorbers)
    const int cta_wav = c->general_path ? 256 : 32 * (h.n_abs <= 2 ? 8 : h.n_abs <= 4 ? 4 : 2);
    const int64_t ctas = ((n_wav_sel + cta_wav - 1) / cta_wav) * n_phase;
    // Chord-axis splits: enough CTAs that the last wave is a small fraction of the run (~16 per SM).  Kernel (m) does
    // the same work in every CTA; kernel (c') has one CTA per wavelength window, most of which
    // finish early (thin or culled), so ~16 per SM is plenty -- and with a single split it writes R itself
    // and finalize_kernel (and the partial sums' traffic) is not needed at all.
    // (kernel (m) pays its prologue -- table searches and the V columns of every thread -- once per CTA: few, long CTAs)
    // (the sigma-sorted windows differ in cost by orders of magnitude -- a window in the tau ~ 1 band walks every chord
    // through the exp, most others are decided per chord without one: more, shorter CTAs keep the tail short)
    const bool cells = want_cells(c, n_phase, n_wav_sel);
    const int per_sm = (!c->general_path && want_sorted(c, n_phase, n_wav_sel)) ? 64 : 16;
    int64_t want = ((int64_t)c->sm_count * per_sm + ctas - 1) / ctas;
    // at least 4 tiles per CTA of kernel (c'), 2 of kernel (m) (whose tiles are ~30 us of work each: with mirror pairing
    // C2 has only 6 tiles per phase, and an eighth of its wavelengths is 517 CTAs -- 1.16 waves -- unless the chords split)
    const int min_tiles_per_cta = c->general_path ? 2 : 4;
    if (want > n_tiles / min_tiles_per_cta) want = n_tiles / min_tiles_per_cta;
    if (want < 1) want = 1;
    if (cells) want = 1;    // cell_eval_kernel writes R itself: no partial sums, no finalize_kernel
    prm.tiles_per_split = (int)((n_tiles + want - 1) / want);
    prm.n_split = (n_tiles + prm.tiles_per_split - 1) / prm.tiles_per_split;
    const int n_cs = cells ? cell_chord_splits(c->sm_count, n_phase, n_tiles) : 0;
    {
        prm.min_tiles = (int)c->opt[PB200_OPT_MIN_TILES];   // tuning experiments only
        prm.culled = c->general_path ? 0 : 1;
        prm.thin_v = c->opt[PB200_OPT_THIN] ? 0.125 : 0.0;   // 0 (measurement only): every active sample through the table exp
        prm.centre = c->opt[PB200_OPT_THIN] ? (int)c->opt[PB200_OPT_CENTRE] : 0;
        if (prm.min_tiles < 1) prm.min_tiles = 1;
    }

    const size_t r_bytes = R_in_scratch ? (size_t)((n_out * sizeof(double) + 255) & ~255ll) : 0;
    const size_t part_bytes = prm.n_split > 1 ? (size_t)prm.n_split * n_out * sizeof(double) : 0;
    const size_t nsig_bytes = c->general_path ? 0 : (size_t)h.n_abs * n_out * sizeof(double);
    prm.win = 32 * (h.n_abs <= 2 ? 8 : h.n_abs <= 4 ? 4 : 2);
    prm.n_win = (int)((n_wav_sel + prm.win - 1) / prm.win);
    const size_t order_bytes = c->general_path ? 0 : (((size_t)n_phase * prm.n_win * sizeof(int) + 255) & ~size_t(255));
    const size_t winmax_bytes = c->general_path ? 0 : (size_t)h.n_abs * n_phase * prm.n_win * sizeof(double);
    const bool sorted = !cells && want_sorted(c, n_phase, n_wav_sel);
    const size_t cell_bytes = cells ? ((cell_scratch_bytes(n_phase, n_cs) + 255) & ~size_t(255)) : 0;
    const size_t al = 255;
    const size_t sort_temp = sorted ? ((sigma_sort_temp_bytes(n_out) + al) & ~al) : 0;
    const size_t sort_bytes = sorted ? (size_t)n_out * (8 + 8 + 8 + 4 + 4) + 5 * 256 + sort_temp : 0;
    int rc = ensure_scratch(c, r_bytes + part_bytes + nsig_bytes + order_bytes + winmax_bytes + sort_bytes + cell_bytes + 1024);
    if (rc) return rc;
    prm.R = R_in_scratch ? c->scratch : R_dev;
    prm.partial = reinterpret_cast<double*>(reinterpret_cast<char*>(c->scratch) + r_bytes);
    prm.nsig = reinterpret_cast<double*>(reinterpret_cast<char*>(c->scratch) + r_bytes + part_bytes);
    int* order = reinterpret_cast<int*>(reinterpret_cast<char*>(c->scratch) + r_bytes + part_bytes + nsig_bytes);
    prm.order = (cells || sorted) ? nullptr : order;   // compact_kernel sorts the window keys only for the kernel that reads them
    prm.winmax = reinterpret_cast<double*>(reinterpret_cast<char*>(c->scratch) + r_bytes + part_bytes + nsig_bytes + order_bytes);
    prm.perm = nullptr;
    prm.cell_range = nullptr; prm.cell_count = nullptr;
    char* sort_base = reinterpret_cast<char*>(c->scratch) + ((r_bytes + part_bytes + nsig_bytes + order_bytes + winmax_bytes + al) & ~al);

    if (cells) cell_bind(prm, sort_base);
    PB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
    cudaStream_t sig_stream = c->stream;
    if (!c->buckets_ready && h.n_atoms > 0) {   // after the (possibly sharded + all-gathered) tables are complete on the device
        launch_bucket_build(c->d_scene, c->sm_count, c->stream);
        c->launches += 1;
        c->buckets_ready = true;
    }
    if (!c->general_path) {   // sigma(p, w) of the tile and its per-phase maxima, needed by the culling below
        if (h.n_atoms < 1) return fail(c, PB200_ERR_INVALID, "scene has no absorbing constituent");
        if (h.n_abs > PB200_MAX_ABSORBERS)   // before any kernel that sizes shared memory by the absorber count
            return fail(c, PB200_ERR_UNSUPPORTED, "%d absorbers (scenarios with atomic constituents); kernel (c') is compiled for <= %d", h.n_abs, PB200_MAX_ABSORBERS);
        // smax is left clean by compact_kernel of the previous run (and by pb200_set_scene), the run
        // statistics are reset by sigma_phase_kernel: no memset nodes in the run
        // sigma_phase (+ the sort) and the column kernel do not depend on each other (the column kernel bounds sigma by
        // the tables' maxima): fork -- sigma on the auxiliary stream, column on the main one -- and join before compaction
        // (sigma cells: no sigma_phase_kernel at all -- the compaction and the cell sums bound sigma by the tables' maxima,
        //  cell_eval_kernel looks sigma(p, lambda) up where it uses it)
        if (!cells) {
            PB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
            PB_CUDA(c, cudaStreamWaitEvent(c->aux, c->ev_fork, 0));
            sig_stream = c->aux;
            launch_sigma_phase(c->d_scene, prm, sig_stream);
            c->launches += 1;
        }
        if (sorted) {
            auto take = [&](size_t bytes) { char* p0 = sort_base; sort_base += (bytes + al) & ~al; return p0; };
            double* nsig_sorted = reinterpret_cast<double*>(take((size_t)n_out * 8));
            unsigned long long* k_in = reinterpret_cast<unsigned long long*>(take((size_t)n_out * 8));
            unsigned long long* k_out = reinterpret_cast<unsigned long long*>(take((size_t)n_out * 8));
            int* v_in = reinterpret_cast<int*>(take((size_t)n_out * 4));
            int* perm = reinterpret_cast<int*>(take((size_t)n_out * 4));
            void* temp = take(sort_temp);
            cudaError_t e = launch_sigma_sort(prm.nsig, n_phase, n_wav_sel, k_in, k_out, v_in, perm, temp, sort_temp, nsig_sorted, sig_stream);
            if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "sigma sort: %s", cudaGetErrorString(e));
            c->launches += 3;   // keys, gather + the radix sort's own passes
            prm.nsig = nsig_sorted;
            prm.perm = perm;
            c->last_perm = perm; c->last_nsig_sorted = nsig_sorted; c->last_sort_n = n_out;
        }
    }
    // PB200_OPT_CULL = 0 (measurement only): every unblocked chord gets its exact sums and a record, as in the
    // reference's loop
    const bool cull = c->opt[PB200_OPT_CULL] != 0;
    launch_column(c->d_scene, h.Mp, phase_lo, n_phase, c->d_oob, c->sm_count, c->stream, h.n_mols > 0,
                  c->general_path ? 0 : (cull ? 1 : 2));
    c->launches += 1;
    if (!c->general_path) {   // chord records of kernel (c'), transparent chords culled
        if (!cells) {
            PB_CUDA(c, cudaEventRecord(c->ev_join, c->aux));
            PB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
        }
        PB_CUDA(c, launch_compact(c->d_scene, prm, cull, h.M, c->stream));
        c->launches += 1;
    }
    PB_CUDA(c, cudaGetLastError());
    PB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
    // Kernel (m) runs in moment form unless the tables are too steep for it (more than 64 sub-intervals per pressure
    // interval would be needed), its blocks do not fit shared memory, or PB200_OPT_MOL_RECORDS asks for the per-record
    // form (measurement / regression).
    bool moments_done = false;
    if (c->general_path && h.n_mols > 0 && !c->opt[PB200_OPT_MOL_RECORDS]) {
        if (!c->d_mol_stat) PB_CUDA(c, cudaMalloc((void**)&c->d_mol_stat, 256));   // [0..3] see above, [4 + 2 i], [5 + 2 i] node range of molecule i
        unsigned long long st[2] = {0ull, 0ull};
        if (c->mol_bins < 0) {   // d_max of the tables, once per scene
            PB_CUDA(c, cudaMemsetAsync(c->d_mol_stat, 0, 32, c->stream));
            launch_mol_dmax(c->d_scene, c->d_mol_stat, c->sm_count, c->stream);
            c->launches += 1;
            PB_CUDA(c, cudaMemcpyAsync(st, c->d_mol_stat, 8, cudaMemcpyDeviceToHost, c->stream));
            PB_CUDA(c, cudaStreamSynchronize(c->stream));
            double dmax;
            memcpy(&dmax, &st[0], sizeof dmax);
            c->mol_bins = (dmax < 64.0 * 0.38) ? (dmax > 0.38 ? (int)ceil(dmax / 0.38) : 1) : 0;   // 0: too steep (or NaN)
        }
        if (c->mol_bins > 0) {
            // pair slots per phase, rounded up to whole chord tiles (the splits of the chord axis are whole tiles too)
            const int n_chord_items = h.mirror ? n_pairs : h.M;
            const int n_slots = ((n_chord_items + kChordTile - 1) / kChordTile) * kChordTile;
            // The item count of a chord depends on the scene (density rows, pressure grids, bins), not on the wavelength
            // tile: counted once per scene over ALL its phases, so later runs of the scene have no host round trip.
            if (c->mol_max_items < 0) {
                unsigned long long init[2 + 2 * kMaxMols];
                for (int i = 0; i < kMaxMols; ++i) { init[2 + 2 * i] = ~0ull; init[3 + 2 * i] = 0ull; }   // atomicMin / atomicMax targets
                PB_CUDA(c, cudaMemcpyAsync(c->d_mol_stat + 4, init + 2, 2 * kMaxMols * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
                PB_CUDA(c, cudaMemsetAsync(c->d_mol_stat + 1, 0, 8, c->stream));
                if (!(phase_lo == 0 && n_phase == h.n_orb)) {   // the count needs every phase's density rows
                    launch_column(c->d_scene, h.Mp, 0, h.n_orb, c->d_oob, c->sm_count, c->stream, true, 0);
                    c->launches += 1;
                }
                launch_mol_moments(c->d_scene, n_slots, 0, h.n_orb, c->mol_bins, nullptr, 0, c->d_mol_stat, true, MolLayout{}, c->stream);
                unsigned long long back[4 + 2 * kMaxMols];
                PB_CUDA(c, cudaMemcpyAsync(back, c->d_mol_stat, sizeof back, cudaMemcpyDeviceToHost, c->stream));
                PB_CUDA(c, cudaStreamSynchronize(c->stream));
                c->launches += 1;
                c->mol_max_items = back[1] > 0 ? (int)back[1] : 1;
                for (int i = 0; i < kMaxMols; ++i)
                    if (back[5 + 2 * i] == 0ull) back[4 + 2 * i] = 0ull;   // no in-grid sample of this molecule anywhere
                c->mol_lay = mol_layout(h, back + 4, 1);
            }
            const int max_items = c->mol_max_items;
            const int stride_d = mol_block_stride(h.n_abs, max_items);
            const size_t need = (size_t)n_phase * n_slots * stride_d * sizeof(double);
            if (need > c->mol_items_bytes) {
                if (c->mol_items) PB_CUDA(c, cudaFree(c->mol_items));
                c->mol_items = nullptr; c->mol_items_bytes = 0;
                PB_CUDA(c, cudaMalloc((void**)&c->mol_items, need));
                c->mol_items_bytes = need;
            }
            PB_CUDA(c, cudaMemsetAsync(c->d_mol_stat + 2, 0, 16, c->stream));
            launch_mol_moments(c->d_scene, n_slots, phase_lo, n_phase, c->mol_bins, c->mol_items, stride_d, c->d_mol_stat, false, c->mol_lay, c->stream);
            c->launches += 1;
            int n_launch = 0;
            cudaError_t e = launch_transit_moments(c->d_scene, h, prm, c->mol_items, stride_d, n_slots, c->mol_lay, c->stream, &n_launch);
            if (e == cudaSuccess) { c->launches += n_launch; moments_done = true; c->mol_stats_pending = true; }
            else if (e != cudaErrorInvalidValue) return fail(c, PB200_ERR_CUDA, "transit (moments) launch: %s", cudaGetErrorString(e));
            else cudaGetLastError();   // does not fit shared memory: the per-record form below
        }
    }
    if (moments_done) {
        // kernel (m) has run in moment form
    } else if (c->general_path) {
        int n_launch = 0;
        cudaError_t e = launch_transit_general(c->d_scene, h, prm, c->sm_count, c->stream, &n_launch);
        if (e == cudaErrorInvalidValue)
            return fail(c, PB200_ERR_UNSUPPORTED, "molecular tables need more shared memory than one SM has (pressure nodes x molecules > 64)");
        if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "transit (general) launch: %s", cudaGetErrorString(e));
        c->launches += n_launch;
    } else if (cells) {
        cudaError_t e = launch_cell_moments(c->d_scene, prm, n_cs, sort_base, c->stream);
        if (e == cudaSuccess) e = launch_cell_eval(c->d_scene, prm, n_cs, sort_base, c->stream);
        if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "sigma cells launch: %s", cudaGetErrorString(e));
        c->launches += 2;
    } else {
        cudaError_t e = launch_transit(c->d_scene, h.n_abs, prm, c->stream);
        if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "transit launch: %s", cudaGetErrorString(e));
        c->launches += 1;
    }
    PB_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
    if (prm.n_split > 1) {
        launch_finalize(c->d_scene, prm, c->sm_count, c->stream);
        c->launches += 1;
        PB_CUDA(c, cudaGetLastError());
    }
    PB_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
    c->last = pb200_timing{};
    c->last.upload_ms = c->upload_ms;
    c->last.launches = c->launches - launches0;
    c->last.samples = (int64_t)n_phase * h.M * n_wav_sel;
    c->last.executed_samples = c->general_path ? c->last.samples : -1;   // known after the next synchronisation
    c->last.exp_samples = c->last.executed_samples;
    c->last_wpt = prm.win / 32;   // wavelengths per thread of kernel (c')
    c->last_n_wav_sel = n_wav_sel;
    // where the run leaves R (pb200_lightcurve with R_dev == NULL reads it there; the caller keeps it alive)
    c->lastR_full = !cyc_block && !prm.R_row_stride && wav_lo == 0 && n_wav_sel == h.n_wav;
    c->lastR_phase_lo = phase_lo; c->lastR_n_phase = n_phase; c->lastR_ptr = prm.R;
    return PB200_OK;
}

int collect_timing(pb200_ctx* c) {
    float a = 0, b = 0, d = 0;
    PB_CUDA(c, cudaEventElapsedTime(&a, c->ev[2], c->ev[3]));
    PB_CUDA(c, cudaEventElapsedTime(&b, c->ev[3], c->ev[4]));
    PB_CUDA(c, cudaEventElapsedTime(&d, c->ev[4], c->ev[5]));
    c->last.column_ms = a; c->last.transit_ms = b; c->last.finalize_ms = d;
    if (c->upload_ms < 0.0) {
        float u = 0;
        PB_CUDA(c, cudaEventElapsedTime(&u, c->ev[6], c->ev[7]));
        c->upload_ms = u;
    }
    c->last.upload_ms = c->upload_ms;
    return PB200_OK;
}

int check_oob(pb200_ctx* c) {
    unsigned long long st[5] = {0, 0, 0, 0, 0};
    PB_CUDA(c, cudaMemcpyAsync(st, c->d_oob, sizeof st, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    const int flag = (int)(st[0] & 0xffffffffull);
    if (c->last.executed_samples < 0) {
        c->last.executed_samples = (int64_t)st[1] * c->last_n_wav_sel;
        c->last.exp_samples = (int64_t)st[2] * 32 * c->last_wpt   // a warp owns 32 * WPT wavelengths
                              + (int64_t)st[4];                    // sigma cells / narrow sorted windows: one exp per (chord, cell)
        c->last.centre_pairs = (int64_t)st[4];
    }
    if (c->mol_stats_pending) {   // kernel (m), moment form: (live chord-pair blocks, items) x wavelengths
        unsigned long long ms[2] = {0, 0};
        PB_CUDA(c, cudaMemcpy(ms, c->d_mol_stat + 2, sizeof ms, cudaMemcpyDeviceToHost));
        c->last.executed_samples = (int64_t)ms[1] * c->last_n_wav_sel;
        c->last.exp_samples = (int64_t)ms[0] * c->last_n_wav_sel;
        c->mol_stats_pending = false;
    }
    if (flag) {
        PB_CUDA(c, cudaMemsetAsync(c->d_oob, 0, sizeof(int), c->stream));
        return fail(c, PB200_ERR_BOUNDS, "tabulated density queried outside its grid (the reference raises ValueError)");
    }
    return PB200_OK;
}

}  // namespace

static int need_ctx_early(pb200_ctx* c) {
    if (!c) return PB200_ERR_INVALID;
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return PB200_OK;
}

extern "C" int pb200_run(pb200_ctx* c, int phase_lo, int phase_hi, int64_t wav_lo, int64_t wav_hi, double* R_host,
                         pb200_timing* timing) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!R_host) return fail(c, PB200_ERR_INVALID, "R_host == NULL");
    const size_t bytes = (size_t)(phase_hi - phase_lo) * (size_t)(wav_hi - wav_lo) * sizeof(double);
    // A result array in page-locked memory of pb200_host_alloc / pb200_host_register is written by the kernels
    // themselves (zero-copy stores over PCIe as the windows finish) instead of device scratch + a D2H copy
    // afterwards; PB200_OPT_ZEROCOPY = 0 (measurement) keeps the copy.
    double* mapped = c->opt[PB200_OPT_ZEROCOPY] ? device_view(R_host, bytes) : nullptr;
    // the sorted kernel scatters R through a permutation: 8-byte stores over PCIe would be slow, so it writes to
    // device scratch and R crosses PCIe as one copy
    if (mapped && wav_lo >= 0 && wav_hi > wav_lo && phase_hi > phase_lo && want_sorted(c, phase_hi - phase_lo, wav_hi - wav_lo) &&
        !want_cells(c, phase_hi - phase_lo, wav_hi - wav_lo)) mapped = nullptr;   // (the cell path stores R in wavelength order)
    rc = mapped ? enqueue_run(c, phase_lo, phase_hi, wav_lo, wav_hi, mapped, false)
                : enqueue_run(c, phase_lo, phase_hi, wav_lo, wav_hi, nullptr, true);
    if (rc) return rc;
    cudaEvent_t e0 = c->ev[0], e1 = c->ev[1];
    PB_CUDA(c, cudaEventRecord(e0, c->stream));
    if (!mapped) PB_CUDA(c, cudaMemcpyAsync(R_host, c->scratch, bytes, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(c, cudaEventRecord(e1, c->stream));
    rc = check_oob(c);   // synchronises the stream
    if (rc) return rc;
    PB_CUDA(c, cudaGetLastError());
    rc = collect_timing(c);
    if (rc) return rc;
    float ms = 0;
    PB_CUDA(c, cudaEventElapsedTime(&ms, e0, e1));
    c->last.download_ms = ms;
    if (timing) *timing = c->last;
    return PB200_OK;
}

extern "C" int pb200_run_device(pb200_ctx* c, int phase_lo, int phase_hi, int64_t wav_lo, int64_t wav_hi,
                                double* R_dev, pb200_timing* timing) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!R_dev) return fail(c, PB200_ERR_INVALID, "R_dev == NULL");
    rc = enqueue_run(c, phase_lo, phase_hi, wav_lo, wav_hi, R_dev, false);
    if (rc) return rc;
    if (timing) {   // the caller asked for timings: this needs a synchronisation
        rc = check_oob(c);
        if (rc) return rc;
        rc = collect_timing(c);
        if (rc) return rc;
        *timing = c->last;
    }
    return PB200_OK;
}

// number of wavelengths of rank `rank` in a block-cyclic split of n_wav into blocks of `block` over n_ranks
static int64_t cyclic_count(int64_t n_wav, int64_t block, int n_ranks, int rank) {
    const int64_t n_blocks = (n_wav + block - 1) / block;
    int64_t n = 0;
    for (int64_t b = rank; b < n_blocks; b += n_ranks) n += (b + 1) * block <= n_wav ? block : n_wav - b * block;
    return n;
}

extern "C" int64_t pb200_cyclic_count(int64_t n_wav, int64_t wav_block, int n_ranks, int rank) {
    if (n_wav < 0 || wav_block <= 0 || n_ranks <= 0 || rank < 0 || rank >= n_ranks) return -1;
    return cyclic_count(n_wav, wav_block, n_ranks, rank);
}

extern "C" int pb200_run_device_cyclic(pb200_ctx* c, int phase_lo, int phase_hi, int64_t wav_block, int n_ranks, int rank,
                                       double* R_dev, pb200_timing* timing) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!R_dev) return fail(c, PB200_ERR_INVALID, "R_dev == NULL");
    if (wav_block <= 0 || n_ranks <= 0 || rank < 0 || rank >= n_ranks)
        return fail(c, PB200_ERR_INVALID, "block-cyclic split: block %lld, rank %d of %d", (long long)wav_block, rank, n_ranks);
    const int64_t n_local = cyclic_count(c->h_scene.n_wav, wav_block, n_ranks, rank);
    if (n_local == 0) return PB200_OK;   // more ranks than blocks: nothing to do
    rc = enqueue_run(c, phase_lo, phase_hi, 0, n_local, R_dev, false, wav_block, n_ranks, rank);
    if (rc) return rc;
    if (timing) {
        rc = check_oob(c);
        if (rc) return rc;
        rc = collect_timing(c);
        if (rc) return rc;
        *timing = c->last;
    }
    return PB200_OK;
}

// waits for the context's stream; reports a tabulated-density out-of-grid condition of the runs since the last check
extern "C" int pb200_sync(pb200_ctx* c) {
    int rc = need_ctx_early(c);
    if (rc) return rc;
    if (!c->has_scene) { PB_CUDA(c, cudaStreamSynchronize(c->stream)); return PB200_OK; }
    rc = check_oob(c);
    if (rc) return rc;
    PB_CUDA(c, cudaGetLastError());
    return PB200_OK;
}

// The rank's block-cyclic tile written IN PLACE into the caller's full result: R_full is row phase_lo of an array
// with row_stride doubles per phase that the device can address -- mapped page-locked host memory
// (pb200_host_alloc / pb200_host_register: the D2H transfer is then the kernel's own stores, every rank over its
// own PCIe link) or a peer GPU's buffer (stores over NVLink).  n_ranks == 1: the whole wavelength axis.
extern "C" int pb200_run_scatter(pb200_ctx* c, int phase_lo, int phase_hi, int64_t wav_block, int n_ranks, int rank,
                                 double* R_full, int64_t row_stride, pb200_timing* timing) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!R_full) return fail(c, PB200_ERR_INVALID, "R_full == NULL");
    if (n_ranks <= 0 || rank < 0 || rank >= n_ranks || (n_ranks > 1 && wav_block <= 0) || row_stride < c->h_scene.n_wav)
        return fail(c, PB200_ERR_INVALID, "pb200_run_scatter: block %lld, rank %d of %d, row stride %lld", (long long)wav_block, rank, n_ranks, (long long)row_stride);
    double* dst = R_full;
    if (double* m = device_view(R_full, sizeof(double))) dst = m;   // a host address of a page-locked block: use its device view
    if (n_ranks == 1) {
        rc = enqueue_run(c, phase_lo, phase_hi, 0, c->h_scene.n_wav, dst, false, 0, 1, 0, row_stride);
    } else {
        const int64_t n_local = cyclic_count(c->h_scene.n_wav, wav_block, n_ranks, rank);
        if (n_local == 0) return PB200_OK;
        rc = enqueue_run(c, phase_lo, phase_hi, 0, n_local, dst, false, wav_block, n_ranks, rank, row_stride);
    }
    if (rc) return rc;
    if (timing) {
        rc = check_oob(c);
        if (rc) return rc;
        rc = collect_timing(c);
        if (rc) return rc;
        *timing = c->last;
    }
    return PB200_OK;
}

// =================================================================================================
// inner seams (materialised) -- API parity with the methods sumOverChords calls
// =================================================================================================
namespace {

// temporary device buffers for host-pointer entry points
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 8); }
    template <typename T> T* as() { return static_cast<T*>(p); }
};

int upload(pb200_ctx* c, DevBuf& b, const void* src, size_t bytes) {
    PB_CUDA(c, b.alloc(bytes));
    PB_CUDA(c, cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return PB200_OK;
}

int download(pb200_ctx* c, void* dst, const void* src, size_t bytes) {
    PB_CUDA(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    PB_CUDA(c, cudaGetLastError());
    return PB200_OK;
}

int need_ctx(pb200_ctx* c) {
    if (!c) return PB200_ERR_INVALID;
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return PB200_OK;
}

}  // namespace

// =================================================================================================
// retrieval batches: models stacked along the phase axis, light-curve epilogue
// =================================================================================================
extern "C" int pb200_set_models(pb200_ctx* c, int n_models, const pb200_model* models) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (n_models < 0 || (n_models > 0 && !models)) return fail(c, PB200_ERR_INVALID, "pb200_set_models: n_models=%d", n_models);
    const DevScene& b = c->base;
    DevScene& h = c->h_scene;
    c->lastR_full = false;
    if (n_models == 0) {   // back to the scene as uploaded
        h = b;
    } else {
        const int nb = b.n_orb;
        const int64_t n_tot64 = (int64_t)n_models * nb;
        if (n_tot64 > (1 << 24)) return fail(c, PB200_ERR_INVALID, "batch of %d models x %d phases is too large", n_models, nb);
        const int n_tot = (int)n_tot64;
        for (int m = 0; m < n_models; ++m) {
            const pb200_model& q = models[m];
            if (!q.scenarios || !q.planet_y || (b.n_moons && (!q.moon_R || !q.moon_y)))
                return fail(c, PB200_ERR_INVALID, "model %d has NULL array pointers", m);
            for (int i = 0; i < b.n_scen; ++i) {
                const pb200_scenario& qs = q.scenarios[i];
                if (qs.type != b.scen[i].type || qs.n_constituents != b.scen[i].n_cons || qs.first_constituent != b.scen[i].first_cons)
                    return fail(c, PB200_ERR_INVALID, "model %d scenario %d: type / constituents differ from the resident scene", m, i);
                if (!qs.shift || (qs.type != PB200_SCN_SERPENS && (!qs.centre_x || !qs.centre_y)))
                    return fail(c, PB200_ERR_INVALID, "model %d scenario %d: per-phase arrays are NULL", m, i);
            }
        }
        // layout of the model arena, in doubles
        size_t off = 0;
        auto take = [&](size_t n) { const size_t o = off; off += (n + 31) & ~size_t(31); return o; };
        const size_t o_py = take(n_tot), o_pR = take(n_tot);
        const size_t o_mR = b.n_moons ? take((size_t)b.n_moons * n_tot) : 0;
        const size_t o_my = b.n_moons ? take((size_t)b.n_moons * n_tot) : 0;
        struct SO { size_t cx, cy, sh, amp, pp; bool has_centre, has_amp; };
        SO so[kMaxScenarios];
        for (int i = 0; i < b.n_scen; ++i) {
            so[i].has_centre = b.scen[i].type != PB200_SCN_SERPENS;
            so[i].has_amp = b.scen[i].amplitude != nullptr;
            for (int m = 0; m < n_models; ++m) so[i].has_amp |= models[m].scenarios[i].amplitude != nullptr;
            so[i].cx = so[i].has_centre ? take(n_tot) : 0;
            so[i].cy = so[i].has_centre ? take(n_tot) : 0;
            so[i].sh = take(n_tot);
            so[i].amp = so[i].has_amp ? take(n_tot) : 0;
            so[i].pp = take((size_t)n_tot * 8);
        }
        const size_t bytes = off * sizeof(double);
        PB_CUDA(c, cudaSetDevice(c->device));
        if (c->mstage_busy) { PB_CUDA(c, cudaStreamSynchronize(c->stream)); c->mstage_busy = false; }
        if (bytes > c->mstage_bytes) {
            if (c->mstage) cudaFreeHost(c->mstage);
            c->mstage = nullptr; c->mstage_bytes = 0;
            PB_CUDA(c, cudaMallocHost(&c->mstage, bytes));
            c->mstage_bytes = bytes;
        }
        if (bytes > c->marena_bytes) {
            if (c->marena) PB_CUDA(c, cudaFree(c->marena));
            c->marena = nullptr; c->marena_bytes = 0;
            PB_CUDA(c, cudaMalloc(&c->marena, bytes));
            c->marena_bytes = bytes;
        }
        double* st = static_cast<double*>(c->mstage);
        const size_t row = (size_t)nb * sizeof(double);
        for (int m = 0; m < n_models; ++m) {
            const pb200_model& q = models[m];
            const size_t r0 = (size_t)m * nb;
            memcpy(st + o_py + r0, q.planet_y, row);
            for (int p = 0; p < nb; ++p) st[o_pR + r0 + p] = q.planet_R > 0.0 ? q.planet_R : b.planet_R;
            for (int k = 0; k < b.n_moons; ++k) {
                memcpy(st + o_my + (size_t)k * n_tot + r0, q.moon_y + (size_t)k * nb, row);
                for (int p = 0; p < nb; ++p) st[o_mR + (size_t)k * n_tot + r0 + p] = q.moon_R[k];
            }
            for (int i = 0; i < b.n_scen; ++i) {
                const pb200_scenario& qs = q.scenarios[i];
                if (so[i].has_centre) { memcpy(st + so[i].cx + r0, qs.centre_x, row); memcpy(st + so[i].cy + r0, qs.centre_y, row); }
                memcpy(st + so[i].sh + r0, qs.shift, row);
                if (so[i].has_amp)
                    for (int p = 0; p < nb; ++p) st[so[i].amp + r0 + p] = qs.amplitude ? qs.amplitude[p] : 1.0;
                for (int p = 0; p < nb; ++p) {
                    memcpy(st + so[i].pp + (r0 + p) * 8, qs.p, 7 * sizeof(double));
                    st[so[i].pp + (r0 + p) * 8 + 7] = qs.T;   // pb_scn_T
                }
            }
        }
        PB_CUDA(c, cudaMemcpyAsync(c->marena, c->mstage, bytes, cudaMemcpyHostToDevice, c->stream));
        c->mstage_busy = true;
        const double* A = static_cast<const double*>(c->marena);
        h = b;
        h.n_orb = n_tot; h.n_orb_model = nb;
        h.planet_y = A + o_py;
        h.planet_R_phase = A + o_pR;
        if (b.n_moons) { h.moon_R = A + o_mR; h.moon_y = A + o_my; h.moon_R_per_phase = 1; }
        for (int i = 0; i < b.n_scen; ++i) {
            DevScenario& d = h.scen[i];
            if (so[i].has_centre) { d.centre_x = A + so[i].cx; d.centre_y = A + so[i].cy; }
            d.shift = A + so[i].sh;
            d.amplitude = so[i].has_amp ? A + so[i].amp : nullptr;
            d.p_phase = A + so[i].pp;
            memcpy(d.p, models[0].scenarios[i].p, sizeof d.p);
        }
    }
    rc = plan_derived(c);
    if (rc) return rc;
    c->mol_max_items = -1;   // other densities: other item counts
    PB_CUDA(c, cudaMemcpyAsync(c->d_scene, &h, sizeof(DevScene), cudaMemcpyHostToDevice, c->stream));
    launch_sky_cells(c->d_scene, h.Mp, c->stream);   // the derived arena was laid out afresh
    launch_fout_scalar(c->d_scene, c->stream);
    c->launches += 2;
    PB_CUDA(c, cudaGetLastError());
    return PB200_OK;
}

namespace {

int ensure_band(pb200_ctx* c, size_t doubles) {
    const size_t bytes = doubles * sizeof(double);
    if (bytes <= c->band_bytes) return PB200_OK;
    if (c->d_band) { PB_CUDA(c, cudaFree(c->d_band)); c->d_band = nullptr; c->band_bytes = 0; }
    PB_CUDA(c, cudaMalloc((void**)&c->d_band, bytes));
    c->band_bytes = bytes;
    return PB200_OK;
}

// light curves of rows [phase_lo, phase_hi) of R_rows (device) into d_band + n_orb_model; not synchronised
int enqueue_lightcurve(pb200_ctx* c, const pb200_band* band, int phase_lo, int phase_hi, const double* R_rows) {
    const DevScene& h = c->h_scene;
    if (!band || band->n_lines < 1 || band->n_lines > 16 || !band->line_centre || !band->shift)
        return fail(c, PB200_ERR_INVALID, "pb200_band: 1..16 line centres and a shift array are required");
    if (phase_lo < 0 || phase_hi > h.n_orb || phase_lo >= phase_hi)
        return fail(c, PB200_ERR_INVALID, "rows [%d,%d) outside [0,%d)", phase_lo, phase_hi, h.n_orb);
    const int nb = h.n_orb_model, n_rows = phase_hi - phase_lo;
    int rc = ensure_band(c, (size_t)nb + n_rows);
    if (rc) return rc;
    PB_CUDA(c, cudaMemcpyAsync(c->d_band, band->shift, (size_t)nb * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    cudaError_t e = launch_lightcurve(R_rows, h.wavelength, h.n_wav, n_rows, phase_lo, nb, *band, c->d_band, c->d_band + nb, c->stream);
    if (e != cudaSuccess) return fail(c, PB200_ERR_CUDA, "lightcurve launch: %s", cudaGetErrorString(e));
    c->launches += 1;
    return PB200_OK;
}

}  // namespace

extern "C" int pb200_lightcurve(pb200_ctx* c, const pb200_band* band, int phase_lo, int phase_hi, const double* R_dev,
                                double* lightcurve_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!lightcurve_host) return fail(c, PB200_ERR_INVALID, "lightcurve_host == NULL");
    const double* rows = R_dev;
    if (!rows) {
        if (!c->lastR_full || phase_lo < c->lastR_phase_lo || phase_hi > c->lastR_phase_lo + c->lastR_n_phase)
            return fail(c, PB200_ERR_INVALID, "no resident R covers rows [%d,%d): run pb200_run over the whole wavelength grid first", phase_lo, phase_hi);
        rows = c->lastR_ptr + (int64_t)(phase_lo - c->lastR_phase_lo) * c->h_scene.n_wav;
    }
    rc = enqueue_lightcurve(c, band, phase_lo, phase_hi, rows);
    if (rc) return rc;
    return download(c, lightcurve_host, c->d_band + c->h_scene.n_orb_model, (size_t)(phase_hi - phase_lo) * sizeof(double));
}

extern "C" int pb200_run_batch(pb200_ctx* c, int n_models, const pb200_model* models, const pb200_band* band,
                               double* lightcurve_host, double* R_host, pb200_timing* timing) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (n_models < 1) return fail(c, PB200_ERR_INVALID, "pb200_run_batch: n_models=%d", n_models);
    if (!lightcurve_host && !R_host) return fail(c, PB200_ERR_INVALID, "pb200_run_batch: nothing to return");
    if (lightcurve_host && !band) return fail(c, PB200_ERR_INVALID, "pb200_run_batch: a band is needed for light curves");
    PB_CUDA(c, cudaEventRecord(c->ev[6], c->stream));
    rc = pb200_set_models(c, n_models, models);
    if (rc) return rc;
    PB_CUDA(c, cudaEventRecord(c->ev[7], c->stream));
    c->upload_ms = -1.0;
    const DevScene& h = c->h_scene;
    rc = enqueue_run(c, 0, h.n_orb, 0, h.n_wav, nullptr, true);
    if (rc) return rc;
    PB_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    if (lightcurve_host) {
        rc = enqueue_lightcurve(c, band, 0, h.n_orb, c->scratch);
        if (rc) return rc;
        PB_CUDA(c, cudaMemcpyAsync(lightcurve_host, c->d_band + h.n_orb_model, (size_t)h.n_orb * sizeof(double),
                                   cudaMemcpyDeviceToHost, c->stream));
    }
    if (R_host)
        PB_CUDA(c, cudaMemcpyAsync(R_host, c->scratch, (size_t)h.n_orb * h.n_wav * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
    rc = check_oob(c);   // synchronises the stream
    c->mstage_busy = false;
    if (rc) return rc;
    PB_CUDA(c, cudaGetLastError());
    rc = collect_timing(c);
    if (rc) return rc;
    float ms = 0;
    PB_CUDA(c, cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last.download_ms = ms;   // light-curve kernel + D2H
    if (timing) *timing = c->last;
    return PB200_OK;
}


extern "C" int pb200_number_density(pb200_ctx* c, int scenario, double* n_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (scenario < 0 || scenario >= h.n_scen || !n_host) return fail(c, PB200_ERR_INVALID, "bad scenario index %d", scenario);
    const int64_t n_c = (int64_t)h.M * h.n_orb;
    DevBuf out;
    PB_CUDA(c, out.alloc((size_t)n_c * h.n_x * sizeof(double)));
    launch_number_density_grid(c->d_scene, n_c, scenario, out.as<double>(), c->d_oob, c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    rc = check_oob(c);
    if (rc) return rc;
    return download(c, n_host, out.p, (size_t)n_c * h.n_x * sizeof(double));
}

// the same into device memory, not synchronised: the bandwidth measurement of kernel (a) (bench.py)
extern "C" int pb200_number_density_device(pb200_ctx* c, int scenario, double* n_dev) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (scenario < 0 || scenario >= h.n_scen || !n_dev) return fail(c, PB200_ERR_INVALID, "bad scenario index %d", scenario);
    if (h.scen[scenario].type == PB200_SCN_SERPENS)
        launch_number_density_grid(c->d_scene, (int64_t)h.M * h.n_orb, scenario, n_dev, c->d_oob, c->sm_count, c->stream);
    else
        launch_number_density_thread(c->d_scene, (int64_t)h.M * h.n_orb, h.n_x, scenario, n_dev, c->sm_count,
                                     reinterpret_cast<unsigned long long*>(c->d_oob) + 3, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    return PB200_OK;
}

extern "C" int pb200_number_density_chords(pb200_ctx* c, int scenario, int64_t n, const double* y, const double* z,
                                           const double* cx, const double* cy, const double* amp, double* n_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (scenario < 0 || scenario >= h.n_scen || n < 0 || !y || !z || !n_host)
        return fail(c, PB200_ERR_INVALID, "pb200_number_density_chords: bad arguments");
    if (n == 0) return PB200_OK;
    const bool serp = h.scen[scenario].type == PB200_SCN_SERPENS;
    if (!serp && (!cx || !cy)) return fail(c, PB200_ERR_INVALID, "centre arrays required");
    DevBuf dy, dz, dcx, dcy, damp, out;
    const size_t nb = (size_t)n * sizeof(double);
    if ((rc = upload(c, dy, y, nb)) || (rc = upload(c, dz, z, nb))) return rc;
    if (cx && (rc = upload(c, dcx, cx, nb))) return rc;
    if (cy && (rc = upload(c, dcy, cy, nb))) return rc;
    if (amp && (rc = upload(c, damp, amp, nb))) return rc;
    PB_CUDA(c, out.alloc(nb * h.n_x));
    launch_number_density_chords(c->d_scene, scenario, n, dy.as<double>(), dz.as<double>(), dcx.as<double>(),
                                 dcy.as<double>(), amp ? damp.as<double>() : nullptr, out.as<double>(), c->d_oob,
                                 c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    rc = check_oob(c);
    if (rc) return rc;
    return download(c, n_host, out.p, nb * h.n_x);
}

// phase-major device rows -> the reference's chord order c = m*n_orb + p
static void to_chord_order(const double* src, int n_orb, int M, int Mp, double* dst) {
    for (int p = 0; p < n_orb; ++p)
        for (int m = 0; m < M; ++m) dst[(int64_t)m * n_orb + p] = src[(int64_t)p * Mp + m];
}

extern "C" int pb200_column_density(pb200_ctx* c, double* ncol_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (!ncol_host) return fail(c, PB200_ERR_INVALID, "ncol_host == NULL");
    launch_column(c->d_scene, h.Mp, 0, h.n_orb, c->d_oob, c->sm_count, c->stream, h.n_mols > 0, 0);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    rc = check_oob(c);
    if (rc) return rc;
    const size_t per_atom = (size_t)h.n_orb * h.Mp;
    std::vector<double> tmp(per_atom * (h.n_atoms ? h.n_atoms : 1));
    rc = download(c, tmp.data(), h.ncol, per_atom * h.n_atoms * sizeof(double));
    if (rc) return rc;
    for (int a = 0; a < h.n_atoms; ++a)
        to_chord_order(tmp.data() + a * per_atom, h.n_orb, h.M, h.Mp, ncol_host + (size_t)a * h.n_orb * h.M);
    return PB200_OK;
}

extern "C" int pb200_chord_masks(pb200_ctx* c, uint8_t* blocked_host, double* weight_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    launch_column(c->d_scene, h.Mp, 0, h.n_orb, c->d_oob, c->sm_count, c->stream, h.n_mols > 0, 0);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    std::vector<double> w_in((size_t)h.n_orb * h.Mp), w_out(h.Mp);
    if ((rc = download(c, w_in.data(), h.w_in, w_in.size() * sizeof(double)))) return rc;
    if ((rc = download(c, w_out.data(), h.w_out, w_out.size() * sizeof(double)))) return rc;
    for (int p = 0; p < h.n_orb; ++p)
        for (int m = 0; m < h.M; ++m) {
            const int64_t cidx = (int64_t)m * h.n_orb + p;
            // every rho midpoint is > 0 and clv > 0 for physical u1,u2, so w_in == 0 <=> blocked
            if (blocked_host) blocked_host[cidx] = (w_in[(size_t)p * h.Mp + m] == 0.0 && w_out[m] != 0.0) ? 1 : 0;
            if (weight_host) weight_host[cidx] = w_out[m];
        }
    return PB200_OK;
}

extern "C" int pb200_optical_depth(pb200_ctx* c, int64_t wav_lo, int64_t wav_hi, double* tau_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (wav_lo < 0 || wav_hi > h.n_wav || wav_lo >= wav_hi || !tau_host) return fail(c, PB200_ERR_INVALID, "bad wavelength range");
    const int64_t n_wav_sel = wav_hi - wav_lo, n_c = (int64_t)h.M * h.n_orb;
    DevBuf sig, tau;
    PB_CUDA(c, sig.alloc((size_t)(h.n_atoms ? h.n_atoms : 1) * h.n_orb * n_wav_sel * sizeof(double)));
    PB_CUDA(c, tau.alloc((size_t)n_c * n_wav_sel * sizeof(double)));
    launch_column(c->d_scene, h.Mp, 0, h.n_orb, c->d_oob, c->sm_count, c->stream, h.n_mols > 0, 0);
    if (h.n_atoms) launch_sigma_u(c->d_scene, h.n_atoms, h.n_orb, wav_lo, n_wav_sel, sig.as<double>(), c->sm_count, c->stream);
    launch_optical_depth(c->d_scene, sig.as<double>(), n_c, wav_lo, n_wav_sel, tau.as<double>(), c->stream);
    c->launches += 2 + (h.n_atoms ? 1 : 0);
    PB_CUDA(c, cudaGetLastError());
    rc = check_oob(c);
    if (rc) return rc;
    return download(c, tau_host, tau.p, (size_t)n_c * n_wav_sel * sizeof(double));
}

// Atmosphere.getLOSopticalDepth_Batch for caller-supplied chords (gasProperties.py:885-956): tau[n][wav_hi - wav_lo].
// y, z [n]; cx, cy, shift [n_scenarios][n] (cx, cy ignored for a tabulated density); amp NULL or [n_scenarios][n].
extern "C" int pb200_optical_depth_chords(pb200_ctx* c, int64_t n, const double* y, const double* z, const double* cx,
                                          const double* cy, const double* shift, const double* amp, int64_t wav_lo,
                                          int64_t wav_hi, double* tau_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    const DevScene& h = c->h_scene;
    if (n < 0 || wav_lo < 0 || wav_hi > h.n_wav || wav_lo >= wav_hi || (n && (!y || !z || !cx || !cy || !shift || !tau_host)))
        return fail(c, PB200_ERR_INVALID, "pb200_optical_depth_chords: bad arguments");
    if (n == 0) return PB200_OK;
    if (h.n_orb_model != h.n_orb) return fail(c, PB200_ERR_INVALID, "pb200_optical_depth_chords: not with a stacked retrieval batch");
    const int64_t n_wav_sel = wav_hi - wav_lo;
    const size_t nb = (size_t)n * sizeof(double), sb = nb * h.n_scen;
    DevBuf dy, dz, dcx, dcy, dsh, damp, tau;
    if ((rc = upload(c, dy, y, nb)) || (rc = upload(c, dz, z, nb)) || (rc = upload(c, dcx, cx, sb)) || (rc = upload(c, dcy, cy, sb)) ||
        (rc = upload(c, dsh, shift, sb)))
        return rc;
    if (amp && (rc = upload(c, damp, amp, sb))) return rc;
    PB_CUDA(c, tau.alloc((size_t)n * n_wav_sel * sizeof(double)));
    if (!c->buckets_ready && h.n_atoms > 0) {
        launch_bucket_build(c->d_scene, c->sm_count, c->stream);
        c->launches += 1;
        c->buckets_ready = true;
    }
    int n_slots = 0;
    for (int i = 0; i < h.n_scen; ++i) n_slots += h.scen[i].density_slot >= 0 ? 1 : 0;
    cudaError_t e = launch_optical_depth_chords(c->d_scene, h.n_x, n_slots, n, dy.as<double>(), dz.as<double>(), dcx.as<double>(),
                                                dcy.as<double>(), dsh.as<double>(), amp ? damp.as<double>() : nullptr, wav_lo,
                                                n_wav_sel, tau.as<double>(), c->d_oob, c->stream);
    if (e != cudaSuccess) return fail(c, e == cudaErrorInvalidValue ? PB200_ERR_UNSUPPORTED : PB200_ERR_CUDA,
                                      "pb200_optical_depth_chords: %s", cudaGetErrorString(e));
    c->launches += 1;
    rc = check_oob(c);
    if (rc) return rc;
    return download(c, tau_host, tau.p, (size_t)n * n_wav_sel * sizeof(double));
}

// =================================================================================================
// cross sections
// =================================================================================================
extern "C" int pb200_sigma_lookup(pb200_ctx* c, const double* tab_x, const double* tab_y, int64_t n_tab, double offset,
                                  const double* wl, int64_t n, double* out_host) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (!tab_x || !tab_y || n_tab < 1 || n < 0 || (n && (!wl || !out_host))) return fail(c, PB200_ERR_INVALID, "pb200_sigma_lookup: bad arguments");
    if (n == 0) return PB200_OK;
    DevBuf tx, ty, dw, out;
    if ((rc = upload(c, tx, tab_x, n_tab * sizeof(double))) || (rc = upload(c, ty, tab_y, n_tab * sizeof(double))) ||
        (rc = upload(c, dw, wl, n * sizeof(double))))
        return rc;
    PB_CUDA(c, out.alloc(n * sizeof(double)));
    launch_sigma_lookup(tx.as<double>(), ty.as<double>(), n_tab, offset, dw.as<double>(), n, out.as<double>(), c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    return download(c, out_host, out.p, n * sizeof(double));
}

extern "C" int pb200_voigt_sigma(pb200_ctx* c, const double* line_wl, const double* line_gamma, const double* line_f,
                                 int64_t n_lines, double sigma_v, const double* wl, int64_t n, double offset,
                                 double* out_sigma, double* out_log10) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (n_lines < 0 || n < 0 || (n_lines && (!line_wl || !line_gamma || !line_f)) || (n && !wl))
        return fail(c, PB200_ERR_INVALID, "pb200_voigt_sigma: bad arguments");
    if (n == 0) return PB200_OK;
    DevBuf lw, lg, lf, dw, os, ol;
    const size_t lb = (size_t)n_lines * sizeof(double), nb = (size_t)n * sizeof(double);
    if ((rc = upload(c, lw, line_wl, lb)) || (rc = upload(c, lg, line_gamma, lb)) || (rc = upload(c, lf, line_f, lb)) ||
        (rc = upload(c, dw, wl, nb)))
        return rc;
    PB_CUDA(c, os.alloc(nb));
    PB_CUDA(c, ol.alloc(nb));
    PB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
    launch_voigt_sigma(lw.as<double>(), lg.as<double>(), lf.as<double>(), n_lines, sigma_v, dw.as<double>(), n, offset,
                       os.as<double>(), ol.as<double>(), c->stream);
    PB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    if (out_sigma && (rc = download(c, out_sigma, os.p, nb))) return rc;
    if (out_log10 && (rc = download(c, out_log10, ol.p, nb))) return rc;
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0;
    PB_CUDA(c, cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]));
    c->last = pb200_timing{};
    c->last.transit_ms = ms;   // kernel (b2) time, read back through pb200_last_timing
    c->last.launches = 1;
    c->last.samples = n * n_lines;
    return PB200_OK;
}

// out[r][c] = log10(values[r][reverse ? n_cols-1-c : c] + offset): the elementwise part of the table producers
extern "C" int pb200_log_table(pb200_ctx* c, const double* values, int64_t n_rows, int64_t n_cols, int reverse_cols,
                               double offset, double* out_host) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (n_rows < 0 || n_cols < 0 || (n_rows * n_cols && (!values || !out_host))) return fail(c, PB200_ERR_INVALID, "pb200_log_table: bad arguments");
    if (n_rows * n_cols == 0) return PB200_OK;
    DevBuf in, out;
    const size_t nb = (size_t)n_rows * n_cols * sizeof(double);
    if ((rc = upload(c, in, values, nb))) return rc;
    PB_CUDA(c, out.alloc(nb));
    launch_log_table(in.as<double>(), n_rows, n_cols, reverse_cols ? 1 : 0, offset, out.as<double>(), c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    return download(c, out_host, out.p, nb);
}

extern "C" int pb200_voigt_profile(pb200_ctx* c, const double* x, const double* sigma, const double* gamma, int64_t n,
                                   double* out_host) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (n < 0 || (n && (!x || !sigma || !gamma || !out_host))) return fail(c, PB200_ERR_INVALID, "pb200_voigt_profile: bad arguments");
    if (n == 0) return PB200_OK;
    DevBuf dx, ds, dg, out;
    const size_t nb = (size_t)n * sizeof(double);
    if ((rc = upload(c, dx, x, nb)) || (rc = upload(c, ds, sigma, nb)) || (rc = upload(c, dg, gamma, nb))) return rc;
    PB_CUDA(c, out.alloc(nb));
    launch_voigt_profile(dx.as<double>(), ds.as<double>(), dg.as<double>(), n, out.as<double>(), c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    return download(c, out_host, out.p, nb);
}

extern "C" int pb200_molecular_sigma(pb200_ctx* c, const double* grid_P, const double* grid_T, const double* grid_w,
                                     const double* values, const int32_t dims[3], double fill, double offset,
                                     const double* P, double T, const double* wl, int64_t n_c, int64_t n_x,
                                     int64_t n_wav, double* out_host) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (!grid_P || !grid_T || !grid_w || !values || !dims || dims[0] < 2 || dims[1] < 2 || dims[2] < 2 || !P || !wl ||
        !out_host || n_c < 0 || n_x < 0 || n_wav < 0)
        return fail(c, PB200_ERR_INVALID, "pb200_molecular_sigma: bad arguments");
    if (n_c * n_x * n_wav == 0) return PB200_OK;
    DevBuf gP, gT, gW, val, dP, dw, out;
    const size_t D = sizeof(double);
    if ((rc = upload(c, gP, grid_P, dims[0] * D)) || (rc = upload(c, gT, grid_T, dims[1] * D)) ||
        (rc = upload(c, gW, grid_w, dims[2] * D)) || (rc = upload(c, val, values, (size_t)dims[0] * dims[1] * dims[2] * D)) ||
        (rc = upload(c, dP, P, (size_t)n_c * n_x * D)) || (rc = upload(c, dw, wl, (size_t)n_c * n_wav * D)))
        return rc;
    PB_CUDA(c, out.alloc((size_t)n_c * n_x * n_wav * D));
    MolTable t;
    t.gP = gP.as<double>(); t.gT = gT.as<double>(); t.gW = gW.as<double>(); t.values = val.as<double>();
    t.nP = dims[0]; t.nT = dims[1]; t.nW = dims[2]; t.fill = fill;
    launch_molecular_sigma(t, offset, dP.as<double>(), T, dw.as<double>(), n_c, n_x, n_wav, out.as<double>(), c->sm_count, c->stream);
    c->launches += 1;
    PB_CUDA(c, cudaGetLastError());
    return download(c, out_host, out.p, (size_t)n_c * n_x * n_wav * D);
}

// =================================================================================================
// SerpensExosphere.addInterpolatedDensity (gasProperties.py:548-583): particle histogram + Gaussian smoothing
// =================================================================================================
extern "C" int pb200_serpens_density(pb200_ctx* c, const double* pos, int64_t n, const double* xb, int32_t nxb,
                                     const double* yb, int32_t nyb, const double* zb, int32_t nzb, double N, double denom,
                                     const double* weights, int32_t radius, const double* y_points, const double* z_points,
                                     double planet_R, double star_R, double* out_host, double* sums_host) {
    int rc = need_ctx(c);
    if (rc) return rc;
    if (n < 0 || (n && !pos) || !xb || !yb || !zb || nxb < 2 || nyb < 2 || nzb < 2 || !out_host || radius < 0 ||
        (radius > 0 && !weights) || (sums_host && (!y_points || !z_points)))
        return fail(c, PB200_ERR_INVALID, "pb200_serpens_density: bad arguments");
    const int nx = nxb - 1, ny = nyb - 1, nz = nzb - 1;
    const int64_t cells = (int64_t)nx * ny * nz;
    DevBuf dpos, dxb, dyb, dzb, dcnt, da, db, dw, dyp, dzp, dsum;
    const size_t D = sizeof(double);
    if ((rc = upload(c, dpos, pos, (size_t)n * 3 * D)) || (rc = upload(c, dxb, xb, nxb * D)) || (rc = upload(c, dyb, yb, nyb * D)) ||
        (rc = upload(c, dzb, zb, nzb * D)))
        return rc;
    PB_CUDA(c, dcnt.alloc((size_t)cells * sizeof(unsigned int)));
    PB_CUDA(c, da.alloc((size_t)cells * D));
    PB_CUDA(c, cudaMemsetAsync(dcnt.p, 0, (size_t)cells * sizeof(unsigned int), c->stream));
    if (n) launch_serpens_histogram(dpos.as<double>(), n, dxb.as<double>(), nxb, dyb.as<double>(), nyb, dzb.as<double>(), nzb,
                                    dcnt.as<unsigned int>(), c->sm_count, c->stream);
    launch_serpens_scale(dcnt.as<unsigned int>(), cells, N, denom, da.as<double>(), c->stream);
    c->launches += n ? 2 : 1;
    double* result = da.as<double>();
    if (radius > 0) {   // gaussian_filter: one correlate1d pass per axis, x then y then z
        PB_CUDA(c, db.alloc((size_t)cells * D));
        if ((rc = upload(c, dw, weights, (size_t)(2 * radius + 1) * D))) return rc;
        const int len[3] = {nx, ny, nz};
        const int64_t stride[3] = {(int64_t)ny * nz, nz, 1};
        double* src = da.as<double>();
        double* dst = db.as<double>();
        for (int ax = 0; ax < 3; ++ax) {
            launch_serpens_filter_axis(src, dst, cells, len[ax], stride[ax], dw.as<double>(), radius, c->stream);
            double* t = src; src = dst; dst = t;
        }
        c->launches += 3;
        result = src;
    }
    if (sums_host) {
        if ((rc = upload(c, dyp, y_points, ny * D)) || (rc = upload(c, dzp, z_points, nz * D))) return rc;
        PB_CUDA(c, dsum.alloc(2 * D));
        launch_serpens_sums(result, nx, ny, nz, dyp.as<double>(), dzp.as<double>(), planet_R * planet_R, star_R * star_R,
                            dsum.as<double>(), c->stream);
        c->launches += 1;
        PB_CUDA(c, cudaMemcpyAsync(sums_host, dsum.p, 2 * D, cudaMemcpyDeviceToHost, c->stream));
    }
    PB_CUDA(c, cudaGetLastError());
    return download(c, out_host, result, (size_t)cells * D);
}

// Diagnostics: the order in which the last run of this context walked its wavelengths (sigma-sorted windows of kernel
// (c'), pb200_sort.cu).  perm[pi * n_wav_sel + i] = local wavelength index at sorted position i of phase pi,
// nsig_sorted the matching -sigma * 4096/ln2.  *n receives n_phase * n_wav_sel of that run, 0 if it was not sorted;
// the arrays (may be NULL) must hold that many entries.  Valid until the next run of the context.
extern "C" int pb200_last_sort_order(pb200_ctx* c, int64_t* n, int32_t* perm_host, double* nsig_sorted_host) {
    int rc = check_scene(c);
    if (rc) return rc;
    if (!n) return fail(c, PB200_ERR_INVALID, "pb200_last_sort_order: n == NULL");
    const int64_t cap = *n;
    *n = c->last_sort_n;
    if (c->last_sort_n == 0 || (!perm_host && !nsig_sorted_host)) return PB200_OK;
    if (cap < c->last_sort_n) return fail(c, PB200_ERR_INVALID, "pb200_last_sort_order: room for %lld entries, the run had %lld", (long long)cap, (long long)c->last_sort_n);
    if (perm_host) PB_CUDA(c, cudaMemcpyAsync(perm_host, c->last_perm, (size_t)c->last_sort_n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (nsig_sorted_host) PB_CUDA(c, cudaMemcpyAsync(nsig_sorted_host, c->last_nsig_sorted, (size_t)c->last_sort_n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    PB_CUDA(c, cudaStreamSynchronize(c->stream));
    return PB200_OK;
}

// accessors for entry points that live in other translation units (pb200_serpens.cu: pb200_serpens_histogram_static)
namespace pb200 {
int ctx_fail(pb200_ctx* c, int code, const char* msg) { return fail(c, code, "%s", msg); }
int ctx_info(pb200_ctx* c, int* device, cudaStream_t* stream, int* sm_count) {
    if (!c) return PB200_ERR_INVALID;
    *device = c->device; *stream = c->stream; *sm_count = c->sm_count;
    return PB200_OK;
}
void ctx_add_launches(pb200_ctx* c, int n) { c->launches += n; }
}  // namespace pb200

extern "C" int pb200_last_timing(pb200_ctx* c, pb200_timing* t) {
    if (!c || !t) return PB200_ERR_INVALID;
    *t = c->last;
    return PB200_OK;
}

// =================================================================================================
// machine probes
// =================================================================================================
extern "C" int pb200_probe_fp64(pb200_ctx* c, double* dfma_tflops, double* exp_gops) {
    int rc = need_ctx(c);
    if (rc) return rc;
    const int blocks = c->sm_count * 8, iters = 1 << 15;
    DevBuf out;
    PB_CUDA(c, out.alloc((size_t)blocks * 256 * sizeof(double)));
    float best_f = 1e30f, best_e = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        PB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
        launch_probe_dfma(out.as<double>(), blocks, iters, c->stream);
        PB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
        launch_probe_exp(out.as<double>(), blocks, iters / 8, c->d_exp_tab, c->stream);
        PB_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
        PB_CUDA(c, cudaStreamSynchronize(c->stream));
        float a, b;
        PB_CUDA(c, cudaEventElapsedTime(&a, c->ev[2], c->ev[3]));
        PB_CUDA(c, cudaEventElapsedTime(&b, c->ev[3], c->ev[4]));
        if (rep > 0) { best_f = a < best_f ? a : best_f; best_e = b < best_e ? b : best_e; }
    }
    c->launches += 8;
    const double n_thr = (double)blocks * 256;
    if (dfma_tflops) *dfma_tflops = n_thr * iters * 8 * 2 / (best_f * 1e-3) / 1e12;
    if (exp_gops) *exp_gops = n_thr * (iters / 8) * 8 / (best_e * 1e-3) / 1e9;
    return PB200_OK;
}
