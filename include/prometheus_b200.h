/*
 * prometheus_b200.h -- C ABI of libprometheus_b200.so
 *
 * B200 (sm_100a) implementation of the PROMETHEUS transit forward-model hot path.
 * The reference has no FFI layer: its seam is the Python method
 *     Transit.sumOverChords(self, max_memory_gb=2.0) -> ndarray (n_orb, n_wav)
 * (reference: pythonScripts/gasProperties.py:1160-1258) and the methods it calls.  Each entry
 * point below names the reference method it stands in for; the ctypes binding a maintainer
 * adds on the reference side is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain C: pointers + sizes, float64 everywhere (the reference computes in float64).
 *   - every function returns PB200_OK (0) or a negative pb200_status; the text of the last
 *     failure is available from pb200_last_error().
 *   - host pointers are caller-owned and only read during the call; device memory is owned
 *     by the context.  One context per GPU; a context is not thread-safe.
 *   - there is no CPU fallback: without a usable CUDA device pb200_create fails.
 *   - chord order is the reference's: flat chord index c = (i_phi*n_rho + i_rho)*n_orb + i_orb
 *     (reference: pythonScripts/geometryHandler.py:188-208, meshgrid 'ij' + flatten).
 */
#ifndef PROMETHEUS_B200_H
#define PROMETHEUS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB200_ABI_VERSION 4

typedef enum {
    PB200_OK = 0,
    PB200_ERR_INVALID = -1,      /* bad argument / inconsistent scene                    */
    PB200_ERR_CUDA = -2,         /* CUDA runtime error (text in pb200_last_error)        */
    PB200_ERR_NO_DEVICE = -3,    /* no CUDA device: there is no CPU fallback             */
    PB200_ERR_UNSUPPORTED = -4,  /* scenario / constituent kind not implemented          */
    PB200_ERR_NO_SCENE = -5,     /* pb200_run & co. called before pb200_set_scene        */
    PB200_ERR_BOUNDS = -6        /* tabulated density queried outside its grid (the
                                    reference raises ValueError there, gasProperties.py:600) */
} pb200_status;

/* density scenarios -- reference classes in pythonScripts/gasProperties.py */
typedef enum {
    PB200_SCN_BAROMETRIC = 0,    /* BarometricAtmosphere   :122-161  p = {n0, R, H}          */
    PB200_SCN_HYDROSTATIC = 1,   /* HydrostaticAtmosphere  :164-204  p = {n0, R, G*mu*M, kB*T} */
    PB200_SCN_POWERLAW = 2,      /* PowerLawAtmosphere :207-244, PowerLawExosphere :291-330,
                                    MoonExosphere :333-374, TidallyHeatedMoon :377-461
                                                                      p = {n0, R, q}          */
    PB200_SCN_TORUS = 3,         /* TorusExosphere         :464-516  p = {n0, a_torus, H_torus} */
    PB200_SCN_SERPENS = 4        /* SerpensExosphere       :519-601  trilinear table          */
} pb200_scenario_type;

typedef struct {
    int32_t type;                /* pb200_scenario_type                                      */
    int32_t n_constituents;      /* constituents [first, first+n) belong to this scenario    */
    int32_t first_constituent;
    int32_t reserved;
    double T;                    /* temperature for P = n*kB*T (molecular lookups), :927     */
    double p[8];                 /* scenario parameters, see pb200_scenario_type             */
    /* per-phase arrays, length n_orb, evaluated by the host code with the reference's own
       expressions (celestialBodies.py:370-397, :499-550; constants.py:31-45)                */
    const double* centre_x;      /* x of the density centre (planet or moon)                 */
    const double* centre_y;      /* y of the density centre                                  */
    const double* shift;         /* Doppler factor of the absorber frame, :916               */
    const double* amplitude;     /* NULL or per-phase multiplier of n0 (TidallyHeatedMoon)   */
    /* serpens only: trilinear table on (gx, gy, gz), values C-ordered [ix][iy][iz], :580    */
    const double* gx; const double* gy; const double* gz; const double* values;
    int32_t dims[3];
    int32_t reserved2;
} pb200_scenario;

typedef struct {
    int32_t kind;                /* 0 = atom/ion (AtmosphericConstituent :609-735),
                                    1 = molecule (MolecularConstituent :738-818)             */
    int32_t reserved;
    double chi;                  /* mixing ratio (1 for evaporative scenarios, :274)         */
    double offset;               /* lookupOffset, 1e-50 (:638, :761)                         */
    /* atoms: log10(sigma+offset) on a sorted wavelength table (interp1d .x/.y, :713)        */
    const double* tab_x; const double* tab_y; int64_t n_tab;
    /* molecules: log10(sigma+offset) on (P, T, lambda), C-ordered [iP][iT][iw] (:779)       */
    const double* grid_P; const double* grid_T; const double* grid_w; const double* values;
    int32_t dims[3];
    int32_t reserved2;
    double fill;                 /* RegularGridInterpolator fill_value = log10(offset)       */
} pb200_constituent;

typedef struct {
    int32_t abi_version;         /* PB200_ABI_VERSION                                        */
    int32_t n_x, n_rho, n_phi, n_orb;
    int32_t n_scenarios, n_constituents, n_moons;
    int64_t n_wav;
    /* axes as produced by the host code (Grid.construct*Axis, geometryHandler.py:112-186;
       WavelengthGrid.constructWavelengthGrid, gasProperties.py:1050-1071)                  */
    const double* x;             /* [n_x]   line-of-sight midpoints                          */
    const double* rho;           /* [n_rho]                                                  */
    const double* sin_phi;       /* [n_phi] sin / cos of the phi midpoints                   */
    const double* cos_phi;       /* [n_phi]                                                  */
    const double* cos_phi_rot;   /* [n_phi] cos(phi - phiStarrot), gasProperties.py:1182     */
    const double* wavelength;    /* [n_wav] cm, ascending                                    */
    double delta_x;              /* Grid.getDeltaX, geometryHandler.py:88-94                 */
    /* star (celestialBodies.py:59-74) */
    double star_R, clv_u1, clv_u2, vsini;
    const double* fstar_x; const double* fstar_y; int64_t n_fstar;   /* NULL/0: F* == 1     */
    /* opaque bodies (gasProperties.py:1224-1230) */
    double planet_R;
    const double* planet_y;      /* [n_orb] a*sin(orbphase)                                  */
    const double* moon_R;        /* [n_moons]                                                */
    const double* moon_y;        /* [n_moons*n_orb]                                          */
    const pb200_scenario* scenarios;
    const pb200_constituent* constituents;
} pb200_scene;

typedef struct {
    double upload_ms;            /* H2D of the scene (pb200_set_scene)                       */
    double column_ms;            /* kernel (a'): density + column density                    */
    double transit_ms;           /* kernel (c'): tau + exp(-tau) + disk sum                  */
    double finalize_ms;          /* split reduction + division                               */
    double download_ms;          /* D2H of R                                                 */
    int64_t launches;            /* kernels launched by the call                             */
    int64_t samples;             /* chord*lambda samples of the call (nominal)               */
    int64_t executed_samples;    /* samples kernel (c') evaluated: transparent chords (tau <
                                    2^-54 at every wavelength, exp(-tau) == 1.0) and blocked
                                    chords are summed without an exp; -1 = not yet known     */
    int64_t exp_samples;         /* of those, samples that went through the table exp (the
                                    others were optically thin for their warp: tau < 2.1e-5,
                                    second-order expansion summed per chord); upper bound.
                                    Narrow sigma-sorted windows (relative half spread of sigma
                                    <= 1/40) count ONE per (chord, window) -- or per (chord,
                                    32-wavelength group): one table exp feeds the nine Taylor
                                    sums about the centre.
                                    Scenes with molecules (kernel (m), moment form): executed_samples =
                                    unblocked chord-pair blocks x wavelengths, exp_samples = moment
                                    items x wavelengths (one table exp + 8 FMAs each)         */
    int64_t centre_pairs;        /* (chord, sigma cell) or (chord, narrow window / group) pairs that
                                    fed Taylor sums about a centre (one table exp + 18 FP64
                                    instructions each); they are also counted in exp_samples   */
} pb200_timing;

/* ---- retrieval batches ------------------------------------------------
 * The consumer sketched by mainRetrieval.py:18-93: many forward models of ONE setup that differ only in
 * free parameters (mainRetrieval.py:18-21: moon phase, orbit, radius; equally N, q, T, R ...).  Axes, the
 * wavelength grid and every table (sigma, F*, molecules, serpens) of the resident scene stay in HBM; a
 * model carries what a free parameter can move: the scenario scalars and the per-phase arrays the host
 * code evaluates with the reference's own expressions.  The models of a batch are stacked along the
 * orbital-phase axis ("virtual phases", model-major: row m*n_orb + p), so that ONE set of launches does
 * the whole batch. */
typedef struct {
    const pb200_scenario* scenarios; /* [n_scenarios of the resident scene], same types and constituent
                                        ranges; read: T (P = n kB T and the temperature axis of molecular
                                        tables), p[8], centre_x, centre_y, shift, amplitude ([n_orb] each;
                                        amplitude NULL = 1)                                          */
    const double* planet_y;      /* [n_orb]                                                          */
    const double* moon_R;        /* [n_moons]                                                        */
    const double* moon_y;        /* [n_moons*n_orb]                                                  */
    double planet_R;             /* radius of the opaque planet disk of this model (blocking mask,
                                    gasProperties.py:1224); <= 0: the resident scene's               */
} pb200_model;

/* band-averaged light curve (mainRetrieval.py:72-93): for phase p the filter selects the wavelengths with
   centre_l*shift[p] - bandwidth/2 <= lambda <= centre_l*shift[p] + bandwidth/2 for any line l;
   lightcurve[p] = mean(R[p, selected]) / max(R[p, :])   (nan if nothing is selected, as np.mean)        */
typedef struct {
    int32_t n_lines;             /* <= 16                                                             */
    int32_t reserved;
    const double* line_centre;   /* [n_lines] rest wavelengths, cm                                    */
    const double* shift;         /* [n_orb] Doppler factor of the band centres (planet LOS velocity,
                                    mainRetrieval.py:77-78); the same for every model of a batch      */
    double bandwidth;            /* filterBandwidth, cm                                               */
} pb200_band;

typedef struct pb200_ctx pb200_ctx;

/* ---- run-time options ---------------------------------------------------------------------
 * Every default is the production path.  The switches exist for measurement (the brute-force launch the
 * roofline is quoted on) and for regression tests (each reduction of the work must leave R unchanged); they are
 * per context and the library never reads the process environment.  MIRROR and GROUP are read by
 * pb200_set_scene (they shape the resident scene); the others by the next run.                          */
typedef enum {
    PB200_OPT_CULL = 0,          /* 1: chords with tau < 2^-54 at every wavelength are summed without a record   */
    PB200_OPT_MIRROR = 1,        /* 1: chords (phi, rho) and (pi - phi, rho) share one record (set_scene)         */
    PB200_OPT_THIN = 2,          /* 1: optically thin (chord, window) pairs as a per-chord expansion              */
    PB200_OPT_GROUP = 3,         /* 1: one absorber per scenario, 0: one per atomic constituent (set_scene)       */
    PB200_OPT_SORT = 4,          /* -1: automatic, 0 / 1: sigma-sorted windows of kernel (c') forced off / on     */
    PB200_OPT_MOL_RECORDS = 5,   /* 1: per-record form of kernel (m) instead of the moment form                   */
    PB200_OPT_ZEROCOPY = 6,      /* 1: pb200_run stores R straight into a page-locked result                      */
    PB200_OPT_MIN_TILES = 7,     /* least number of active chord tiles per working CTA of kernel (c') (>= 1)      */
    PB200_OPT_CENTRE = 8,        /* one absorber, where the sigma-sorted windows apply: 3 (default) no sort at all -- R(p, .)
                                    tabulated per phase on a geometric grid of sigma cells (Taylor sums about each cell's
                                    centre), one polynomial per wavelength; 0..2 sorted windows: 0 without, 1 with Taylor
                                    sums about the centre of narrow windows (256 wavelengths), 2 also per 32-wavelength group */
    PB200_OPT_COUNT = 9
} pb200_option;
int pb200_set_option(pb200_ctx* ctx, int option, int64_t value);
int pb200_get_option(pb200_ctx* ctx, int option, int64_t* value);
/* compiled limits: PB200_ERR_INVALID / PB200_ERR_UNSUPPORTED beyond them */
#define PB200_MAX_SCENARIOS 8
#define PB200_MAX_CONSTITUENTS 32
#define PB200_MAX_MOONS 8
#define PB200_MAX_ABSORBERS 12   /* absorbers of the atomic run path = scenarios with atomic constituents (every
                                    atomic constituent with PB200_OPT_GROUP = 0); kernel (c') is compiled for 1..12 */

/* ---- life cycle ------------------------------------------------------------------------ */
int pb200_abi_version(void);
int pb200_device_count(void);
int pb200_create(int device, pb200_ctx** out);
void pb200_destroy(pb200_ctx* ctx);
const char* pb200_last_error(pb200_ctx* ctx);          /* ctx may be NULL (create errors)   */
/* run all work of this context on an existing cudaStream_t (e.g. torch's current stream). */
int pb200_set_stream(pb200_ctx* ctx, void* cuda_stream);
int64_t pb200_launch_count(pb200_ctx* ctx);            /* kernels launched since create     */
/* page-locked host memory for result arrays: pb200_run writes R at DMA speed into such a buffer,
   and the caller avoids first-touch page faults on a fresh (n_orb, n_wav) array every call
   (the reference allocates F_in_sum/F_out_sum afresh per call, gasProperties.py:1202-1203).      */
int pb200_host_alloc(size_t bytes, void** out);
int pb200_host_free(void* p);
/* page-lock and map memory the caller owns (e.g. a POSIX shared-memory segment that several ranks attach to);
   *dev_ptr receives the address the device uses for it.  Such memory, like that of pb200_host_alloc, can be the
   target of pb200_run (written without a staging copy) and of pb200_run_scatter.                         */
int pb200_host_register(void* p, size_t bytes, void** dev_ptr);
int pb200_host_unregister(void* p);

/* ---- scene ------------------------------------------------------------------------------ */
/* Uploads everything Transit.sumOverChords reads through `self`.             */
int pb200_set_scene(pb200_ctx* ctx, const pb200_scene* scene);
/* The same for rank `rank` of n_ranks processes that upload the SAME scene to the GPUs of one node: of every
   large atomic sigma table only this rank's n_ranks-th crosses PCIe.  Before the first run the ranks complete
   the tables with one all-gather per region over NVLink: region i is n_ranks*chunk_doubles[i] doubles at device
   address base[i], rank r owns [r*chunk, (r+1)*chunk) (ncclAllGather in place; prometheus_b200/distributed.py).
   The reference re-reads its tables from host RAM per chunk (gasProperties.py:948-952).                      */
int pb200_set_scene_sharded(pb200_ctx* ctx, const pb200_scene* scene, int n_ranks, int rank);

/* Optional, before flattening the next scene: re-issue the host-to-device copies of the large page-locked tables of the
   LAST pb200_set_scene at the places they had, so that the transfer runs while the host prepares the small arrays.
   The next pb200_set_scene skips the copies that are already on the stream (same source pointer, size and place) and
   makes every other one as usual.  The tables must not be written between this call and that pb200_set_scene.
   No counterpart in the reference.                                                                                 */
int pb200_prefetch_tables(pb200_ctx* ctx);
int pb200_table_regions(pb200_ctx* ctx, int max_regions, void** base, int64_t* chunk_doubles, int* n_regions);

/* ---- Transit.sumOverChords (gasProperties.py:1160-1258) ---------------------------------
 * R[(p-phase_lo)*(wav_hi-wav_lo) + (w-wav_lo)] for p in [phase_lo,phase_hi), w in [wav_lo,wav_hi).
 * pb200_run writes to host memory; pb200_run_device leaves the tile in device memory
 * (R_dev is a device pointer, e.g. a torch tensor) and does not synchronise.               */
int pb200_run(pb200_ctx* ctx, int phase_lo, int phase_hi, int64_t wav_lo, int64_t wav_hi,
              double* R_host, pb200_timing* timing);
int pb200_run_device(pb200_ctx* ctx, int phase_lo, int phase_hi, int64_t wav_lo, int64_t wav_hi,
                     double* R_dev, pb200_timing* timing);
/* Multi-GPU tile with balanced work: the wavelength axis is cut into blocks of wav_block (a multiple of
 * 256) and rank r of n_ranks takes blocks r, r + n_ranks, ... (line cores and continuum are then spread
 * evenly over the ranks, unlike with contiguous chunks).  R_dev receives [n_phase][n_local] in the order
 * of the rank's wavelengths, n_local = pb200_cyclic_count(n_wav, wav_block, n_ranks, rank).  Atomic scenes
 * only (PB200_ERR_UNSUPPORTED with molecules / a per-chord stellar spectrum).  Replaces, together with the
 * gather in prometheus_b200/distributed.py, memoryHandler.calculate_optimal_chunk_size (memoryHandler.py:55-66). */
int pb200_run_device_cyclic(pb200_ctx* ctx, int phase_lo, int phase_hi, int64_t wav_block, int n_ranks,
                            int rank, double* R_dev, pb200_timing* timing);
int64_t pb200_cyclic_count(int64_t n_wav, int64_t wav_block, int n_ranks, int rank);
/* The same block-cyclic tile written IN PLACE into the full result: R_full is row phase_lo of an array with
 * row_stride (>= n_wav) doubles per phase that the device can address -- mapped page-locked host memory
 * (pb200_host_alloc / pb200_host_register; the device-to-host transfer is then the kernel's own stores, each
 * rank over its own PCIe link, no gather) or a peer GPU's buffer (stores over NVLink).  n_ranks == 1: the whole
 * wavelength axis (any scene); n_ranks > 1: atomic scenes.  Not synchronised unless timing != NULL.        */
int pb200_run_scatter(pb200_ctx* ctx, int phase_lo, int phase_hi, int64_t wav_block, int n_ranks, int rank,
                      double* R_full, int64_t row_stride, pb200_timing* timing);
/* ---- retrieval batches -------------------------------------------------------------------
 * pb200_set_models: stacks n_models models on the resident scene (pb200_set_scene must have been called;
 * its own per-phase arrays are model-independent inputs such as the grid and are replaced).  Afterwards
 * the context has n_models*n_orb (virtual) phases: pb200_run*, pb200_column_density ... address row
 * m*n_orb + p.  pb200_set_models(ctx, 0, NULL) returns to the scene as uploaded.
 * pb200_lightcurve: light curve of rows [phase_lo, phase_hi) of R (device pointer to
 * [phase_hi-phase_lo][n_wav], or NULL = the R of the last pb200_run / pb200_run_batch of this context).
 * pb200_run_batch = pb200_set_models + one run over all rows and wavelengths + pb200_lightcurve;
 * lightcurve_host[n_models*n_orb]; R_host NULL or [n_models*n_orb*n_wav].                             */
int pb200_set_models(pb200_ctx* ctx, int n_models, const pb200_model* models);
int pb200_lightcurve(pb200_ctx* ctx, const pb200_band* band, int phase_lo, int phase_hi,
                     const double* R_dev, double* lightcurve_host);
int pb200_run_batch(pb200_ctx* ctx, int n_models, const pb200_model* models, const pb200_band* band,
                    double* lightcurve_host, double* R_host, pb200_timing* timing);
/* waits for everything enqueued on the context's stream (pb200_run_device*, pb200_run_scatter do not) and returns
   PB200_ERR_BOUNDS if a tabulated density was queried outside its grid since the last check            */
int pb200_sync(pb200_ctx* ctx);
/* Diagnostics (tests pick their oracle columns with it): the order in which the last run walked its wavelengths when
   it used sigma-sorted windows -- perm[pi*n_wav_sel + i] = local wavelength index at sorted position i of phase pi
   (a window of kernel (c') is 256 consecutive positions), nsig_sorted = -sigma(p, lambda) * 4096/ln2 in that order.
   In: *n = capacity of the arrays (either may be NULL); out: *n = n_phase*n_wav_sel of the run, 0 if not sorted.
   Valid until the next run of the context.  No counterpart in the reference.                                  */
int pb200_last_sort_order(pb200_ctx* ctx, int64_t* n, int32_t* perm_host, double* nsig_sorted_host);
/* timing record of the last pb200_run / pb200_voigt_sigma call of this context */
int pb200_last_timing(pb200_ctx* ctx, pb200_timing* timing);

/* ---- inner seams, same arithmetic as the fused path, materialised ------------------------ */
/* <Scenario>.calculateNumberDensity on the whole chord grid -> out[n_c*n_x] (chord-major)   */
int pb200_number_density(pb200_ctx* ctx, int scenario, double* n_host);
/* the same into device memory (n_dev[n_c*n_x]), not synchronised: kernel (a) on its own, the form the
   north star calls bandwidth-bound                                                              */
int pb200_number_density_device(pb200_ctx* ctx, int scenario, double* n_dev);
/* same for caller-supplied chords: y,z = rho*sin(phi), rho*cos(phi); cx,cy = density centre;
   amp = NULL or per-chord multiplier.  out[n*n_x]                                           */
int pb200_number_density_chords(pb200_ctx* ctx, int scenario, int64_t n, const double* y,
                                const double* z, const double* cx, const double* cy,
                                const double* amp, double* n_host);
/* column density sum_x(n*chi)*delta_x of every atomic constituent (gasProperties.py:940),
   out[n_atoms*n_c] in constituent order, chord order as the reference                       */
int pb200_column_density(pb200_ctx* ctx, double* ncol_host);
/* blocked mask (gasProperties.py:1224-1230) out[n_c] (1 = blocked) and
   flux weights rho*clv (gasProperties.py:1178-1180, :1232) out[n_c]                         */
int pb200_chord_masks(pb200_ctx* ctx, uint8_t* blocked_host, double* weight_host);
/* Atmosphere.getLOSopticalDepth_Batch on the whole chord grid (gasProperties.py:885-956):
   tau[n_c*(wav_hi-wav_lo)], blocked chords included (the reference evaluates only active
   chords; their rows are simply not used).  Testing / bandwidth probe, memory-hungry.       */
int pb200_optical_depth(pb200_ctx* ctx, int64_t wav_lo, int64_t wav_hi, double* tau_host);

/* The same method for CALLER-SUPPLIED chords, as the reference's signature has it
   (getLOSopticalDepth_Batch(x_grid, phi_batch, rho_batch, orbphase_batch, wavelength, delta_x), gasProperties.py:885):
   the host evaluates what depends on a chord's orbital phase with the objects' own methods and passes it per chord --
   y, z = rho*sin(phi), rho*cos(phi) [n]; per scenario the density centre cx, cy, the Doppler factor shift and, or
   NULL, the amplitude: [n_scenarios][n] each.  x axis, delta_x, wavelengths and tables are the resident scene's.
   tau_host[n*(wav_hi-wav_lo)].                                                                          */
int pb200_optical_depth_chords(pb200_ctx* ctx, int64_t n, const double* y, const double* z, const double* cx,
                               const double* cy, const double* shift, const double* amp, int64_t wav_lo,
                               int64_t wav_hi, double* tau_host);

/* ---- cross sections --------------------------------------------------------------------- */
/* n_interp_log (gasProperties.py:34-51) == AtmosphericConstituent.getSigmaAbs (:727-735):
   out[i] = 10**interp(wl[i]; tab_x, tab_y) - offset, np.interp end clamping                 */
int pb200_sigma_lookup(pb200_ctx* ctx, const double* tab_x, const double* tab_y, int64_t n_tab,
                       double offset, const double* wl, int64_t n, double* out_host);
/* AtmosphericConstituent.calculateVoigtProfile (:672-692): sigma(wl) = sum over lines of
   pi e^2/(m_e c) f V(c/wl - c/wl0; sigma_v/wl0, gamma);  if out_log10 != NULL it also gets
   log10(sigma + offset), the table constructLookupFunction stores (:713-714)                */
int pb200_voigt_sigma(pb200_ctx* ctx, const double* line_wl, const double* line_gamma,
                      const double* line_f, int64_t n_lines, double sigma_v, const double* wl,
                      int64_t n, double offset, double* out_sigma, double* out_log10);
/* scipy.special.voigt_profile(x, sigma, gamma) element-wise (validation of the Faddeeva code) */
int pb200_voigt_profile(pb200_ctx* ctx, const double* x, const double* sigma,
                        const double* gamma, int64_t n, double* out_host);
/* MolecularConstituent.getSigmaAbs (:789-818): P[n_c*n_x] (clipped at 1e-4 inside), scalar T,
   wl[n_c*n_wav] -> out[n_c*n_x*n_wav]                                                       */
int pb200_molecular_sigma(pb200_ctx* ctx, const double* grid_P, const double* grid_T,
                          const double* grid_w, const double* values, const int32_t dims[3],
                          double fill, double offset, const double* P, double T, const double* wl,
                          int64_t n_c, int64_t n_x, int64_t n_wav, double* out_host);

/* Elementwise part of the table producers: out[r][c] = log10(values[r][reverse_cols ? n_cols-1-c : c] + offset).
   MolecularConstituent.constructLookupFunction (gasProperties.py:773-781: sigma_abs = xsecarr[:, :, ::-1],
   log10(sigma_abs + lookupOffset); n_rows = n_P*n_T, n_cols = n_lambda) and Star.addFstarFunction
   (celestialBodies.py:263: log10(F_0); n_rows = 1, offset 0).  Reading the HDF5 / FITS files stays host code. */
int pb200_log_table(pb200_ctx* ctx, const double* values, int64_t n_rows, int64_t n_cols, int reverse_cols,
                    double offset, double* out_host);

/* ---- tabulated-density producer ----------------------------------------------------------- */
/* SerpensExosphere.addInterpolatedDensity (gasProperties.py:548-583), the part after np.loadtxt:
   out[nx][ny][nz] = histogramdd(pos[n][3]; edges xb, yb, zb) * N / denom   (denom = n * cellVolume, formed by
   the caller as the reference does), then, if radius > 0, scipy.ndimage.gaussian_filter with the caller's
   symmetric weights[2*radius+1] (scipy's _gaussian_kernel1d; boundary mode 'reflect') along x, y, z.
   Binning is bit-exact with numpy (right-open bins, last edge inclusive, outliers dropped).
   sums_host (or NULL) receives the two sums the reference prints: all cells, and the cells with
   planet_R^2 < y^2 + z^2 < star_R^2 on the cell midpoints y_points[ny], z_points[nz].              */
int pb200_serpens_density(pb200_ctx* ctx, const double* pos, int64_t n, const double* xb, int32_t nxb,
                          const double* yb, int32_t nyb, const double* zb, int32_t nzb, double N,
                          double denom, const double* weights, int32_t radius, const double* y_points,
                          const double* z_points, double planet_R, double star_R, double* out_host,
                          double* sums_host);

/* Allocation-free variant of the particle histogram of pb200_serpens_density (no smoothing):
   out_dev[nx][ny][nz] = histogramdd(pos_dev[n][3]; edges) * N / denom, all pointers DEVICE pointers owned by the
   caller, enqueued on the context's stream, not synchronised.  NO COUNTERPART IN THE REFERENCE: added
   for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that
   forbid allocation).  The integer count grid -- the kernel's scratch -- is a static device array of
   PB200_SERPENS_STATIC_MAX_CELLS cells (8 MiB, held by every GPU that loads the library until exit); the call does
   no cudaMalloc / cudaFree and fails with PB200_ERR_INVALID when (nxb-1)*(nyb-1)*(nzb-1) exceeds it.  Every call on
   a device shares that array (it is cleared at the start of each call): calls that overlap on the GPU -- two
   contexts or two streams of one device -- race; serialise them.  The first call on a device looks the array's
   address up (cudaGetSymbolAddress): make it before a stream capture begins.                               */
#define PB200_SERPENS_STATIC_MAX_CELLS (1 << 21)
int pb200_serpens_histogram_static(pb200_ctx* ctx, const double* pos_dev, int64_t n, const double* xb_dev, int32_t nxb,
                                   const double* yb_dev, int32_t nyb, const double* zb_dev, int32_t nzb, double N,
                                   double denom, double* out_dev);

/* ---- result file ------------------------------------------------------------------------- */
/* Byte-identical replacement of the reference's writer (prometheus.py:142-156): first row
   "nan orbphase/(2 pi) ...", then one row per wavelength "lambda R(phase_0, lambda) ...", np.savetxt
   defaults (fmt %.18e, ' ' delimiter, header lines prefixed "# ").  R is [n_orb][n_wav] as returned by
   pb200_run.  Host code, n_threads <= 0: all hardware threads.                                   */
int pb200_write_output(const char* path, const char* header, const double* wavelength,
                       const double* orbphase, const double* R, int n_orb, int64_t n_wav, int n_threads);

/* ---- machine probes (roofline denominators that MEASURED_PEAKS.json does not hold) ------- */
/* DFMA-chain FP64 peak [TFLOP/s] and fp64 exp throughput [G exp/s] on this context's GPU    */
int pb200_probe_fp64(pb200_ctx* ctx, double* dfma_tflops, double* exp_gops);

#ifdef __cplusplus
}
#endif
#endif /* PROMETHEUS_B200_H */
