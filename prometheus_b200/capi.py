"""ctypes binding of libprometheus_b200.so (include/prometheus_b200.h).

This is the only door from Python to the compute path.  If the shared library is missing
or no CUDA device is usable the calls raise: there is no CPU implementation to fall back on.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libprometheus_b200.so")
ABI_VERSION = 4

c_double_p = C.POINTER(C.c_double)

# every symbol include/prometheus_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "pb200_abi_version", "pb200_device_count", "pb200_create", "pb200_destroy", "pb200_last_error",
    "pb200_set_stream", "pb200_launch_count", "pb200_host_alloc", "pb200_host_free", "pb200_set_scene", "pb200_run", "pb200_run_device", "pb200_run_device_cyclic", "pb200_cyclic_count",
    "pb200_last_timing", "pb200_number_density", "pb200_number_density_chords", "pb200_column_density",
    "pb200_chord_masks", "pb200_optical_depth", "pb200_sigma_lookup", "pb200_voigt_sigma",
    "pb200_voigt_profile", "pb200_molecular_sigma", "pb200_write_output", "pb200_probe_fp64",
    "pb200_set_models", "pb200_lightcurve", "pb200_run_batch",
    "pb200_host_register", "pb200_host_unregister", "pb200_run_scatter", "pb200_sync",
    "pb200_set_scene_sharded", "pb200_table_regions", "pb200_number_density_device", "pb200_prefetch_tables",
    "pb200_serpens_density", "pb200_set_option", "pb200_get_option", "pb200_serpens_histogram_static",
    "pb200_last_sort_order", "pb200_optical_depth_chords", "pb200_log_table",
]
SERPENS_STATIC_MAX_CELLS = 1 << 21

# pb200_option (include/prometheus_b200.h)
OPT = {"cull": 0, "mirror": 1, "thin": 2, "group": 3, "sort": 4, "mol_records": 5, "zerocopy": 6, "min_tiles": 7, "centre": 8}
OPT_DEFAULTS = {"cull": 1, "mirror": 1, "thin": 1, "group": 1, "sort": -1, "mol_records": 0, "zerocopy": 1, "min_tiles": 1, "centre": 3}

SCN_BAROMETRIC, SCN_HYDROSTATIC, SCN_POWERLAW, SCN_TORUS, SCN_SERPENS = range(5)

STATUS = {0: "PB200_OK", -1: "PB200_ERR_INVALID", -2: "PB200_ERR_CUDA", -3: "PB200_ERR_NO_DEVICE",
          -4: "PB200_ERR_UNSUPPORTED", -5: "PB200_ERR_NO_SCENE", -6: "PB200_ERR_BOUNDS"}


class Scenario(C.Structure):
    _fields_ = [
        ("type", C.c_int32), ("n_constituents", C.c_int32), ("first_constituent", C.c_int32),
        ("reserved", C.c_int32), ("T", C.c_double), ("p", C.c_double * 8),
        ("centre_x", c_double_p), ("centre_y", c_double_p), ("shift", c_double_p), ("amplitude", c_double_p),
        ("gx", c_double_p), ("gy", c_double_p), ("gz", c_double_p), ("values", c_double_p),
        ("dims", C.c_int32 * 3), ("reserved2", C.c_int32),
    ]


class Constituent(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("reserved", C.c_int32), ("chi", C.c_double), ("offset", C.c_double),
        ("tab_x", c_double_p), ("tab_y", c_double_p), ("n_tab", C.c_int64),
        ("grid_P", c_double_p), ("grid_T", c_double_p), ("grid_w", c_double_p), ("values", c_double_p),
        ("dims", C.c_int32 * 3), ("reserved2", C.c_int32), ("fill", C.c_double),
    ]


class Scene(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("n_x", C.c_int32), ("n_rho", C.c_int32), ("n_phi", C.c_int32), ("n_orb", C.c_int32),
        ("n_scenarios", C.c_int32), ("n_constituents", C.c_int32), ("n_moons", C.c_int32),
        ("n_wav", C.c_int64),
        ("x", c_double_p), ("rho", c_double_p), ("sin_phi", c_double_p), ("cos_phi", c_double_p),
        ("cos_phi_rot", c_double_p), ("wavelength", c_double_p),
        ("delta_x", C.c_double),
        ("star_R", C.c_double), ("clv_u1", C.c_double), ("clv_u2", C.c_double), ("vsini", C.c_double),
        ("fstar_x", c_double_p), ("fstar_y", c_double_p), ("n_fstar", C.c_int64),
        ("planet_R", C.c_double),
        ("planet_y", c_double_p), ("moon_R", c_double_p), ("moon_y", c_double_p),
        ("scenarios", C.POINTER(Scenario)), ("constituents", C.POINTER(Constituent)),
    ]


class Model(C.Structure):
    """pb200_model: what a free parameter of a retrieval can move (mainRetrieval.py:18-21)."""
    _fields_ = [("scenarios", C.POINTER(Scenario)), ("planet_y", c_double_p), ("moon_R", c_double_p),
                ("moon_y", c_double_p), ("planet_R", C.c_double)]


class Band(C.Structure):
    """pb200_band: the light-curve filter of mainRetrieval.py:72-93."""
    _fields_ = [("n_lines", C.c_int32), ("reserved", C.c_int32), ("line_centre", c_double_p),
                ("shift", c_double_p), ("bandwidth", C.c_double)]


def struct_dtype(struct) -> np.dtype:
    """numpy structured dtype with the exact layout of a ctypes Structure (pointers as uint64): arrays of
    pb200_model / pb200_scenario for a whole retrieval batch are then filled with vectorised assignments
    instead of one ctypes attribute store per field and model."""
    names, formats, offsets = [], [], []
    for name, ct in struct._fields_:
        if isinstance(ct, type) and issubclass(ct, C.Array):
            fmt = (np.dtype(ct._type_), ct._length_)
        elif isinstance(ct, type) and issubclass(ct, C._Pointer):
            fmt = np.dtype(np.uint64)
        else:
            fmt = np.dtype(ct)
        names.append(name); formats.append(fmt); offsets.append(getattr(struct, name).offset)
    return np.dtype({"names": names, "formats": formats, "offsets": offsets, "itemsize": C.sizeof(struct)})


class Timing(C.Structure):
    _fields_ = [("upload_ms", C.c_double), ("column_ms", C.c_double), ("transit_ms", C.c_double),
                ("finalize_ms", C.c_double), ("download_ms", C.c_double),
                ("launches", C.c_int64), ("samples", C.c_int64), ("executed_samples", C.c_int64),
                ("exp_samples", C.c_int64), ("centre_pairs", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Pb200Error(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"{STATUS.get(code, code)}: {text}")
        self.code = code


_lib = None


def library_path() -> str:
    return LIB_PATH


def load():
    """dlopen the shared library and declare the prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m prometheus_b200.build` "
            "(needs nvcc; there is no CPU fallback for the transit hot path)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_double
    sig = {
        "pb200_abi_version": (C.c_int, []),
        "pb200_device_count": (C.c_int, []),
        "pb200_create": (C.c_int, [i32, C.POINTER(vp)]),
        "pb200_destroy": (None, [vp]),
        "pb200_last_error": (C.c_char_p, [vp]),
        "pb200_set_stream": (C.c_int, [vp, vp]),
        "pb200_launch_count": (i64, [vp]),
        "pb200_set_option": (C.c_int, [vp, i32, i64]),
        "pb200_get_option": (C.c_int, [vp, i32, C.POINTER(i64)]),
        "pb200_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(vp)]),
        "pb200_host_free": (C.c_int, [vp]),
        "pb200_set_scene": (C.c_int, [vp, C.POINTER(Scene)]),
        "pb200_run": (C.c_int, [vp, i32, i32, i64, i64, c_double_p, C.POINTER(Timing)]),
        "pb200_run_device": (C.c_int, [vp, i32, i32, i64, i64, vp, C.POINTER(Timing)]),
        "pb200_run_device_cyclic": (C.c_int, [vp, i32, i32, i64, i32, i32, vp, C.POINTER(Timing)]),
        "pb200_cyclic_count": (C.c_int64, [i64, i64, i32, i32]),
        "pb200_last_timing": (C.c_int, [vp, C.POINTER(Timing)]),
        "pb200_number_density": (C.c_int, [vp, i32, c_double_p]),
        "pb200_number_density_device": (C.c_int, [vp, i32, vp]),
        "pb200_number_density_chords": (C.c_int, [vp, i32, i64] + [c_double_p] * 6),
        "pb200_column_density": (C.c_int, [vp, c_double_p]),
        "pb200_chord_masks": (C.c_int, [vp, C.POINTER(C.c_uint8), c_double_p]),
        "pb200_optical_depth": (C.c_int, [vp, i64, i64, c_double_p]),
        "pb200_optical_depth_chords": (C.c_int, [vp, i64] + [c_double_p] * 6 + [i64, i64, c_double_p]),
        "pb200_log_table": (C.c_int, [vp, c_double_p, i64, i64, i32, dbl, c_double_p]),
        "pb200_sigma_lookup": (C.c_int, [vp, c_double_p, c_double_p, i64, dbl, c_double_p, i64, c_double_p]),
        "pb200_voigt_sigma": (C.c_int, [vp, c_double_p, c_double_p, c_double_p, i64, dbl, c_double_p, i64, dbl,
                                        c_double_p, c_double_p]),
        "pb200_voigt_profile": (C.c_int, [vp, c_double_p, c_double_p, c_double_p, i64, c_double_p]),
        "pb200_molecular_sigma": (C.c_int, [vp, c_double_p, c_double_p, c_double_p, c_double_p,
                                            C.POINTER(C.c_int32), dbl, dbl, c_double_p, dbl, c_double_p,
                                            i64, i64, i64, c_double_p]),
        "pb200_write_output": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, c_double_p, i32, i64, i32]),
        "pb200_probe_fp64": (C.c_int, [vp, c_double_p, c_double_p]),
        "pb200_sync": (C.c_int, [vp]),
        "pb200_serpens_density": (C.c_int, [vp, c_double_p, i64, c_double_p, i32, c_double_p, i32, c_double_p, i32, dbl, dbl,
                                            c_double_p, i32, c_double_p, c_double_p, dbl, dbl, c_double_p, c_double_p]),
        "pb200_serpens_histogram_static": (C.c_int, [vp, vp, i64, vp, i32, vp, i32, vp, i32, dbl, dbl, vp]),
        "pb200_last_sort_order": (C.c_int, [vp, C.POINTER(i64), C.POINTER(C.c_int32), c_double_p]),
        "pb200_set_scene_sharded": (C.c_int, [vp, C.POINTER(Scene), i32, i32]),
        "pb200_prefetch_tables": (C.c_int, [vp]),
        "pb200_table_regions": (C.c_int, [vp, i32, C.POINTER(vp), C.POINTER(i64), C.POINTER(i32)]),
        "pb200_host_register": (C.c_int, [vp, C.c_size_t, C.POINTER(vp)]),
        "pb200_host_unregister": (C.c_int, [vp]),
        "pb200_run_scatter": (C.c_int, [vp, i32, i32, i64, i32, i32, vp, i64, C.POINTER(Timing)]),
        "pb200_set_models": (C.c_int, [vp, i32, C.POINTER(Model)]),
        "pb200_lightcurve": (C.c_int, [vp, C.POINTER(Band), i32, i32, vp, c_double_p]),
        "pb200_run_batch": (C.c_int, [vp, i32, C.POINTER(Model), C.POINTER(Band), c_double_p, c_double_p,
                                      C.POINTER(Timing)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.pb200_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI {lib.pb200_abi_version()} != binding {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def as_f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def dptr(a: np.ndarray):
    return a.ctypes.data_as(c_double_p)


class _PinnedBlock:
    """A page-locked host block that goes back to the pool when the last numpy view of it dies."""

    def __init__(self, pool, ptr, nbytes):
        self.pool, self.ptr, self.nbytes = pool, ptr, nbytes
        self.__array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}

    def __del__(self):
        pool = self.pool
        if pool is not None:
            pool._give_back(self.ptr, self.nbytes)


class PinnedPool:
    """Recycles pinned result buffers: `R = transit.sumOverChords()` in a loop reuses two blocks
    instead of faulting in a fresh 13 MB array (8 ms of page faults at C4) on every call."""

    GRANULE = 1 << 21
    MAX_CACHED_BYTES = 1 << 30

    def __init__(self):
        self.free = {}          # nbytes -> [ptr]
        self.cached = 0

    def empty(self, shape, dtype=np.float64) -> np.ndarray:
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        nbytes = max(self.GRANULE, -(-n // self.GRANULE) * self.GRANULE)
        lst = self.free.get(nbytes)
        if lst:
            ptr = lst.pop()
            self.cached -= nbytes
        else:
            p = C.c_void_p()
            rc = load().pb200_host_alloc(nbytes, C.byref(p))
            if rc != 0:
                raise Pb200Error(rc, load().pb200_last_error(None).decode())
            ptr = p.value
        block = _PinnedBlock(self, ptr, nbytes)
        return np.asarray(block)[:n].view(dtype).reshape(shape)

    def _give_back(self, ptr, nbytes):
        try:
            if self.cached + nbytes <= self.MAX_CACHED_BYTES:
                self.free.setdefault(nbytes, []).append(ptr)
                self.cached += nbytes
            else:
                load().pb200_host_free(C.c_void_p(ptr))
        except Exception:
            pass


pinned_pool = PinnedPool()


class Context:
    """One pb200_ctx: one GPU, one stream.  Not thread-safe (mirrors the C contract)."""

    def __init__(self, device: int = 0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.pb200_create(int(device), C.byref(h))
        if rc != 0:
            raise Pb200Error(rc, self.lib.pb200_last_error(None).decode())
        self.handle = h
        self.device = device
        self._keep = None   # numpy arrays referenced by the last scene struct

    def check(self, rc: int):
        if rc != 0:
            raise Pb200Error(rc, self.lib.pb200_last_error(self.handle).decode())

    def close(self):
        if getattr(self, "handle", None):
            self.lib.pb200_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- thin wrappers ---------------------------------------------------------------------------
    def set_stream(self, cuda_stream: int):
        self.check(self.lib.pb200_set_stream(self.handle, C.c_void_p(cuda_stream)))

    def launch_count(self) -> int:
        return int(self.lib.pb200_launch_count(self.handle))

    def set_option(self, name: str, value: int):
        """pb200_set_option by name (capi.OPT).  MIRROR and GROUP take effect at the next set_scene."""
        self.check(self.lib.pb200_set_option(self.handle, OPT[name], int(value)))

    def get_option(self, name: str) -> int:
        v = C.c_int64()
        self.check(self.lib.pb200_get_option(self.handle, OPT[name], C.byref(v)))
        return int(v.value)

    def options(self, **kw):
        """Context manager: the given options for the duration of the block, the previous values afterwards."""
        ctx = self

        class _Scope:
            def __enter__(self):
                self.saved = {k: ctx.get_option(k) for k in kw}
                for k, v in kw.items():
                    ctx.set_option(k, v)
                return ctx

            def __exit__(self, *exc):
                for k, v in self.saved.items():
                    ctx.set_option(k, v)

        return _Scope()

    def prefetch_tables(self):
        """Start the H2D copies of the last scene's large page-locked tables again (pb200_prefetch_tables)."""
        self.check(self.lib.pb200_prefetch_tables(self.handle))

    def set_scene(self, scene: Scene, keep):
        self._keep = keep
        self.check(self.lib.pb200_set_scene(self.handle, C.byref(scene)))

    def set_scene_sharded(self, scene: Scene, keep, n_ranks: int, rank: int):
        """Upload this rank's share of the large tables; returns [(device pointer, doubles per rank)] -- the regions
        the ranks must all-gather (in place, rank r owns slice r) before running."""
        self._keep = keep
        self.check(self.lib.pb200_set_scene_sharded(self.handle, C.byref(scene), n_ranks, rank))
        n = C.c_int32(0)
        base = (C.c_void_p * 64)()
        chunk = (C.c_int64 * 64)()
        self.check(self.lib.pb200_table_regions(self.handle, 64, base, chunk, C.byref(n)))
        return [(int(base[i]), int(chunk[i])) for i in range(min(n.value, 64))]

    def run(self, phase_lo, phase_hi, wav_lo, wav_hi, out: np.ndarray | None = None):
        n = (phase_hi - phase_lo, wav_hi - wav_lo)
        if out is None:
            out = pinned_pool.empty(n)
        assert out.shape == n and out.dtype == np.float64 and out.flags.c_contiguous
        t = Timing()
        self.check(self.lib.pb200_run(self.handle, phase_lo, phase_hi, wav_lo, wav_hi, dptr(out), C.byref(t)))
        return out, t.as_dict()

    def run_device(self, phase_lo, phase_hi, wav_lo, wav_hi, dev_ptr: int, want_timing=False):
        t = Timing()
        self.check(self.lib.pb200_run_device(self.handle, phase_lo, phase_hi, wav_lo, wav_hi, C.c_void_p(dev_ptr),
                                             C.byref(t) if want_timing else None))
        return t.as_dict() if want_timing else None

    def run_device_cyclic(self, phase_lo, phase_hi, wav_block, n_ranks, rank, dev_ptr: int, want_timing=False):
        t = Timing()
        self.check(self.lib.pb200_run_device_cyclic(self.handle, phase_lo, phase_hi, wav_block, n_ranks, rank,
                                                    C.c_void_p(dev_ptr), C.byref(t) if want_timing else None))
        return t.as_dict() if want_timing else None

    def set_models(self, models, n_models: int, keep=None):
        """Stack `n_models` pb200_model structs on the resident scene (n_models == 0: back to the scene)."""
        self._keep_models = keep
        self.check(self.lib.pb200_set_models(self.handle, n_models, models))

    def lightcurve(self, band: Band, phase_lo: int, phase_hi: int, dev_ptr: int | None = None) -> np.ndarray:
        out = np.empty(phase_hi - phase_lo, dtype=np.float64)
        self.check(self.lib.pb200_lightcurve(self.handle, C.byref(band), phase_lo, phase_hi,
                                             C.c_void_p(dev_ptr) if dev_ptr else None, dptr(out)))
        return out

    def run_batch(self, models, n_models: int, band: Band | None, n_orb: int, n_wav: int, want_R: bool = False,
                  keep=None):
        self._keep_models = keep
        lc = np.empty((n_models, n_orb), dtype=np.float64) if band is not None else None
        R = pinned_pool.empty((n_models, n_orb, n_wav)) if want_R else None
        t = Timing()
        self.check(self.lib.pb200_run_batch(self.handle, n_models, models, C.byref(band) if band is not None else None,
                                            dptr(lc) if lc is not None else None, dptr(R) if R is not None else None,
                                            C.byref(t)))
        return lc, R, t.as_dict()

    def run_scatter(self, phase_lo, phase_hi, wav_block, n_ranks, rank, full_ptr: int, row_stride: int, want_timing=False):
        """This rank's block-cyclic tile written in place into the full result at `full_ptr` (mapped host memory
        or device memory), see pb200_run_scatter."""
        t = Timing()
        self.check(self.lib.pb200_run_scatter(self.handle, phase_lo, phase_hi, wav_block, n_ranks, rank,
                                              C.c_void_p(full_ptr), row_stride, C.byref(t) if want_timing else None))
        return t.as_dict() if want_timing else None

    def synchronize(self):
        """Wait for everything enqueued on this context (a pb200_last_timing round trip is not needed for that)."""
        self.check(self.lib.pb200_sync(self.handle))

    def number_density_device(self, scenario: int, dev_ptr: int):
        self.check(self.lib.pb200_number_density_device(self.handle, scenario, C.c_void_p(dev_ptr)))

    def serpens_density(self, pos, x_edges, y_edges, z_edges, N, denom, weights, y_points, z_points, planet_R, star_R):
        """histogramdd * N / denom (+ Gaussian smoothing with `weights`) -> ((nx, ny, nz) array, (sum all, sum selected))."""
        pos = as_f64(pos).reshape(-1, 3)
        xb, yb, zb = as_f64(x_edges), as_f64(y_edges), as_f64(z_edges)
        w = as_f64(weights) if weights is not None and len(weights) > 1 else None
        radius = (w.size - 1) // 2 if w is not None else 0
        yp, zp = as_f64(y_points), as_f64(z_points)
        out = np.empty((xb.size - 1, yb.size - 1, zb.size - 1), dtype=np.float64)
        sums = np.empty(2, dtype=np.float64)
        self.check(self.lib.pb200_serpens_density(self.handle, dptr(pos), pos.shape[0], dptr(xb), xb.size, dptr(yb), yb.size,
                                                  dptr(zb), zb.size, float(N), float(denom), dptr(w) if w is not None else None,
                                                  radius, dptr(yp), dptr(zp), float(planet_R), float(star_R), dptr(out), dptr(sums)))
        return out, sums

    def serpens_histogram_static(self, pos_ptr: int, n: int, xb_ptr: int, nxb: int, yb_ptr: int, nyb: int, zb_ptr: int,
                                 nzb: int, N: float, denom: float, out_ptr: int):
        """pb200_serpens_histogram_static: device pointers in and out, no allocation inside, not synchronised."""
        self.check(self.lib.pb200_serpens_histogram_static(self.handle, C.c_void_p(pos_ptr), int(n), C.c_void_p(xb_ptr), nxb,
                                                           C.c_void_p(yb_ptr), nyb, C.c_void_p(zb_ptr), nzb, float(N),
                                                           float(denom), C.c_void_p(out_ptr)))

    def last_sort_order(self, n_phase: int):
        """(perm [n_phase][n_wav_sel] int32, nsig_sorted [n_phase][n_wav_sel]) of the last run, or None if it did not
        use sigma-sorted windows (pb200_last_sort_order)."""
        n = C.c_int64(0)
        self.check(self.lib.pb200_last_sort_order(self.handle, C.byref(n), None, None))
        if n.value == 0:
            return None
        perm = np.empty(n.value, dtype=np.int32)
        nsig = np.empty(n.value, dtype=np.float64)
        self.check(self.lib.pb200_last_sort_order(self.handle, C.byref(n), perm.ctypes.data_as(C.POINTER(C.c_int32)), dptr(nsig)))
        return perm.reshape(n_phase, -1), nsig.reshape(n_phase, -1)

    def log_table(self, values, offset: float = 0.0, reverse_last_axis: bool = False) -> np.ndarray:
        """log10(values + offset), optionally with the last axis reversed (pb200_log_table); page-locked result."""
        values = as_f64(values)
        n_cols = values.shape[-1] if values.ndim else 1
        out = pinned_pool.empty(values.shape)
        self.check(self.lib.pb200_log_table(self.handle, dptr(values), values.size // max(n_cols, 1), n_cols,
                                            1 if reverse_last_axis else 0, float(offset), dptr(out)))
        return out

    def last_timing(self) -> dict:
        t = Timing()
        self.check(self.lib.pb200_last_timing(self.handle, C.byref(t)))
        return t.as_dict()

    def probe_fp64(self):
        a, b = C.c_double(), C.c_double()
        self.check(self.lib.pb200_probe_fp64(self.handle, C.byref(a), C.byref(b)))
        return {"dfma_tflops": a.value, "exp_gops": b.value}

    def sigma_lookup(self, tab_x, tab_y, offset, wl):
        tab_x, tab_y = as_f64(tab_x), as_f64(tab_y)
        wl = as_f64(wl)
        out = np.empty(wl.shape, dtype=np.float64)
        self.check(self.lib.pb200_sigma_lookup(self.handle, dptr(tab_x), dptr(tab_y), tab_x.size, float(offset),
                                               dptr(wl), wl.size, dptr(out)))
        return out

    def voigt_sigma(self, line_wl, line_gamma, line_f, sigma_v, wl, offset=1e-50, want_log=False):
        lw, lg, lf, wl = as_f64(line_wl), as_f64(line_gamma), as_f64(line_f), as_f64(wl)
        # page-locked results: a table that goes back into pb200_set_scene is then uploaded without a staging copy
        sig = pinned_pool.empty(wl.shape)
        lg10 = pinned_pool.empty(wl.shape) if want_log else None
        self.check(self.lib.pb200_voigt_sigma(self.handle, dptr(lw), dptr(lg), dptr(lf), lw.size, float(sigma_v),
                                              dptr(wl), wl.size, float(offset), dptr(sig),
                                              dptr(lg10) if want_log else None))
        return (sig, lg10) if want_log else sig

    def voigt_profile(self, x, sigma, gamma):
        x = as_f64(x)
        sigma = as_f64(np.broadcast_to(sigma, x.shape))
        gamma = as_f64(np.broadcast_to(gamma, x.shape))
        out = np.empty(x.shape, dtype=np.float64)
        self.check(self.lib.pb200_voigt_profile(self.handle, dptr(x), dptr(sigma), dptr(gamma), x.size, dptr(out)))
        return out

    def molecular_sigma(self, grid, values, fill, offset, P, T, wl):
        gP, gT, gW = (as_f64(g) for g in grid)
        values = as_f64(values)
        P, wl = as_f64(P), as_f64(wl)
        n_c, n_x = P.shape
        n_wav = wl.shape[1]
        dims = (C.c_int32 * 3)(gP.size, gT.size, gW.size)
        out = np.empty((n_c, n_x, n_wav), dtype=np.float64)
        self.check(self.lib.pb200_molecular_sigma(self.handle, dptr(gP), dptr(gT), dptr(gW), dptr(values), dims,
                                                  float(fill), float(offset), dptr(P), float(T), dptr(wl),
                                                  n_c, n_x, n_wav, dptr(out)))
        return out
