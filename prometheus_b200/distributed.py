"""Multi-GPU Transit.sumOverChords: one process per GPU, R(phase, lambda) tiled over ranks.

R[i_orb, lambda] depends only on the chords of that phase and on that wavelength
(reference: gasProperties.py:1243-1258), so the output tiles are independent units: there is no
exchange step on the data path.  This replaces memoryHandler's RAM-bounded chord batching
(memoryHandler.py:55-66) and the multiprocessing the reference's README advertises.  The one
collective is the final all-gather of the fp64 R tiles (NCCL over NVLink; gloo in the CPU tests).

Partition: contiguous wavelength chunks of equal padded width (every rank then does the same
kernel (c') work; kernel (a') -- n_c*n_x, < 1 % of n_c*n_wav -- is replicated), or -- `axis="cyclic"`,
atomic scenes -- blocks of 256 wavelengths dealt round-robin to the ranks, which spreads line cores (table
exp on most chords) and continuum (thin path) evenly where contiguous chunks leave one rank with all the
lines; or contiguous phase blocks when asked for (`axis="phase"`: partitions kernel (a') as well; useful for dense
phase sweeps such as C3 where n_orb >> world).
"""
from __future__ import annotations

import numpy as np


def partition(n: int, world: int) -> list[tuple[int, int]]:
    """Contiguous [lo, hi) blocks of equal padded width ceil(n/world); trailing blocks may be
    short or empty."""
    per = -(-n // world)
    return [(min(r * per, n), min((r + 1) * per, n)) for r in range(world)]


def padded_width(n: int, world: int) -> int:
    return -(-n // world)


CYCLIC_BLOCK = 256   # wavelengths per block of the block-cyclic split: one window of kernel (c')


def cyclic_indices(n_wav: int, world: int, rank: int, block: int = CYCLIC_BLOCK) -> np.ndarray:
    """Global wavelength indices of `rank` in the block-cyclic split (blocks rank, rank + world, ...), in the
    order pb200_run_device_cyclic writes them."""
    n_blocks = -(-n_wav // block)
    parts = [np.arange(b * block, min((b + 1) * block, n_wav)) for b in range(rank, n_blocks, world)]
    return np.concatenate(parts) if parts else np.empty(0, dtype=np.int64)


def cyclic_width(n_wav: int, world: int, block: int = CYCLIC_BLOCK) -> int:
    """Largest per-rank wavelength count of the block-cyclic split (rank 0 has it)."""
    return int(cyclic_indices(n_wav, world, 0, block).size)


def unpack_tiles(flat: np.ndarray, n_orb: int, n_wav: int, world: int, axis: str = "wavelength") -> np.ndarray:
    """[world][n_orb*per] (axis="wavelength") or [world][per*n_wav] (axis="phase") gathered tile
    buffers -> (n_orb, n_wav).  Each rank's tile is C-contiguous [n_phase_sel][n_wav_sel] at the
    start of its slot."""
    R = np.empty((n_orb, n_wav), dtype=np.float64)
    if axis == "cyclic":
        per = cyclic_width(n_wav, world)
        flat = flat.reshape(world, n_orb * per)
        for r in range(world):
            idx = cyclic_indices(n_wav, world, r)
            if idx.size:
                R[:, idx] = flat[r, : n_orb * idx.size].reshape(n_orb, idx.size)
        return R
    if axis == "wavelength":
        per = padded_width(n_wav, world)
        flat = flat.reshape(world, n_orb * per)
        for r, (lo, hi) in enumerate(partition(n_wav, world)):
            if hi > lo:
                R[:, lo:hi] = flat[r, : n_orb * (hi - lo)].reshape(n_orb, hi - lo)
    else:
        per = padded_width(n_orb, world)
        flat = flat.reshape(world, per * n_wav)
        for r, (lo, hi) in enumerate(partition(n_orb, world)):
            if hi > lo:
                R[lo:hi, :] = flat[r, : (hi - lo) * n_wav].reshape(hi - lo, n_wav)
    return R


class ShardedTransit:
    """Keeps the per-rank state of a sharded run: context, tile bounds, gather buffers.

    `compute_tile(phase_lo, phase_hi, wav_lo, wav_hi, out_tensor)` defaults to the CUDA library
    (pb200_run_device into `out_tensor`'s memory); tests on CPU inject a host function.
    """

    def __init__(self, n_orb: int, n_wav: int, group=None, axis: str = "wavelength", device=None, compute_tile=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_orb, self.n_wav, self.axis = int(n_orb), int(n_wav), axis
        if axis == "wavelength":
            self.lo, self.hi = partition(self.n_wav, self.world)[self.rank]
            self.tile = (0, self.n_orb, self.lo, self.hi)
            self.slot = self.n_orb * padded_width(self.n_wav, self.world)
        elif axis == "cyclic":
            self.lo, self.hi = 0, int(cyclic_indices(self.n_wav, self.world, self.rank).size)   # local wavelength count
            self.tile = (0, self.n_orb, 0, self.hi)
            self.slot = self.n_orb * cyclic_width(self.n_wav, self.world)
        elif axis == "phase":
            self.lo, self.hi = partition(self.n_orb, self.world)[self.rank]
            self.tile = (self.lo, self.hi, 0, self.n_wav)
            self.slot = padded_width(self.n_orb, self.world) * self.n_wav
        else:
            raise ValueError("axis must be 'wavelength', 'cyclic' or 'phase'")
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.local = torch.zeros(self.slot, dtype=torch.float64, device=self.device)
        self.gathered = torch.empty(self.world * self.slot, dtype=torch.float64, device=self.device) if self.world > 1 else self.local
        self.host = None
        self.compute_tile = compute_tile

    @property
    def samples_tile_shape(self):
        p0, p1, w0, w1 = self.tile
        return p1 - p0, w1 - w0

    def run_device(self, gather: bool = True):
        """Enqueue this rank's tile and (gather=True) the all-gather; returns the gathered device buffer, or
        this rank's tile buffer when the tiles are left where they were computed (the output tiles are
        independent units: nothing on the data path needs them in one place)."""
        p0, p1, w0, w1 = self.tile
        if p1 > p0 and w1 > w0:
            self.compute_tile(p0, p1, w0, w1, self.local)
        if not gather:
            return self.local
        if self.world > 1:
            self.dist.all_gather_into_tensor(self.gathered, self.local, group=self.group)
        return self.gathered

    def unpack_device(self):
        """Gathered slots -> one contiguous (n_orb, n_wav) device tensor (device-side copies only)."""
        torch = self.torch
        if self.world == 1:
            return self.local[: self.n_orb * self.n_wav].view(self.n_orb, self.n_wav)
        if getattr(self, "full", None) is None:
            self.full = torch.empty((self.n_orb, self.n_wav), dtype=torch.float64, device=self.device)
        slots = self.gathered.view(self.world, self.slot)
        if self.axis == "cyclic":
            if getattr(self, "_cyc_idx", None) is None:
                self._cyc_idx = [torch.from_numpy(cyclic_indices(self.n_wav, self.world, r)).to(self.device)
                                 for r in range(self.world)]
            for r, idx in enumerate(self._cyc_idx):
                if idx.numel():
                    self.full.index_copy_(1, idx, slots[r, : self.n_orb * idx.numel()].view(self.n_orb, idx.numel()))
        elif self.axis == "wavelength":
            for r, (lo, hi) in enumerate(partition(self.n_wav, self.world)):
                if hi > lo:
                    self.full[:, lo:hi] = slots[r, : self.n_orb * (hi - lo)].view(self.n_orb, hi - lo)
        else:
            for r, (lo, hi) in enumerate(partition(self.n_orb, self.world)):
                if hi > lo:
                    self.full[lo:hi, :] = slots[r, : (hi - lo) * self.n_wav].view(hi - lo, self.n_wav)
        return self.full

    def to_host(self, dst: int | None = 0):
        """D2H of R (on `dst` only, or on every rank if dst is None) -> (n_orb, n_wav) float64.
        On CUDA the tiles are put in place on the device and land in a pooled page-locked array."""
        torch = self.torch
        if dst is not None and self.rank != dst:
            return None
        if self.device.type == "cuda":
            from . import capi
            R = capi.pinned_pool.empty((self.n_orb, self.n_wav))
            torch.from_numpy(R).copy_(self.unpack_device(), non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            return R
        return unpack_tiles(self.gathered.numpy(), self.n_orb, self.n_wav, self.world, self.axis)


def tile_function(ctx, axis: str, world: int, rank: int, want_timing: bool = False, sink: list | None = None):
    """compute_tile callable of ShardedTransit for the CUDA library: contiguous tiles through pb200_run_device,
    axis="cyclic" through pb200_run_device_cyclic (the tile bounds are then the local wavelength count)."""
    def run(p0, p1, w0, w1, out):
        if axis == "cyclic":
            t = ctx.run_device_cyclic(p0, p1, CYCLIC_BLOCK, world, rank, out.data_ptr(), want_timing)
        else:
            t = ctx.run_device(p0, p1, w0, w1, out.data_ptr(), want_timing)
        if sink is not None and t is not None:
            sink.append(t)
    return run


def sum_over_chords(transit, group=None, axis: str = "wavelength", device: int | None = None, dst: int | None = 0,
                    state: dict | None = None):
    """Sharded replacement of Transit.sumOverChords: every rank uploads the (replicated) scene,
    computes its tile on its GPU, the tiles are all-gathered, rank `dst` returns R (n_orb, n_wav).

    Must be called with torch's current CUDA stream being the one the context runs on
    (pass `state` from a previous call to reuse buffers and the context binding)."""
    import torch
    from . import host, scene
    if device is None:
        device = torch.cuda.current_device()
    ctx = host.get_context(device)
    ctx.set_stream(torch.cuda.current_stream(device).cuda_stream)
    world = torch.distributed.get_world_size(group) if torch.distributed.is_initialized() else 1
    rank = torch.distributed.get_rank(group) if torch.distributed.is_initialized() else 0
    wav_range = partition(int(np.size(transit.wavelength)), world)[rank] if (axis == "wavelength" and world > 1) else None
    sc, keep, info = scene.flatten(transit, wav_range=wav_range)
    ctx.set_scene(sc, keep)
    st = state.get("sharded") if state else None
    if st is None or (st.n_orb, st.n_wav, st.axis) != (info["n_orb"], info["n_wav"], axis):
        st = ShardedTransit(info["n_orb"], info["n_wav"], group, axis, torch.device("cuda", device),
                            compute_tile=tile_function(ctx, axis, world, rank))
        if state is not None:
            state["sharded"] = st
    else:
        st.compute_tile = tile_function(ctx, axis, world, rank)
    st.run_device()
    if state is not None:
        state["info"] = info
    return st.to_host(dst)


# ---------------------------------------------------------------------------------------------------
# node-local result without a gather: every rank's kernels store their tiles straight into ONE page-locked
# POSIX shared-memory array that all ranks of the node have mapped (pb200_run_scatter) -- the device-to-host
# transfer runs over every GPU's own PCIe link at once instead of funnelling through rank 0's.
# ---------------------------------------------------------------------------------------------------
class SharedResult:
    """Pool of shared, page-locked (n_orb, n_wav) result arrays plus the few control words the ranks need.

    Control block (int64): [0] sequence number of the current call, [1] segment index chosen by rank 0,
    [2] its size in bytes, [4 + r] sequence number rank r has completed.  x86 keeps the stores in order; the
    waits are spins on these words (microseconds: every rank has just synchronised its stream)."""

    CTL_WORDS = 4

    def __init__(self, world: int, rank: int, tag: str, register=None, unregister=None, timeout: float = 120.0):
        self.world, self.rank, self.tag, self.timeout = int(world), int(rank), str(tag), timeout
        self._register, self._unregister = register, unregister
        self.seq = 0
        self.segments = {}          # index -> (SharedMemory, device pointer)
        self.free = []              # rank 0: indices not handed out
        self.sizes = {}
        self._n_created = 0
        self.ctl_shm = self._open(f"pb200_{self.tag}_ctl", 8 * (self.CTL_WORDS + self.world), create=(self.rank == 0))
        self.ctl = np.ndarray((self.CTL_WORDS + self.world,), dtype=np.int64, buffer=self.ctl_shm.buf)
        if self.rank == 0:
            self.ctl[:] = 0

    def _open(self, name, size, create):
        """mmap of /dev/shm/<name> (plain files: no resource tracker involved); the creator sizes it, the others wait
        until it exists with its full size."""
        import mmap
        import os
        import time
        path = os.path.join("/dev/shm", name)
        if create:
            fd = os.open(path, os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
            os.ftruncate(fd, size)
        else:
            t0 = time.monotonic()
            while True:
                try:
                    fd = os.open(path, os.O_RDWR)
                    if os.fstat(fd).st_size >= size:
                        break
                    os.close(fd)
                except FileNotFoundError:
                    pass
                if time.monotonic() - t0 > self.timeout:
                    raise TimeoutError(f"SharedResult: {path} did not appear")
                time.sleep(0.0005)
        try:
            mm = mmap.mmap(fd, size)
        finally:
            os.close(fd)
        return _Segment(mm, path, size, create)

    def _spin(self, cond, what):
        import time
        t0 = time.monotonic()
        n = 0
        while not cond():
            n += 1
            if n % 4096 == 0 and time.monotonic() - t0 > self.timeout:
                raise TimeoutError(f"SharedResult: rank {self.rank} waited {self.timeout}s for {what}")

    def _segment(self, k, size):
        if k not in self.segments:
            shm = self._open(f"pb200_{self.tag}_{k}", size, create=(self.rank == 0))
            addr = C_addressof(shm.buf)
            dev = self._register(addr, shm.size) if self._register else addr
            self.segments[k] = (shm, dev, addr)
            self.sizes[k] = size
        return self.segments[k]

    def begin(self, n_orb: int, n_wav: int):
        """-> (segment index, (n_orb, n_wav) float64 view, device pointer) for this call; collective over the ranks."""
        self.seq += 1
        nbytes = int(n_orb) * int(n_wav) * 8
        if self.rank == 0:
            k = next((i for i in self.free if self.sizes[i] >= nbytes), None)
            if k is None:
                k = self._n_created
                self._n_created += 1
            else:
                self.free.remove(k)
            size = self.sizes.get(k, nbytes)
            shm, dev, _ = self._segment(k, size)
            self.ctl[1], self.ctl[2] = k, size
            self.ctl[0] = self.seq
        else:
            self._spin(lambda: self.ctl[0] >= self.seq, f"call {self.seq}")
            k, size = int(self.ctl[1]), int(self.ctl[2])
            shm, dev, _ = self._segment(k, size)
        view = np.ndarray((n_orb, n_wav), dtype=np.float64, buffer=shm.buf)
        return k, view, dev

    def end(self, k: int, view):
        """Call after this rank's stream is synchronised.  Rank 0 waits for every rank and gets the array; the others
        return None.  The segment goes back to the pool when the LAST array that looks into it dies: the returned
        array and every slice or view taken from it have one `_ResultToken` as their ultimate base (numpy collapses
        `.base` chains onto it), and the give-back hangs on that token, not on the outermost array object."""
        self.ctl[self.CTL_WORDS + self.rank] = self.seq
        if self.rank != 0:
            return None
        flags = self.ctl[self.CTL_WORDS:]
        self._spin(lambda: bool(np.all(flags >= self.seq)), "the other ranks' tiles")
        token = _ResultToken(self.segments[k][2], view.nbytes, self.free, k, view)
        return np.asarray(token).view(np.float64).reshape(view.shape)

    def close(self):
        for k, (shm, dev, addr) in list(self.segments.items()):
            try:
                if self._unregister:
                    self._unregister(addr)
            except Exception:
                pass
            shm.close()
        self.segments.clear()
        if self.ctl_shm is not None:
            self.ctl = None
            self.ctl_shm.close()
            self.ctl_shm = None


class _ResultToken:
    """Lifetime token of one handed-out result: exposes the segment's memory through __array_interface__, so every
    numpy array derived from the result keeps it alive; returns the segment index to the pool's free list when the
    last of them is gone (capi._PinnedBlock does the same for page-locked blocks)."""

    def __init__(self, addr: int, nbytes: int, free_list: list, k: int, keep):
        self._free, self._k, self._keep = free_list, k, keep      # keep: the mmap-backed view (the mapping stays valid)
        self.__array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(addr), False), "version": 3}

    def __del__(self):
        try:
            self._free.append(self._k)
        except Exception:
            pass


class _Segment:
    def __init__(self, mm, path, size, owner):
        self.buf, self.path, self.size, self.owner = mm, path, size, owner

    def close(self):
        import os
        try:
            self.buf.close()
        except BufferError:
            pass              # views handed to the caller are still alive: the mapping goes with the process
        if self.owner:
            try:
                os.unlink(self.path)
            except FileNotFoundError:
                pass


def C_addressof(buf) -> int:
    import ctypes
    return ctypes.addressof(ctypes.c_char.from_buffer(buf))


def shared_result_for(group=None, state: dict | None = None) -> SharedResult:
    """The SharedResult of this process group (created on first use; rank 0's pid and clock name the segments)."""
    import os
    import time
    import torch.distributed as dist
    if state is not None and state.get("shared") is not None:
        return state["shared"]
    from . import capi
    lib = capi.load()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    tag = [f"{os.getpid()}_{time.monotonic_ns()}"]
    if world > 1:
        dist.broadcast_object_list(tag, src=0, group=group)

    def register(addr, nbytes):
        import ctypes as C
        dev = C.c_void_p()
        rc = lib.pb200_host_register(C.c_void_p(addr), nbytes, C.byref(dev))
        if rc:
            raise capi.Pb200Error(rc, (lib.pb200_last_error(None) or b"").decode())
        return int(dev.value)

    def unregister(addr):
        import ctypes as C
        lib.pb200_host_unregister(C.c_void_p(addr))

    sr = SharedResult(world, rank, tag[0], register, unregister)
    import atexit
    atexit.register(sr.close)
    if state is not None:
        state["shared"] = sr
    return sr


class _DevArray:
    """A device pointer of the library presented through __cuda_array_interface__ (for torch.as_tensor)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


def _device_tensor(torch, ptr: int, n: int, device: int, state: dict | None):
    cache = state.setdefault("dev_tensors", {}) if state is not None else {}
    key = (ptr, n, device)
    if key not in cache:
        cache[key] = torch.as_tensor(_DevArray(ptr, n), device=torch.device("cuda", device))
    return cache[key]


def sum_over_chords_shared(transit, group=None, device: int | None = None, state: dict | None = None):
    """Sharded Transit.sumOverChords for the ranks of ONE node without a gather: block-cyclic wavelength tiles
    (atomic scenes), each rank's kernels store into the shared page-locked result (pb200_run_scatter).  Rank 0
    returns R (n_orb, n_wav) -- a view of pooled shared memory, valid until it is garbage collected -- the others
    None.  Replaces memoryHandler's chord batching (memoryHandler.py:55-66) on a multi-GPU node."""
    import torch
    from . import host, scene
    if device is None:
        device = torch.cuda.current_device()
    ctx = host.get_context(device)
    ctx.set_stream(torch.cuda.current_stream(device).cuda_stream)
    world = torch.distributed.get_world_size(group) if torch.distributed.is_initialized() else 1
    rank = torch.distributed.get_rank(group) if torch.distributed.is_initialized() else 0
    import time
    sr = shared_result_for(group, state)
    timeline = state is not None and state.get("timeline")       # CUDA events on the context's stream (bench.py asks)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if timeline else None
    t0 = time.perf_counter()
    sc, keep, info = scene.flatten(transit)
    t1 = time.perf_counter()
    if ev:
        ev[0].record()
    if world > 1:
        # every rank sends one world-th of each large sigma table over its PCIe link; the rest arrives from the
        # peers over NVLink (one in-place all-gather per table)
        for ptr, chunk in ctx.set_scene_sharded(sc, keep, world, rank):
            info["h2d_bytes"] -= chunk * 8 * (world - 1)          # the peers' slices do not cross this rank's PCIe link
            info["nvlink_allgather_bytes"] = info.get("nvlink_allgather_bytes", 0) + chunk * 8 * world
            full = _device_tensor(torch, ptr, chunk * world, device, state)
            torch.distributed.all_gather_into_tensor(full, full[rank * chunk:(rank + 1) * chunk], group=group)
    else:
        ctx.set_scene(sc, keep)
    if ev:
        ev[1].record()
    t2 = time.perf_counter()
    k, view, dev = sr.begin(info["n_orb"], info["n_wav"])
    t3 = time.perf_counter()
    ctx.run_scatter(0, info["n_orb"], CYCLIC_BLOCK, world, rank, dev, info["n_wav"])
    if ev:
        ev[2].record()
    t3b = time.perf_counter()
    ctx.synchronize()
    t4 = time.perf_counter()
    out = sr.end(k, view)
    t5 = time.perf_counter()
    if state is not None:
        state["info"] = info
        state["host_ms"] = {"flatten": (t1 - t0) * 1e3, "set_scene_call": (t2 - t1) * 1e3, "begin": (t3 - t2) * 1e3,
                            "run_enqueue": (t3b - t3) * 1e3, "sync_wait": (t4 - t3b) * 1e3, "end_wait": (t5 - t4) * 1e3}
        if ev:   # device side: scene upload + table all-gather | sigma, column, compaction, transit + the result stores
            state["host_ms"]["dev_upload_allgather"] = ev[0].elapsed_time(ev[1])
            state["host_ms"]["dev_kernels_and_stores"] = ev[1].elapsed_time(ev[2])
    return out
