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
