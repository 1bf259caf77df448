"""world_size > 1 on CPU (gloo): partitioning, the single all-gather and the unpacking of the R tiles
(prometheus_b200/distributed.py).  The tile computation is injected (a host function that copies
the tile out of a known R), so no GPU is needed; the GPU path is the same code with
compute_tile = pb200_run_device."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from prometheus_b200 import distributed as pbd


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _truth(n_orb, n_wav):
    return (np.arange(n_orb)[:, None] * 1000.0 + np.arange(n_wav)[None, :] + 0.25)


def _worker(rank, world, port, n_orb, n_wav, axis, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        R_true = _truth(n_orb, n_wav)
        calls = []

        def compute_tile(p0, p1, w0, w1, out):
            if axis == "cyclic":      # what pb200_run_device_cyclic writes: the rank's blocks, in order
                idx = pbd.cyclic_indices(n_wav, world, rank)
                assert (w0, w1) == (0, idx.size)
                calls.append(("cyclic", p0, p1, idx.tolist()))
                tile = np.ascontiguousarray(R_true[p0:p1][:, idx]).reshape(-1)
            else:
                calls.append((p0, p1, w0, w1))
                tile = np.ascontiguousarray(R_true[p0:p1, w0:w1]).reshape(-1)
            out[: tile.size] = torch.from_numpy(tile)

        st = pbd.ShardedTransit(n_orb, n_wav, None, axis, "cpu", compute_tile)
        st.run_device()
        R0 = st.to_host(0)
        Rall = st.to_host(None)
        ok = np.array_equal(Rall, R_true) and ((R0 is None) if rank else np.array_equal(R0, R_true))
        q.put((rank, bool(ok), calls))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_orb,n_wav,axis", [(2, 5, 219, "wavelength"), (2, 5, 219, "phase"),
                                                    (3, 4, 10, "wavelength"), (3, 2, 7, "phase"), (2, 1, 1, "wavelength"),
                                                    (2, 3, 1000, "cyclic"), (3, 2, 300, "cyclic")])
def test_sharded_gather(world, n_orb, n_wav, axis):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_orb, n_wav, axis, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    covered = np.zeros((n_orb, n_wav), dtype=int)
    for rank, ok, calls in results:
        assert ok, f"rank {rank}"
        for call in calls:
            if call[0] == "cyclic":
                _, p0, p1, idx = call
                covered[p0:p1, idx] += 1
            else:
                p0, p1, w0, w1 = call
                covered[p0:p1, w0:w1] += 1
    assert np.all(covered == 1), "tiles must partition the output exactly once"


def test_partition_properties():
    for n in (0, 1, 7, 219, 31969, 300306):
        for world in (1, 2, 3, 4, 8):
            parts = pbd.partition(n, world)
            assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            assert all(0 <= hi - lo <= pbd.padded_width(n, world) for lo, hi in parts)


def test_unpack_single_rank():
    R = _truth(3, 11)
    assert np.array_equal(pbd.unpack_tiles(R.reshape(-1).copy(), 3, 11, 1, "wavelength"), R)
    assert np.array_equal(pbd.unpack_tiles(R.reshape(-1).copy(), 3, 11, 1, "phase"), R)


def test_cyclic_indices_match_the_library_count():
    from prometheus_b200 import capi
    lib = capi.load()
    for n in (1, 255, 256, 257, 31969, 300306):
        for world in (1, 2, 3, 8):
            seen = np.concatenate([pbd.cyclic_indices(n, world, r) for r in range(world)])
            assert np.array_equal(np.sort(seen), np.arange(n))
            for r in range(world):
                assert lib.pb200_cyclic_count(n, pbd.CYCLIC_BLOCK, world, r) == pbd.cyclic_indices(n, world, r).size


# ---- SharedResult: the gather-free result path (control words + pooled shared segments), no GPU needed ------
def _shared_worker(rank, world, tag, n_orb, n_wav, n_calls, q):
    sr = pbd.SharedResult(world, rank, tag, timeout=60.0)
    try:
        ok = True
        held = []
        for call in range(n_calls):
            k, view, dev = sr.begin(n_orb, n_wav)
            idx = pbd.cyclic_indices(n_wav, world, rank)
            view[:, idx] = _truth(n_orb, n_wav)[:, idx] + call            # what pb200_run_scatter stores
            R = sr.end(k, view)
            if rank == 0:
                ok &= bool(np.array_equal(R, _truth(n_orb, n_wav) + call))
                if call == 0:
                    held.append(R)                                         # a result the caller keeps ...
            else:
                ok &= R is None
        if rank == 0:
            ok &= bool(np.array_equal(held[0], _truth(n_orb, n_wav)))      # ... is not overwritten by later calls
            ok &= len(sr.segments) <= 3                                    # one held, the others recycled
        q.put((rank, ok))
        # leave only after everyone is done with the segments
        sr.ctl[sr.CTL_WORDS + rank] = 10 ** 9
        if rank == 0:
            sr._spin(lambda: bool(np.all(sr.ctl[sr.CTL_WORDS:] >= 10 ** 9)), "exit")
    finally:
        held = None
        sr.close()


@pytest.mark.parametrize("world,n_orb,n_wav", [(2, 5, 1000), (3, 2, 300)])
def test_shared_result_protocol(world, n_orb, n_wav):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    tag = f"test{os.getpid()}_{world}_{n_wav}"
    procs = [ctx.Process(target=_shared_worker, args=(r, world, tag, n_orb, n_wav, 4, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in results), results


def test_shared_result_segment_lives_as_long_as_any_slice_of_it():
    """A caller that keeps only a slice of the returned array (R[0], R[:, lo:hi]) still owns the segment: it must not
    go back to the pool -- and be overwritten by the next call's kernels -- while that slice is alive."""
    import gc
    sr = pbd.SharedResult(1, 0, f"slice{os.getpid()}", timeout=10.0)
    try:
        k0, view, _ = sr.begin(4, 100)
        view[...] = 7.0
        R = sr.end(k0, view)
        row = R[1]                      # numpy collapses .base: the slice does not reference R itself
        part = R[:, 10:20]
        del R, view
        gc.collect()
        assert sr.free == [], "segment given back while slices of the result are alive"
        k1, view1, _ = sr.begin(4, 100)
        view1[...] = -1.0               # what the next call's kernels would store
        assert k1 != k0 and np.all(row == 7.0) and np.all(part == 7.0)
        R1 = sr.end(k1, view1)
        del row
        gc.collect()
        assert sr.free == []            # `part` still holds segment k0
        del part
        gc.collect()
        assert sr.free == [k0]
        k2, view2, _ = sr.begin(4, 100)
        assert k2 == k0                 # recycled only now
        sr.end(k2, view2)
        del R1, view1, view2
    finally:
        gc.collect()
        sr.close()
