"""Aggregate host<->device bandwidth of the node, all ranks at once -- the ceiling of the e2e figure at N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 profiles/host_link_probe.py

Every rank moves `MB` megabytes between page-locked host memory and its GPU, all ranks between the same barriers:
D2H by copy engine, D2H by zero-copy stores of a kernel (what pb200_run_scatter does), H2D by copy engine, and both
directions at once.  Prints one JSON line: per-rank mean and aggregate GB/s."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

MB = 13


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = MB * (1 << 20) // 8
    host_a = torch.empty(n, dtype=torch.float64).pin_memory()
    host_b = torch.empty(n, dtype=torch.float64).pin_memory()
    d_a = torch.randn(n, dtype=torch.float64, device=dev)
    d_b = torch.empty(n, dtype=torch.float64, device=dev)
    s2 = torch.cuda.Stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, reps=20):
        out = []
        for i in range(reps + 3):
            barrier()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize(dev)
            dt = time.perf_counter() - t0
            if i >= 3:
                out.append(dt)
        t = torch.tensor([sum(out) / len(out)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def both():
        host_a.copy_(d_a, non_blocking=True)
        with torch.cuda.stream(s2):
            d_b.copy_(host_b, non_blocking=True)
        s2.synchronize()

    res = {}
    res["d2h_copy"] = timed(lambda: host_a.copy_(d_a, non_blocking=True))
    res["h2d_copy"] = timed(lambda: d_b.copy_(host_b, non_blocking=True))
    res["both_copy"] = timed(both)
    # zero-copy stores: a device kernel writing mapped page-locked memory (torch exposes pinned memory to the device)
    try:
        import ctypes
        cudart = ctypes.CDLL("libcudart.so")
        dptr = ctypes.c_void_p()
        rc = cudart.cudaHostGetDevicePointer(ctypes.byref(dptr), ctypes.c_void_p(host_a.data_ptr()), 0)
        if rc == 0 and dptr.value:
            class _Arr:
                def __init__(self, p, k):
                    self.__cuda_array_interface__ = {"shape": (k,), "typestr": "<f8", "data": (p, False), "version": 2}
            mapped = torch.as_tensor(_Arr(dptr.value, n), device=dev)
            res["d2h_zero_copy_stores"] = timed(lambda: mapped.copy_(d_a))
    except Exception as exc:       # the probe still reports the copy-engine figures
        res["d2h_zero_copy_stores_error"] = str(exc)[:120]
    if rank == 0:
        gb = MB * (1 << 20) / 1e9
        out = {"n_gpus": world, "mb_per_rank": MB}
        for k, t in res.items():
            if isinstance(t, float):
                mult = 2.0 if k == "both_copy" else 1.0
                out[k] = {"ms": t * 1e3, "per_rank_GBs": mult * gb / t, "aggregate_GBs": mult * gb * world / t}
            else:
                out[k] = t
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
