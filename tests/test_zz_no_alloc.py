"""pb200_serpens_histogram_static: the allocation-free variant of the particle histogram (static device scratch;
no counterpart in the reference) against the existing entry point
pb200_serpens_density, through the C ABI, on two calls in a row.  Binning is index work: bit-exact.

Collected last on purpose (file name)."""
import numpy as np
import pytest

from prometheus_b200 import capi, host, workloads

pytestmark = pytest.mark.gpu


def _edges(grid):
    x_mid, x_b, n_x, rho_b, n_rho = grid[0], grid[1], int(grid[2]), grid[3], int(grid[4])
    xb = np.linspace(x_mid - x_b, x_mid + x_b, n_x + 1)
    yb = np.linspace(-rho_b, rho_b, 2 * n_rho + 1)
    return xb, yb, yb.copy()


def test_static_histogram_equals_the_allocating_entry_point_twice():
    import torch
    ctx = host.get_context(0)
    spec = workloads.get("SERs")
    xb, yb, zb = _edges(spec["grid"])
    cell = (xb[1] - xb[0]) * (yb[1] - yb[0]) * (zb[1] - zb[0])
    yp = 0.5 * (yb[1:] + yb[:-1])
    dev = torch.device("cuda", 0)
    d_xb, d_yb, d_zb = (torch.from_numpy(e).to(dev) for e in (xb, yb, zb))
    out = torch.full((xb.size - 1, yb.size - 1, zb.size - 1), -1.0, dtype=torch.float64, device=dev)
    # second call: fewer, different particles -- counts left over from the first call would show
    for n_part, seed, N in ((20000, 3, 1e33), (777, 9, 2.5e31)):
        pos = np.ascontiguousarray(workloads.synthetic_serpens_particles(spec["grid"], n=n_part, seed=seed)[:, 0:3] * 1e2)
        denom = pos.shape[0] * cell
        ref, _ = ctx.serpens_density(pos, xb, yb, zb, N, denom, None, yp, yp, 1.0, 2.0)     # existing entry point
        d_pos = torch.from_numpy(pos).to(dev)
        torch.cuda.synchronize()                      # the inputs are in place before the context's stream reads them
        ctx.serpens_histogram_static(d_pos.data_ptr(), pos.shape[0], d_xb.data_ptr(), xb.size, d_yb.data_ptr(), yb.size,
                                     d_zb.data_ptr(), zb.size, N, denom, out.data_ptr())
        ctx.synchronize()
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert ref.sum() > 0 and np.array_equal(got, ref)


def test_static_histogram_rejects_grids_beyond_its_scratch():
    import torch
    ctx = host.get_context(0)
    buf = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    p = buf.data_ptr()
    n_edges = 130                                      # 129^3 cells > 2^21
    assert (n_edges - 1) ** 3 > capi.SERPENS_STATIC_MAX_CELLS
    with pytest.raises(capi.Pb200Error) as ei:
        ctx.serpens_histogram_static(p, 0, p, n_edges, p, n_edges, p, n_edges, 1.0, 1.0, p)
    assert ei.value.code == -1 and "PB200_SERPENS_STATIC_MAX_CELLS" in str(ei.value)
