"""pb200_serpens_density against the oracle (numpy histogramdd + scipy gaussian_filter) at a realistic size."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "oracle"))
import prometheus_oracle as oracle
from prometheus_b200 import gas, host, workloads
spec = workloads.get("C4")
grid = spec["grid"]
n = 4_000_000
pos = workloads.synthetic_serpens_particles(grid, n=n)[:, 0:3] * 1e2
N, sigma = 1e33, 1.5
x_mid, x_b, n_x, rho_b, n_rho = grid[0], grid[1], int(grid[2]), grid[3], int(grid[4])
xb = np.linspace(x_mid - x_b, x_mid + x_b, n_x + 1); yb = np.linspace(-rho_b, rho_b, 2 * n_rho + 1); zb = yb.copy()
cv = (xb[1] - xb[0]) * (yb[1] - yb[0]) * (zb[1] - zb[0])
yp = np.linspace(-rho_b, rho_b, 2 * n_rho, endpoint=False) + 2. * rho_b / (4. * n_rho)
ctx = host.get_context(0)
w = gas.gaussian_weights(sigma)
for i in range(3):
    t0 = time.perf_counter()
    vals, sums = ctx.serpens_density(pos, xb, yb, zb, N, n * cv, w, yp, yp, 8.5e9, 7.2e10)
    t_gpu = time.perf_counter() - t0
t0 = time.perf_counter()
_, ref = oracle.serpens_density(pos, grid, N, sigma)
t_cpu = time.perf_counter() - t0
print(f"{n} particles -> {vals.shape} cells: GPU call {t_gpu * 1e3:.1f} ms (incl. H2D of {pos.nbytes / 1e6:.0f} MB, D2H of {vals.nbytes / 1e6:.0f} MB), "
      f"numpy/scipy {t_cpu * 1e3:.0f} ms, max |diff| / max = {np.max(np.abs(vals - ref)) / ref.max():.2e}, sum diff {abs(sums[0] - ref.sum()) / ref.sum():.1e}")
