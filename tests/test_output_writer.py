"""pb200_write_output against the reference's np.savetxt call (prometheus.py:142-156): the files must
be byte-identical.  Host code, no GPU needed."""
import time

import numpy as np
import pytest

from prometheus_b200 import io as pbio


def both(tmp_path, wl, orb, R, **kw):
    a, b = tmp_path / "ours.txt", tmp_path / "ref.txt"
    pbio.save_output(str(a), wl, orb, R, **kw)
    pbio.save_output_reference(str(b), wl, orb, R)
    return a.read_bytes(), b.read_bytes()


@pytest.mark.parametrize("name", ["C1", "C3s", "C5s"])
def test_golden_R_files_are_byte_identical(name, golden, tmp_path):
    g = golden(name)
    ours, ref = both(tmp_path, g["wavelength"], g["orb"], g["R"])
    assert ours == ref
    assert ours.startswith(b"# Prometheus output file.\n# First row: Orbital phases [1]\n# All other rows")


def test_awkward_values_and_shapes(tmp_path):
    rng = np.random.default_rng(1)
    for n_orb, n_wav in [(1, 1), (3, 7), (21, 9000), (5, 4097)]:
        wl = np.sort(rng.uniform(1e-5, 1e-4, n_wav))
        orb = np.linspace(-0.3, 0.3, n_orb)
        R = 1.0 - 10 ** rng.uniform(-320, 0, (n_orb, n_wav))       # denormals, exact 1.0, tiny excesses
        special = [0.0, 1.0, 5e-324, 1e-304, np.nan, np.inf, -np.inf, -0.0, 0.1, 1e100, -1.5e-7, 2.2250738585072014e-308]
        idx = rng.choice(R.size, size=min(R.size, len(special)), replace=False)
        R.flat[idx] = special[: idx.size]
        for threads in (0, 1, 3):
            ours, ref = both(tmp_path, wl, orb, R, threads=threads)
            assert ours == ref, (n_orb, n_wav, threads)


def test_shape_mismatch_raises(tmp_path):
    with pytest.raises(ValueError):
        pbio.save_output(str(tmp_path / "x.txt"), np.ones(4), np.ones(2), np.ones((4, 2)))


def test_writer_is_much_faster_than_savetxt(tmp_path):
    rng = np.random.default_rng(2)
    n_orb, n_wav = 21, 30000
    wl, orb, R = np.sort(rng.uniform(2e-5, 8e-5, n_wav)), np.linspace(-0.2, 0.2, n_orb), rng.uniform(0.9, 1.0, (n_orb, n_wav))
    t0 = time.perf_counter(); pbio.save_output(str(tmp_path / "a.txt"), wl, orb, R); t_ours = time.perf_counter() - t0
    t0 = time.perf_counter(); pbio.save_output_reference(str(tmp_path / "b.txt"), wl, orb, R); t_ref = time.perf_counter() - t0
    assert (tmp_path / "a.txt").read_bytes() == (tmp_path / "b.txt").read_bytes()
    print(f"writer {t_ours:.3f} s, np.savetxt {t_ref:.3f} s")
    assert t_ours < t_ref
