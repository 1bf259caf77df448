"""The drop-in in ONE piece: JSON setup file (the reference's schema) -> the reference's own `prometheus.py` ->
prometheus_b200.install() -> GPU -> output/<name>.txt, against the file the un-patched reference writes
(prometheus.py:56-156).  Needs the reference copy oracle/_ref (oracle/make_ref.py); skipped cleanly without it."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from helpers import ROOT, excess_depth_ok, oracle, oracle_inputs
from prometheus_b200 import workloads

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_ref  # noqa: E402

needs_ref = pytest.mark.skipif(not make_ref.available(), reason="oracle/_ref is absent (run python oracle/make_ref.py)")
RUNNER = os.path.join(ROOT, "tests", "cli_runner.py")


def _run_cli(name, spec, gpu):
    setup = os.path.join(make_ref.REF_ROOT, "setupFiles", f"{name}.txt")
    out = os.path.join(make_ref.REF_ROOT, "output", f"{name}.txt")
    with open(setup, "w") as fh:
        json.dump(workloads.to_setup_file(spec), fh)           # what setup.createSetupFile writes (setup.py:576-578)
    if os.path.exists(out):
        os.unlink(out)
    cmd = [sys.executable, RUNNER, name] + (["--gpu"] if gpu else [])
    pr = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert pr.returncode == 0, pr.stdout[-2000:] + pr.stderr[-2000:]
    assert "PROMETHEUS finished" in pr.stdout
    text = open(out).read()
    os.unlink(out)
    return text, pr.stdout


def _parse(text):
    lines = text.split("\n")
    header = [l for l in lines if l.startswith("#")]
    table = np.loadtxt(lines)
    return header, table[0], table[1:, 0], table[1:, 1:].T          # header, first row, wavelength, R (n_orb, n_wav)


def test_setup_file_round_trip():
    for name in ("C1", "C3", "C4", "C3s", "C4s", "C2a"):
        spec = workloads.cli_twin(name)
        for scn in spec["scenarios"]:
            if scn.get("molecules") == []:
                del scn["molecules"]                            # C2a: "no molecules" has no place in the file
        back = workloads.from_setup_file(json.loads(json.dumps(workloads.to_setup_file(spec))), name)
        for key in ("scenarios", "wavelength", "grid", "doppler"):
            assert back[key] == spec[key], (name, key)
        assert back["bodies"]["planet"] == spec["bodies"]["planet"]
    for name in ("C2", "C5", "PLNs", "THMs", "FSTs"):
        with pytest.raises(ValueError):
            workloads.to_setup_file(workloads.get(name))


@needs_ref
def test_reference_cli_reproduces_the_golden_R(golden):
    """The un-patched CLI on the C1 setup file gives the R of tests/golden/C1.npz bit for bit: the JSON twin, the
    runner and the body scalars in workloads.BODIES describe the same scene as the reference's catalogue."""
    text, _ = _run_cli("pb200_c1_cpu", workloads.cli_twin("C1"), gpu=False)
    header, first, wl, R = _parse(text)
    g = golden("C1")
    assert header[0] == "# Prometheus output file." and len(header) == 3
    assert np.isnan(first[0]) and np.array_equal(first[1:], g["orb"] / (2.0 * np.pi))
    assert np.array_equal(wl, g["wavelength"]) and np.array_equal(R, g["R"])


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C1", "C3s"])
def test_cli_through_install_matches_the_unpatched_run(name, nist_rows):
    spec = workloads.cli_twin(name)
    ref_text, _ = _run_cli(f"pb200_{name}_ref", spec, gpu=False)
    gpu_text, log = _run_cli(f"pb200_{name}_gpu", spec, gpu=True)
    h_ref, first_ref, wl_ref, R_ref = _parse(ref_text)
    h_gpu, first_gpu, wl_gpu, R_gpu = _parse(gpu_text)
    assert h_gpu == h_ref
    assert gpu_text.split("\n")[3] == ref_text.split("\n")[3]          # first row: "nan phase/2pi ...", byte for byte
    assert np.array_equal(wl_gpu, wl_ref)                               # %.18e round-trips float64: bit-exact grid
    assert R_gpu.shape == R_ref.shape
    m = oracle.build_model(spec, oracle_inputs(spec, nist_rows))
    R0 = oracle.sum_over_chords(m, opaque_only=True)
    ok, worst = excess_depth_ok(R_gpu, R_ref, R0)
    assert ok, f"{name}: worst excess-depth error / bound = {worst:.3g}"
    assert float((R0 - R_ref).max()) > 1e-3
