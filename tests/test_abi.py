"""CPU-side checks of the boundary: the shared library loads, exports every symbol the header
declares, the ctypes structs have the C layout, and the host mirror reproduces the reference's
axes / wavelength grids bit-exactly (no compute call is made: there is no GPU here)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import prometheus_b200 as pb
from prometheus_b200 import builder, capi, workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "prometheus_b200.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pb200_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built_and_loads():
    assert os.path.exists(capi.LIB_PATH), "run python -m prometheus_b200.build"
    lib = capi.load()
    assert lib.pb200_abi_version() == capi.ABI_VERSION


def test_every_header_symbol_is_exported_and_bound():
    syms = header_symbols()
    assert len(syms) >= 20
    lib = capi.load()
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/prometheus_b200.h but not exported"
    assert sorted(capi.EXPORTS) == syms, "capi.EXPORTS and the header disagree"


def test_exported_symbols_are_unmangled_c():
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    names = {line.split()[-1] for line in out.splitlines() if line.strip()}
    for s in header_symbols():
        assert s in names


def test_struct_layout_matches_header(tmp_path):
    """sizeof/offsetof from a C translation unit that includes the header == ctypes layout."""
    src = tmp_path / "layout.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "prometheus_b200.h"\n'
        "int main(void){\n"
        'printf("%zu %zu %zu %zu\\n", sizeof(pb200_scenario), sizeof(pb200_constituent), sizeof(pb200_scene), sizeof(pb200_timing));\n'
        'printf("%zu %zu %zu\\n", offsetof(pb200_scenario, centre_x), offsetof(pb200_scenario, dims), offsetof(pb200_constituent, fill));\n'
        'printf("%zu %zu %zu\\n", offsetof(pb200_scene, wavelength), offsetof(pb200_scene, planet_y), offsetof(pb200_scene, constituents));\n'
        'printf("%zu %zu %zu %zu\\n", sizeof(pb200_model), sizeof(pb200_band), offsetof(pb200_model, moon_y), offsetof(pb200_band, bandwidth));\n'
        "return 0;}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    rows = [list(map(int, l.split())) for l in subprocess.run([str(exe)], capture_output=True, text=True).stdout.splitlines()]
    assert rows[0] == [C.sizeof(capi.Scenario), C.sizeof(capi.Constituent), C.sizeof(capi.Scene), C.sizeof(capi.Timing)]
    assert rows[1] == [capi.Scenario.centre_x.offset, capi.Scenario.dims.offset, capi.Constituent.fill.offset]
    assert rows[2] == [capi.Scene.wavelength.offset, capi.Scene.planet_y.offset, capi.Scene.constituents.offset]
    assert rows[3] == [C.sizeof(capi.Model), C.sizeof(capi.Band), capi.Model.moon_y.offset, capi.Band.bandwidth.offset]


def test_no_gpu_means_loud_failure():
    lib = capi.load()
    if lib.pb200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(capi.Pb200Error) as ei:
        capi.Context(0)
    assert ei.value.code == -3 and "no CPU fallback" in str(ei.value)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "prometheus_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "prometheus_oracle" not in text and "import oracle" not in text, f


@pytest.mark.parametrize("name", workloads.GOLDEN_CASES)
def test_mirror_axes_and_wavelength_grid_bit_exact(name, nist_rows, golden):
    """Grid.construct*Axis / WavelengthGrid.constructWavelengthGrid of the host mirror against the
    arrays the reference produced (geometryHandler.py:112-208, gasProperties.py:990-1071)."""
    g = golden(name)
    tr = builder.build_transit(workloads.get(name), nist_rows=nist_rows, build_tables=False)
    grid = tr.spatialGrid
    assert np.array_equal(grid.constructXaxis(), g["x"])
    assert np.array_equal(grid.constructRhoAxis(), g["rho"])
    assert np.array_equal(grid.constructPhiAxis(), g["phi"])
    assert np.array_equal(grid.constructOrbphaseAxis(), g["orb"])
    assert grid.getDeltaX() == float(g["delta_x"])
    assert np.array_equal(tr.wavelength, g["wavelength"])
    chords = grid.getChordGrid()
    assert np.array_equal(chords[:64], g["chord_grid_head"])
    assert np.array_equal(chords[-64:], g["chord_grid_tail"])


def test_full_size_workload_shapes(nist_rows):
    """The five BASELINE configs instantiate with the shapes SURVEY 8 lists."""
    expect = {"C1": (12000, 219), "C3": (962400, 4048), "C4": (459000, 31969)}
    for name, (n_c, n_wav) in expect.items():
        tr = builder.build_transit(workloads.get(name), nist_rows=nist_rows, build_tables=False)
        assert workloads.shape(workloads.get(name))["n_chords"] == n_c
        assert tr.wavelength.size == n_wav


def test_flatten_axes_cache_follows_the_grid(nist_rows):
    """scene.flatten keeps the grid axes per (class, constructor arguments, phiStarrot): a changed grid or rotation
    angle must give new axes, an unchanged one the very same arrays, all equal to what the grid's methods return."""
    from prometheus_b200 import scene
    tr = builder.build_transit(workloads.get("C1"), nist_rows=nist_rows, build_tables=False)
    for d in tr.atmosphere.densityDistributionList:
        for con in d.constituents:
            con.lookupFunction = pb.host.TableFunction1D(np.linspace(5.88e-5, 5.91e-5, 64), np.zeros(64))
    g = tr.spatialGrid
    a = scene._axes(g, 0.0)
    assert scene._axes(g, 0.0)[0] is a[0]
    assert np.array_equal(a[0], g.constructXaxis()) and np.array_equal(a[1], g.constructRhoAxis())
    assert np.array_equal(a[2], g.constructOrbphaseAxis()) and np.array_equal(a[3], np.sin(g.constructPhiAxis()))
    g.x_steps = int(g.x_steps) + 3
    b = scene._axes(g, 0.0)
    assert b[0].size == a[0].size + 3 and np.array_equal(b[0], g.constructXaxis()) and b[6] == g.getDeltaX()
    c = scene._axes(g, 0.25)
    assert np.array_equal(c[5], np.cos(g.constructPhiAxis() - 0.25)) and not np.array_equal(c[5], b[5])
    sc, keep, info = scene.flatten(tr)
    assert info["n_x"] == b[0].size and sc.n_x == b[0].size
