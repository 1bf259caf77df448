"""Runs the UNMODIFIED reference entry point `python prometheus.py <setup_name>` from oracle/_ref/Prometheus,
optionally after prometheus_b200.install() has rebound the hot path to the CUDA library:

    python tests/cli_runner.py <setup_name> [--gpu] [--max-memory G]

The setup file is read from oracle/_ref/setupFiles/<setup_name>.txt and the result written to
oracle/_ref/output/<setup_name>.txt by the reference's own code (prometheus.py:56, :156).  TEST INFRASTRUCTURE."""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import make_ref  # noqa: E402


def main(argv):
    gpu = "--gpu" in argv
    args = [a for a in argv if a != "--gpu"]
    gasprop = make_ref.load()[0]               # puts the reference directory (and stubs if needed) on sys.path
    if gpu:
        import prometheus_b200
        prometheus_b200.install(gasprop)       # Transit.sumOverChords + calculateVoigtProfile -> libprometheus_b200.so
    script = os.path.join(make_ref.REF_DIR, "prometheus.py")
    sys.argv = [script] + args
    runpy.run_path(script, run_name="__main__")


if __name__ == "__main__":
    main(sys.argv[1:])
