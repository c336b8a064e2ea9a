"""The library's own bounds checker (compute-sanitizer is closed on the B200 pool: "runs under it have left GPUs needing a
reset"): librtb200_checked.so (-DRTB_CHECKED) compares every index the kernels form with the size of what it indexes and
counts violations.  The workload of tools/gpu_sanitize.py - every kernel of the library - must finish with all counters zero,
and the counting path itself must work (a deliberate violation is seen)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "3360-ray-tracer_b200" / "librtb200_checked.so"

CODE = r"""
import runpy, sys, numpy as np
sys.path.insert(0, %r)
import __graft_entry__ as g
m = g.load_package()
checked, c0 = m.check_failures()
assert checked, "not a checked build"
runpy.run_path(%r, run_name="__main__")
checked, c = m.check_failures()
print("CHECK_FAILURES", " ".join(str(int(x)) for x in c))
probe = np.zeros(8, np.uint64); probe[7] = 0x5e1f7e57
lib = m.load()
assert lib.rtb_debug_check_failures(probe.ctypes.data) == 1
print("SELFTEST", int(probe[7]) - int(c[7]))
"""


def test_every_kernel_runs_clean_under_the_checked_build(tmp_path):
    if not LIB.exists():
        pytest.fail(f"{LIB} missing: run __graft_entry__.build()")
    r = subprocess.run([sys.executable, "-c", CODE % (str(ROOT), str(ROOT / "tools" / "gpu_sanitize.py"))], capture_output=True, text=True,
                       cwd=ROOT, timeout=900, env=dict(os.environ, RTB_LIB=LIB.name))
    assert r.returncode == 0, (r.stdout[-600:], r.stderr[-1200:])
    out = {ln.split()[0]: ln.split()[1:] for ln in r.stdout.split("\n") if ln.startswith(("CHECK_FAILURES", "SELFTEST"))}
    assert "sanitize workload done" in r.stdout
    assert out["CHECK_FAILURES"] == ["0"] * 8, f"bounds violations per class: {out['CHECK_FAILURES']}"
    assert out["SELFTEST"] == ["1"], "the checker did not see a deliberate violation"
    log = ROOT / "gpurun_out"
    log.mkdir(exist_ok=True)
    (log / "checked_build.txt").write_text(r.stdout)
