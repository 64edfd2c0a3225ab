"""Memory-safety evidence without compute-sanitizer (closed on this GPU pool).

``libgol_b200_debug.so`` is the library compiled with ``-DGOL_DEBUG_BOUNDS=1``: every global
load / store / atomic of every kernel checks its address range against a registry of the buffers
the library owns (each slab buffer, recv buffer, flag-in-data buffer, flag, scratch ...), and every
buffer is fenced by guard rows that ``sync()`` verifies.  The kernel test-suites - the hypothesis
property test, the benchmark-width oracle tests, the per-kernel tests - are run against that
build in a subprocess and must pass with zero trips; a deliberate bad access proves the detector
fires.  (SURVEY.md section 4 "Sanitizers"; VERDICT r1 task 8.)"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_py(code: str, timeout=900):
    env = dict(os.environ, GOL_B200_LIB="debug", PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_detector_fires_on_a_bad_address_and_on_a_trampled_guard_row():
    code = r"""
import numpy as np, torch
from gol_b200 import World, GolError
from gol_b200 import _native as n
assert n.debug_bounds_enabled()
w = World(512, 64, nranks=2, backend="cuda", mode="fused", k=4, ndevices=1); w.connect(); w.init_pattern(6, 1, 0.5)
w.step(20); w.sync()
assert n.debug_bounds_report()["violations"] == 0
# 1. a checked load of an address that belongs to no registered buffer (inside a guard row)
stride_bytes = w.stride * 8
# device_ptr is interior row 0 and G = 32 margin rows precede it inside the buffer: the word right
# in front of those lies in the guard row
n.check(n.lib.gol_debug_bounds_poke(w.device_ptr(0) - 32 * stride_bytes - 8))
rep = n.debug_bounds_report()
assert rep["violations"] >= 1 and rep["first_bad_address"] != 0, rep
try:
    w.sync(); raise SystemExit("sync() did not report the out-of-bounds access")
except GolError as e:
    assert "out-of-bounds" in str(e), e
n.debug_bounds_report(reset=True)
w.sync()
# 2. trample a guard row from the host side: the next sync() must notice
guard_addr = w.device_ptr(0) - 32 * stride_bytes - 16     # inside the guard row in front of the buffer
torch.cuda.synchronize()
# write through a raw pointer with a tiny torch kernel: wrap the address as a tensor
class V:  # __cuda_array_interface__ carrier
    def __init__(s, p): s.__cuda_array_interface__ = {"shape": (2,), "typestr": "<i8", "data": (p, False), "version": 3, "strides": None}
t = torch.as_tensor(V(guard_addr), device="cuda"); t.fill_(0); torch.cuda.synchronize()
try:
    w.sync(); raise SystemExit("sync() did not notice the trampled guard row")
except GolError as e:
    assert "guard row" in str(e), e
print("DETECTOR_OK")
"""
    r = run_py(code)
    assert r.returncode == 0 and "DETECTOR_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_kernel_suites_pass_with_every_access_checked():
    env = dict(os.environ, GOL_B200_LIB="debug")
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                        os.path.join(ROOT, "tests", "test_gpu_property.py"),
                        os.path.join(ROOT, "tests", "test_gpu_wide_oracle.py"),
                        os.path.join(ROOT, "tests", "test_gpu_kernels.py"),
                        "--deselect", "tests/test_gpu_kernels.py::test_slab_larger_than_2_pow_31_cells",
                        "--deselect", "tests/test_gpu_kernels.py::test_native_cli_on_gpu"],
                       capture_output=True, text=True, timeout=1800, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


def test_flag_in_data_path_with_every_access_checked():
    env = dict(os.environ, GOL_B200_LIB="debug", GOL_FORCE_LL="1")
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                        os.path.join(ROOT, "tests", "test_gpu_property.py")],
                       capture_output=True, text=True, timeout=1800, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
