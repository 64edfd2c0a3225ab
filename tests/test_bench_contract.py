"""bench.py contract pieces that can be checked without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line_and_exits_zero_without_a_gpu():
    env = dict(os.environ, GOL_REF_BUDGET_S="5")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1"],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference"
    # on a CPU-only box the reference cannot run: the arm must say so instead of failing
    assert "unavailable" in line or "value" in line


def test_reference_arm_does_not_load_the_product_library():
    """The reference arm must not map libgol_b200.so (the clock sampler import dragged
    the package in).  Run its Python part in a process with an import hook on gol_b200."""
    code = (
        "import sys, os, json\n"
        "sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1']\n"
        "os.environ['GOL_REF_BUDGET_S'] = '5'\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "import runpy\n"
        "try:\n"
        "    runpy.run_path(os.path.join(%r, 'bench.py'), run_name='__main__')\n"
        "except SystemExit:\n"
        "    pass\n"
        "maps = open('/proc/self/maps').read()\n"
        "print('LOADED' if 'libgol_b200' in maps or any(m.startswith('gol_b200') for m in sys.modules) else 'CLEAN')\n"
    ) % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.stdout.strip().splitlines()[-1] == "CLEAN", r.stdout[-500:] + r.stderr[-500:]


def test_strong_scaling_geometry_and_labels():
    sys.path.insert(0, ROOT)
    import importlib

    bench = importlib.import_module("bench")

    class A:
        scaling, size, rows = "strong", 0, 0

    assert bench.geometry(A, 8) == (65536, 8192, "strong")
    A.scaling = "weak"
    assert bench.geometry(A, 8) == (16384, 16384, "weak")


def test_bench_help_lists_the_driver_flags():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True)
    assert r.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl", "--scaling"):
        assert flag in r.stdout


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g

    assert callable(g.build) and callable(g.smoke)
