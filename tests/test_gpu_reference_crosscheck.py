"""Cross-check against the UNMODIFIED reference binary (baseline/_ref/bin/gol_ref, built by
baseline/build_ref.sh against the minimpi shim).

Where the reference is correct (generation 1 at R >= 3, pattern 0,
pattern 3 at R >= 3) its output files must equal ours byte for byte.  Elsewhere its files must
equal what ``reference_quirk_run`` predicts (stale send rows, ghost rows swapped at R <= 2) and
differ from true Life - which pins down both the reference's behaviour and the shim.

Tiny worlds only: as shipped the reference spends ~35-65 us per cell update in device printf.
"""
import os
import subprocess

import numpy as np
import pytest

from gol_b200 import World
from gol_b200 import _native as n
from gol_b200.models.patterns import global_pattern, pattern_bytes
from gol_b200.ops.oracle import reference_quirk_run
from gol_b200.utils.io import read_rank_file, read_world

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "bin", "gol_ref")
MPIRUN = os.path.join(ROOT, "baseline", "_ref", "bin", "mpirun")


def run_reference(tmp, pattern, size, gens, R):
    if not (os.path.exists(REF) and os.path.exists(MPIRUN)):
        pytest.skip("baseline/_ref/bin is not built (run baseline/build_ref.sh)")
    with open(os.path.join(tmp, "stdout.bin"), "wb") as fo:
        p = subprocess.run([MPIRUN, "-n", str(R), REF, str(pattern), str(size), str(gens), "512", "1"],
                           cwd=tmp, stdout=fo, stderr=subprocess.PIPE, timeout=600)
    assert p.returncode == 0, p.stderr[-500:]
    return read_world(tmp, R)


def run_ours(tmp, pattern, size, gens, R):
    w = World(size, size, nranks=R, backend="cuda", ndevices=1)
    w.connect()
    w.init_pattern(pattern)
    w.step(gens)
    w.sync()
    for i in range(R):
        w.dump_text(i, tmp)
    w.close()
    return [open(os.path.join(tmp, f"Rank_{r}_of_{R}.txt"), "rb").read() for r in range(R)]


@pytest.mark.parametrize("pattern,size,gens,R", [(4, 32, 1, 3), (0, 32, 4, 3), (3, 32, 5, 3), (4, 64, 1, 4),
                                                (3, 64, 3, 4)])
def test_files_are_byte_identical_where_the_reference_is_correct(tmp_path, pattern, size, gens, R):
    ref_dir, our_dir = tmp_path / "ref", tmp_path / "ours"
    ref_dir.mkdir()
    our_dir.mkdir()
    ref_world = run_reference(str(ref_dir), pattern, size, gens, R)
    ours = run_ours(str(our_dir), pattern, size, gens, R)
    for r in range(R):
        assert (ref_dir / f"Rank_{r}_of_{R}.txt").read_bytes() == ours[r]
    assert (ref_world == n.oracle_run(global_pattern(pattern, size, size, R), gens)).all()


@pytest.mark.parametrize("pattern,size,gens,R", [(4, 32, 3, 2), (1, 32, 3, 4), (4, 32, 2, 1), (4, 64, 3, 3)])
def test_reference_defects_match_the_quirk_model(tmp_path, pattern, size, gens, R):
    ref_world = run_reference(str(tmp_path), pattern, size, gens, R)
    slabs = [pattern_bytes(pattern, size, size, r, R) for r in range(R)]
    predicted = np.concatenate(reference_quirk_run(slabs, gens))
    assert (ref_world == predicted).all(), "the reference binary does not behave as SURVEY.md 2f describes"
    truth = n.oracle_run(np.concatenate(slabs), gens)
    assert not (ref_world == truth).all(), "expected the reference to diverge from true Life here"
