"""Byte-exact output contract (gol-main.c:17-28,64-73,124-139)."""
import os
import subprocess

import numpy as np

from gol_b200 import _native as n

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = "#" * 25 + " FINAL WORLD IN RANK %d IS " + "#" * 31 + "\n"


def expected_text(cells, rank):
    H, W = cells.shape
    out = [HEADER % rank]
    for i in range(H):
        out.append("Row %2d: " % (H * rank + i) + "".join("%u " % v for v in cells[i]) + "\n")
    return "".join(out)


def test_rank_file_golden(tmp_path):
    rng = np.random.default_rng(0)
    cells = (rng.random((12, 7)) < 0.5).astype(np.uint8)
    for rank in (0, 1, 9):
        p = tmp_path / n.rank_file_name(rank, 10)
        n.write_rank_file(str(p), cells, rank)
        assert p.read_text() == expected_text(cells, rank)
    assert n.rank_file_name(3, 8) == "Rank_3_of_8.txt"


def test_row_label_is_global_and_padded_to_two():
    cells = np.zeros((4, 2), dtype=np.uint8)
    cells[0, 0] = 1
    txt = expected_text(cells, 3)
    assert txt.splitlines()[1] == "Row 12: 1 0 "
    assert expected_text(cells, 0).splitlines()[1] == "Row  0: 1 0 "


def test_report_lines():
    s = n.format_report(1.5, 4, 64, 64, 10)
    assert s == ("TOTAL DURATION : 1.50000, number of cell updates = 163840\n"
                 "This is the Game of Life running in parallel on a GPU on multiple ranks.\n")


def test_cpu_world_dump_matches_writer(tmp_path):
    from gol_b200 import World

    w = World(70, 9, nranks=2, backend="cpu")
    w.init_pattern(6, seed=5, density=0.4)
    w.step(3)
    for i in range(2):
        w.dump_text(i, str(tmp_path))
        got = (tmp_path / f"Rank_{i}_of_2.txt").read_text()
        assert got == expected_text(w.read_bytes(i), i)


def test_native_cli_usage_and_errors(tmp_path):
    exe = os.path.join(ROOT, "bin", "gol")
    r = subprocess.run([exe, "1", "2", "3"], capture_output=True, text=True)
    assert r.returncode == 255
    assert r.stdout == ("GOL requires 5 arguments: pattern number, sq size of the world and the number of "
                        "itterations, threads per block and output-on-off e.g. ./gol 0 32 2 512 0 \n")
    r = subprocess.run([exe, "9", "32", "2", "512", "0"], capture_output=True, text=True)
    assert r.returncode == 255 and r.stdout == "Pattern 9 has not been implemented \n"
    r = subprocess.run([exe, "4", "64", "16", "256", "1", "--cpu", "--ranks", "2", "--check",
                        "--outdir", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    assert lines[0].startswith("TOTAL DURATION : ") and lines[0].endswith("number of cell updates = 131072")
    assert lines[1] == "This is the Game of Life running in parallel on a GPU on multiple ranks."
    assert "CHECK: OK" in r.stdout
    assert sorted(os.listdir(tmp_path)) == ["Rank_0_of_2.txt", "Rank_1_of_2.txt"]


def test_python_cli_usage():
    from gol_b200 import cli

    assert cli.main(["1", "2"]) == 255
    assert cli.main(["8", "32", "2", "512", "0"]) == 255


def test_rank_file_reader_round_trip(tmp_path):
    from gol_b200.utils.io import read_rank_file, read_world

    rng = np.random.default_rng(1)
    world = (rng.random((12, 9)) < 0.5).astype(np.uint8)
    for r in range(3):
        n.write_rank_file(str(tmp_path / n.rank_file_name(r, 3)), world[4 * r:4 * r + 4], r)
    rank, first, cells = read_rank_file(str(tmp_path / "Rank_2_of_3.txt"))
    assert (rank, first) == (2, 8) and (cells == world[8:]).all()
    assert (read_world(str(tmp_path), 3) == world).all()
