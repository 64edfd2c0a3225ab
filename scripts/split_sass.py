#!/usr/bin/env python
"""Per-kernel SASS listings for profiles/sass/ (cuobjdump -sass of the built library, no GPU needed).

    python scripts/split_sass.py            # writes profiles/sass/<kernel>.sass + profiles/sass/INDEX.md

Templated step kernels are kept for K = 4 (the default) and K = 1 only; every other kernel is kept.
INDEX.md counts the mnemonics that matter per kernel (LOP3 / SHF / SHFL / LDG / STG / LDS / STS /
UBLKCP / UTMALDG / SYNCS / RED / ATOM / MEMBAR / UCGABAR).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "game-of-life---mpi-cuda_b200", "_lib", "libgol_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
COLS = ["LOP3", "SHF", "SHFL", "LDG", "STG", "LD", "ST", "LDS", "STS", "UBLKCP", "UTMALDG", "SYNCS", "RED", "ATOMG",
        "MEMBAR", "UCGABAR_ARV", "BAR"]


def short(mangled: str) -> str:
    t = re.search(r"(life_\w+?_kernel)I((?:Li\d+E)+)", mangled)
    if t:
        args = re.findall(r"Li(\d+)E", t.group(2))
        return f"{t.group(1)}_K{args[0]}" + ("_" + "_".join(args[1:]) if len(args) > 1 else "")
    m = re.search(r"_cu_[0-9a-f]{8}(\d+)", mangled)
    if m:
        n = int(m.group(1))
        return mangled[m.end():m.end() + n]
    return re.sub(r"\W", "_", mangled)[:60]


def main() -> int:
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    rows = []
    for part in parts[1:]:
        mangled, _, body = part.partition("\n")
        name = short(mangled.strip())
        k = re.search(r"_K(\d+)", name)
        if k and k.group(1) not in ("1", "4"):
            continue
        ins = []
        for line in body.splitlines():
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?)\s*;", line)
            if m:
                ins.append(f"{m.group(1)}  {m.group(2)} ;")
        with open(os.path.join(OUT, name + ".sass"), "w") as f:
            f.write(f"// {mangled.strip()}\n// {len(ins)} instructions, sm_100a, from libgol_b200.so\n")
            f.write("\n".join(ins) + "\n")
        c = collections.Counter()
        for i in ins:
            op = i.split()[1]
            if op.startswith("@"):
                op = i.split()[2]
            c[op.split(".")[0]] += 1
        rows.append((name, len(ins), c))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# SASS listings (cuobjdump -sass libgol_b200.so, sm_100a)\n\n")
        f.write("One file per kernel (`<kernel>.sass`); templated step kernels for K = 1 and K = 4.\n\n")
        f.write("| kernel | instrs | " + " | ".join(COLS) + " |\n|---|---|" + "---|" * len(COLS) + "\n")
        for name, n, c in sorted(rows):
            f.write(f"| `{name}` | {n} | " + " | ".join(str(c.get(k, 0)) for k in COLS) + " |\n")
    print(f"wrote {len(rows)} listings to {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
