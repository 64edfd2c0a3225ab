#!/usr/bin/env python
"""Turn an .ncu-rep into the short markdown summary committed under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep "title" > profiles/ncu_xxx.md
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "kernel duration"),
    ("launch__grid_size", "grid (CTAs)"),
    ("launch__block_size", "block (threads)"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_static", "static smem / CTA"),
    ("launch__waves_per_multiprocessor", "waves / SM"),
    ("launch__occupancy_limit_registers", "occupancy limit (registers), CTAs/SM"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy (% of 64 warps)"),
    ("gpc__cycles_elapsed.avg.per_second", "GPC clock"),
    ("sm__cycles_elapsed.max", "SM cycles elapsed"),
    ("sm__cycles_active.avg", "SM cycles active (avg)"),
    ("smsp__inst_executed.sum", "warp instructions executed"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy (% while active)"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "ALU pipe busy (% while active)"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "ALU pipe busy (% of elapsed)"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe busy (% while active)"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe busy (% while active)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe busy"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput (% of peak)"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput (% of peak)"),
    ("dram__bytes_read.sum", "DRAM bytes read"),
    ("dram__bytes_write.sum", "DRAM bytes written"),
    ("lts__t_bytes.sum", "L2 bytes moved"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("smsp__warps_eligible.avg.per_cycle_active", "eligible warps / scheduler / cycle"),
]


def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout


TABLE = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "ALU %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
]


def table_of_all(rep, title):
    """--all: one row per distinct kernel of the report (first launch of each)."""
    raw = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units = raw[0], raw[1]
    print(f"# {title}\n")
    print(f"Source: `{rep.split('/')[-1]}` (ncu --set full --clock-control none), B200, one GPU; first launch of each kernel.\n")
    print("| kernel | " + " | ".join(l for _, l in TABLE) + " |")
    print("|---|" + "---|" * len(TABLE))
    seen = set()
    ik = hdr.index("Kernel Name")
    for row in raw[2:]:
        name = row[ik].split("(")[0].split("::")[-1]
        if name in seen:
            continue
        seen.add(name)
        cells = []
        for key, _ in TABLE:
            if key in hdr:
                i = hdr.index(key)
                v = row[i]
                try:
                    v = f"{float(v):.4g}"
                except ValueError:
                    pass
                cells.append(f"{v} {units[i]}".strip())
            else:
                cells.append("-")
        print(f"| `{name}` | " + " | ".join(cells) + " |")


def main():
    if len(sys.argv) > 3 and sys.argv[3] == "--all":
        return table_of_all(sys.argv[1], sys.argv[2])
    rep, title = sys.argv[1], sys.argv[2]
    raw = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units = raw[0], raw[1]
    print(f"# {title}\n")
    print(f"Source: `{rep.split('/')[-1]}` (ncu --set full --clock-control none --import-source on), B200, one GPU.\n")
    for row in raw[2:3]:
        print(f"Kernel: `{row[hdr.index('Kernel Name')]}`\n")
        print("| metric | value |\n|---|---|")
        for key, label in KEYS:
            if key in hdr:
                i = hdr.index(key)
                print(f"| {label} (`{key}`) | {row[i]} {units[i]} |")
        print("\nWarp stall reasons (average warps stalled per issue-active cycle, > 0.05):\n")
        print("| reason | ratio |\n|---|---|")
        for i, h in enumerate(hdr):
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    v = float(row[i])
                except ValueError:
                    continue
                if v > 0.05:
                    print(f"| {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} | {v:.3f} |")
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv"]))))
    h2, data = None, []
    for r in src:
        if len(r) > 4 and r[0] == "Address":
            if h2 is not None:
                break
            h2 = r
            continue
        if h2 and len(r) == len(h2):
            data.append(r)
    if h2:
        i_src, i_ex = h2.index("Source"), h2.index("Instructions Executed")
        i_st = h2.index("Warp Stall Sampling (All Samples)")
        mix, st = collections.Counter(), collections.Counter()
        for r in data:
            toks = [t for t in r[i_src].strip().split() if not t.startswith("@")]
            if not toks:
                continue
            op = toks[0].split(".")[0]
            mix[op] += int(r[i_ex] or 0)
            st[op] += int(r[i_st] or 0)
        tot = sum(mix.values()) or 1
        print("\nExecuted SASS instruction mix (warp-level, top 12):\n")
        print("| opcode | executed | share | stall samples |\n|---|---|---|---|")
        for op, n in mix.most_common(12):
            print(f"| {op} | {n} | {100.0 * n / tot:.1f} % | {st[op]} |")
        evid = {k: mix[k] for k in ("UTMALDG", "UBLKCP", "SYNCS", "LDS", "LDG", "STG", "SHFL", "LOP3", "SHF", "RED", "ATOM") if mix.get(k)}
        print(f"\nMemory / sync opcodes executed: {evid}")


if __name__ == "__main__":
    main()
