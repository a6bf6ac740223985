#!/usr/bin/env python3
"""Summarise one `ncu --set full --import-source on` capture (.ncu-rep) as markdown for profiles/.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep "title / command" > profiles/rN_x_summary.md

Sections: headline raw metrics, stall reasons, executed-instruction share per source function (with the
average number of active lanes per issued instruction, i.e. where divergence costs), and the SASS opcode
mix weighted by execution count."""
import collections
import csv
import io
import os
import re
import subprocess
import sys

RAW = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__cycles_active.avg", "smsp__average_warp_latency_per_inst_issued.ratio",
]
FUNC = re.compile(r"^(?:B2C_HD|__global__|static|extern|__device__|__host__|k_\w+\()[^;]*?\b(\w+)\s*\(")
CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                    "risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200", "csrc")


def line_to_function(basename):
    """line number -> enclosing function name, from the repo's copy of the source (column-0 definitions)."""
    path = os.path.join(CSRC, basename)
    out = {}
    if not os.path.exists(path):
        return out
    cur = "(file scope)"
    for n, line in enumerate(open(path), 1):
        m = FUNC.match(line)
        if m and not line.rstrip().endswith(";"):
            cur = m.group(1) if not line.startswith("k_") else line.split("(")[0]
        out[n] = cur
    return out


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, "--csv", *args], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    kid = sys.argv[3] if len(sys.argv) > 3 else None      # which profiled launch (row index), default last
    raw = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw"))))
    h, units, rows = raw[0], raw[1], raw[2:]
    row = rows[int(kid)] if kid is not None else rows[-1]
    print(f"# {title}\n")
    print(f"Kernel `{row[h.index('Kernel Name')]}` (launch {rows.index(row) + 1} of {len(rows)} captured); "
          "numbers read with `ncu -i ... --page raw --csv` / `--page source --csv`.\n")
    print("| metric | value |\n|---|---|")
    for k in RAW:
        if k in h:
            print(f"| {k} | {row[h.index(k)]} {units[h.index(k)]} |")
    print("\n## warp stall reasons (warps stalled per issue-active cycle)\n\n| reason | ratio |\n|---|---|")
    st = []
    for i, k in enumerate(h):
        m = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active.ratio", k)
        if m:
            try:
                st.append((float(row[i]), m.group(1)))
            except ValueError:
                pass
    for v, k in sorted(st, reverse=True)[:8]:
        print(f"| {k} | {v:.3f} |")

    src = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--print-source", "cuda,sass"))))
    fname, func, l2f = None, "?", {}
    by_func = collections.defaultdict(lambda: [0, 0, 0])
    ops = collections.Counter()
    tot = [0, 0, 0]
    col = {}
    seen_kernel = 0
    for r in src:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
            l2f = line_to_function(fname)
            continue
        if len(r) == 2 and r[0] == "Function Name":
            seen_kernel += 1
            continue
        if r and r[0] == "Line No":
            col = {k: i for i, k in enumerate(r)}
            continue
        if not col or len(r) < len(col):
            continue
        i_s, i_i, i_t = col["# Samples"], col["Instructions Executed"], col["Thread Instructions Executed"]
        if r[0] not in ("", "Line No"):                       # a CUDA source line
            try:
                func = l2f.get(int(r[0]), "?")
                s, ie, te = int(r[i_s]), int(r[i_i]), int(r[i_t])
            except ValueError:
                continue
            key = f"{fname}:{func}"
            by_func[key][0] += s; by_func[key][1] += ie; by_func[key][2] += te
            tot[0] += s; tot[1] += ie; tot[2] += te
        else:                                                  # a SASS line under it
            try:
                ie = int(r[i_i])
            except ValueError:
                continue
            op = re.sub(r"^@!?U?P\d+\s+", "", r[3].strip()).split()
            if op:
                ops[op[0].split(".")[0]] += ie
    if tot[1]:
        print(f"\n## executed instructions per source function (all launches of the kernel in the capture)\n\n"
              f"total: {tot[1]:,} warp instructions, {tot[2] / tot[1]:.1f} of 32 lanes active on average, "
              f"{tot[0]:,} stall samples\n\n| function | samples | instructions | active lanes |\n|---|---|---|---|")
        for k, v in sorted(by_func.items(), key=lambda kv: -kv[1][1]):
            if v[1] * 200 < tot[1]:
                continue
            print(f"| {k} | {100 * v[0] / max(tot[0], 1):.1f}% | {100 * v[1] / tot[1]:.1f}% | {v[2] / max(v[1], 1):.1f} |")
        n = sum(ops.values())
        print("\n## SASS opcode mix (weighted by executions)\n\n| opcode | share |\n|---|---|")
        for op, c in ops.most_common(14):
            print(f"| {op} | {100 * c / n:.1f}% |")


if __name__ == "__main__":
    main()
