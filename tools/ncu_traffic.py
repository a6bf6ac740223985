#!/usr/bin/env python3
"""DRAM traffic per launch of the dominant kernel from one `ncu --set full` capture -> profiles/<tag>_slots_traffic.json,
stamped with the hash of the kernel sources it was taken from (bench.py drops the figure when the hash differs).

    python tools/ncu_traffic.py gpurun_out/x.ncu-rep profiles/r2_slots_traffic.json [instances_per_launch]"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rep, out = sys.argv[1], sys.argv[2]
    B = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h, units = rows[0], rows[1]
    recs = []
    for v in rows[2:]:
        name = v[h.index("Kernel Name")]
        if "k_mpc_slots" not in name and "k_mpc<" not in name:
            continue
        def val(metric):
            i = h.index(metric)
            x, u = float(v[i]), units[i].lower()
            return x * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        recs.append(dict(kernel=name, dram_read=val("dram__bytes_read.sum"), dram_write=val("dram__bytes_write.sum"),
                         duration_us=float(v[h.index("gpu__time_duration.sum")])))
    assert recs, "no K3 launch in the capture"
    r = recs[-1]
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    d = {"kernel": r["kernel"], "instances_per_launch": B, "dram_bytes_read": r["dram_read"], "dram_bytes_write": r["dram_write"],
         "dram_bytes_per_launch": r["dram_read"] + r["dram_write"], "algorithmic_bytes_per_launch": 1000.0 * B,
         "ratio_to_algorithmic": (r["dram_read"] + r["dram_write"]) / (1000.0 * B), "duration_us_under_ncu": r["duration_us"],
         "capture": os.path.basename(rep), "kernel_src_sha": bench.kernel_src_sha(), "git_head_at_capture": head,
         "how": "ncu --set full --clock-control none --import-source on -k regex:k_mpc_slots -s 4 -c 1 python bench.py --steps 3 "
                "--warmup 3 --no-cpu --no-secondary --no-config5 (dram__bytes_read.sum + dram__bytes_write.sum)"}
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d))


if __name__ == "__main__":
    main()
