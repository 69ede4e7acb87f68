#!/usr/bin/env python
"""Condense an `ncu --set full` report into the handful of numbers DESIGN.md and bench.py quote.

    tools/ncu_summary.py report.ncu-rep [label] > profiles/rNN_<kernel>.txt

Reads the report with `ncu -i ... --page raw --csv` (no GPU needed) and prints, per captured kernel:
duration, DRAM bytes read/written, DRAM throughput, issue-slot utilisation, instruction count,
registers, shared memory, occupancy and the leading warp-stall reasons. With --json the DRAM bytes are
emitted in the layout of profiles/traffic.json.
"""
import csv
import io
import json
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
    ("sm__inst_executed.sum", "warp instructions"),
    ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor memory (TMEM) active %"),
    ("sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active", "tcgen05.ld/st pipe %"),
    ("smsp__cycles_active.avg", "SMSP active cycles"),
]


def to_float(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return None


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    as_json = "--json" in sys.argv
    rep = args[0]
    label = args[1] if len(args) > 1 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(hdr)}
    traffic = {}
    if not as_json:
        print("# %s  (ncu --set full --clock-control none; cold-cache, serialised replays)" % label)
    for r in data:
        name = r[col["Kernel Name"]].split("(")[0]
        if not as_json:
            print("\n## %s" % name[:150])
        for key, what in WANT:
            if key in col and r[col[key]] != "":
                if not as_json:
                    print("  %-26s %s %s" % (what, r[col[key]], units[col[key]]))
        rd, wr = (to_float(r[col[k]]) if k in col else None for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        if rd is not None and wr is not None:
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            rd *= scale.get(units[col["dram__bytes_read.sum"]], 1)
            wr *= scale.get(units[col["dram__bytes_write.sum"]], 1)
            traffic[name] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": rd + wr}
            if not as_json:
                print("  %-26s %.1f MB" % ("dram read+write", (rd + wr) / 1e6))
        stalls = []
        for k, i in col.items():
            if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") or \
               k.startswith("smsp__average_warp_latency_issue_stalled_") and k.endswith(".ratio"):
                v = to_float(r[i])
                if v:
                    stalls.append((v, k.split("stalled_")[1].replace("_per_issue_active.ratio", "").replace(".ratio", "")))
        if stalls and not as_json:
            stalls.sort(reverse=True)
            print("  top stalls (warps per issue): " + ", ".join("%s %.2f" % (n, v) for v, n in stalls[:6]))
    if as_json:
        print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
