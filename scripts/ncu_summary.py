#!/usr/bin/env python
"""Condense an .ncu-rep into the few numbers the roofline discussion needs (reads it with `ncu -i`, no GPU).

  python scripts/ncu_summary.py gpurun_out/prof_train2.ncu-rep [--regions 150]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
nreg = int(sys.argv[sys.argv.index("--regions") + 1]) if "--regions" in sys.argv else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, unit = rows[0], rows[1]
all_vals = rows[2:]
val = all_vals[0]
if "--all" in sys.argv:      # one block per captured kernel launch
    keep = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
    for v in all_vals:
        if len(v) != len(hdr):
            continue
        for i, h in enumerate(hdr):
            if h in keep:
                print(f"{h:75s} {v[i][:90]:>30s} {unit[i]}")
        print()
    sys.exit(0)
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warps_issue_stalled"]
for i, h in enumerate(hdr):
    if any(h == w or (w.endswith("stalled") and h.startswith(w)) or (w == "sm__inst_executed_pipe_tensor" and h.startswith(w))
           for w in want):
        try:
            if h.startswith("smsp__average_warps_issue_stalled") and float(val[i]) < 0.3:
                continue
        except ValueError:
            pass
        print(f"{h:95s} {val[i]:>18s} {unit[i]}")
if nreg:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(src)))
    h, data = r[1], r[2:]
    iS, iI = h.index("# Samples"), h.index("Instructions Executed")
    tot = sum(int(x[iS]) for x in data) or 1
    toti = sum(int(x[iI]) for x in data) or 1
    print(f"\nstall samples by SASS region ({nreg} instructions each; total {tot} samples, {toti} warp-instructions)")
    for k in range(0, len(data), nreg):
        seg = data[k:k + nreg]
        s_ = sum(int(x[iS]) for x in seg)
        if s_ * 100 / tot < 1.0:
            continue
        ops = {}
        for x in seg:
            t = x[1].split()
            op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
            ops[op] = ops.get(op, 0) + int(x[iS])
        top = ", ".join(f"{o}:{n}" for o, n in sorted(ops.items(), key=lambda kv: -kv[1])[:4])
        print(f"  sass {k:5d}-{k + nreg:5d}: {100 * s_ / tot:5.1f}% of samples, {100 * sum(int(x[iI]) for x in seg) / toti:5.1f}% of instructions   [{top}]")
