"""Summarise an `ncu --set full` report into the CSV + level-traffic JSON kept under profiles/.

usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/r1_ncu_full_c2_b512.csv [profiles/r1_level_traffic.json]

One row per launch, in launch order. Times are cold-cache and serialised (ncu replays each launch), so they
are only good for shares; the DRAM byte counts are what bench.py quotes as `roofline.traffic`.
"""
import csv
import io
import json
import subprocess
import sys

COLS = [
    ("time_ms", "gpu__time_duration.sum", 1e-6),            # ns -> ms (ncu may print us/ms: unit handled below)
    ("dram_read_MB", "dram__bytes_read.sum", 1e-6),
    ("dram_write_MB", "dram__bytes_write.sum", 1e-6),
    ("warp_inst", "smsp__inst_executed.sum", 1.0),
    ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    ("alu_pipe_pct", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1.0),
    ("fma_pipe_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", 1.0),
    ("lsu_pipe_pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", 1.0),
    ("tensor_pipe_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    ("regs", "launch__registers_per_thread", 1.0),
    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("smem_bank_conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", 1.0),
]
UNIT_SCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
              "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def main():
    rep, out_csv = sys.argv[1], sys.argv[2]
    out_json = sys.argv[3] if len(sys.argv) > 3 else None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(hdr)}
    kname = col["Kernel Name"]
    out, level_rd, level_wr, level_inst, n_level = [], 0.0, 0.0, 0.0, 0
    i_inst = [c[0] for c in COLS].index("warp_inst")
    for r in body:
        name = r[kname]
        short = name.split("(")[0].replace("void ", "")
        vals = []
        for label, metric, scale in COLS:
            if metric not in col or r[col[metric]] in ("", "n/a"):
                vals.append("")
                continue
            v = float(r[col[metric]].replace(",", ""))
            u = units[col[metric]]
            if label == "time_ms" or label.endswith("_MB"):
                v *= UNIT_SCALE.get(u, scale)
            vals.append(v)
        if "vs_level_kernel" in name:
            short = "vs_level_kernel/level%d" % (n_level % 8)
            n_level += 1
            if n_level <= 8:
                level_rd += vals[1] * 1e6
                level_wr += vals[2] * 1e6
                level_inst += float(vals[i_inst] or 0)
        out.append([short] + [("%.4g" % v if isinstance(v, float) else v) for v in vals])
    with open(out_csv, "w") as f:
        f.write("# ncu --set full --clock-control none, bench.py --steps 2 --warmup 1 --no-cpu (C2, 512 frames/step)\n")
        f.write("# one row per launch; times are cold-cache/serialised (use shares, not absolutes)\n")
        f.write(",".join(["kernel"] + [c[0] for c in COLS]) + "\n")
        for r in out:
            f.write(",".join(str(x) for x in r) + "\n")
    if out_json and n_level >= 8:
        json.dump({"workload": "c2", "frames_per_step": 512, "kernel": "vs_level_kernel", "launches": 8,
                   "dram_read_bytes": int(level_rd), "dram_write_bytes": int(level_wr),
                   "warp_instructions": int(level_inst),
                   "source": "%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum summed over the 8 "
                             "level launches of one step)" % out_csv}, open(out_json, "w"), indent=1)
    print("wrote", out_csv, len(out), "launches")


if __name__ == "__main__":
    main()
