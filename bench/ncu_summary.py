#!/usr/bin/env python
"""Markdown summary of an .ncu-rep (one section per profiled kernel) for profiles/.

    python bench/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_step.md "title"
"""
import csv
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid size"),
    ("launch__block_size", "block size"),
    ("launch__cluster_size", "cluster size"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("sm__cycles_elapsed.max", "SM cycles elapsed (max)"),
    ("smsp__inst_executed.sum", "warp instructions executed"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor pipe instructions"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts"),
]
STALLS = "smsp__average_warps_issue_stalled_{}_per_issue_active.ratio"
STALL_NAMES = ["barrier", "long_scoreboard", "short_scoreboard", "mio_throttle", "wait", "membar", "math_pipe_throttle",
               "no_instruction", "branch_resolving", "not_selected", "lg_throttle", "dispatch_stall"]


def main():
    rep, out, title = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else sys.argv[1])
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    lines = [f"# {title}", "",
             f"Source: `{rep.split('/')[-1]}` captured with `ncu --set full --clock-control none --import-source on` on one "
             "B200 (each kernel is replayed ~40x with caches flushed between passes, so durations here are colder and "
             "longer than in the timed benchmark; use them for shares and stall reasons, not as bench values).", ""]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        name = d.get("Kernel Name", "?")
        lines += [f"## `{name[:110]}`", "", "| metric | value |", "|---|---|"]
        for key, label in METRICS:
            if key in d and d[key] != "":
                lines.append(f"| {label} | {d[key]} {u.get(key, '')} |")
        st = []
        for n in STALL_NAMES:
            k = STALLS.format(n)
            if k in d and d[k] not in ("", "0"):
                st.append((float(d[k]), n))
        st.sort(reverse=True)
        lines += ["", "Warp stall reasons (warps stalled per issue slot, top 6): " +
                  ", ".join(f"{n} {v:.2f}" for v, n in st[:6]), ""]
    src = subprocess.run([sys.executable, "bench/ncu_lines.py", rep, "14"], capture_output=True, text=True).stdout
    lines += ["## Hottest source lines (sampled stalls, both kernels merged)", "", "```", src.strip()[:6000], "```", ""]
    open(out, "w").write("\n".join(lines))
    print("\n".join(lines)[:2500])


if __name__ == "__main__":
    main()
