#!/usr/bin/env python
"""Summarise an .ncu-rep source page per CUDA source line: samples, instructions, top stall reasons.

    python bench/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]
"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur_file, hdr = None, None
    lines = []
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and len(r) == len(hdr) and r[0] != "":
            d = dict(zip(hdr, r))
            try:
                samples = int(d["# Samples"])
                inst = int(d["Instructions Executed"])
            except ValueError:
                continue
            stalls = {k[6:]: int(v) for k, v in d.items()
                      if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v) > 0}
            bank = d.get("L1 Wavefronts Shared Excessive", "0")
            lines.append((samples, inst, cur_file, d["Line No"], r[1].strip()[:90], stalls, bank))
    tot_s = sum(x[0] for x in lines) or 1
    tot_i = sum(x[1] for x in lines) or 1
    print(f"total samples {tot_s}, total warp-instructions {tot_i}")
    for s, i, f, ln, src, st, bank in sorted(lines, key=lambda x: (x[0], x[1]), reverse=True)[:top]:
        tops = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{100*s/tot_s:5.1f}% smp {100*i/tot_i:5.1f}% ins  {f}:{ln:>4}  bankx={bank:>6} [{tops}]  {src}")


if __name__ == "__main__":
    main()
