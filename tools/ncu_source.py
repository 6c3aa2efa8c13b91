#!/usr/bin/env python
"""Per-source-line instruction counts of a kernel from an .ncu-rep captured with
--import-source on (binary compiled with -lineinfo).

    python tools/ncu_source.py gpurun_out/r2_fold.ncu-rep [top N]

Prints the N source lines that executed the most warp instructions, with their
share and stall samples, then the SASS mnemonic histogram."""
import csv
import io
import re
import subprocess
import sys
from collections import Counter


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    hdr = rows[hdr_i]
    inst_c, samp_c, addr_c = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Address")
    thr_c = hdr.index("Thread Instructions Executed")
    total = 0
    ops, per_line, stall = Counter(), Counter(), Counter()
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr):
            continue
        try:
            n = int(r[inst_c])
        except ValueError:
            continue
        if r[addr_c] == "-":  # a CUDA source line: the sum of its SASS rows
            key = f"{r[0]:>5s}: {r[1].strip()}"
            per_line[key] += n
            stall[key] += int(r[samp_c] or 0)
        else:                 # a SASS instruction
            total += n
            m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[addr_c + 1].strip())
            if m:
                ops[m.group(2).split(".")[0]] += n
    print(f"# {rep}: {total} warp instructions")
    print("\n## mnemonics")
    for op, n in ops.most_common(25):
        print(f"  {op:12s} {n:12d} {100.0 * n / total:5.1f} %")
    print("\n## source lines (warp instructions, share, share of stall samples)")
    tot_s = sum(stall.values()) or 1
    for key, n in per_line.most_common(top):
        print(f"  {n:12d} {100.0 * n / total:5.1f} %  stall {100.0 * stall[key] / tot_s:5.1f} %  {key[:120]}")


if __name__ == "__main__":
    main()
