#!/usr/bin/env python3
"""Per CUDA source line: warp-stall samples and warp instructions of one kernel of an .ncu-rep (needs -lineinfo and --import-source on).

    python tools/ncu_lines.py report.ncu-rep kernel_name [top]
"""
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern], capture_output=True, text=True).stdout
    allrows = list(csv.reader(io.StringIO(out)))
    rows = [r for r in allrows if len(r) > 10 or (len(r) == 2 and r[0] == "File Name")]
    hdr = [r for r in rows if len(r) > 10][0]
    i_s, i_e = hdr.index("# Samples"), hdr.index("Instructions Executed")
    lines = []
    files = set(); fname = "?"
    for r in rows:
        if r[0] == "File Name":
            fname = r[1].split("/")[-1]
            if fname in files:
                break                                  # first launch of that kernel only (its files come round again for the next)
            files.add(fname)
            continue
        if r[0] == "Line No":
            continue
        if r[0] == "":
            continue                                   # SASS row under a source line
        try:
            lines.append((int(r[i_s] or 0), int(r[i_e] or 0), fname[5:12] + ":" + r[0], r[1].strip()))
        except ValueError:
            pass
    tot = sum(l[0] for l in lines) or 1
    print(f"== {kern}: {tot} samples, {sum(l[1] for l in lines)} warp instructions")
    for s, e, no, src in sorted(lines, key=lambda l: -l[0])[:top]:
        print(f"  {100.0 * s / tot:5.1f}%  {e:10d}  {no:>13}  {src[:140]}")


if __name__ == "__main__":
    main()
