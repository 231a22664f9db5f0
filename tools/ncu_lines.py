#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > X.csv`.

  python tools/ncu_lines.py X.csv [min_percent]
"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
    fname, hdr, out = "", None, []
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and r and r[0].isdigit():
            ie = hdr.index("Instructions Executed")
            ws = hdr.index("Warp Stall Sampling (All Samples)")
            te = hdr.index("Thread Instructions Executed")
            try:
                out.append((int(r[ie]), int(r[ws] or 0), int(r[te]), fname, r[0], r[1]))
            except ValueError:
                pass
    tot = sum(o[0] for o in out) or 1
    tots = sum(o[1] for o in out) or 1
    print(f"total warp instructions {tot}, stall samples {tots}")
    for n, w, t, f, l, s in out:
        if n > tot * thr / 100 or w > tots * thr / 100:
            print(f"{n / tot * 100:5.1f}% inst {w / tots * 100:5.1f}% stall  lanes {t / max(n, 1):4.1f}  {f}:{l}: {s.strip()[:100]}")


if __name__ == "__main__":
    main()
