"""Per-source-line instruction and stall-sample shares from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:NAME > lines.csv`.
usage: python profiles/ncu_lines.py lines.csv [top_n]"""
import csv
import sys


def main(path, top=45):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hdr_i]
    ii = hdr.index("Instructions Executed")
    si = hdr.index("# Samples")
    lines = []
    for r in rows[hdr_i + 1:]:
        if not r or r[0] == "" or len(r) <= ii:
            continue
        try:
            lines.append((int(r[0]), r[1], int(r[ii]), int(r[si])))
        except ValueError:
            pass
    ti = sum(x[2] for x in lines) or 1
    ts = sum(x[3] for x in lines) or 1
    print(f"total warp instructions {ti}, samples {ts}")
    print(" inst%  samp%  line  source")
    for ln, src, n, s in sorted(lines, key=lambda x: -x[3])[:top]:
        print(f"{100 * n / ti:6.2f} {100 * s / ts:6.2f} {ln:5d}  {src.strip()[:120]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
