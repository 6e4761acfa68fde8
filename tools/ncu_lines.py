"""Aggregate an `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` dump by CUDA source line.
usage: python tools/ncu_lines.py dump.csv [top]"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out, tot, cur, hdr = [], 0, None, None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0] == "Function Name":
            continue
        elif hdr and r[2] == "-" and r[0]:
            si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
            try:
                n = int(r[si])
            except ValueError:
                continue
            out.append((n, cur, r[0], r[1].strip()[:100], r[ie]))
            tot += n
    print("total samples", tot)
    byfile = {}
    for n, f, *_ in out:
        byfile[f] = byfile.get(f, 0) + n
    for f, n in sorted(byfile.items(), key=lambda kv: -kv[1]):
        print("  %5.1f%%  %s" % (100.0 * n / tot, f))
    for n, f, l, src, ie in sorted(out, reverse=True)[:top]:
        print("%5.1f%% %s:%s  [%s] %s" % (100.0 * n / tot, f, l, ie, src))


if __name__ == "__main__":
    main()
