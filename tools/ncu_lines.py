"""Per-source-line summary of an ncu `--page source --csv --print-source cuda,sass` dump (run here, no GPU).
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > dump.csv; python tools/ncu_lines.py dump.csv [kernel-substring]"""
import csv
import sys
from collections import defaultdict

csv.field_size_limit(10 ** 9)


def main(path, want=None, top=45):
    cur_file = cur_fn = None
    hdr = None
    agg = defaultdict(lambda: [0, 0, 0, ""])  # (file, line) -> [samples, warp inst, thread inst, text]
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r[0] == "Function Name":
            cur_fn = r[1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0] and hdr and (want is None or want in (cur_fn or "")):
            try:
                ln = int(r[0])
            except ValueError:
                continue
            d = dict(zip(hdr[4:], r[4:]))
            a = agg[(cur_file, ln)]
            a[0] += int(d.get("# Samples", "0") or 0)
            a[1] += int(d.get("Instructions Executed", "0") or 0)
            a[2] += int(d.get("Thread Instructions Executed", "0") or 0)
            a[3] = r[1].strip()[:90]
    ts = sum(a[0] for a in agg.values()) or 1
    ti = sum(a[1] for a in agg.values()) or 1
    print("total samples %d, warp instructions %d" % (ts, ti))
    print("%-14s %5s %7s %7s %5s  %s" % ("file", "line", "smp%", "inst%", "thr", "source"))
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%-14s %5d %7.2f %7.2f %5.1f  %s" % (f, ln, 100.0 * a[0] / ts, 100.0 * a[1] / ti,
                                                    a[2] / a[1] if a[1] else 0, a[3]))
    return agg


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
