"""profiles/*.md from an ncu report's raw + source pages: key metrics, per-function and per-line breakdown.
usage: python tools/make_ncu_md.py <raw.csv> <src.csv> <out.md> <title> <notes-file or ''>"""
import contextlib
import io
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_lines
import ncu_summary

raw, src, out, title = sys.argv[1:5]
notes = open(sys.argv[5]).read() if len(sys.argv) > 5 and sys.argv[5] else ""
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    ncu_summary.main(raw)
summary = buf.getvalue()
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    agg = ncu_lines.main(src, None, 32)
lines = buf.getvalue()
code = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "smc_physics_b200", "csrc", "physics.cuh")).read().split("\n")


def fn_of(line):
    for k in range(line - 1, 0, -1):
        t = code[k - 1]
        if "__device__" in t and "(" in t and not t.strip().startswith("//"):
            return t.split("(")[0].split()[-1]
    return "?"


reg = {}
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
for (f, ln), a in agg.items():
    key = "physics.cuh:" + fn_of(ln) if f == "physics.cuh" else f
    r = reg.setdefault(key, [0, 0, 0])
    r[0] += a[0]; r[1] += a[1]; r[2] += a[2]
table = "".join("%-44s smp %5.1f%%  inst %5.1f%%  lanes %4.1f\n" % (k, 100 * r[0] / ts, 100 * r[1] / ti, r[2] / max(r[1], 1))
                for k, r in sorted(reg.items(), key=lambda kv: -kv[1][0]) if r[0] / ts > 0.004)
open(out, "w").write("# %s\n\n%s\n```\n%s```\n\nBy function (stall samples / executed warp instructions / mean active lanes):\n\n```\n%s```\n\n"
                     "Hottest source lines:\n\n```\n%s```\n" % (title, notes, summary, table, lines))
