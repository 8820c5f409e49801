"""Turn one tools/gpu_check.sh run (gpurun_out/*_<tag>.*) into tracked summaries under profiles/ (run here, no GPU).
usage: python tools/make_profile_summary.py <tag> "<one-line description of the build>" """
import collections
import contextlib
import csv
import gzip
import io
import os
import shutil
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_lines  # noqa: E402
import ncu_summary  # noqa: E402


def launches(tag, desc):
    path = "gpurun_out/launches_%s.csv" % tag
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, mi, ui = (hdr.index(x) for x in ("Kernel Name", "Metric Value", "Metric Name", "Metric Unit"))
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if r[mi] != "gpu__time_duration.sum":
            continue
        v = float(r[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[r[ui]]
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open("profiles/%s_launches_summary.md" % tag, "w") as f:
        f.write("# %s — ncu launch list (%s)\n\n" % (tag, desc))
        f.write("Command: `ncu --metrics gpu__time_duration.sum --clock-control none -s 2900 -c 1000 --csv python bench.py "
                "--steps 1 --warmup 3 --no-cpu-baseline` (1000 launches ~ one full run_smc of 2^20 sphere particles x 120 "
                "steps, R=5).  Per-launch times are cold-cache and serialised: compare shares.  Full list: "
                "`profiles/%s_launches.csv.gz`.\n\n| kernel | launches | total ms | share |\n|---|---|---|---|\n" % tag)
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| %s | %d | %.2f | %.3f |\n" % (k, a[0], a[1], a[1] / tot))
        try:
            import json
            b = json.load(open("gpurun_out/bench_%s.json" % tag))
            f.write("\nbench.py (default command, no ncu, CUDA events): %.1f M particle-steps/s, %.1f ms per run_smc; "
                    "share_of_step %s\n" % (b["value"] / 1e6, b["ms_per_step"],
                                            {k: round(v, 3) for k, v in b["roofline"]["share_of_step"].items()}))
        except (OSError, ValueError, KeyError):
            pass
    with open(path, "rb") as src, gzip.open("profiles/%s_launches.csv.gz" % tag, "wb") as dst:
        shutil.copyfileobj(src, dst)
    for name in ("bench_%s.json" % tag, "bench_ref_%s.json" % tag):
        if os.path.exists("gpurun_out/" + name):
            shutil.copy("gpurun_out/" + name, "profiles/" + name.replace("bench_ref_%s" % tag, "%s_bench_reference" % tag)
                        .replace("bench_%s" % tag, "%s_bench" % tag))


def full(tag, desc, rep=None):
    rep = rep or "gpurun_out/rejuv_%s.ncu-rep" % tag
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open("/tmp/_raw.csv", "w").write(raw)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    open("/tmp/_cs.csv", "w").write(src)
    b = io.StringIO()
    with contextlib.redirect_stdout(b):
        print("# %s — `ncu --set full` of k_rejuvenate<sphere> (%s)\n" % (tag, desc))
        print("Command: `ncu --set full --clock-control none --import-source on -k regex:k_rejuvenate -s 60 -c 2 python "
              "bench.py --steps 1 --warmup 3 --no-cpu-baseline` (PF steps t = 60, 61 of the first run; 2^20 sphere "
              "particles, R = 5, S = 5).\n\n```")
        ncu_summary.main("/tmp/_raw.csv")
        print("```\n\nHottest source lines (stall samples / executed warp instructions / mean active threads):\n\n```")
        ncu_lines.main("/tmp/_cs.csv", None, 30)
        print("```")
    open("profiles/%s_k_rejuvenate_ncu.md" % tag, "w").write(b.getvalue())


if __name__ == "__main__":
    tag, desc = sys.argv[1], sys.argv[2]
    os.makedirs("profiles", exist_ok=True)
    if os.path.exists("gpurun_out/launches_%s.csv" % tag):
        launches(tag, desc)
    if os.path.exists("gpurun_out/rejuv_%s.ncu-rep" % tag):
        full(tag, desc)
