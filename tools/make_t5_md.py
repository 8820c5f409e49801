"""profiles/r02_t5.md from the two JSON files tools/t5_exp1.py writes (systematic / multinomial resampler)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
out = ["# %s - T5: the reference's published Exp1 numbers on the CUDA path\n" % tag,
       "`python tools/t5_exp1.py --seeds 16 [--resampler 1]` on one B200 (`tools/gpu_final.sh`).  150 RealFlow Exp1 trials",
       "(`tests/golden/Data/RealFlowData/Exp1`, data only), run as `main.jl:24-31,54` runs them (S = 9); judgment = unweighted mean",
       "elasticity of the final particles (`Analysis/utils.jl:102-124`); r = Pearson correlation with the ground truth",
       "(`Analysis/plots.jl:51`).  The reference is one unseeded draw per number; here: mean +- sd over 16 seeds.\n"]
for name, f in (("systematic resampling (default)", "%s_t5_systematic.json" % tag),
                ("multinomial resampling (Gen.maybe_resample!)", "%s_t5_multinomial.json" % tag)):
    d = json.load(open(os.path.join(ROOT, "profiles", f)))
    out += ["## %s\n" % name,
            "| setting | published r | r (mean +- sd) [min, max] | mean bias (estimate - GT) | stimuli above GT |", "|---|---|---|---|---|"]
    for k, v in d.items():
        out.append("| %s | %s | %.3f +- %.3f [%.3f, %.3f] | %+.3f | %.0f %% |" % (
            k, v["published_r"], v["r_mean"], v["r_sd"], v["r_min"], v["r_max"], v["bias_mean"], 100 * v["frac_above_gt"]))
    k = "Sphere_Par50_Rej5"
    out += ["", "Mean estimate per ground-truth level 0.0 ... 0.9 (%s): %s" % (
        k, " ".join("%.2f" % d[k]["judgment_by_gt"]["%.1f" % (i / 10)] for i in range(10))),
        "(read by eye off the reference's scatter plot of the same setting: 0.18 0.33 0.42 0.60 0.62 0.72 0.82 0.85 0.90 0.93)\n"]
out += ["The float64 CPU oracle on the same experiment (`tests/test_oracle_t5.py`, two seeds each): Sphere Par20_Rej1 0.845 / 0.787,",
        "Sphere Par50_Rej5 0.877 / 0.885, Cube Par20_Rej1 0.693 / 0.659."]
open(os.path.join(ROOT, "profiles", "%s_t5.md" % tag), "w").write("\n".join(out) + "\n")
