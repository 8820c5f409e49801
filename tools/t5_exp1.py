"""T5 anchor: the reference's only published numbers, reproduced on the CUDA path.

The reference prints Pearson r between the per-stimulus model judgment and the ground-truth elasticity on Exp1
(150 RealFlow trials: Ela0..Ela9 x Var1..15; Analysis/plots.jl:51).  The judgment is the UNWEIGHTED mean of the final
particles' elasticities (Analysis/utils.jl:102-124), the ground truth is the `Ela<k>` token of the file name / the
Elasticity column of the init CSV (k / 10), the driver is main.jl:24-31,54 with S = 9 substeps (step_dur 1/30,
Inference/run_inference.jl:40).  Published (plots under Analysis/Exp1/Modelv5/...):
    Sphere  Par50_Rej5 0.869   Par50_Rej1 0.861   Par20_Rej5 0.847   Par20_Rej1 0.82   Par100_Rej1 0.869
    Cube    Par20_Rej1 0.649
all with the model's estimates biased ABOVE the ground truth.  The reference seeds nothing (one unseeded draw per
number), so the comparison is with the seed average of r.

Run on the GPU box:  python tools/t5_exp1.py [--seeds 16] [--resampler 0|1] [--out gpurun_out/t5.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PUBLISHED = {("Sphere", 50, 5): 0.869, ("Sphere", 50, 1): 0.861, ("Sphere", 20, 5): 0.847, ("Sphere", 20, 1): 0.82,
             ("Sphere", 100, 1): 0.869, ("Cube", 20, 1): 0.649}
EXP1 = os.path.join(ROOT, "tests", "golden", "Data", "RealFlowData", "Exp1")


def load_exp1(d=EXP1):
    """-> list of (name, gt, init_pos, init_vel, obs[T,3]) in the reference's trial order (fileio.jl:74-99)."""
    from smc_physics_b200 import host
    names = sorted(host.filter_unwanted_filenames(os.listdir(d)), key=host.trial_order)
    out = []
    for f in names:
        ip, _, iv, obs, T = host.read_obs_file(os.path.join(d, f))
        gt = int("".join(c for c in f.split("_")[1] if c.isdigit())) / 10.0
        out.append((f, gt, np.asarray(ip, float), np.asarray(iv, float), obs))
    return out


def judgments(trials, target, N, R, seed, device=0, resampler=0, substeps=9):
    """One pass of main.jl over the experiment: every trial's filter in flight at once (its own handle + stream),
    then the unweighted mean elasticity of each trial's final particles."""
    import smc_physics_b200 as sp
    shape = sp.SPHERE if target.lower() == "sphere" else sp.CUBE
    pfs = []
    for k, (_, _, ip, iv, obs) in enumerate(trials):
        kw = dict(num_particles=N, shape=shape, substeps=substeps, max_timesteps=max(obs.shape[0], 1),
                  rejuvenation_moves=R, device=device, seed=(seed * 1000003 + k * 7919 + 17) & 0xFFFFFFFFFFFF)
        if resampler:
            kw["resampler"] = resampler
        pf = sp.ParticleFilter(**kw)
        pf.set_trial(ip, iv, obs)
        pf.run_smc(R, want_stats=False)
        pfs.append(pf)
    j = np.zeros(len(trials))
    for k, pf in enumerate(pfs):
        j[k] = pf.get_particles(want_score=False)["e"].mean()
        pf.close()
    return j


def pearson(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    a, b = a - a.mean(), b - b.mean()
    return float((a * b).sum() / np.sqrt((a * a).sum() * (b * b).sum()))


def run(settings=None, seeds=16, device=0, resampler=0, verbose=True):
    trials = load_exp1()
    gt = np.array([t[1] for t in trials])
    res = {}
    for key in (settings or list(PUBLISHED)):
        target, N, R = key
        t0 = time.time()
        rs, bias, J = [], [], []
        for s in range(seeds):
            j = judgments(trials, target, N, R, seed=s + 1, device=device, resampler=resampler)
            rs.append(pearson(gt, j))
            bias.append(float((j - gt).mean()))
            J.append(j)
        J = np.stack(J)
        levels = sorted(set(gt.tolist()))
        res["%s_Par%d_Rej%d" % key] = dict(
            published_r=PUBLISHED.get(key), r_mean=float(np.mean(rs)), r_sd=float(np.std(rs, ddof=1)) if seeds > 1 else 0.0,
            r_min=float(np.min(rs)), r_max=float(np.max(rs)), bias_mean=float(np.mean(bias)),
            frac_above_gt=float((J.mean(axis=0) > gt).mean()),
            judgment_by_gt={("%.1f" % g): float(J[:, gt == g].mean()) for g in levels},
            seeds=seeds, resampler=resampler, seconds=time.time() - t0)
        if verbose:
            v = res["%s_Par%d_Rej%d" % key]
            print("%-20s r = %.3f +- %.3f [%.3f, %.3f]  (published %s)  bias %+.3f  above-GT %.2f  %.1fs" % (
                "%s_Par%d_Rej%d" % key, v["r_mean"], v["r_sd"], v["r_min"], v["r_max"], v["published_r"], v["bias_mean"],
                v["frac_above_gt"], v["seconds"]), flush=True)
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=16)
    ap.add_argument("--resampler", type=int, default=0)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import __graft_entry__ as entry
    entry.build()
    out = run(seeds=a.seeds, resampler=a.resampler)
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)
