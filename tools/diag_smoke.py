"""Diagnostic (GPU box): why does a GPU run differ from the float32 oracle run with the SAME seed and Philox counters?
Round-1 smoke printed posterior means 0.4355 (GPU) vs 0.4888 (oracle) at N = 512, R = 2, T = 40, seed 7.
Prints, per PF step, whether the resampling decisions agree, the relative ESS difference and the accept-rate
difference, the step of first divergence, and the seed-to-seed spread of the posterior mean on both sides."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import smc_physics_b200 as sp
from oracle import oracle as orc

IP, IV, T, N, R = [0.1, -0.2, 0.6], [3.0, -2.0, 1.5], 40, 512, 2
truth = orc.macro_rollout(orc.SPHERE, 0.6, orc.state13(IP, vel=IV), T)[:, 0:3]
path = np.vstack([np.asarray(IP)[None], truth[:-1]])
with sp.ParticleFilter(num_particles=N, shape=0, substeps=5, max_timesteps=T, seed=7) as pf:
    pf.set_trial(IP, IV, path)
    st = pf.run_smc(R=R)
    e_g = pf.get_particles()["e"]
ref = orc.run_smc(0, N, R, 5, IP, IV, path, seed=7, f32=True)
first = None
for t in range(T):
    d_ess = abs(st[t]["ess"] - ref["stats"][t, 0]) / ref["stats"][t, 0]
    d_acc = st[t]["accept_rate"] - ref["stats"][t, 2]
    same_res = int(st[t]["resampled"]) == int(ref["stats"][t, 3])
    if first is None and (d_ess > 1e-9 or abs(d_acc) > 0 or not same_res):
        first = t
    print("t=%2d resampled gpu %d oracle %d  ESS rel diff %.2e  accept diff %+.4f" % (
        t, st[t]["resampled"], ref["stats"][t, 3], d_ess, d_acc))
print("first step at which the two runs differ at all: %s" % first)
print("same seed: gpu mean %.4f oracle(f32) mean %.4f; particles equal to 1e-12: %.3f" % (
    e_g.mean(), ref["e"].mean(), (np.abs(np.sort(e_g) - np.sort(ref["e"])) < 1e-12).mean()))
g, o = [], []
for sd in range(16):
    with sp.ParticleFilter(num_particles=N, shape=0, substeps=5, max_timesteps=T, seed=100 + sd) as pf:
        pf.set_trial(IP, IV, path)
        pf.run_smc(R=R, want_stats=False)
        g.append(pf.get_particles()["e"].mean())
    o.append(orc.run_smc(0, N, R, 5, IP, IV, path, seed=200 + sd, f32=False, threads=16)["e"].mean())
g, o = np.asarray(g), np.asarray(o)
print("16 seeds: gpu mean %.4f (sd between seeds %.4f) | oracle f64 mean %.4f (sd between seeds %.4f)" % (
    g.mean(), g.std(ddof=1), o.mean(), o.std(ddof=1)))
