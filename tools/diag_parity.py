"""Diagnostic (GPU box): where do GPU and float32-oracle sphere trajectories part ways?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import smc_physics_b200 as sp
from oracle import oracle as orc
from smc_physics_b200.workloads import synthetic_trial
IP, IV = [0.1, -0.2, 0.6], [3.0, -2.0, 1.5]
rng = np.random.default_rng(15)
n, T, S = 4096, 120, 5
mu = orc.macro_rollout(0, 0.6, orc.state13(IP, vel=IV), T, S=S)[:, 0:3]
path = np.vstack([np.asarray(IP)[None], mu[:-1]])
e = rng.uniform(0, 1, n).astype(np.float32)
for sig in (0.03, 0.075):
    xt = (path[:, :, None] + sig * rng.standard_normal((T, 3, n))).astype(np.float32)
    mg, lg, sg = sp.simulate(0, S, IP, IV, e, xt)
    mo, lo, so = orc.simulate(0, S, IP, IV, e, xt, f32=True, threads=16)
    m64 = orc.simulate(0, S, IP, IV, e, xt, f32=False, threads=16)[0]
    errt = np.abs(mg.astype(np.float64) - mo).max(axis=1)
    err = errt.max(axis=0)
    yard = np.abs(mo.astype(np.float64) - m64).max(axis=(0, 1))
    print("sigma %.3f: agree(1e-4) gpu-vs-f32 %.4f   f32-vs-f64 %.4f   median err %.2e  p99 %.2e max %.2e" % (
        sig, (err < 1e-4).mean(), (yard < 1e-4).mean(), np.median(err), np.quantile(err, 0.99), err.max()))
    bad = np.where(err >= 1e-4)[0][:6]
    for i in bad:
        k = int(np.argmax(errt[:, i] >= 1e-4))
        print("  particle %d e=%.3f first step %d xt=%s gpu=%s orc=%s" % (i, e[i], k, xt[k, :, i], mg[k, :, i], mo[k, :, i]))
for e_true in (0.8, 0.2):
    ip, iv, obs = synthetic_trial(sp.SPHERE, T=60, S=5, e_true=e_true, init_vel=(1.0, -0.5, 0.0))
    with sp.ParticleFilter(num_particles=4096, shape=0, substeps=5, max_timesteps=60, seed=5) as pf:
        pf.set_trial(ip, iv, obs)
        pf.run_smc(R=2, want_stats=False)
        ee = pf.get_particles(want_score=False)["e"]
    ref = orc.run_smc(0, 1024, 2, 5, ip, iv, obs, seed=5, f32=True, threads=16)["e"]
    print("e_true %.1f: gpu mean %.3f sd %.3f | oracle mean %.3f sd %.3f" % (e_true, ee.mean(), ee.std(), ref.mean(), ref.std()))
