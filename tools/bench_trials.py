"""BASELINE.json configs[2] shape: K independent sphere trials of 100k particles each (S = 9, R = 5, T varies like
the RealFlow trials; `27 1000 cube 5 120` = configs[1], the 27 BulletTest_Cube trials at 1000 particles), one after the other vs all in flight at once on one GPU (one handle + stream per trial).
Prints particle-steps/s for both."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import smc_physics_b200 as sp  # noqa: E402
from smc_physics_b200 import workloads as wl  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
SHAPE = sp.CUBE if len(sys.argv) > 3 and sys.argv[3] == "cube" else sp.SPHERE
S = int(sys.argv[4]) if len(sys.argv) > 4 else 9      # 9 = RealFlow (1/30 s), 5 = BulletTest (1/60 s)
TFIX = int(sys.argv[5]) if len(sys.argv) > 5 else 0   # 0: RealFlow-like 30-70 frames; else that many frames
Ts = [TFIX or 30 + (17 * k) % 40 for k in range(K)]   # RealFlow trials have 20-70 frames (mean 45)
trials = [wl.synthetic_trial(SHAPE, T=T, S=S, e_true=0.1 + 0.8 * k / max(K - 1, 1)) for k, T in enumerate(Ts)]


def run(concurrent):
    pfs = [sp.ParticleFilter(num_particles=N, shape=SHAPE, substeps=S, max_timesteps=T, rejuvenation_moves=5, seed=11 + k)
           for k, T in enumerate(Ts)]
    for pf, tr in zip(pfs, trials):
        pf.set_trial(*tr)
    for pf in pfs:
        pf.synchronize()
    t0 = time.perf_counter()
    means = []
    if concurrent:
        for pf in pfs:
            pf.run_smc(5, want_stats=False)
        for pf in pfs:
            means.append(float(pf.get_particles(want_score=False)["e"].mean()))
    else:
        for pf in pfs:
            pf.run_smc(5, want_stats=False)
            means.append(float(pf.get_particles(want_score=False)["e"].mean()))
    dt = time.perf_counter() - t0
    for pf in pfs:
        pf.close()
    return dt, means


run(True)  # warm-up
ts, ms = run(False)
tc, mc = run(True)
assert ms == mc, "concurrent trials must give the same posteriors"
steps = N * sum(Ts)
print(json.dumps({"shape": "cube" if SHAPE == sp.CUBE else "sphere", "substeps": S, "trials": K, "particles_per_trial": N, "frames": sum(Ts), "sequential_s": ts, "concurrent_s": tc,
                  "sequential_particle_steps_per_s": steps / ts, "concurrent_particle_steps_per_s": steps / tc}))
