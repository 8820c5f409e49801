"""BASELINE.json configs[2] shape: K independent sphere trials of 100k particles each (S = 9, R = 5, T varies like
the RealFlow trials), one after the other vs all in flight at once on one GPU (one handle + stream per trial).
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
Ts = [30 + (17 * k) % 40 for k in range(K)]          # RealFlow trials have 20-70 frames (mean 45)
trials = [wl.synthetic_trial(sp.SPHERE, T=T, S=9, e_true=0.1 + 0.8 * k / max(K - 1, 1)) for k, T in enumerate(Ts)]


def run(concurrent):
    pfs = [sp.ParticleFilter(num_particles=N, shape=sp.SPHERE, substeps=9, max_timesteps=T, rejuvenation_moves=5, seed=11 + k)
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
print(json.dumps({"trials": K, "particles_per_trial": N, "frames": sum(Ts), "sequential_s": ts, "concurrent_s": tc,
                  "sequential_particle_steps_per_s": steps / ts, "concurrent_particle_steps_per_s": steps / tc}))
