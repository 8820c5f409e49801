"""Per-call latency of the stateless PhySMC.forward_step entry (smcphys_sim_step on one body) and of a getter: what a host
loop over PhySMC.step pays per step (GPU box)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import smc_physics_b200 as sp
st = np.zeros((1, 13), np.float32); st[0, 2] = 0.6; st[0, 6] = 1; st[0, 7:10] = (3, -2, 1.5)
for _ in range(5):
    st = sp.sim_step(sp.SPHERE, st, 0.6)
t0 = time.perf_counter()
for _ in range(2000):
    st = sp.sim_step(sp.SPHERE, st, 0.6)
print("smcphys_sim_step, 1 body: %.1f us per call" % ((time.perf_counter() - t0) / 2000 * 1e6))
lw = np.random.default_rng(0).normal(0, 2, 4096)
for _ in range(5):
    sp.resample_systematic(lw, 0.3)
t0 = time.perf_counter()
for _ in range(1000):
    sp.resample_systematic(lw, 0.3)
print("smcphys_resample_systematic, n = 4096: %.1f us per call" % ((time.perf_counter() - t0) / 1000 * 1e6))
