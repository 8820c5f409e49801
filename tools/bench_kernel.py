"""Lineage-controlled timing of ONE k_rejuvenate launch (kernel A/B tool, GPU box).

A full bench run is chaotic in its cost: a rounding-level change in the chain code flips an accept decision, the
lineage changes, and which retained positions survive into the old steps moves the launch time by +-3 %.  Here the
filter runs all T steps WITHOUT rejuvenation (extend + resample only: that code is not what is being tuned, so the
retained-position histories are identical across builds), then the rejuvenation launch of the last step is timed.
usage: python tools/bench_kernel.py [shape] [particles] [T]      (SMCPHYS_SO selects a variant build)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import smc_physics_b200 as sp
from smc_physics_b200 import workloads as wl

shape = sp.CUBE if len(sys.argv) > 1 and sys.argv[1] == "cube" else sp.SPHERE
n = int(sys.argv[2]) if len(sys.argv) > 2 else (1 << 20 if shape == sp.SPHERE else 1 << 18)
T = int(sys.argv[3]) if len(sys.argv) > 3 else 120
trial = wl.synthetic_trial(shape, T=T, S=5)
with sp.ParticleFilter(num_particles=n, shape=shape, substeps=5, max_timesteps=T, seed=1234) as pf:
    pf.set_trial(*trial)
    pf.pf_init()
    for t in range(T):
        pf.pf_extend(t)
        pf.pf_resample_local(t)
    pf.synchronize()
    xt_sum = float(np.abs(pf.get_particles(want_score=True)["score"]).sum())   # fingerprint of the histories
    pf.set_profiling(True)
    times = []
    for k in range(4):
        pf.kernel_time(sp.K_REJUVENATE)
        pf.pf_rejuvenate(T - 1, 5, move0=5 * k)
        times.append(pf.kernel_time(sp.K_REJUVENATE)[0])
    e = pf.get_particles(want_score=False)["e"]
    import ctypes as C
    L = C.CDLL(os.environ.get("SMCPHYS_SO") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "smc_physics_b200", "libsmcphys.so"))
    if hasattr(L, "smcphys_debug_step_clocks"):          # -DRS_STEP_CLOCK build: warp cycles per macro step of the chain
        buf = (C.c_uint64 * T)()
        L.smcphys_debug_step_clocks(buf, T)              # clear what the warm-up launches accumulated
        pf.pf_rejuvenate(T - 1, 5, move0=100)
        L.smcphys_debug_step_clocks(buf, T)
        clk = np.array(list(buf), dtype=np.float64)
        print("share of chain cycles per macro step k (x1000):")
        for k0 in range(0, T, 10):
            print("  k=%3d.. " % k0 + " ".join("%5.1f" % (1000 * x / clk.sum()) for x in clk[k0:k0 + 10]))
print("%s n=%d T=%d: k_rejuvenate(t=%d, R=5) ms: %s  | history fingerprint %.6e  mean e %.5f" % (
    "cube" if shape == sp.CUBE else "sphere", n, T, T - 1, " ".join("%.3f" % x for x in times), xt_sum, e.mean()))
