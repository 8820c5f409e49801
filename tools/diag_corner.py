"""Diagnostic (GPU box): one macro step from corner / under-wall poses: coop kernel vs generic kernel vs oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import smc_physics_b200 as sp
from oracle import oracle as orc
rng = np.random.default_rng(3)
n = 4096
for name, lo, hi, vel in (("corner +x/-y", [0.44, -0.48, 0.3], [0.47, -0.44, 0.9], [2.9, -2.0, 0.4]),
                          ("floor + -x wall", [-0.47, -0.2, 0.03], [-0.44, 0.2, 0.07], [-0.6, 0.25, -0.3]),
                          ("under -x wall", [-0.53, -0.2, -0.04], [-0.47, 0.2, 0.02], [-0.6, 0.25, -0.3])):
    x = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    e = rng.uniform(0, 1, n).astype(np.float32)
    xt = np.ascontiguousarray(x.T[None])                       # [1,3,n]
    mg, _, sg = sp.simulate(0, 5, x[0], vel, e, xt)             # coop path (first row teleports to x)
    st = np.zeros((n, 13), np.float32); st[:, 0:3] = x; st[:, 6] = 1; st[:, 7:10] = vel
    gen = sp.sim_step(0, st, e, S=5)                           # generic per-thread path
    ref = np.stack([orc.macro_rollout(0, float(e[i]), st[i].astype(np.float64), 1, S=5, f32=True)[0] for i in range(n)])
    for lab, a in (("coop", sg), ("generic", gen)):
        err = np.abs(a[:, 0:3] - ref[:, 0:3]).max(axis=1)
        ev = np.abs(a[:, 7:13] - ref[:, 7:13]).max(axis=1)
        print("%-16s %-8s pos: agree(1e-5) %.4f p99 %.2e max %.2e | vel max %.2e" % (name, lab, (err < 1e-5).mean(), np.quantile(err, .99), err.max(), ev.max()))
    i = int(np.argmax(np.abs(sg[:, 0:3] - ref[:, 0:3]).max(axis=1)))
    print("   worst coop: x=%s e=%.3f coop=%s generic=%s oracle=%s" % (x[i], e[i], sg[i, [0,1,2,7,8,9]], gen[i, [0,1,2,7,8,9]], ref[i, [0,1,2,7,8,9]]))
