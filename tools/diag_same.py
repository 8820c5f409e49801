import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import smc_physics_b200 as sp
from oracle import oracle as orc
IP, IV = [0.1, -0.2, 0.6], [3.0, -2.0, 1.5]
for e_gt in (0.5, 0.6):
    mu = orc.macro_rollout(0, e_gt, orc.state13(IP, vel=IV), 12, S=5)[:, 0:3]
    path = np.vstack([np.asarray(IP)[None], mu[:-1]])
    for seed in (3, 99):
        for R in (1, 2):
            with sp.ParticleFilter(num_particles=64, shape=0, substeps=5, max_timesteps=12, seed=seed) as pf:
                pf.set_trial(IP, IV, path)
                st = pf.run_smc(R=R)
                out = pf.get_particles(want_traj=True)
            ref = orc.run_smc(0, 64, R, 5, IP, IV, path, seed=seed, f32=True, want_traj=True)
            res_g = [int(s["resampled"]) for s in st]; res_o = [int(r) for r in ref["stats"][:, 3]]
            acc_g = [round(s["accept_rate"], 3) for s in st]; acc_o = [round(a, 3) for a in ref["stats"][:, 2]]
            print(e_gt, seed, R, "same e", (out["e"] == ref["e"]).mean(), "res equal", res_g == res_o, "first acc diff", next((i for i in range(12) if acc_g[i] != acc_o[i]), None))
