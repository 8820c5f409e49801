"""Synthetic bouncing-object trials of the shapes BASELINE.json names.

The observed path is a noise-free rollout of the library's own simulator (PhySMC.forward_step on a batch of one
body), row 1 = the initial position as in the RealFlow stimuli (Data/RealFlowData/Exp1/*_observed.csv)."""
import numpy as np

from . import api

INIT_POS = (0.1, -0.2, 0.6)
INIT_VEL = (3.0, -2.0, 1.5)
E_TRUE = 0.6


def synthetic_trial(shape=api.SPHERE, T=120, S=5, e_true=E_TRUE, init_pos=INIT_POS, init_vel=INIT_VEL, device=0):
    """-> (init_pos[3], init_vel[3], obs[T,3]) float64."""
    st = np.zeros((1, 13), np.float32)
    st[0, 0:3] = init_pos
    st[0, 6] = 1.0
    st[0, 7:10] = init_vel
    obs = np.zeros((T, 3))
    for t in range(T):
        obs[t] = st[0, 0:3]
        st = api.sim_step(shape, st, e_true, S=S, device=device)
    return np.asarray(init_pos, float), np.asarray(init_vel, float), obs


# Algorithmic FLOP per kernel() execution on the re-simulation path (counting rule of SURVEY.md section 8d:
# add/sub/mul/cmp/min/max = 1, FMA = 2, div/sqrt = 1).  Free-flight substep of a sphere: linear drag 13 + angular
# drag 12 + interior test 7 + position update 6 = 38 -- SURVEY's 86 also counts the 48-FLOP quaternion update, which
# the rejuvenation kernel does not execute for a sphere (its orientation never feeds back into the motion), so it
# is not claimed here.  Cube: 139 (rotation matrix, AABB extents and the quaternion update are needed).  Plus 13 per
# kernel step for the transition log density.  Contact detection / solver, integer and Philox work are NOT counted,
# so achieved fractions are lower bounds.
def flop_per_kernel_step(shape, S):
    free = 38 if shape == api.SPHERE else 139
    return S * free + 13


def flop_rejuvenate(shape, S, n, R, t):
    """one k_rejuvenate launch at PF step t (0-based): n particles x R moves x (t+1) kernel steps + MH overhead."""
    return n * R * ((t + 1) * flop_per_kernel_step(shape, S) + 40)


def kernel_steps(n, T, R):
    """kernel() executions of one SMC run without Gen's redundant re-extension: n (T + R T (T+1) / 2)."""
    return n * (T + R * T * (T + 1) // 2)
