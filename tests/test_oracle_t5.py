"""T5 for the oracle itself (CPU): the reference's published Exp1 correlation, reproduced by the float64 restatement.

This is the pin of the SPHERE contact path, the box walls and the Gen bookkeeping as a whole: the reference ships no
sphere fixture (Tests/SimulatorComparison/test_elasticity_sphere.csv is degenerate), but it publishes r(model, GT) on
the 150 Exp1 stimuli (Analysis/plots.jl:51 over Analysis/utils.jl:102-124) together with the scatter
plot those numbers are printed on.  The oracle, run as main.jl:24-31,54 runs the experiment (S = 9, unweighted mean
of the final particles), lands on the published r, on the published bias direction (estimates above the ground
truth) and on the per-level means read off the published scatter plot.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

from oracle import oracle as orc  # noqa: E402
import t5_exp1 as t5  # noqa: E402

# mean model estimate per ground-truth level 0.0 ... 0.9, read by eye off the reference's scatter plot
# Analysis/Exp1/Modelv5/Sphere/Init00_Obs05_Tra075/SMC/Par50_Rej5/Plots/TestJudgments/sim_vs_gt_individual_stimuli.png
# (15 markers per level; reading error about +-0.05)
PLOT_LEVEL_MEANS = [0.18, 0.33, 0.42, 0.60, 0.62, 0.72, 0.82, 0.85, 0.90, 0.93]


def test_exp1_has_the_150_reference_trials():
    trials = t5.load_exp1()
    assert len(trials) == 150
    assert sorted(set(t[1] for t in trials)) == [k / 10 for k in range(10)]
    T = [t[4].shape[0] for t in trials]
    assert min(T) == 20 and max(T) == 120
    # row 1 of the observed file is the initial position (RealFlow convention)
    for _, _, ip, _, obs in trials:
        assert np.abs(obs[0] - ip).max() < 1e-5


def test_oracle_reproduces_published_exp1_sphere_correlation():
    trials = t5.load_exp1()
    gt = np.array([t[1] for t in trials])
    rs, J = [], []
    for seed in range(2):
        j = np.array([orc.run_smc(orc.SPHERE, 20, 1, 9, ip, iv, obs, seed=seed * 1000003 + k * 7919 + 17, f32=False,
                                  threads=os.cpu_count() or 1)["e"].mean() for k, (_, _, ip, iv, obs) in enumerate(trials)])
        rs.append(t5.pearson(gt, j))
        J.append(j)
    J = np.stack(J).mean(axis=0)
    # published: 0.82 for Sphere Par20_Rej1 (one unseeded draw); the reference's own settings span 0.82 ... 0.869
    assert 0.74 <= np.mean(rs) <= 0.90, rs
    assert (J - gt).mean() > 0.08                       # biased above the ground truth, as every Exp1 plot shows
    assert (J > gt).mean() > 0.75
    levels = np.array([J[gt == g].mean() for g in sorted(set(gt))])
    assert np.abs(levels - np.array(PLOT_LEVEL_MEANS)).max() < 0.12, levels
