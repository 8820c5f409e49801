"""T5 on the CUDA path: the reference's published Exp1 numbers, the only results the reference
itself holds for this path.  r(model, GT) = Pearson correlation of the per-stimulus unweighted mean elasticity
(Analysis/utils.jl:102-124) with the ground truth (Analysis/plots.jl:51) over the 150 Exp1 trials, run as main.jl:24-31,54
runs them (RealFlow => S = 9), averaged over seeds because the reference is one unseeded draw per number.
The same experiment on the float64 oracle is tests/test_oracle_t5.py."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

SEEDS = 16
# published r, and the band the seed average must fall in: sphere settings span 0.82 ... 0.869 in the reference's plots
# ("in the 0.8-0.9 band"); the cube is quoted once, 0.649 (+-0.1).
BANDS = {("Sphere", 50, 5): (0.80, 0.90), ("Sphere", 50, 1): (0.80, 0.90), ("Sphere", 20, 5): (0.80, 0.90),
         ("Sphere", 20, 1): (0.78, 0.90), ("Cube", 20, 1): (0.549, 0.749)}


@pytest.fixture(scope="module")
def t5():
    import __graft_entry__ as entry
    entry.build()
    import t5_exp1
    return t5_exp1


@pytest.mark.parametrize("key", sorted(BANDS))
def test_exp1_correlation_in_published_band(t5, key):
    res = t5.run(settings=[key], seeds=SEEDS, verbose=True)["%s_Par%d_Rej%d" % key]
    lo, hi = BANDS[key]
    assert lo <= res["r_mean"] <= hi, res
    assert abs(res["r_mean"] - res["published_r"]) < 0.06 + 3 * res["r_sd"] / np.sqrt(SEEDS), res
    # every Exp1 plot of the reference shows the model above the ground truth (restitutions multiply: e x 0.5)
    assert res["bias_mean"] > 0.08 and res["frac_above_gt"] > 0.75, res
    lv = [res["judgment_by_gt"]["%.1f" % (k / 10)] for k in range(10)]
    assert lv[9] > lv[5] > lv[0]
