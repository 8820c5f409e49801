"""Pins the CPU oracle against the reference's own PyBullet golden vectors.

Fixtures (copied verbatim from the reference repo, data only):
  tests/golden/test_elasticity_cube.csv   <- Tests/SimulatorComparison/test_elasticity_cube.csv
      written by test_elasticity (Tests/SimulatorComparison/tests.py:108-145): cube h=.05 dropped from (0,0,1),
      200 raw stepSimulation() calls per restitution 0.0..0.9, float64.
  tests/golden/test_orientation_cube.csv  <- Tests/SimulatorComparison/test_orientation_cube.csv
      test_orientation (tests.py:148-196): 10 pitch angles, values rounded to 3 dp.
"""
import collections
import csv
import os

import numpy as np
import pytest

from oracle import oracle as orc

G = os.path.join(os.path.dirname(__file__), "golden")


def _load(name, key_col=0):
    rows = list(csv.reader(open(os.path.join(G, name))))[1:]
    runs = collections.OrderedDict()
    for r in rows:
        runs.setdefault(r[key_col], []).append([float(x) for x in r[1:]])
    return {float(k): np.array(v) for k, v in runs.items()}


def _drop(e, quat=(0, 0, 0, 1), f32=False):
    st0 = orc.state13([0, 0, 1.0], quat)
    # the fixture's plane restitution is 0.3 (tests.py:20), not the model's 0.5
    # that scene has no ceiling and its walls are 0.45 m from the cube: plane only (nbox=0)
    out = orc.sim_raw(orc.CUBE, e, st0, 200, static_rest=0.3, f32=f32, nbox=0)
    return np.vstack([st0[None, :], out])


def test_free_flight_exact():
    fx = _load("test_elasticity_cube.csv")
    for e, ref in fx.items():
        o = _drop(e)
        # frames 0..106 are contact-free and identical for every restitution (SURVEY [FIX])
        assert np.abs(o[:107, 0:3] - ref[:107]).max() < 1e-9


def test_first_contact_frame_and_impact_response():
    fx = _load("test_elasticity_cube.csv")
    for e, ref in fx.items():
        o = _drop(e)
        # contact is registered at frame 106 (gap 1.57 mm < 0.02*sqrt(3)*h) and first acts in frame 107
        assert np.abs(o[107:112, 0:3] - ref[107:112]).max() < 1e-9, e


@pytest.mark.parametrize("e,tol", [(0.0, 1e-8), (0.1, 1e-8), (0.3, 1e-8), (0.4, 1e-8), (0.5, 1e-8), (0.6, 1e-8),
                                   (0.7, 1e-8), (0.8, 1e-8),
                                   # solver early-exit threshold flip at frame 117 (residual 8.7e-8 vs 1e-7)
                                   (0.2, 1e-6)])
def test_full_200_frames(e, tol):
    fx = _load("test_elasticity_cube.csv")
    ref = fx[min(fx, key=lambda k: abs(k - e))]
    o = _drop(e)
    assert np.abs(o[:, 0:3] - ref).max() < tol


def test_e09_until_corner_balance():
    fx = _load("test_elasticity_cube.csv")
    ref = fx[min(fx, key=lambda k: abs(k - 0.9))]
    o = _drop(0.9)
    err = np.abs(o[:, 0:3] - ref).max(axis=1)
    assert err[:157].max() < 1e-8       # bit-level agreement through the bounce and the second landing
    assert err.max() < 2e-3             # then the cube balances on one corner and round-off is amplified
    # post-bounce apex: 0.08995 m
    assert abs(o[108:160, 2].max() - ref[108:160, 2].max()) < 1e-8


def test_apex_heights_and_rest_height():
    fx = _load("test_elasticity_cube.csv")
    for e, ref in fx.items():
        o = _drop(e)
        assert abs(o[112:, 2].max() - ref[112:, 2].max()) < 1e-5
        assert abs(o[-1, 2] - ref[-1, 2]) < 1e-3


def _q_from_pitch(th):
    return (0.0, np.sin(th / 2), 0.0, np.cos(th / 2))


def test_orientation_fixture():
    fx = _load("test_orientation_cube.csv")
    for th, ref in fx.items():
        # the body in that script had restitution 0.9 (tests.py:76 pattern) on a 0.3 plane
        o = _drop(0.9, _q_from_pitch(th))
        err = np.abs(o[:190, 0:3] - ref[:190, 0:3]).max()
        assert err < 1.0e-3, (th, err)   # fixture is rounded to 1e-3


def test_f32_oracle_tracks_f64_through_first_bounce():
    for e in (0.0, 0.5, 0.9):
        a = _drop(e)
        b = _drop(e, f32=True)
        assert np.abs(a[:112, 0:3] - b[:112, 0:3]).max() < 2e-5


def test_substep_count_rule():
    # `while dt <= step_dur` (PhyBullet.jl:82-88): 5 substeps at 1/60, 9 at 1/30
    assert orc.lib().oracle_substeps_for(1 / 60) == 5
    assert orc.lib().oracle_substeps_for(1 / 30) == 9
