"""CPU checks of the oracle legs added in round 2 (the checker must be right before it checks the kernels):
multinomial parents (Gen.maybe_resample!, Inference/particle_filter.jl:55), predict (particle_filter.jl:84-102),
get_score (Utilities/fileio.jl:267) and the Gaussian-drift MCMC chain (Inference/mcmc.jl:53-93)."""
import numpy as np

from oracle import oracle as orc

IP, IV = [0.1, -0.2, 0.6], [3.0, -2.0, 1.5]


def _path(T, e=0.6, shape=0):
    mu = orc.macro_rollout(shape, e, orc.state13(IP, vel=IV), T, S=5)[:, 0:3]
    return np.vstack([np.asarray(IP)[None], mu[:-1]])


def test_multinomial_parents_follow_the_weights():
    rng = np.random.default_rng(0)
    n = 60000
    lw = rng.normal(0, 1.0, n)
    w = np.exp(lw - lw.max())
    w /= w.sum()
    idx = orc.resample_multinomial(lw, 5, 2)
    assert idx.min() >= 0 and idx.max() < n
    assert np.array_equal(idx, orc.resample_multinomial(lw, 5, 2))            # counter based: reproducible
    assert not np.array_equal(idx, orc.resample_multinomial(lw, 5, 3))        # another step, other parents
    counts = np.bincount(idx, minlength=n)
    blocks = np.array_split(np.arange(n), 40)
    z = [(counts[b].sum() - n * w[b].sum()) / np.sqrt(n * w[b].sum() * (1 - w[b].sum())) for b in blocks]
    assert np.abs(z).max() < 4.5 and abs(np.mean(z)) < 0.7
    # parents are i.i.d.: not sorted, and the variance of the offspring counts is binomial, not systematic's < 1
    assert not np.all(np.diff(idx) >= 0)
    heavy = w * n > 3
    assert counts[heavy].var() > 1.5
    # degenerate weights
    lw2 = np.full(100, -np.inf)
    lw2[17] = 0.0
    assert np.all(orc.resample_multinomial(lw2, 1, 0) == 17)


def test_filter_with_multinomial_resampling_has_the_same_posterior_as_systematic():
    T = 25
    path = _path(T, 0.7)
    a = [orc.run_smc_ex(0, 256, 1, 5, IP, IV, path, seed=s, resampler=0, threads=8)["e"].mean() for s in range(8)]
    b = [orc.run_smc_ex(0, 256, 1, 5, IP, IV, path, seed=100 + s, resampler=1, threads=8)["e"].mean() for s in range(8)]
    se = np.sqrt(np.var(a, ddof=1) / 8 + np.var(b, ddof=1) / 8)
    assert abs(np.mean(a) - np.mean(b)) < 3 * se + 1e-3, (np.mean(a), np.mean(b), se)


def test_score_is_the_log_joint_of_the_trace():
    """score - ltrans = first-step transition term + observation terms: bounded above by the density maxima."""
    T = 10
    path = _path(T)
    r = orc.run_smc_ex(0, 64, 1, 5, IP, IV, path, seed=4, want_score=True)
    s_obs, s_tr = float(np.float32(0.05)), float(np.float32(0.075))
    top = -3 * T * (np.log(s_obs) + 0.9189385332046727) - 3 * (np.log(s_tr) + 0.9189385332046727)
    rest = r["score"] - r["ltrans"]
    assert np.isfinite(r["score"]).all() and np.all(rest <= top + 1e-9)
    assert np.all(r["ltrans"] <= -3 * (T - 1) * (np.log(s_tr) + 0.9189385332046727) + 1e-9)


def test_predict_rolls_forward_with_model_noise():
    T, P = 8, 15
    r = orc.run_smc_ex(0, 400, 1, 5, IP, IV, _path(T), seed=9, n_pred=P)
    pos, obs = r["pred_pos"].astype(np.float64), r["pred_obs"].astype(np.float64)
    assert pos.shape == (P, 3, 400) and np.isfinite(pos).all()
    assert abs((obs - pos).std() - 0.05) < 0.004                     # observation noise sigma_obs
    assert np.array_equal(pos, orc.run_smc_ex(0, 400, 1, 5, IP, IV, _path(T), seed=9, n_pred=P)["pred_pos"])


def test_mcmc_chain_without_resampling_moves_towards_the_data():
    T = 30
    path = _path(T, 0.8)
    r = orc.run_smc_ex(0, 48, 0, 5, IP, IV, path, seed=2, no_resample=True, mcmc_runs=60, threads=8)
    assert r["stats"][:, 3].sum() == 0                                # generate(): never resampled
    tr = r["e_trace"]
    assert tr.shape == (60, 48) and tr.min() >= 0 and tr.max() <= 1
    assert np.mean([np.unique(tr[:, c]).size for c in range(48)]) > 5  # chains accept moves
    prior = r["e_trace"][0]
    assert abs(tr[-20:].mean() - 0.8) < abs(prior.mean() - 0.8) + 0.05
