"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the CPU restatement in smc_oracle.hpp).  Imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by smc_physics_b200.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

SPHERE, CUBE = 0, 1


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("smc_oracle_capi.cpp", "smc_oracle.hpp", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        dp, fp, ip = C.POINTER(C.c_double), C.POINTER(C.c_float), C.POINTER(C.c_int64)
        up, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
        L.oracle_sim_raw.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, dp, C.c_int, dp]
        L.oracle_macro_rollout.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, dp, C.c_int, dp]
        L.oracle_simulate.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, dp, dp, fp, fp, C.c_int64,
                                      C.c_int, fp, dp, dp, C.c_int]
        L.oracle_run_smc.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                     C.c_uint64, C.c_double, C.c_int, dp, dp, dp, C.c_int, dp, dp, dp, fp, dp, dp]
        L.oracle_run_smc_ex.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                        C.c_uint64, C.c_double, C.c_int, C.c_int, C.c_int, dp, dp, dp, C.c_int, dp, dp, dp,
                                        fp, dp, dp, dp, C.c_int, fp, fp, C.c_int, dp]
        L.oracle_resample_multinomial.argtypes = [dp, C.c_int64, C.c_uint64, C.c_int, ip]
        L.oracle_resample_systematic.argtypes = [dp, C.c_int64, C.c_double, ip]
        L.oracle_resample_systematic_f64.argtypes = [dp, C.c_int64, C.c_double, ip]
        L.oracle_weight_stats.argtypes = [dp, C.c_int64, dp, up]
        L.oracle_weight_q.argtypes = [dp, C.c_int64, C.c_double, up]
        L.oracle_sys_threshold.argtypes = [C.c_uint64] * 4
        L.oracle_sys_threshold.restype = C.c_uint64
        L.oracle_det_exp.argtypes = [C.c_double]
        L.oracle_det_exp.restype = C.c_double
        L.oracle_philox.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, u32p]
        L.oracle_u01.argtypes = [C.c_uint32]
        L.oracle_u01.restype = C.c_double
        L.oracle_box_muller.argtypes = [C.c_uint32, C.c_uint32, dp]
        L.oracle_truncnorm_logpdf.argtypes = [C.c_double] * 5
        L.oracle_truncnorm_logpdf.restype = C.c_double
        L.oracle_truncnorm_sample.argtypes = [C.c_double] * 5
        L.oracle_truncnorm_sample.restype = C.c_double
        L.oracle_normal3_logpdf.argtypes = [dp, dp, C.c_double]
        L.oracle_normal3_logpdf.restype = C.c_double
        L.oracle_substeps_for.argtypes = [C.c_double]
        L.oracle_substeps_for.restype = C.c_int
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def state13(pos, quat=(0, 0, 0, 1), vel=(0, 0, 0), omega=(0, 0, 0)):
    return np.ascontiguousarray(np.concatenate([pos, quat, vel, omega]), dtype=np.float64)


def sim_raw(shape, e, st0, nsub, static_rest=0.5, ceiling_half=0.01, f32=False, nbox=5):
    """nsub raw 1/240 s substeps with a persistent contact cache; returns [nsub,13]."""
    out = np.zeros((nsub, 13))
    lib().oracle_sim_raw(int(f32), shape, e, static_rest, ceiling_half, nbox, _dp(st0), nsub, _dp(out))
    return out


def macro_rollout(shape, e, st0, nsteps, S=5, ceiling_half=0.01, f32=False):
    """Noise-free PhySMC.step rollout; returns [nsteps,13]."""
    out = np.zeros((nsteps, 13))
    lib().oracle_macro_rollout(int(f32), shape, e, ceiling_half, S, _dp(st0), nsteps, _dp(out))
    return out


def simulate(shape, S, init_pos, init_vel, e, xt, sigma_tr=float(np.float32(0.075)), ceiling_half=0.01, f32=False,
             threads=1):
    """Re-simulate n particles over retained noisy positions xt[T,3,n] (float32).

    Returns mu[T,3,n] float32 (simulator output after each step), logp[n] (sum_{s>=2} log N(xt_s; mu_s)),
    state[n,13]."""
    xt = np.ascontiguousarray(xt, dtype=np.float32)
    e = np.ascontiguousarray(e, dtype=np.float32)
    T, _, n = xt.shape
    mu = np.zeros((T, 3, n), np.float32)
    logp = np.zeros(n)
    st = np.zeros((n, 13))
    ip = np.ascontiguousarray(init_pos, dtype=np.float64)
    iv = np.ascontiguousarray(init_vel, dtype=np.float64)
    lib().oracle_simulate(int(f32), shape, S, ceiling_half, sigma_tr, _dp(ip), _dp(iv), _fp(e), _fp(xt), n, T, _fp(mu),
                          _dp(logp), _dp(st), threads)
    return mu, logp, st


def run_smc(shape, N, R, S, init_pos, init_vel, obs, seed=1234, sigma_obs=float(np.float32(0.05)),
            sigma_tr=float(np.float32(0.075)), prop_sigma=0.1, ceiling_half=0.01, f32=False, threads=1, want_traj=False):
    obs = np.ascontiguousarray(obs, dtype=np.float64)
    T = obs.shape[0]
    e = np.zeros(N)
    lw = np.zeros(N)
    lt = np.zeros(N)
    stats = np.zeros((T, 4))
    logml = np.zeros(1)
    traj = np.zeros((T, 3, N), np.float32) if want_traj else None
    ip = np.ascontiguousarray(init_pos, dtype=np.float64)
    iv = np.ascontiguousarray(init_vel, dtype=np.float64)
    lib().oracle_run_smc(int(f32), shape, N, R, S, sigma_obs, sigma_tr, prop_sigma, seed, ceiling_half, threads, _dp(ip),
                         _dp(iv), _dp(obs), T, _dp(e), _dp(lw), _dp(lt), _fp(traj) if want_traj else None, _dp(stats),
                         _dp(logml))
    return dict(e=e, log_w=lw, ltrans=lt, stats=stats, log_ml=float(logml[0]), traj=traj)


def run_smc_ex(shape, N, R, S, init_pos, init_vel, obs, seed=1234, sigma_obs=float(np.float32(0.05)),
               sigma_tr=float(np.float32(0.075)), prop_sigma=0.1, ceiling_half=0.01, f32=False, threads=1, resampler=0,
               no_resample=False, want_traj=False, want_score=False, n_pred=0, mcmc_runs=0):
    """run_smc plus the optional legs: multinomial resampling, get_score, predict, Gaussian-drift MCMC moves at the last
    step (with no_resample=True, R=0: Inference/mcmc.jl:53-93)."""
    obs = np.ascontiguousarray(obs, dtype=np.float64)
    T = obs.shape[0]
    e, lw, lt = np.zeros(N), np.zeros(N), np.zeros(N)
    stats, logml = np.zeros((T, 4)), np.zeros(1)
    traj = np.zeros((T, 3, N), np.float32) if want_traj else None
    score = np.zeros(N) if want_score else None
    ppos = np.zeros((n_pred, 3, N), np.float32) if n_pred else None
    pobs = np.zeros((n_pred, 3, N), np.float32) if n_pred else None
    etr = np.zeros((mcmc_runs, N)) if mcmc_runs else None
    ip = np.ascontiguousarray(init_pos, dtype=np.float64)
    iv = np.ascontiguousarray(init_vel, dtype=np.float64)
    lib().oracle_run_smc_ex(int(f32), shape, N, R, S, sigma_obs, sigma_tr, prop_sigma, seed, ceiling_half, threads,
                            int(resampler), int(no_resample), _dp(ip), _dp(iv), _dp(obs), T, _dp(e), _dp(lw), _dp(lt),
                            _fp(traj) if want_traj else None, _dp(stats), _dp(logml),
                            _dp(score) if want_score else None, n_pred, _fp(ppos) if n_pred else None,
                            _fp(pobs) if n_pred else None, mcmc_runs, _dp(etr) if mcmc_runs else None)
    return dict(e=e, log_w=lw, ltrans=lt, stats=stats, log_ml=float(logml[0]), traj=traj, score=score, pred_pos=ppos,
                pred_obs=pobs, e_trace=etr)


def resample_multinomial(lw, seed, t):
    lw = np.ascontiguousarray(lw, dtype=np.float64)
    idx = np.zeros(lw.shape[0], np.int64)
    lib().oracle_resample_multinomial(_dp(lw), lw.shape[0], seed, t, idx.ctypes.data_as(C.POINTER(C.c_int64)))
    return idx


def resample_systematic(lw, u0, f64_textbook=False):
    lw = np.ascontiguousarray(lw, dtype=np.float64)
    idx = np.zeros(lw.shape[0], np.int64)
    fn = lib().oracle_resample_systematic_f64 if f64_textbook else lib().oracle_resample_systematic
    fn(_dp(lw), lw.shape[0], u0, idx.ctypes.data_as(C.POINTER(C.c_int64)))
    return idx


def weight_stats(lw):
    lw = np.ascontiguousarray(lw, dtype=np.float64)
    out = np.zeros(3)
    iout = np.zeros(4, np.uint64)
    lib().oracle_weight_stats(_dp(lw), lw.shape[0], _dp(out), iout.ctypes.data_as(C.POINTER(C.c_uint64)))
    return dict(mx=out[0], ess=out[1], lse=out[2], W=int(iout[0]), S1=int(iout[1]), S2=int(iout[2]), below_half=bool(iout[3]))


def weight_q(lw, mx):
    lw = np.ascontiguousarray(lw, dtype=np.float64)
    q = np.zeros(lw.shape[0], np.uint64)
    lib().oracle_weight_q(_dp(lw), lw.shape[0], mx, q.ctypes.data_as(C.POINTER(C.c_uint64)))
    return q


def philox(seed, c0, c1, c2, c3):
    out = np.zeros(4, np.uint32)
    lib().oracle_philox(seed, c0, c1, c2, c3, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    return out
