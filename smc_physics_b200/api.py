"""numpy-level wrappers over the C ABI (host buffers in, host buffers out)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import CUBE, SPHERE, Config, StepStats, check, lib  # noqa: F401

F32 = np.float32
SIGMA_OBS = float(F32(0.05))    # args.jl:14 (Float32 literal promoted to Float64)
SIGMA_TR = float(F32(0.075))    # args.jl:15


def _dp(a):
    return a.ctypes.data_as(_lib._dp) if a is not None else None


def _fp(a):
    return a.ctypes.data_as(_lib._fp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(_lib._ip) if a is not None else None


def substeps_for(step_dur, pb_timestep=1.0 / 240.0):
    """`while dt <= sim.step_dur` of PhySMC/PhyBullet/PhyBullet.jl:82-88: 5 at 1/60, 9 at 1/30."""
    n, dt = 0, 0.0
    while dt <= step_dur:
        n += 1
        dt += pb_timestep
    return n


def sim_step(shape, state13, restitution, S=5, ceiling_half=0.01, device=0):
    """PhySMC.forward_step on a batch: state13 [n,13] (pos, quat xyzw, vel, omega) -> new [n,13] float32."""
    st = np.array(state13, dtype=F32, order="C", copy=True).reshape(-1, 13)
    e = np.ascontiguousarray(np.broadcast_to(np.asarray(restitution, dtype=F32), (st.shape[0],)))
    check(lib().smcphys_sim_step(device, shape, S, ceiling_half, st.shape[0], _fp(st), _fp(e)))
    return st


def simulate(shape, S, init_pos, init_vel, e, xt, sigma_tr=SIGMA_TR, ceiling_half=0.01, device=0, want_mu=True):
    """Re-simulate n particles from t=0 over retained noisy positions xt[T,3,n].

    Returns mu[T,3,n] float32, logp[n] float64, state[n,13] float32 (see smcphys_simulate)."""
    xt = np.ascontiguousarray(xt, dtype=F32)
    e = np.ascontiguousarray(e, dtype=F32)
    T, three, n = xt.shape
    assert three == 3 and e.shape == (n,)
    mu = np.zeros((T, 3, n), F32) if want_mu else None
    logp = np.zeros(n)
    st = np.zeros((n, 13), F32)
    ip = np.ascontiguousarray(init_pos, dtype=np.float64)
    iv = np.ascontiguousarray(init_vel, dtype=np.float64)
    check(lib().smcphys_simulate(device, shape, S, ceiling_half, sigma_tr, _dp(ip), _dp(iv), _fp(e), _fp(xt), n, T,
                                 _fp(mu), _dp(logp), _fp(st)))
    return mu, logp, st


def resample_systematic(log_w, u0, device=0):
    lw = np.ascontiguousarray(log_w, dtype=np.float64)
    idx = np.zeros(lw.shape[0], np.int64)
    check(lib().smcphys_resample_systematic(device, _dp(lw), lw.shape[0], float(u0), _ip(idx)))
    return idx


def resample_multinomial(log_w, seed, t, device=0):
    lw = np.ascontiguousarray(log_w, dtype=np.float64)
    idx = np.zeros(lw.shape[0], np.int64)
    check(lib().smcphys_resample_multinomial(device, _dp(lw), lw.shape[0], int(seed), int(t), _ip(idx)))
    return idx


def weight_stats(log_w, device=0):
    lw = np.ascontiguousarray(log_w, dtype=np.float64)
    out = np.zeros(3)
    sums = np.zeros(3, np.uint64)
    check(lib().smcphys_weight_stats(device, _dp(lw), lw.shape[0], _dp(out), sums.ctypes.data_as(_lib._up)))
    W, S1, S2 = (int(x) for x in sums)
    return dict(mx=out[0], ess=out[1], lse=out[2], W=W, S1=S1, S2=S2,
                below_half=bool(lib().smcphys_ess_below(S1, S2, lw.shape[0], 1, 2)))


def default_config(**kw):
    cfg = Config()
    check(lib().smcphys_config_default(C.byref(cfg)))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError("smcphys_config has no field %r" % k)
        setattr(cfg, k, v)
    if "global_particles" not in kw:
        cfg.global_particles = cfg.num_particles
    return cfg


SYSTEMATIC, MULTINOMIAL = 0, 1


class ParticleFilter:
    """One smcphys handle: the per-trial filter state run_inference.jl:33-48 sets up.

    devices=[d0, d1, ...] makes it ONE filter sharded over several devices of this process (smcphys_create_multi,
    num_particles per device); comm_init() joins the handles of several processes into one filter (NCCL)."""

    def __init__(self, devices=None, **cfg):
        self.cfg = default_config(**cfg)
        self._h = _lib._H()
        if devices is None:
            check(lib().smcphys_create(C.byref(self.cfg), C.byref(self._h)))
        else:
            devs = (C.c_int32 * len(devices))(*[int(d) for d in devices])
            check(lib().smcphys_create_multi(C.byref(self.cfg), len(devices), devs, C.byref(self._h)))
        self.T = 0

    @staticmethod
    def comm_unique_id():
        """128 bytes that rank 0 makes and ships to the other ranks (any transport) before comm_init."""
        buf = C.create_string_buffer(128)
        check(lib().smcphys_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, unique_id, nranks, rank):
        """Collective: NCCL communicator + CUDA IPC mapping of the peers' buffers; run_smc is collective afterwards."""
        check(lib().smcphys_comm_init_rank(self._h, C.create_string_buffer(bytes(unique_id), 128), nranks, rank))

    def close(self):
        if getattr(self, "_h", None):
            lib().smcphys_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def n(self):
        return int(lib().smcphys_total_particles(self._h))

    def set_stream(self, cuda_stream_ptr):
        check(lib().smcphys_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def set_trial(self, init_pos, init_vel, obs):
        ip = np.ascontiguousarray(init_pos, dtype=np.float64)
        iv = np.ascontiguousarray(init_vel, dtype=np.float64)
        ob = np.ascontiguousarray(obs, dtype=np.float64).reshape(-1, 3)
        self.T = ob.shape[0]
        check(lib().smcphys_set_trial(self._h, _dp(ip), _dp(iv), _dp(ob), self.T))

    def run_smc(self, R=-1, want_stats=True):
        stats = (StepStats * max(self.T, 1))() if want_stats else None
        check(lib().smcphys_run_smc(self._h, R, stats))
        if not want_stats:
            return None
        return [dict(ess=s.ess, lse=s.lse, accept_rate=s.accept_rate, resampled=bool(s.resampled))
                for s in stats[:self.T]]

    def set_profiling(self, on=True):
        check(lib().smcphys_set_profiling(self._h, int(bool(on))))

    def kernel_time(self, kclass):
        """(total ms, launch groups) of kernel class `kclass` since the last query (synchronises)."""
        ms, cnt = C.c_double(), C.c_int64()
        check(lib().smcphys_kernel_time(self._h, kclass, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def synchronize(self):
        check(lib().smcphys_synchronize(self._h))

    def get_particles(self, want_score=True, want_traj=False, steps=None):
        # the library writes one trajectory row per PF step applied so far; `steps` is kept for the callers' sake
        n, T = self.n, int(lib().smcphys_steps_done(self._h))
        if steps is not None and steps != T:
            raise ValueError("get_particles(steps=%d): the filter has applied %d steps" % (steps, T))
        e, lw = np.zeros(n), np.zeros(n)
        score = np.zeros(n) if want_score else None
        traj = np.zeros((T, 3, n), F32) if want_traj else None
        check(lib().smcphys_get_particles(self._h, _dp(e), _dp(lw), _dp(score), _fp(traj)))
        return dict(e=e, log_w=lw, score=score, traj=traj)

    def get_states(self):
        st = np.zeros((self.n, 13), F32)
        check(lib().smcphys_get_states(self._h, _fp(st)))
        return st

    def get_log_ml(self):
        v = C.c_double()
        check(lib().smcphys_get_log_ml(self._h, C.byref(v)))
        return v.value

    def predict(self, n_steps, want_obs=True):
        pos = np.zeros((n_steps, 3, self.n), F32)
        obs = np.zeros((n_steps, 3, self.n), F32) if want_obs else None
        check(lib().smcphys_predict(self._h, n_steps, _fp(pos), _fp(obs)))
        return pos, obs

    # ---- phases (sharded filter); d_* are raw device pointers (ints)
    def pf_init(self):
        check(lib().smcphys_pf_init(self._h))

    def pf_extend(self, t, d_local_max=None):
        check(lib().smcphys_pf_extend(self._h, t, C.c_void_p(d_local_max)))

    def pf_weight_sums(self, d_global_max, d_sums3):
        check(lib().smcphys_pf_weight_sums(self._h, C.c_void_p(d_global_max), C.c_void_p(d_sums3)))

    def pf_payload_words(self, t):
        return int(lib().smcphys_pf_payload_words(self._h, t))

    def pf_pack(self, t, d_global_max, W_before, W_total, u0, k_lo, k_hi, slots_per_rank, d_send):
        check(lib().smcphys_pf_pack(self._h, t, C.c_void_p(d_global_max), W_before, W_total, u0, k_lo, k_hi,
                                    slots_per_rank, C.c_void_p(d_send)))

    def pf_gather_local(self, t, d_global_max, W_before, W_total, u0, k_lo, k_hi, slot0):
        check(lib().smcphys_pf_gather_local(self._h, t, C.c_void_p(d_global_max), W_before, W_total, u0, k_lo, k_hi,
                                            slot0))

    def pf_unpack(self, t, d_recv, counts, slot0, word_off):
        c = np.ascontiguousarray(counts, dtype=np.int64)
        s = np.ascontiguousarray(slot0, dtype=np.int64)
        w = np.ascontiguousarray(word_off, dtype=np.int64)
        check(lib().smcphys_pf_unpack(self._h, t, C.c_void_p(d_recv), len(c), _ip(c), _ip(s), _ip(w)))

    def pf_resample_local(self, t):
        check(lib().smcphys_pf_resample_local(self._h, t))

    def pf_rejuvenate(self, t, R, d_accept=None, move0=0):
        check(lib().smcphys_pf_rejuvenate_from(self._h, t, R, move0, C.c_void_p(d_accept)))


K_EXTEND, K_WEIGHTS, K_GATHER, K_REJUVENATE = 0, 1, 2, 3


def fp32_peak(device=0):
    """Measured FP32 FMA throughput (TFLOP/s) of the device: roofline denominator of the simulate kernels."""
    v = C.c_double()
    check(lib().smcphys_fp32_peak(device, C.byref(v)))
    return v.value


def offspring_range(W_before, W_local, W_total, N, u0):
    lo, hi = C.c_int64(), C.c_int64()
    check(lib().smcphys_pf_offspring_range(W_before, W_local, W_total, N, u0, C.byref(lo), C.byref(hi)))
    return lo.value, hi.value
