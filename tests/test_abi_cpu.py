"""CPU-only checks of the boundary: the library loads, exports every symbol include/smcphys.h declares, and refuses
to compute without a device (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import __graft_entry__ as entry

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ECUDA = -2


@pytest.fixture(scope="module")
def L():
    entry.build()
    from smc_physics_b200 import _lib
    return _lib.lib()


def test_every_header_symbol_is_exported_and_bound(L):
    from smc_physics_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "smcphys.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(smcphys_[a-z_0-9]+)\s*\(", hdr))
    assert len(declared) >= 30
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(L, name), name


def test_config_default_mirrors_args_jl(L):
    from smc_physics_b200 import _lib
    cfg = _lib.Config()
    assert L.smcphys_config_default(C.byref(cfg)) == 0
    assert cfg.struct_size == C.sizeof(_lib.Config)
    assert cfg.num_particles == 20 and cfg.rejuvenation_moves == 1            # args.jl:19-20
    assert cfg.observation_noise == float(np.float32(0.05))                    # args.jl:14
    assert cfg.transition_noise == float(np.float32(0.075))                    # args.jl:15
    assert cfg.proposal_sigma == 0.1 and (cfg.ess_num, cfg.ess_den) == (1, 2)  # particle_filter.jl:23,55


def test_host_side_integer_helpers_match_oracle(L):
    from oracle import oracle as orc
    rng = np.random.default_rng(3)
    for n in (1, 2, 7, 1000, 4097):
        lw = rng.normal(0, 4, n)
        ws = orc.weight_stats(lw)
        assert bool(L.smcphys_ess_below(ws["S1"], ws["S2"], n, 1, 2)) == ws["below_half"]
        assert L.smcphys_ess_value(ws["S1"], ws["S2"]) == ws["ess"]
        assert L.smcphys_lse_value(ws["mx"], ws["W"]) == ws["lse"]
    # resampling offset = first Philox word of (seed; 0, t, 0, ST_RESAMPLE) / 2^32
    for seed, t in ((1234, 0), (7, 55), (2**40 + 3, 119)):
        assert L.smcphys_resample_offset(seed, t) == orc.philox(seed, 0, t, 0, 3)[0] / 2.0**32


def test_offspring_ranges_tile_the_slots(L):
    """Shard k-ranges from smcphys_pf_offspring_range must agree with the oracle's global indices."""
    from oracle import oracle as orc
    from smc_physics_b200 import offspring_range
    rng = np.random.default_rng(5)
    for N, G in ((64, 2), (1000, 4), (4096, 8)):
        lw = rng.normal(0, 3, N)
        u0 = float(rng.uniform())
        idx = orc.resample_systematic(lw, u0)
        q = orc.weight_q(lw, lw.max()).astype(object)
        per = N // G
        Wl = [int(sum(q[r * per:(r + 1) * per])) for r in range(G)]
        Wt = sum(Wl)
        prev = 0
        for r in range(G):
            lo, hi = offspring_range(sum(Wl[:r]), Wl[r], Wt, N, u0)
            assert lo == prev
            owners = idx[lo:hi] // per
            assert np.all(owners == r)
            prev = hi
        assert prev == N


@pytest.mark.skipif(os.environ.get("SMCPHYS_EXPECT_GPU") == "1", reason="GPU box")
def test_no_cpu_fallback(L):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    import smc_physics_b200 as sp
    with pytest.raises(sp.SmcphysError) as ei:
        sp.ParticleFilter(num_particles=8)
    assert ei.value.code == -2
    with pytest.raises(sp.SmcphysError):
        sp.resample_systematic(np.zeros(4), 0.5)
    with pytest.raises(sp.SmcphysError):
        sp.sim_step(sp.SPHERE, np.zeros((1, 13), np.float32), 0.5)


def test_bad_arguments_are_reported_not_thrown(L):
    from smc_physics_b200 import _lib
    cfg = _lib.Config()
    L.smcphys_config_default(C.byref(cfg))
    cfg.struct_size = 4
    h = _lib._H()
    assert L.smcphys_create(C.byref(cfg), C.byref(h)) == -1
    assert b"size mismatch" in L.smcphys_last_error()
    L.smcphys_config_default(C.byref(cfg))
    cfg.shape = 5
    assert L.smcphys_create(C.byref(cfg), C.byref(h)) == -1
    assert L.smcphys_destroy(None) == 0
    assert L.smcphys_set_trial(None, None, None, None, 0) == -1


def test_multi_gpu_entries_fail_cleanly_without_a_device(L):
    """The in-library sharded filter (smcphys_comm_* / smcphys_create_multi): argument checks and no CPU path."""
    from smc_physics_b200 import _lib
    cfg = _lib.Config()
    L.smcphys_config_default(C.byref(cfg))
    h = _lib._H()
    devs = (C.c_int32 * 2)(0, 0)
    rc = L.smcphys_create_multi(C.byref(cfg), 2, devs, C.byref(h))
    assert rc in (_ECUDA, 0)
    if rc == 0:
        L.smcphys_destroy(h)
    assert L.smcphys_create_multi(C.byref(cfg), 0, devs, C.byref(h)) == -1            # n_shards out of range
    assert L.smcphys_create_multi(C.byref(cfg), 65, devs, C.byref(h)) == -1
    assert L.smcphys_comm_init_rank(None, b"\0" * 128, 2, 0) == -1                       # NULL handle
    assert L.smcphys_comm_unique_id(None) == -1
    cfg.resampler = 7
    assert L.smcphys_create(C.byref(cfg), C.byref(h)) == -1                              # unknown resampler
    assert L.smcphys_total_particles(None) == 0 and L.smcphys_steps_done(None) == -1


def test_nccl_is_bound_at_run_time(L):
    """libsmcphys.so must load without NCCL on the link line (dlopen at first use)."""
    import subprocess
    so = os.path.join(ROOT, "smc_physics_b200", "libsmcphys.so")
    needed = subprocess.run(["readelf", "-d", so], capture_output=True, text=True).stdout
    assert "libnccl" not in needed
    buf = C.create_string_buffer(128)
    rc = L.smcphys_comm_unique_id(buf)
    assert rc in (0, _ECUDA)          # 0 where a libnccl.so.2 is installed; a clean error where it is not
    if rc == 0:
        assert any(b != 0 for b in buf.raw)
