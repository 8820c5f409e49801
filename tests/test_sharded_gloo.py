"""N > 1 path on CPU: the sharded driver's host logic (ESS decision from the gathered integer sums, offspring
ranges, chunk planning, the all-reduce / all-gather / all-to-all exchanges) over gloo with world_size 2 and 3,
each rank holding a CPU stand-in shard (tests/_cpu_shard.py).  The resampled lineage must be bit-identical to
the single-shard oracle resampling."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, T, seed, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from _cpu_shard import CpuShard
        from smc_physics_b200.sharded import PhaseShardedFilter
        f = PhaseShardedFilter(particles_per_rank=n, seed=seed, shard=CpuShard(n, rank, world))
        f.T = T
        stats = f.run_smc(R=0)
        sh = f.shard
        q.put((rank, sh.ident.copy(), sh.e.copy(), sh.lw.copy(), [s["resampled"] for s in stats], f.migrated_particles,
               f.log_ml))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 300), (3, 257)])
def test_sharded_host_logic_matches_single_shard(world, n):
    from oracle import oracle as orc  # noqa: F401  (builds liboracle.so before the workers race for it)
    orc.lib()
    import smc_physics_b200 as sp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _cpu_shard import serial_reference
    T, seed = 14, 77
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, T, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = sorted([q.get(timeout=180) for _ in procs], key=lambda o: o[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ident = np.concatenate([o[1] for o in outs])
    e = np.concatenate([o[2] for o in outs])
    lw = np.concatenate([o[3] for o in outs])
    ref_ident, ref_e, ref_lw, ref_res = serial_reference(n * world, T, seed,
                                                         lambda s, t: sp.lib().smcphys_resample_offset(s, t))
    assert sum(ref_res) >= 2, "the synthetic weights must trigger resampling for the test to mean anything"
    for o in outs:
        assert o[4] == ref_res                    # every rank took the same decisions as the single shard
        assert abs(o[6] - outs[0][6]) == 0.0      # and holds the same marginal-likelihood estimate
    assert np.array_equal(ident, ref_ident)       # bit-identical lineage
    assert np.array_equal(e, ref_e)
    assert np.array_equal(lw, ref_lw)
    assert sum(o[5] for o in outs) > 0            # particles really crossed ranks


def test_plan_migration_tiles_every_slot():
    from smc_physics_b200.sharded import plan_migration
    rng = np.random.default_rng(0)
    for G in (1, 2, 4, 8):
        spr = 100
        N = G * spr
        cuts = np.sort(rng.integers(0, N + 1, G - 1))
        bounds = [0, *cuts.tolist(), N]
        ranges = [(bounds[i], bounds[i + 1]) for i in range(G)]
        plans = [plan_migration(ranges, spr, r) for r in range(G)]
        for d in range(G):
            send_to_d = [plans[s][0][d] for s in range(G)]
            assert send_to_d == plans[d][1]                    # what s sends to d is what d expects from s
            assert sum(plans[d][1]) == spr                     # every local slot is filled exactly once
            filled = np.zeros(spr, int)
            for c, s0 in zip(plans[d][1], plans[d][2]):
                filled[s0:s0 + c] += 1
            assert np.all(filled == 1)
