"""bench.py — particle-steps/s (incl. rejuvenation) of the SMC hot path on N B200s, one JSON line on rank 0.

A "step" is one complete SMC run (run_smc, Inference/particle_filter.jl:37-81) over one synthetic trial:
T PF steps, each = extend + weight, ESS-triggered resampling, R MH rejuvenation moves per particle.
Workload (BASELINE.json configs[3]): synthetic sphere bounce, 1M particles per GPU x 120 timesteps, R = 5, S = 5.
`value`  = N_global * T / time, trial resident on the device, timed with CUDA events on the library's stream.
`e2e`    = the same through the public host-buffer API: set_trial (H2D of the observations) + run_smc +
           get_particles (D2H of elasticities and log weights) inside the timed region.
`--impl reference` times the CPU restatement (oracle/, all host threads) on a bounded particle sample of the
same trial; the reference's own Julia + Gen.jl + PyBullet stack cannot be installed in this image.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "particle-steps/sec incl. rejuvenation"
UNIT = "particle-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--shape", default="sphere", choices=["sphere", "cube"])
    ap.add_argument("--particles", type=int, default=1 << 20, help="particles per GPU")
    ap.add_argument("--timesteps", type=int, default=120)
    ap.add_argument("--moves", type=int, default=5)
    ap.add_argument("--substeps", type=int, default=5)
    ap.add_argument("--cpu-particles", type=int, default=0, help="particle sample of the CPU legs (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thr.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_leg(args, shape_id, trial, steps, warmup):
    """The CPU restatement (oracle/) on a bounded particle sample of the same trial, all host threads.
    The only place outside tests/ and smoke() that executes oracle/ — as the reported baseline, never the product."""
    from oracle import oracle as orc
    orc.build()
    cores = os.cpu_count() or 1
    ip, iv, obs = trial
    T = obs.shape[0]
    n = args.cpu_particles
    if n <= 0:
        # calibrate: a 64-particle-per-thread run, then size the sample for ~8 s per step
        n0 = 64 * cores
        t0 = time.perf_counter()
        orc.run_smc(shape_id, n0, args.moves, args.substeps, ip, iv, obs, seed=99, threads=cores)
        dt = max(time.perf_counter() - t0, 1e-3)
        n = int(max(n0, min(1 << 16, n0 * 8.0 / dt)))
        n -= n % cores
    times = []
    for k in range(warmup + steps):
        t0 = time.perf_counter()
        orc.run_smc(shape_id, n, args.moves, args.substeps, ip, iv, obs, seed=1234 + k, threads=cores)
        times.append(time.perf_counter() - t0)
        if k >= warmup and sum(times) > 150:  # keep the whole leg within a few minutes
            break
    timed = times[warmup:] or times[-1:]
    sec = sum(timed) / len(timed)
    return {"value": n * T / sec, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d particles x %d timesteps, R=%d, S=%d, same trial; oracle/smc_oracle.hpp float64, %d threads "
                      "over particles; %d timed runs" % (n, T, args.moves, args.substeps, cores, len(timed)),
            "ms_per_step": sec * 1e3, "steps": len(timed)}


def config_of(args, n_gpus):
    return {"workload": "synthetic %s bounce, %d particles/GPU x %d timesteps, rejuvenation_moves=%d, substeps=%d "
                        "(BASELINE.json configs[%d])" % (args.shape, args.particles, args.timesteps, args.moves,
                                                        args.substeps, 3 if args.shape == "sphere" else 4),
            "particles_per_gpu": args.particles, "global_particles": args.particles * n_gpus,
            "timesteps": args.timesteps, "rejuvenation_moves": args.moves, "substeps": args.substeps,
            "shape": args.shape, "sharding": "particles over ranks" if n_gpus > 1 else "single shard",
            "l2": "working set (retained-position history, %.2f GB per GPU) exceeds the 126 MB L2" %
                  (args.particles * args.timesteps * 12 / 1e9)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    shape_id = 0 if args.shape == "sphere" else 1
    from oracle import oracle as orc
    orc.build()
    ip, iv = np.array([0.1, -0.2, 0.6]), np.array([3.0, -2.0, 1.5])
    truth = orc.macro_rollout(shape_id, 0.6, orc.state13(ip, vel=iv), args.timesteps, S=args.substeps)[:, 0:3]
    obs = np.vstack([ip[None], truth[:-1]])
    cb = cpu_leg(args, shape_id, (ip, iv, obs), args.steps, min(args.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": cb["steps"], "warmup": min(args.warmup, 1), "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(args, args.gpus),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference Julia+Gen.jl+PyBullet stack is not installable here (no julia, no pybullet); this is "
                    "the CPU restatement of the same algorithm on the host cores"}
    print(json.dumps(line), flush=True)


def gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import smc_physics_b200 as sp
    from smc_physics_b200 import workloads as wl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d"
                         % (args.gpus, world, args.gpus))
    if sp.lib().smcphys_device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: libsmcphys has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    shape_id = sp.SPHERE if args.shape == "sphere" else sp.CUBE
    n, T, R, S = args.particles, args.timesteps, args.moves, args.substeps
    trial = wl.synthetic_trial(shape_id, T=T, S=S, device=local)
    stream = torch.cuda.Stream(device=dev)

    if world == 1:
        pf = sp.ParticleFilter(num_particles=n, shape=shape_id, substeps=S, max_timesteps=T, rejuvenation_moves=R,
                               device=local, seed=1234)
        core = pf
    else:
        from smc_physics_b200.sharded import ShardedFilter
        pf = ShardedFilter(particles_per_rank=n, shape=shape_id, substeps=S, max_timesteps=T, rejuvenation_moves=R,
                           device=local, seed=1234, stream=stream)
        core = pf.pf
    core.set_stream(stream.cuda_stream)
    lib = sp.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- correctness guard of the multi-GPU path: a small fixed GLOBAL run (2^16 particles, 25 steps, R = 2) through
    # the same code path as the timed runs; its particles must be bit-identical for every N (integer weights, global
    # Philox counters).  For N > 1 rank 0 also runs it unsharded on its own GPU and the bench FAILS on a mismatch.
    import zlib
    CHK_N, CHK_T, CHK_R = 1 << 16, 25, 2
    chk_trial = wl.synthetic_trial(shape_id, T=CHK_T, S=S, device=local)

    def fingerprint(e, lw):
        return "%08x%08x" % (zlib.crc32(np.ascontiguousarray(e).tobytes()), zlib.crc32(np.ascontiguousarray(lw).tobytes()))

    if world == 1:
        with sp.ParticleFilter(num_particles=CHK_N, shape=shape_id, substeps=S, max_timesteps=CHK_T, device=local,
                               seed=1234) as cpf:
            cpf.set_trial(*chk_trial)
            cpf.run_smc(CHK_R, want_stats=False)
            o = cpf.get_particles(want_score=False)
        check = {"global_particles": CHK_N, "timesteps": CHK_T, "moves": CHK_R, "checksum": fingerprint(o["e"], o["log_w"]),
                 "matches_unsharded": True}
    else:
        from smc_physics_b200.sharded import ShardedFilter
        cf = ShardedFilter(particles_per_rank=CHK_N // world, shape=shape_id, substeps=S, max_timesteps=CHK_T,
                           rejuvenation_moves=CHK_R, device=local, seed=1234)
        cf.set_trial(*chk_trial)
        cf.run_smc(CHK_R, want_stats=False)
        o = cf.get_particles(want_score=False)
        parts = [torch.zeros(2, CHK_N // world, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(np.stack([o["e"], o["log_w"]])).to(dev))
        cf.close()
        check = None
        if rank == 0:
            allp = torch.cat(parts, dim=1).cpu().numpy()
            with sp.ParticleFilter(num_particles=CHK_N, shape=shape_id, substeps=S, max_timesteps=CHK_T, device=local,
                                   seed=1234) as cpf:
                cpf.set_trial(*chk_trial)
                cpf.run_smc(CHK_R, want_stats=False)
                o1 = cpf.get_particles(want_score=False)
            check = {"global_particles": CHK_N, "timesteps": CHK_T, "moves": CHK_R,
                     "checksum": fingerprint(allp[0], allp[1]),
                     "matches_unsharded": fingerprint(allp[0], allp[1]) == fingerprint(o1["e"], o1["log_w"])}
            if not check["matches_unsharded"]:
                raise SystemExit("bench.py: the %d-GPU filter is NOT bit-identical to the unsharded one on the check run "
                                 "(%s vs %s)" % (world, check["checksum"], fingerprint(o1["e"], o1["log_w"])))

    # ---- warm-up (also gives the per-step resampling record used for the gather bytes)
    pf.set_trial(*trial)
    stats = None
    for k in range(max(args.warmup, 3)):
        if k == 0:
            stats = pf.run_smc(R, want_stats=True)
        else:
            pf.run_smc(R, want_stats=False)
    core.synchronize()

    # ---- device-resident timing: K runs, CUDA events on the library's stream
    core.set_profiling(True)
    for c in range(4):
        core.kernel_time(c)  # drop anything recorded so far
    clocks = ClockSampler(local)
    barrier()
    if rank == 0:
        clocks.start()
    l0 = lib.smcphys_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        pf.run_smc(R, want_stats=False)
    ev1.record(stream)
    barrier()
    launches = lib.smcphys_launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    ktime = [core.kernel_time(c) for c in range(4)]
    core.set_profiling(False)
    ms_per_step = ms_total / args.steps
    n_global = n * world
    value = n_global * T / (ms_per_step * 1e-3)

    # ---- end to end through the host-buffer API
    barrier()
    t0 = time.perf_counter()
    e_mean = 0.0
    for _ in range(args.steps):
        pf.set_trial(*trial)
        pf.run_smc(R, want_stats=False)
        out = pf.get_particles(want_score=False)
        e_mean = float(out["e"].mean())
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / args.steps
    h2d = world * (trial[2].size * 8 + 6 * 8)
    d2h = world * (out["e"].nbytes + out["log_w"].nbytes)
    e_mean = sum_over_ranks(e_mean) / world

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (k_rejuvenate): algorithmic FLOP / measured launch time
    peak_tf = sp.fp32_peak(local)
    rej_ms, rej_groups = ktime[sp.K_REJUVENATE]
    flop_run = sum(wl.flop_rejuvenate(shape_id, S, n, R, t) for t in range(T))
    rej_tf = flop_run * args.steps / (rej_ms * 1e-3) / 1e12 if rej_ms > 0 else 0.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # measured DRAM traffic of the mean k_rejuvenate launch, from the committed ncu capture of this same workload
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_rejuvenate"]
        c = tr["config"]
        if (c["shape"], c["particles"], c["timesteps"], c["moves"], c["substeps"]) == (args.shape, n, T, R, S):
            traffic = tr["dram_bytes"]
    except (OSError, ValueError, KeyError):
        pass
    res_steps = [t for t, s in enumerate(stats or []) if s["resampled"]]
    gather_bytes = sum(n * (2 * (68 + 12 * (t + 1)) + 12) for t in res_steps) * args.steps
    g_ms = ktime[sp.K_GATHER][0]
    shares = {name: ktime[c][0] / ms_total for name, c in
              (("extend", sp.K_EXTEND), ("weights", sp.K_WEIGHTS), ("gather", sp.K_GATHER), ("rejuvenate", sp.K_REJUVENATE))}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": config_of(args, world), "clocks": clk,
        "e2e": {"value": n_global * T / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_s * 1e3},
        "gpu_launches": int(launches),
        "roofline": {"kernel": "k_rejuvenate", "bound": "fp32", "achieved": rej_tf, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": rej_tf / peak_tf if peak_tf else None, "traffic": traffic,
                     "traffic_note": "dram__bytes_read+write of the t=60 launch (ncu --set full, profiles/); algorithmic "
                                     "bytes of that launch = R*12*(t+1)*n = %.3g" % (R * 12 * 61 * n),
                     "peak_source": "FP32 FMA probe run in this process (smcphys_fp32_peak); not in MEASURED_PEAKS.json",
                     "launches": int(rej_groups), "avg_launch_ms": rej_ms / max(rej_groups, 1),
                     "flop_per_launch_avg": flop_run / T, "share_of_step": shares,
                     "hbm_gather": {"kernel": "k_gather+k_gather_hist", "bound": "hbm",
                                    "achieved": gather_bytes / (g_ms * 1e-3) / 1e9 if g_ms > 0 and gather_bytes else None,
                                    "peak": hbm_peak, "unit": "GB/s", "resampled_steps": len(res_steps),
                                    "note": "group time includes the skipped (non-resampling) launches"}},
        "kernel_steps_per_s": wl.kernel_steps(n_global, T, R) / (ms_per_step * 1e-3),
        "posterior_e_mean": e_mean,
        "check": check,
    }
    if not args.no_cpu_baseline and world == 1:
        cb = cpu_leg(args, shape_id, trial, 1, 0)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        reference_arm(a)
    else:
        gpu_arm(a)
