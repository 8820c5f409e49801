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
    ap.add_argument("--config", type=int, default=-1, choices=[-1, 0, 1, 2, 3, 4],
                    help="BASELINE.json configs[k]: 0 BulletTest_Sphere N=20 R=2 (27 trials), 1 BulletTest_Cube N=1000 R=5 "
                         "(27 trials), 2 RealFlow Exp1-4 sphere 100k particles/trial (668 trials, S=9), 3 synthetic sphere "
                         "2^20 x 120 R=5 (the default), 4 synthetic cube 2^21 per GPU (16M on 8)")
    ap.add_argument("--trial-particles", type=int, default=0, help="configs 0-2: particles per trial (0 = the config's)")
    ap.add_argument("--max-trials", type=int, default=0, help="configs 0-2: first K trials only (0 = all)")
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


def cpu_leg(args, shape_id, trial, steps, warmup, threads=0):
    """The CPU restatement (oracle/) on a bounded particle sample of the same trial, all host threads (threads=1: one
    thread in the reference's loop order, Inference/particle_filter.jl:48-76).
    The only place outside tests/ and smoke() that executes oracle/ — as the reported baseline, never the product."""
    from oracle import oracle as orc
    orc.build()
    cores = threads or os.cpu_count() or 1
    ip, iv, obs = trial
    T = obs.shape[0]
    n = args.cpu_particles
    if n <= 0:
        # calibrate: a 64-particle-per-thread run, then size the sample for ~8 s per step
        n0 = 64 * cores
        t0 = time.perf_counter()
        orc.run_smc(shape_id, n0, args.moves, args.substeps, ip, iv, obs, seed=99, threads=cores)
        dt = max(time.perf_counter() - t0, 1e-3)
        n = int(max(n0, min(1 << 16, n0 * (8.0 if not threads else 4.0) / dt)))
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
            "ms_per_step": sec * 1e3, "steps": len(timed), "sample_particles": n}


def kernel_source_hash():
    """sha256 (first 16 hex) of the code k_rejuvenate is compiled from: physics.cuh, rng.cuh and smcphys.cu up to the end
    of k_rejuvenate (the section marker that follows it; #include lines left out).  Stamps ncu-derived numbers so that a
    capture of another kernel body is not reported, while changes elsewhere (entry points, scratch management, the
    kernels of the sharded exchange) leave the stamp valid."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "smc_physics_b200", "csrc")
    for f in ("physics.cuh", "rng.cuh"):
        h.update(open(os.path.join(d, f), "rb").read())
    kern = open(os.path.join(d, "smcphys.cu")).read().split("// ---------------------------------------------------------------- getters / stateless kernels")[0]
    h.update("\n".join(l for l in kern.split("\n") if not l.startswith("#include")).encode())
    return h.hexdigest()[:16]


def config_of(args, n_gpus):
    return {"workload": "synthetic %s bounce, %d particles/GPU x %d timesteps, rejuvenation_moves=%d, substeps=%d "
                        "(BASELINE.json configs[%d])" % (args.shape, args.particles, args.timesteps, args.moves,
                                                        args.substeps, 3 if args.shape == "sphere" else 4),
            "particles_per_gpu": args.particles, "global_particles": args.particles * n_gpus,
            "timesteps": args.timesteps, "rejuvenation_moves": args.moves, "substeps": args.substeps,
            "shape": args.shape, "sharding": "particles over ranks" if n_gpus > 1 else "single shard",
            "l2": "working set (retained-position history, %.2f GB per GPU) exceeds the 126 MB L2" %
                  (args.particles * args.timesteps * 12 / 1e9),
            "cpu_sample_particles": getattr(args, "cpu_sample_particles", None)}


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
    args.cpu_sample_particles = cb["sample_particles"]
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
    # measured DRAM traffic of one k_rejuvenate launch (t = 60), from the committed ncu capture of this same workload.
    # The entry is stamped with a hash of the kernel sources it was captured from: a stale capture gives null.
    traffic, traffic_note = None, "no ncu capture of this workload under profiles/traffic.json"
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_rejuvenate_" + args.shape]
        c = tr["config"]
        if (c["particles"], c["timesteps"], c["moves"], c["substeps"]) == (n, T, R, S):
            if tr.get("src_sha") == kernel_source_hash():
                traffic = tr["dram_bytes"]
                traffic_note = ("dram__bytes_read+write of the t=%d launch (ncu --set full, %s); algorithmic bytes of that "
                                "launch = R*12*(t+1)*n = %.3g" % (tr["launch_t"], tr["source"], R * 12 * (tr["launch_t"] + 1) * n))
            else:
                traffic_note = "profiles/traffic.json was captured from other kernel sources (%s): not reported" % tr.get("src_sha")
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
                     "traffic_note": traffic_note,
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
        line["config"]["cpu_sample_particles"] = cb["sample_particles"]
        c1 = cpu_leg(args, shape_id, trial, 1, 0, threads=1)
        line["cpu_baseline_1thread"] = {k: c1[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------- BASELINE.json configs[0..2]: many independent trials
def load_trials(args, sp, local):
    """-> (name, shape_id, S, R, N, [(init_pos, init_vel, obs[T,3])]) of configs 0-2.  configs 0/1: the 27 BulletTest
    stimuli generated by Tests/generate_trajectories.jl on the CUDA simulator (host.generate_trajectories, seed
    20240830, S = 5); config 2: the 668 RealFlow trials of Exp1-4 (tests/golden/Data, data only; S = 9)."""
    import tempfile
    import numpy as np
    from smc_physics_b200 import host
    if args.config in (0, 1):
        shape_name = "Sphere" if args.config == 0 else "Cube"
        N, R, S = (20, 2, 5) if args.config == 0 else (1000, 5, 5)
        old = host.DATA_ROOT
        host.DATA_ROOT = tempfile.mkdtemp(prefix="smcphys_bench_")
        try:
            names = host.generate_trajectories(shape_name, seed=20240830, device=local)
            d = os.path.join(host.DATA_ROOT, "Data/BulletStimulus/New", shape_name)
            trials = []
            for f in names:
                ip, _, iv, obs, T = host.read_obs_file(os.path.join(d, f), "Bullet")
                trials.append((np.asarray(ip, float), np.asarray(iv, float), obs))
        finally:
            host.DATA_ROOT = old
        name = "BulletTest_%s, Modelv5 %s, SMC num_particles=%d, rejuvenation_moves=%d, 27 generated trials in flight" % (
            shape_name, shape_name, N, R)
        return name, (sp.SPHERE if args.config == 0 else sp.CUBE), S, R, N, trials
    N, R, S = 100000, 5, 9
    trials = []
    for ex in ("Exp1", "Exp2", "Exp3", "Exp4"):
        d = os.path.join(ROOT, "tests", "golden", "Data", "RealFlowData", ex)
        for f in sorted(host.filter_unwanted_filenames(os.listdir(d)), key=host.trial_order):
            ip, _, iv, obs, T = host.read_obs_file(os.path.join(d, f))
            trials.append((np.asarray(ip, float), np.asarray(iv, float), obs))
    return "RealFlow Exp1-4 batch, Sphere, 100k particles per trial, all trials concurrent", sp.SPHERE, S, R, N, trials


def trials_arm(args):
    """configs 0-2: a step = every trial of the experiment through run_smc once, all trials of a wave in flight (one
    handle + stream per trial; the analogue of the reference's pmap over trial files, main.jl:54).  value = sum N T /
    time.  Over several GPUs the trial list is dealt round robin to the ranks, no communication ("replicas")."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import smc_physics_b200 as sp
    from smc_physics_b200 import workloads as wl
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    name, shape_id, S, R, N, trials = load_trials(args, sp, local)
    if args.trial_particles:
        N = args.trial_particles
    if args.max_trials:
        trials = trials[:args.max_trials]
    mine = trials[rank::world]
    lib = sp.lib()
    # waves sized by device memory: two history buffers + state per particle
    budget, waves, cur, used = 120e9, [], [], 0.0
    for tr in mine:
        need = N * (24.0 * max(tr[2].shape[0], 1) + 300.0)
        if cur and used + need > budget:
            waves.append(cur); cur, used = [], 0.0
        cur.append(tr); used += need
    if cur:
        waves.append(cur)

    def one_pass(e2e, prof=False):
        """-> (seconds on the device clock of this rank, posterior means, k_rejuvenate ms)"""
        t_dev, means, rej_ms = 0.0, [], 0.0
        for wave in waves:
            pfs = [sp.ParticleFilter(num_particles=N, shape=shape_id, substeps=S, max_timesteps=max(tr[2].shape[0], 1),
                                     rejuvenation_moves=R, device=local, seed=1234 + 7919 * k) for k, tr in enumerate(wave)]
            if not e2e:
                for pf, tr in zip(pfs, wave):
                    pf.set_trial(*tr)
                    pf.set_profiling(prof)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for pf, tr in zip(pfs, wave):
                if e2e:
                    pf.set_trial(*tr)
                pf.run_smc(R, want_stats=False)
            if e2e:
                for pf in pfs:
                    means.append(float(pf.get_particles(want_score=False)["e"].mean()))
            else:
                for pf in pfs:
                    pf.synchronize()
            t_dev += time.perf_counter() - t0
            for pf in pfs:
                if prof:
                    rej_ms += pf.kernel_time(sp.K_REJUVENATE)[0]
                pf.close()
        return t_dev, means, rej_ms

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=op)
        return float(t.item())

    for _ in range(max(args.warmup, 1)):
        one_pass(False)
    clocks = ClockSampler(local)
    barrier()
    if rank == 0:
        clocks.start()
    l0 = lib.smcphys_launch_count()
    t_res = 0.0
    for _ in range(args.steps):
        t_res += one_pass(False)[0]
    barrier()
    launches = lib.smcphys_launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    t_res = reduce(t_res / args.steps, dist.ReduceOp.MAX if world > 1 else None)
    rej_ms = one_pass(False, prof=True)[2]
    barrier()
    t_e2e, means = 0.0, []
    for _ in range(args.steps):
        dt, means, _ = one_pass(True)
        t_e2e += dt
    barrier()
    t_e2e = reduce(t_e2e / args.steps, dist.ReduceOp.MAX if world > 1 else None)
    units = float(N * sum(tr[2].shape[0] for tr in trials))
    kst = float(sum(wl.kernel_steps(N, tr[2].shape[0], R) for tr in trials))
    flop = float(sum(sum(wl.flop_rejuvenate(shape_id, S, N, R, t) for t in range(tr[2].shape[0])) for tr in mine))
    h2d = sum(tr[2].size * 8 + 48 for tr in trials)
    d2h = len(trials) * N * 16
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak_tf = sp.fp32_peak(local)
    ach = flop / t_res / 1e12
    line = {"metric": METRIC, "value": units / t_res, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": t_res * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic" if args.config < 2 else "reference stimuli (RealFlow Exp1-4)",
            "config": {"workload": name + " (BASELINE.json configs[%d])" % args.config, "trials": len(trials),
                       "particles_per_trial": N, "frames_total": int(sum(tr[2].shape[0] for tr in trials)),
                       "rejuvenation_moves": R, "substeps": S, "shape": "sphere" if shape_id == sp.SPHERE else "cube",
                       "sharding": "trials dealt to ranks, no communication (replicas)" if world > 1 else "single GPU",
                       "waves": len(waves), "l2": "every run re-creates its filters; histories %s the 126 MB L2" % (
                           "exceed" if N * 24 * 60 * len(waves[0]) > 126e6 else "fit in")},
            "clocks": clk,
            "e2e": {"value": units / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": t_e2e * 1e3},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "k_rejuvenate", "bound": "fp32", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": ach / peak_tf if peak_tf else None, "traffic": None,
                         "note": "trials overlap on the device, so per-kernel times are not additive: achieved = algorithmic "
                                 "FLOP of every k_rejuvenate launch / device time of the whole step (a lower bound); the sum "
                                 "of the per-launch event times is rejuvenate_event_ms",
                         "rejuvenate_event_ms": rej_ms,
                         "peak_source": "FP32 FMA probe run in this process (smcphys_fp32_peak)"},
            "kernel_steps_per_s": kst / t_res, "posterior_e_mean": float(np.mean(means)) if means else None}
    if not args.no_cpu_baseline and world == 1:
        # bounded sample: the first trials of the list with the config's particle count (capped), both thread modes
        from oracle import oracle as orc
        orc.build()
        cores = os.cpu_count() or 1
        ncpu = min(N, 2000)
        for key, th in (("cpu_baseline", cores), ("cpu_baseline_1thread", 1)):
            t0, done, k = time.perf_counter(), 0.0, 0
            while k < len(trials) and time.perf_counter() - t0 < (12.0 if th > 1 else 8.0):
                ip, iv, obs = trials[k]
                orc.run_smc(shape_id, ncpu, R, S, ip, iv, obs, seed=1234 + k, threads=th)
                done += ncpu * obs.shape[0]
                k += 1
            sec = time.perf_counter() - t0
            line[key] = {"value": done / sec, "unit": UNIT, "cores": th, "kind": "port",
                         "sample": "first %d trials at %d particles each, R=%d, S=%d; oracle/smc_oracle.hpp float64, %d "
                                   "thread(s) over particles" % (k, ncpu, R, S, th)}
        line["config"]["cpu_sample_particles"] = ncpu
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.config == 3:
        a.shape, a.particles, a.timesteps, a.moves, a.substeps = "sphere", 1 << 20, 120, 5, 5
    if a.config == 4:
        a.shape, a.particles, a.timesteps, a.moves, a.substeps = "cube", 1 << 21, 120, 5, 5
    if a.impl == "reference":
        reference_arm(a)
    elif a.config in (0, 1, 2):
        trials_arm(a)
    else:
        gpu_arm(a)
