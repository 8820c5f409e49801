"""Host-side mirror of the reference's Julia interfaces for the particle-filter path.

Julia is not present in this image, so the host layer above the C ABI is written in Python with the reference's
names, argument meaning, side effects and file formats; julia/SmcPhysB200.jl carries the same layer as `ccall`
bindings for a Julia host.  What is mirrored, and from where (paths in the reference repo):

  Args                                   args.jl:5-27 (field names, types, defaults)
  read_obs_file / write_to_csv           Utilities/fileio.jl:20-66, 255-278 (CSV schemas, 5-digit truncation)
  generate_noise_id / ..._param_id       Utilities/fileio.jl:175-202
  filter_unwanted_filenames, trial_order Utilities/fileio.jl:12-17, 74-99
  run_smc, predict                       Inference/particle_filter.jl:37-81, 84-102
  run_inference                          Inference/run_inference.jl:19-124
  PhySim / PhyState / step / sync! / forward_step, BulletSim / RigidBodyState
                                         PhySMC/PhySMC.jl:13-49, PhySMC/PhyBullet/PhyBullet.jl:23-103,
                                         PhySMC/PhyBullet/Elements/rigid_body.jl:16-27

All arithmetic is done by libsmcphys on the GPU; nothing here simulates or weighs anything.
"""
import csv
import math
import os
from dataclasses import dataclass, field

import numpy as np

from . import api

F32 = np.float32


# ------------------------------------------------------------------ args.jl:5-27
@dataclass
class Args:
    expt_id: str = "BulletTest_Sphere"        # BulletTest_Cube, BulletTest_Sphere, Exp1, Exp2, ...
    model_id: str = "Modelv5"
    target_id: str = "Sphere"
    init_vel_noise: float = float(F32(0.0))
    observation_noise: float = float(F32(0.05))
    transition_noise: float = float(F32(0.075))
    algorithm: str = "SMC"                    # MCMC, SMC, or DEBUG
    num_particles: int = 20
    rejuvenation_moves: int = 1
    save_particles: bool = True
    predict: bool = False
    prediction_timesteps: int = 90
    # not in the reference struct: it reseeds nothing, so its runs are not reproducible
    seed: int = 1234
    device: int = 0


# ------------------------------------------------------------------ PhySMC interface
class PhySim:
    """PhySMC.PhySim (PhySMC/PhySMC.jl:13)."""


@dataclass
class CudaSim(PhySim):
    """Stands in for BulletSim (PhyBullet.jl:23-30): same field vocabulary, no PyBullet client."""
    step_dur: float = 1.0 / 60.0
    pb_timestep: float = 1.0 / 240.0
    shape: int = api.SPHERE
    ceiling_half: float = 0.01
    device: int = 0

    @property
    def substeps(self):
        return api.substeps_for(self.step_dur, self.pb_timestep)


BulletSim = CudaSim  # drop-in name


@dataclass
class RigidBodyState:
    """rigid_body.jl:16-27, batched: arrays of n bodies."""
    position: np.ndarray       # [n,3]
    orientation: np.ndarray    # [n,4] xyzw
    linear_vel: np.ndarray     # [n,3]
    angular_vel: np.ndarray    # [n,3]
    aabb: np.ndarray = None    # [n,2,3], filled by forward_step


@dataclass
class CudaState:
    """BulletState (PhyBullet.jl:54-65): elements, latents, kinematics - one rigid body per batch entry."""
    kinematics: RigidBodyState
    latents: dict = field(default_factory=lambda: {"restitution": 0.9, "mass": 1.0, "lateralFriction": 0.5})
    elements: tuple = ("target",)


BulletState = CudaState


def _aabb(shape, pos, quat):
    h = 0.05
    if shape == api.SPHERE:
        ext = np.full_like(pos, h)
    else:
        x, y, z, w = quat.T
        R = np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                      np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                      np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1)
        ext = h * np.abs(R).sum(-1)
    return np.stack([pos - ext, pos + ext], 1)


def sync(sim, st):
    """PhySMC.sync! (PhyBullet.jl:68-76).  Stateless here: the state travels with every forward_step call."""
    return None


def forward_step(sim, st):
    """PhySMC.forward_step (PhyBullet.jl:79-103): S substeps of 1/240 s on the whole batch, new state returned."""
    k = st.kinematics
    n = np.atleast_2d(k.position).shape[0]
    s13 = np.concatenate([np.atleast_2d(k.position), np.atleast_2d(k.orientation), np.atleast_2d(k.linear_vel),
                          np.atleast_2d(k.angular_vel)], axis=1).astype(F32)
    e = np.broadcast_to(np.asarray(st.latents["restitution"], F32), (n,))
    out = api.sim_step(sim.shape, s13, e, S=sim.substeps, ceiling_half=sim.ceiling_half, device=sim.device)
    kin = RigidBodyState(out[:, 0:3], out[:, 3:7], out[:, 7:10], out[:, 10:13], _aabb(sim.shape, out[:, 0:3], out[:, 3:7]))
    return CudaState(kinematics=kin, latents=dict(st.latents), elements=st.elements)


def step(sim, st):
    """PhySMC.step = sync! ; forward_step (PhySMC.jl:32-35)."""
    sync(sim, st)
    return forward_step(sim, st)


def init_target_state(sim, shape, init_position, init_velocity):
    """Model/bouncing_object.jl:63-86: identity orientation (:75), zero spin, restitution placeholder 0.9 (:80)."""
    sim.shape = api.SPHERE if str(shape).lower().startswith("sphere") else api.CUBE
    kin = RigidBodyState(np.asarray(init_position, float)[None], np.array([[0.0, 0.0, 0.0, 1.0]]),
                         np.asarray(init_velocity, float)[None], np.zeros((1, 3)))
    return CudaState(kinematics=kin)


# ------------------------------------------------------------------ Utilities/fileio.jl
DATA_ROOT = ""   # prefix of the reference's relative "Data/..." paths (tests point it at fixtures)


def filter_unwanted_filenames(fnames):
    """fileio.jl:12-17."""
    for bad in ("predicted", "observed", "batch", "Store"):
        fnames = [f for f in fnames if bad not in f]
    return fnames


def trial_order(x):
    """Sort key equivalent to the comparator of fileio.jl:74-99: by elasticity then variation number."""
    t = os.path.basename(x).split("_")
    ela = int("".join(c for c in t[1] if c.isdigit()))
    var = int("".join(c for c in t[2].split(".")[0] if c.isdigit()))
    return (ela, var)


def _frac_id(v):
    s = ("%s" % np.float32(v)).split(".")[1]
    return s + "0" if len(s) < 2 else s


def generate_noise_id(args):
    """fileio.jl:175-193, e.g. Init00_Obs05_Tra075 (digits after the point of the Float32 printout)."""
    return "Init%s_Obs%s_Tra%s" % (_frac_id(args.init_vel_noise), _frac_id(args.observation_noise),
                                   _frac_id(args.transition_noise))


def generate_inference_param_id(args):
    """fileio.jl:196-202."""
    return "Par%d_Rej%d" % (args.num_particles, args.rejuvenation_moves)


def make_directories(output_path, base="Analysis"):
    """fileio.jl:205-240: Analysis/<expt>/<model>/<target>/<noise>/<alg>/<params>/Data/ (created as needed)."""
    tokens = output_path.split("/")
    d = os.path.join(DATA_ROOT, base, *tokens[:6], "Data") + "/"
    os.makedirs(d, exist_ok=True)
    return d


def read_obs_file(fname, source="RealFlow", alg="SMC"):
    """fileio.jl:20-66.  Row 1 of the init CSV: position columns 5:7, orientation 8:10, velocity 11:13 (1-based);
    the trajectory comes from <name>_observed.csv (x,y,z per frame).  The reference reads an undefined `obs_fname`
    at :44; the evident intent (and Tests/generate_trajectories.jl:208) is the _observed file for both sources."""
    with open(fname, newline="") as f:
        rows = list(csv.reader(f))
    r1 = rows[1]
    init_position = [float(v) for v in r1[4:7]]
    init_orientation = [float(v) for v in r1[7:10]]
    init_velocity = [float(v) for v in r1[10:13]]
    head, tail = fname.rsplit(".", 1)
    obs_fname = head + "_observed." + tail
    with open(obs_fname, newline="") as f:
        rows = list(csv.reader(f))[1:]
    obs = np.array([[float(r[0]), float(r[1]), float(r[2])] for r in rows if r], dtype=np.float64).reshape(-1, 3)
    return init_position, init_orientation, init_velocity, obs, obs.shape[0]


def _trunc5(v):
    return math.trunc(v * 1e5) / 1e5


def _jl_float(v):
    """Julia's shortest float printout (CSV.write): 0.5, 1.0, -12.34567, 1.0e-5."""
    s = repr(float(v))
    if "e" in s:
        m, e = s.split("e")
        if "." not in m:
            m += ".0"
        s = "%se%d" % (m, int(e))
    return s


class Particles:
    """What run_smc returns in place of Gen traces: per particle the elasticity
    (trace[:latents=>1=>:restitution]), score (get_score) and the simulator outputs per frame."""

    def __init__(self, e, score, traj, log_w=None, pf=None):
        self.e, self.score, self.traj, self.log_w, self.pf = e, score, traj, log_w, pf

    def __len__(self):
        return len(self.e)


def write_to_csv(particles, fname=os.path.join(os.getcwd(), "test.csv")):
    """fileio.jl:255-278: columns particle,elasticity,weight,frame,x,y,z; floats truncated to 5 digits."""
    T = particles.traj.shape[0]
    with open(fname, "w", newline="") as f:
        f.write("particle,elasticity,weight,frame,x,y,z\n")
        for p in range(len(particles)):
            ela, w = _jl_float(_trunc5(particles.e[p])), _jl_float(_trunc5(particles.score[p]))
            for fr in range(T):
                x, y, z = (_jl_float(_trunc5(float(c))) for c in particles.traj[fr, :, p])
                f.write("%d,%s,%s,%d,%s,%s,%s\n" % (p + 1, ela, w, fr + 1, x, y, z))


# ------------------------------------------------------------------ Inference/particle_filter.jl
def generate_trajectory(*a, **k):
    """Placeholder for the Gen model handle (Model/bouncing_object.jl:148-162) that run_smc's signature carries."""
    raise RuntimeError("the generative model runs inside libsmcphys; pass this object only as run_smc's `model`")


def run_smc(fname, model, gm_args, obs, rejuvenation_moves=1, num_particles=20, save_particles=False,
            output_path=None, w2=None, args=None):
    """particle_filter.jl:37-81.  gm_args = (sim, init_state, T); obs = [T,3] observed positions.
    Returns (particles, log_weights) like (state.traces, get_log_weights(state))."""
    sim, init_state, T = gm_args
    args = args or Args()
    tokens = os.path.basename(fname).split("_")
    csvp = tokens[2].split(".")
    obs = np.asarray(obs, dtype=np.float64).reshape(-1, 3)[:T]
    pf = api.ParticleFilter(num_particles=num_particles, shape=sim.shape, substeps=sim.substeps, max_timesteps=max(T, 1),
                            rejuvenation_moves=rejuvenation_moves, observation_noise=Args().observation_noise,
                            transition_noise=Args().transition_noise, ceiling_half=sim.ceiling_half, device=sim.device,
                            seed=args.seed)
    # the model reads its noise from a default-constructed Args (bouncing_object.jl:21), not from the caller's
    k = init_state.kinematics
    pf.set_trial(np.asarray(k.position, float).reshape(3), np.asarray(k.linear_vel, float).reshape(3), obs)
    if not save_particles:
        pf.run_smc(rejuvenation_moves, want_stats=False)
    else:
        pf.pf_init()
        for t in range(T):
            pf.pf_extend(t)
            pf.pf_resample_local(t)
            pf.pf_rejuvenate(t, rejuvenation_moves)
            out = pf.get_particles(want_traj=True, steps=t + 1)
            name = "%s_%s_%d.%s" % (tokens[1], csvp[0], t + 1, csvp[1])
            write_to_csv(Particles(out["e"], out["score"], out["traj"]), os.path.join(output_path, name))
    out = pf.get_particles(want_traj=True, steps=T)
    return Particles(out["e"], out["score"], out["traj"], out["log_w"], pf), out["log_w"]


def predict(particles, T):
    """particle_filter.jl:84-102: every particle rolled T steps further with sampled transition and observation
    noise; the returned particles carry the extended trajectory (sampled positions of the new frames)."""
    pos, _ = particles.pf.predict(T, want_obs=True)
    return Particles(particles.e, particles.score, np.concatenate([particles.traj, pos], axis=0), particles.log_w,
                     particles.pf)


# ------------------------------------------------------------------ Tests/generate_trajectories.jl
def _quat_from_euler(roll, pitch, yaw):
    """PyBullet getQuaternionFromEuler (xyzw; rotation = Rz(yaw) Ry(pitch) Rx(roll))."""
    cr, sr, cp, sp_, cy, sy = (math.cos(roll / 2), math.sin(roll / 2), math.cos(pitch / 2), math.sin(pitch / 2),
                               math.cos(yaw / 2), math.sin(yaw / 2))
    return [sr * cp * cy - cr * sp_ * sy, cr * sp_ * cy + sr * cp * sy, cr * cp * sy - sr * sp_ * cy,
            cr * cp * cy + sr * sp_ * sy]


def generate_trajectories(obj_type="Sphere", seed=20240830, device=0):
    """Tests/generate_trajectories.jl:107-212 on the CUDA simulator: the 27 stimuli (elasticity 0.1 ... 0.9, three
    variations each) of the BulletTest_<obj_type> experiments that main.jl:26 expects under
    Data/BulletStimulus/New/<obj_type>/ and the reference repository does not ship.  Initial position U[-0.4,0.4]^2 x
    [0.1,0.9], Euler orientation U[0,2pi]^3, velocity U[-5,5]^3 (:73-89); one PhySMC.step (1/60 s) per frame until
    120 frames or every velocity component is within 0.1 (:184); positions truncated to 5 digits (:175,:208); the
    ceiling slab is 0.02 thick here (:52).  The reference draws from an unseeded RNG; `seed` makes the set
    reproducible.  Writes <obj>_Ela<e>_Var<v>.csv (initial state, the column layout read_obs_file expects) next to
    <obj>_Ela<e>_Var<v>_observed.csv, and the initial-state file also where the reference puts it (one level up)."""
    rng = np.random.default_rng(seed)
    sim = CudaSim(step_dur=1 / 60, device=device, ceiling_half=0.02)
    sim.shape = api.SPHERE if obj_type.lower().startswith("sphere") else api.CUBE
    d_init = os.path.join(DATA_ROOT, "Data/BulletStimulus", obj_type)
    d_new = os.path.join(DATA_ROOT, "Data/BulletStimulus/New", obj_type)
    os.makedirs(d_init, exist_ok=True)
    os.makedirs(d_new, exist_ok=True)
    names, pos, quat, vel, ela = [], [], [], [], []
    for i in range(3, 30):
        e_idx, var = i // 3, i % 3 + 1
        p = [rng.uniform(-0.4, 0.4), rng.uniform(-0.4, 0.4), rng.uniform(0.1, 0.9)]
        eul = [rng.uniform(0, 2 * math.pi) for _ in range(3)]
        v = [rng.uniform(-5, 5) for _ in range(3)]
        name = "%s_Ela%d_Var%d" % (obj_type, e_idx, var)
        header = "Date,SceneID,Elasticity,Variation,PositionX,PositionY,PositionZ,RotationX,RotationY,RotationZ," \
                 "VelocityX,VelocityY,VelocityZ\n"
        row = "2024-08-30,Cube_Ela%dVar_%d,%s,%d,%s\n" % (e_idx, var, _jl_float(0.1 * e_idx), var,
                                                         ",".join(_jl_float(x) for x in p + eul + v))
        for d in (d_init, d_new):
            with open(os.path.join(d, name + ".csv"), "w") as f:
                f.write(header + row)
        names.append(name); pos.append(p); quat.append(_quat_from_euler(*eul)); vel.append(v); ela.append(0.1 * e_idx)
    n = len(names)
    kin = RigidBodyState(np.array(pos), np.array(quat), np.array(vel), np.zeros((n, 3)))
    st = CudaState(kinematics=kin, latents={"restitution": np.array(ela, F32), "mass": 1.0, "lateralFriction": 0.5})
    alive = np.ones(n, bool)
    traj = [[] for _ in range(n)]
    v_now = np.array(vel)
    for t in range(120):
        alive &= (np.abs(v_now) > 0.1).any(axis=1)
        if not alive.any():
            break
        st = step(sim, st)                  # all 27 bodies in one batch; a stopped body's further steps are ignored
        v_now = st.kinematics.linear_vel
        for b in np.where(alive)[0]:
            traj[b].append(st.kinematics.position[b].copy())
    for name, rows in zip(names, traj):
        with open(os.path.join(d_new, name + "_observed.csv"), "w") as f:
            f.write("x,y,z\n")
            for r in rows:
                f.write(",".join(_jl_float(_trunc5(float(c))) for c in r) + "\n")
    return [nm + ".csv" for nm in names]


# ------------------------------------------------------------------ Inference/mcmc.jl
def gaussian_drift_mcmc(model, gm_args, obs, runs=1000, keep=100, chains=1, args=None):
    """mcmc.jl:53-93: one trace generated from the prior and conditioned on the observations (`generate`, :56), then
    `runs` Metropolis-Hastings moves on the elasticity with the particle filter's Gaussian-drift proposal
    (`mh(trace, proposal, ())`, :68), each a full re-simulation of the T steps; returns (mean elasticity of the last
    `keep` moves, the last `keep` states as Particles: one row per recorded move).  `chains` > 1 runs that many
    independent chains side by side (the GPU's way of using the hardware; the reference runs one) and averages them.
    On the device this is the rejuvenation kernel at the fixed step T: the first runs - keep moves in one launch, then
    one launch per recorded move (smcphys_pf_rejuvenate_from keeps the Philox move numbering)."""
    sim, init_state, T = gm_args
    args = args or Args()
    obs = np.asarray(obs, dtype=np.float64).reshape(-1, 3)[:T]
    pf = api.ParticleFilter(num_particles=chains, shape=sim.shape, substeps=sim.substeps, max_timesteps=max(T, 1),
                            rejuvenation_moves=1, observation_noise=Args().observation_noise,
                            transition_noise=Args().transition_noise, ceiling_half=sim.ceiling_half, device=sim.device,
                            seed=args.seed)
    k = init_state.kinematics
    pf.set_trial(np.asarray(k.position, float).reshape(3), np.asarray(k.linear_vel, float).reshape(3), obs)
    pf.pf_init()
    for t in range(T):                      # generate(): every step extended from the prior, no resampling
        pf.pf_extend(t)
    burn = max(runs - keep, 0)
    if burn:
        pf.pf_rejuvenate(T - 1, burn, move0=0)
    es, scores, trajs = [], [], []
    for i in range(runs - burn):
        pf.pf_rejuvenate(T - 1, 1, move0=burn + i)
        out = pf.get_particles(want_traj=True, steps=T)
        es.append(out["e"].copy()); scores.append(out["score"].copy()); trajs.append(out["traj"].copy())
    pf.close()
    e = np.stack(es)                        # [keep, chains]
    avg_last = float(e.mean())
    results = Particles(e.reshape(-1), np.stack(scores).reshape(-1), np.concatenate(trajs, axis=2))
    return avg_last, results


# ------------------------------------------------------------------ Inference/run_inference.jl
def run_inference(fname, args, output_path, w1=None, w2=None):
    """run_inference.jl:19-124: one trial -> <output_path>inferences/<Ela>_<Var>.csv (+ intermediate/ dumps)."""
    if "Bullet" in args.expt_id:
        source, sim = "Bullet", CudaSim(step_dur=1 / 60, device=args.device)
        bullet_shape = args.expt_id.split("_")[1]
        fname = os.path.join(DATA_ROOT, "Data/BulletStimulus/New/", bullet_shape, fname)
    else:
        source, sim = "RealFlow", CudaSim(step_dur=1 / 30, device=args.device)
        fname = os.path.join(DATA_ROOT, "Data/RealFlowData/", args.expt_id, fname)
    init_position, _, init_velocity, obs, t_s = read_obs_file(fname, source, args.algorithm)
    init_state = init_target_state(sim, args.target_id, init_position, init_velocity)
    sim_args = (sim, init_state, t_s)
    if args.algorithm == "MCMC":            # run_inference.jl:56-68
        print("Initializing MCMC")
        avg_last_hundred, results = gaussian_drift_mcmc(generate_trajectory, sim_args, obs, args=args)
        print("Elasticity estimate: ", avg_last_hundred)
        tokens = os.path.basename(fname).split("_")
        write_to_csv(results, os.path.join(output_path, "%s_%s" % (tokens[1], tokens[2])))
        return results
    if args.algorithm != "SMC":
        raise NotImplementedError("algorithm=%r: the DEBUG mode opens PyBullet's GUI (unavailable on this box)"
                                  % args.algorithm)
    print("Initializing Particle Filter")
    particle_dir = os.path.join(output_path, "intermediate/")
    os.makedirs(particle_dir, exist_ok=True)
    results, _ = run_smc(fname, generate_trajectory, sim_args, obs, args.rejuvenation_moves, args.num_particles,
                         args.save_particles, particle_dir, w2, args=args)
    result_dir = os.path.join(output_path, "inferences/")
    os.makedirs(result_dir, exist_ok=True)
    tokens = os.path.basename(fname).split("_")
    output_fname = "%s_%s" % (tokens[1], tokens[2])
    print("Writing to: ", result_dir, output_fname)
    write_to_csv(results, os.path.join(result_dir, output_fname))
    if args.predict:
        ppd = predict(results, args.prediction_timesteps)
        pred_dir = os.path.join(output_path, "Predictions/")
        os.makedirs(pred_dir, exist_ok=True)
        write_to_csv(ppd, os.path.join(pred_dir, "predictions_%s_%s" % (tokens[1], tokens[2])))
    results.pf.close()
    results.pf = None
    return results


def run_inference_batch(fnames, args, output_path, max_concurrent=None, memory_budget=96e9):
    """Many trials at once on one GPU - the stand-in for the reference's `pmap` over 15 worker processes
    (main.jl:7,54) and BASELINE.json's "all trials concurrent" configuration.  Every trial gets its own filter
    handle, and a handle owns its own CUDA stream; `run_smc` only enqueues work (the resampling decision is taken
    on the device), so enqueueing all trials of a wave before reading any result lets their kernels overlap and
    fill the GPU even when one trial's particles do not.  Waves are sized by device memory.  Outputs are identical
    to calling run_inference on each trial in turn (same seeds, same files); per-step particle dumps
    (save_particles) need a host round trip per step and fall back to the sequential path."""
    if args.algorithm != "SMC":
        raise NotImplementedError("algorithm=%r: only the SMC path is on the B200 hot path" % args.algorithm)
    if args.save_particles:
        return [run_inference(f, args, output_path) for f in fnames]
    bullet = "Bullet" in args.expt_id
    trials = []
    for f in fnames:
        if bullet:
            sim = CudaSim(step_dur=1 / 60, device=args.device)
            path = os.path.join(DATA_ROOT, "Data/BulletStimulus/New/", args.expt_id.split("_")[1], f)
        else:
            sim = CudaSim(step_dur=1 / 30, device=args.device)
            path = os.path.join(DATA_ROOT, "Data/RealFlowData/", args.expt_id, f)
        init_position, _, init_velocity, obs, t_s = read_obs_file(path, "Bullet" if bullet else "RealFlow", args.algorithm)
        init_target_state(sim, args.target_id, init_position, init_velocity)
        trials.append((path, sim, init_position, init_velocity, obs, t_s))
    per_trial = lambda T: args.num_particles * (24.0 * max(T, 1) + 260.0)   # two history buffers + state, bytes
    result_dir = os.path.join(output_path, "inferences/")
    os.makedirs(os.path.join(output_path, "intermediate/"), exist_ok=True)
    os.makedirs(result_dir, exist_ok=True)
    results, i = [], 0
    while i < len(trials):
        wave, used = [], 0.0
        while i < len(trials) and (not wave or (used + per_trial(trials[i][5]) <= memory_budget and
                                               (max_concurrent is None or len(wave) < max_concurrent))):
            used += per_trial(trials[i][5])
            wave.append(trials[i])
            i += 1
        pfs = []
        for path, sim, ip, iv, obs, t_s in wave:       # enqueue everything ...
            pf = api.ParticleFilter(num_particles=args.num_particles, shape=sim.shape, substeps=sim.substeps,
                                    max_timesteps=max(t_s, 1), rejuvenation_moves=args.rejuvenation_moves,
                                    observation_noise=Args().observation_noise, transition_noise=Args().transition_noise,
                                    ceiling_half=sim.ceiling_half, device=sim.device, seed=args.seed)
            pf.set_trial(np.asarray(ip, float), np.asarray(iv, float), np.asarray(obs, float).reshape(-1, 3)[:t_s])
            pf.run_smc(args.rejuvenation_moves, want_stats=False)
            pfs.append(pf)
        for (path, sim, ip, iv, obs, t_s), pf in zip(wave, pfs):   # ... then collect
            out = pf.get_particles(want_traj=True, steps=t_s)
            res = Particles(out["e"], out["score"], out["traj"], out["log_w"], pf)
            tokens = os.path.basename(path).split("_")
            write_to_csv(res, os.path.join(result_dir, "%s_%s" % (tokens[1], tokens[2])))
            if args.predict:
                ppd = predict(res, args.prediction_timesteps)
                os.makedirs(os.path.join(output_path, "Predictions/"), exist_ok=True)
                write_to_csv(ppd, os.path.join(output_path, "Predictions/", "predictions_%s_%s" % (tokens[1], tokens[2])))
            pf.close()
            res.pf = None
            results.append(res)
    return results


def main(args=None, fnames=None, concurrent=False):
    """main.jl:18-85 on one process: list + sort the trials, build the output directory, run every trial
    (concurrent=True: all trials of a wave in flight at once, run_inference_batch)."""
    args = args or Args()
    if "Bullet" in args.expt_id:
        d = os.path.join(DATA_ROOT, "Data/BulletStimulus/New/", args.expt_id.split("_")[1])
        names = [f for f in os.listdir(d) if "observed" not in f]
    else:
        d = os.path.join(DATA_ROOT, "Data/RealFlowData/", args.expt_id)
        names = filter_unwanted_filenames(os.listdir(d))
    names = sorted(fnames if fnames is not None else names, key=trial_order)
    out = make_directories("/".join([args.expt_id, args.model_id, args.target_id, generate_noise_id(args),
                                     args.algorithm, generate_inference_param_id(args), "Data"]))
    print("Output Filepath... ", out)
    if concurrent:
        return run_inference_batch(names, args, out)
    return [run_inference(f, args, out) for f in names]
