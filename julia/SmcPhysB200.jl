# SmcPhysB200.jl — Julia host layer over libsmcphys.so (include/smcphys.h).
#
# NOT EXECUTED IN THIS REPO'S CI: neither build machine has Julia.  The file is kept mechanical on purpose — every
# function is one or two `ccall`s; all logic lives behind the C ABI and is tested through the identical Python
# ctypes binding (smc_physics_b200/_lib.py, api.py, host.py).  See INTEGRATION.md for where each piece plugs into
# the reference (paths below are in jackhdewey/smc-physics).
#
#   CudaSim <: PhySim, CudaState <: PhyState{CudaSim}     PhySMC/PhySMC.jl:13-16, PhyBullet.jl:23-30,54-65
#   PhySMC.sync!, PhySMC.forward_step                      PhySMC/PhySMC.jl:32-49, PhyBullet.jl:68-103
#   run_smc(fname, model, gm_args, obs, R, N, save, out, w2)   Inference/particle_filter.jl:37-81
#   predict(particles, T)                                  Inference/particle_filter.jl:84-102
#   write_to_csv(particles, fname)                         Utilities/fileio.jl:255-278
module SmcPhysB200

using Libdl

export CudaSim, CudaState, CudaRigidBodyState, Particles, run_smc, predict, write_to_csv, resample_systematic

const LIB = Ref{String}(get(ENV, "SMCPHYS_LIB", joinpath(@__DIR__, "..", "smc_physics_b200", "libsmcphys.so")))

const SPHERE = Cint(0)
const CUBE = Cint(1)

# ---- include/smcphys.h: smcphys_config (field order and types must match the C struct exactly)
mutable struct Config
    struct_size::Int32
    device::Int32
    shape::Int32
    substeps::Int32
    num_particles::Int64
    global_particles::Int64
    particle_offset::Int64
    max_timesteps::Int32
    rejuvenation_moves::Int32
    observation_noise::Float64
    transition_noise::Float64
    proposal_sigma::Float64
    ceiling_half::Float64
    ess_num::Int64
    ess_den::Int64
    seed::UInt64
    resampler::Int32          # SYSTEMATIC | MULTINOMIAL (what Gen.maybe_resample! draws, particle_filter.jl:55)
    pad_::Int32
    Config() = new()
end

const SYSTEMATIC = Int32(0)
const MULTINOMIAL = Int32(1)

struct StepStats
    ess::Float64
    lse::Float64
    accept_rate::Float64
    resampled::Int32
    pad_::Int32
end

last_error() = unsafe_string(ccall((:smcphys_last_error, LIB[]), Cstring, ()))

function check(rc::Cint)
    rc == 0 || error("libsmcphys error $(rc): $(last_error())")   # Julia exception, like the reference's own failures
    nothing
end

function default_config()
    c = Config()
    check(ccall((:smcphys_config_default, LIB[]), Cint, (Ref{Config},), c))
    c
end

# ---------------------------------------------------------------- PhySMC simulator interface
# The module is loaded next to the reference's PhySMC (Main.PhySMC after `include("PhySMC/PhySMC.jl")`).
import Main.PhySMC
import Main.PhySMC: PhySim, PhyState

"Stands in for BulletSim (PhyBullet.jl:23-30): same field vocabulary, a CUDA device instead of a PyBullet client."
Base.@kwdef struct CudaSim <: PhySim
    device::Int32 = 0
    pb_timestep::Float64 = 1 / 240
    step_dur::Float64 = 1 / 60
    shape::Int32 = SPHERE
    ceiling_half::Float64 = 0.01
end

"`while dt <= step_dur` with floating accumulation (PhyBullet.jl:82-88): 5 at 1/60, 9 at 1/30."
function substeps(sim::CudaSim)
    dt = 0.0; s = 0
    while dt <= sim.step_dur
        s += 1; dt += sim.pb_timestep
    end
    Int32(s)
end

"rigid_body.jl:16-27 for a batch of n bodies: row i of `state13` = position(3) orientation xyzw(4) linear(3) angular(3)."
struct CudaRigidBodyState
    state13::Matrix{Float32}          # 13 × n (column = body ⇒ C row-major [n][13])
end

struct CudaState <: PhyState{CudaSim}
    kinematics::CudaRigidBodyState
    restitution::Vector{Float32}      # latents (rigid_body.jl:116-126): one elasticity per body
end

PhySMC.sync!(sim::CudaSim, st::CudaState) = nothing   # stateless: the state travels with forward_step

function PhySMC.forward_step(sim::CudaSim, st::CudaState)
    s = copy(st.kinematics.state13)
    n = size(s, 2)
    check(ccall((:smcphys_sim_step, LIB[]), Cint,
                (Int32, Int32, Int32, Float64, Int64, Ptr{Float32}, Ptr{Float32}),
                sim.device, sim.shape, substeps(sim), sim.ceiling_half, n, s, st.restitution))
    CudaState(CudaRigidBodyState(s), st.restitution)
end

# ---------------------------------------------------------------- run_smc
"What run_smc returns in place of Gen traces."
struct Particles
    elasticity::Vector{Float64}       # trace[:latents => 1 => :restitution]
    score::Vector{Float64}            # get_score(trace)
    traj::Array{Float32,3}            # n × 3 × T (C layout [T][3][n]): simulator output after each step
    handle::Ptr{Cvoid}
    T::Int
end
Base.length(p::Particles) = length(p.elasticity)

"""
    run_smc(fname, model, gm_args, obs, rejuvenation_moves=1, num_particles=20, save_particles=false,
            output_path=nothing, w2=nothing; seed=1234) -> (particles, log_weights)

Same positional signature as Inference/particle_filter.jl:37.  `gm_args = (sim::CudaSim, init_state, T)` where
`init_state` has `position` and `linear_vel` (3-vectors); `obs` is a T×3 matrix or a vector of 3-vectors.
`model` is accepted and ignored (the generative model is fixed inside the library).
"""
function run_smc(fname::String, model, gm_args::Tuple, obs, rejuvenation_moves=1, num_particles::Int=20,
                 save_particles=false, output_path=nothing, w2=nothing; seed::Integer=1234,
                 observation_noise=Float64(0.05f0), transition_noise=Float64(0.075f0),
                 resampler::Integer=SYSTEMATIC, devices::Union{Nothing,Vector{<:Integer}}=nothing)
    sim, init_state, T = gm_args
    tokens = split(basename(fname), "_")
    csv = split(tokens[3], ".")
    c = default_config()
    c.device = sim.device; c.shape = sim.shape; c.substeps = substeps(sim)
    c.num_particles = num_particles; c.global_particles = num_particles; c.particle_offset = 0
    c.max_timesteps = max(T, 1); c.rejuvenation_moves = rejuvenation_moves
    c.observation_noise = observation_noise; c.transition_noise = transition_noise
    c.ceiling_half = sim.ceiling_half; c.seed = seed; c.resampler = resampler
    h = Ref{Ptr{Cvoid}}(C_NULL)
    if devices === nothing
        check(ccall((:smcphys_create, LIB[]), Cint, (Ref{Config}, Ref{Ptr{Cvoid}}), c, h))
    else
        # ONE filter over several GPUs of this process (smcphys_create_multi): num_particles is split evenly over the
        # listed devices; everything below (set_trial, run_smc, get_particles) is unchanged - still one ccall each.
        # save_particles needs the phase entries and stays single-device.
        num_particles % length(devices) == 0 || error("num_particles must be a multiple of the device count")
        save_particles && error("save_particles: per-step dumps are single-device")
        c.num_particles = num_particles ÷ length(devices)
        check(ccall((:smcphys_create_multi, LIB[]), Cint, (Ref{Config}, Int32, Ptr{Int32}, Ref{Ptr{Cvoid}}),
                    c, length(devices), Int32.(devices), h))
    end
    obsm = obs isa AbstractMatrix ? Matrix{Float64}(obs') : reduce(hcat, [Float64.(o) for o in obs])  # 3 × T
    check(ccall((:smcphys_set_trial, LIB[]), Cint, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32),
                h[], Float64.(collect(init_state.position)), Float64.(collect(init_state.linear_vel)), obsm, T))
    if !save_particles
        check(ccall((:smcphys_run_smc, LIB[]), Cint, (Ptr{Cvoid}, Int32, Ptr{StepStats}), h[], rejuvenation_moves, C_NULL))
    else
        check(ccall((:smcphys_pf_init, LIB[]), Cint, (Ptr{Cvoid},), h[]))
        for t in 0:T-1
            check(ccall((:smcphys_pf_extend, LIB[]), Cint, (Ptr{Cvoid}, Int32, Ptr{Float64}), h[], t, C_NULL))
            check(ccall((:smcphys_pf_resample_local, LIB[]), Cint, (Ptr{Cvoid}, Int32), h[], t))
            check(ccall((:smcphys_pf_rejuvenate, LIB[]), Cint, (Ptr{Cvoid}, Int32, Int32, Ptr{UInt64}), h[], t,
                        rejuvenation_moves, C_NULL))
            p, _ = get_particles(h[], num_particles, T)   # frames beyond t+1 are not yet simulated
            write_to_csv(p, string(output_path, tokens[2], "_", csv[1], "_", t + 1, ".", csv[2]); frames=t + 1)
        end
    end
    get_particles(h[], num_particles, T)
end

function get_particles(h::Ptr{Cvoid}, n::Int, T::Int)
    e = Vector{Float64}(undef, n); lw = similar(e); sc = similar(e)
    traj = Array{Float32,3}(undef, n, 3, T)
    check(ccall((:smcphys_get_particles, LIB[]), Cint,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float32}), h, e, lw, sc, traj))
    Particles(e, sc, traj, h, T), lw
end

close(p::Particles) = check(ccall((:smcphys_destroy, LIB[]), Cint, (Ptr{Cvoid},), p.handle))

"particle_filter.jl:84-102: every particle rolled T steps further with sampled transition + observation noise."
function predict(p::Particles, T::Int)
    n = length(p)
    pos = Array{Float32,3}(undef, n, 3, T)
    check(ccall((:smcphys_predict, LIB[]), Cint, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}), p.handle, T, pos, C_NULL))
    Particles(p.elasticity, p.score, cat(p.traj, pos; dims=3), p.handle, p.T + T)
end

trunc5(v) = trunc(Float64(v); digits=5)

"fileio.jl:255-278: header particle,elasticity,weight,frame,x,y,z; floats truncated to 5 digits."
function write_to_csv(p::Particles, fname; frames::Int=p.T)
    open(fname, "w") do io
        println(io, "particle,elasticity,weight,frame,x,y,z")
        for i in 1:length(p), t in 1:frames
            println(io, i, ",", trunc5(p.elasticity[i]), ",", trunc5(p.score[i]), ",", t, ",",
                    trunc5(p.traj[i, 1, t]), ",", trunc5(p.traj[i, 2, t]), ",", trunc5(p.traj[i, 3, t]))
        end
    end
end

# ---------------------------------------------------------------- one filter over the GPUs of several processes
# (Julia Distributed workers, one per GPU - the reference's addprocs layout, main.jl:5-9): every worker creates its
# shard (Config with num_particles = N / nranks, global_particles = N, particle_offset = rank * num_particles), rank 0
# makes the id and ships it (e.g. remotecall_fetch), every worker calls comm_init_rank; smcphys_run_smc is collective.
function comm_unique_id()
    id = Vector{UInt8}(undef, 128)
    check(ccall((:smcphys_comm_unique_id, LIB[]), Cint, (Ptr{UInt8},), id))
    id
end
comm_init_rank(h::Ptr{Cvoid}, id::Vector{UInt8}, nranks::Integer, rank::Integer) =
    check(ccall((:smcphys_comm_init_rank, LIB[]), Cint, (Ptr{Cvoid}, Ptr{UInt8}, Int32, Int32), h, id, nranks, rank))

"Parity entry: systematic resampling indices (1-based) for log weights `lw` and uniform offset `u0`."
function resample_systematic(lw::Vector{Float64}, u0::Float64; device::Integer=0)
    idx = Vector{Int64}(undef, length(lw))
    check(ccall((:smcphys_resample_systematic, LIB[]), Cint, (Int32, Ptr{Float64}, Int64, Float64, Ptr{Int64}),
                device, lw, length(lw), u0, idx))
    idx .+ 1
end

end # module
