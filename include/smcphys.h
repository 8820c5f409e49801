/* libsmcphys — C ABI of the B200 (sm_100a) particle-filter hot path of smc-physics.
 *
 * The reference (jackhdewey/smc-physics) has no FFI of its own: its seams are Julia-level.  Each entry point
 * below names the reference interface it stands in for, so a maintainer can bind it with `ccall` (see
 * INTEGRATION.md).  Conventions: every function returns 0 on success or a negative SMCPHYS_E* code and never
 * throws or aborts across the boundary; smcphys_last_error() gives the message (thread local).  The caller owns
 * all host buffers; a handle owns its device memory and stream.  A handle is not thread safe; distinct handles
 * are independent.  There is no CPU fallback: without a CUDA device every compute entry returns SMCPHYS_ECUDA.
 */
#ifndef SMCPHYS_H
#define SMCPHYS_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMCPHYS_OK 0
#define SMCPHYS_EINVAL (-1) /* bad argument */
#define SMCPHYS_ECUDA (-2)  /* CUDA runtime error / no device */
#define SMCPHYS_ESTATE (-3) /* call out of order (e.g. run before set_trial) */
#define SMCPHYS_ENOMEM (-4)

#define SMCPHYS_RESAMPLE_SYSTEMATIC 0
#define SMCPHYS_RESAMPLE_MULTINOMIAL 1
#define SMCPHYS_MAX_RANKS 64 /* shards of one filter */
#define SMCPHYS_UNIQUE_ID_BYTES 128

#define SMCPHYS_SPHERE 0 /* Model/bouncing_object.jl:71  sphere r = 0.05 */
#define SMCPHYS_CUBE 1   /* Model/bouncing_object.jl:68  box half extent 0.05 */

typedef struct smcphys_handle smcphys_handle;

/* Mirrors args.jl:5-27 (num_particles :19, rejuvenation_moves :20, observation_noise :14, transition_noise :15)
 * plus what the reference hard-codes: substeps per PhySMC.step (PhyBullet.jl:82-88: 5 at step_dur 1/60, 9 at
 * 1/30), the ESS threshold N/2 (Inference/particle_filter.jl:55), the proposal width 0.1 (:23), the ceiling
 * half height 0.01 (Model/bouncing_object.jl:53). */
typedef struct smcphys_config {
  int32_t struct_size;        /* sizeof(smcphys_config), checked */
  int32_t device;             /* CUDA device ordinal */
  int32_t shape;              /* SMCPHYS_SPHERE | SMCPHYS_CUBE */
  int32_t substeps;           /* S */
  int64_t num_particles;      /* particles held by THIS handle (the local shard) */
  int64_t global_particles;   /* N over all shards; = num_particles on one GPU */
  int64_t particle_offset;    /* global index of local particle 0 */
  int32_t max_timesteps;      /* capacity of the retained-position history */
  int32_t rejuvenation_moves; /* R */
  double observation_noise;   /* sigma_obs, Float32(0.05) promoted */
  double transition_noise;    /* sigma_tr, Float32(0.075) promoted */
  double proposal_sigma;      /* 0.1 */
  double ceiling_half;        /* 0.01 */
  int64_t ess_num, ess_den;   /* resample when ESS < N * ess_num / ess_den (1/2) */
  uint64_t seed;
  int32_t resampler;          /* SMCPHYS_RESAMPLE_SYSTEMATIC (north_star) | SMCPHYS_RESAMPLE_MULTINOMIAL (what
                                 Gen.maybe_resample! draws in Gen 0.4.7, Inference/particle_filter.jl:55) */
  int32_t pad_;
} smcphys_config;

/* per PF step diagnostics (the reference discards its own @elapsed, particle_filter.jl:49,75) */
typedef struct smcphys_step_stats {
  double ess;         /* effective sample size before the resampling decision */
  double lse;         /* log-sum-exp of the log weights */
  double accept_rate; /* accepted / (N*R) rejuvenation moves */
  int32_t resampled;
  int32_t pad_;
} smcphys_step_stats;

int smcphys_config_default(smcphys_config* cfg);
const char* smcphys_last_error(void);
int smcphys_device_count(void);
/* number of kernel launches this library issued in this process (bench.py's gpu_launches) */
int64_t smcphys_launch_count(void);

/* ---- per-trial filter state: what run_inference.jl:33-48 sets up (sim, init_state, observations) */
int smcphys_create(const smcphys_config* cfg, smcphys_handle** out);
int smcphys_destroy(smcphys_handle* h);
/* run all work of this handle on an existing CUDA stream (cudaStream_t passed as void*); 0 = own stream */
int smcphys_set_stream(smcphys_handle* h, void* cuda_stream);
/* init position / velocity (Utilities/fileio.jl:27,35; orientation is identity, bouncing_object.jl:75) and the
 * T observed positions, row-major T x 3 (fileio.jl:46-60). */
int smcphys_set_trial(smcphys_handle* h, const double init_pos[3], const double init_vel[3], const double* obs,
                      int32_t T);

/* run_smc (Inference/particle_filter.jl:37-81) on one GPU: prior draw, then for t = 1..T extend + weight,
 * ESS-triggered systematic resampling, N x R Metropolis-Hastings rejuvenation moves.  Asynchronous on the
 * handle's stream until a getter is called.  per_step may be NULL, else T entries. */
int smcphys_run_smc(smcphys_handle* h, int32_t R, smcphys_step_stats* per_step);
/* final particles: elasticity (trace[:latents=>1=>:restitution]), PF log weights (particle_filter.jl:79),
 * score = log joint of the trace (Utilities/fileio.jl:267), traj = simulator output after each step
 * (fileio.jl:264-265), layout [T][3][num_particles] float32.  Any pointer may be NULL. */
int smcphys_get_particles(smcphys_handle* h, double* e, double* log_w, double* score, float* traj);
/* final body state per particle, [num_particles][13] = position(3) orientation xyzw(4) linear(3) angular(3)
 * (PhySMC/PhyBullet/Elements/rigid_body.jl:16-27) */
int smcphys_get_states(smcphys_handle* h, float* state13);
int smcphys_get_log_ml(smcphys_handle* h, double* log_ml);
/* predict (particle_filter.jl:84-102): roll every final particle forward n_steps with sampled transition and
 * observation noise; out layouts [n_steps][3][num_particles] float32 (either may be NULL). */
int smcphys_predict(smcphys_handle* h, int32_t n_steps, float* pos_out, float* obs_out);

/* ---- one filter over several GPUs, inside the library (the multi-GPU drop-in for main.jl:5-9,54: the host makes
 * the same set_trial / run_smc / get_particles calls, the particles of a trial are sharded over the GPUs).
 * Per PF step, all stream ordered with no host synchronisation: local max -> all-reduce(max) -> integer weight sums
 * -> all-gather -> ESS decision, CDF offset and offspring slot range on the device -> ancestors -> the offspring are
 * STORED into the destination rank's idle buffers through peer memory (NVLink) -> barrier -> rejuvenation.  (Multinomial
 * resampler: barrier -> every rank draws the parents of its own slots, searches the owning rank's CDF and PULLS them.)
 *
 * (a) One process per GPU (torchrun, MPI, Julia Distributed workers): every rank creates its handle with
 *     num_particles = N / nranks, global_particles = N, particle_offset = rank * num_particles, rank 0 calls
 *     smcphys_comm_unique_id and ships the 128 bytes to the others (any transport), every rank calls
 *     smcphys_comm_init_rank (NCCL communicator + CUDA IPC mapping of the peers' buffers).  smcphys_run_smc is then
 *     collective: every rank calls it.  NCCL is bound at run time (dlopen libnccl.so.2).
 * (b) One process, several devices (a single ccall-ing Julia process): smcphys_create_multi returns ONE handle that
 *     owns a shard per listed device (cfg->num_particles particles each; a device may be listed more than once);
 *     set_trial / run_smc / get_particles / get_log_ml / synchronize / destroy act on the whole filter. */
int smcphys_comm_unique_id(uint8_t id[SMCPHYS_UNIQUE_ID_BYTES]);
int smcphys_comm_init_rank(smcphys_handle* h, const uint8_t id[SMCPHYS_UNIQUE_ID_BYTES], int32_t nranks, int32_t rank);
int smcphys_create_multi(const smcphys_config* cfg, int32_t n_shards, const int32_t* devices, smcphys_handle** out);
/* particles of all shards of a smcphys_create_multi handle (= num_particles of an ordinary handle) */
int64_t smcphys_total_particles(smcphys_handle* h);
/* simulator outputs of the first `steps` PF steps actually applied (the `traj` rows smcphys_get_particles writes) */
int32_t smcphys_steps_done(smcphys_handle* h);

/* ---- the same filter as phases, for particles sharded over several handles / GPUs (one per process).  The
 * host does the exchanges between phases (NCCL through torch.distributed or ccall'd NCCL).  Pointers named
 * d_* are DEVICE pointers on the handle's device; everything is enqueued on the handle's stream. */
int smcphys_pf_init(smcphys_handle* h);                                   /* prior e ~ U(0,1), log w = 0 */
int smcphys_pf_extend(smcphys_handle* h, int32_t t, double* d_local_max); /* kernel(): noise, weight, step */
int smcphys_pf_weight_sums(smcphys_handle* h, const double* d_global_max, uint64_t* d_sums3 /* W,S1,S2 */);
/* exact integer test  ESS < N*num/den  from the global sums (any host may call it; no device work) */
int smcphys_ess_below(uint64_t S1, uint64_t S2, int64_t N, int64_t num, int64_t den);
double smcphys_ess_value(uint64_t S1, uint64_t S2);
double smcphys_lse_value(double global_max, uint64_t W);
/* uniform offset of step t's systematic resampling (Philox, identical on every rank) */
double smcphys_resample_offset(uint64_t seed, int32_t t);
/* Offspring of this shard's particles fill the global slots [k_lo, k_hi); both are returned.  W_before = sum of
 * the integer weights of all lower ranks, W_total over all ranks. */
int smcphys_pf_offspring_range(uint64_t W_before, uint64_t W_local, uint64_t W_total, int64_t N, double u0,
                               int64_t* k_lo, int64_t* k_hi);
/* words (4-byte) per migrating particle at step t */
int64_t smcphys_pf_payload_words(smcphys_handle* h, int32_t t);
/* pack the offspring [k_lo,k_hi) of this shard into d_send: consecutive chunks, one per destination rank
 * (slots_per_rank = N / G), each chunk SoA [word][count]. */
int smcphys_pf_pack(smcphys_handle* h, int32_t t, const double* d_global_max, uint64_t W_before, uint64_t W_total,
                    double u0, int64_t k_lo, int64_t k_hi, int64_t slots_per_rank, float* d_send);
/* The part of this shard's offspring that stays on this shard - global slots [k_lo, k_hi), a sub-range of both its
 * offspring range and its own slot range - gathered straight into the local slots slot0.. (= k_lo - particle_offset)
 * of the idle buffer, without a trip through the send / receive buffers.  Call before pf_unpack of the same step;
 * pf_pack then only handles the ranges that leave the shard. */
int smcphys_pf_gather_local(smcphys_handle* h, int32_t t, const double* d_global_max, uint64_t W_before,
                            uint64_t W_total, double u0, int64_t k_lo, int64_t k_hi, int64_t slot0);
/* unpack n_chunks received chunks: chunk c holds counts[c] particles for local slots starting at slot0[c],
 * at word offset word_off[c] of d_recv (host arrays).  Together with pf_gather_local the chunks must fill every
 * local slot exactly once.  Resets log weights to 0 and switches buffers. */
int smcphys_pf_unpack(smcphys_handle* h, int32_t t, const float* d_recv, int32_t n_chunks, const int64_t* counts,
                      const int64_t* slot0, const int64_t* word_off);
/* single-shard resampling step with the decision taken on the device (no host synchronisation) */
int smcphys_pf_resample_local(smcphys_handle* h, int32_t t);
int smcphys_pf_rejuvenate(smcphys_handle* h, int32_t t, int32_t R, uint64_t* d_accept_count /* may be NULL */);
/* The same with the moves numbered move0 .. move0+R-1 in the Philox counter, so that a long chain of MH moves at a
 * fixed step can be issued in pieces with results read in between: gaussian_drift_mcmc (Inference/mcmc.jl:53-93,
 * 1000 moves at t = T, the last 100 recorded). */
int smcphys_pf_rejuvenate_from(smcphys_handle* h, int32_t t, int32_t R, int32_t move0, uint64_t* d_accept_count);
int smcphys_synchronize(smcphys_handle* h);

/* ---- stateless entries (host pointers), used for parity checks and the PhySMC simulator interface */
/* PhySMC.forward_step on a batch (PhySMC/PhySMC.jl:32-35, PhyBullet.jl:79-103): n bodies, state13 rows as
 * above, restitution per body, S substeps; in place. */
int smcphys_sim_step(int32_t device, int32_t shape, int32_t S, double ceiling_half, int64_t n, float* state13,
                     const float* restitution);
/* re-simulate n particles from t=0 over retained noisy positions xt [T][3][n] with elasticities e[n]
 * (what Gen.update does inside Gen.mh, particle_filter.jl:60).  mu_out [T][3][n] = simulator output after
 * each step, logp_out[n] = sum_{s>=2} log N(xt_s; mu_s, sigma_tr), state_out [n][13].  Any output may be NULL. */
int smcphys_simulate(int32_t device, int32_t shape, int32_t S, double ceiling_half, double sigma_tr,
                     const double init_pos[3], const double init_vel[3], const float* e, const float* xt, int64_t n,
                     int32_t T, float* mu_out, double* logp_out, float* state_out);
/* systematic resampling of n log weights with uniform offset u0 in [0,1): idx_out[k] = ancestor of slot k.
 * Stands in for Gen.maybe_resample! (particle_filter.jl:55); see DESIGN.md for why it is systematic. */
int smcphys_resample_systematic(int32_t device, const double* log_w, int64_t n, double u0, int64_t* idx_out);
/* multinomial parents as Gen.maybe_resample! draws them (particle_filter.jl:55; Gen 0.4.7): slot k takes 64 Philox bits
 * u_k of counter (k, t) and the ancestor min{ i : C_i > floor(u_k W / 2^64) } on the same integer CDF. */
int smcphys_resample_multinomial(int32_t device, const double* log_w, int64_t n, uint64_t seed, int32_t t,
                                 int64_t* idx_out);
/* Allocation-free variant of smcphys_resample_systematic.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  d_log_w and
 * d_idx_out are DEVICE pointers on `device` (n doubles in, n int64 out); the work is enqueued on cuda_stream
 * (cudaStream_t as void*, NULL = default stream) and NOT synchronised.  All scratch (integer CDF, ancestors, tile sums,
 * control block: 12.6 MB) lives in one static device array of the library, so n <= SMCPHYS_NOALLOC_MAX_N (larger n is
 * rejected with SMCPHYS_EINVAL), every GPU that loads the library holds the array until exit, and every call shares it:
 * calls that overlap on one GPU race - order them on one stream.  Calls no cudaMalloc / cudaFree / allocating routine.
 * Same indices as smcphys_resample_systematic, bit for bit. */
#define SMCPHYS_NOALLOC_MAX_N (1 << 20)
int smcphys_resample_systematic_noalloc(int32_t device, const double* d_log_w, int64_t n, double u0, int64_t* d_idx_out,
                                        void* cuda_stream);
/* out3 = {max, ESS, LSE}; sums3 = {W, S1, S2} */
int smcphys_weight_stats(int32_t device, const double* log_w, int64_t n, double* out3, uint64_t* sums3);

/* ---- per-kernel device timing (CUDA events on the handle's stream around each launch of a kernel class) */
#define SMCPHYS_K_EXTEND 0     /* k_extend: noise draw + observation weight + PhySMC.step */
#define SMCPHYS_K_WEIGHTS 1    /* k_weight_sums + k_cdf + k_index: LSE / ESS / integer CDF / systematic indices */
#define SMCPHYS_K_GATHER 2     /* k_gather + k_gather_hist (or k_pack / k_unpack on a sharded filter) */
#define SMCPHYS_K_REJUVENATE 3 /* k_rejuvenate: re-simulate + Metropolis-Hastings */
#define SMCPHYS_K_CLASSES 4
int smcphys_set_profiling(smcphys_handle* h, int32_t on);
/* synchronises, then adds the elapsed time of every recorded launch group of `kclass` since the last call */
int smcphys_kernel_time(smcphys_handle* h, int32_t kclass, double* ms_total, int64_t* groups);

/* measured FP32 FMA throughput of the device in TFLOP/s: the roofline denominator of the simulate kernels */
int smcphys_fp32_peak(int32_t device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
