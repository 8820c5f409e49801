// ORACLE — TEST INFRASTRUCTURE ONLY (see smc_oracle.hpp).  C entry points for ctypes.
#include "smc_oracle.hpp"
#include <cstdio>
using namespace oracle;

template <class T>
static void sim_raw(int shape, double e, double static_rest, double ceiling_half, int nbox, const double* st0, int nsub, double* out) {
  Scene<T> sc((T)ceiling_half, (T)static_rest);
  sc.nbox = nbox;
  State<T> s;
  s.x = {(T)st0[0], (T)st0[1], (T)st0[2]};
  s.q = {(T)st0[3], (T)st0[4], (T)st0[5], (T)st0[6]};
  s.v = {(T)st0[7], (T)st0[8], (T)st0[9]};
  s.w = {(T)st0[10], (T)st0[11], (T)st0[12]};
  PlaneManifold<T> pm;
  for (int k = 0; k < nsub; k++) {
    substep(sc, shape, (T)e, s, pm);
    double* o = out + 13 * k;
    o[0] = s.x.x; o[1] = s.x.y; o[2] = s.x.z; o[3] = s.q.x; o[4] = s.q.y; o[5] = s.q.z; o[6] = s.q.w;
    o[7] = s.v.x; o[8] = s.v.y; o[9] = s.v.z; o[10] = s.w.x; o[11] = s.w.y; o[12] = s.w.z;
  }
}

template <class T>
static void macro_rollout(int shape, double e, double ceiling_half, int S, const double* st0, int nsteps, double* out) {
  Scene<T> sc((T)ceiling_half);
  State<T> s;
  s.x = {(T)st0[0], (T)st0[1], (T)st0[2]};
  s.q = {(T)st0[3], (T)st0[4], (T)st0[5], (T)st0[6]};
  s.v = {(T)st0[7], (T)st0[8], (T)st0[9]};
  s.w = {(T)st0[10], (T)st0[11], (T)st0[12]};
  for (int k = 0; k < nsteps; k++) {
    macro_step(sc, shape, (T)e, S, s);
    double* o = out + 13 * k;
    o[0] = s.x.x; o[1] = s.x.y; o[2] = s.x.z; o[3] = s.q.x; o[4] = s.q.y; o[5] = s.q.z; o[6] = s.q.w;
    o[7] = s.v.x; o[8] = s.v.y; o[9] = s.v.z; o[10] = s.w.x; o[11] = s.w.y; o[12] = s.w.z;
  }
}

// batch re-simulation with retained noisy positions: xt is [T][3][n] float32, e is [n]
template <class T>
static void simulate(int shape, int S, double ceiling_half, double sigma_tr, const double* init_pos, const double* init_vel,
                     const float* e, const float* xt, int64_t n, int Tn, float* mu_out, double* logp_out, double* state_out,
                     int threads) {
  Scene<T> sc((T)ceiling_half);
  SmcConfig cfg;
  cfg.shape = shape; cfg.S = S; cfg.sigma_tr = sigma_tr;
  Particles<T> P;
  P.N = n; P.T_ = Tn;
  P.xt.resize((size_t)Tn * 3 * n);
  for (size_t k = 0; k < P.xt.size(); k++) P.xt[k] = (T)xt[k];
  parallel_for(n, threads, 16, [&](int64_t i) {
    std::vector<T> traj((size_t)Tn * 3);
    State<T> s;
    double lt = resimulate(sc, cfg, init_pos, init_vel, (double)e[i], P, i, Tn, s, traj.data());
    if (logp_out) logp_out[i] = lt;
    if (mu_out)
      for (int k = 0; k < Tn; k++)
        for (int c = 0; c < 3; c++) mu_out[((size_t)k * 3 + c) * n + i] = (float)traj[k * 3 + c];
    if (state_out) {
      double* o = state_out + 13 * i;
      o[0] = s.x.x; o[1] = s.x.y; o[2] = s.x.z; o[3] = s.q.x; o[4] = s.q.y; o[5] = s.q.z; o[6] = s.q.w;
      o[7] = s.v.x; o[8] = s.v.y; o[9] = s.v.z; o[10] = s.w.x; o[11] = s.w.y; o[12] = s.w.z;
    }
  });
}

// extras of one run: score[N]; predict (n_pred steps: pos/obs [n_pred][3][N]); mcmc (runs moves at the last step,
// e_trace [runs][N]) - all optional
struct Extras { double* score; int n_pred; float* pred_pos; float* pred_obs; int mcmc_runs; double* e_trace; };

template <class T>
static void smc(const SmcConfig& cfg, const double* init_pos, const double* init_vel, const double* obs, int Tn,
                double* e_out, double* logw_out, double* ltrans_out, float* traj_out, double* stats_out, double* log_ml,
                const Extras* ex = nullptr) {
  Particles<T> P;
  std::vector<StepStats> st(Tn);
  run_smc<T>(cfg, init_pos, init_vel, obs, Tn, P, st.data(), log_ml);
  if (ex && ex->mcmc_runs > 0) mcmc_moves<T>(cfg, init_pos, init_vel, Tn, P, 0, ex->mcmc_runs, ex->e_trace);
  if (ex && ex->score)
    for (int64_t i = 0; i < cfg.N; i++) ex->score[i] = trace_score<T>(cfg, init_pos, obs, Tn, P, i);
  if (ex && ex->n_pred > 0) predict<T>(cfg, P, Tn, ex->n_pred, ex->pred_pos, ex->pred_obs);
  for (int64_t i = 0; i < cfg.N; i++) {
    if (e_out) e_out[i] = P.e[i];
    if (logw_out) logw_out[i] = P.logw[i];
    if (ltrans_out) ltrans_out[i] = P.ltrans[i];
  }
  if (traj_out) {
    Scene<T> sc((T)cfg.ceiling_half);
    parallel_for(cfg.N, cfg.threads, 16, [&](int64_t i) {
      std::vector<T> traj((size_t)Tn * 3);
      State<T> s;
      resimulate(sc, cfg, init_pos, init_vel, P.e[i], P, i, Tn, s, traj.data());
      for (int k = 0; k < Tn; k++)
        for (int c = 0; c < 3; c++) traj_out[((size_t)k * 3 + c) * cfg.N + i] = (float)traj[k * 3 + c];
    });
  }
  if (stats_out)
    for (int t = 0; t < Tn; t++) {
      stats_out[4 * t + 0] = st[t].ess; stats_out[4 * t + 1] = st[t].lse;
      stats_out[4 * t + 2] = st[t].accept_rate; stats_out[4 * t + 3] = st[t].resampled;
    }
}

extern "C" {
// raw 1/240 s substeps with a persistent manifold (what `for _ in range(200): p.stepSimulation()` does,
// Tests/SimulatorComparison/tests.py:132-134).  st0/out rows: x(3) q_xyzw(4) v(3) w(3).
void oracle_sim_raw(int prec32, int shape, double e, double static_rest, double ceiling_half, int nbox, const double* st0,
                    int nsub, double* out) {
  if (prec32) sim_raw<float>(shape, e, static_rest, ceiling_half, nbox, st0, nsub, out);
  else sim_raw<double>(shape, e, static_rest, ceiling_half, nbox, st0, nsub, out);
}
// noise-free PhySMC.step rollout (Tests/generate_trajectories.jl:184-205)
void oracle_macro_rollout(int prec32, int shape, double e, double ceiling_half, int S, const double* st0, int nsteps,
                          double* out) {
  if (prec32) macro_rollout<float>(shape, e, ceiling_half, S, st0, nsteps, out);
  else macro_rollout<double>(shape, e, ceiling_half, S, st0, nsteps, out);
}
void oracle_simulate(int prec32, int shape, int S, double ceiling_half, double sigma_tr, const double* init_pos,
                     const double* init_vel, const float* e, const float* xt, int64_t n, int Tn, float* mu_out,
                     double* logp_out, double* state_out, int threads) {
  if (prec32) simulate<float>(shape, S, ceiling_half, sigma_tr, init_pos, init_vel, e, xt, n, Tn, mu_out, logp_out, state_out, threads);
  else simulate<double>(shape, S, ceiling_half, sigma_tr, init_pos, init_vel, e, xt, n, Tn, mu_out, logp_out, state_out, threads);
}
void oracle_run_smc(int prec32, int shape, int64_t N, int R, int S, double sigma_obs, double sigma_tr, double prop_sigma,
                    uint64_t seed, double ceiling_half, int threads, const double* init_pos, const double* init_vel,
                    const double* obs, int Tn, double* e_out, double* logw_out, double* ltrans_out, float* traj_out,
                    double* stats_out, double* log_ml) {
  SmcConfig cfg;
  cfg.shape = shape; cfg.N = N; cfg.R = R; cfg.S = S; cfg.sigma_obs = sigma_obs; cfg.sigma_tr = sigma_tr;
  cfg.prop_sigma = prop_sigma; cfg.seed = seed; cfg.ceiling_half = ceiling_half; cfg.threads = threads > 0 ? threads : 1;
  if (prec32) smc<float>(cfg, init_pos, init_vel, obs, Tn, e_out, logw_out, ltrans_out, traj_out, stats_out, log_ml);
  else smc<double>(cfg, init_pos, init_vel, obs, Tn, e_out, logw_out, ltrans_out, traj_out, stats_out, log_ml);
}
// run_smc with the optional legs: resampler (0 systematic, 1 multinomial), no_resample (mcmc.jl generate), score,
// predict, MCMC moves at the last step
void oracle_run_smc_ex(int prec32, int shape, int64_t N, int R, int S, double sigma_obs, double sigma_tr, double prop_sigma,
                       uint64_t seed, double ceiling_half, int threads, int resampler, int no_resample,
                       const double* init_pos, const double* init_vel, const double* obs, int Tn, double* e_out,
                       double* logw_out, double* ltrans_out, float* traj_out, double* stats_out, double* log_ml,
                       double* score_out, int n_pred, float* pred_pos, float* pred_obs, int mcmc_runs, double* e_trace) {
  SmcConfig cfg;
  cfg.shape = shape; cfg.N = N; cfg.R = R; cfg.S = S; cfg.sigma_obs = sigma_obs; cfg.sigma_tr = sigma_tr;
  cfg.prop_sigma = prop_sigma; cfg.seed = seed; cfg.ceiling_half = ceiling_half; cfg.threads = threads > 0 ? threads : 1;
  cfg.resampler = resampler; cfg.no_resample = no_resample;
  Extras ex{score_out, n_pred, pred_pos, pred_obs, mcmc_runs, e_trace};
  if (prec32) smc<float>(cfg, init_pos, init_vel, obs, Tn, e_out, logw_out, ltrans_out, traj_out, stats_out, log_ml, &ex);
  else smc<double>(cfg, init_pos, init_vel, obs, Tn, e_out, logw_out, ltrans_out, traj_out, stats_out, log_ml, &ex);
}
void oracle_resample_multinomial(const double* lw, int64_t n, uint64_t seed, int t, int64_t* idx) { resample_multinomial(lw, n, seed, t, idx); }
void oracle_resample_systematic(const double* lw, int64_t n, double u0, int64_t* idx) { resample_systematic(lw, n, u0, idx); }
void oracle_resample_systematic_f64(const double* lw, int64_t n, double u0, int64_t* idx) { resample_systematic_f64(lw, n, u0, idx); }
// out: mx, ess, lse ; iout: W, S1, S2, below_half
void oracle_weight_stats(const double* lw, int64_t n, double* out, uint64_t* iout) {
  WeightStats s = weight_stats(lw, n);
  out[0] = s.mx; out[1] = ess_of(s); out[2] = lse_of(s);
  iout[0] = s.W; iout[1] = s.S1; iout[2] = s.S2; iout[3] = ess_below(s, n, 2) ? 1 : 0;
}
void oracle_weight_q(const double* lw, int64_t n, double mx, uint64_t* q) { for (int64_t i = 0; i < n; i++) q[i] = weight_q(lw[i], mx); }
uint64_t oracle_sys_threshold(uint64_t U0, uint64_t k, uint64_t W, uint64_t N) { return sys_threshold(U0, k, W, N); }
double oracle_det_exp(double x) { return det_exp(x); }
void oracle_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t* out) { Philox::gen(seed, c0, c1, c2, c3, out); }
double oracle_u01(uint32_t bits) { return u01(bits); }
void oracle_box_muller(uint32_t a, uint32_t b, double* z) { box_muller(a, b, z[0], z[1]); }
double oracle_truncnorm_logpdf(double x, double mu, double sigma, double lo, double hi) { return truncnorm_logpdf(x, mu, sigma, lo, hi); }
double oracle_truncnorm_sample(double u, double mu, double sigma, double lo, double hi) { return truncnorm_sample(u, mu, sigma, lo, hi); }
double oracle_normal3_logpdf(const double* x, const double* mu, double sigma) { return normal3_logpdf(x, mu, sigma); }
int oracle_substeps_for(double step_dur) { return substeps_for(step_dur); }
}
