// libsmcphys: CUDA kernels (sm_100a) and the C ABI of include/smcphys.h.
//
// Data layout in HBM (n = particles of this shard, all SoA so that a warp touches 128 consecutive bytes):
//   xt[b]   float  [Tmax][3][n]   retained noisy positions x~_s  (:trajectory=>s=>:state=>1=>:position)
//   st[b]   float  [13][n]        simulator output after the last step: position, quaternion xyzw, v, omega
//   e[b]    double [n]            elasticity latent (:latents=>1=>:restitution)
//   qs[b]   double [n]            sum_{s>=2} |x~_s - mu_s|^2  (transition log density up to constants)
//   logw    double [n]            particle-filter log weights
//   cdf     uint64 [n]            inclusive integer CDF of the weights;  idx int32 [n] ancestors
// b in {0,1}: resampling gathers from buffer `cur` into the other one and flips `cur` on the device, so the
// host never has to wait for the ESS decision.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/smcphys.h"
#include "nccl_dl.h"
#include "physics.cuh"
#include "rng.cuh"

using namespace smcphys;
typedef unsigned long long u64;

// ---------------------------------------------------------------- errors, launch accounting
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches(0);

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CK(expr)                                                                                   \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return fail(SMCPHYS_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));              \
  } while (0)
#define LAUNCH(kernel, grid, block, stream, ...)                                                   \
  do {                                                                                             \
    kernel<<<grid, block, 0, stream>>>(__VA_ARGS__);                                               \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                            \
    cudaError_t _e = cudaGetLastError();                                                           \
    if (_e != cudaSuccess) return fail(SMCPHYS_ECUDA, std::string(#kernel) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// ---------------------------------------------------------------- device control block
struct Ctl {
  u64 max_enc;        // order-preserving encoding of the largest local log weight
  u64 W, S1, S2;      // integer weight sums of this shard
  double gmax;        // the maximum the integer weights are taken against
  double log_ml;      // running log marginal likelihood estimate
  unsigned int done;  // last-block-done counter
  int cur;            // buffer holding the particles
  int resampled;      // decision of the current step
  int pad;
  // sharded filter: written by k_decide from the all-gathered sums, identical decisions on every rank
  u64 W_before, W_total;  // integer weight of all lower ranks / of all ranks
  u64 U0;                 // uniform offset of this step's systematic resampling, 32-bit fixed point
  long long k_lo, k_hi;   // global slots [k_lo, k_hi) filled by the offspring of this shard's particles
  double lmax;            // this shard's largest log weight, read by the other shards of a one-process group
};

// device-visible buffers of one rank of a sharded filter (own pointers, peer-mapped or IPC-opened ones)
struct Peer {
  float* xt[2];
  float* st[2];
  double* e[2];
  double* qs[2];
  u64* cdf;
  struct Ctl* ctl;
  u64* sums;  // W, S1, S2 of that rank (k_weight_sums)
};

struct TrialPar {
  float x0[3], v0[3];
  double sigma_obs, sigma_tr;
};

struct Bufs {
  float* xt[2];
  float* st[2];
  double* e[2];
  double* qs[2];
  double* logw;
  u64* cdf;
  u64* tile_sum;
  u64* tile_off;
  int* idx;
};

__device__ __forceinline__ u64 enc_max(double d) {
  u64 b = (u64)__double_as_longlong(d);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dec_max(u64 b) {
  b = (b >> 63) ? (b & 0x7fffffffffffffffull) : ~b;
  return __longlong_as_double((long long)b);
}

__device__ __forceinline__ u64 warp_sum(u64 v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ void load_body(const float* st, int64_t n, int64_t i, Body& s) {
  s.x = mk3(st[0 * n + i], st[1 * n + i], st[2 * n + i]);
  s.q.x = st[3 * n + i]; s.q.y = st[4 * n + i]; s.q.z = st[5 * n + i]; s.q.w = st[6 * n + i];
  s.v = mk3(st[7 * n + i], st[8 * n + i], st[9 * n + i]);
  s.w = mk3(st[10 * n + i], st[11 * n + i], st[12 * n + i]);
}
__device__ __forceinline__ void store_body(float* st, int64_t n, int64_t i, const Body& s) {
  st[0 * n + i] = s.x.x; st[1 * n + i] = s.x.y; st[2 * n + i] = s.x.z;
  st[3 * n + i] = s.q.x; st[4 * n + i] = s.q.y; st[5 * n + i] = s.q.z; st[6 * n + i] = s.q.w;
  st[7 * n + i] = s.v.x; st[8 * n + i] = s.v.y; st[9 * n + i] = s.v.z;
  st[10 * n + i] = s.w.x; st[11 * n + i] = s.w.y; st[12 * n + i] = s.w.z;
}
__device__ __forceinline__ void init_body(const TrialPar& tp, Body& s) {
  s.x = mk3(tp.x0[0], tp.x0[1], tp.x0[2]);
  s.q.x = 0.f; s.q.y = 0.f; s.q.z = 0.f; s.q.w = 1.f;  // bouncing_object.jl:75
  s.v = mk3(tp.v0[0], tp.v0[1], tp.v0[2]);
  s.w = mk3(0.f, 0.f, 0.f);
}

#define PF_BLOCK 128
#define TILE_THREADS 256
#define TILE_ITEMS 4
#define TILE (TILE_THREADS * TILE_ITEMS)

// ---------------------------------------------------------------- K0: prior + control
// initialize_particle_filter with T = 0 (Inference/particle_filter.jl:44): e ~ U(0,1), log w = 0.
__global__ void k_init(Bufs B, Ctl* ctl, TrialPar tp, int64_t n, int64_t goff, u64 seed) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    ctl->cur = 0; ctl->resampled = 0; ctl->done = 0; ctl->log_ml = 0.0;
    ctl->W = ctl->S1 = ctl->S2 = 0; ctl->max_enc = 0; ctl->gmax = 0.0;
  }
  if (i >= n) return;
  uint32_t r[4];
  philox4x32_10(seed, (uint32_t)(goff + i), 0u, 0u, (uint32_t)ST_PRIOR, r);
  B.e[0][i] = u01(r[0]);
  B.qs[0][i] = 0.0;
  B.logw[i] = 0.0;
  Body s;
  init_body(tp, s);
  store_body(B.st[0], n, i, s);
}

// start of a PF step: commit the buffer switch of the previous step, clear the reductions
__global__ void k_step_begin(Ctl* ctl) {
  ctl->cur ^= ctl->resampled;
  ctl->resampled = 0;
  ctl->max_enc = 0;  // below enc_max(-inf)
  ctl->W = ctl->S1 = ctl->S2 = 0;
  ctl->done = 0;
}

// ---------------------------------------------------------------- K1: extend + weight + step
// One particle per thread.  kernel() of Model/bouncing_object.jl:132-145 under Gen.particle_filter_step!
// (Inference/particle_filter.jl:52): x~_t ~ N(mu_t, sigma_tr) (:135, resample_state :109-120), log w +=
// log N(y_t; x~_t, sigma_obs) (:139, sample_observation :123-128), then PhySMC.step from x~_t (:142).
// The block maximum of the new log weights goes to ctl->max_enc; the last block publishes it as ctl->gmax.
template <int SHAPE>
__global__ void __launch_bounds__(PF_BLOCK) k_extend(Bufs B, Ctl* ctl, TrialPar tp, ScenePar sp, int64_t n,
                                                     int64_t goff, u64 seed, int t, int S, double y0, double y1,
                                                     double y2) {
  __shared__ double smax[PF_BLOCK / 32];
  const int cur = ctl->cur;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double lw = -INFINITY;
  if (i < n) {
    float* st = B.st[cur];
    Body s;
    load_body(st, n, i, s);
    uint32_t r[4];
    philox4x32_10(seed, (uint32_t)(goff + i), (uint32_t)t, 0u, (uint32_t)ST_TRANSITION, r);
    double z[4];
    box_muller(r[0], r[1], z[0], z[1]);
    box_muller(r[2], r[3], z[2], z[3]);
    const double mu[3] = {(double)s.x.x, (double)s.x.y, (double)s.x.z};
    const double y[3] = {y0, y1, y2};
    float xf[3];
    double d2 = 0.0, o2 = 0.0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
      xf[c] = (float)(mu[c] + tp.sigma_tr * z[c]);
      double d = (double)xf[c] - mu[c];
      d2 += d * d;
      double zo = (y[c] - (double)xf[c]) / tp.sigma_obs;
      o2 += -0.5 * zo * zo;
    }
    float* xt = B.xt[cur];
#pragma unroll
    for (int c = 0; c < 3; c++) xt[((int64_t)t * 3 + c) * n + i] = xf[c];
    if (t >= 1) B.qs[cur][i] += d2;
    lw = B.logw[i] + (o2 - 3.0 * log(tp.sigma_obs) - 3.0 * SMC_HALF_LOG_2PI);
    B.logw[i] = lw;
    s.x = mk3(xf[0], xf[1], xf[2]);
    // a sphere's orientation never feeds back into its motion and no reader uses the stored one (the state getter
    // re-simulates with orientation tracking): not integrated here either
    macro_step<SHAPE, SHAPE == SHAPE_CUBE>(sp, (float)B.e[cur][i], S, s);
    store_body(st, n, i, s);
  }
  lw = warp_max(lw);
  if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = lw;
  __syncthreads();
  if (threadIdx.x == 0) {
    double m = smax[0];
#pragma unroll
    for (int w = 1; w < PF_BLOCK / 32; w++) m = fmax(m, smax[w]);
    atomicMax(&ctl->max_enc, enc_max(m));
    __threadfence();
    unsigned int d = atomicAdd(&ctl->done, 1u);
    if (d == gridDim.x - 1) {
      ctl->gmax = dec_max(atomicMax(&ctl->max_enc, 0ull));
      ctl->done = 0;
    }
  }
}

// ---------------------------------------------------------------- K2: integer weights, sums, decision
// q_i = floor(exp(lw_i - max) 2^39); W = sum q, S1 = sum (q>>20), S2 = sum (q>>20)^2 (integer => the same for
// any reduction order).  Stands in for the normalisation / effective_sample_size inside Gen.maybe_resample!
// (Inference/particle_filter.jl:55).  Each block also leaves the sum of its tile for the CDF scan; the last
// block scans the tile sums and (decide != 0, single shard) takes the ESS < N*num/den decision on the device.
__global__ void __launch_bounds__(TILE_THREADS) k_weight_sums(Bufs B, Ctl* ctl, const double* gmax_ptr, int64_t n,
                                                              int ntiles, int decide, int64_t ess_num,
                                                              int64_t ess_den, int t, double* st_ess, double* st_lse,
                                                              int* st_res, u64* sums_out) {
  __shared__ u64 sh[3][TILE_THREADS / 32];
  __shared__ int is_last;
  const double gmax = *gmax_ptr;
  u64 w = 0, s1 = 0, s2 = 0;
  int64_t base = (int64_t)blockIdx.x * TILE;
#pragma unroll
  for (int k = 0; k < TILE_ITEMS; k++) {
    int64_t i = base + k * TILE_THREADS + threadIdx.x;
    if (i < n) {
      u64 q = weight_q(B.logw[i], gmax), q19 = q >> 20;
      w += q; s1 += q19; s2 += q19 * q19;
    }
  }
  w = warp_sum(w); s1 = warp_sum(s1); s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = w; sh[1][threadIdx.x >> 5] = s1; sh[2][threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    w = s1 = s2 = 0;
#pragma unroll
    for (int k = 0; k < TILE_THREADS / 32; k++) { w += sh[0][k]; s1 += sh[1][k]; s2 += sh[2][k]; }
    B.tile_sum[blockIdx.x] = w;
    atomicAdd(&ctl->W, w); atomicAdd(&ctl->S1, s1); atomicAdd(&ctl->S2, s2);
    __threadfence();
    unsigned int d = atomicAdd(&ctl->done, 1u);
    is_last = (d == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // exclusive scan of the tile sums by this block (ntiles <= 16384)
  __shared__ u64 carry;
  __shared__ u64 wtot[TILE_THREADS / 32];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < ntiles; b0 += TILE_THREADS) {
    int j = b0 + threadIdx.x;
    u64 v = (j < ntiles) ? ((volatile u64*)B.tile_sum)[j] : 0, x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      u64 y = __shfl_up_sync(0xffffffffu, x, o);
      if ((threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = x;
    __syncthreads();
    u64 woff = 0;
    for (int k = 0; k < (threadIdx.x >> 5); k++) woff += wtot[k];
    u64 c = carry;
    if (j < ntiles) B.tile_off[j] = c + woff + x - v;
    __syncthreads();
    if (threadIdx.x == TILE_THREADS - 1) carry = c + woff + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    u64 W = atomicAdd(&ctl->W, 0ull), S1 = atomicAdd(&ctl->S1, 0ull), S2 = atomicAdd(&ctl->S2, 0ull);
    ctl->done = 0;
    if (sums_out) { sums_out[0] = W; sums_out[1] = S1; sums_out[2] = S2; }
    if (decide) {
      unsigned __int128 lhs = (unsigned __int128)S1 * S1 * (u64)ess_den;
      unsigned __int128 rhs = (unsigned __int128)S2 * (u64)n * (u64)ess_num;
      int res = lhs < rhs;
      double lse = gmax + log((double)W / 549755813888.0);
      ctl->resampled = res;
      if (res) ctl->log_ml += lse - log((double)n);
      if (st_ess) {
        st_ess[t] = S2 ? ((double)S1 * (double)S1) / (double)S2 : 0.0;
        st_lse[t] = lse;
        st_res[t] = res;
      }
    }
  }
}

// ---------------------------------------------------------------- K3: integer CDF and systematic indices
// C_i = W_before + inclusive prefix sum of q.  Skipped when the step does not resample (gate != 0 reads the
// device decision).
__global__ void __launch_bounds__(TILE_THREADS) k_cdf(Bufs B, const Ctl* ctl, const double* gmax_ptr, int64_t n,
                                                      u64 W_before, int gate, const u64* W_before_ptr = nullptr) {
  if (gate && !ctl->resampled) return;
  if (W_before_ptr) W_before = *W_before_ptr;
  __shared__ u64 wtot[TILE_THREADS / 32];
  const double gmax = *gmax_ptr;
  // blocked arrangement: thread owns TILE_ITEMS consecutive particles
  int64_t base = (int64_t)blockIdx.x * TILE + (int64_t)threadIdx.x * TILE_ITEMS;
  u64 q[TILE_ITEMS], tsum = 0;
#pragma unroll
  for (int k = 0; k < TILE_ITEMS; k++) {
    int64_t i = base + k;
    q[k] = (i < n) ? weight_q(B.logw[i], gmax) : 0;
    tsum += q[k];
  }
  u64 x = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    u64 y = __shfl_up_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) >= o) x += y;
  }
  if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = x;
  __syncthreads();
  u64 off = W_before + B.tile_off[blockIdx.x] + x - tsum;
  for (int k = 0; k < (threadIdx.x >> 5); k++) off += wtot[k];
#pragma unroll
  for (int k = 0; k < TILE_ITEMS; k++) {
    int64_t i = base + k;
    off += q[k];
    if (i < n) B.cdf[i] = off;
  }
}

// idx[j] = min{ i : C_i > tau_k }, k = k_lo + j; tau_k = floor((U0 + k 2^32) W / (N 2^32)).
__global__ void k_index(Bufs B, const Ctl* ctl, int64_t n, int64_t Nglob, u64 U0, const u64* W_ptr, u64 W_val,
                        int64_t k_lo, int64_t count, int gate) {
  if (gate && !ctl->resampled) return;
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= count) return;
  const u64 W = W_ptr ? *W_ptr : W_val;
  u64 tau = sys_threshold(U0, (u64)(k_lo + j), W, (u64)Nglob);
  int64_t lo = 0, hi = n - 1;  // answer in [lo, hi]; clamps to n-1 like the sequential definition
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (B.cdf[mid] > tau) hi = mid; else lo = mid + 1;
  }
  B.idx[j] = (int)lo;
}

// ---------------------------------------------------------------- K4: gather offspring
// traces := traces[parents]; log weights := 0 (Gen.maybe_resample!, particle_filter.jl:55).
// Offspring j = 0..count-1 (ancestor idx[j]) go to the local slots slot0 + j.  gate != 0: skip unless the device-side
// decision of this step says "resample" (single shard); the sharded filter decides on the host and passes 0.
__global__ void k_gather(Bufs B, const Ctl* ctl, int64_t n, int64_t count, int64_t slot0, int gate) {
  if (gate && !ctl->resampled) return;
  const int cur = ctl->cur, nxt = cur ^ 1;
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= count) return;
  const int64_t a = B.idx[j], k = slot0 + j;
  B.e[nxt][k] = B.e[cur][a];
  B.qs[nxt][k] = B.qs[cur][a];
  B.logw[k] = 0.0;
#pragma unroll
  for (int c = 0; c < 13; c++) B.st[nxt][c * n + k] = B.st[cur][c * n + a];
}
// history rows 0..t.  Systematic resampling keeps the ancestors sorted, so a warp reads (almost) consecutive
// addresses; each thread moves GH_ROWS x 3 floats of one particle with all loads in flight before the first store,
// and reads its ancestor index once per GH_ROWS rows.  blockIdx.y = group of GH_ROWS rows.
#define GH_ROWS 8
__global__ void __launch_bounds__(256) k_gather_hist(Bufs B, const Ctl* ctl, int64_t n, int rows, int64_t count,
                                                     int64_t slot0, int gate) {
  if (gate && !ctl->resampled) return;
  const int cur = ctl->cur, nxt = cur ^ 1;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= count) return;
  const int64_t a = B.idx[j], k = slot0 + j;
  const int c0 = (int)blockIdx.y * (GH_ROWS * 3), c1 = rows * 3;
  const float* __restrict__ src = B.xt[cur] + (int64_t)c0 * n + a;
  float* __restrict__ dst = B.xt[nxt] + (int64_t)c0 * n + k;
  float v[GH_ROWS * 3];
#pragma unroll
  for (int c = 0; c < GH_ROWS * 3; c++)
    if (c0 + c < c1) v[c] = __ldg(src + (int64_t)c * n);
#pragma unroll
  for (int c = 0; c < GH_ROWS * 3; c++)
    if (c0 + c < c1) __stcs(dst + (int64_t)c * n, v[c]);
}

// ---------------------------------------------------------------- re-simulation from t = 0
// What Gen.update does to a trace when the elasticity changes (Gen.mh at Inference/particle_filter.jl:60):
// generate_trajectory (Model/bouncing_object.jl:148-162) re-runs kernel() for s = 1..t with the retained x~_s:
// the body is teleported to x~_s, keeps the velocity / orientation the simulator produced, and is stepped.
// Returns sum_{s>=2} |x~_s - mu'_s|^2.  The next row of x~ is fetched before the current macro step so the
// load latency hides behind the arithmetic.
#ifdef RS_STEP_CLOCK
// diagnostic build only (tools/bench_kernel.py): SM cycles the warps of a launch spend in each macro step of the chain
__device__ unsigned long long g_step_clk[4096];
#endif
template <int SHAPE, bool TRACK_Q, bool TRAJ>
__device__ __forceinline__ double resimulate(const ScenePar& sp, const TrialPar& tp, int S, float e,
                                             const float* __restrict__ xt, int64_t n, int64_t i, int steps, Body& s,
                                             float* traj) {
  init_body(tp, s);
  double qs = 0.0;
  // one pointer walks the rows (3 n floats apart); the row of the next step is pulled into L1 while the current macro
  // step runs and is loaded at the top of the next iteration (held in registers across the macro step it was spilled
  // right after the load, and the spill store waited for the load: 11 % of the stall samples, profiles/r01c)
  const float* pk = xt + i;
  float* o = TRAJ ? traj + i : nullptr;
  const int64_t row = 3 * n;
#pragma unroll 1
  for (int k = steps; k > 0; k--) {
    const float x = pk[0], y = pk[n], z = pk[2 * n];
    pk += row;
    if (k > 1) {
      asm volatile("prefetch.global.L1 [%0];" ::"l"(pk));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(pk + n));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(pk + 2 * n));
    }
    if (k != steps) {
      float dx = x - s.x.x, dy = y - s.x.y, dz = z - s.x.z;
      qs += (double)dx * (double)dx + (double)dy * (double)dy + (double)dz * (double)dz;
    }
    s.x = mk3(x, y, z);
#ifdef RS_STEP_CLOCK
    const long long clk0 = clock64();
#endif
    macro_step<SHAPE, TRACK_Q>(sp, e, S, s);
#ifdef RS_STEP_CLOCK
    __syncwarp();
    if ((threadIdx.x & 31) == 0) atomicAdd(&g_step_clk[(steps - k) & 4095], (unsigned long long)(clock64() - clk0));
#endif
    if (TRAJ) {
      o[0] = s.x.x; o[n] = s.x.y; o[2 * n] = s.x.z;
      o += row;
    }
  }
  return qs;
}

// The sphere chain on one struct of loop state (physics.cuh, SphereChain): rows walked by one pointer, the row of the
// next step loaded while the current macro step runs (the loop itself needs ~30 registers, so holding the row costs
// nothing; an L1 prefetch instead left the load of the row as the top stall of the loop).  TRAJ: also write the
// simulator output after every step (the CSV trajectory).  Orientation is not integrated (see physics.cuh).
template <bool TRAJ>
__device__ __forceinline__ SphereChain sphere_chain_run(float ceil_half, f3 x0, f3 v0, int S, float e,
                                                        const float* __restrict__ xt, int64_t n, int64_t i, int steps,
                                                        float* traj) {
  SphereChain c;
  c.v = v0; c.w = mk3(0.f, 0.f, 0.f);
  c.restitution = e; c.ceil_half = ceil_half;
  c.zc = 1.01f - ceil_half - (PH_H + PH_THR);
  c.pk = xt + i; c.n = n; c.qs = 0.0; c.S = S;
  c.x = x0;
  c.nx = x0;
  float* o = TRAJ ? traj + i : nullptr;
  if (steps > 0) {
    c.x = mk3(c.pk[0], c.pk[c.n], c.pk[2 * c.n]);
#pragma unroll 1
    for (c.k = steps; ; ) {
      c.pk += 3 * c.n;
#ifdef RS_L1_PREFETCH
      if (c.k > 1) {
        asm volatile("prefetch.global.L1 [%0];" ::"l"(c.pk));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(c.pk + c.n));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(c.pk + 2 * c.n));
      }
#else
      if (c.k > 1) c.nx = mk3(c.pk[0], c.pk[c.n], c.pk[2 * c.n]);
#endif
#ifdef RS_STEP_CLOCK
      const long long clk0 = clock64();
#endif
#ifdef RS_UNROLL5
      if (c.S == 5) {
        c.sub = 0;
#pragma unroll
        for (int u = 0; u < 5; u++) sphere_chain_substep(c);
      } else
#endif
#pragma unroll 1
      for (c.sub = c.S; c.sub > 0; c.sub--) sphere_chain_substep(c);
#ifdef RS_STEP_CLOCK
      __syncwarp();
      if ((threadIdx.x & 31) == 0) atomicAdd(&g_step_clk[(steps - c.k) & 4095], (unsigned long long)(clock64() - clk0));
#endif
      if (TRAJ) {
        o[0] = c.x.x; o[c.n] = c.x.y; o[2 * c.n] = c.x.z;
        o += 3 * c.n;
      }
      if (--c.k == 0) break;
#ifdef RS_L1_PREFETCH
      c.nx = mk3(c.pk[0], c.pk[c.n], c.pk[2 * c.n]);
#endif
      const float dx = c.nx.x - c.x.x, dy = c.nx.y - c.x.y, dz = c.nx.z - c.x.z;
      c.qs += (double)dx * (double)dx + (double)dy * (double)dy + (double)dz * (double)dz;
      c.x = c.nx;
    }
  }
  return c;
}

// The sphere chain of the rejuvenation kernel as a call: inside, the register allocation sees nothing but the
// chain (body, elasticity, the prefetched row), while the Metropolis-Hastings state of the caller (five doubles,
// counters, pointers) is parked by the calling convention once per chain instead of being spilled in the substep loop.
struct ChainOut {
  double qs;
  f3 x, v, w;
  f4 q;
};
template <int SHAPE>
__device__ __noinline__ ChainOut resimulate_call(float ceil_half, f3 x0, f3 v0, int S, float e,
                                                 const float* __restrict__ xt, int64_t n, int64_t i, int steps) {
  ChainOut o;
  if (SHAPE == SHAPE_SPHERE) {
    SphereChain c = sphere_chain_run<false>(ceil_half, x0, v0, S, e, xt, n, i, steps, nullptr);
    o.qs = c.qs;
    o.x = c.x; o.v = c.v; o.w = c.w;
    o.q.x = o.q.y = o.q.z = 0.f; o.q.w = 1.f;
    return o;
  }
  ScenePar sp;
  sp.ceil_half = ceil_half;
  TrialPar tp;
  tp.x0[0] = x0.x; tp.x0[1] = x0.y; tp.x0[2] = x0.z;
  tp.v0[0] = v0.x; tp.v0[1] = v0.y; tp.v0[2] = v0.z;
  Body s;
  o.qs = resimulate<SHAPE, SHAPE == SHAPE_CUBE, false>(sp, tp, S, e, xt, n, i, steps, s, nullptr);
  o.x = s.x; o.v = s.v; o.w = s.w; o.q = s.q;
  return o;
}

#ifndef RJ_MINB_CUBE
#define RJ_MINB_CUBE 8  // cube: 32 warps at 64 registers.  With the contact loops rolled (physics.cuh) the kernel is
                        // bound by local-memory and arithmetic latency, which warps hide: 4/5/6/7/8 blocks per SM =
                        // 9.18 / 9.44 / 9.96 / 10.36 / 10.47 M particle-steps/s (2^18 particles)
#endif
#ifndef RJ_BLOCK
#define RJ_BLOCK 128
#endif
#ifndef RJ_MINB
#define RJ_MINB 8  // blocks per SM the register allocation must allow (32 warps at 64 registers)
#endif

// ---------------------------------------------------------------- K5: rejuvenation (re-simulate + MH)
// for i = 1:N, j = 1:R: Gen.mh(trace_i, proposal, ())  (Inference/particle_filter.jl:58-62):
//   e' ~ TruncNorm(e, 0.1, 0, 1) (proposal :17-29, Utilities/truncatednorm.jl:11-16); re-simulate steps 1..t+1
//   with the retained x~; accept iff log u < [sum_s logN(x~_s; mu'_s) - sum_s logN(x~_s; mu_s)] - log q(e'|e)
//   + log q(e|e').  The observation terms and the s = 1 term cancel; the PF log weight is untouched.
// One thread per particle, the R moves of a particle form a chain and run back to back in registers.  Warps are
// independent (no shared memory, no barriers): contact response that needs no iteration is in line, the rest is
// an out-of-line call (physics.cuh).  A sphere's orientation is not integrated here: it never feeds back.
template <int SHAPE>
__global__ void __launch_bounds__(RJ_BLOCK, SHAPE == SHAPE_SPHERE ? RJ_MINB : RJ_MINB_CUBE)
    k_rejuvenate(Bufs B, const Ctl* ctl, TrialPar tp, ScenePar sp, int64_t n, int64_t goff, u64 seed, int t, int S,
                 int R, int move0, double prop_sigma, u64* accept_out) {
  const int cur = ctl->cur ^ ctl->resampled;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned int acc = 0;
  if (i < n) {
    double e_old = B.e[cur][i], q_old = B.qs[cur][i];
    const double inv2s2 = 0.5 / (tp.sigma_tr * tp.sigma_tr);
    float* st = B.st[cur];
    const float* xt_cur = B.xt[cur];
    TruncZ z_old = truncnorm_z(e_old, prop_sigma, 0.0, 1.0);  // normaliser of the current elasticity, renewed on accept
    for (int j = 0; j < R; j++) {
      uint32_t r[4];
      philox4x32_10(seed, (uint32_t)(goff + i), (uint32_t)t, (uint32_t)(move0 + j), (uint32_t)ST_PROPOSAL, r);
      double log_qratio;  // log q(e | e') - log q(e' | e)
      TruncZ z_new;
      const double e_new = truncnorm_propose_z(u01(r[0]), e_old, z_old, prop_sigma, 0.0, 1.0, &z_new, &log_qratio);
      const double log_u = log(u01(r[1]));
      const ChainOut o = resimulate_call<SHAPE>(sp.ceil_half, mk3(tp.x0[0], tp.x0[1], tp.x0[2]),
                                                mk3(tp.v0[0], tp.v0[1], tp.v0[2]), S, (float)e_new, xt_cur, n, i, t + 1);
      const double q_new = o.qs;
      Body s;
      s.x = o.x; s.v = o.v; s.w = o.w; s.q = o.q;
      const double alpha = -(q_new - q_old) * inv2s2 + log_qratio;
      if (log_u < alpha) {
        // accepted: the simulator state goes out right away, so nothing but (e, q) stays live across the moves
        e_old = e_new;
        q_old = q_new;
        z_old = z_new;
        acc++;
        st[0 * n + i] = s.x.x; st[1 * n + i] = s.x.y; st[2 * n + i] = s.x.z;
        if (SHAPE == SHAPE_CUBE) {  // a sphere keeps its stored quaternion (see above)
          st[3 * n + i] = s.q.x; st[4 * n + i] = s.q.y; st[5 * n + i] = s.q.z; st[6 * n + i] = s.q.w;
        }
        st[7 * n + i] = s.v.x; st[8 * n + i] = s.v.y; st[9 * n + i] = s.v.z;
        st[10 * n + i] = s.w.x; st[11 * n + i] = s.w.y; st[12 * n + i] = s.w.z;
      }
    }
    if (acc) {
      B.e[cur][i] = e_old;
      B.qs[cur][i] = q_old;
    }
  }
  if (accept_out) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(accept_out, (u64)acc);
  }
}

// ---------------------------------------------------------------- getters / stateless kernels
template <int SHAPE>
__global__ void __launch_bounds__(RJ_BLOCK)
    k_traj(const double* e, const float* xt, TrialPar tp, ScenePar sp, int64_t n, int S, int T, float* mu_out,
           double* q_out, float* state_out) {
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= n) return;
  Body s;
  double qs;
  if (SHAPE == SHAPE_SPHERE && !state_out) {
    // the trajectory and the transition density of a sphere come from the SAME chain code the rejuvenation kernel
    // runs (so the trajectory parity tests pin that code); only the state getter needs the orientation
    const f3 x0 = mk3(tp.x0[0], tp.x0[1], tp.x0[2]), v0 = mk3(tp.v0[0], tp.v0[1], tp.v0[2]);
    const SphereChain c = mu_out ? sphere_chain_run<true>(sp.ceil_half, x0, v0, S, (float)e[i0], xt, n, i0, T, mu_out)
                                 : sphere_chain_run<false>(sp.ceil_half, x0, v0, S, (float)e[i0], xt, n, i0, T, nullptr);
    if (q_out) q_out[i0] = c.qs;
    return;
  }
  if (mu_out) qs = resimulate<SHAPE, true, true>(sp, tp, S, (float)e[i0], xt, n, i0, T, s, mu_out);
  else qs = resimulate<SHAPE, true, false>(sp, tp, S, (float)e[i0], xt, n, i0, T, s, nullptr);
  if (q_out) q_out[i0] = qs;
  if (state_out) store_body(state_out, n, i0, s);
}

// score = log joint of the trace: sum_s logN(x~_s; mu_s, sigma_tr) + sum_s logN(y_s; x~_s, sigma_obs); the
// U(0,1) prior contributes 0.  (get_score, written to the CSV `weight` column, Utilities/fileio.jl:267)
__global__ void k_score(const double* qs, const float* xt, const double* obs, TrialPar tp, int64_t n, int T,
                        double* score) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double q = qs[i], o = 0.0;
  for (int s = 0; s < T; s++) {
    for (int c = 0; c < 3; c++) {
      double x = (double)xt[((int64_t)s * 3 + c) * n + i];
      if (s == 0) { double d = x - (double)tp.x0[c]; q += d * d; }
      double z = obs[s * 3 + c] - x;
      o += z * z;
    }
  }
  double lt = -0.5 * q / (tp.sigma_tr * tp.sigma_tr) - 3.0 * T * (log(tp.sigma_tr) + SMC_HALF_LOG_2PI);
  double lo = -0.5 * o / (tp.sigma_obs * tp.sigma_obs) - 3.0 * T * (log(tp.sigma_obs) + SMC_HALF_LOG_2PI);
  score[i] = lt + lo;
}

// PhySMC.forward_step on a batch of bodies (state13 SoA [13][n])
template <int SHAPE>
__global__ void __launch_bounds__(PF_BLOCK) k_sim_step(float* st, const float* rest, ScenePar sp, int64_t n, int S) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Body s;
  load_body(st, n, i, s);
  macro_step<SHAPE, true>(sp, rest[i], S, s);
  store_body(st, n, i, s);
}

// predict (Inference/particle_filter.jl:84-102): extend every particle n_steps further, sampling transition and
// observation noise (no constraints).
template <int SHAPE>
__global__ void __launch_bounds__(PF_BLOCK) k_predict(const double* e, const float* st_in, TrialPar tp, ScenePar sp,
                                                      int64_t n, int64_t goff, u64 seed, int S, int t0, int n_steps,
                                                      float* pos_out, float* obs_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Body s;
  load_body(st_in, n, i, s);
  const float ef = (float)e[i];
  for (int k = 0; k < n_steps; k++) {
    uint32_t r[4], r2[4];
    philox4x32_10(seed, (uint32_t)(goff + i), (uint32_t)(t0 + k), 0u, (uint32_t)ST_PREDICT, r);
    philox4x32_10(seed, (uint32_t)(goff + i), (uint32_t)(t0 + k), 1u, (uint32_t)ST_PREDICT, r2);
    double z[4], zo[4];
    box_muller(r[0], r[1], z[0], z[1]);
    box_muller(r[2], r[3], z[2], z[3]);
    box_muller(r2[0], r2[1], zo[0], zo[1]);
    box_muller(r2[2], r2[3], zo[2], zo[3]);
    float x[3] = {(float)((double)s.x.x + tp.sigma_tr * z[0]), (float)((double)s.x.y + tp.sigma_tr * z[1]),
                  (float)((double)s.x.z + tp.sigma_tr * z[2])};
    for (int c = 0; c < 3; c++) {
      if (pos_out) pos_out[((int64_t)k * 3 + c) * n + i] = x[c];
      if (obs_out) obs_out[((int64_t)k * 3 + c) * n + i] = (float)((double)x[c] + tp.sigma_obs * zo[c]);
    }
    s.x = mk3(x[0], x[1], x[2]);
    macro_step<SHAPE, true>(sp, ef, S, s);
  }
}

// AoS [n][13] <-> SoA [13][n]
__global__ void k_aos_to_soa(const float* aos, float* soa, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int c = 0; c < 13; c++) soa[c * n + i] = aos[i * 13 + c];
}
__global__ void k_soa_to_aos(const float* soa, float* aos, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int c = 0; c < 13; c++) aos[i * 13 + c] = soa[c * n + i];
}
__global__ void k_f2d(const float* a, double* b, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = (double)a[i];
}
__global__ void k_block_max(const double* lw, int64_t n, Ctl* ctl) {
  __shared__ double smax[8];
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double v = (i < n) ? lw[i] : -INFINITY;
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) v = fmax(v, smax[w]);
    atomicMax(&ctl->max_enc, enc_max(v));
  }
}
__global__ void k_publish_max(Ctl* ctl, double* out) {
  double m = dec_max(ctl->max_enc);
  ctl->gmax = m;
  if (out) *out = m;
}
__global__ void k_idx_to_i64(const int* idx, int64_t* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = idx[i];
}

// ---------------------------------------------------------------- migration payload (sharded filter)
// word layout of one particle: [0,1] e  [2,3] qs  [4..16] state13  [17 + 3 s + c] x~_s, s = 0..t
__device__ __forceinline__ void pack_chunk(int64_t k, int64_t k_lo, int64_t k_hi, int64_t spr, int64_t words,
                                           int64_t& base, int64_t& cnt, int64_t& p) {
  int64_t r = k / spr;
  int64_t cfirst = max(k_lo, r * spr), clast = min(k_hi, (r + 1) * spr);
  cnt = clast - cfirst;
  base = (cfirst - k_lo) * words;
  p = k - cfirst;
}
// offspring off..off+c-1 of the range [k_lo, k_lo+count); idx[j] is the ancestor of offspring off + j
__global__ void k_pack(Bufs B, const Ctl* ctl, int64_t n, int64_t k_lo, int64_t count, int64_t off, int64_t c,
                       int64_t spr, int64_t words, float* send) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= c) return;
  const int cur = ctl->cur;
  int64_t base, cnt, p;
  pack_chunk(k_lo + off + j, k_lo, k_lo + count, spr, words, base, cnt, p);
  int64_t a = B.idx[j];
  float* o = send + base + p;
  if (blockIdx.y == 0) {
    double e = B.e[cur][a], q = B.qs[cur][a];
    o[0 * cnt] = __int_as_float(__double2loint(e)); o[1 * cnt] = __int_as_float(__double2hiint(e));
    o[2 * cnt] = __int_as_float(__double2loint(q)); o[3 * cnt] = __int_as_float(__double2hiint(q));
#pragma unroll
    for (int c = 0; c < 13; c++) o[(4 + c) * cnt] = B.st[cur][c * n + a];
  } else {
    int64_t s = blockIdx.y - 1;
#pragma unroll
    for (int c = 0; c < 3; c++) o[(17 + 3 * s + c) * cnt] = B.xt[cur][(s * 3 + c) * n + a];
  }
}
__global__ void k_unpack(Bufs B, const Ctl* ctl, int64_t n, const float* chunk, int64_t cnt, int64_t slot0) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= cnt) return;
  const int nxt = ctl->cur ^ 1;
  const float* o = chunk + p;
  int64_t k = slot0 + p;
  if (blockIdx.y == 0) {
    B.e[nxt][k] = __hiloint2double(__float_as_int(o[1 * cnt]), __float_as_int(o[0 * cnt]));
    B.qs[nxt][k] = __hiloint2double(__float_as_int(o[3 * cnt]), __float_as_int(o[2 * cnt]));
    B.logw[k] = 0.0;
#pragma unroll
    for (int c = 0; c < 13; c++) B.st[nxt][c * n + k] = o[(4 + c) * cnt];
  } else {
    int64_t s = blockIdx.y - 1;
#pragma unroll
    for (int c = 0; c < 3; c++) B.xt[nxt][(s * 3 + c) * n + k] = o[(17 + 3 * s + c) * cnt];
  }
}
__global__ void k_set_resampled(Ctl* ctl, int v) { ctl->resampled = v; }

// ---------------------------------------------------------------- sharded filter, decided and exchanged on the device
// #{k in [0,N) : tau_k < C}; tau is non-decreasing in k
__host__ __device__ static inline int64_t count_below(u64 C, u64 U0, u64 W, int64_t N) {
  int64_t lo = 0, hi = N;  // first k with tau_k >= C
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (sys_threshold(U0, (u64)mid, W, (u64)N) < C) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// One thread: from the all-gathered integer sums (W, S1, S2 of every rank) take the ESS decision of
// Gen.maybe_resample! (Inference/particle_filter.jl:55) - identical on every rank, integers - and derive this rank's
// CDF offset and the global slot range its offspring fill (systematic resampling preserves order).
__global__ void k_decide(Ctl* ctl, const u64* all, const Peer* peers, int G, int rank, int64_t N, int64_t ess_num,
                         int64_t ess_den, u64 U0, int t, double* st_ess, double* st_lse, int* st_res) {
  u64 W = 0, S1 = 0, S2 = 0, Wb = 0, Wl = 0;
  for (int r = 0; r < G; r++) {
    // NCCL all-gather result, or (one-process group) the other shards' sums through peer memory
    const volatile u64* q = all ? all + 3 * r : peers[r].sums;
    const u64 w = q[0];
    if (r < rank) Wb += w;
    if (r == rank) Wl = w;
    W += w; S1 += q[1]; S2 += q[2];
  }
  unsigned __int128 lhs = (unsigned __int128)S1 * S1 * (u64)ess_den;
  unsigned __int128 rhs = (unsigned __int128)S2 * (u64)N * (u64)ess_num;
  const int res = lhs < rhs;
  const double lse = ctl->gmax + log((double)W / 549755813888.0);
  ctl->resampled = res;
  if (res) ctl->log_ml += lse - log((double)N);
  ctl->W_before = Wb; ctl->W_total = W; ctl->U0 = U0;
  ctl->k_lo = res ? count_below(Wb, U0, W, N) : 0;
  ctl->k_hi = res ? count_below(Wb + Wl, U0, W, N) : 0;
  if (st_ess) {
    st_ess[t] = S2 ? ((double)S1 * (double)S1) / (double)S2 : 0.0;
    st_lse[t] = lse;
    st_res[t] = res;
  }
}

// one-process group: global max of the shards' local maxima through peer memory (stands in for the NCCL all-reduce)
__global__ void k_copy_lmax(Ctl* ctl) { ctl->lmax = ctl->gmax; }
__global__ void k_local_max(Ctl* ctl, const Peer* peers, int G) {
  double m = -INFINITY;
  for (int r = 0; r < G; r++) m = fmax(m, ((const volatile Ctl*)peers[r].ctl)->lmax);
  ctl->gmax = m;
}

// ancestors of the offspring k_lo + j, j < k_hi - k_lo, in this shard's CDF (which starts at W_before).  A shard's
// offspring can outnumber its particles (up to N): grid-stride over j, the grid is sized for the usual count ~ n.
__global__ void k_index_sh(Bufs B, const Ctl* ctl, int64_t n, int64_t Nglob) {
  if (!ctl->resampled) return;
  const int64_t k_lo = ctl->k_lo, count = ctl->k_hi - k_lo;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += (int64_t)gridDim.x * blockDim.x) {
    const u64 tau = sys_threshold(ctl->U0, (u64)(k_lo + j), ctl->W_total, (u64)Nglob);
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (B.cdf[mid] > tau) hi = mid; else lo = mid + 1;
    }
    B.idx[j] = (int)lo;
  }
}

// Offspring k_lo + j of this shard goes to global slot K = k_lo + j = rank d = K / n, local slot K - d n: stored
// straight into that rank's idle buffers through peer memory (NVLink stores; the own rank is just another entry of
// the table).  Consecutive offspring go to consecutive slots, so a warp writes 128 consecutive bytes per row.
__global__ void k_push(Bufs B, const Ctl* ctl, const Peer* __restrict__ peers, int64_t n) {
  if (!ctl->resampled) return;
  const int64_t k_lo = ctl->k_lo, count = ctl->k_hi - k_lo;
  const int cur = ctl->cur, nxt = cur ^ 1;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t K = k_lo + j, d = K / n, k = K - d * n, a = B.idx[j];
    const Peer P = peers[d];
    P.e[nxt][k] = B.e[cur][a];
    P.qs[nxt][k] = B.qs[cur][a];
#pragma unroll
    for (int c = 0; c < 13; c++) P.st[nxt][c * n + k] = B.st[cur][c * n + a];
  }
}
__global__ void __launch_bounds__(256) k_push_hist(Bufs B, const Ctl* ctl, const Peer* __restrict__ peers, int64_t n,
                                                   int rows) {
  if (!ctl->resampled) return;
  const int64_t k_lo = ctl->k_lo, count = ctl->k_hi - k_lo;
  const int cur = ctl->cur, nxt = cur ^ 1;
  const int c0 = (int)blockIdx.y * (GH_ROWS * 3), c1 = rows * 3;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t K = k_lo + j, d = K / n, k = K - d * n, a = B.idx[j];
    const float* __restrict__ src = B.xt[cur] + (int64_t)c0 * n + a;
    float* __restrict__ dst = peers[d].xt[nxt] + (int64_t)c0 * n + k;
    float v[GH_ROWS * 3];
#pragma unroll
    for (int c = 0; c < GH_ROWS * 3; c++)
      if (c0 + c < c1) v[c] = __ldg(src + (int64_t)c * n);
#pragma unroll
    for (int c = 0; c < GH_ROWS * 3; c++)
      if (c0 + c < c1) dst[(int64_t)c * n] = v[c];
  }
}
// ---- multinomial parents on a sharded filter.  The parents are i.i.d., not sorted, so there is no offspring range to
// push: every rank draws the thresholds of ITS OWN slots (Philox counter = global slot, as on one shard), finds the
// rank whose integer weight interval holds the threshold, searches that rank's CDF through peer memory and PULLS the
// ancestor's payload into its idle buffers.  CDFs hold global values (own prefix sum + W_before), so the ancestors are
// the single-shard ones, bit for bit.  `all` = NCCL all-gather of the sums, or NULL: read the peers' sums directly.
__global__ void k_index_multi_sh(Bufs B, const Ctl* ctl, const u64* all, const Peer* __restrict__ peers, int G,
                                 int64_t n, int64_t goff, u64 seed, int t) {
  if (!ctl->resampled) return;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  uint32_t r[4];
  philox4x32_10(seed, (uint32_t)(goff + j), (uint32_t)t, 0u, (uint32_t)ST_RESAMPLE_MULTI, r);
  const u64 u = ((u64)r[0] << 32) | (u64)r[1];
  const u64 tau = multi_threshold(u, ctl->W_total);
  int owner = G - 1;
  u64 before = 0;
  for (int q = 0; q < G; q++) {
    const u64 w = all ? all[3 * q] : ((const volatile u64*)peers[q].sums)[0];
    if (tau < before + w) { owner = q; break; }
    before += w;
  }
  const u64* __restrict__ cdf = peers[owner].cdf;
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (cdf[mid] > tau) hi = mid; else lo = mid + 1;
  }
  B.idx[j] = (int)lo;
  B.idx[n + j] = owner;  // the ancestor scratch of a sharded filter holds N >= 2 n entries
}
__global__ void k_pull(Bufs B, const Ctl* ctl, const Peer* __restrict__ peers, int64_t n) {
  if (!ctl->resampled) return;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int cur = ctl->cur, nxt = cur ^ 1;  // every rank holds the same `cur`: identical decisions
  const int64_t a = B.idx[j];
  const Peer P = peers[B.idx[n + j]];
  B.e[nxt][j] = P.e[cur][a];
  B.qs[nxt][j] = P.qs[cur][a];
#pragma unroll
  for (int c = 0; c < 13; c++) B.st[nxt][c * n + j] = P.st[cur][c * n + a];
}
__global__ void __launch_bounds__(256) k_pull_hist(Bufs B, const Ctl* ctl, const Peer* __restrict__ peers, int64_t n,
                                                   int rows) {
  if (!ctl->resampled) return;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int cur = ctl->cur, nxt = cur ^ 1;
  const int64_t a = B.idx[j];
  const int c0 = (int)blockIdx.y * (GH_ROWS * 3), c1 = rows * 3;
  const float* __restrict__ src = peers[B.idx[n + j]].xt[cur] + (int64_t)c0 * n + a;
  float* __restrict__ dst = B.xt[nxt] + (int64_t)c0 * n + j;
  float v[GH_ROWS * 3];
#pragma unroll
  for (int c = 0; c < GH_ROWS * 3; c++)
    if (c0 + c < c1) v[c] = src[(int64_t)c * n];
#pragma unroll
  for (int c = 0; c < GH_ROWS * 3; c++)
    if (c0 + c < c1) __stcs(dst + (int64_t)c * n, v[c]);
}

// after the exchange: log weights := 0 (Gen.maybe_resample!).  Runs after the barrier that follows the pushes; the
// log weights are single buffered and a rank's CDF is built from them, so nobody may reset them earlier.
__global__ void k_post_resample(Bufs B, const Ctl* ctl, int64_t n) {
  if (!ctl->resampled) return;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) B.logw[i] = 0.0;
}

// Multinomial parents (what Gen.maybe_resample! draws in Gen 0.4.7: rand(Categorical(w)) per slot): slot k takes
// tau_k = floor(u_k W / 2^64) with u_k 64 Philox bits of counter (global slot, step), against the same integer CDF.
__global__ void k_index_multi(Bufs B, const Ctl* ctl, int64_t n, const u64* W_ptr, u64 seed, int t, int64_t goff,
                              int64_t count, int gate) {
  if (gate && !ctl->resampled) return;
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= count) return;
  uint32_t r[4];
  philox4x32_10(seed, (uint32_t)(goff + j), (uint32_t)t, 0u, (uint32_t)ST_RESAMPLE_MULTI, r);
  const u64 u = ((u64)r[0] << 32) | (u64)r[1];
  const u64 tau = multi_threshold(u, *W_ptr);
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (B.cdf[mid] > tau) hi = mid; else lo = mid + 1;
  }
  B.idx[j] = (int)lo;
}

// FP32 FMA peak probe: the roofline denominator of the simulate kernels (not in MEASURED_PEAKS.json)
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x[8];
#pragma unroll
  for (int k = 0; k < 8; k++) x[k] = (float)(threadIdx.x + k);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 8; k++) x[k] = fmaf(x[k], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; k++) s += x[k];
  if (s == 12345.678f) out[0] = s;
}

// ---------------------------------------------------------------- handle
struct smcphys_handle {
  smcphys_config cfg;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  Bufs B;
  Ctl* ctl = nullptr;
  double* d_obs = nullptr;
  double *d_ess = nullptr, *d_lse = nullptr;
  int* d_res = nullptr;
  u64* d_acc = nullptr;
  TrialPar tp;
  ScenePar sp;
  std::vector<double> obs;
  int T = 0;
  bool trial_set = false;
  int steps_done = -1;  // PF steps applied to the particles; -1 = not initialised
  int ntiles = 0;
  int64_t local_gathered = 0;  // slots of the current resampling step already filled by pf_gather_local
  std::vector<void*> allocs;
  // sharded filter inside the library: NCCL communicator (one process per GPU) or a one-process group of shards
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  Peer* d_peers = nullptr;          // device table, nranks entries
  u64 *d_sums3 = nullptr, *d_all = nullptr;
  int* d_bar = nullptr;
  std::vector<void*> ipc_opened;    // peer buffers opened with cudaIpcOpenMemHandle
  std::vector<smcphys_handle*> kids;  // group handle: the shards it owns (smcphys_create_multi)
  smcphys_handle* parent = nullptr;
  cudaEvent_t ev_sync = nullptr;    // one-process group: this shard's progress marker
  bool prof = false;
  struct Ev { int cls; cudaEvent_t a, b; };
  std::vector<Ev> evs;
  std::vector<cudaEvent_t> ev_pool;
};

// event pair around one launch group of a kernel class (only when profiling is on)
struct ProfScope {
  smcphys_handle* h;
  cudaEvent_t b = nullptr;
  ProfScope(smcphys_handle* h_, int cls) : h(h_) {
    if (!h->prof) return;
    cudaEvent_t a = nullptr;
    auto take = [&](cudaEvent_t& e) {
      if (!h->ev_pool.empty()) { e = h->ev_pool.back(); h->ev_pool.pop_back(); }
      else cudaEventCreate(&e);
    };
    take(a); take(b);
    cudaEventRecord(a, h->stream);
    h->evs.push_back({cls, a, b});
  }
  ~ProfScope() { if (b) cudaEventRecord(b, h->stream); }
};

static inline unsigned int nblk(int64_t n, int b) { return (unsigned int)((n + b - 1) / b); }

// Device scratch of the getters and the stateless entries.  Pieces are carved from ONE grow-only pool per device, so a
// host loop over PhySMC.forward_step (smcphys_sim_step per step) or over the getters does not pay cudaMalloc / cudaFree
// per call: the first call that needs more than the pool holds takes separate blocks (freed on scope exit) and leaves
// the pool grown for the next one.  A scope holds the device's pool mutex from its first get() to its end - every
// user synchronises its stream before it returns, so the pool is idle when the next scope takes it.  Nothing leaks on
// the early returns of CK / LAUNCH.
struct DevPool {
  std::mutex mu;
  unsigned char* base = nullptr;
  size_t cap = 0;
};
static DevPool g_pool[SMCPHYS_MAX_RANKS];
struct DevScratch {
  DevPool* pool = nullptr;
  std::unique_lock<std::mutex> lk;
  size_t used = 0, need = 0;
  std::vector<void*> extra;
  ~DevScratch() {
    for (void* q : extra) cudaFree(q);
    if (pool && need > pool->cap) {  // grow for the next call (with headroom); the pool is idle here
      if (pool->base) cudaFree(pool->base);
      pool->base = nullptr;
      pool->cap = 0;
      void* q = nullptr;
      const size_t want = need + need / 4 + 4096;
      if (cudaMalloc(&q, want) == cudaSuccess) { pool->base = (unsigned char*)q; pool->cap = want; }
      else cudaGetLastError();
    }
  }
  template <class T> cudaError_t get(T** out, size_t count) {
    if (!pool) {
      int dev = 0;
      cudaError_t e = cudaGetDevice(&dev);
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < SMCPHYS_MAX_RANKS) {
        pool = &g_pool[dev];
        lk = std::unique_lock<std::mutex>(pool->mu);
      }
    }
    const size_t bytes = ((count * sizeof(T) + 16) + 255) & ~(size_t)255;
    need += bytes;
    if (pool && used + bytes <= pool->cap) {
      *out = (T*)(pool->base + used);
      used += bytes;
      return cudaSuccess;
    }
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e == cudaSuccess) extra.push_back(q);
    *out = (T*)q;
    return e;
  }
};

template <class T>
static int dalloc(smcphys_handle* h, T** p, size_t count) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, count * sizeof(T) + 16);
  if (e != cudaSuccess) return fail(SMCPHYS_ENOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e));
  h->allocs.push_back(q);
  *p = (T*)q;
  return 0;
}

extern "C" {

const char* smcphys_last_error(void) { return g_err.c_str(); }
int64_t smcphys_launch_count(void) { return g_launches.load(); }
int smcphys_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int smcphys_config_default(smcphys_config* c) {
  if (!c) return fail(SMCPHYS_EINVAL, "cfg is NULL");
  memset(c, 0, sizeof(*c));
  c->struct_size = (int32_t)sizeof(*c);
  c->device = 0;
  c->shape = SMCPHYS_SPHERE;
  c->substeps = 5;
  c->num_particles = 20;      // args.jl:19
  c->global_particles = 20;
  c->particle_offset = 0;
  c->max_timesteps = 120;
  c->rejuvenation_moves = 1;  // args.jl:20
  c->observation_noise = (double)0.05f;   // args.jl:14
  c->transition_noise = (double)0.075f;   // args.jl:15
  c->proposal_sigma = 0.1;                // particle_filter.jl:23
  c->ceiling_half = 0.01;                 // bouncing_object.jl:53
  c->ess_num = 1; c->ess_den = 2;         // particle_filter.jl:55
  c->seed = 1234;
  c->resampler = SMCPHYS_RESAMPLE_SYSTEMATIC;  // north_star; SMCPHYS_RESAMPLE_MULTINOMIAL = Gen.maybe_resample!
  return 0;
}

int smcphys_create(const smcphys_config* cfg, smcphys_handle** out) {
  if (!cfg || !out) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (cfg->struct_size != (int32_t)sizeof(smcphys_config)) return fail(SMCPHYS_EINVAL, "smcphys_config size mismatch");
  if (cfg->shape != SMCPHYS_SPHERE && cfg->shape != SMCPHYS_CUBE) return fail(SMCPHYS_EINVAL, "shape must be 0 or 1");
  if (cfg->num_particles < 1 || cfg->num_particles > (1ll << 24) || cfg->global_particles > (1ll << 24))
    return fail(SMCPHYS_EINVAL, "num_particles must be in [1, 2^24]");
  if (cfg->global_particles < cfg->num_particles || cfg->particle_offset < 0 ||
      cfg->particle_offset + cfg->num_particles > cfg->global_particles)
    return fail(SMCPHYS_EINVAL, "shard [offset, offset+num_particles) must lie inside global_particles");
  if (cfg->substeps < 1 || cfg->substeps > 64) return fail(SMCPHYS_EINVAL, "substeps must be in [1, 64]");
  if (cfg->max_timesteps < 1 || cfg->max_timesteps > 4096) return fail(SMCPHYS_EINVAL, "max_timesteps must be in [1, 4096]");
  if (!(cfg->observation_noise > 0) || !(cfg->transition_noise > 0) || !(cfg->proposal_sigma > 0))
    return fail(SMCPHYS_EINVAL, "noise parameters must be positive");
  if (cfg->ess_num < 0 || cfg->ess_den < 1) return fail(SMCPHYS_EINVAL, "bad ESS fraction");
  if (cfg->resampler != SMCPHYS_RESAMPLE_SYSTEMATIC && cfg->resampler != SMCPHYS_RESAMPLE_MULTINOMIAL)
    return fail(SMCPHYS_EINVAL, "resampler must be 0 (systematic) or 1 (multinomial)");
  int ndev = smcphys_device_count();
  if (ndev <= 0) return fail(SMCPHYS_ECUDA, "no CUDA device (libsmcphys has no CPU path)");
  if (cfg->device < 0 || cfg->device >= ndev) return fail(SMCPHYS_EINVAL, "device ordinal out of range");
  CK(cudaSetDevice(cfg->device));
  smcphys_handle* h = new smcphys_handle();
  h->cfg = *cfg;
  h->sp.ceil_half = (float)cfg->ceiling_half;
  const int64_t n = cfg->num_particles;
  const size_t Tm = (size_t)cfg->max_timesteps;
  h->ntiles = (int)((n + TILE - 1) / TILE);
  int rc = 0;
  for (int b = 0; b < 2 && !rc; b++) {
    rc = rc ? rc : dalloc(h, &h->B.xt[b], Tm * 3 * n);
    rc = rc ? rc : dalloc(h, &h->B.st[b], (size_t)13 * n);
    rc = rc ? rc : dalloc(h, &h->B.e[b], (size_t)n);
    rc = rc ? rc : dalloc(h, &h->B.qs[b], (size_t)n);
  }
  rc = rc ? rc : dalloc(h, &h->B.logw, (size_t)n);
  rc = rc ? rc : dalloc(h, &h->B.cdf, (size_t)n);
  rc = rc ? rc : dalloc(h, &h->B.tile_sum, (size_t)h->ntiles);
  rc = rc ? rc : dalloc(h, &h->B.tile_off, (size_t)h->ntiles);
  const bool sharded = cfg->global_particles > cfg->num_particles;
  // a shard's offspring can outnumber its particles (up to N): the ancestor scratch holds N entries when sharded
  rc = rc ? rc : dalloc(h, &h->B.idx, (size_t)(sharded ? cfg->global_particles : n));
  rc = rc ? rc : dalloc(h, &h->ctl, 1);
  rc = rc ? rc : dalloc(h, &h->d_sums3, (size_t)4);
  if (sharded) {
    rc = rc ? rc : dalloc(h, &h->d_peers, (size_t)SMCPHYS_MAX_RANKS);
    rc = rc ? rc : dalloc(h, &h->d_all, (size_t)3 * SMCPHYS_MAX_RANKS);
    rc = rc ? rc : dalloc(h, &h->d_bar, (size_t)4);
  }
  rc = rc ? rc : dalloc(h, &h->d_obs, Tm * 3);
  rc = rc ? rc : dalloc(h, &h->d_ess, Tm);
  rc = rc ? rc : dalloc(h, &h->d_lse, Tm);
  rc = rc ? rc : dalloc(h, &h->d_res, Tm);
  rc = rc ? rc : dalloc(h, &h->d_acc, Tm);
  if (rc) { smcphys_destroy(h); return rc; }
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
    smcphys_destroy(h);
    return fail(SMCPHYS_ECUDA, "cudaStreamCreate failed");
  }
  h->own_stream = true;
  cudaMemsetAsync(h->ctl, 0, sizeof(Ctl), h->stream);
  if (sharded) cudaMemsetAsync(h->d_bar, 0, sizeof(int) * 4, h->stream);
  *out = h;
  return 0;
}

int smcphys_destroy(smcphys_handle* h) {
  if (!h) return 0;
  if (!h->kids.empty()) {
    for (auto* k : h->kids) smcphys_synchronize(k);
    for (auto* k : h->kids) { k->parent = nullptr; smcphys_destroy(k); }
    delete h;
    return 0;
  }
  cudaSetDevice(h->cfg.device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  if (h->comm && nccl().lib) nccl().CommDestroy(h->comm);
  for (void* p : h->ipc_opened) cudaIpcCloseMemHandle(p);
  if (h->ev_sync) cudaEventDestroy(h->ev_sync);
  if (h->stream) cudaStreamSynchronize(h->stream);
  for (void* p : h->allocs) cudaFree(p);
  for (auto& e : h->evs) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
  for (auto e : h->ev_pool) cudaEventDestroy(e);
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

int smcphys_set_stream(smcphys_handle* h, void* cuda_stream) {
  if (!h) return fail(SMCPHYS_EINVAL, "NULL handle");
  if (!h->kids.empty()) return fail(SMCPHYS_ESTATE, "a multi-device handle runs every shard on its own stream");
  CK(cudaSetDevice(h->cfg.device));
  if (h->stream) CK(cudaStreamSynchronize(h->stream));
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  h->own_stream = false;
  h->stream = (cudaStream_t)cuda_stream;
  return 0;
}

int smcphys_set_profiling(smcphys_handle* h, int32_t on) {
  if (!h) return fail(SMCPHYS_EINVAL, "NULL handle");
  if (!h->kids.empty()) h = h->kids[0];
  h->prof = on != 0;
  return 0;
}

int smcphys_kernel_time(smcphys_handle* h, int32_t kclass, double* ms_total, int64_t* groups) {
  if (!h || !ms_total) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (kclass < 0 || kclass >= SMCPHYS_K_CLASSES) return fail(SMCPHYS_EINVAL, "unknown kernel class");
  if (!h->kids.empty()) h = h->kids[0];
  CK(cudaSetDevice(h->cfg.device));
  CK(cudaStreamSynchronize(h->stream));
  double tot = 0;
  int64_t cnt = 0;
  std::vector<smcphys_handle::Ev> rest;
  for (auto& e : h->evs) {
    if (e.cls != kclass) { rest.push_back(e); continue; }
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e.a, e.b));
    tot += ms;
    cnt++;
    h->ev_pool.push_back(e.a);
    h->ev_pool.push_back(e.b);
  }
  h->evs.swap(rest);
  *ms_total = tot;
  if (groups) *groups = cnt;
  return 0;
}

int smcphys_synchronize(smcphys_handle* h) {
  if (!h) return fail(SMCPHYS_EINVAL, "NULL handle");
  if (!h->kids.empty()) {
    for (auto* k : h->kids) {
      int rc = smcphys_synchronize(k);
      if (rc) return rc;
    }
    return 0;
  }
  CK(cudaSetDevice(h->cfg.device));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int smcphys_set_trial(smcphys_handle* h, const double init_pos[3], const double init_vel[3], const double* obs,
                      int32_t T) {
  if (!h || !init_pos || !init_vel || (!obs && T > 0)) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (T < 0 || T > h->cfg.max_timesteps) return fail(SMCPHYS_EINVAL, "T exceeds max_timesteps");
  if (!h->kids.empty()) {
    for (auto* k : h->kids) {
      int rc = smcphys_set_trial(k, init_pos, init_vel, obs, T);
      if (rc) return rc;
    }
    h->T = T;
    h->trial_set = true;
    return 0;
  }
  CK(cudaSetDevice(h->cfg.device));
  CK(cudaStreamSynchronize(h->stream));  // the previous trial's copy may still read h->obs
  for (int c = 0; c < 3; c++) { h->tp.x0[c] = (float)init_pos[c]; h->tp.v0[c] = (float)init_vel[c]; }
  h->tp.sigma_obs = h->cfg.observation_noise;
  h->tp.sigma_tr = h->cfg.transition_noise;
  h->obs.assign(obs, obs + (size_t)T * 3);
  h->T = T;
  h->steps_done = -1;
  h->trial_set = true;
  if (T > 0) {
    CK(cudaMemcpyAsync(h->d_obs, h->obs.data(), sizeof(double) * 3 * T, cudaMemcpyHostToDevice, h->stream));
  }
  return 0;
}

// ---------------------------------------------------------------- filter phases
#define NEED_TRIAL(h)                                                                     \
  if (!(h)) return fail(SMCPHYS_EINVAL, "NULL handle");                                   \
  if (!(h)->trial_set) return fail(SMCPHYS_ESTATE, "no trial: call smcphys_set_trial first"); \
  CK(cudaSetDevice((h)->cfg.device));

int smcphys_pf_init(smcphys_handle* h) {
  NEED_TRIAL(h);
  const int64_t n = h->cfg.num_particles;
  LAUNCH(k_init, nblk(n, 256), 256, h->stream, h->B, h->ctl, h->tp, n, h->cfg.particle_offset, h->cfg.seed);
  CK(cudaMemsetAsync(h->d_acc, 0, sizeof(u64) * h->cfg.max_timesteps, h->stream));
  h->steps_done = 0;
  return 0;
}

int smcphys_pf_extend(smcphys_handle* h, int32_t t, double* d_local_max) {
  NEED_TRIAL(h);
  if (h->steps_done < 0) return fail(SMCPHYS_ESTATE, "pf_extend before pf_init / set_trial");
  if (t != h->steps_done || t >= h->T) return fail(SMCPHYS_ESTATE, "pf_extend: steps must run in order 0..T-1");
  const int64_t n = h->cfg.num_particles;
  const double* y = &h->obs[(size_t)t * 3];
  LAUNCH(k_step_begin, 1, 1, h->stream, h->ctl);
  ProfScope ps(h, SMCPHYS_K_EXTEND);
  if (h->cfg.shape == SMCPHYS_SPHERE)
    LAUNCH(k_extend<SHAPE_SPHERE>, nblk(n, PF_BLOCK), PF_BLOCK, h->stream, h->B, h->ctl, h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, t, h->cfg.substeps, y[0], y[1], y[2]);
  else
    LAUNCH(k_extend<SHAPE_CUBE>, nblk(n, PF_BLOCK), PF_BLOCK, h->stream, h->B, h->ctl, h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, t, h->cfg.substeps, y[0], y[1], y[2]);
  if (d_local_max)
    CK(cudaMemcpyAsync(d_local_max, &h->ctl->gmax, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  h->steps_done = t + 1;
  return 0;
}

static int weight_sums(smcphys_handle* h, const double* d_gmax, int decide, int t, u64* d_sums) {
  const int64_t n = h->cfg.num_particles;
  ProfScope ps(h, SMCPHYS_K_WEIGHTS);
  LAUNCH(k_weight_sums, h->ntiles, TILE_THREADS, h->stream, h->B, h->ctl, d_gmax, n, h->ntiles, decide,
         h->cfg.ess_num, h->cfg.ess_den, t, h->d_ess, h->d_lse, h->d_res, d_sums);
  return 0;
}

int smcphys_pf_weight_sums(smcphys_handle* h, const double* d_global_max, uint64_t* d_sums3) {
  NEED_TRIAL(h);
  if (h->steps_done < 1) return fail(SMCPHYS_ESTATE, "pf_weight_sums before pf_extend");
  if (!d_global_max || !d_sums3) return fail(SMCPHYS_EINVAL, "NULL device pointer");
  return weight_sums(h, d_global_max, 0, h->steps_done - 1, (u64*)d_sums3);
}

int smcphys_pf_resample_local(smcphys_handle* h, int32_t t) {
  NEED_TRIAL(h);
  if (t != h->steps_done - 1) return fail(SMCPHYS_ESTATE, "pf_resample_local: t is not the current step");
  if (h->cfg.global_particles != h->cfg.num_particles)
    return fail(SMCPHYS_ESTATE, "pf_resample_local needs an unsharded filter; use pf_pack / pf_unpack");
  const int64_t n = h->cfg.num_particles;
  int rc = weight_sums(h, &h->ctl->gmax, 1, t, nullptr);
  if (rc) return rc;
  uint32_t r[4];
  philox4x32_10(h->cfg.seed, 0u, (uint32_t)t, 0u, (uint32_t)ST_RESAMPLE, r);
  {
    ProfScope ps(h, SMCPHYS_K_WEIGHTS);
    LAUNCH(k_cdf, h->ntiles, TILE_THREADS, h->stream, h->B, h->ctl, &h->ctl->gmax, n, (u64)0, 1, (const u64*)nullptr);
    if (h->cfg.resampler == SMCPHYS_RESAMPLE_MULTINOMIAL)
      LAUNCH(k_index_multi, nblk(n, 256), 256, h->stream, h->B, h->ctl, n, &h->ctl->W, h->cfg.seed, t, (int64_t)0, n, 1);
    else
      LAUNCH(k_index, nblk(n, 256), 256, h->stream, h->B, h->ctl, n, n, (u64)r[0], &h->ctl->W, (u64)0, (int64_t)0, n, 1);
  }
  ProfScope ps(h, SMCPHYS_K_GATHER);
  LAUNCH(k_gather, nblk(n, 256), 256, h->stream, h->B, h->ctl, n, n, (int64_t)0, 1);
  LAUNCH(k_gather_hist, dim3(nblk(n, 256), (unsigned)((t + GH_ROWS) / GH_ROWS)), 256, h->stream, h->B, h->ctl, n, t + 1,
         n, (int64_t)0, 1);
  return 0;
}

int smcphys_pf_rejuvenate(smcphys_handle* h, int32_t t, int32_t R, uint64_t* d_accept_count) {
  return smcphys_pf_rejuvenate_from(h, t, R, 0, d_accept_count);
}
int smcphys_pf_rejuvenate_from(smcphys_handle* h, int32_t t, int32_t R, int32_t move0, uint64_t* d_accept_count) {
  NEED_TRIAL(h);
  if (t != h->steps_done - 1) return fail(SMCPHYS_ESTATE, "pf_rejuvenate: t is not the current step");
  if (R < 0 || move0 < 0) return fail(SMCPHYS_EINVAL, "R < 0 or move0 < 0");
  if (R == 0) return 0;
  const int64_t n = h->cfg.num_particles;
  u64* acc = d_accept_count ? (u64*)d_accept_count : h->d_acc + t;
  ProfScope ps(h, SMCPHYS_K_REJUVENATE);
  if (h->cfg.shape == SMCPHYS_SPHERE)
    LAUNCH(k_rejuvenate<SHAPE_SPHERE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B, h->ctl, h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, t, h->cfg.substeps, R, move0, h->cfg.proposal_sigma, acc);
  else
    LAUNCH(k_rejuvenate<SHAPE_CUBE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B, h->ctl, h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, t, h->cfg.substeps, R, move0, h->cfg.proposal_sigma, acc);
  return 0;
}

// ---------------------------------------------------------------- sharded run (NCCL ranks or a one-process group)
#define NCCLCK(expr)                                                                                         \
  do {                                                                                                       \
    ncclResult_t _r = (expr);                                                                                \
    if (_r != ncclSuccess) return fail(SMCPHYS_ECUDA, std::string(#expr) + ": " + nccl().GetErrorString(_r)); \
  } while (0)

// one-process group: every shard's stream waits until all shards have reached this point
static int group_sync(std::vector<smcphys_handle*>& S) {
  if (S.size() < 2) return 0;
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    CK(cudaEventRecord(s->ev_sync, s->stream));
  }
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    for (auto* o : S)
      if (o != s) CK(cudaStreamWaitEvent(s->stream, o->ev_sync, 0));
  }
  return 0;
}

// One PF step of a sharded filter: S = the shards this process drives (one with an NCCL communicator, or all
// shards of a one-process group).  Nothing here waits on the host.
static int sharded_step(std::vector<smcphys_handle*>& S, int t, int R) {
  const bool use_nccl = S[0]->comm != nullptr;
  int rc = 0;
  for (auto* s : S) {
    rc = smcphys_pf_extend(s, t, nullptr);  // local max -> ctl->gmax
    if (rc) return rc;
    if (!use_nccl) LAUNCH(k_copy_lmax, 1, 1, s->stream, s->ctl);
  }
  // global max of the log weights
  if (use_nccl) {
    for (auto* s : S) NCCLCK(nccl().AllReduce(&s->ctl->gmax, &s->ctl->gmax, 1, ncclDouble, ncclMax, s->comm, s->stream));
  } else {
    rc = group_sync(S);
    if (rc) return rc;
    for (auto* s : S) {
      CK(cudaSetDevice(s->cfg.device));
      LAUNCH(k_local_max, 1, 1, s->stream, s->ctl, s->d_peers, s->nranks);
    }
  }
  // integer weight sums of every shard, visible to every shard
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    const int64_t n = s->cfg.num_particles;
    ProfScope ps(s, SMCPHYS_K_WEIGHTS);
    LAUNCH(k_weight_sums, s->ntiles, TILE_THREADS, s->stream, s->B, s->ctl, &s->ctl->gmax, n, s->ntiles, 0,
           s->cfg.ess_num, s->cfg.ess_den, t, s->d_ess, s->d_lse, s->d_res, s->d_sums3);
  }
  if (use_nccl) {
    for (auto* s : S) NCCLCK(nccl().AllGather(s->d_sums3, s->d_all, 3, ncclUint64, s->comm, s->stream));
  } else {
    rc = group_sync(S);
    if (rc) return rc;
  }
  uint32_t r[4];
  philox4x32_10(S[0]->cfg.seed, 0u, (uint32_t)t, 0u, (uint32_t)ST_RESAMPLE, r);
  const bool multinomial = S[0]->cfg.resampler == SMCPHYS_RESAMPLE_MULTINOMIAL;
  // decision and the shard's CDF (global values: own prefix sums + the weight of all lower ranks)
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    const int64_t n = s->cfg.num_particles, N = s->cfg.global_particles;
    ProfScope ps(s, SMCPHYS_K_WEIGHTS);
    LAUNCH(k_decide, 1, 1, s->stream, s->ctl, use_nccl ? (const u64*)s->d_all : (const u64*)nullptr, s->d_peers,
           s->nranks, s->rank, N, s->cfg.ess_num, s->cfg.ess_den, (u64)r[0], t, s->d_ess, s->d_lse, s->d_res);
    LAUNCH(k_cdf, s->ntiles, TILE_THREADS, s->stream, s->B, s->ctl, &s->ctl->gmax, n, (u64)0, 1,
           (const u64*)&s->ctl->W_before);
  }
  if (!multinomial) {
    // systematic: ancestors of this shard's offspring range, stored into their destination shard (push)
    for (auto* s : S) {
      CK(cudaSetDevice(s->cfg.device));
      const int64_t n = s->cfg.num_particles, N = s->cfg.global_particles;
      {
        ProfScope ps(s, SMCPHYS_K_WEIGHTS);
        LAUNCH(k_index_sh, nblk(n, 256), 256, s->stream, s->B, s->ctl, n, N);
      }
      ProfScope ps(s, SMCPHYS_K_GATHER);
      LAUNCH(k_push, nblk(n, 256), 256, s->stream, s->B, s->ctl, s->d_peers, n);
      LAUNCH(k_push_hist, dim3(nblk(n, 256), (unsigned)((t + GH_ROWS) / GH_ROWS)), 256, s->stream, s->B, s->ctl,
             s->d_peers, n, t + 1);
    }
  } else {
    // multinomial: every shard's CDF must be complete before anyone searches it ...
    if (use_nccl) {
      for (auto* s : S) NCCLCK(nccl().AllReduce(s->d_bar, s->d_bar, 1, ncclInt, ncclSum, s->comm, s->stream));
    } else {
      rc = group_sync(S);
      if (rc) return rc;
    }
    // ... then each shard draws the parents of its own slots and pulls them through the peer table
    for (auto* s : S) {
      CK(cudaSetDevice(s->cfg.device));
      const int64_t n = s->cfg.num_particles;
      {
        ProfScope ps(s, SMCPHYS_K_WEIGHTS);
        LAUNCH(k_index_multi_sh, nblk(n, 256), 256, s->stream, s->B, s->ctl,
               use_nccl ? (const u64*)s->d_all : (const u64*)nullptr, s->d_peers, s->nranks, n, s->cfg.particle_offset,
               s->cfg.seed, t);
      }
      ProfScope ps(s, SMCPHYS_K_GATHER);
      LAUNCH(k_pull, nblk(n, 256), 256, s->stream, s->B, s->ctl, s->d_peers, n);
      LAUNCH(k_pull_hist, dim3(nblk(n, 256), (unsigned)((t + GH_ROWS) / GH_ROWS)), 256, s->stream, s->B, s->ctl,
             s->d_peers, n, t + 1);
    }
  }
  // every shard's stores / loads of this step have completed before anyone reads its new buffer or reuses the old one
  if (use_nccl) {
    for (auto* s : S) NCCLCK(nccl().AllReduce(s->d_bar, s->d_bar, 1, ncclInt, ncclSum, s->comm, s->stream));
  } else {
    rc = group_sync(S);
    if (rc) return rc;
  }
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    const int64_t n = s->cfg.num_particles;
    {
      ProfScope ps(s, SMCPHYS_K_GATHER);
      LAUNCH(k_post_resample, nblk(n, 256), 256, s->stream, s->B, s->ctl, n);
    }
    rc = smcphys_pf_rejuvenate(s, t, R, nullptr);
    if (rc) return rc;
  }
  return 0;
}

static int copy_stats(smcphys_handle* h, int32_t R, smcphys_step_stats* per_step) {
  const int T = h->T;
  std::vector<double> ess(T), lse(T);
  std::vector<int> res(T);
  std::vector<u64> acc(T);
  CK(cudaSetDevice(h->cfg.device));
  CK(cudaMemcpyAsync(ess.data(), h->d_ess, sizeof(double) * T, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(lse.data(), h->d_lse, sizeof(double) * T, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(res.data(), h->d_res, sizeof(int) * T, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(acc.data(), h->d_acc, sizeof(u64) * T, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  for (int t = 0; t < T; t++) {
    per_step[t].ess = ess[t];
    per_step[t].lse = lse[t];
    per_step[t].resampled = res[t];
    per_step[t].pad_ = 0;
    per_step[t].accept_rate = R > 0 ? (double)acc[t] / ((double)h->cfg.num_particles * R) : 0.0;
  }
  return 0;
}

int smcphys_run_smc(smcphys_handle* h, int32_t R, smcphys_step_stats* per_step) {
  if (!h) return fail(SMCPHYS_EINVAL, "NULL handle");
  if (h->kids.size() == 1) return smcphys_run_smc(h->kids[0], R, per_step);  // a group of one shard is an ordinary filter
  if (!h->kids.empty() || h->cfg.global_particles != h->cfg.num_particles) {
    std::vector<smcphys_handle*> S;
    if (!h->kids.empty()) S = h->kids;
    else S.push_back(h);
    for (auto* s : S) {
      if (!s->trial_set) return fail(SMCPHYS_ESTATE, "no trial: call smcphys_set_trial first");
      if (!s->comm && (!s->parent || s->nranks != (int)S.size()))
        return fail(SMCPHYS_ESTATE, "sharded handle without a communicator: call smcphys_comm_init_rank (one process "
                                    "per GPU) or use smcphys_create_multi, or drive the smcphys_pf_* phases yourself");
    }
    if (R < 0) R = S[0]->cfg.rejuvenation_moves;
    int rc = 0;
    for (auto* s : S) {
      rc = smcphys_pf_init(s);
      if (rc) return rc;
    }
    const int T = S[0]->T;
    for (int t = 0; t < T && !rc; t++) rc = sharded_step(S, t, R);
    if (rc) return rc;
    if (per_step && T > 0) return copy_stats(S[0], R, per_step);
    return 0;
  }
  NEED_TRIAL(h);
  if (R < 0) R = h->cfg.rejuvenation_moves;
  int rc = smcphys_pf_init(h);
  for (int t = 0; t < h->T && !rc; t++) {
    rc = smcphys_pf_extend(h, t, nullptr);
    rc = rc ? rc : smcphys_pf_resample_local(h, t);
    rc = rc ? rc : smcphys_pf_rejuvenate(h, t, R, nullptr);
  }
  if (rc) return rc;
  if (per_step && h->T > 0) return copy_stats(h, R, per_step);
  return 0;
}

// ---------------------------------------------------------------- communicator / group set-up
static void fill_own_peer(smcphys_handle* h, Peer& p) {
  for (int b = 0; b < 2; b++) { p.xt[b] = h->B.xt[b]; p.st[b] = h->B.st[b]; p.e[b] = h->B.e[b]; p.qs[b] = h->B.qs[b]; }
  p.cdf = h->B.cdf;
  p.ctl = h->ctl;
  p.sums = h->d_sums3;
}

int smcphys_comm_unique_id(uint8_t id[SMCPHYS_UNIQUE_ID_BYTES]) {
  if (!id) return fail(SMCPHYS_EINVAL, "NULL id");
  if (!nccl().load()) return fail(SMCPHYS_ECUDA, nccl().err);
  static_assert(sizeof(ncclUniqueId) == SMCPHYS_UNIQUE_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId u;
  NCCLCK(nccl().GetUniqueId(&u));
  memcpy(id, &u, sizeof(u));
  return 0;
}

int smcphys_comm_init_rank(smcphys_handle* h, const uint8_t id[SMCPHYS_UNIQUE_ID_BYTES], int32_t nranks, int32_t rank) {
  if (!h || !id) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (!h->kids.empty() || h->comm) return fail(SMCPHYS_ESTATE, "handle already spans several shards");
  const int64_t n = h->cfg.num_particles;
  if (nranks < 2 || nranks > SMCPHYS_MAX_RANKS || rank < 0 || rank >= nranks)
    return fail(SMCPHYS_EINVAL, "comm_init_rank: need 2 <= nranks <= 64 and 0 <= rank < nranks");
  if (h->cfg.global_particles != n * nranks || h->cfg.particle_offset != n * rank)
    return fail(SMCPHYS_EINVAL, "comm_init_rank: the handle must hold the slots [rank n, (rank+1) n) of N = nranks n");
  if (!nccl().load()) return fail(SMCPHYS_ECUDA, nccl().err);
  CK(cudaSetDevice(h->cfg.device));
  ncclUniqueId u;
  memcpy(&u, id, sizeof(u));
  NCCLCK(nccl().CommInitRank(&h->comm, nranks, u, rank));
  h->nranks = nranks;
  h->rank = rank;
  // exchange CUDA IPC handles of the buffers the peers store into / read from
  const int NB = 11;
  void* mine[NB] = {h->B.xt[0], h->B.xt[1], h->B.st[0], h->B.st[1], h->B.e[0], h->B.e[1], h->B.qs[0], h->B.qs[1],
                    h->B.cdf, h->ctl, h->d_sums3};
  std::vector<cudaIpcMemHandle_t> hs(NB), all((size_t)NB * nranks);
  for (int b = 0; b < NB; b++) CK(cudaIpcGetMemHandle(&hs[b], mine[b]));
  DevScratch m;
  char *d_send = nullptr, *d_recv = nullptr;
  const size_t bytes = sizeof(cudaIpcMemHandle_t) * NB;
  CK(m.get(&d_send, bytes));
  CK(m.get(&d_recv, bytes * nranks));
  CK(cudaMemcpyAsync(d_send, hs.data(), bytes, cudaMemcpyHostToDevice, h->stream));
  NCCLCK(nccl().AllGather(d_send, d_recv, bytes, ncclChar, h->comm, h->stream));
  CK(cudaMemcpyAsync(all.data(), d_recv, bytes * nranks, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  std::vector<Peer> peers(nranks);
  for (int r = 0; r < nranks; r++) {
    if (r == rank) { fill_own_peer(h, peers[r]); continue; }
    void* p[NB];
    for (int b = 0; b < NB; b++) {
      cudaError_t e = cudaIpcOpenMemHandle(&p[b], all[(size_t)r * NB + b], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess)
        return fail(SMCPHYS_ECUDA, std::string("cudaIpcOpenMemHandle (peer buffers of rank ") + std::to_string(r) +
                                       "): " + cudaGetErrorString(e));
      h->ipc_opened.push_back(p[b]);
    }
    Peer& q = peers[r];
    q.xt[0] = (float*)p[0]; q.xt[1] = (float*)p[1]; q.st[0] = (float*)p[2]; q.st[1] = (float*)p[3];
    q.e[0] = (double*)p[4]; q.e[1] = (double*)p[5]; q.qs[0] = (double*)p[6]; q.qs[1] = (double*)p[7];
    q.cdf = (u64*)p[8]; q.ctl = (Ctl*)p[9]; q.sums = (u64*)p[10];
  }
  CK(cudaMemcpyAsync(h->d_peers, peers.data(), sizeof(Peer) * nranks, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int smcphys_create_multi(const smcphys_config* cfg, int32_t n_shards, const int32_t* devices, smcphys_handle** out) {
  if (!cfg || !out || !devices) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (n_shards < 1 || n_shards > SMCPHYS_MAX_RANKS) return fail(SMCPHYS_EINVAL, "n_shards must be in [1, 64]");
  if (cfg->struct_size != (int32_t)sizeof(smcphys_config)) return fail(SMCPHYS_EINVAL, "smcphys_config size mismatch");
  smcphys_handle* g = new smcphys_handle();
  g->cfg = *cfg;
  g->cfg.global_particles = cfg->num_particles * n_shards;
  int rc = 0;
  for (int r = 0; r < n_shards && !rc; r++) {
    smcphys_config c = *cfg;
    c.device = devices[r];
    c.global_particles = cfg->num_particles * n_shards;
    c.particle_offset = cfg->num_particles * r;
    smcphys_handle* k = nullptr;
    rc = smcphys_create(&c, &k);
    if (rc) break;
    k->parent = g; k->nranks = n_shards; k->rank = r;
    g->kids.push_back(k);
    if (n_shards > 1 && cudaEventCreateWithFlags(&k->ev_sync, cudaEventDisableTiming) != cudaSuccess)
      rc = fail(SMCPHYS_ECUDA, "cudaEventCreate failed");
  }
  // peer access between the devices of the group, then every shard's table of everyone's buffers
  for (int a = 0; a < n_shards && !rc && n_shards > 1; a++)
    for (int b = 0; b < n_shards && !rc; b++) {
      if (devices[a] == devices[b]) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
      if (!can) { rc = fail(SMCPHYS_ECUDA, "devices of the group cannot access each other's memory (no P2P)"); break; }
      cudaSetDevice(devices[a]);
      cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = fail(SMCPHYS_ECUDA, cudaGetErrorString(e));
      cudaGetLastError();
    }
  if (!rc && n_shards > 1) {
    std::vector<Peer> peers(n_shards);
    for (int r = 0; r < n_shards; r++) fill_own_peer(g->kids[r], peers[r]);
    for (auto* k : g->kids) {
      cudaSetDevice(k->cfg.device);
      if (cudaMemcpyAsync(k->d_peers, peers.data(), sizeof(Peer) * n_shards, cudaMemcpyHostToDevice, k->stream) != cudaSuccess ||
          cudaStreamSynchronize(k->stream) != cudaSuccess) {
        rc = fail(SMCPHYS_ECUDA, "peer table upload failed");
        break;
      }
    }
  }
  if (rc) { smcphys_destroy(g); return rc; }
  *out = g;
  return 0;
}

#ifdef RS_STEP_CLOCK
// diagnostic build only: read and clear the per-macro-step cycle counters of the sphere chain
int smcphys_debug_step_clocks(unsigned long long* out, int32_t n) {
  if (!out || n < 1 || n > 4096) return fail(SMCPHYS_EINVAL, "bad arguments");
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(out, g_step_clk, sizeof(unsigned long long) * n));
  std::vector<unsigned long long> z(4096, 0);
  CK(cudaMemcpyToSymbol(g_step_clk, z.data(), sizeof(unsigned long long) * 4096));
  return 0;
}
#endif

int64_t smcphys_total_particles(smcphys_handle* h) {
  if (!h) return 0;
  return h->kids.empty() ? h->cfg.num_particles : h->cfg.num_particles * (int64_t)h->kids.size();
}
int32_t smcphys_steps_done(smcphys_handle* h) {
  if (!h) return -1;
  return h->kids.empty() ? h->steps_done : h->kids[0]->steps_done;
}

// current buffer index (synchronises)
static int current_buffer(smcphys_handle* h, int* cur) {
  Ctl c;
  CK(cudaMemcpyAsync(&c, h->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  *cur = c.cur ^ c.resampled;
  return 0;
}

int smcphys_get_particles(smcphys_handle* h, double* e, double* log_w, double* score, float* traj) {
  if (h && !h->kids.empty()) {
    // shard r holds the global slots [r n, (r+1) n): scalars concatenate, trajectory rows interleave
    const int64_t n = h->cfg.num_particles, G = (int64_t)h->kids.size();
    const int T = h->kids[0]->steps_done;
    std::vector<float> tmp;
    if (traj && T > 0) tmp.resize((size_t)3 * T * n);
    for (int64_t r = 0; r < G; r++) {
      int rc = smcphys_get_particles(h->kids[r], e ? e + r * n : nullptr, log_w ? log_w + r * n : nullptr,
                                     score ? score + r * n : nullptr, tmp.empty() ? nullptr : tmp.data());
      if (rc) return rc;
      if (!tmp.empty())
        for (int64_t row = 0; row < (int64_t)3 * T; row++)
          memcpy(traj + row * (n * G) + r * n, tmp.data() + row * n, sizeof(float) * n);
    }
    return 0;
  }
  NEED_TRIAL(h);
  if (h->steps_done < 0) return fail(SMCPHYS_ESTATE, "get_particles before the filter ran");
  const int64_t n = h->cfg.num_particles;
  const int T = h->steps_done;
  int cur = 0, rc = current_buffer(h, &cur);
  if (rc) return rc;
  if (e) CK(cudaMemcpyAsync(e, h->B.e[cur], sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  if (log_w) CK(cudaMemcpyAsync(log_w, h->B.logw, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  if (score) {
    double* d = (double*)h->B.cdf;  // scratch: the CDF is dead between steps
    LAUNCH(k_score, nblk(n, 256), 256, h->stream, h->B.qs[cur], h->B.xt[cur], h->d_obs, h->tp, n, T, d);
    CK(cudaMemcpyAsync(score, d, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  }
  if (traj && T > 0) {
    // the other history buffer is free between steps: rows 0..T-1 of it receive the simulator outputs
    float* d = h->B.xt[cur ^ 1];
    if (h->cfg.shape == SMCPHYS_SPHERE)
      LAUNCH(k_traj<SHAPE_SPHERE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B.e[cur], h->B.xt[cur], h->tp, h->sp, n,
             h->cfg.substeps, T, d, (double*)nullptr, (float*)nullptr);
    else
      LAUNCH(k_traj<SHAPE_CUBE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B.e[cur], h->B.xt[cur], h->tp, h->sp, n,
             h->cfg.substeps, T, d, (double*)nullptr, (float*)nullptr);
    CK(cudaMemcpyAsync(traj, d, sizeof(float) * 3 * (size_t)T * n, cudaMemcpyDeviceToHost, h->stream));
  }
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int smcphys_get_states(smcphys_handle* h, float* state13) {
  if (h && !h->kids.empty()) return fail(SMCPHYS_ESTATE, "get_states: not available on a multi-device handle");
  NEED_TRIAL(h);
  if (h->steps_done < 0) return fail(SMCPHYS_ESTATE, "get_states before the filter ran");
  if (!state13) return fail(SMCPHYS_EINVAL, "NULL output");
  const int64_t n = h->cfg.num_particles;
  int cur = 0, rc = current_buffer(h, &cur);
  if (rc) return rc;
  // re-simulated with orientation tracking (the rejuvenation kernel does not carry a sphere's quaternion)
  float* soa = h->B.st[cur ^ 1];
  DevScratch m;
  float* aos = nullptr;
  CK(m.get(&aos, (size_t)13 * n));
  if (h->steps_done == 0) {
    soa = h->B.st[cur];
  } else if (h->cfg.shape == SMCPHYS_SPHERE) {
    LAUNCH(k_traj<SHAPE_SPHERE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B.e[cur], h->B.xt[cur], h->tp, h->sp, n,
           h->cfg.substeps, h->steps_done, (float*)nullptr, (double*)nullptr, soa);
  } else {
    LAUNCH(k_traj<SHAPE_CUBE>, nblk(n, RJ_BLOCK), RJ_BLOCK, h->stream, h->B.e[cur], h->B.xt[cur], h->tp, h->sp, n,
           h->cfg.substeps, h->steps_done, (float*)nullptr, (double*)nullptr, soa);
  }
  LAUNCH(k_soa_to_aos, nblk(n, 256), 256, h->stream, soa, aos, n);
  CK(cudaMemcpyAsync(state13, aos, sizeof(float) * 13 * n, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

// log marginal likelihood estimate = sum over resampling steps of (LSE - log N) (kept on the device, Ctl::log_ml) plus
// the same term of the current weights: global max and integer sum W by the kernels of the resampling step (and the
// collectives of the sharded filter; on NCCL ranks this call is collective).
int smcphys_get_log_ml(smcphys_handle* h, double* log_ml) {
  if (!h || !log_ml) return fail(SMCPHYS_EINVAL, "NULL argument");
  std::vector<smcphys_handle*> S;
  if (!h->kids.empty()) S = h->kids;
  else S.push_back(h);
  for (auto* s : S)
    if (!s->trial_set || s->steps_done < 0) return fail(SMCPHYS_ESTATE, "get_log_ml before the filter ran");
  const bool use_nccl = S[0]->comm != nullptr;
  if (!use_nccl && S[0]->cfg.global_particles != S[0]->cfg.num_particles && S.size() == 1)
    return fail(SMCPHYS_ESTATE, "get_log_ml: a phase-driven sharded filter keeps its estimate on the host");
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    const int64_t n = s->cfg.num_particles;
    LAUNCH(k_step_begin, 1, 1, s->stream, s->ctl);
    LAUNCH(k_block_max, nblk(n, 256), 256, s->stream, s->B.logw, n, s->ctl);
    LAUNCH(k_publish_max, 1, 1, s->stream, s->ctl, (double*)nullptr);
    if (!use_nccl && S.size() > 1) LAUNCH(k_copy_lmax, 1, 1, s->stream, s->ctl);
  }
  if (use_nccl) {
    for (auto* s : S) NCCLCK(nccl().AllReduce(&s->ctl->gmax, &s->ctl->gmax, 1, ncclDouble, ncclMax, s->comm, s->stream));
  } else if (S.size() > 1) {
    int rc = group_sync(S);
    if (rc) return rc;
    for (auto* s : S) {
      CK(cudaSetDevice(s->cfg.device));
      LAUNCH(k_local_max, 1, 1, s->stream, s->ctl, s->d_peers, s->nranks);
    }
  }
  u64 W = 0;
  Ctl c;
  for (auto* s : S) {
    CK(cudaSetDevice(s->cfg.device));
    const int64_t n = s->cfg.num_particles;
    LAUNCH(k_weight_sums, s->ntiles, TILE_THREADS, s->stream, s->B, s->ctl, &s->ctl->gmax, n, s->ntiles, 0,
           s->cfg.ess_num, s->cfg.ess_den, 0, (double*)nullptr, (double*)nullptr, (int*)nullptr, s->d_sums3);
    u64 sums[3 * SMCPHYS_MAX_RANKS];
    if (use_nccl) {
      NCCLCK(nccl().AllGather(s->d_sums3, s->d_all, 3, ncclUint64, s->comm, s->stream));
      CK(cudaMemcpyAsync(sums, s->d_all, sizeof(u64) * 3 * s->nranks, cudaMemcpyDeviceToHost, s->stream));
    } else {
      CK(cudaMemcpyAsync(sums, s->d_sums3, sizeof(u64) * 3, cudaMemcpyDeviceToHost, s->stream));
    }
    CK(cudaMemcpyAsync(&c, s->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    for (int r = 0; r < (use_nccl ? s->nranks : 1); r++) W += sums[3 * r];
  }
  *log_ml = c.log_ml + c.gmax + std::log((double)W / 549755813888.0) - std::log((double)S[0]->cfg.global_particles);
  return 0;
}

int smcphys_predict(smcphys_handle* h, int32_t n_steps, float* pos_out, float* obs_out) {
  if (h && !h->kids.empty()) return fail(SMCPHYS_ESTATE, "predict: not available on a multi-device handle");
  NEED_TRIAL(h);
  if (h->steps_done < 0) return fail(SMCPHYS_ESTATE, "predict before the filter ran");
  if (n_steps < 1) return fail(SMCPHYS_EINVAL, "n_steps < 1");
  const int64_t n = h->cfg.num_particles;
  int cur = 0, rc = current_buffer(h, &cur);
  if (rc) return rc;
  DevScratch m;
  float *d_pos = nullptr, *d_obs = nullptr;
  size_t bytes = sizeof(float) * 3 * (size_t)n_steps * n;
  if (pos_out) CK(m.get(&d_pos, (size_t)3 * n_steps * n));
  if (obs_out) CK(m.get(&d_obs, (size_t)3 * n_steps * n));
  if (h->cfg.shape == SMCPHYS_SPHERE)
    LAUNCH(k_predict<SHAPE_SPHERE>, nblk(n, PF_BLOCK), PF_BLOCK, h->stream, h->B.e[cur], h->B.st[cur], h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, h->cfg.substeps, h->steps_done, n_steps, d_pos, d_obs);
  else
    LAUNCH(k_predict<SHAPE_CUBE>, nblk(n, PF_BLOCK), PF_BLOCK, h->stream, h->B.e[cur], h->B.st[cur], h->tp, h->sp, n,
           h->cfg.particle_offset, h->cfg.seed, h->cfg.substeps, h->steps_done, n_steps, d_pos, d_obs);
  if (pos_out) CK(cudaMemcpyAsync(pos_out, d_pos, bytes, cudaMemcpyDeviceToHost, h->stream));
  if (obs_out) CK(cudaMemcpyAsync(obs_out, d_obs, bytes, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

// ---------------------------------------------------------------- sharded filter: host-side integer helpers
int smcphys_ess_below(uint64_t S1, uint64_t S2, int64_t N, int64_t num, int64_t den) {
  unsigned __int128 lhs = (unsigned __int128)S1 * S1 * (u64)den;
  unsigned __int128 rhs = (unsigned __int128)S2 * (u64)N * (u64)num;
  return lhs < rhs ? 1 : 0;
}
double smcphys_ess_value(uint64_t S1, uint64_t S2) { return S2 ? ((double)S1 * (double)S1) / (double)S2 : 0.0; }
double smcphys_lse_value(double global_max, uint64_t W) { return global_max + std::log((double)W / 549755813888.0); }
double smcphys_resample_offset(uint64_t seed, int32_t t) {
  uint32_t r[4];
  philox4x32_10(seed, 0u, (uint32_t)t, 0u, (uint32_t)ST_RESAMPLE, r);
  return (double)r[0] / 4294967296.0;
}
int smcphys_pf_offspring_range(uint64_t W_before, uint64_t W_local, uint64_t W_total, int64_t N, double u0,
                               int64_t* k_lo, int64_t* k_hi) {
  if (!k_lo || !k_hi || N < 1 || W_total == 0 || W_before + W_local > W_total || !(u0 >= 0.0 && u0 < 1.0))
    return fail(SMCPHYS_EINVAL, "pf_offspring_range: bad arguments");
  u64 U0 = (u64)(u0 * 4294967296.0);
  *k_lo = count_below(W_before, U0, W_total, N);
  *k_hi = count_below(W_before + W_local, U0, W_total, N);
  return 0;
}
int64_t smcphys_pf_payload_words(smcphys_handle* h, int32_t t) { return h ? 17 + 3 * (int64_t)(t + 1) : 0; }

int smcphys_pf_pack(smcphys_handle* h, int32_t t, const double* d_global_max, uint64_t W_before, uint64_t W_total,
                    double u0, int64_t k_lo, int64_t k_hi, int64_t slots_per_rank, float* d_send) {
  NEED_TRIAL(h);
  if (t != h->steps_done - 1) return fail(SMCPHYS_ESTATE, "pf_pack: t is not the current step");
  if (!d_global_max || k_hi < k_lo || slots_per_rank < 1) return fail(SMCPHYS_EINVAL, "pf_pack: bad arguments");
  const int64_t n = h->cfg.num_particles, count = k_hi - k_lo;
  if (count == 0) return 0;
  if (!d_send) return fail(SMCPHYS_EINVAL, "pf_pack: NULL send buffer");
  u64 U0 = (u64)(u0 * 4294967296.0);
  ProfScope ps(h, SMCPHYS_K_GATHER);
  LAUNCH(k_cdf, h->ntiles, TILE_THREADS, h->stream, h->B, h->ctl, d_global_max, n, (u64)W_before, 0, (const u64*)nullptr);
  const int64_t words = 17 + 3 * (int64_t)(t + 1);
  // a shard's offspring can outnumber its particles; the index scratch holds n, so go in slabs of n
  for (int64_t off = 0; off < count; off += n) {
    int64_t c = std::min(n, count - off);
    LAUNCH(k_index, nblk(c, 256), 256, h->stream, h->B, h->ctl, n, h->cfg.global_particles, U0, (const u64*)nullptr,
           (u64)W_total, k_lo + off, c, 0);
    LAUNCH(k_pack, dim3(nblk(c, 256), (unsigned)(t + 2)), 256, h->stream, h->B, h->ctl, n, k_lo, count, off, c,
           slots_per_rank, words, d_send);
  }
  return 0;
}

int smcphys_pf_gather_local(smcphys_handle* h, int32_t t, const double* d_global_max, uint64_t W_before,
                            uint64_t W_total, double u0, int64_t k_lo, int64_t k_hi, int64_t slot0) {
  NEED_TRIAL(h);
  if (t != h->steps_done - 1) return fail(SMCPHYS_ESTATE, "pf_gather_local: t is not the current step");
  const int64_t n = h->cfg.num_particles, count = k_hi - k_lo;
  if (!d_global_max || count < 0 || slot0 < 0 || slot0 + count > n)
    return fail(SMCPHYS_EINVAL, "pf_gather_local: bad arguments");
  h->local_gathered = count;
  if (count == 0) return 0;
  u64 U0 = (u64)(u0 * 4294967296.0);
  ProfScope ps(h, SMCPHYS_K_GATHER);
  LAUNCH(k_cdf, h->ntiles, TILE_THREADS, h->stream, h->B, h->ctl, d_global_max, n, (u64)W_before, 0, (const u64*)nullptr);
  LAUNCH(k_index, nblk(count, 256), 256, h->stream, h->B, h->ctl, n, h->cfg.global_particles, U0, (const u64*)nullptr,
         (u64)W_total, k_lo, count, 0);
  LAUNCH(k_gather, nblk(count, 256), 256, h->stream, h->B, h->ctl, n, count, slot0, 0);
  LAUNCH(k_gather_hist, dim3(nblk(count, 256), (unsigned)((t + GH_ROWS) / GH_ROWS)), 256, h->stream, h->B, h->ctl, n,
         t + 1, count, slot0, 0);
  return 0;
}
int smcphys_pf_unpack(smcphys_handle* h, int32_t t, const float* d_recv, int32_t n_chunks, const int64_t* counts,
                      const int64_t* slot0, const int64_t* word_off) {
  NEED_TRIAL(h);
  if (t != h->steps_done - 1) return fail(SMCPHYS_ESTATE, "pf_unpack: t is not the current step");
  if (n_chunks < 0 || (n_chunks > 0 && (!d_recv || !counts || !slot0 || !word_off)))
    return fail(SMCPHYS_EINVAL, "pf_unpack: bad arguments");
  const int64_t n = h->cfg.num_particles;
  int64_t total = 0;
  for (int c = 0; c < n_chunks; c++) {
    if (counts[c] < 0 || slot0[c] < 0 || slot0[c] + counts[c] > n)
      return fail(SMCPHYS_EINVAL, "pf_unpack: chunk outside the shard");
    total += counts[c];
  }
  if (total + h->local_gathered != n)
    return fail(SMCPHYS_EINVAL, "pf_unpack: chunks (+ pf_gather_local) must fill every local slot exactly once");
  h->local_gathered = 0;
  ProfScope ps(h, SMCPHYS_K_GATHER);
  for (int c = 0; c < n_chunks; c++) {
    if (counts[c] == 0) continue;
    LAUNCH(k_unpack, dim3(nblk(counts[c], 256), (unsigned)(t + 2)), 256, h->stream, h->B, h->ctl, n,
           d_recv + word_off[c], counts[c], slot0[c]);
  }
  LAUNCH(k_set_resampled, 1, 1, h->stream, h->ctl, 1);
  return 0;
}

// ---------------------------------------------------------------- stateless entries (host pointers)
static int use_device(int32_t device) {
  int ndev = smcphys_device_count();
  if (ndev <= 0) return fail(SMCPHYS_ECUDA, "no CUDA device (libsmcphys has no CPU path)");
  if (device < 0 || device >= ndev) return fail(SMCPHYS_EINVAL, "device ordinal out of range");
  CK(cudaSetDevice(device));
  return 0;
}

int smcphys_sim_step(int32_t device, int32_t shape, int32_t S, double ceiling_half, int64_t n, float* state13,
                     const float* restitution) {
  if (n < 0 || (n > 0 && (!state13 || !restitution))) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (shape != SMCPHYS_SPHERE && shape != SMCPHYS_CUBE) return fail(SMCPHYS_EINVAL, "shape must be 0 or 1");
  if (S < 1 || S > 64) return fail(SMCPHYS_EINVAL, "substeps must be in [1, 64]");
  int rc = use_device(device);
  if (rc) return rc;
  if (n == 0) return 0;
  DevScratch m;
  float *aos, *soa, *r;
  CK(m.get(&aos, (size_t)13 * n)); CK(m.get(&soa, (size_t)13 * n)); CK(m.get(&r, (size_t)n));
  cudaStream_t st = 0;
  ScenePar sp; sp.ceil_half = (float)ceiling_half;
  CK(cudaMemcpyAsync(aos, state13, sizeof(float) * 13 * n, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(r, restitution, sizeof(float) * n, cudaMemcpyHostToDevice, st));
  LAUNCH(k_aos_to_soa, nblk(n, 256), 256, st, aos, soa, n);
  if (shape == SMCPHYS_SPHERE) LAUNCH(k_sim_step<SHAPE_SPHERE>, nblk(n, PF_BLOCK), PF_BLOCK, st, soa, r, sp, n, S);
  else LAUNCH(k_sim_step<SHAPE_CUBE>, nblk(n, PF_BLOCK), PF_BLOCK, st, soa, r, sp, n, S);
  LAUNCH(k_soa_to_aos, nblk(n, 256), 256, st, soa, aos, n);
  CK(cudaMemcpyAsync(state13, aos, sizeof(float) * 13 * n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

int smcphys_simulate(int32_t device, int32_t shape, int32_t S, double ceiling_half, double sigma_tr,
                     const double init_pos[3], const double init_vel[3], const float* e, const float* xt, int64_t n,
                     int32_t T, float* mu_out, double* logp_out, float* state_out) {
  if (n < 0 || T < 0 || !init_pos || !init_vel || (n > 0 && !e) || (n > 0 && T > 0 && !xt))
    return fail(SMCPHYS_EINVAL, "NULL or negative argument");
  if (shape != SMCPHYS_SPHERE && shape != SMCPHYS_CUBE) return fail(SMCPHYS_EINVAL, "shape must be 0 or 1");
  if (S < 1 || S > 64) return fail(SMCPHYS_EINVAL, "substeps must be in [1, 64]");
  if (!(sigma_tr > 0)) return fail(SMCPHYS_EINVAL, "sigma_tr must be positive");
  int rc = use_device(device);
  if (rc) return rc;
  if (n == 0) return 0;
  DevScratch m;
  cudaStream_t st = 0;
  float *d_e, *d_xt, *d_mu = nullptr, *d_soa, *d_aos;
  double *d_ed, *d_q;
  const size_t hist = (size_t)std::max(T, 1) * 3 * n;
  CK(m.get(&d_e, (size_t)n)); CK(m.get(&d_ed, (size_t)n)); CK(m.get(&d_xt, hist)); CK(m.get(&d_q, (size_t)n));
  CK(m.get(&d_soa, (size_t)13 * n)); CK(m.get(&d_aos, (size_t)13 * n));
  if (mu_out) CK(m.get(&d_mu, hist));
  TrialPar tp;
  for (int c = 0; c < 3; c++) { tp.x0[c] = (float)init_pos[c]; tp.v0[c] = (float)init_vel[c]; }
  tp.sigma_tr = sigma_tr; tp.sigma_obs = 1.0;
  ScenePar sp; sp.ceil_half = (float)ceiling_half;
  CK(cudaMemcpyAsync(d_e, e, sizeof(float) * n, cudaMemcpyHostToDevice, st));
  if (T > 0) CK(cudaMemcpyAsync(d_xt, xt, sizeof(float) * 3 * (size_t)T * n, cudaMemcpyHostToDevice, st));
  LAUNCH(k_f2d, nblk(n, 256), 256, st, d_e, d_ed, n);
  if (shape == SMCPHYS_SPHERE) {
    // trajectory + transition density through the rejuvenation kernel's chain code, the final state (with the
    // orientation the chain does not carry) through the general substep: the parity tests see both
    LAUNCH(k_traj<SHAPE_SPHERE>, nblk(n, RJ_BLOCK), RJ_BLOCK, st, d_ed, d_xt, tp, sp, n, S, T, d_mu, d_q, (float*)nullptr);
    LAUNCH(k_traj<SHAPE_SPHERE>, nblk(n, RJ_BLOCK), RJ_BLOCK, st, d_ed, d_xt, tp, sp, n, S, T, (float*)nullptr,
           (double*)nullptr, d_soa);
  } else
    LAUNCH(k_traj<SHAPE_CUBE>, nblk(n, RJ_BLOCK), RJ_BLOCK, st, d_ed, d_xt, tp, sp, n, S, T, d_mu, d_q, d_soa);
  if (mu_out && T > 0) CK(cudaMemcpyAsync(mu_out, d_mu, sizeof(float) * 3 * (size_t)T * n, cudaMemcpyDeviceToHost, st));
  std::vector<double> q;
  if (logp_out) {
    q.resize(n);
    CK(cudaMemcpyAsync(q.data(), d_q, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
  }
  if (state_out) {
    LAUNCH(k_soa_to_aos, nblk(n, 256), 256, st, d_soa, d_aos, n);
    CK(cudaMemcpyAsync(state_out, d_aos, sizeof(float) * 13 * n, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));
  if (logp_out) {
    const double cst = 3.0 * (std::log(sigma_tr) + SMC_HALF_LOG_2PI) * (double)std::max(T - 1, 0);
    for (int64_t i = 0; i < n; i++) logp_out[i] = -0.5 * q[i] / (sigma_tr * sigma_tr) - cst;
  }
  return 0;
}

// shared by the two weight entries: max -> integer sums (+ tile offsets) on a scratch handle-less Bufs
static int weights_on_device(const double* log_w, int64_t n, DevScratch& m, Bufs& B, Ctl** ctl_out, cudaStream_t st) {
  Ctl* ctl;
  int ntiles = (int)((n + TILE - 1) / TILE);
  memset(&B, 0, sizeof(B));
  CK(m.get(&B.logw, (size_t)n)); CK(m.get(&B.cdf, (size_t)n)); CK(m.get(&B.idx, (size_t)n));
  CK(m.get(&B.tile_sum, (size_t)ntiles)); CK(m.get(&B.tile_off, (size_t)ntiles)); CK(m.get(&ctl, 1));
  CK(cudaMemsetAsync(ctl, 0, sizeof(Ctl), st));
  CK(cudaMemcpyAsync(B.logw, log_w, sizeof(double) * n, cudaMemcpyHostToDevice, st));
  LAUNCH(k_block_max, nblk(n, 256), 256, st, B.logw, n, ctl);
  LAUNCH(k_publish_max, 1, 1, st, ctl, (double*)nullptr);
  LAUNCH(k_weight_sums, ntiles, TILE_THREADS, st, B, ctl, &ctl->gmax, n, ntiles, 0, (int64_t)1, (int64_t)2, 0,
         (double*)nullptr, (double*)nullptr, (int*)nullptr, (u64*)nullptr);
  *ctl_out = ctl;
  return 0;
}

int smcphys_fp32_peak(int32_t device, double* tflops) {
  if (!tflops) return fail(SMCPHYS_EINVAL, "NULL output");
  int rc = use_device(device);
  if (rc) return rc;
  DevScratch m;
  float* d;
  CK(m.get(&d, 4));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, iters = 1 << 16;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    CK(cudaEventRecord(e0, 0));
    LAUNCH(k_fma_peak, blocks, 256, 0, d, iters, 0.999f, 0.001f);
    CK(cudaEventRecord(e1, 0));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    double tf = 2.0 * 8.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *tflops = best;
  return 0;
}

int smcphys_weight_stats(int32_t device, const double* log_w, int64_t n, double* out3, uint64_t* sums3) {
  if (n < 1 || n > (1ll << 24) || !log_w) return fail(SMCPHYS_EINVAL, "n must be in [1, 2^24] and log_w non-NULL");
  int rc = use_device(device);
  if (rc) return rc;
  DevScratch m;
  Bufs B;
  Ctl *ctl, c;
  rc = weights_on_device(log_w, n, m, B, &ctl, 0);
  if (rc) return rc;
  CK(cudaMemcpyAsync(&c, ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, 0));
  CK(cudaStreamSynchronize(0));
  if (out3) { out3[0] = c.gmax; out3[1] = smcphys_ess_value(c.S1, c.S2); out3[2] = smcphys_lse_value(c.gmax, c.W); }
  if (sums3) { sums3[0] = c.W; sums3[1] = c.S1; sums3[2] = c.S2; }
  return 0;
}

int smcphys_resample_systematic(int32_t device, const double* log_w, int64_t n, double u0, int64_t* idx_out) {
  if (n < 0 || n > (1ll << 24)) return fail(SMCPHYS_EINVAL, "n must be in [0, 2^24]");
  if (n == 0) return 0;
  if (!log_w || !idx_out) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (!(u0 >= 0.0 && u0 < 1.0)) return fail(SMCPHYS_EINVAL, "u0 must be in [0, 1)");
  int rc = use_device(device);
  if (rc) return rc;
  DevScratch m;
  Bufs B;
  Ctl* ctl;
  cudaStream_t st = 0;
  rc = weights_on_device(log_w, n, m, B, &ctl, st);
  if (rc) return rc;
  int64_t* d_out;
  CK(m.get(&d_out, (size_t)n));
  int ntiles = (int)((n + TILE - 1) / TILE);
  LAUNCH(k_cdf, ntiles, TILE_THREADS, st, B, ctl, &ctl->gmax, n, (u64)0, 0, (const u64*)nullptr);
  LAUNCH(k_index, nblk(n, 256), 256, st, B, ctl, n, n, (u64)(u0 * 4294967296.0), &ctl->W, (u64)0, (int64_t)0, n, 0);
  LAUNCH(k_idx_to_i64, nblk(n, 256), 256, st, B.idx, d_out, n);
  CK(cudaMemcpyAsync(idx_out, d_out, sizeof(int64_t) * n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}


int smcphys_resample_multinomial(int32_t device, const double* log_w, int64_t n, uint64_t seed, int32_t t,
                                 int64_t* idx_out) {
  if (n < 0 || n > (1ll << 24)) return fail(SMCPHYS_EINVAL, "n must be in [0, 2^24]");
  if (n == 0) return 0;
  if (!log_w || !idx_out) return fail(SMCPHYS_EINVAL, "NULL argument");
  int rc = use_device(device);
  if (rc) return rc;
  DevScratch m;
  Bufs B;
  Ctl* ctl;
  cudaStream_t st = 0;
  rc = weights_on_device(log_w, n, m, B, &ctl, st);
  if (rc) return rc;
  int64_t* d_out;
  CK(m.get(&d_out, (size_t)n));
  int ntiles = (int)((n + TILE - 1) / TILE);
  LAUNCH(k_cdf, ntiles, TILE_THREADS, st, B, ctl, &ctl->gmax, n, (u64)0, 0, (const u64*)nullptr);
  LAUNCH(k_index_multi, nblk(n, 256), 256, st, B, ctl, n, &ctl->W, (u64)seed, (int)t, (int64_t)0, n, 0);
  LAUNCH(k_idx_to_i64, nblk(n, 256), 256, st, B.idx, d_out, n);
  CK(cudaMemcpyAsync(idx_out, d_out, sizeof(int64_t) * n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}


// ---------------------------------------------------------------- allocation-free entry
// smcphys_resample_systematic on DEVICE buffers with every byte of scratch in a file-scope device array: nothing here
// calls cudaMalloc / cudaFree or any allocating routine, so it can be used where those are forbidden (CUDA graph
// capture, tight serving loops).  Scratch layout (256-byte aligned pieces): integer CDF u64[MAX_N], ancestors
// i32[MAX_N], tile sums / offsets u64[MAX_N / 1024] each, control block.  Every GPU that loads the library holds the
// array until exit; every call shares it, so calls that overlap on one GPU race.
#define NA_MAX_N SMCPHYS_NOALLOC_MAX_N
#define NA_TILES ((NA_MAX_N + TILE - 1) / TILE)
#define NA_ALIGN(x) (((x) + 255) & ~(size_t)255)
#define NA_OFF_CDF ((size_t)0)
#define NA_OFF_IDX (NA_OFF_CDF + NA_ALIGN(sizeof(u64) * (size_t)NA_MAX_N))
#define NA_OFF_TSUM (NA_OFF_IDX + NA_ALIGN(sizeof(int) * (size_t)NA_MAX_N))
#define NA_OFF_TOFF (NA_OFF_TSUM + NA_ALIGN(sizeof(u64) * (size_t)NA_TILES))
#define NA_OFF_CTL (NA_OFF_TOFF + NA_ALIGN(sizeof(u64) * (size_t)NA_TILES))
#define NA_BYTES (NA_OFF_CTL + NA_ALIGN(sizeof(Ctl)))
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

}  // extern "C"
#include <mutex>
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, taken once per device with that device current.
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[SMCPHYS_MAX_RANKS] = {nullptr};
static int noalloc_scratch(int device, unsigned char** out) {
  if (device < 0 || device >= SMCPHYS_MAX_RANKS) return fail(SMCPHYS_EINVAL, "device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // the caller has made `device` current
    g_noalloc_ptr[device] = (unsigned char*)p;
  }
  *out = g_noalloc_ptr[device];
  return 0;
}
extern "C" {

int smcphys_resample_systematic_noalloc(int32_t device, const double* d_log_w, int64_t n, double u0, int64_t* d_idx_out,
                                        void* cuda_stream) {
  if (n < 0) return fail(SMCPHYS_EINVAL, "n must be >= 0");
  if (n > (int64_t)NA_MAX_N)
    return fail(SMCPHYS_EINVAL, "resample_systematic_noalloc: n = " + std::to_string(n) + " exceeds SMCPHYS_NOALLOC_MAX_N = " +
                                    std::to_string((long long)NA_MAX_N) + " (the static scratch array); use "
                                    "smcphys_resample_systematic for larger inputs");
  if (n == 0) return 0;
  if (!d_log_w || !d_idx_out) return fail(SMCPHYS_EINVAL, "NULL argument");
  if (!(u0 >= 0.0 && u0 < 1.0)) return fail(SMCPHYS_EINVAL, "u0 must be in [0, 1)");
  int rc = use_device(device);
  if (rc) return rc;
  unsigned char* base = nullptr;
  rc = noalloc_scratch(device, &base);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  Bufs B;
  memset(&B, 0, sizeof(B));
  B.logw = const_cast<double*>(d_log_w);  // read only by the kernels below
  B.cdf = (u64*)(base + NA_OFF_CDF);
  B.idx = (int*)(base + NA_OFF_IDX);
  B.tile_sum = (u64*)(base + NA_OFF_TSUM);
  B.tile_off = (u64*)(base + NA_OFF_TOFF);
  Ctl* ctl = (Ctl*)(base + NA_OFF_CTL);
  const int ntiles = (int)((n + TILE - 1) / TILE);
  // the array keeps the last call's contents: the control block (running max, integer sums, block counter) must start
  // at zero; CDF, indices and tile sums are fully rewritten for [0, n)
  CK(cudaMemsetAsync(ctl, 0, sizeof(Ctl), st));
  LAUNCH(k_block_max, nblk(n, 256), 256, st, B.logw, n, ctl);
  LAUNCH(k_publish_max, 1, 1, st, ctl, (double*)nullptr);
  LAUNCH(k_weight_sums, ntiles, TILE_THREADS, st, B, ctl, &ctl->gmax, n, ntiles, 0, (int64_t)1, (int64_t)2, 0,
         (double*)nullptr, (double*)nullptr, (int*)nullptr, (u64*)nullptr);
  LAUNCH(k_cdf, ntiles, TILE_THREADS, st, B, ctl, &ctl->gmax, n, (u64)0, 0, (const u64*)nullptr);
  LAUNCH(k_index, nblk(n, 256), 256, st, B, ctl, n, n, (u64)(u0 * 4294967296.0), &ctl->W, (u64)0, (int64_t)0, n, 0);
  LAUNCH(k_idx_to_i64, nblk(n, 256), 256, st, B.idx, d_idx_out, n);
  return 0;
}

}  // extern "C"
