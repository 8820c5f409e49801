// Device-side rigid-body transition of the bouncing-object model (float32, one body per thread).
//
// Replaces, for one particle: PhySMC.step = sync! ; forward_step  (PhySMC/PhySMC.jl:32-35,
// PhySMC/PhyBullet/PhyBullet.jl:68-103, Elements/rigid_body.jl:33-76,129-139) and the Bullet
// stepSimulation() it drives, for the scene of Model/bouncing_object.jl:24-86.
// The scene is compile-time constant data; the body state lives in registers.  The contact path (narrow
// phase + projected Gauss-Seidel rows) is out of line and only entered by threads that are near a static.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace smcphys {

enum { SHAPE_SPHERE = 0, SHAPE_CUBE = 1 };

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float comp3(f3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
__device__ __forceinline__ void setc3(f3& a, int i, float v) {
  if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v;
}
struct f4 { float x, y, z, w; };  // quaternion xyzw (rigid_body.jl:21)

struct Body {
  f3 x;  // position
  f4 q;  // orientation
  f3 v;  // linear velocity
  f3 w;  // angular velocity, world frame
};

// ---- scene constants (Model/bouncing_object.jl:24-60) and Bullet solver constants
#define PH_DT (1.0f / 240.0f)             // pb_timestep, PhyBullet.jl:27
#define PH_H 0.05f                        // sphere radius / cube half extent, bouncing_object.jl:68,71
#define PH_G (-9.81f)                     // bouncing_object.jl:58
#define PH_STATIC_REST 0.5f               // restitution of every static, bouncing_object.jl:29-55
#define PH_FRICTION 0.25f                 // lateralFriction 0.5 (body) x 0.5 (static)
#define PH_DAMP 0.04f                     // btMultiBody linear = angular damping
#define PH_ERP 0.08f                      // erp2
#define PH_SLOP 1e-5f                     // linearSlop
#define PH_REST_THRESH 0.2f               // restitutionVelocityThreshold
#define PH_RESID 1e-7f                    // leastSquaresResidualThreshold
#ifndef PH_ITERS
#define PH_ITERS 50                       // numIterations
#endif
#define PH_STALL 1e-6f                     // early exit of an infeasible solve: velocity change per sweep
#define PH_BOX_MARGIN 0.001f              // collision margin of a 0.01-thin btBoxShape
#define PH_ANG_THRESH 0.78539816339744830962f
#define PH_THR (0.02f * 1.7320508075688772f * PH_H)  // contact breaking threshold (0.02 x AABB radius)
#define PH_MAXC 8
#define PH_INV_M 1.0f                                // mass 1 (bouncing_object.jl:63)
#define PH_INV_I_SPHERE (1.0f / (0.4f * PH_H * PH_H))          // 2/5 m r^2
#define PH_INV_I_CUBE (6.0f / ((2.0f * PH_H) * (2.0f * PH_H)))  // m (2h)^2 / 6

struct ScenePar {
  float ceil_half;  // 0.01 (bouncing_object.jl:53); Tests/generate_trajectories.jl:52 uses 0.02
};

// the five static boxes in creation order (bouncing_object.jl:34-55)
__device__ __forceinline__ void static_box(int b, float ceil_half, f3& c, f3& h) {
  switch (b) {
    case 0: c = mk3(-0.51f, 0.f, 0.51f); h = mk3(0.01f, 0.5f, 0.5f); break;
    case 1: c = mk3(0.f, 0.51f, 0.51f); h = mk3(0.5f, 0.01f, 0.5f); break;
    case 2: c = mk3(0.51f, 0.f, 0.51f); h = mk3(0.01f, 0.5f, 0.5f); break;
    case 3: c = mk3(0.f, -0.51f, 0.51f); h = mk3(0.5f, 0.01f, 0.5f); break;
    default: c = mk3(0.f, 0.f, 1.01f); h = mk3(0.5f, 0.5f, ceil_half); break;
  }
}

struct Contact {
  f3 r;  // arm: contact point on the body minus centre
  f3 n;  // normal, static -> body
  float dist;
};

// persistent plane manifold of the cube, reset at every macro step (sync! re-poses the body)
struct Manifold {
  int n;
  f3 la[4];  // body-local corner
  f3 wb[4];  // world point on the plane
};

__device__ __forceinline__ void rotmat(const f4& q, float R[3][3]) {
  float d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  float s = 2.0f / d;
  float xs = q.x * s, ys = q.y * s, zs = q.z * s;
  float wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  float xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
  float yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  R[0][0] = 1.0f - (yy + zz); R[0][1] = xy - wz; R[0][2] = xz + wy;
  R[1][0] = xy + wz; R[1][1] = 1.0f - (xx + zz); R[1][2] = yz - wx;
  R[2][0] = xz - wy; R[2][1] = yz + wx; R[2][2] = 1.0f - (xx + yy);
}
__device__ __forceinline__ f3 rotv(const float R[3][3], f3 v) {
  return mk3(R[0][0] * v.x + R[0][1] * v.y + R[0][2] * v.z, R[1][0] * v.x + R[1][1] * v.y + R[1][2] * v.z,
             R[2][0] * v.x + R[2][1] * v.y + R[2][2] * v.z);
}

// A body whose centre is inside this region touches nothing (every static is farther than the bounding
// radius + margin + breaking threshold), so the narrow phase and the solver are skipped and a cube's
// manifold empties (all of its cached points are beyond the breaking threshold).
template <int SHAPE>
__device__ __forceinline__ bool body_clear(const f3& x, float ceil_half) {
  const float rad = (SHAPE == SHAPE_SPHERE) ? PH_H : 1.7320508075688772f * PH_H;
  const float pad = rad + 2.0f * PH_THR;
  return fabsf(x.x) < 0.5f - pad && fabsf(x.y) < 0.5f - pad && x.z > pad && x.z < 1.01f - ceil_half - pad;
}

// ---------------------------------------------------------------- narrow phase
// cube: one deepest support vertex per substep into the persistent plane manifold (keyed on the body-local
// corner, refreshed and pruned every substep), then face-vertex contacts against the five boxes along the
// axis of least overlap (the documented modelling choice of the oracle; Bullet's btBoxBoxDetector is not
// restated anywhere in the reference tree).
// The loops over the manifold slots and the face vertices are kept rolled: the cube kernel is bound by instruction
// fetch, and unrolling them added a thousand instructions to the substep for no arithmetic saved.
__device__ __forceinline__ int collide_cube(const ScenePar& sp, const Body& s, Manifold& pm, Contact* out) {
  const float h = PH_H, thr = PH_THR;
  int nc = 0;
  float R[3][3];
  rotmat(s.q, R);
  {
    f3 vl = mk3((-R[2][0]) >= 0.f ? h : -h, (-R[2][1]) >= 0.f ? h : -h, (-R[2][2]) >= 0.f ? h : -h);
    f3 vw = s.x + rotv(R, vl);
    float d = vw.z;
    if (d < thr) {
      int best = -1;
      float sd = thr * thr;
#pragma unroll 1
      for (int i = 0; i < pm.n; i++) {
        f3 dd = pm.la[i] - vl;
        float l2 = dot3(dd, dd);
        if (l2 < sd) { sd = l2; best = i; }
      }
      f3 wb = mk3(vw.x, vw.y, 0.f);
      if (best >= 0) { pm.la[best] = vl; pm.wb[best] = wb; }
      else if (pm.n < 4) { pm.la[pm.n] = vl; pm.wb[pm.n] = wb; pm.n++; }
    }
    float dist[4];
    f3 wa[4];
#pragma unroll 1
    for (int i = pm.n - 1; i >= 0; i--) {
      wa[i] = s.x + rotv(R, pm.la[i]);
      dist[i] = wa[i].z - pm.wb[i].z;
    }
#pragma unroll 1
    for (int i = pm.n - 1; i >= 0; i--) {
      bool drop = false;
      if (dist[i] > thr) drop = true;
      else {
        float dx = pm.wb[i].x - wa[i].x, dy = pm.wb[i].y - wa[i].y;
        if (dx * dx + dy * dy > thr * thr) drop = true;
      }
      if (drop) {
        int last = pm.n - 1;
        if (i != last) { pm.la[i] = pm.la[last]; pm.wb[i] = pm.wb[last]; wa[i] = wa[last]; dist[i] = dist[last]; }
        pm.n--;
      }
    }
#pragma unroll 1
    for (int i = 0; i < pm.n; i++) { out[nc].r = wa[i] - s.x; out[nc].n = mk3(0.f, 0.f, 1.f); out[nc].dist = dist[i]; nc++; }
  }
  // only a box whose inner face is within the bounding radius of the centre can overlap the cube (the exact
  // overlap test below would reject the others anyway): usually none when the cube is on the floor
  const float rad = 1.7320508075688772f * h + 1e-6f;
  unsigned int near = 0;
  if (s.x.x < -0.5f + rad) near |= 1u;
  if (s.x.y > 0.5f - rad) near |= 2u;
  if (s.x.x > 0.5f - rad) near |= 4u;
  if (s.x.y < -0.5f + rad) near |= 8u;
  if (s.x.z > 1.01f - sp.ceil_half - rad) near |= 16u;
#pragma unroll 1
  for (int b = 0; b < 5; b++) {
    if (!((near >> b) & 1u)) continue;
    f3 c, bh;
    static_box(b, sp.ceil_half, c, bh);
    f3 d = s.x - c;
    float ex[3];
#pragma unroll
    for (int i = 0; i < 3; i++) ex[i] = h * (fabsf(R[i][0]) + fabsf(R[i][1]) + fabsf(R[i][2]));
    float ov[3] = {bh.x + ex[0] - fabsf(d.x), bh.y + ex[1] - fabsf(d.y), bh.z + ex[2] - fabsf(d.z)};
    if (ov[0] <= 0.f || ov[1] <= 0.f || ov[2] <= 0.f) continue;
    int ax = 0;
    if (ov[1] < ov[ax]) ax = 1;
    if (ov[2] < ov[ax]) ax = 2;
    float sgn = comp3(d, ax) >= 0.f ? 1.f : -1.f;
    f3 n = mk3(0.f, 0.f, 0.f);
    setc3(n, ax, sgn);
    float face = comp3(c, ax) + sgn * comp3(bh, ax);
    int j = 0;
    if (fabsf(R[ax][1]) > fabsf(R[ax][j])) j = 1;
    if (fabsf(R[ax][2]) > fabsf(R[ax][j])) j = 2;
    float js = (R[ax][j] * sgn > 0.f) ? -1.f : 1.f;
    int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    int a1 = (ax + 1) % 3, a2 = (ax + 2) % 3;
#pragma unroll 1
    for (int k = 0; k < 4; k++) {
      f3 vl = mk3(0.f, 0.f, 0.f);
      setc3(vl, j, js * h);
      setc3(vl, j1, (k & 1) ? h : -h);
      setc3(vl, j2, (k & 2) ? h : -h);
      f3 rw = rotv(R, vl);
      f3 vw = s.x + rw;
      float gap = sgn * (comp3(vw, ax) - face);
      if (gap > 0.f) continue;
      if (fabsf(comp3(vw, a1) - comp3(c, a1)) > comp3(bh, a1)) continue;
      if (fabsf(comp3(vw, a2) - comp3(c, a2)) > comp3(bh, a2)) continue;
      if (nc < PH_MAXC) { out[nc].r = rw; out[nc].n = n; out[nc].dist = gap; nc++; }
    }
  }
  return nc;
}

// btPlaneSpace1
__device__ __forceinline__ void plane_space(f3 n, f3& p, f3& q) {
  if (fabsf(n.z) > 0.70710678118654752440f) {
    float a = n.y * n.y + n.z * n.z;
    float k = 1.0f / sqrtf(a);
    p = mk3(0.f, -n.z * k, n.y * k);
    q = mk3(a * k, -n.x * p.z, n.x * p.y);
  } else {
    float a = n.x * n.x + n.y * n.y;
    float k = 1.0f / sqrtf(a);
    p = mk3(-n.y * k, n.x * k, 0.f);
    q = mk3(-n.z * p.y, n.z * p.x, a * k);
  }
}

// ---------------------------------------------------------------- contact rows + projected Gauss-Seidel
// One normal row and two friction rows (implicit cone, mu = 0.25) per contact; <= 50 sweeps, normal rows
// first, early exit when the largest squared velocity change of a sweep is <= 1e-7; no warm starting.
// Restitution e_body * 0.5 applies only when |v_n| >= 0.2; penetration is recovered at erp 0.08.
#ifdef PH_ROLL_PGS
#define PH_PGS_ROLL _Pragma("unroll 1")
#else
#define PH_PGS_ROLL
#endif
template <int SHAPE>
__device__ __forceinline__ void solve_contacts(const Contact* c, int nc, float restitution, f3& v, f3& w) {
  const float inv_m = PH_INV_M;
  const float inv_i = (SHAPE == SHAPE_SPHERE) ? PH_INV_I_SPHERE : PH_INV_I_CUBE;
  const float crest = restitution * PH_STATIC_REST;
  f3 dir[PH_MAXC][3], rxn[PH_MAXC][3];
  float dinv[PH_MAXC][3], rhs[PH_MAXC][3], app[PH_MAXC][3];
  PH_PGS_ROLL
  for (int i = 0; i < nc; i++) {
    dir[i][0] = c[i].n;
    plane_space(c[i].n, dir[i][1], dir[i][2]);
    float relv[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      rxn[i][k] = cross3(c[i].r, dir[i][k]);
      dinv[i][k] = 1.0f / (inv_m + dot3(rxn[i][k], rxn[i][k]) * inv_i);
      relv[k] = dot3(dir[i][k], v) + dot3(rxn[i][k], w);
      app[i][k] = 0.f;
    }
    float rest = 0.f;
    if (fabsf(relv[0]) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * relv[0]);
    float distance = c[i].dist + PH_SLOP;
    float verr = rest - relv[0], perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
    rhs[i][0] = (perr + verr) * dinv[i][0];
    rhs[i][1] = -relv[1] * dinv[i][1];
    rhs[i][2] = -relv[2] * dinv[i][2];
  }
  f3 dv = mk3(0.f, 0.f, 0.f), dw = mk3(0.f, 0.f, 0.f);
  for (int it = 0; it < PH_ITERS; it++) {
    float resid = 0.f;
    PH_PGS_ROLL
    for (int i = 0; i < nc; i++) {
      float dimp = rhs[i][0] - (dot3(dir[i][0], dv) + dot3(rxn[i][0], dw)) * dinv[i][0];
      float sum = app[i][0] + dimp;
      if (sum < 0.f) { dimp = -app[i][0]; app[i][0] = 0.f; }
      else app[i][0] = sum;
      dv = dv + dir[i][0] * (dimp * inv_m);
      dw = dw + rxn[i][0] * (dimp * inv_i);
      float dvel = dimp / dinv[i][0];
      resid = fmaxf(resid, dvel * dvel);
    }
    PH_PGS_ROLL
    for (int i = 0; i < nc; i++) {
      float lim = PH_FRICTION * app[i][0];
      float sm[2];
#pragma unroll
      for (int k = 1; k <= 2; k++) {
        float dimp = rhs[i][k] - (dot3(dir[i][k], dv) + dot3(rxn[i][k], dw)) * dinv[i][k];
        sm[k - 1] = app[i][k] + dimp;
      }
      float l2 = sm[0] * sm[0] + sm[1] * sm[1];
      if (l2 > lim * lim) {
        float scl = lim / sqrtf(l2);
        sm[0] *= scl;
        sm[1] *= scl;
      }
#pragma unroll
      for (int k = 1; k <= 2; k++) {
        float dimp = sm[k - 1] - app[i][k];
        app[i][k] = sm[k - 1];
        dv = dv + dir[i][k] * (dimp * inv_m);
        dw = dw + rxn[i][k] * (dimp * inv_i);
        float dvel = dimp / dinv[i][k];
        resid = fmaxf(resid, dvel * dvel);
      }
    }
    if (resid <= PH_RESID) break;
  }
  v = v + dv;
  w = w + dw;
}

// ---------------------------------------------------------------- cube: plane-manifold contacts, rows reduced to arms
// The cube's most common contact set is 1-4 points of the persistent plane manifold and nothing else (resting or
// tumbling on the floor away from the walls).  All of them share n = +z, t1 = (0,-1,0), t2 = (1,0,0) (btPlaneSpace1
// of +z), so a row is its arm r alone: r x n = (r.y, -r.x, 0), r x t1 = (r.z, 0, -r.x), r x t2 = (0, r.z, -r.y).
// Same rows, same sweep order, same exit rule as solve_contacts - the terms multiplied by the zero components of
// n, t1, t2 are simply not computed: 13 floats per contact instead of the general solver's 27, a third of its
// arithmetic per row.
struct VelPair { f3 v, w; };
__device__ __noinline__ VelPair solve_plane_contacts(const Contact* c, int nc, float restitution, f3 v, f3 w) {
  const float inv_m = PH_INV_M, inv_i = PH_INV_I_CUBE;
  const float crest = restitution * PH_STATIC_REST;
  // The loops over the manifold slots stay rolled: unrolled over the four slots a sweep is 296 instructions against
  // 80 for the rolled pair of loops, and the kernel was bound by instruction fetch (ncu: no_instruction 38 % of the
  // stall samples, one per 128-byte line of straight-line code; 2.5 % rolled).  The rows then live in local memory.
  f3 r[4];
  float den[4][3], dinv[4][3], rhs[4][3], app[4][3];
#pragma unroll 1
  for (int i = 0; i < nc; i++) {
    r[i] = c[i].r;
    const float rx2 = r[i].x * r[i].x, ry2 = r[i].y * r[i].y, rz2 = r[i].z * r[i].z;
    den[i][0] = inv_m + (ry2 + rx2) * inv_i;
    den[i][1] = inv_m + (rz2 + rx2) * inv_i;
    den[i][2] = inv_m + (rz2 + ry2) * inv_i;
    const float relv0 = v.z + (r[i].y * w.x - r[i].x * w.y);
    const float relv1 = -v.y + (r[i].z * w.x - r[i].x * w.z);
    const float relv2 = v.x + (r[i].z * w.y - r[i].y * w.z);
    float rest = 0.f;
    if (fabsf(relv0) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * relv0);
    const float distance = c[i].dist + PH_SLOP;
    float verr = rest - relv0, perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
#pragma unroll
    for (int k = 0; k < 3; k++) { dinv[i][k] = 1.0f / den[i][k]; app[i][k] = 0.f; }
    rhs[i][0] = (perr + verr) * dinv[i][0];
    rhs[i][1] = -relv1 * dinv[i][1];
    rhs[i][2] = -relv2 * dinv[i][2];
  }
  f3 dv = mk3(0.f, 0.f, 0.f), dw = mk3(0.f, 0.f, 0.f);
  for (int it = 0; it < PH_ITERS; it++) {
    float resid = 0.f;
#pragma unroll 1
    for (int i = 0; i < nc; i++) {
      float dimp = rhs[i][0] - (dv.z + (r[i].y * dw.x - r[i].x * dw.y)) * dinv[i][0];
      const float sum = app[i][0] + dimp;
      if (sum < 0.f) { dimp = -app[i][0]; app[i][0] = 0.f; }
      else app[i][0] = sum;
      dv.z += dimp * inv_m;
      const float q = dimp * inv_i;
      dw.x += r[i].y * q;
      dw.y -= r[i].x * q;
      const float dvel = dimp * den[i][0];
      resid = fmaxf(resid, dvel * dvel);
    }
#pragma unroll 1
    for (int i = 0; i < nc; i++) {
      const float lim = PH_FRICTION * app[i][0];
      float s1 = app[i][1] + (rhs[i][1] - (-dv.y + (r[i].z * dw.x - r[i].x * dw.z)) * dinv[i][1]);
      float s2 = app[i][2] + (rhs[i][2] - (dv.x + (r[i].z * dw.y - r[i].y * dw.z)) * dinv[i][2]);
      const float l2 = s1 * s1 + s2 * s2;
      if (l2 > lim * lim) {
        const float scl = lim / sqrtf(l2);
        s1 *= scl;
        s2 *= scl;
      }
      const float d1 = s1 - app[i][1], d2 = s2 - app[i][2];
      app[i][1] = s1;
      app[i][2] = s2;
      dv.y -= d1 * inv_m;
      dv.x += d2 * inv_m;
      const float q1 = d1 * inv_i, q2 = d2 * inv_i;
      dw.x += r[i].z * q1;
      dw.y += r[i].z * q2;
      dw.z -= r[i].x * q1 + r[i].y * q2;
      const float e1 = d1 * den[i][1], e2 = d2 * den[i][2];
      resid = fmaxf(resid, fmaxf(e1 * e1, e2 * e2));
    }
    if (resid <= PH_RESID) break;
  }
  VelPair o;
  o.v = v + dv;
  o.w = w + dw;
  return o;
}

// The same solve with the rows of NC contacts in REGISTERS (every loop over the contacts unrolled): r, the three
// effective masses den and the right-hand sides / applied impulses, 12 floats per contact; 1 / den is re-taken by the
// approximate reciprocal where the rolled form reads its stored copy (same instruction, same bits).  The rolled form
// keeps its rows in local memory, and with 1.6 KB of stack per thread those loads miss L1: `long_scoreboard` was the
// top stall of the cube kernel and this loop half of its instructions (profiles/r02_cube_*).  Results are bit
// identical to solve_plane_contacts.  One instantiation per contact count, dispatched by a switch: a single unrolled
// body with every slot guarded by i < nc (so that lanes with different counts share instructions) was measured and is
// slower (lineage-controlled A/B of the t = 119 launch: 40-41.5 ms against 38.5 ms) - the lanes of a warp are few in here (5.9 of 32) because most
// contact sets are separating and leave after their first sweep, not because the counts differ.
template <int NC>
__device__ __forceinline__ VelPair solve_plane_reg(const Contact* c, float restitution, f3 v, f3 w) {
  const float inv_m = PH_INV_M, inv_i = PH_INV_I_CUBE;
  const float crest = restitution * PH_STATIC_REST;
  f3 r[NC];
  float den[NC][3], rhs[NC][3], app[NC][3];
#pragma unroll
  for (int i = 0; i < NC; i++) {
    r[i] = c[i].r;
    const float rx2 = r[i].x * r[i].x, ry2 = r[i].y * r[i].y, rz2 = r[i].z * r[i].z;
    den[i][0] = inv_m + (ry2 + rx2) * inv_i;
    den[i][1] = inv_m + (rz2 + rx2) * inv_i;
    den[i][2] = inv_m + (rz2 + ry2) * inv_i;
    const float relv0 = v.z + (r[i].y * w.x - r[i].x * w.y);
    const float relv1 = -v.y + (r[i].z * w.x - r[i].x * w.z);
    const float relv2 = v.x + (r[i].z * w.y - r[i].y * w.z);
    float rest = 0.f;
    if (fabsf(relv0) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * relv0);
    const float distance = c[i].dist + PH_SLOP;
    float verr = rest - relv0, perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
#pragma unroll
    for (int k = 0; k < 3; k++) app[i][k] = 0.f;
    rhs[i][0] = (perr + verr) * (1.0f / den[i][0]);
    rhs[i][1] = -relv1 * (1.0f / den[i][1]);
    rhs[i][2] = -relv2 * (1.0f / den[i][2]);
  }
  f3 dv = mk3(0.f, 0.f, 0.f), dw = mk3(0.f, 0.f, 0.f);
#pragma unroll 1
  for (int it = 0; it < PH_ITERS; it++) {
    float resid = 0.f;
#pragma unroll
    for (int i = 0; i < NC; i++) {
      float dimp = rhs[i][0] - (dv.z + (r[i].y * dw.x - r[i].x * dw.y)) * (1.0f / den[i][0]);
      const float sum = app[i][0] + dimp;
      if (sum < 0.f) { dimp = -app[i][0]; app[i][0] = 0.f; }
      else app[i][0] = sum;
      dv.z += dimp * inv_m;
      const float q = dimp * inv_i;
      dw.x += r[i].y * q;
      dw.y -= r[i].x * q;
      const float dvel = dimp * den[i][0];
      resid = fmaxf(resid, dvel * dvel);
    }
#pragma unroll
    for (int i = 0; i < NC; i++) {
      const float lim = PH_FRICTION * app[i][0];
      float s1 = app[i][1] + (rhs[i][1] - (-dv.y + (r[i].z * dw.x - r[i].x * dw.z)) * (1.0f / den[i][1]));
      float s2 = app[i][2] + (rhs[i][2] - (dv.x + (r[i].z * dw.y - r[i].y * dw.z)) * (1.0f / den[i][2]));
      const float l2 = s1 * s1 + s2 * s2;
      if (l2 > lim * lim) {
        const float scl = lim / sqrtf(l2);
        s1 *= scl;
        s2 *= scl;
      }
      const float d1 = s1 - app[i][1], d2 = s2 - app[i][2];
      app[i][1] = s1;
      app[i][2] = s2;
      dv.y -= d1 * inv_m;
      dv.x += d2 * inv_m;
      const float q1 = d1 * inv_i, q2 = d2 * inv_i;
      dw.x += r[i].z * q1;
      dw.y += r[i].z * q2;
      dw.z -= r[i].x * q1 + r[i].y * q2;
      const float e1 = d1 * den[i][1], e2 = d2 * den[i][2];
      resid = fmaxf(resid, fmaxf(e1 * e1, e2 * e2));
    }
    if (resid <= PH_RESID) break;
  }
  VelPair o;
  o.v = v + dv;
  o.w = w + dw;
  return o;
}
#ifndef PH_PLANE_ROLLED
__device__ __noinline__ VelPair solve_plane_contacts_reg(const Contact* c, int nc, float restitution, f3 v, f3 w) {
  switch (nc) {
    case 1: return solve_plane_reg<1>(c, restitution, v, w);
    case 2: return solve_plane_reg<2>(c, restitution, v, w);
    case 3: return solve_plane_reg<3>(c, restitution, v, w);
    default: return solve_plane_reg<4>(c, restitution, v, w);
  }
}
#endif

// general contact set of the cube (walls, ceiling, more than the plane manifold), out of line
__device__ __noinline__ VelPair solve_contacts_cube_call(const Contact* c, int nc, float restitution, f3 v, f3 w) {
  solve_contacts<SHAPE_CUBE>(c, nc, restitution, v, w);
  VelPair o;
  o.v = v;
  o.w = w;
  return o;
}

// ---------------------------------------------------------------- sphere: single-contact fast path
// For a sphere the contact arm is r = -h n, so r x n = 0 and the three rows of one contact (normal, two
// tangents) are mutually decoupled: J_n M^-1 J_t^T = n.t / m = 0 and J_t1 M^-1 J_t2^T = h^2 (n x t1).(n x t2) / I = 0.
// The projected Gauss-Seidel sweep therefore converges in its first iteration (the oracle's second sweep only
// confirms a residual at rounding level and exits), and the solve is closed form:
//   lambda_n = max(0, rhs_n),  (lambda_t1, lambda_t2) = cone-clamp(-relv_t / (1/m + h^2/I), mu lambda_n).
// Two or more simultaneous contacts (edges, corners of the arena) couple through friction and take the general
// solver, out of line.
struct SphereHit {
  int nc;
  f3 n;     // first contact
  float dist;
  f3 n2;    // second contact (valid when nc >= 2)
  float dist2;
};

__device__ __forceinline__ void sphere_vs_box(const f3& x, const f3& c, const f3& h, SphereHit& hit) {
  f3 cb = x - c;
  f3 hc = mk3(h.x - PH_BOX_MARGIN, h.y - PH_BOX_MARGIN, h.z - PH_BOX_MARGIN);
  f3 p = mk3(fminf(fmaxf(cb.x, -hc.x), hc.x), fminf(fmaxf(cb.y, -hc.y), hc.y), fminf(fmaxf(cb.z, -hc.z), hc.z));
  f3 df = cb - p;
  float d2 = dot3(df, df);
  f3 n;
  float dist;
  if (d2 > 0.f) {
    float dl = sqrtf(d2);
    n = df * (1.0f / dl);
    dist = dl - PH_H - PH_BOX_MARGIN;
  } else {
    int ax = 0;
    float best = h.x - fabsf(cb.x);
    float dy = h.y - fabsf(cb.y), dz = h.z - fabsf(cb.z);
    if (dy < best) { best = dy; ax = 1; }
    if (dz < best) { best = dz; ax = 2; }
    n = mk3(0.f, 0.f, 0.f);
    setc3(n, ax, comp3(cb, ax) >= 0.f ? 1.f : -1.f);
    dist = -(best + PH_H);
  }
  if (dist < PH_THR) {
    if (hit.nc == 0) { hit.n = n; hit.dist = dist; }
    else if (hit.nc == 1) { hit.n2 = n; hit.dist2 = dist; }
    hit.nc++;
  }
}

// Contacts of the sphere: count, and the contact itself when there is exactly one.  Only the wall on the
// body's side of each axis can be within reach (the opposite one is > 0.9 m away), so three box tests suffice.
__device__ __forceinline__ void sphere_hits(const ScenePar& sp, const f3& x, SphereHit& hit) {
  const float reach = PH_H + PH_THR;  // a box farther than this along any axis cannot touch
  hit.nc = 0;
  float d = x.z - PH_H;
  if (d < PH_THR) { hit.nc = 1; hit.n = mk3(0.f, 0.f, 1.f); hit.dist = d; }
  if (fabsf(x.x) > 0.5f - reach) sphere_vs_box(x, mk3(x.x >= 0.f ? 0.51f : -0.51f, 0.f, 0.51f), mk3(0.01f, 0.5f, 0.5f), hit);
  if (fabsf(x.y) > 0.5f - reach) sphere_vs_box(x, mk3(0.f, x.y >= 0.f ? 0.51f : -0.51f, 0.51f), mk3(0.5f, 0.01f, 0.5f), hit);
  if (x.z > 1.01f - sp.ceil_half - reach) sphere_vs_box(x, mk3(0.f, 0.f, 1.01f), mk3(0.5f, 0.5f, sp.ceil_half), hit);
}

// right-hand side of the normal row of one sphere contact (its effective mass is 1: r x n = 0, m = 1), i.e. the
// normal impulse the contact would get on its own
__device__ __forceinline__ float sphere_rhs_n(const f3& n, float dist, float restitution, const f3& v) {
  const float vn = dot3(n, v);
  float rest = 0.f;
  if (fabsf(vn) >= PH_REST_THRESH) rest = fmaxf(0.f, -(restitution * PH_STATIC_REST) * vn);
  const float distance = dist + PH_SLOP;
  float verr = rest - vn, perr = 0.f;
  if (distance > 0.f) verr -= distance / PH_DT;
  else perr = -distance * PH_ERP / PH_DT;
  return perr + verr;
}

// The two friction rows are written as one tangential impulse vector: with u = v + w x r the velocity of the contact
// point and u_t its tangential part, (l1, l2) = -(u.t1, u.t2) / 3.5 clamped to the cone is J = -u_t / 3.5 clamped in
// norm, for ANY orthonormal tangent basis - so btPlaneSpace1 is not needed, and floor, wall and edge contacts run
// the same instructions.  v += n lambda_n + J,  w += (r x J) / I.
__device__ __forceinline__ void sphere_solve_one(const SphereHit& hit, float restitution, f3& v, f3& w) {
  const float dinv_t = 1.0f / (PH_INV_M + PH_H * PH_H * PH_INV_I_SPHERE);  // 1 / 3.5
  const f3 n = hit.n;
  const f3 r = n * (-PH_H);
  const float vn = dot3(n, v);
  const f3 u = v + cross3(w, r);
  const f3 ut = u - n * dot3(n, u);
  float rest = 0.f;
  if (fabsf(vn) >= PH_REST_THRESH) rest = fmaxf(0.f, -(restitution * PH_STATIC_REST) * vn);
  const float distance = hit.dist + PH_SLOP;
  float verr = rest - vn, perr = 0.f;
  if (distance > 0.f) verr -= distance / PH_DT;
  else perr = -distance * PH_ERP / PH_DT;
  const float ln = fmaxf(0.f, perr + verr);  // 1 / (1/m) = 1
  const float lim = PH_FRICTION * ln;
  f3 J = ut * (-dinv_t);
  const float m2 = dot3(J, J);
  if (m2 > lim * lim) J = J * (lim / sqrtf(m2));
  v = v + n * ln + J;
  w = w + cross3(r, J) * PH_INV_I_SPHERE;
}

// Contact response that needs no iteration, in line: one contact (closed form above), and the two-contact cases
// the projected Gauss-Seidel solver leaves after one or two sweeps anyway (85-90 % of them in a filter run):
//  * neither normal row pushes (rhs_n <= 0 for both: contacts within the breaking threshold but separating):
//    sweep 1 clamps both normal impulses to 0, both friction cones are empty, the residual is 0 -> no impulse;
//  * exactly one pushes and the other stays idle: sweep 1 gives the active contact its single-contact closed form
//    (the idle contact's rows are all clamped to 0), sweep 2 repeats it to rounding and exits.  "Stays idle" is
//    checked exactly as the solver would: the idle normal row against the active normal impulse alone (sweep 1)
//    and against the full velocity change including friction (sweep 2).  If either check fails the velocities are
//    restored and the body goes to the iterative solver.
// The order of the two contacts does not matter in these cases.  Differences to the iterative result are at
// rounding level (< 1e-6 m/s, tools/diag in DESIGN.md section 3.1).  Returns true when the body still needs the
// iterative solver; otherwise v, w are final.
__device__ __forceinline__ bool sphere_contacts_inline(SphereHit& hit, float restitution, f3& v, f3& w) {
  bool many = hit.nc > 1, screen = false;
  float r_idle = 0.f;
  f3 n_idle = mk3(0.f, 0.f, 0.f);
  if (hit.nc == 2) {
    const float r0 = sphere_rhs_n(hit.n, hit.dist, restitution, v), r1 = sphere_rhs_n(hit.n2, hit.dist2, restitution, v);
    if (r0 <= 0.f && r1 <= 0.f) return false;
    if (r0 <= 0.f || r1 <= 0.f) {
      float lam;
      if (r0 > 0.f) { n_idle = hit.n2; r_idle = r1; lam = r0; }
      else { n_idle = hit.n; r_idle = r0; lam = r1; hit.n = hit.n2; hit.dist = hit.dist2; }
      screen = r_idle - dot3(n_idle, hit.n) * lam <= 0.f;
    }
  }
  const f3 v0 = v, w0 = w;
  if (hit.nc == 1 || screen) sphere_solve_one(hit, restitution, v, w);
  if (screen) {
    if (r_idle - dot3(n_idle, v - v0) <= 0.f) many = false;
    else { v = v0; w = w0; }
  }
  return many;
}

struct Body;
__device__ __forceinline__ void apply_forces(Body& s);
__device__ __forceinline__ void integrate_q(Body& s);
// ---------------------------------------------------------------- sphere: multi-contact solve (per thread, out of line)
// Contacts of the sphere in the order the oracle / Bullet visits them: plane, then the boxes in creation order
// (-x, +y, +x, -y, ceiling).  Only (n, dist) are kept; the arm is always -h n.
__device__ __forceinline__ int sphere_contacts_ordered(const ScenePar& sp, const f3& x, f3* n, float* dist) {
  const float reach = PH_H + PH_THR;
  int nc = 0;
  SphereHit h1;
  float d = x.z - PH_H;
  if (d < PH_THR) { n[nc] = mk3(0.f, 0.f, 1.f); dist[nc] = d; nc++; }
  const bool xpos = x.x >= 0.f, ypos = x.y >= 0.f;
  const bool nearx = fabsf(x.x) > 0.5f - reach, neary = fabsf(x.y) > 0.5f - reach;
  const bool yfirst = xpos && ypos;  // box1 (+y) precedes box2 (+x)
#pragma unroll
  for (int pass = 0; pass < 2; pass++) {
    const bool do_y = (pass == 0) == yfirst;
    if (do_y ? neary : nearx) {
      h1.nc = 0;
      if (do_y) sphere_vs_box(x, mk3(0.f, ypos ? 0.51f : -0.51f, 0.51f), mk3(0.5f, 0.01f, 0.5f), h1);
      else sphere_vs_box(x, mk3(xpos ? 0.51f : -0.51f, 0.f, 0.51f), mk3(0.01f, 0.5f, 0.5f), h1);
      if (h1.nc) { n[nc] = h1.n; dist[nc] = h1.dist; nc++; }
    }
  }
  if (x.z > 1.01f - sp.ceil_half - reach) {
    h1.nc = 0;
    sphere_vs_box(x, mk3(0.f, 0.f, 1.01f), mk3(0.5f, 0.5f, sp.ceil_half), h1);
    if (h1.nc) { n[nc] = h1.n; dist[nc] = h1.dist; nc++; }
  }
  return nc;
}

// Projected Gauss-Seidel for exactly two sphere contacts, all rows in registers.  Same sweep order and exit rule
// as solve_contacts (normal rows of both contacts, then the friction pairs; stop when the largest squared
// velocity change of a sweep is <= 1e-7; at most 50 sweeps).  Uses r = -h n: r x n = 0, r x t1 = -h t2,
// r x t2 = +h t1 for the btPlaneSpace1 basis (t2 = n x t1).
__device__ __forceinline__ void sphere_solve_two(const f3* nn, const float* dd, float restitution, f3& v, f3& w) {
  const float inv_i = PH_INV_I_SPHERE, h = PH_H;
  const float dinv_t = 1.0f / (PH_INV_M + PH_H * PH_H * PH_INV_I_SPHERE), kt = 1.0f / dinv_t;
  const float crest = restitution * PH_STATIC_REST;
  f3 n[2], t1[2], t2[2];
  float rhs[2][3], app[2][3];
#pragma unroll
  for (int i = 0; i < 2; i++) {
    n[i] = nn[i];
    plane_space(n[i], t1[i], t2[i]);
    const float vn = dot3(n[i], v);
    const float vt1 = dot3(t1[i], v) - h * dot3(t2[i], w), vt2 = dot3(t2[i], v) + h * dot3(t1[i], w);
    float rest = 0.f;
    if (fabsf(vn) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * vn);
    const float distance = dd[i] + PH_SLOP;
    float verr = rest - vn, perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
    rhs[i][0] = perr + verr;
    rhs[i][1] = -vt1 * dinv_t;
    rhs[i][2] = -vt2 * dinv_t;
    app[i][0] = app[i][1] = app[i][2] = 0.f;
  }
  f3 dv = mk3(0.f, 0.f, 0.f), dw = mk3(0.f, 0.f, 0.f);
  for (int it = 0; it < PH_ITERS; it++) {
    float resid = 0.f;
    const f3 pdv = dv, pdw = dw;
    float dn[2];
    bool steady = true;  // no projection touched this sweep and no friction row outgrows its cone
#pragma unroll
    for (int i = 0; i < 2; i++) {
      float dimp = rhs[i][0] - dot3(n[i], dv);
      const float sum = app[i][0] + dimp;
      if (sum < 0.f) { dimp = -app[i][0]; app[i][0] = 0.f; steady = false; }
      else app[i][0] = sum;
      dn[i] = dimp;
      dv = dv + n[i] * dimp;
      resid = fmaxf(resid, dimp * dimp);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float lim = PH_FRICTION * app[i][0];
      const float a1 = dot3(t1[i], dv) - h * dot3(t2[i], dw), a2 = dot3(t2[i], dv) + h * dot3(t1[i], dw);
      float s1 = app[i][1] + (rhs[i][1] - a1 * dinv_t), s2 = app[i][2] + (rhs[i][2] - a2 * dinv_t);
      const float l2 = s1 * s1 + s2 * s2;
      if (l2 > lim * lim) {
        const float scl = lim / sqrtf(l2);
        s1 *= scl;
        s2 *= scl;
        steady = false;
      }
      const float d1 = s1 - app[i][1], d2 = s2 - app[i][2];
      app[i][1] = s1;
      app[i][2] = s2;
      dv = dv + t1[i] * d1 + t2[i] * d2;
      dw = dw + (t1[i] * d2 - t2[i] * d1) * (h * inv_i);
      const float e1 = d1 * kt, e2 = d2 * kt;
      resid = fmaxf(resid, fmaxf(e1 * e1, e2 * e2));
      const float grow = PH_FRICTION * dn[i];
      steady = steady && (d1 * d1 + d2 * d2 <= grow * grow);
    }
    if (resid <= PH_RESID) break;
    // A body squeezed between (nearly) opposing faces - teleported under a wall, into the 1 cm gap between the
    // wall boxes and the floor - has no feasible solution.  The reference keeps sweeping to its 50-sweep cap:
    // the impulses of the conflicting rows grow linearly for ever while (dv, dw), the only thing that leaves
    // the solver, converged geometrically (ratio ~0.2) within ~10 sweeps.  An unprojected sweep is a fixed
    // affine map of (dv, dw) alone, so once a sweep (a) leaves (dv, dw) where they were, (b) projected nothing
    // (no normal impulse clamped at 0, no friction pair scaled back to its cone) and (c) grows every friction
    // pair no faster than its cone mu * lambda_n grows, every later sweep repeats it exactly: stop.  A friction
    // row that outgrows its cone (floor friction holding against a wall's push) will saturate in a later sweep
    // and release the body, so that case keeps sweeping like the reference.
    if (steady) {
      const f3 ddv = dv - pdv, ddw = dw - pdw;
      const float mv = fmaxf(fmaxf(fabsf(ddv.x), fabsf(ddv.y)), fabsf(ddv.z));
      const float mw = fmaxf(fmaxf(fabsf(ddw.x), fabsf(ddw.y)), fabsf(ddw.z)) * PH_H;
      if (fmaxf(mv, mw) < PH_STALL) break;
    }
  }
  v = v + dv;
  w = w + dw;
}

// The two-contact set a filter run actually meets is the floor plus one wall face (a ball rolling along a wall or
// wedged in the 1 cm slot under it): n0 = +z and n1 = s e_x or s e_y, whose btPlaneSpace1 tangents are signed unit
// axes as well (floor: t1 = -e_y, t2 = e_x; wall x: t1 = s e_y, t2 = e_z; wall y: t1 = -s e_x, t2 = e_z).  Every dot
// product of sphere_solve_two then picks one component, so the same sweep - same rows, order, exit rules - is
// written on scalars: ~40 % fewer instructions per sweep.  The wall normal is taken as exactly +-1 (the narrow
// phase normalises it with an approximate reciprocal; the oracle's float64 division gives exactly +-1).
template <int AX>
__device__ __forceinline__ void sphere_solve_floor_wall(float s, float d_floor, float d_wall, float restitution, f3& v,
                                                        f3& w) {
  const float h = PH_H, c = PH_H * PH_INV_I_SPHERE;
  const float dinv_t = 1.0f / (PH_INV_M + PH_H * PH_H * PH_INV_I_SPHERE), kt = 1.0f / dinv_t;
  const float crest = restitution * PH_STATIC_REST;
  float rhs[2][3], app[2][3];
#pragma unroll
  for (int i = 0; i < 2; i++) {
    float vn, vt1, vt2;
    if (i == 0) { vn = v.z; vt1 = -v.y - h * w.x; vt2 = v.x - h * w.y; }
    else if (AX == 0) { vn = s * v.x; vt1 = s * v.y - h * w.z; vt2 = v.z + h * (s * w.y); }
    else { vn = s * v.y; vt1 = -(s * v.x) - h * w.z; vt2 = v.z - h * (s * w.x); }
    float rest = 0.f;
    if (fabsf(vn) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * vn);
    const float distance = (i == 0 ? d_floor : d_wall) + PH_SLOP;
    float verr = rest - vn, perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
    rhs[i][0] = perr + verr;
    rhs[i][1] = -vt1 * dinv_t;
    rhs[i][2] = -vt2 * dinv_t;
    app[i][0] = app[i][1] = app[i][2] = 0.f;
  }
  f3 dv = mk3(0.f, 0.f, 0.f), dw = mk3(0.f, 0.f, 0.f);
  for (int it = 0; it < PH_ITERS; it++) {
    float resid = 0.f;
    const f3 pdv = dv, pdw = dw;
    float dn[2];
    bool steady = true;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float ndv = (i == 0) ? dv.z : s * (AX == 0 ? dv.x : dv.y);
      float dimp = rhs[i][0] - ndv;
      const float sum = app[i][0] + dimp;
      if (sum < 0.f) { dimp = -app[i][0]; app[i][0] = 0.f; steady = false; }
      else app[i][0] = sum;
      dn[i] = dimp;
      if (i == 0) dv.z += dimp;
      else if (AX == 0) dv.x += s * dimp;
      else dv.y += s * dimp;
      resid = fmaxf(resid, dimp * dimp);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float lim = PH_FRICTION * app[i][0];
      float a1, a2;
      if (i == 0) { a1 = -dv.y - h * dw.x; a2 = dv.x - h * dw.y; }
      else if (AX == 0) { a1 = s * dv.y - h * dw.z; a2 = dv.z + h * (s * dw.y); }
      else { a1 = -(s * dv.x) - h * dw.z; a2 = dv.z - h * (s * dw.x); }
      float s1 = app[i][1] + (rhs[i][1] - a1 * dinv_t), s2 = app[i][2] + (rhs[i][2] - a2 * dinv_t);
      const float l2 = s1 * s1 + s2 * s2;
      if (l2 > lim * lim) {
        const float scl = lim / sqrtf(l2);
        s1 *= scl;
        s2 *= scl;
        steady = false;
      }
      const float d1 = s1 - app[i][1], d2 = s2 - app[i][2];
      app[i][1] = s1;
      app[i][2] = s2;
      if (i == 0) { dv.y -= d1; dv.x += d2; dw.x -= d1 * c; dw.y -= d2 * c; }
      else if (AX == 0) { dv.y += s * d1; dv.z += d2; dw.y += (s * d2) * c; dw.z -= d1 * c; }
      else { dv.x -= s * d1; dv.z += d2; dw.x -= (s * d2) * c; dw.z -= d1 * c; }
      const float e1 = d1 * kt, e2 = d2 * kt;
      resid = fmaxf(resid, fmaxf(e1 * e1, e2 * e2));
      const float grow = PH_FRICTION * dn[i];
      steady = steady && (d1 * d1 + d2 * d2 <= grow * grow);
    }
    if (resid <= PH_RESID) break;
    if (steady) {  // the steady-state exit of sphere_solve_two
      const f3 ddv = dv - pdv, ddw = dw - pdw;
      const float mv = fmaxf(fmaxf(fabsf(ddv.x), fabsf(ddv.y)), fabsf(ddv.z));
      const float mw = fmaxf(fmaxf(fabsf(ddw.x), fabsf(ddw.y)), fabsf(ddw.z)) * PH_H;
      if (fmaxf(mv, mw) < PH_STALL) break;
    }
  }
  v = v + dv;
  w = w + dw;
}

// three or four simultaneous contacts (arena corners): general solver, out of line
__device__ __noinline__ void sphere_solve_generic(const f3* n, const float* dist, int nc, float restitution, f3* v,
                                                  f3* w) {
  Contact c[PH_MAXC];
  for (int i = 0; i < nc; i++) { c[i].n = n[i]; c[i].r = n[i] * (-PH_H); c[i].dist = dist[i]; }
  f3 vv = *v, ww = *w;
  solve_contacts<SHAPE_SPHERE>(c, nc, restitution, vv, ww);
  *v = vv;
  *w = ww;
}

// Contacts that do need the iterative solver (10-15 % of the two-contact substeps of a filter run, and every
// corner): out of line, so that the registers of the hot loop are sized by the free-flight / single-contact path.
// Everything travels by value: taking the address of the caller's Body would pin it to local memory.
__device__ __noinline__ VelPair sphere_solve_iter(float ceil_half, float restitution, f3 x, f3 v, f3 w) {
  ScenePar sp;
  sp.ceil_half = ceil_half;
  f3 n[4];
  float dist[4];
  const int nc = sphere_contacts_ordered(sp, x, n, dist);
  if (nc == 2) {
    // floor first (sphere_contacts_ordered), then a wall face: its normal has one non-zero component, x or y.
    // Which form runs depends on the body's own contacts only, never on its neighbours in the warp: the two forms
    // agree to rounding, not to the bit (measured: voting on the warp's composition made results lane dependent).
    const bool floor_wall = n[0].z == 1.0f && n[1].z == 0.f && (n[1].x == 0.f) != (n[1].y == 0.f);
    if (floor_wall && n[1].y == 0.f)
      sphere_solve_floor_wall<0>(copysignf(1.0f, n[1].x), dist[0], dist[1], restitution, v, w);
    else if (floor_wall)
      sphere_solve_floor_wall<1>(copysignf(1.0f, n[1].y), dist[0], dist[1], restitution, v, w);
    else
      sphere_solve_two(n, dist, restitution, v, w);
  } else if (nc > 0) sphere_solve_generic(n, dist, nc, restitution, &v, &w);
  VelPair r;
  r.v = v;
  r.w = w;
  return r;
}

// ---------------------------------------------------------------- one 1/240 s substep (Bullet stepSimulation)
// Order (pinned by the reference's PyBullet cube fixture, see oracle/smc_oracle.hpp): narrow phase at the
// current pose; v += dt (g - k v (1 + |v|)), same drag on omega; contact solve; x += dt v; q <- exp(w dt/2) q.
// TRACK_Q = false drops the orientation update: a sphere's orientation never feeds back into its motion,
// so the particle-filter kernels do not carry it (PhySMC.forward_step on a batch does).
__device__ __forceinline__ void apply_forces(Body& s) {
  const float dt = PH_DT;
  float vn = sqrtf(dot3(s.v, s.v));
  float f = PH_DAMP * (1.0f + vn);
  s.v = mk3(s.v.x + dt * (-f * s.v.x), s.v.y + dt * (-f * s.v.y), s.v.z + dt * (PH_G - f * s.v.z));
  float wn = sqrtf(dot3(s.w, s.w));
  float fw = PH_DAMP * (1.0f + wn);
  s.w = mk3(s.w.x + dt * (-fw * s.w.x), s.w.y + dt * (-fw * s.w.y), s.w.z + dt * (-fw * s.w.z));
}

__device__ __forceinline__ void integrate_q(Body& s) {
  const float dt = PH_DT;
  float ang = sqrtf(dot3(s.w, s.w));
  if (ang * dt > PH_ANG_THRESH) ang = 0.5f * 1.57079632679489661923f / dt;
  // half angle hh <= pi/8 after the clamp: sin(hh)/ang = (dt/2) sinc(hh) and cos(hh) as even polynomials in hh
  // (truncation < 1e-11, i.e. below float rounding; Bullet's own small-angle branch is the first two terms), which
  // is branch free and keeps libm's range-reduction slow paths out of the kernels' instruction footprint
  const float hh = 0.5f * ang * dt, h2 = hh * hh;
  const float sinc = 1.0f + h2 * (-1.0f / 6.0f + h2 * (1.0f / 120.0f + h2 * (-1.0f / 5040.0f + h2 * (1.0f / 362880.0f))));
  const float cw = 1.0f + h2 * (-0.5f + h2 * (1.0f / 24.0f + h2 * (-1.0f / 720.0f + h2 * (1.0f / 40320.0f))));
  const f3 ax = s.w * (0.5f * dt * sinc);
  f4 q = s.q, r;
  r.x = cw * q.x + ax.x * q.w + ax.y * q.z - ax.z * q.y;
  r.y = cw * q.y - ax.x * q.z + ax.y * q.w + ax.z * q.x;
  r.z = cw * q.z + ax.x * q.y - ax.y * q.x + ax.z * q.w;
  r.w = cw * q.w - ax.x * q.x - ax.y * q.y - ax.z * q.z;
  float inv = 1.0f / sqrtf(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w);
  s.q.x = r.x * inv; s.q.y = r.y * inv; s.q.z = r.z * inv; s.q.w = r.w * inv;
}

// Sphere on the floor, away from every box: sphere_solve_one for n = +z, r = (0, 0, -h), on scalars.  With these n, r
// the contact-point velocity is u = (v.x - h w.y, v.y + h w.x, v.z), its tangential part (u.x, u.y, 0), and
// r x J = (h J.y, -h J.x, 0): the same closed form in a third of the instructions.  This is the contact a filter run
// meets in most substeps (the ball rests or rolls on the floor from frame ~36 on and is teleported into it by the
// transition noise every other step).
__device__ __forceinline__ void sphere_solve_floor(float d, float restitution, f3& v, f3& w) {
  const float dinv_t = 1.0f / (PH_INV_M + PH_H * PH_H * PH_INV_I_SPHERE);  // 1 / 3.5
  const float vn = v.z;
  float rest = 0.f;
  if (fabsf(vn) >= PH_REST_THRESH) rest = fmaxf(0.f, -(restitution * PH_STATIC_REST) * vn);
  const float distance = d + PH_SLOP;
  float verr = rest - vn, perr = 0.f;
  if (distance > 0.f) verr -= distance / PH_DT;
  else perr = -distance * PH_ERP / PH_DT;
  const float ln = fmaxf(0.f, perr + verr);
  const float lim = PH_FRICTION * ln;
  float jx = (v.x - PH_H * w.y) * (-dinv_t), jy = (v.y + PH_H * w.x) * (-dinv_t);
  const float m2 = jx * jx + jy * jy;
  if (m2 > lim * lim) {
    const float scl = lim / sqrtf(m2);
    jx *= scl;
    jy *= scl;
  }
  v = mk3(v.x + jx, v.y + jy, v.z + ln);
  w = mk3(w.x + (PH_H * PH_INV_I_SPHERE) * jy, w.y - (PH_H * PH_INV_I_SPHERE) * jx, w.z);
}

// Sphere within reach of a wall or the ceiling (possibly the floor as well): the general narrow phase, the in-line
// screening and, when needed, the iterative solve.  Out of line, and everything the substep loop keeps live travels
// THROUGH the call by value (position and the chain's loop state included): nothing of the caller is live across it,
// so the calling convention gives the register allocator no reason to park the hot loop's values in local memory
// (with v, w alone returned it spilled the position and the loop counters for every substep of every lane).
#ifdef RS_NEAR_OUTLINE
#define RS_NEAR_ATTR __noinline__
#else
#define RS_NEAR_ATTR __forceinline__
#endif
struct SphereChain {
  f3 x, v, w;
  float restitution, ceil_half;
  float zc;         // centre height above which the ceiling is within reach
  f3 nx;            // the next retained position, loaded one macro step ahead
  const float* pk;  // row pointer of the caller's loop
  long long n;      // row stride
  double qs;        // accumulator of the caller
  int k, sub, S;    // loop counters of the caller
};
__device__ __noinline__ SphereChain sphere_contacts_near_box(SphereChain s) {
  ScenePar sp;
  sp.ceil_half = s.ceil_half;
  SphereHit hit;
  sphere_hits(sp, s.x, hit);
  if (hit.nc > 0 && sphere_contacts_inline(hit, s.restitution, s.v, s.w)) {
    const VelPair r = sphere_solve_iter(s.ceil_half, s.restitution, s.x, s.v, s.w);
    s.v = r.v;
    s.w = r.w;
  }
  return s;
}

// One face contact of the sphere in the contact's own frame: (n, a, b) a cyclic permutation of (x, y, z), normal
// s e_n with s = +-1, arm r = -h s e_n.  sphere_solve_one on scalars: u_t = (v_a - h s w_b, v_b + h s w_a),
// r x J = (0, h s J_b, -h s J_a).  Floor: (n, a, b) = (z, x, y), s = +1; x walls: (x, y, z); y walls: (y, z, x).
__device__ __forceinline__ float face_rhs(float s, float dist, float crest, float vn) {
  const float vnn = s * vn;
  float rest = 0.f;
  if (fabsf(vnn) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * vnn);
  const float distance = dist + PH_SLOP;
  float verr = rest - vnn, perr = 0.f;
  if (distance > 0.f) verr -= distance / PH_DT;
  else perr = -distance * PH_ERP / PH_DT;
  return perr + verr;
}
__device__ __forceinline__ void face_solve(float s, float rhs_n, float& vn, float& va, float& vb, float& wa, float& wb) {
  const float dinv_t = 1.0f / (PH_INV_M + PH_H * PH_H * PH_INV_I_SPHERE);  // 1 / 3.5
  const float ln = fmaxf(0.f, rhs_n);
  const float lim = PH_FRICTION * ln;
  const float hs = PH_H * s;
  float ja = (va - hs * wb) * (-dinv_t), jb = (vb + hs * wa) * (-dinv_t);
  const float m2 = ja * ja + jb * jb;
  if (m2 > lim * lim) {
    const float scl = lim / sqrtf(m2);
    ja *= scl;
    jb *= scl;
  }
  vn += s * ln;
  va += ja;
  vb += jb;
  const float k = (PH_H * PH_INV_I_SPHERE) * s;
  wa += k * jb;
  wb -= k * ja;
}

// The near-box tiers of sphere_chain_substep, optionally out of line (-DRS_NEAR_OUTLINE): the substep loop then holds
// free flight and the floor only.  Pose-derived values are recomputed here; velocities come after the forces.
__device__ RS_NEAR_ATTR SphereChain sphere_near_box_tiers(SphereChain c) {
  const float reach = PH_H + PH_THR;
  const float crest = c.restitution * PH_STATIC_REST;
  const float ax = fabsf(c.x.x), ay = fabsf(c.x.y);
  const bool nearx = ax > 0.5f - reach, neary = ay > 0.5f - reach;
  const bool nearc = c.x.z > c.zc;
  const float d = c.x.z - PH_H;
  {
    bool slow = true;
    const float aw = nearx ? ax : ay;  // distance of the centre from the arena centre across the wall
    // face region of exactly one wall: closest point of the wall's core box (half thickness 0.01 - margin, spanning
    // z in [0.011, 1.009]) lies on its inner face and the centre is outside the core
    if (nearx != neary && !nearc && c.x.z >= 0.011f && aw < 0.501f) {
      const float cb = aw - 0.51f;                                // sphere_vs_box on the mirrored coordinate
      const float dl = -(cb + (0.01f - PH_BOX_MARGIN));
      const float dist = dl - PH_H - PH_BOX_MARGIN;
      const float sw = (nearx ? c.x.x : c.x.y) >= 0.f ? -1.0f : 1.0f;  // normal: from the wall into the arena
      // the wall's frame (n, a, b): x wall (x, y, z), y wall (y, z, x)
      float vn = nearx ? c.v.x : c.v.y, va = nearx ? c.v.y : c.v.z, vb = nearx ? c.v.z : c.v.x;
      float wa = nearx ? c.w.y : c.w.z, wb = nearx ? c.w.z : c.w.x;
      const bool hit_w = dist < PH_THR, hit_f = d < PH_THR;
      bool iterate = false;  // floor + wall face whose rows couple: the scalar projected Gauss-Seidel, in line
      slow = false;
      if (hit_w && hit_f) {
        const float r0 = face_rhs(1.0f, d, crest, c.v.z), r1 = face_rhs(sw, dist, crest, vn);
        if (r0 > 0.f && r1 > 0.f) iterate = true;
        else if (r0 > 0.f) {
          // floor active, wall idle: the wall's normal row must stay idle under the floor's full velocity change
          const f3 v0 = c.v, w0 = c.w;
          face_solve(1.0f, r0, c.v.z, c.v.x, c.v.y, c.w.x, c.w.y);
          const float dvn = (nearx ? c.v.x - v0.x : c.v.y - v0.y) * sw;
          if (r1 - dvn > 0.f) { c.v = v0; c.w = w0; iterate = true; }
        } else if (r1 > 0.f) {
          const float vz0 = c.v.z;
          face_solve(sw, r1, vn, va, vb, wa, wb);
          // in the wall's frame the floor normal is b (x wall) or a (y wall)
          const float vz1 = nearx ? vb : va;
          if (r0 - (vz1 - vz0) > 0.f) iterate = true;
          else {
            if (nearx) { c.v = mk3(vn, va, vb); c.w = mk3(c.w.x, wa, wb); }
            else { c.v = mk3(vb, vn, va); c.w = mk3(wb, c.w.y, wa); }
          }
        }
      } else if (hit_w) {
        face_solve(sw, face_rhs(sw, dist, crest, vn), vn, va, vb, wa, wb);
        if (nearx) { c.v = mk3(vn, va, vb); c.w = mk3(c.w.x, wa, wb); }
        else { c.v = mk3(vb, vn, va); c.w = mk3(wb, c.w.y, wa); }
      } else if (hit_f) {
        face_solve(1.0f, face_rhs(1.0f, d, crest, c.v.z), c.v.z, c.v.x, c.v.y, c.w.x, c.w.y);
      }
      // A ball resting in the corner of the floor and a wall is pushed by both (both normal rows active): the state
      // of the whole population for the ~10 steps in which the benchmark's ball rolls to rest against a wall, and
      // 60 % of a late launch when it took the call (narrow phase twice, screening, dispatch: ~350 instructions
      // around ~7 sweeps of 120).
      if (iterate) {
        if (nearx) sphere_solve_floor_wall<0>(sw, d, dist, c.restitution, c.v, c.w);
        else sphere_solve_floor_wall<1>(sw, d, dist, c.restitution, c.v, c.w);
      }
    }
    if (slow) c = sphere_contacts_near_box(c);
  }
  return c;
}

// One 1/240 s substep of the sphere on the chain state (orientation not tracked: it never feeds back into the motion).
// In line, on scalars: free flight, the floor, the face of ONE wall (its face region: the centre between the wall's
// edges, outside its core), and floor + wall face when the solver would not iterate (sphere_contacts_inline's
// screening: neither normal row pushes / one pushes and the other stays idle).  These are 80 % of the substeps spent
// within reach of a box in a filter run and all of those away from the boxes; the rest - wall edges, a centre inside
// a wall, two walls, the ceiling, pairs that iterate - takes the call, with the whole chain state passed through it.
__device__ __forceinline__ void sphere_chain_substep(SphereChain& c) {
  const float reach = PH_H + PH_THR;
  const float crest = c.restitution * PH_STATIC_REST;
  const float ax = fabsf(c.x.x), ay = fabsf(c.x.y);
  const bool nearx = ax > 0.5f - reach, neary = ay > 0.5f - reach;
  const bool nearc = c.x.z > c.zc;  // zc = 1.01 - ceil_half - reach, kept in the chain state
  const float d = c.x.z - PH_H;
  {
    // v += dt (g - k v (1 + |v|))  written as  v <- v (1 - dt k (1 + |v|)) + dt g: one multiplier per vector instead of
    // a product and an FMA per component (7 instructions less per substep); the same number up to the last bit
    const float a = 1.0f - PH_DT * PH_DAMP, b = -(PH_DT * PH_DAMP);
    const float gv = fmaf(sqrtf(dot3(c.v, c.v)), b, a);
    c.v = mk3(c.v.x * gv, c.v.y * gv, fmaf(c.v.z, gv, PH_DT * PH_G));
    const float gw = fmaf(sqrtf(dot3(c.w, c.w)), b, a);
    c.w = mk3(c.w.x * gw, c.w.y * gw, c.w.z * gw);
  }
  if (!(nearx || neary || nearc)) {
    if (d < PH_THR) face_solve(1.0f, face_rhs(1.0f, d, crest, c.v.z), c.v.z, c.v.x, c.v.y, c.w.x, c.w.y);
  } else {
    c = sphere_near_box_tiers(c);
  }
  c.x = c.x + c.v * PH_DT;
}

// true when at least one normal row of the contact set has a positive right-hand side (see substep)
__device__ __forceinline__ bool contacts_push(const Contact* c, int nc, float crest, const f3& v, const f3& w) {
  bool push = false;
#pragma unroll 1
  for (int i = 0; i < nc; i++) {
    const f3 n = c[i].n, r = c[i].r;
    const float relv = dot3(n, v) + dot3(cross3(r, n), w);
    float rest = 0.f;
    if (fabsf(relv) >= PH_REST_THRESH) rest = fmaxf(0.f, -crest * relv);
    const float distance = c[i].dist + PH_SLOP;
    float verr = rest - relv, perr = 0.f;
    if (distance > 0.f) verr -= distance / PH_DT;
    else perr = -distance * PH_ERP / PH_DT;
    push = push || (perr + verr > 0.f);
  }
  return push;
}

template <int SHAPE, bool TRACK_Q>
__device__ __forceinline__ void substep(const ScenePar& sp, float restitution, Body& s, Manifold& pm) {
  const float dt = PH_DT;
  if (SHAPE == SHAPE_SPHERE) {
    // the narrow phase depends on the pose only: taken before the forces, like the oracle
    const float reach = PH_H + PH_THR;
    const bool near_box = fmaxf(fabsf(s.x.x), fabsf(s.x.y)) > 0.5f - reach || s.x.z > 1.01f - sp.ceil_half - reach;
    const float d = s.x.z - PH_H;
    apply_forces(s);
    if (near_box) {
      SphereChain a;
      a.x = s.x; a.v = s.v; a.w = s.w; a.restitution = restitution; a.ceil_half = sp.ceil_half;
      a.zc = 0.f; a.pk = nullptr; a.n = 0; a.qs = 0.0; a.k = a.sub = a.S = 0;
      a = sphere_contacts_near_box(a);
      s.v = a.v;
      s.w = a.w;
    } else if (d < PH_THR) {
      sphere_solve_floor(d, restitution, s.v, s.w);
    }
  } else {
    const bool clear = body_clear<SHAPE>(s.x, sp.ceil_half);
    Contact c[PH_MAXC];
    int nc = 0;
    if (!clear) nc = collide_cube(sp, s, pm, c);
    else pm.n = 0;
    const int npl = pm.n;  // the plane-manifold points come first in c[]
    apply_forces(s);
#ifndef PH_NO_PUSH_SCREEN
    // No normal row pushes (every contact is within the breaking threshold but separating, or hovering inside the
    // margin): sweep 1 of the solver clamps every normal impulse to 0, every friction cone is then empty, the residual
    // is 0 and it returns the velocities unchanged - exactly, for any contact set.  Checked here on the normal rows
    // alone (10 instructions per contact instead of the ~60 of a row set-up + sweep); this is most of the contact
    // substeps of a cube at rest (the cold manifold of every macro step re-collects its points one per substep).
    if (nc > 0 && !contacts_push(c, nc, restitution * PH_STATIC_REST, s.v, s.w)) nc = 0;
#endif
    if (nc > 0) {
#ifdef PH_PLANE_ROLLED
      const VelPair r = (nc == npl) ? solve_plane_contacts(c, nc, restitution, s.v, s.w)
                                    : solve_contacts_cube_call(c, nc, restitution, s.v, s.w);
#else
      const VelPair r = (nc == npl) ? solve_plane_contacts_reg(c, nc, restitution, s.v, s.w)
                                    : solve_contacts_cube_call(c, nc, restitution, s.v, s.w);
#endif
      s.v = r.v;
      s.w = r.w;
    }
  }
  s.x = s.x + s.v * dt;
  if (TRACK_Q || SHAPE == SHAPE_CUBE) integrate_q(s);
}

// PhySMC.step = sync! ; forward_step (PhySMC.jl:32-35): the body is re-posed, so the contact cache starts
// cold, then S substeps run (S = 5 at step_dur 1/60, 9 at 1/30: `while dt <= step_dur`, PhyBullet.jl:82-88).
template <int SHAPE, bool TRACK_Q>
__device__ __forceinline__ void macro_step(const ScenePar& sp, float restitution, int S, Body& s) {
  Manifold pm;
  pm.n = 0;
  for (int k = 0; k < S; k++) substep<SHAPE, TRACK_Q>(sp, restitution, s, pm);
}

}  // namespace smcphys
