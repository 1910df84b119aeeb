// geom.cuh — per-lane narrow-phase collision for the batched Adroit simulator (sm_100a).
//
// One lane handles one candidate geom pair that survived the bounding-sphere cull.  All routines work in a
// frame translated so that geom1's centre is the origin (keeps fp32 magnitudes at link scale); the caller adds the
// offset back.  Semantics follow MuJoCo 2.1's collision table for the primitives the Adroit scenes contain
// (SURVEY.md App. A.4 / App. B): analytic plane-X, sphere/capsule/box combinations, and a Minkowski-portal-
// refinement query on margin-inflated shapes for every pair that involves a cylinder (single contact).
#pragma once
#include <cuda_runtime.h>

namespace adroit {

enum { G_PLANE = 0, G_SPHERE = 2, G_CAPSULE = 3, G_CYLINDER = 5, G_BOX = 6 };

template <typename T> struct V3 { T x, y, z; };
#define DI __device__ __forceinline__
// phase functions are inlined into their phase kernel (one call site each): address spaces and model-table loads
// resolve statically (LDS / LDG / LDC instead of generic loads); SUB_FN marks helpers with several call sites
#define PHASE_FN __device__ __forceinline__
#define SUB_FN __device__ __noinline__

template <typename T> DI V3<T> mk(T x, T y, T z) { V3<T> r; r.x = x; r.y = y; r.z = z; return r; }
template <typename T> DI V3<T> operator+(V3<T> a, V3<T> b) { return mk<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> DI V3<T> operator-(V3<T> a, V3<T> b) { return mk<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> DI V3<T> operator*(V3<T> a, T s) { return mk<T>(a.x * s, a.y * s, a.z * s); }
template <typename T> DI V3<T> neg(V3<T> a) { return mk<T>(-a.x, -a.y, -a.z); }
template <typename T> DI T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> DI V3<T> cross(V3<T> a, V3<T> b) { return mk<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
template <typename T> DI T norm(V3<T> a) { return sqrt(dot(a, a)); }
template <typename T> DI T comp(V3<T> a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
template <typename T> DI void setc(V3<T>& a, int i, T v) { if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v; }
template <typename T> DI T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> DI T tmax(T a, T b) { return a > b ? a : b; }
template <typename T> DI T tclamp(T x, T lo, T hi) { return x < lo ? lo : (x > hi ? hi : x); }
template <typename T> DI T tabs(T x) { return x < 0 ? -x : x; }
template <typename T> DI T tiny();
template <> DI float tiny<float>() { return 1e-30f; }
template <> DI double tiny<double>() { return 1e-300; }
template <typename T> DI T eps_rel();   // "is zero" threshold for unit-scale quantities
template <> DI float eps_rel<float>() { return 1e-7f; }
template <> DI double eps_rel<double>() { return 1e-14; }

template <typename T> DI V3<T> normalized(V3<T> a) {
  T n = norm(a);
  if (n < tiny<T>()) return mk<T>(1, 0, 0);
  return a * (T(1) / n);
}
DI double inv_sqrt(double x) { return rsqrt(x); }
DI float inv_sqrt(float x) { return rsqrtf(x); }
// fp64 (the MPR portal refinement, a serial chain the whole block waits for): one reciprocal square root instead of a square root
// and a division; differs from 1/sqrt by an ulp or two
template <> DI V3<double> normalized<double>(V3<double> a) {
  double n2 = dot(a, a);
  if (n2 < tiny<double>()) return mk<double>(1, 0, 0);
  return a * rsqrt(n2);
}

// 3x3 rotation, row-major; columns are the body axes in the outer frame
template <typename T> struct M3 { T m[9]; };
template <typename T> DI V3<T> mulv(const M3<T>& R, V3<T> v) {
  return mk<T>(R.m[0] * v.x + R.m[1] * v.y + R.m[2] * v.z, R.m[3] * v.x + R.m[4] * v.y + R.m[5] * v.z,
               R.m[6] * v.x + R.m[7] * v.y + R.m[8] * v.z);
}
template <typename T> DI V3<T> multv(const M3<T>& R, V3<T> v) {
  return mk<T>(R.m[0] * v.x + R.m[3] * v.y + R.m[6] * v.z, R.m[1] * v.x + R.m[4] * v.y + R.m[7] * v.z,
               R.m[2] * v.x + R.m[5] * v.y + R.m[8] * v.z);
}
template <typename T> DI V3<T> col(const M3<T>& R, int c) { return mk<T>(R.m[c], R.m[3 + c], R.m[6 + c]); }

// contact candidates produced by one pair (at most 8: box-box face clipping)
template <typename T> struct ConOut {
  int n;
  T dist[8];
  V3<T> pos[8];
  V3<T> nrm[8];
  DI void add(T d, V3<T> p, V3<T> nn, T margin) {
    if (d < margin && n < 8) { dist[n] = d; pos[n] = p; nrm[n] = nn; n++; }
  }
};

// ---------------------------------------------------------------- sphere-type primitives
template <typename T> DI void c_plane_sphere(ConOut<T>& o, T margin, V3<T> pn, V3<T> c /*rel. plane point*/, T r) {
  T d = dot(c, pn) - r;
  o.add(d, c - pn * (r + T(0.5) * d), pn, margin);
}
template <typename T> DI void c_sphere_sphere(ConOut<T>& o, T margin, V3<T> c1, T r1, V3<T> c2, T r2) {
  V3<T> n = c2 - c1;
  T len = norm(n);
  n = len < tiny<T>() ? mk<T>(1, 0, 0) : n * (T(1) / len);
  T d = len - r1 - r2;
  o.add(d, c1 + n * (r1 + T(0.5) * d), n, margin);
}
template <typename T> DI void c_plane_capsule(ConOut<T>& o, T margin, V3<T> pn, V3<T> c, V3<T> axis, T r, T h) {
  c_plane_sphere(o, margin, pn, c + axis * h, r);
  c_plane_sphere(o, margin, pn, c - axis * h, r);
}
template <typename T> DI void c_plane_cylinder(ConOut<T>& o, T margin, V3<T> n, V3<T> c, const M3<T>& R, T r, T h) {
  V3<T> axis = col(R, 2);
  T dist0 = dot(c, n), prjaxis = dot(n, axis);
  if (prjaxis > 0) { axis = neg(axis); prjaxis = -prjaxis; }
  V3<T> vec = axis * prjaxis - n;
  T len = norm(vec);
  if (len < T(1e-12)) vec = col(R, 0) * r; else vec = vec * (r / len);
  T prjvec = dot(vec, n);
  axis = axis * h;
  prjaxis *= h;
  T d = dist0 + prjaxis + prjvec;
  if (d > margin) return;
  o.add(d, c + vec + axis - n * (d * T(0.5)), n, margin);
  d = dist0 - prjaxis + prjvec;
  if (d <= margin) o.add(d, c + vec - axis - n * (d * T(0.5)), n, margin);
  d = dist0 + prjaxis - T(0.5) * prjvec;
  if (d <= margin) {
    V3<T> v1 = normalized(cross(vec, axis)) * (r * T(0.8660254037844386));
    o.add(d, c + axis - vec * T(0.5) + v1 - n * (d * T(0.5)), n, margin);
    o.add(d, c + axis - vec * T(0.5) - v1 - n * (d * T(0.5)), n, margin);
  }
}
template <typename T> DI void c_plane_box(ConOut<T>& o, T margin, V3<T> n, V3<T> c, const M3<T>& R, V3<T> s) {
  int cnt = 0;
  for (int k = 0; k < 8 && cnt < 4; k++) {
    V3<T> v = mk<T>((k & 1) ? s.x : -s.x, (k & 2) ? s.y : -s.y, (k & 4) ? s.z : -s.z);
    V3<T> w = mulv(R, v) + c;
    T d = dot(w, n);
    if (d > margin) continue;
    int before = o.n;
    o.add(d, w - n * (T(0.5) * d), n, margin);
    cnt += o.n - before;
  }
}
template <typename T> DI void c_sphere_capsule(ConOut<T>& o, T margin, V3<T> c1, T r1, V3<T> c2, V3<T> axis2, T r2, T h2) {
  T x = tclamp(dot(c1 - c2, axis2), -h2, h2);
  c_sphere_sphere(o, margin, c1, r1, c2 + axis2 * x, r2);
}
template <typename T> DI void c_capsule_capsule(ConOut<T>& o, T margin, V3<T> c1, V3<T> a1, T r1, T h1, V3<T> c2, V3<T> a2, T r2, T h2) {
  V3<T> dif = c1 - c2;
  T mb = -dot(a1, a2), u = -dot(a1, dif), v = dot(a2, dif), det = T(1) - mb * mb;
  if (tabs(det) >= T(1e-12)) {
    T x1 = (u - mb * v) / det, x2 = (v - mb * u) / det;
    if (x1 > h1) { x1 = h1; x2 = v - mb * h1; }
    else if (x1 < -h1) { x1 = -h1; x2 = v + mb * h1; }
    if (x2 > h2) { x2 = h2; x1 = tclamp(u - mb * h2, -h1, h1); }
    else if (x2 < -h2) { x2 = -h2; x1 = tclamp(u + mb * h2, -h1, h1); }
    c_sphere_sphere(o, margin, c1 + a1 * x1, r1, c2 + a2 * x2, r2);
    return;
  }
  int n0 = o.n;
  for (int s = 1; s >= -1 && o.n - n0 < 2; s -= 2) {
    T x1 = s * h1, x2 = v - mb * x1;
    if (x2 > h2 || x2 < -h2) continue;
    c_sphere_sphere(o, margin, c1 + a1 * x1, r1, c2 + a2 * x2, r2);
  }
  for (int s = 1; s >= -1 && o.n - n0 < 2; s -= 2) {
    T x2 = s * h2, x1 = u - mb * x2;
    if (x1 >= h1 || x1 <= -h1) continue;
    c_sphere_sphere(o, margin, c1 + a1 * x1, r1, c2 + a2 * x2, r2);
  }
  if (o.n == n0) {
    T x1 = u > 0 ? h1 : -h1, x2 = tclamp(v - mb * x1, -h2, h2);
    c_sphere_sphere(o, margin, c1 + a1 * x1, r1, c2 + a2 * x2, r2);
  }
}

// sphere (geom1) against box (geom2); q = sphere centre in the box frame
template <typename T> DI void sphere_box(V3<T> c, T r, V3<T> bpos, const M3<T>& bR, V3<T> s, T& dist, V3<T>& pos, V3<T>& nrm) {
  V3<T> q = multv(bR, c - bpos);
  V3<T> cl = mk<T>(tclamp(q.x, -s.x, s.x), tclamp(q.y, -s.y, s.y), tclamp(q.z, -s.z, s.z));
  bool inside = (cl.x == q.x) && (cl.y == q.y) && (cl.z == q.z);
  if (!inside) {
    V3<T> dl = cl - q;
    T len = norm(dl);
    nrm = mulv(bR, dl * (T(1) / len));
    dist = len - r;
    pos = c + nrm * (r + T(0.5) * dist);
    return;
  }
  int best = 0;
  T dmin = T(1e30), sgn = 1;
  for (int i = 0; i < 3; i++) {
    T dp = comp(s, i) - comp(q, i), dn = comp(s, i) + comp(q, i);
    if (dp < dmin) { dmin = dp; best = i; sgn = 1; }
    if (dn < dmin) { dmin = dn; best = i; sgn = -1; }
  }
  V3<T> nl = mk<T>(0, 0, 0);
  setc(nl, best, -sgn);
  nrm = mulv(bR, nl);
  dist = -(r + dmin);
  pos = c + nrm * (T(0.5) * (r - dmin));
}

template <typename T> DI T seg_box_d2(V3<T> a, V3<T> u, V3<T> s, T t) {
  T px = tabs(a.x + t * u.x) - s.x, py = tabs(a.y + t * u.y) - s.y, pz = tabs(a.z + t * u.z) - s.z;
  T r = 0;
  if (px > 0) r += px * px;
  if (py > 0) r += py * py;
  if (pz > 0) r += pz * pz;
  return r;
}
// exact minimiser of the convex piecewise-quadratic t -> dist^2(a + t u, box) on [-h, h]
template <typename T> DI T seg_box_argmin(V3<T> a, V3<T> u, V3<T> s, T h) {
  T br[8];
  int nb = 0;
  br[nb++] = -h;
  br[nb++] = h;
  for (int i = 0; i < 3; i++) {
    T ui = comp(u, i);
    if (tabs(ui) > T(1e-14))
      for (int sg = -1; sg <= 1; sg += 2) {
        T t = (sg * comp(s, i) - comp(a, i)) / ui;
        if (t > -h && t < h) br[nb++] = t;
      }
  }
  for (int i = 1; i < nb; i++) {
    T v = br[i];
    int j = i - 1;
    while (j >= 0 && br[j] > v) { br[j + 1] = br[j]; j--; }
    br[j + 1] = v;
  }
  T best_t = -h, best = seg_box_d2(a, u, s, -h);
  for (int k = 0; k + 1 < nb; k++) {
    T t0 = br[k], t1 = br[k + 1], tm = T(0.5) * (t0 + t1), A = 0, B = 0;
    for (int i = 0; i < 3; i++) {
      T p = comp(a, i) + tm * comp(u, i);
      if (tabs(p) > comp(s, i)) {
        T sg = p > 0 ? T(1) : T(-1), c0 = sg * comp(a, i) - comp(s, i), c1 = sg * comp(u, i);
        A += c1 * c1;
        B += c0 * c1;
      }
    }
    T tc = tm;
    if (A > tiny<T>()) tc = tclamp(-B / A, t0, t1);
    T v = seg_box_d2(a, u, s, tc);
    if (v < best) { best = v; best_t = tc; }
    v = seg_box_d2(a, u, s, t1);
    if (v < best) { best = v; best_t = t1; }
  }
  return best_t;
}

// ---------------------------------------------------------------- MPR on margin-inflated convex shapes
template <typename T> struct MprObj { int type; V3<T> pos; M3<T> R; V3<T> size; T margin; };
template <typename T> struct MprSup { V3<T> v, v1, v2; };

template <typename T> DI V3<T> mpr_support_obj(const MprObj<T>& o, V3<T> dir) {
  V3<T> l = multv(o.R, dir), s = mk<T>(0, 0, 0);
  if (o.type == G_SPHERE) s = l * o.size.x;
  else if (o.type == G_CAPSULE) { s = l * o.size.x; s.z += l.z >= 0 ? o.size.y : -o.size.y; }
  else if (o.type == G_CYLINDER) {
    T n2 = l.x * l.x + l.y * l.y;
    if (n2 > T(1e-24)) { const T q = o.size.x * inv_sqrt(n2); s.x = l.x * q; s.y = l.y * q; }   // (one reciprocal square root on the chain instead of a square root and two divisions)
    s.z = l.z >= 0 ? o.size.y : -o.size.y;
  } else {
    s = mk<T>(l.x >= 0 ? o.size.x : -o.size.x, l.y >= 0 ? o.size.y : -o.size.y, l.z >= 0 ? o.size.z : -o.size.z);
  }
  s = s + l * (T(0.5) * o.margin);
  return mulv(o.R, s) + o.pos;
}
template <typename T> DI void mpr_support(const MprObj<T>& o1, const MprObj<T>& o2, V3<T> dir, MprSup<T>& s) {
  s.v1 = mpr_support_obj(o1, dir);
  s.v2 = mpr_support_obj(o2, neg(dir));
  s.v = s.v1 - s.v2;
}
template <typename T> DI bool is0(T x) { return tabs(x) < eps_rel<T>() * T(1e-3); }

template <typename T> DI T point_tri_dist2(V3<T> x0, V3<T> B, V3<T> C, V3<T>& wit) {  // from the origin
  V3<T> d1 = B - x0, d2 = C - x0, a = x0;
  T v = dot(d1, d1), w = dot(d2, d2), p = dot(a, d1), q = dot(a, d2), r = dot(d1, d2);
  T den = w * v - r * r, s = -1, t = -1;
  if (tabs(den) > tiny<T>()) { s = (q * r - w * p) / den; t = (-s * r - q) / w; }
  const T e = eps_rel<T>();
  if (s >= -e && s <= 1 + e && t >= -e && t <= 1 + e && t + s <= 1 + e) {
    wit = x0 + d1 * s + d2 * t;
    return dot(wit, wit);
  }
  T best = T(1e30);
  for (int ed = 0; ed < 3; ed++) {
    V3<T> E0 = ed == 2 ? B : x0, E1 = ed == 0 ? B : C;
    V3<T> ab = E1 - E0;
    T l2 = dot(ab, ab), tt = l2 > 0 ? tclamp(-dot(E0, ab) / l2, T(0), T(1)) : T(0);
    V3<T> c = E0 + ab * tt;
    T dd = dot(c, c);
    if (dd < best) { best = dd; wit = c; }
  }
  return best;
}

template <typename T> DI void mpr_expand(MprSup<T>* P, const MprSup<T>& v4) {
  V3<T> v4v0 = cross(v4.v, P[0].v);
  if (dot(P[1].v, v4v0) > 0) { if (dot(P[2].v, v4v0) > 0) P[1] = v4; else P[3] = v4; }
  else { if (dot(P[3].v, v4v0) > 0) P[2] = v4; else P[1] = v4; }
}
template <typename T> DI V3<T> portal_dir(const MprSup<T>* P) { return normalized(cross(P[2].v - P[1].v, P[3].v - P[1].v)); }

// returns true and fills depth/dir/pos if the inflated shapes intersect
#ifndef ADROIT_MPR_INLINE
#define ADROIT_MPR_INLINE 1   // measured: inlined -10 % (hammer) / -14 % (pen) of the collide kernel -- the objects stay in registers instead of the caller's local memory
#endif
#if ADROIT_MPR_INLINE
#define MPR_FN __device__ __forceinline__
#else
#define MPR_FN __device__ __noinline__
#endif
template <typename T>
MPR_FN bool mpr_penetration(const MprObj<T>& o1, const MprObj<T>& o2, int max_iter, T tol, T& depth, V3<T>& dir, V3<T>& pos) {
  MprSup<T> P[4], v4;
  V3<T> d;
  T dt;
  P[0].v1 = o1.pos; P[0].v2 = o2.pos; P[0].v = o1.pos - o2.pos;
  if (is0(P[0].v.x) && is0(P[0].v.y) && is0(P[0].v.z)) P[0].v.x += T(1e-6);
  d = normalized(neg(P[0].v));
  mpr_support(o1, o2, d, P[1]);
  dt = dot(P[1].v, d);
  if (is0(dt) || dt < 0) return false;
  d = cross(P[0].v, P[1].v);
  if (is0(dot(d, d))) {
    pos = (P[1].v1 + P[1].v2) * T(0.5);
    if (is0(dot(P[1].v, P[1].v))) { depth = 0; dir = mk<T>(0, 0, 0); return true; }
    depth = norm(P[1].v);
    dir = normalized(P[1].v);
    return true;
  }
  d = normalized(d);
  mpr_support(o1, o2, d, P[2]);
  dt = dot(P[2].v, d);
  if (is0(dt) || dt < 0) return false;
  d = normalized(cross(P[1].v - P[0].v, P[2].v - P[0].v));
  if (dot(d, P[0].v) > 0) { MprSup<T> t = P[1]; P[1] = P[2]; P[2] = t; d = neg(d); }
  for (int guard = 0;; guard++) {
    if (guard > 100) return false;
    mpr_support(o1, o2, d, P[3]);
    dt = dot(P[3].v, d);
    if (is0(dt) || dt < 0) return false;
    bool cont = false;
    dt = dot(cross(P[1].v, P[3].v), P[0].v);
    if (dt < 0 && !is0(dt)) { P[2] = P[3]; cont = true; }
    if (!cont) {
      dt = dot(cross(P[3].v, P[2].v), P[0].v);
      if (dt < 0 && !is0(dt)) { P[1] = P[3]; cont = true; }
    }
    if (!cont) break;
    d = normalized(cross(P[1].v - P[0].v, P[2].v - P[0].v));
  }
  for (int guard = 0;; guard++) {
    if (guard > 200) return false;
    d = portal_dir(P);
    dt = dot(d, P[1].v);
    if (is0(dt) || dt > 0) break;
    mpr_support(o1, o2, d, v4);
    dt = dot(v4.v, d);
    if (!(is0(dt) || dt > 0)) return false;
    T dv4 = dot(v4.v, d);
    T mn = tmin(tmin(dv4 - dot(P[1].v, d), dv4 - dot(P[2].v, d)), dv4 - dot(P[3].v, d));
    if (mn <= tol) return false;
    mpr_expand(P, v4);
  }
  for (int it = 0;; it++) {
    d = portal_dir(P);
    mpr_support(o1, o2, d, v4);
    T dv4 = dot(v4.v, d);
    T mn = tmin(tmin(dv4 - dot(P[1].v, d), dv4 - dot(P[2].v, d)), dv4 - dot(P[3].v, d));
    if (mn <= tol || it > max_iter) {
      V3<T> w = mk<T>(0, 0, 0);
      T d2 = point_tri_dist2(P[1].v, P[2].v, P[3].v, w);
      depth = sqrt(d2);
      dir = is0(dot(w, w)) ? d : normalized(w);
      T b0 = dot(cross(P[1].v, P[2].v), P[3].v), b1 = dot(cross(P[3].v, P[2].v), P[0].v);
      T b2 = dot(cross(P[0].v, P[1].v), P[3].v), b3 = dot(cross(P[2].v, P[1].v), P[0].v);
      T sum = b0 + b1 + b2 + b3;
      if (is0(sum) || sum < 0) {
        b0 = 0;
        b1 = dot(cross(P[2].v, P[3].v), d);
        b2 = dot(cross(P[3].v, P[1].v), d);
        b3 = dot(cross(P[1].v, P[2].v), d);
        sum = b1 + b2 + b3;
      }
      V3<T> q1 = P[0].v1 * b0 + P[1].v1 * b1 + P[2].v1 * b2 + P[3].v1 * b3;
      V3<T> q2 = P[0].v2 * b0 + P[1].v2 * b1 + P[2].v2 * b2 + P[3].v2 * b3;
      pos = (q1 + q2) * (T(0.5) / sum);
      return true;
    }
    mpr_expand(P, v4);
  }
}

// The portal refinement is evaluated in fp64 in BOTH builds: for shallow (margin-level) penetrations near a cylinder rim
// the portal that survives is decided by differences of ~1e-6-scale support values, and in fp32 the surviving portal --
// hence the contact normal -- flips (measured on B200: every fp32 single-step error above 1e-3 in pen-v0 was an MPR
// normal, tools/gpu_worst.py).  B200 runs fp64 at half the fp32 rate and MPR pairs are few after the broadphase.
// the MPR contact of one pair: false = no contact found; otherwise dist / pos / nrm as the narrow phase reports them
template <typename T>
DI bool mpr_contact(T margin, int t1, V3<T> p1, const M3<T>& R1, V3<T> s1, int t2, V3<T> p2, const M3<T>& R2, V3<T> s2, int max_iter, T tol,
                    T& dist, V3<T>& cpos, V3<T>& cnrm) {
  MprObj<double> o1, o2;
  o1.type = t1; o1.pos = mk<double>(p1.x, p1.y, p1.z); o1.size = mk<double>(s1.x, s1.y, s1.z); o1.margin = margin;
  o2.type = t2; o2.pos = mk<double>(p2.x, p2.y, p2.z); o2.size = mk<double>(s2.x, s2.y, s2.z); o2.margin = margin;
  for (int e = 0; e < 9; e++) { o1.R.m[e] = R1.m[e]; o2.R.m[e] = R2.m[e]; }
  double depth;
  V3<double> dir, pos;
  if (!mpr_penetration<double>(o1, o2, max_iter, (double)tol, depth, dir, pos)) return false;
  if (dot(dir, dir) < 0.5) return false;
  dist = margin - (T)depth;
  cpos = mk<T>((T)pos.x, (T)pos.y, (T)pos.z);
  cnrm = mk<T>((T)dir.x, (T)dir.y, (T)dir.z);
  return true;
}
template <typename T>
DI void c_mpr(ConOut<T>& o, T margin, int t1, V3<T> p1, const M3<T>& R1, V3<T> s1, int t2, V3<T> p2, const M3<T>& R2, V3<T> s2,
              int max_iter, T tol) {
  T dist;
  V3<T> pos, nrm;
  if (mpr_contact(margin, t1, p1, R1, s1, t2, p2, R2, s2, max_iter, tol, dist, pos, nrm)) o.add(dist, pos, nrm, margin);
}
// does collide_pair send this type pair to MPR (its `default:` case)?  Everything but the analytic routines listed there.
DI bool is_mpr_pair(int t1, int t2) {
  switch (t1 * 8 + t2) {
    case G_PLANE * 8 + G_SPHERE: case G_PLANE * 8 + G_CAPSULE: case G_PLANE * 8 + G_CYLINDER: case G_PLANE * 8 + G_BOX:
    case G_SPHERE * 8 + G_SPHERE: case G_SPHERE * 8 + G_CAPSULE: case G_SPHERE * 8 + G_BOX:
    case G_CAPSULE * 8 + G_CAPSULE: case G_CAPSULE * 8 + G_BOX: case G_BOX * 8 + G_BOX: return false;
    default: return true;
  }
}

template <typename T>
DI void c_capsule_box(ConOut<T>& o, T margin, V3<T> c, const M3<T>& cR, T r, T h, V3<T> bpos, const M3<T>& bR, V3<T> bs,
                      int max_iter, T tol) {
  V3<T> axis = col(cR, 2);
  V3<T> a = multv(bR, c - bpos), u = multv(bR, axis);
  T ts = seg_box_argmin(a, u, bs, h);
  T d2s = seg_box_d2(a, u, bs, ts);
  if (d2s <= (sizeof(T) == 4 ? T(1e-14) : T(1e-24))) {  // the core segment pierces the box: deep penetration
    c_mpr(o, margin, (int)G_CAPSULE, c, cR, mk<T>(r, h, 0), (int)G_BOX, bpos, bR, bs, max_iter, tol);
    return;
  }
  T lim = margin + r, ds = sqrt(d2s);
  T d0 = sqrt(seg_box_d2(a, u, bs, -h)), d1 = sqrt(seg_box_d2(a, u, bs, h));
  T dmin_end = tmin(d0, d1);
  if (ds >= dmin_end - (sizeof(T) == 4 ? T(1e-6) : T(1e-9)) * (h + r)) ts = d0 <= d1 ? -h : h;
  T dist;
  V3<T> pos, n;
  sphere_box(c + axis * ts, r, bpos, bR, bs, dist, pos, n);
  o.add(dist, pos, n, margin);
  for (int s = -1; s <= 1; s += 2) {
    T de = s < 0 ? d0 : d1;
    if (de >= lim || de <= 0 || tabs(s * h - ts) <= T(1e-6) * h) continue;
    sphere_box(c + axis * (s * h), r, bpos, bR, bs, dist, pos, n);
    o.add(dist, pos, n, margin);
  }
}

// ---------------------------------------------------------------- box-box (SAT + reference-face clipping)
template <typename T> DI int clip_poly(V3<T>* poly, int n, V3<T> pn, T pd) {
  V3<T> out[16];
  int no = 0;
  for (int i = 0; i < n; i++) {
    V3<T> A = poly[i], B = poly[(i + 1) % n];
    T da = dot(pn, A) - pd, db = dot(pn, B) - pd;
    if (da <= 0) out[no++] = A;
    if ((da < 0 && db > 0) || (da > 0 && db < 0)) out[no++] = A + (B - A) * (da / (da - db));
    if (no >= 15) break;
  }
  for (int i = 0; i < no; i++) poly[i] = out[i];
  return no;
}
template <typename T>
__device__ __noinline__ void c_box_box(ConOut<T>& o, T margin, V3<T> p1, const M3<T>& R1, V3<T> s1, V3<T> p2, const M3<T>& R2, V3<T> s2) {
  V3<T> ax1[3], ax2[3];
  for (int i = 0; i < 3; i++) { ax1[i] = col(R1, i); ax2[i] = col(R2, i); }
  V3<T> dp = p2 - p1;
  T best = T(-1e30);
  V3<T> bestn = mk<T>(1, 0, 0);
  int bestcode = -1;
  for (int code = 0; code < 6; code++) {
    V3<T> L = code < 3 ? ax1[code] : ax2[code - 3];
    T ra = 0, rb = 0;
    for (int i = 0; i < 3; i++) { ra += comp(s1, i) * tabs(dot(ax1[i], L)); rb += comp(s2, i) * tabs(dot(ax2[i], L)); }
    T dl = dot(dp, L), sep = tabs(dl) - ra - rb;
    if (sep > margin) return;
    if (sep > best) { best = sep; bestcode = code; bestn = L * (dl >= 0 ? T(1) : T(-1)); }
  }
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      V3<T> L = cross(ax1[i], ax2[j]);
      T len = norm(L);
      if (len < T(1e-6)) continue;
      L = L * (T(1) / len);
      T ra = 0, rb = 0;
      for (int k = 0; k < 3; k++) { ra += comp(s1, k) * tabs(dot(ax1[k], L)); rb += comp(s2, k) * tabs(dot(ax2[k], L)); }
      T dl = dot(dp, L), sep = tabs(dl) - ra - rb;
      if (sep > margin) return;
      if (sep > best + T(1e-6) + T(0.02) * tabs(best)) { best = sep; bestcode = 6 + 3 * i + j; bestn = L * (dl >= 0 ? T(1) : T(-1)); }
    }
  if (bestcode >= 6) {
    int i = (bestcode - 6) / 3, j = (bestcode - 6) % 3;
    V3<T> pa = p1, pb = p2;
    for (int k = 0; k < 3; k++) {
      if (k != i) pa = pa + ax1[k] * (dot(ax1[k], bestn) > 0 ? comp(s1, k) : -comp(s1, k));
      if (k != j) pb = pb + ax2[k] * (dot(ax2[k], bestn) > 0 ? -comp(s2, k) : comp(s2, k));
    }
    V3<T> w = pa - pb;
    T b = dot(ax1[i], ax2[j]), dd = dot(ax1[i], w), e = dot(ax2[j], w), den = 1 - b * b;
    T x = den > T(1e-12) ? (b * e - dd) / den : T(0), y = den > T(1e-12) ? (e - b * dd) / den : T(0);
    x = tclamp(x, -comp(s1, i), comp(s1, i));
    y = tclamp(y, -comp(s2, j), comp(s2, j));
    o.add(best, ((pa + ax1[i] * x) + (pb + ax2[j] * y)) * T(0.5), bestn, margin);
    return;
  }
  bool ref1 = bestcode < 3;
  int ra = ref1 ? bestcode : bestcode - 3;
  V3<T>* RA = ref1 ? ax1 : ax2;
  V3<T>* IA = ref1 ? ax2 : ax1;
  V3<T> rp = ref1 ? p1 : p2, ip = ref1 ? p2 : p1, rs = ref1 ? s1 : s2, is = ref1 ? s2 : s1;
  V3<T> nref = bestn * (ref1 ? T(1) : T(-1));
  int ia = 0;
  T mn = T(1e30), isgn = 1;
  for (int k = 0; k < 3; k++) {
    T v = dot(IA[k], nref);
    if (v < mn) { mn = v; ia = k; isgn = 1; }
    if (-v < mn) { mn = -v; ia = k; isgn = -1; }
  }
  int i1 = (ia + 1) % 3, i2 = (ia + 2) % 3;
  V3<T> poly[16];
  V3<T> fc = ip + IA[ia] * (isgn * comp(is, ia));
  for (int k = 0; k < 4; k++) {
    T sa = (k == 0 || k == 3) ? T(1) : T(-1), sb = (k < 2) ? T(1) : T(-1);
    poly[k] = fc + IA[i1] * (sa * comp(is, i1)) + IA[i2] * (sb * comp(is, i2));
  }
  int np = 4, r1 = (ra + 1) % 3, r2 = (ra + 2) % 3;
  for (int k = 0; k < 4 && np > 0; k++) {
    int ax = k < 2 ? r1 : r2;
    V3<T> pn = RA[ax] * ((k & 1) ? T(-1) : T(1));
    np = clip_poly(poly, np, pn, dot(pn, rp) + comp(rs, ax));
  }
  T plane_d = dot(nref, rp) + comp(rs, ra);
  for (int k = 0; k < np && o.n < 8; k++) {
    T depth = dot(nref, poly[k]) - plane_d;
    if (depth >= margin) continue;
    o.add(depth, poly[k] - nref * (T(0.5) * depth), bestn, margin);
  }
}

// dispatch one candidate pair; geom1 is translated to the origin by the caller (p1 = 0 unless plane)
template <typename T>
__device__ __noinline__ void collide_pair(ConOut<T>& o, T margin, int t1, V3<T> p1, const M3<T>& R1, V3<T> s1, int t2, V3<T> p2,
                                          const M3<T>& R2, V3<T> s2, int mpr_iter, T mpr_tol) {
  o.n = 0;
  switch (t1 * 8 + t2) {
    case G_PLANE * 8 + G_SPHERE: c_plane_sphere(o, margin, col(R1, 2), p2 - p1, s2.x); break;
    case G_PLANE * 8 + G_CAPSULE: c_plane_capsule(o, margin, col(R1, 2), p2 - p1, col(R2, 2), s2.x, s2.y); break;
    case G_PLANE * 8 + G_CYLINDER: c_plane_cylinder(o, margin, col(R1, 2), p2 - p1, R2, s2.x, s2.y); break;
    case G_PLANE * 8 + G_BOX: c_plane_box(o, margin, col(R1, 2), p2 - p1, R2, s2); break;
    case G_SPHERE * 8 + G_SPHERE: c_sphere_sphere(o, margin, p1, s1.x, p2, s2.x); break;
    case G_SPHERE * 8 + G_CAPSULE: c_sphere_capsule(o, margin, p1, s1.x, p2, col(R2, 2), s2.x, s2.y); break;
    case G_SPHERE * 8 + G_BOX: {
      T dist; V3<T> pos, n;
      sphere_box(p1, s1.x, p2, R2, s2, dist, pos, n);
      o.add(dist, pos, n, margin);
    } break;
    case G_CAPSULE * 8 + G_CAPSULE: c_capsule_capsule(o, margin, p1, col(R1, 2), s1.x, s1.y, p2, col(R2, 2), s2.x, s2.y); break;
    case G_CAPSULE * 8 + G_BOX: c_capsule_box(o, margin, p1, R1, s1.x, s1.y, p2, R2, s2, mpr_iter, mpr_tol); break;
    case G_BOX * 8 + G_BOX: c_box_box(o, margin, p1, R1, s1, p2, R2, s2); break;
    default: c_mpr(o, margin, t1, p1, R1, s1, t2, p2, R2, s2, mpr_iter, mpr_tol); break;
  }
  // plane routines work relative to the plane point: shift back
  if (t1 == G_PLANE)
    for (int k = 0; k < o.n; k++) o.pos[k] = o.pos[k] + p1;
}

}  // namespace adroit
