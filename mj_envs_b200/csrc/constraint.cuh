// constraint.cuh — collision driver, constraint rows, Newton solver, noslip pass, touch sensor, Euler step.
//
// Restates MuJoCo 2.1's soft-constraint pipeline for the row types the Adroit scenes produce:
// dof friction loss, joint limits, fixed-tendon limits and pyramidal contacts (condim 1/3/4), solved in primal
// space by Newton's method with an exact line search, followed by the noslip Gauss-Seidel pass on friction rows.
//
// Representation (B200-first, not MuJoCo's dense efc_J):
//   * "simple" rows (friction / limits) are never materialised: J is a signed unit vector (2 entries for a tendon);
//   * a contact stores its 4-row contact-frame Jacobian Jf (normal, tangent1, tangent2, torsion) once; the 1/4/6
//     pyramid edges are the combinations Jn +- mu_k Jt_k, evaluated on the fly from Jf*x (4 dot products per contact
//     instead of 6 rows), and the Hessian contribution is Jf^T W Jf with a 4x4 weight W built from the edge states;
//   * the noslip pass is matrix-free in acceleration space: with u = M^-1 J^T f the residual of a row is J_i u + b_i,
//     so A = J M^-1 J^T (nefc^2 entries) is never formed.
#pragma once
#include "dynamics.cuh"

namespace adroit {

// (noinline: the general-power branch pulls in pow(); one copy instead of four keeps the solve kernel's code small)
template <typename T> SUB_FN T impedance(const DModel<T>& m, T pos, T margin) {
  T s0 = tclamp(m.solimp[0], T(0.0001), T(0.9999)), s1 = tclamp(m.solimp[1], T(0.0001), T(0.9999)), width = m.solimp[2];
  T mid = tclamp(m.solimp[3], T(0.0001), T(0.9999)), power = m.solimp[4] < 1 ? T(1) : m.solimp[4];
  if (s0 == s1 || width <= T(1e-15)) return T(0.5) * (s0 + s1);
  T x = tabs((pos - margin) / width);
  if (x >= 1) return s1;
  if (x <= 0) return s0;
  T y;
  if (power == T(1)) y = x;
  else if (power == T(2)) y = x <= mid ? x * x / mid : 1 - (1 - x) * (1 - x) / (1 - mid);
  else if (x <= mid) y = pow(x, power) / pow(mid, power - 1);
  else y = 1 - pow(1 - x, power) / pow(1 - mid, power - 1);
  return s0 + y * (s1 - s0);
}
template <typename T> DI void sol_KB(const DModel<T>& m, T& K, T& B) {
  T tc = tmax(m.solref[0], 2 * m.timestep), dmax = tclamp(m.solimp[1], T(0.0001), T(0.9999));
  K = T(1) / tmax(T(1e-15), dmax * dmax * tc * tc * m.solref[1] * m.solref[1]);
  B = T(2) / tmax(T(1e-15), dmax * tc);
}

// lower bound of the distance from a point (given relative to the geom centre, world axes) to the geom's surface
template <typename T> DI T shape_lower_bound(int type, const T* R, const T* size, T px, T py, T pz) {
  T lx = R[0] * px + R[3] * py + R[6] * pz, ly = R[1] * px + R[4] * py + R[7] * pz, lz = R[2] * px + R[5] * py + R[8] * pz;
  if (type == G_BOX) {
    T ex = tmax(tabs(lx) - size[0], T(0)), ey = tmax(tabs(ly) - size[1], T(0)), ez = tmax(tabs(lz) - size[2], T(0));
    return sqrt(ex * ex + ey * ey + ez * ez);
  }
  if (type == G_CAPSULE || type == G_CYLINDER) {   // a cylinder lies inside the capsule of the same radius and half length
    T ez = tmax(tabs(lz) - size[1], T(0));
    return sqrt(lx * lx + ly * ly + ez * ez) - size[0];
  }
  if (type == G_SPHERE) return sqrt(lx * lx + ly * ly + lz * lz) - size[0];
  return T(-1e30);
}

// ---------------------------------------------------------------- collision over the static candidate list
// Three steps (collide_kernel, adroit_b200.cu):
//   1. collision_broad  (per warp = env): bounding-sphere + shape-aware cull of the static candidate list, survivors
//      compacted into a 16-bit list (lane = pair, ballot); the survivors whose type pair goes to MPR (anything with a
//      cylinder but plane-cylinder) are also pushed, in survivor order, into the warp's queue in shared memory;
//   2. collision_mpr_pool (whole block): the queues of the block's envs are worked off together, one lane per queued pair --
//      MPR (fp64 portal refinement, ~40 % of the kernel) then runs once per block in full lanes instead of once per env in
//      one or two lanes; results go to the env's G_MPRRES slots;
//   3. collision_narrow (per warp): the analytic routines, lane = survivor; queued pairs pick up their MPR result; ordered
//      compaction keeps MuJoCo's contact order (candidate-list order).
template <class CX> struct CollideGeom {
  using T = typename CX::real;
  const DModel<T>& m;
  const T *gpos, *gmat, *par;
  int var_geom;
  // hammer "size" variation (hammer_v0.py:123-125): geom_size of the head is a per-env parameter; geom_rbound stays stale
  // as in MuJoCo (the model is not recompiled), except in this file's own extra culls, which must stay conservative
  __device__ CollideGeom(const DModel<T>& mm, const T* gp, const T* gm, const T* pr)
      : m(mm), gpos(gp), gmat(gm), par(pr), var_geom(mm.task_int[8] == VAR_SIZE ? mm.task_int[10] : -1) {}
  DI T gsize(int g, int k) const { return (g == var_geom && k < 2) ? par[3 + k] : m.geom_size[3 * g + k]; }
  DI T rb_safe(int g) const {
    T rb = m.geom_rbound[g];
    if (g == var_geom) rb = tmax(rb, sqrt(par[3] * par[3] + par[4] * par[4]));
    return rb;
  }
  // pair k with geom1 translated to the origin
  DI void pair(int k, int& g1, int& g2, M3<T>& R1, M3<T>& R2, V3<T>& origin, V3<T>& p2, V3<T>& s1, V3<T>& s2) const {
    g1 = m.pair_geom[2 * k];
    g2 = m.pair_geom[2 * k + 1];
    for (int e = 0; e < 9; e++) { R1.m[e] = gmat[9 * g1 + e]; R2.m[e] = gmat[9 * g2 + e]; }
    origin = mk<T>(gpos[3 * g1], gpos[3 * g1 + 1], gpos[3 * g1 + 2]);
    p2 = mk<T>(gpos[3 * g2], gpos[3 * g2 + 1], gpos[3 * g2 + 2]) - origin;
    s1 = mk<T>(gsize(g1, 0), gsize(g1, 1), gsize(g1, 2));
    s2 = mk<T>(gsize(g2, 0), gsize(g2, 1), gsize(g2, 2));
  }
};

// returns the number of survivors (list in I_SURV); qk / qn: this warp's MPR queue (pair indices, count)
template <class CX> PHASE_FN int collision_broad(CX& c, unsigned long long* overflow, unsigned short* qk, int* qn) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const CollideGeom<CX> G(m, c.template G<G_GPOS>(), c.template G<G_GMAT>(), c.template S<O_PAR>());
  const T *gpos = G.gpos, *gmat = G.gmat;
  unsigned short* surv = reinterpret_cast<unsigned short*>(c.template I<I_SURV>());   // pair indices (< 65536), 16 bits each: MAXSURV / 2 words
  const unsigned lt = (1u << c.lane) - 1u;
  int nsurv = 0, nmpr = 0;
  for (int base = 0; base < m.npair; base += 32) {
    int k = base + c.lane;
    bool hit = false, mpr = false;
    if (k < m.npair) {
      const int w = m.pair_word[k];
      const T bound = m.pair_bound[k];
      const int g1 = w & 255, g2 = (w >> 8) & 255, t1 = (w >> 16) & 255, t2 = (w >> 24) & 255;
      T dx = gpos[3 * g2] - gpos[3 * g1], dy = gpos[3 * g2 + 1] - gpos[3 * g1 + 1], dz = gpos[3 * g2 + 2] - gpos[3 * g1 + 2];
      if (t1 == G_PLANE) {
        const T* R = gmat + 9 * g1;
        hit = !(dx * R[2] + dy * R[5] + dz * R[8] > bound);
      } else {
        hit = !(dx * dx + dy * dy + dz * dz > bound * bound);
        const T margin = hit ? m.pair_margin[k] : T(0);
        // second, tighter (still conservative) test: the other geom's bounding sphere against this geom's own shape
        // (large boxes -- table, wall, door -- have bounding spheres that swallow the whole hand)
        T z1[3] = {G.gsize(g1, 0), G.gsize(g1, 1), G.gsize(g1, 2)}, z2[3] = {G.gsize(g2, 0), G.gsize(g2, 1), G.gsize(g2, 2)};
        if (hit) hit = !(shape_lower_bound(t1, gmat + 9 * g1, z1, dx, dy, dz) > G.rb_safe(g2) + margin);
        if (hit) hit = !(shape_lower_bound(t2, gmat + 9 * g2, z2, -dx, -dy, -dz) > G.rb_safe(g1) + margin);
      }
      mpr = hit && is_mpr_pair(t1, t2);
    }
    unsigned mask = __ballot_sync(0xffffffffu, hit);
    int p = nsurv + __popc(mask & lt);
    if (hit && p < MAXSURV) surv[p] = (unsigned short)k;
    // (a survivor beyond MAXSURV is dropped and counted below; it is not queued either)
    unsigned mm = __ballot_sync(0xffffffffu, mpr && p < MAXSURV);
    int j = nmpr + __popc(mm & lt);
    if (mpr && p < MAXSURV && j < MPRQ) qk[j] = (unsigned short)k;
    nsurv += __popc(mask);
    nmpr += __popc(mm);
  }
  if (nsurv > MAXSURV) {
    if (c.lane == 0) atomicAdd(overflow, (unsigned long long)(nsurv - MAXSURV));
    nsurv = MAXSURV;
  }
  if (c.lane == 0) *qn = nmpr < MPRQ ? nmpr : MPRQ;
  __syncwarp();
  return nsurv;
}

// one queued MPR pair (all threads of the block call this with their own entry): env_gb = context block of the pair's env
template <class CX> DI void collision_mpr_entry(const DModel<typename CX::real>& m, const typename CX::real* gpos, const typename CX::real* gmat,
                                                typename CX::real* env_gb, int k, int slot) {
  using T = typename CX::real;
  const CollideGeom<CX> G(m, gpos, gmat, env_gb + O_PAR);
  int g1, g2;
  M3<T> R1, R2;
  V3<T> origin, p2, s1, s2;
  G.pair(k, g1, g2, R1, R2, origin, p2, s1, s2);
  T dist = 0;
  V3<T> pos = mk<T>(0, 0, 0), nrm = mk<T>(0, 0, 0);
  const bool hit = mpr_contact(m.pair_margin[k], m.geom_type[g1], mk<T>(0, 0, 0), R1, s1, m.geom_type[g2], p2, R2, s2, m.mpr_iterations,
                               m.mpr_tolerance, dist, pos, nrm);
  T* R = env_gb + SM_T + G_MPRRES + 8 * slot;
  R[0] = hit ? T(1) : T(0);
  R[1] = dist; R[2] = pos.x; R[3] = pos.y; R[4] = pos.z; R[5] = nrm.x; R[6] = nrm.y; R[7] = nrm.z;
}

template <class CX> PHASE_FN void collision_narrow(CX& c, int nsurv, unsigned long long* overflow) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const CollideGeom<CX> G(m, c.template G<G_GPOS>(), c.template G<G_GMAT>(), c.template S<O_PAR>());
  const unsigned short* surv = reinterpret_cast<const unsigned short*>(c.template I<I_SURV>());
  const T* mres = c.template G<G_MPRRES>();
  const unsigned lt = (1u << c.lane) - 1u;
  int ncon = 0, nmpr = 0;
  T* cmu = c.template S<O_CMU>();
  for (int base = 0; base < nsurv; base += 32) {
    ConOut<T> out;
    out.n = 0;
    int i = base + c.lane, k = -1, g1 = 0, g2 = 0;
    V3<T> origin = mk<T>(0, 0, 0);
    bool mpr = false;
    if (i < nsurv) {
      k = surv[i];
      mpr = is_mpr_pair(m.geom_type[m.pair_geom[2 * k]], m.geom_type[m.pair_geom[2 * k + 1]]);
    }
    const unsigned mm = __ballot_sync(0xffffffffu, mpr);
    const int j = nmpr + __popc(mm & lt);       // this survivor's place in the warp's MPR queue (same order as collision_broad)
    nmpr += __popc(mm);
    if (i < nsurv) {
      M3<T> R1, R2;
      V3<T> p2, s1, s2;
      G.pair(k, g1, g2, R1, R2, origin, p2, s1, s2);
      if (mpr && j < MPRQ) {                    // computed by the pooled pass
        const T* R = mres + 8 * j;
        if (R[0] != T(0)) out.add(R[1], mk<T>(R[2], R[3], R[4]), mk<T>(R[5], R[6], R[7]), m.pair_margin[k]);
      } else {
        collide_pair(out, m.pair_margin[k], m.geom_type[g1], mk<T>(0, 0, 0), R1, s1, m.geom_type[g2], p2, R2, s2,
                     m.mpr_iterations, m.mpr_tolerance);
      }
    }
    // ordered compaction: contacts keep candidate-list order (the noslip sweep order depends on it)
    int cnt = out.n, incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (c.lane >= o) incl += t;
    }
    int total = __shfl_sync(0xffffffffu, incl, 31), excl = incl - cnt;
    for (int q = 0; q < cnt; q++) {
      int ci = ncon + excl + q;
      if (ci >= MAXCON) break;
      T* C = c.template G<G_CON>() + 16 * ci;
      V3<T> pos = out.pos[q] + origin, x = normalized(out.nrm[q]);
      V3<T> y = (x.y > T(0.5) || x.y < T(-0.5)) ? mk<T>(0, 0, 1) : mk<T>(0, 1, 0);
      y = normalized(y - x * dot(x, y));
      V3<T> z = cross(x, y);
      C[0] = out.dist[q];
      C[1] = pos.x; C[2] = pos.y; C[3] = pos.z;
      C[4] = x.x; C[5] = x.y; C[6] = x.z; C[7] = y.x; C[8] = y.y; C[9] = y.z; C[10] = z.x; C[11] = z.y; C[12] = z.z;
      C[13] = m.pair_margin[k];
      c.template I<I_CDIM>()[ci] = m.pair_condim[k];
      c.template I<I_CB1>()[ci] = m.geom_body[g1];
      c.template I<I_CB2>()[ci] = m.geom_body[g2];
      cmu[3 * ci] = m.pair_friction[5 * k];
      cmu[3 * ci + 1] = m.pair_friction[5 * k + 1];
      cmu[3 * ci + 2] = m.pair_friction[5 * k + 2];
      c.gi[GI_CON + 4 * ci] = g1;
      c.gi[GI_CON + 4 * ci + 1] = g2;
      c.gi[GI_CON + 4 * ci + 2] = m.pair_condim[k];
    }
    ncon += total;
  }
  if (ncon > MAXCON) {
    if (c.lane == 0) atomicAdd(overflow, (unsigned long long)(ncon - MAXCON));
    ncon = MAXCON;
  }
  c.ncon = ncon;
  __syncwarp();
}

// ---------------------------------------------------------------- constraint rows
template <typename T> DI int n_edges(int dim) { return dim == 1 ? 1 : 2 * (dim - 1); }

template <class CX> PHASE_FN void make_constraint(CX& c, unsigned long long* overflow) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const T *qpos = c.template V<V_QPOS>(), *qvel = c.template V<V_QVEL>();
  int* sinfo = c.template I<I_SINFO>();
  T *sD = c.template S<O_SD>(), *saref = c.template S<O_SAREF>(), *sfl = c.template S<O_SFL>(), *sforce = c.template S<O_SFORCE>();
  const unsigned lt = (1u << c.lane) - 1u;
  T K, B;
  sol_KB(m, K, B);
  int nrow = 0, dropped = 0;
  // friction loss (one row per dof with frictionloss > 0)
  _Pragma("unroll 1") for (int base = 0; base < m.nv; base += 32) {
    int j = base + c.lane;
    bool has = j < m.nv && m.dof_frictionloss[j] > 0;
    unsigned mask = __ballot_sync(0xffffffffu, has);
    int r = nrow + __popc(mask & lt);
    if (has && r < MAXSIMPLE) {
      T imp = impedance(m, T(0), T(0));
      T R = tmax(T(1e-15), (1 - imp) * m.dof_invweight0[j] / imp);
      sinfo[r] = S_FRICTION | (j << 2);
      sD[r] = T(1) / R;
      saref[r] = -B * qvel[j];
      sfl[r] = m.dof_frictionloss[j];
      sforce[r] = 0;
    }
    nrow += __popc(mask);
  }
  // joint limits: q = 2*j + (upper side)
  _Pragma("unroll 1") for (int base = 0; base < 2 * m.nv; base += 32) {
    int q = base + c.lane, j = q >> 1, up = q & 1;
    bool has = false;
    T dist = 0;
    if (j < m.nv && m.jnt_limited[j]) {
      dist = up ? m.jnt_range[2 * j + 1] - qpos[j] : qpos[j] - m.jnt_range[2 * j];
      has = dist < m.jnt_margin[j];
    }
    unsigned mask = __ballot_sync(0xffffffffu, has);
    int r = nrow + __popc(mask & lt);
    if (has && r < MAXSIMPLE) {
      T mg = m.jnt_margin[j], imp = impedance(m, dist, mg);
      T R = tmax(T(1e-15), (1 - imp) * m.dof_invweight0[j] / imp);
      T vel = up ? -qvel[j] : qvel[j];
      sinfo[r] = S_JLIMIT | (j << 2) | (up << 12);
      sD[r] = T(1) / R;
      saref[r] = -B * vel - K * imp * (dist - mg);
      sfl[r] = 0;
      sforce[r] = 0;
    }
    nrow += __popc(mask);
  }
  // tendon limits
  _Pragma("unroll 1") for (int base = 0; base < 2 * m.ntendon; base += 32) {
    int q = base + c.lane, t = q >> 1, up = q & 1;
    bool has = false;
    T dist = 0, vel = 0;
    if (t < m.ntendon && m.tendon_limited[t]) {
      T len = 0;
      for (int k = 0; k < 2; k++) {
        int dj = m.tdof[2 * t + k];
        if (dj >= 0) { len += m.tcoef[2 * t + k] * qpos[dj]; vel += m.tcoef[2 * t + k] * qvel[dj]; }
      }
      dist = up ? m.tendon_range[2 * t + 1] - len : len - m.tendon_range[2 * t];
      has = dist < m.tendon_margin[t];
    }
    unsigned mask = __ballot_sync(0xffffffffu, has);
    int r = nrow + __popc(mask & lt);
    if (has && r < MAXSIMPLE) {
      T mg = m.tendon_margin[t], imp = impedance(m, dist, mg);
      T R = tmax(T(1e-15), (1 - imp) * m.tendon_invweight0[t] / imp);
      sinfo[r] = S_TLIMIT | (t << 2) | (up << 12);
      sD[r] = T(1) / R;
      saref[r] = -B * (up ? -vel : vel) - K * imp * (dist - mg);
      sfl[r] = 0;
      sforce[r] = 0;
    }
    nrow += __popc(mask);
  }
  if (nrow > MAXSIMPLE) { dropped = nrow - MAXSIMPLE; nrow = MAXSIMPLE; }
  if (dropped && c.lane == 0) atomicAdd(overflow + 1, (unsigned long long)dropped);
  c.nsimple = nrow;

  // contacts: contact-frame Jacobian rows (both layouts), frame velocity, impedance
  const T *xanc = c.template S<O_XANC>(), *xaxis = c.template S<O_XAXIS>();
  T *JR = c.template G<G_JR>();
  T *cD = c.template S<O_CD>(), *cmu = c.template S<O_CMU>(), *carefn = c.template S<O_CAREFN>(), *cbt = c.template S<O_CBT>(), *ef = c.template S<O_EF>();
  // Everything that does not depend on the contact is read once, up front: the lane's two dofs (lane, lane + 32) with
  // their joint type, axis, anchor and velocity, and the dof masks of every contact's two bodies (lane = contact).  The
  // contact record of the NEXT contact is fetched while the current one is worked on, so that no model-table or
  // context load sits on the dependent chain of the serial loop over contacts.
  const T* CON = c.template G<G_CON>();
  T *av = c.template S<O_AV>();      // scratch until the first Newton iteration: frame velocity J v of every contact
  int jt[2];
  T ax[2][3], an[2][3], qv[2];
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int j = c.lane + 32 * h;
    const bool ok = j < m.nv;
    jt[h] = ok ? m.jnt_type[j] : -1;
#pragma unroll
    for (int i = 0; i < 3; i++) { ax[h][i] = ok ? xaxis[3 * j + i] : T(0); an[h][i] = ok ? xanc[3 * j + i] : T(0); }
    qv[h] = ok ? qvel[j] : T(0);
  }
  unsigned long long mym1 = 0, mym2 = 0;     // (reloaded at the head of every chunk of 32 contacts)
  T Cn[12];
  if (c.ncon > 0) {
#pragma unroll
    for (int i = 0; i < 12; i++) Cn[i] = CON[1 + i];
  }
  for (int ci = 0; ci < c.ncon; ci++) {
    T C[12];
#pragma unroll
    for (int i = 0; i < 12; i++) C[i] = Cn[i];     // C[i] = record[1 + i]: position (3), frame rows (9)
    if (ci + 1 < c.ncon) {
#pragma unroll
      for (int i = 0; i < 12; i++) Cn[i] = CON[16 * (ci + 1) + 1 + i];
    }
    if ((ci & 31) == 0) {                    // warp-uniform
      mym1 = mym2 = 0;
      if (ci + c.lane < c.ncon) { mym1 = m.body_dofmask[c.template I<I_CB1>()[ci + c.lane]]; mym2 = m.body_dofmask[c.template I<I_CB2>()[ci + c.lane]]; }
    }
    const unsigned long long m1 = __shfl_sync(0xffffffffu, mym1, ci & 31), m2 = __shfl_sync(0xffffffffu, mym2, ci & 31);
    T vsum[4] = {0, 0, 0, 0};
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int j = c.lane + 32 * h;
      if (h == 1 && c.lane >= NVP - 32) break;
      T row[4] = {0, 0, 0, 0};
      if (jt[h] >= 0) {
        int sg = (int)((m2 >> j) & 1ull) - (int)((m1 >> j) & 1ull);
        if (sg != 0) {
          const T* a = ax[h];
          T jp[3], jr[3] = {0, 0, 0};
          if (jt[h] == J_SLIDE) { jp[0] = a[0]; jp[1] = a[1]; jp[2] = a[2]; }
          else {
            T r[3] = {C[0] - an[h][0], C[1] - an[h][1], C[2] - an[h][2]};
            cross3(jp, a, r);
            jr[0] = a[0]; jr[1] = a[1]; jr[2] = a[2];
          }
          T s = (T)sg;
          row[0] = s * (C[3] * jp[0] + C[4] * jp[1] + C[5] * jp[2]);
          row[1] = s * (C[6] * jp[0] + C[7] * jp[1] + C[8] * jp[2]);
          row[2] = s * (C[9] * jp[0] + C[10] * jp[1] + C[11] * jp[2]);
          row[3] = s * (C[3] * jr[0] + C[4] * jr[1] + C[5] * jr[2]);
#pragma unroll
          for (int a4 = 0; a4 < 4; a4++) vsum[a4] += row[a4] * qv[h];
        }
      }
#pragma unroll
      for (int a4 = 0; a4 < 4; a4++) JR[(4 * ci + a4) * NVP + j] = row[a4];
    }
    T t = warp_sum4(vsum[0], vsum[1], vsum[2], vsum[3], c.lane);
    if ((c.lane & 7) == 0) av[4 * ci + (c.lane >> 3)] = t;
  }
  __syncwarp();
  // per-contact scalars, lane = contact
  _Pragma("unroll 1") for (int ci = c.lane; ci < c.ncon; ci += 32) {
    const T* C = CON + 16 * ci;
    const int b1 = c.template I<I_CB1>()[ci], b2 = c.template I<I_CB2>()[ci], dim = c.template I<I_CDIM>()[ci];
    T dist = C[0], mg = C[13], imp = impedance(m, dist, mg);
    T tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
    T mu0 = cmu[3 * ci];
    T diag = dim == 1 ? tran : tran + mu0 * mu0 * tran;
    T R = tmax(T(1e-15), (1 - imp) * diag / imp);
    if (dim > 1) { T mu = mu0 * sqrt(T(1) / m.impratio); R = 2 * mu * mu * R; }
    cD[ci] = T(1) / R;
    carefn[ci] = -B * av[4 * ci] - K * imp * (dist - mg);
    for (int k = 0; k < 3; k++) cbt[3 * ci + k] = (k + 2 <= dim) ? -B * cmu[3 * ci + k] * av[4 * ci + k + 1] : T(0);
    for (int e = 0; e < 6; e++) ef[6 * ci + e] = 0;
  }
  __syncwarp();
}

#ifndef ADROIT_CFM_UNROLL
#define ADROIT_CFM_UNROLL 2
#endif
constexpr int CFM_UNROLL = ADROIT_CFM_UNROLL;
// J_r . x for a simple row
template <typename T> DI T simple_dot(const DModel<T>& m, int info, const T* x) {
  int kind = info & 3, id = (info >> 2) & 1023, up = (info >> 12) & 1;
  if (kind == S_FRICTION) return x[id];
  if (kind == S_JLIMIT) return up ? -x[id] : x[id];
  T s = 0;
  for (int k = 0; k < 2; k++) {
    int dj = m.tdof[2 * id + k];
    if (dj >= 0) s += m.tcoef[2 * id + k] * x[dj];
  }
  return up ? -s : s;
}
// contact-frame products  out[ci][a] = Jf[ci][a] . x   (lanes over dofs: coalesced row reads, the four rows of a
// contact reduced together)
template <class CX> SUB_FN void contact_frame_mul(CX c, const typename CX::real* x, typename CX::real* out) {
  using T = typename CX::real;
  assume_spaces(c);
  __builtin_assume(__isShared(x));
  if constexpr (CX::template in_smem<O_AC>()) __builtin_assume(__isShared(out)); else __builtin_assume(__isGlobal(out));   // O_AC / O_AV
  const T* JR = c.template G<G_JR>();
  const int nv = c.m.nv, lane = c.lane;
  const bool hi = nv > 32 && lane < NVP - 32;
  T x0, x1;
  load_vec(x, nv, lane, x0, x1);
  // the rows of the NEXT contact are in flight while the current one is reduced (J lives in L2)
  T jn[8];
  auto fetch = [&](int ci) {
    const T* J = JR + 4 * ci * NVP;
#pragma unroll
    for (int a = 0; a < 4; a++) jn[a] = J[a * NVP + lane];              // pad columns of J are zero
#pragma unroll
    for (int a = 0; a < 4; a++) jn[4 + a] = hi ? J[a * NVP + 32 + lane] : T(0);
  };
  if (c.ncon > 0) fetch(0);
#pragma unroll 1
  for (int ci = 0; ci < c.ncon; ci++) {
    T v[4];
#pragma unroll
    for (int a = 0; a < 4; a++) v[a] = jn[a] * x0 + jn[4 + a] * x1;
    if (ci + 1 < c.ncon) fetch(ci + 1);
    T t = warp_sum4(v[0], v[1], v[2], v[3], lane);
    if ((lane & 7) == 0) out[4 * ci + (lane >> 3)] = t;
  }
}
// edge e of contact ci: value of (Jn + sign*mu_k*Jt_k).x given the frame product fx[4]
template <typename T> DI T edge_val(const T* fx, const T* mu, int dim, int e, T& sign, int& k) {
  if (dim == 1) { sign = 0; k = 0; return fx[0]; }
  k = e >> 1;
  sign = (e & 1) ? T(-1) : T(1);
  return fx[0] + sign * mu[k] * fx[k + 1];
}

// frame force F[ci][4] (normal, 2 tangents, torsion) of every contact from its pyramid edge forces
template <class CX> DI void frame_forces(CX& c, typename CX::real* F) {
  using T = typename CX::real;
  const T *ef = c.template S<O_EF>(), *cmu = c.template S<O_CMU>();
  _Pragma("unroll 1") for (int ci = c.lane; ci < c.ncon; ci += 32) {
    const int dim = c.template I<I_CDIM>()[ci];
    const T* f = ef + 6 * ci;
    T e[6];
#pragma unroll
    for (int q = 0; q < 6; q++) e[q] = f[q];
    T out[4] = {e[0], 0, 0, 0};
    if (dim > 1) {
      out[0] = e[0] + e[1];
      out[1] = cmu[3 * ci] * (e[0] - e[1]);
      if (dim > 2) { out[0] += e[2] + e[3]; out[2] = cmu[3 * ci + 1] * (e[2] - e[3]); }
      if (dim > 3) { out[0] += e[4] + e[5]; out[3] = cmu[3 * ci + 2] * (e[4] - e[5]); }
    }
#pragma unroll
    for (int q = 0; q < 4; q++) F[4 * ci + q] = out[q];
  }
}
// qfc[j] += sum_r Jf[r][j] F[r]   (lanes over dofs; contacts without force are skipped)
template <class CX> DI void add_jt_force(CX& c, const typename CX::real* F, typename CX::real* qfc) {
  using T = typename CX::real;
  const T* JR = c.template G<G_JR>();
  const int nv = c.m.nv, lane = c.lane;
  T a0 = 0, a1 = 0;
#pragma unroll(CFM_UNROLL)
  for (int ci = 0; ci < c.ncon; ci++) {
    Vec4<T> f = ld4(F + 4 * ci);
    if (f.v[0] == 0 && f.v[1] == 0 && f.v[2] == 0 && f.v[3] == 0) continue;   // warp-uniform (shared-memory values)
    const T* J = JR + 4 * ci * NVP;
    a0 += J[lane] * f.v[0] + J[NVP + lane] * f.v[1] + J[2 * NVP + lane] * f.v[2] + J[3 * NVP + lane] * f.v[3];
    if (nv > 32 && lane < NVP - 32)
      a1 += J[32 + lane] * f.v[0] + J[NVP + 32 + lane] * f.v[1] + J[2 * NVP + 32 + lane] * f.v[2] + J[3 * NVP + 32 + lane] * f.v[3];
  }
  if (lane < nv) qfc[lane] += a0;
  if (lane + 32 < nv) qfc[lane + 32] += a1;
}

// cost, forces and states at the point whose frame accelerations are in O_AC and dof accelerations in `qacc`.
// Returns the constraint cost (all lanes).  Also accumulates J^T f into V_QFC when `want_qfc`.
template <typename T> struct CuResult { T cost; int nchanged, nflipped; };
template <class CX> SUB_FN CuResult<typename CX::real> constraint_update_fn(CX c, const typename CX::real* qacc, bool want_qfc) {
  using T = typename CX::real;
  assume_spaces(c);
  __builtin_assume(__isShared(qacc));
  const DModel<T>& m = c.m;
  int* sinfo = c.template I<I_SINFO>();
  const T *sD = c.template S<O_SD>(), *saref = c.template S<O_SAREF>(), *sfl = c.template S<O_SFL>();
  T *sforce = c.template S<O_SFORCE>(), *qfc = c.template V<V_QFC>();
  T cost = 0;
  bool changed = false;   // did any row / pyramid edge move between the quadratic and the inactive (or linear) zone?
  bool flipped = false;   // did a friction row jump from one linear zone to the other (through the quadratic one)?
  if (want_qfc) {
    _Pragma("unroll 1") for (int j = c.lane; j < m.nv; j += 32) qfc[j] = 0;
    __syncwarp();
  }
  _Pragma("unroll 1") for (int r = c.lane; r < c.nsimple; r += 32) {
    int info = sinfo[r] & 0x1fff, old_state = (sinfo[r] >> 13) & 1, old_side = (sinfo[r] >> 14) & 1;
    T jar = simple_dot(m, info, qacc) - saref[r], D = sD[r], f;
    int state, side = 0;   // side: which of the two linear zones of a friction row (no effect on the Hessian)
    if ((info & 3) == S_FRICTION) {
      T fl = sfl[r], rf = fl / D;
      if (jar <= -rf) { f = fl; state = 0; cost += -T(0.5) * rf * fl - fl * jar; }
      else if (jar >= rf) { f = -fl; state = 0; side = 1; cost += -T(0.5) * rf * fl + fl * jar; }
      else { f = -D * jar; state = 1; cost += T(0.5) * D * jar * jar; }
    } else if (jar < 0) { f = -D * jar; state = 1; cost += T(0.5) * D * jar * jar; }
    else { f = 0; state = 0; }
    sforce[r] = f;
    sinfo[r] = info | (state << 13) | (side << 14);
    changed |= state != old_state;
    flipped |= side != old_side;
    if (want_qfc && f != 0) {
      int kind = info & 3, id = (info >> 2) & 1023, up = (info >> 12) & 1;
      if (kind == S_FRICTION) atomicAdd(qfc + id, f);
      else if (kind == S_JLIMIT) atomicAdd(qfc + id, up ? -f : f);
      else
        for (int k = 0; k < 2; k++) {
          int dj = m.tdof[2 * id + k];
          if (dj >= 0) atomicAdd(qfc + dj, (up ? -f : f) * m.tcoef[2 * id + k]);
        }
    }
  }
  // contacts: one lane per contact walks its edges; builds the 4x4 Hessian weight W and the frame force
  const T *ac = c.template S<O_AC>(), *cD = c.template S<O_CD>(), *cmu = c.template S<O_CMU>(), *carefn = c.template S<O_CAREFN>(), *cbt = c.template S<O_CBT>();
  T *ef = c.template S<O_EF>(), *W = c.template S<O_W>();
  int* cmask = c.template I<I_CMASK>();
  _Pragma("unroll 1") for (int ci = c.lane; ci < c.ncon; ci += 32) {
    const int dim = c.template I<I_CDIM>()[ci], ne = n_edges<T>(dim);
    const T D = cD[ci], an = carefn[ci];
    const Vec4<T> fx = ld4(ac + 4 * ci);
    T w0 = 0, w1[3] = {0, 0, 0}, w4[3] = {0, 0, 0}, fe[6];
    int emask = 0;
#pragma unroll
    for (int e = 0; e < 6; e++) {      // compile-time edge index: every array stays in registers
      const int k = e >> 1;
      T f = 0;
      if (e < ne) {
        const T smu = dim == 1 ? T(0) : ((e & 1) ? -cmu[3 * ci + k] : cmu[3 * ci + k]);
        const T sgn = dim == 1 ? T(0) : ((e & 1) ? T(-1) : T(1));
        const T jar = (fx.v[0] + smu * fx.v[k + 1]) - (an + sgn * cbt[3 * ci + k]);
        if (jar < 0) {
          emask |= 1 << e;
          f = -D * jar;
          cost += T(0.5) * D * jar * jar;
          w0 += D;                                                   // W_nn
          if (dim > 1) { w1[k] += D * smu; w4[k] += D * smu * smu; }  // W_nk, W_kk
        }
      }
      fe[e] = f;
    }
#pragma unroll
    for (int e = 0; e < 6; e++) ef[6 * ci + e] = fe[e];
    T* w = W + 7 * ci;
    w[0] = w0; w[1] = w1[0]; w[2] = w1[1]; w[3] = w1[2]; w[4] = w4[0]; w[5] = w4[1]; w[6] = w4[2];
    changed |= emask != cmask[ci];
    cmask[ci] = emask;
  }
  c.nchanged = __any_sync(0xffffffffu, changed);
  c.nflipped = __any_sync(0xffffffffu, flipped);
  __syncwarp();
  if (want_qfc && c.ncon > 0) {
    // frame forces F[ci][4] into O_AV (scratch use), then qfc[j] += sum_r Jr[r][j] F[r]
    T* F = c.template S<O_AV>();
    frame_forces(c, F);
    __syncwarp();
    add_jt_force(c, F, qfc);
    __syncwarp();
  }
  CuResult<T> res;
  res.cost = warp_sum(cost);
  res.nchanged = c.nchanged;
  res.nflipped = c.nflipped;
  return res;
}
template <class CX> DI typename CX::real constraint_update(CX& c, const typename CX::real* qacc, bool want_qfc) {
  CuResult<typename CX::real> r = constraint_update_fn(c, qacc, want_qfc);
  c.nchanged = r.nchanged;
  c.nflipped = r.nflipped;
  return r.cost;
}

// Exact line search support.  Along qacc + alpha*search every row's residual is x + alpha*y; x, y and the row
// constants are computed ONCE per Newton iteration into registers (lane l owns simple rows l, l+32, l+64 and pyramid
// edges l, l+32, ..., l+160), so that evaluating (f', f'') at a new alpha is pure register arithmetic + two warp sums.
// NE = register sets for the pyramid edges of this kernel's contact capacity (3 for the 16-contact class, 6 for 32)
constexpr int NS_LS = (MAXSIMPLE + 31) / 32;
template <typename T, int NE> struct LsRows {
  T xs[NS_LS], ys[NS_LS], Ds[NS_LS], fs[NS_LS], rf[NS_LS];   // simple rows: D = 0 marks an absent row, fs < 0 a one-sided (limit) row
  T xe[NE], ye[NE], De[NE];                                   // pyramid edges
};
template <class CX> constexpr int ls_sets() { return (6 * (CX::MCAP < 32 ? CX::MCAP : 32) + 31) / 32; }
template <class CX> DI void ls_prepare(CX& c, const typename CX::real* qacc, const typename CX::real* search, LsRows<typename CX::real, ls_sets<CX>()>& R) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const int* sinfo = c.template I<I_SINFO>();
  const T *sD = c.template S<O_SD>(), *saref = c.template S<O_SAREF>(), *sfl = c.template S<O_SFL>();
#pragma unroll
  for (int q = 0; q < NS_LS; q++) {
    int r = c.lane + 32 * q;
    R.Ds[q] = 0; R.xs[q] = 0; R.ys[q] = 0; R.fs[q] = -1; R.rf[q] = 0;
    if (r < c.nsimple) {
      int info = sinfo[r] & 0x1fff;
      T D = sD[r];
      R.xs[q] = simple_dot(m, info, qacc) - saref[r];
      R.ys[q] = simple_dot(m, info, search);
      R.Ds[q] = D;
      if ((info & 3) == S_FRICTION) { R.fs[q] = sfl[r]; R.rf[q] = sfl[r] / D; }
    }
  }
  const T *ac = c.template S<O_AC>(), *av = c.template S<O_AV>(), *cD = c.template S<O_CD>(), *cmu = c.template S<O_CMU>();
  const T *carefn = c.template S<O_CAREFN>(), *cbt = c.template S<O_CBT>();
  const int* cdim = c.template I<I_CDIM>();
#pragma unroll
  for (int q = 0; q < ls_sets<CX>(); q++) {
    int t = c.lane + 32 * q;
    R.De[q] = 0; R.xe[q] = 0; R.ye[q] = 0;
    if (t < 6 * c.ncon) {
      int ci = t / 6, e = t - 6 * ci, dim = cdim[ci];
      if (e < n_edges<T>(dim)) {
        T sign;
        int k;
        T x = edge_val(ac + 4 * ci, cmu + 3 * ci, dim, e, sign, k);
        R.xe[q] = x - (carefn[ci] + sign * cbt[3 * ci + k]);
        R.ye[q] = edge_val(av + 4 * ci, cmu + 3 * ci, dim, e, sign, k);
        R.De[q] = cD[ci];
      }
    }
  }
  if constexpr (CX::MCAP > 32) {   // big class: the edges of contacts 32.. go through a scratch table in the context block
    T* ls = c.template G<G_LS>();
    _Pragma("unroll 1") for (int t = 32 * ls_sets<CX>() + c.lane; t < 6 * c.ncon; t += 32) {
      int ci = t / 6, e = t - 6 * ci, dim = cdim[ci];
      T x = 0, y = 0, D = 0;
      if (e < n_edges<T>(dim)) {
        T sign = 0;
        int k = 0;
        x = edge_val(ac + 4 * ci, cmu + 3 * ci, dim, e, sign, k) - (carefn[ci] + sign * cbt[3 * ci + k]);
        y = edge_val(av + 4 * ci, cmu + 3 * ci, dim, e, sign, k);
        D = cD[ci];
      }
      ls[t] = x; ls[6 * MAXCON + t] = y; ls[12 * MAXCON + t] = D;
    }
    __syncwarp();
  }
}
// the same for the edges kept in the context block (big class only): partial sums of (f', f'') before the warp reduction
template <class CX> DI void ls_eval_tail(CX& c, typename CX::real alpha, typename CX::real& a, typename CX::real& b) {
  using T = typename CX::real;
  if constexpr (CX::MCAP > 32) {
    const T* ls = c.template G<G_LS>();
    _Pragma("unroll 1") for (int t = 32 * ls_sets<CX>() + c.lane; t < 6 * c.ncon; t += 32) {
      T D = ls[12 * MAXCON + t], y = ls[6 * MAXCON + t], z = ls[t] + alpha * y;
      if (D > 0 && z < 0) { a += D * z * y; b += D * y * y; }
    }
  }
}
// derivative pair (f', f'') of the constraint cost at alpha
template <class CX, typename T, int NE> DI void ls_eval(CX& c, const LsRows<T, NE>& R, T alpha, T& d1, T& d2) {
  T a = 0, b = 0;
  ls_eval_tail(c, alpha, a, b);
#pragma unroll
  for (int q = 0; q < NS_LS; q++) {
    T D = R.Ds[q], y = R.ys[q], z = R.xs[q] + alpha * y;
    if (D > 0) {
      if (R.fs[q] >= 0) {
        if (z <= -R.rf[q]) a += -R.fs[q] * y;
        else if (z >= R.rf[q]) a += R.fs[q] * y;
        else { a += D * z * y; b += D * y * y; }
      } else if (z < 0) { a += D * z * y; b += D * y * y; }
    }
  }
#pragma unroll
  for (int q = 0; q < NE; q++) {
    T D = R.De[q], y = R.ye[q], z = R.xe[q] + alpha * y;
    if (D > 0 && z < 0) { a += D * z * y; b += D * y * y; }
  }
  d1 = warp_sum(a);
  d2 = warp_sum(b);
}

// H (packed, O_H) = M + J^T D_active J, using the states / weights left by constraint_update
template <class CX> PHASE_FN void build_hessian(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  T *H = c.template S<O_H>(), *M = c.template S<O_M>();
  for (int p = c.lane; p < NTRI; p += 32) H[p] = M[p];
  __syncwarp();
  const int* sinfo = c.template I<I_SINFO>();
  const T* sD = c.template S<O_SD>();
  _Pragma("unroll 1") for (int r = c.lane; r < c.nsimple; r += 32) {
    int info = sinfo[r];
    if (!((info >> 13) & 1)) continue;
    int kind = info & 3, id = (info >> 2) & 1023;
    T D = sD[r];
    if (kind != S_TLIMIT) atomicAdd(H + tri(id, id), D);
    else {
      int d0 = m.tdof[2 * id], d1 = m.tdof[2 * id + 1];
      T c0 = m.tcoef[2 * id], c1 = m.tcoef[2 * id + 1];
      if (d0 >= 0) atomicAdd(H + tri(d0, d0), D * c0 * c0);
      if (d1 >= 0) atomicAdd(H + tri(d1, d1), D * c1 * c1);
      if (d0 >= 0 && d1 >= 0) atomicAdd(H + (d0 > d1 ? tri(d0, d1) : tri(d1, d0)), D * c0 * c1);
    }
  }
  __syncwarp();
  const T *JR = c.template G<G_JR>(), *W = c.template S<O_W>();
  for (int ci = 0; ci < c.ncon; ci++) {
    const T* w = W + 7 * ci;
    if (w[0] == 0) continue;  // no active edge (warp-uniform: shared memory value)
    T v0[4], v1[4], g0[4], g1[4];
    int j0 = c.lane, j1 = c.lane + 32;
    bool nz0 = false, nz1 = false;
    for (int a = 0; a < 4; a++) {
      v0[a] = JR[(4 * ci + a) * NVP + j0];
      v1[a] = j1 < NVP ? JR[(4 * ci + a) * NVP + j1] : T(0);
      nz0 |= v0[a] != 0;
      nz1 |= v1[a] != 0;
    }
    g0[0] = w[0] * v0[0] + w[1] * v0[1] + w[2] * v0[2] + w[3] * v0[3];
    g1[0] = w[0] * v1[0] + w[1] * v1[1] + w[2] * v1[2] + w[3] * v1[3];
    for (int k = 0; k < 3; k++) {
      g0[1 + k] = w[1 + k] * v0[0] + w[4 + k] * v0[1 + k];
      g1[1 + k] = w[1 + k] * v1[0] + w[4 + k] * v1[1 + k];
    }
    unsigned mk0 = __ballot_sync(0xffffffffu, nz0), mk1 = __ballot_sync(0xffffffffu, nz1);
    while (mk0) {
      int i = __ffs(mk0) - 1;
      mk0 &= mk0 - 1;
      T s = 0;
      for (int a = 0; a < 4; a++) s += __shfl_sync(0xffffffffu, v0[a], i) * g0[a];
      if (j0 <= i && nz0) H[tri(i, j0)] += s;
    }
    while (mk1) {
      int i = __ffs(mk1) - 1;
      mk1 &= mk1 - 1;
      T s0 = 0, s1 = 0;
      for (int a = 0; a < 4; a++) {
        T vi = __shfl_sync(0xffffffffu, v1[a], i);
        s0 += vi * g0[a];
        s1 += vi * g1[a];
      }
      int ii = i + 32;
      if (nz0) H[tri(ii, j0)] += s0;
      if (j1 <= ii && nz1) H[tri(ii, j1)] += s1;
    }
    __syncwarp();
  }
}

#ifndef ADROIT_STOP_ON_QUADRATIC
#define ADROIT_STOP_ON_QUADRATIC 1
#endif
// fp32 stopping rules of the Newton solver (relative line-search gradient tolerance; cost-decrease floor below which an
// improvement is rounding noise)
#ifndef ADROIT_F32_LS_GTOL
#define ADROIT_F32_LS_GTOL 1e-6
#endif
#ifndef ADROIT_F32_COST_FLOOR
#define ADROIT_F32_COST_FLOOR 2e-7   // measured (tools/gpu_fp32_outliers.py): 2e-6 left single-step errors up to 8.7e-5 (hammer) / 4.6e-4 (pen) against
#endif                               // the fp64 oracle, 2e-7 gives 1.0e-6 / 4.6e-5 -- every agreeing env-step inside the north-star 1e-4 -- for 1 % of throughput
#ifndef ADROIT_LOCKSTEP_MASK
#define ADROIT_LOCKSTEP_MASK 15   // which of the four inner barriers of the lockstep iteration are kept (A/B, profiles/R2c_ab_lockstep.txt)
#endif
#ifndef ADROIT_LOCKSTEP
#define ADROIT_LOCKSTEP 2   // 0 = free-running warps, 1 = block barrier per Newton iteration, 2 = block barrier per sub-phase (measured: +10 % hammer)
#endif
// ---------------------------------------------------------------- Newton solver (primal, exact line search)
template <class CX> PHASE_FN void newton(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const int nv = m.nv, lane = c.lane;
  T *qacc = c.template V<V_QACC>(), *Ma = c.template V<V_MA>(), *grad = c.template V<V_GRAD>(), *search = c.template V<V_SEARCH>(), *Mv = c.template V<V_MV>(), *qfc = c.template V<V_QFC>();
  const T *qfs = c.template V<V_QFS>(), *qas = c.template V<V_QAS>(), *warm = c.template V<V_WARM>();
  T *M = c.template S<O_M>(), *H = c.template S<O_H>(), *ac = c.template S<O_AC>(), *av = c.template S<O_AV>();
  const T scale = T(1) / (m.meaninertia * (nv > 1 ? nv : 1));

  // cost at the warm start and at qacc_smooth (Gauss term vanishes there)
  auto gauss = [&](const T* a, const T* Maa) {
    T s = 0;
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) s += T(0.5) * (Maa[j] - qfs[j]) * (a[j] - qas[j]);
    return warp_sum(s);
  };
  // (qacc_smooth first, so that the states / forces / J^T f left behind are those of the warm start, the usual winner)
  contact_frame_mul(c, qas, ac);
  __syncwarp();
  T cost_sm = constraint_update(c, qas, false);
  symv<T, true, true>(M, warm, Ma, nv, lane);
  contact_frame_mul(c, warm, ac);
  __syncwarp();
  T cost = constraint_update(c, warm, true) + gauss(warm, Ma);
  bool use_ws = cost < cost_sm;
  _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) qacc[j] = use_ws ? warm[j] : qas[j];
  __syncwarp();
  if (!use_ws) {
    symv<T, true, true>(M, qacc, Ma, nv, lane);
    contact_frame_mul(c, qacc, ac);
    __syncwarp();
    cost = constraint_update(c, qacc, true) + gauss(qacc, Ma);
  }
  int iters = 0;
  bool refactor = true;
  bool done = false;
#if ADROIT_LOCKSTEP == 2
  // Block-lockstep Newton: the warps of a block walk the iteration in the same order of sub-phases with a block barrier
  // between them, so that the SM's instruction cache holds ONE sub-phase's code at a time instead of every warp's own
  // position in the ~60 KB loop (at full occupancy the kernel stalls 3.4 cycles per issued instruction on instruction
  // fetch, profiles/R2c_solve_hammer_full.summary.txt).  A warp whose env has converged idles at the barriers until the
  // block's last env is done.  Arithmetic and its order are those of the free-running loop.
  for (int iter = 0; iter < m.iterations; iter++) {
    if (!__syncthreads_or(!done)) break;
    bool act = !done;
    done = true;
    if (act) {
      T g2 = 0;
      _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) {
        T g = Ma[j] - qfs[j] - qfc[j];
        grad[j] = g;
        search[j] = -g;
        g2 += g * g;
      }
      g2 = warp_sum(g2);
      if (scale * sqrt(g2) < m.tolerance) act = false;
    }
    if (act && refactor) build_hessian(c);
    if (ADROIT_LOCKSTEP_MASK & 1) __syncthreads();
    if (act && refactor) chol_dense(H, c.template V<V_DINV>(), grad, Mv, nv, lane);
    if (ADROIT_LOCKSTEP_MASK & 2) __syncthreads();
    if (act) {
      chol_solve(H, c.template V<V_DINV>(), search, nv, lane);
      symv<T, true, true>(M, search, Mv, nv, lane);
      contact_frame_mul(c, search, av);
      __syncwarp();
    }
    if (ADROIT_LOCKSTEP_MASK & 4) __syncthreads();
    T alpha = 0;
    if (act) {
      T A0 = 0, B0 = 0;
      _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) { A0 += search[j] * Mv[j]; B0 += search[j] * (Ma[j] - qfs[j]); }
      A0 = warp_sum(A0);
      B0 = warp_sum(B0);
      T d1, d2;
      LsRows<T, ls_sets<CX>()> R;
      ls_prepare(c, qacc, search, R);
      ls_eval(c, R, T(0), d1, d2);
      T f1 = B0 + d1, f2 = A0 + d2;
      if (f1 < 0) {
        const T gtol = (sizeof(T) == 4 ? T(ADROIT_F32_LS_GTOL) : T(1e-13)) * (-f1);
        T lo = 0, hi = -1;  // hi < 0: unbounded
        alpha = -f1 / f2;
        for (int it = 0; it < 40; it++) {
          ls_eval(c, R, alpha, d1, d2);
          T fa = A0 * alpha + B0 + d1, fb = A0 + d2;
          if (tabs(fa) <= gtol) break;
          if (fa < 0) lo = alpha; else hi = alpha;
          T nxt = alpha - fa / fb;
          bool inside = nxt > lo && (hi < 0 || nxt < hi);
          if (!inside) nxt = hi < 0 ? 2 * alpha : T(0.5) * (lo + hi);
          if (hi >= 0 && (hi - lo) <= (sizeof(T) == 4 ? T(1e-7) : T(1e-15)) * hi) { alpha = nxt; break; }
          alpha = nxt;
        }
      }
      if (alpha == 0) act = false;
    }
    if (ADROIT_LOCKSTEP_MASK & 8) __syncthreads();
    if (act) {
      _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) { qacc[j] += alpha * search[j]; Ma[j] += alpha * Mv[j]; }
      _Pragma("unroll 1") for (int r = lane; r < 4 * c.ncon; r += 32) ac[r] += alpha * av[r];
      __syncwarp();
      T old = cost;
      cost = constraint_update(c, qacc, true) + gauss(qacc, Ma);
      refactor = c.nchanged != 0;
      iters = iter + 1;
      T floor_ = sizeof(T) == 4 ? T(ADROIT_F32_COST_FLOOR) * tabs(cost) : T(0);
      bool stop = scale * (old - cost) < m.tolerance || (old - cost) < floor_;
#if ADROIT_STOP_ON_QUADRATIC
      stop = stop || (!refactor && !c.nflipped);
#endif
      if (!stop) done = false;
    }
  }
#else
  for (int iter = 0; iter < m.iterations; iter++) {
#if ADROIT_LOCKSTEP
    // the warps of a block enter every Newton iteration together: one instruction-cache working set per block
    if (!__syncthreads_or(!done)) break;
    if (done) continue;
#else
    if (done) break;
#endif
    done = true;   // cleared at the end of a full iteration; every early exit below leaves it set
    do {
    T g2 = 0;
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) {
      T g = Ma[j] - qfs[j] - qfc[j];
      grad[j] = g;
      search[j] = -g;
      g2 += g * g;
    }
    g2 = warp_sum(g2);
    if (scale * sqrt(g2) < m.tolerance) break;
    if (refactor) {   // an unchanged active set leaves the Hessian (and its factor) of the previous iteration valid
      build_hessian(c);
      chol_dense(H, c.template V<V_DINV>(), grad, Mv, nv, lane);   // grad / Mv: free at this point (column broadcast buffers)
    }
    chol_solve(H, c.template V<V_DINV>(), search, nv, lane);
    symv<T, true, true>(M, search, Mv, nv, lane);
    contact_frame_mul(c, search, av);
    __syncwarp();
    // exact line search on the convex piecewise-quadratic: safeguarded Newton on f'(alpha)
    T A0 = 0, B0 = 0;
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) { A0 += search[j] * Mv[j]; B0 += search[j] * (Ma[j] - qfs[j]); }
    A0 = warp_sum(A0);
    B0 = warp_sum(B0);
    T d1, d2;
    LsRows<T, ls_sets<CX>()> R;
    ls_prepare(c, qacc, search, R);
    ls_eval(c, R, T(0), d1, d2);
    T f1 = B0 + d1, f2 = A0 + d2;
    T alpha = 0;
    if (f1 < 0) {
      const T gtol = (sizeof(T) == 4 ? T(ADROIT_F32_LS_GTOL) : T(1e-13)) * (-f1);
      T lo = 0, hi = -1;  // hi < 0: unbounded
      alpha = -f1 / f2;
      for (int it = 0; it < 40; it++) {
        ls_eval(c, R, alpha, d1, d2);
        T fa = A0 * alpha + B0 + d1, fb = A0 + d2;
        if (tabs(fa) <= gtol) break;
        if (fa < 0) lo = alpha; else hi = alpha;
        T nxt = alpha - fa / fb;
        bool inside = nxt > lo && (hi < 0 || nxt < hi);
        if (!inside) nxt = hi < 0 ? 2 * alpha : T(0.5) * (lo + hi);
        if (hi >= 0 && (hi - lo) <= (sizeof(T) == 4 ? T(1e-7) : T(1e-15)) * hi) { alpha = nxt; break; }
        alpha = nxt;
      }
    }
    if (alpha == 0) break;
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) { qacc[j] += alpha * search[j]; Ma[j] += alpha * Mv[j]; }
    _Pragma("unroll 1") for (int r = lane; r < 4 * c.ncon; r += 32) ac[r] += alpha * av[r];
    __syncwarp();
    T old = cost;
    cost = constraint_update(c, qacc, true) + gauss(qacc, Ma);
    refactor = c.nchanged != 0;
    iters = iter + 1;
    // fp32: an improvement below the rounding noise of the cost itself (~1e-6 relative) cannot be measured; Newton's
    // quadratic convergence makes the step after such an iteration a no-op, so stop there instead of burning a factorisation
    T floor_ = sizeof(T) == 4 ? T(ADROIT_F32_COST_FLOOR) * tabs(cost) : T(0);
    if (scale * (old - cost) < m.tolerance || (old - cost) < floor_) break;
#if ADROIT_STOP_ON_QUADRATIC
    // no row / edge changed state along the step: the cost was exactly quadratic on it and the Newton step landed on
    // its minimiser, so the next iteration could only polish rounding error
    if (!refactor && !c.nflipped) break;
#endif
    done = false;
    } while (0);
  }
#endif
  if (lane == 0) c.gi[GI_NITER] = iters;
}

// ---------------------------------------------------------------- noslip (Gauss-Seidel on friction rows, matrix-free)
// The sweep is MuJoCo's row-by-row Gauss-Seidel, evaluated without walking the rows one at a time:
//   * u = M^-1 J^T f lives in two registers per lane (dof lane, lane + 32) for the whole pass;
//   * friction-loss rows: every lane evaluates the update of ITS row against the current u; the first row (in row
//     order) whose force would change is applied (one column of M^-1 per lane) and the evaluation repeats behind it.
//     Rows in front of the first change are untouched by definition, so this is exactly the serial sweep, at one
//     round per CHANGED row instead of one per row (saturated friction rows -- moving joints -- never change);
//   * contact tangents: a compact list of tangent rows; the Jacobian row and its M^-1 image are prefetched one
//     tangent ahead, so the global loads are off the dependent chain  dot -> warp sum -> clamp -> update.
// a / b on the dependent Gauss-Seidel chain: fp32 multiplies by the reciprocal prepared off the chain (MUFU.RCP + one
// Newton step, ~1 ulp; the IEEE division is a ~10-instruction dependent sequence), fp64 divides
DI float quot(float a, float, float rb) { return a * rb; }
DI double quot(double a, double b, double) { return a / b; }
template <class CX> PHASE_FN void noslip(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const int nv = m.nv, lane = c.lane;
  const unsigned FULL = 0xffffffffu;
  const T *Minv = c.template G<G_MINV>(), *JR = c.template G<G_JR>(), *qas = c.template V<V_QAS>();
  T *u = c.template V<V_GRAD>(), *tmp = c.template V<V_SEARCH>(), *qfc = c.template V<V_QFC>(), *YT = c.template G<G_YT>();
  const int* sinfo = c.template I<I_SINFO>();
  const T *saref = c.template S<O_SAREF>(), *sfl = c.template S<O_SFL>(), *cmu = c.template S<O_CMU>(), *cbt = c.template S<O_CBT>();
  T *sforce = c.template S<O_SFORCE>(), *ef = c.template S<O_EF>();
  int* trow = c.template I<I_SURV>();          // compact tangent list: contact-frame row 4*ci + 1 + k
  // u = M^-1 J^T f  (qfc holds J^T f from the last constraint_update)
  symv(Minv, qfc, u, nv, lane);
  // Y[t] = M^-1 Jt^T and K1[t] = 4 mu^2 Jt.Y[t] for every tangent row t
  T* K1 = c.template S<O_AV>();
  int ntan = 0;
  for (int ci = 0; ci < c.ncon; ci++) {
    int dim = c.template I<I_CDIM>()[ci];
    for (int k = 0; k + 2 <= dim && k < 3; k++) {
      if (lane == 0) trow[ntan] = 4 * ci + 1 + k;
      ntan++;
    }
  }
  __syncwarp();
  {  // the Jacobian row of the NEXT tangent is in flight (L2) while the current one is multiplied
    T jn0 = 0, jn1 = 0;
    if (ntan > 0) load_vec(JR + trow[0] * NVP, nv, lane, jn0, jn1);
    _Pragma("unroll 1") for (int t = 0; t < ntan; t++) {
      const T j0 = jn0, j1 = jn1;
      const int row = trow[t], ci = row >> 2, k = (row & 3) - 1;
      if (t + 1 < ntan) load_vec(JR + trow[t + 1] * NVP, nv, lane, jn0, jn1);
      store_vec(tmp, nv, lane, j0, j1);
      __syncwarp();
      T* Y = YT + t * NVP;
      symv_range(Minv, tmp, Y, nv, lane);
      __syncwarp();
      T dsum = (lane < nv ? Y[lane] * j0 : T(0)) + (lane + 32 < nv ? Y[lane + 32] * j1 : T(0));
      dsum = warp_sum(dsum);
      if (lane == 0) K1[t] = 4 * cmu[3 * ci + k] * cmu[3 * ci + k] * dsum;
    }
  }
  __syncwarp();
  // registers: u, qacc_smooth by dof; the lane's friction rows (row lane -> set 0, row lane + 32 -> set 1)
  T u0, u1, q0, q1;
  load_vec(u, nv, lane, u0, u1);
  load_vec(qas, nv, lane, q0, q1);
  bool val[2];
  int dj[2];
  T qd[2], sa[2], Aii[2], rA[2], fl[2], f[2];   // rA = 1 / Aii (fp32: the Gauss-Seidel chain multiplies, see quot())
#pragma unroll
  for (int s = 0; s < 2; s++) {
    int r = lane + 32 * s;
    int info = r < c.nsimple ? sinfo[r] : -1;
    val[s] = r < c.nsimple && (info & 3) == S_FRICTION;
    dj[s] = val[s] ? (info >> 2) & 1023 : 0;
    qd[s] = qas[dj[s]];
    sa[s] = val[s] ? saref[r] : T(0);
    Aii[s] = Minv[tri(dj[s], dj[s])];
    rA[s] = rcp_t(Aii[s] > tiny<T>() ? Aii[s] : tiny<T>());
    fl[s] = val[s] ? sfl[r] : T(0);
    f[s] = val[s] ? sforce[r] : T(0);
  }
  const bool ident = __all_sync(FULL, (!val[0] || dj[0] == lane) && (!val[1] || dj[1] == lane + 32));   // row r <-> dof r
  const bool any1 = __any_sync(FULL, val[1]);
  const T scale = T(1) / (m.meaninertia * (nv > 1 ? nv : 1));
  int sweeps = 0;
  for (int iter = 0; iter < m.noslip_iterations; iter++) {
    T improvement = 0;
    int start = 0;
    while (true) {
      T res[2] = {0, 0}, fn[2] = {0, 0}, de[2] = {0, 0};
      bool nz[2] = {false, false};
      auto eval = [&](int s) {
        T ud;
        if (ident) ud = s ? u1 : u0;
        else {
          T a = __shfl_sync(FULL, u0, dj[s] & 31), b = __shfl_sync(FULL, u1, dj[s] & 31);
          ud = dj[s] < 32 ? a : b;
        }
        res[s] = ud + qd[s] - sa[s];
        fn[s] = tclamp(f[s] - quot(res[s], Aii[s], rA[s]), -fl[s], fl[s]);
        de[s] = fn[s] - f[s];
        nz[s] = val[s] && lane + 32 * s >= start && de[s] != 0;
      };
      eval(0);
      if (any1) eval(1);
      unsigned b0 = __ballot_sync(FULL, nz[0]), b1 = any1 ? __ballot_sync(FULL, nz[1]) : 0u;
      if (!(b0 | b1)) break;
      const int set = b0 ? 0 : 1, src = __ffs(b0 ? b0 : b1) - 1;
      const T d = __shfl_sync(FULL, set ? de[1] : de[0], src), rs = __shfl_sync(FULL, set ? res[1] : res[0], src);
      const T A = __shfl_sync(FULL, set ? Aii[1] : Aii[0], src);
      const int col = __shfl_sync(FULL, set ? dj[1] : dj[0], src);
      if (lane < nv) u0 += d * sym_get(Minv, lane, col);
      if (lane + 32 < nv) u1 += d * sym_get(Minv, lane + 32, col);
      if (lane == src) { if (set) f[1] = fn[1]; else f[0] = fn[0]; }
      improvement -= T(0.5) * d * d * A + d * rs;
      start = src + 32 * set + 1;
    }
    if (ntan > 0) {
      const T *Jn = JR + trow[0] * NVP;
      T jn0, jn1, yn0, yn1;
      load_vec(Jn, nv, lane, jn0, jn1);
      load_vec(YT, nv, lane, yn0, yn1);
      T k1n = K1[0], rkn = rcp_t(k1n > tiny<T>() ? k1n : tiny<T>());
      _Pragma("unroll 1") for (int t = 0; t < ntan; t++) {
        const T j0 = jn0, j1 = jn1, y0 = yn0, y1 = yn1, k1 = k1n, rk = rkn;
        const int row = trow[t], ci = row >> 2, k = (row & 3) - 1;
        if (t + 1 < ntan) {
          load_vec(JR + trow[t + 1] * NVP, nv, lane, jn0, jn1);
          load_vec(YT + (t + 1) * NVP, nv, lane, yn0, yn1);
          k1n = K1[t + 1];
          rkn = rcp_t(k1n > tiny<T>() ? k1n : tiny<T>());     // the reciprocal is ready before the dependent chain needs it
        }
        T s = j0 * (u0 + q0);
        if (nv > 32) s += j1 * (u1 + q1);
        s = warp_sum(s);
        T mu = cmu[3 * ci + k], g = 2 * mu * s - 2 * cbt[3 * ci + k];
        T f0 = ef[6 * ci + 2 * k], f1 = ef[6 * ci + 2 * k + 1], mid = T(0.5) * (f0 + f1), yold = T(0.5) * (f0 - f1), y = yold;
        if (k1 >= T(1e-15)) y = tclamp(yold - quot(g, k1, rk), -mid, mid);
        else y = 0;
        T dy = y - yold;
        __syncwarp();
        if (dy != 0) {
          u0 += 2 * mu * dy * y0;
          u1 += 2 * mu * dy * y1;
          if (lane == 0) { ef[6 * ci + 2 * k] = mid + y; ef[6 * ci + 2 * k + 1] = mid - y; }
          improvement -= T(0.5) * k1 * dy * dy + g * dy;
          __syncwarp();
        }
      }
    }
    sweeps = iter + 1;
    if (scale * improvement < m.noslip_tolerance) break;
  }
#pragma unroll
  for (int s = 0; s < 2; s++)
    if (val[s]) sforce[lane + 32 * s] = f[s];
  if (lane == 0) c.gi[GI_NSLIP] = sweeps;
  __syncwarp();
}

// qfrc_constraint = J^T f from the stored forces (V_QFC)
template <class CX> PHASE_FN void constraint_force(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  T* qfc = c.template V<V_QFC>();
  // frame-force scratch [ncon][4]: the O_AV region -- unless the kernel stages it too small for that (the finish kernel keeps
  // only the noslip pass's K1[3 * MAXCON_M] there); then the staged, by now dead, saref | floss rows serve (adjacent, 2 * MAXSIMPLE)
  T* F;
  if constexpr (map_rlen<typename CX::MAP>(O_AV) > 0 && map_rlen<typename CX::MAP>(O_AV) < 4 * MAXCON_M) {
    static_assert(CX::template in_smem<O_SAREF>() && CX::template in_smem<O_SFL>() && O_SFL == O_SAREF + MAXSIMPLE && 2 * MAXSIMPLE >= 4 * MAXCON_M, "");
    F = c.template S<O_SAREF>();
  } else {
    F = c.template S<O_AV>();
  }
  const int* sinfo = c.template I<I_SINFO>();
  const T *sforce = c.template S<O_SFORCE>(), *ef = c.template S<O_EF>(), *cmu = c.template S<O_CMU>();
  _Pragma("unroll 1") for (int j = c.lane; j < m.nv; j += 32) qfc[j] = 0;
  __syncwarp();
  _Pragma("unroll 1") for (int r = c.lane; r < c.nsimple; r += 32) {
    T f = sforce[r];
    if (f == 0) continue;
    int info = sinfo[r], kind = info & 3, id = (info >> 2) & 1023, up = (info >> 12) & 1;
    if (kind == S_FRICTION) atomicAdd(qfc + id, f);
    else if (kind == S_JLIMIT) atomicAdd(qfc + id, up ? -f : f);
    else
      for (int k = 0; k < 2; k++) {
        int dj = m.tdof[2 * id + k];
        if (dj >= 0) atomicAdd(qfc + dj, (up ? -f : f) * m.tcoef[2 * id + k]);
      }
  }
  frame_forces(c, F);
  __syncwarp();
  add_jt_force(c, F, qfc);
  __syncwarp();
}

// touch sensor S_nail (hammer_v0.py:102): normal forces of contacts on the site's body whose contact ray hits the site cylinder
template <typename T> DI bool ray_hits_cylinder(const T* spos, const T* smat, const T* size, const T* pnt, const T* vec) {
  T t[3] = {pnt[0] - spos[0], pnt[1] - spos[1], pnt[2] - spos[2]}, o[3], v[3];
  for (int i = 0; i < 3; i++) {
    o[i] = smat[i] * t[0] + smat[3 + i] * t[1] + smat[6 + i] * t[2];
    v[i] = smat[i] * vec[0] + smat[3 + i] * vec[1] + smat[6 + i] * vec[2];
  }
  T r = size[0], h = size[1];
  if (o[0] * o[0] + o[1] * o[1] <= r * r && tabs(o[2]) <= h) return true;
  T a = v[0] * v[0] + v[1] * v[1], b = o[0] * v[0] + o[1] * v[1], cc = o[0] * o[0] + o[1] * o[1] - r * r;
  if (a > tiny<T>()) {
    T det = b * b - a * cc;
    if (det >= 0) {
      T sq = sqrt(det);
      for (int k = -1; k <= 1; k += 2) {
        T x = (-b + k * sq) / a;
        if (x >= 0 && tabs(o[2] + x * v[2]) <= h) return true;
      }
    }
  }
  if (tabs(v[2]) > tiny<T>())
    for (int k = -1; k <= 1; k += 2) {
      T x = (k * h - o[2]) / v[2], px = o[0] + x * v[0], py = o[1] + x * v[1];
      if (x >= 0 && px * px + py * py <= r * r) return true;
    }
  return false;
}
template <class CX> PHASE_FN typename CX::real touch_sensor(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  if (m.touch_site < 0) return T(0);
  int sb = m.site_body[m.touch_site];
  const T *ef = c.template S<O_EF>(), *spos = c.template G<G_SPOS>() + 3 * m.touch_site, *smat = c.template S<O_XMAT>() + 9 * sb;
  T sum = 0;
  _Pragma("unroll 1") for (int ci = c.lane; ci < c.ncon; ci += 32) {
    int b1 = c.template I<I_CB1>()[ci], b2 = c.template I<I_CB2>()[ci], dim = c.template I<I_CDIM>()[ci];
    if (sb != b1 && sb != b2) continue;
    T fn = 0;
    for (int e = 0; e < n_edges<T>(dim); e++) fn += ef[6 * ci + e];
    if (fn <= 0) continue;
    const T* C = c.template G<G_CON>() + 16 * ci;
    T sgn = sb == b2 ? T(-1) : T(1), ray[3] = {sgn * C[4], sgn * C[5], sgn * C[6]};
    if (ray_hits_cylinder(spos, smat, m.touch_size, C + 1, ray)) sum += fn;
  }
  return warp_sum(sum);
}

}  // namespace adroit
