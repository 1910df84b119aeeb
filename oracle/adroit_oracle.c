/* adroit_oracle.c — fp64 CPU restatement of the reference's Adroit step path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library; the product (mj_envs_b200/) never does.
 *
 * PARITY UNPINNED.  The arithmetic of the reference's hot path lives in third-party code that is
 * absent from /root/reference: MuJoCo 2.1.0 (`mujoco210`, Dockerfile:53) driven through mujoco-py
 * >=2.1.2,<2.2 (setup.py:19) and mjrl's MujocoEnv (Dockerfile:80).  None of them can be installed
 * here and the reference holds no tests / golden vectors.  This file restates
 * the published MuJoCo 2.1 semantics that the four Adroit scenes exercise with
 * textbook algorithms (world-frame spatial algebra, dense Cholesky), and executes the reference's
 * task layer (hammer_v0.py, door_v0.py, pen_v0.py, relocate_v0.py, utils/quatmath.py) verbatim in
 * semantics.  Trust is earned through physics invariants (tests/test_oracle_*.py) and through the
 * reference's own weak pins (tests/golden/).
 *
 * Exports the C ABI of include/adroit_b200.h with an `oracle_` prefix and host pointers.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAXV 40
#define MAXB 36
#define MAXG 40
#define MAXS 36
#define MAXU 32
#define MAXT 48
#define MAXCON 100   /* nconmax, DAPG_assets.xml:4 */
#define MAXEFC 500   /* njmax,   DAPG_assets.xml:4 */
#define MAXPAR 8
#define MINVAL 1e-15
#define MINIMP 0.0001
#define MAXIMP 0.9999

enum { G_PLANE = 0, G_SPHERE = 2, G_CAPSULE = 3, G_CYLINDER = 5, G_BOX = 6 };
enum { J_SLIDE = 2, J_HINGE = 3 };
enum { T_HAMMER = 0, T_DOOR = 1, T_PEN = 2, T_RELOCATE = 3 };
/* hammer variations (hammer_v0.py:110-129): task_int[8] = type, [9] = Object body, [10] = head geom, [11] = neck geom; values in par[3], par[4] */
enum { VAR_NONE = 0, VAR_MASS = 1, VAR_POS = 2, VAR_SIZE = 3 };
enum { C_FRICTION = 0, C_LIMIT_JOINT = 1, C_LIMIT_TENDON = 2, C_CONTACT_FRICTIONLESS = 3, C_CONTACT_PYRAMIDAL = 4 };

static __thread char g_err[256];
const char* oracle_last_error(void) { return g_err; }
#define FAIL(...) do { snprintf(g_err, sizeof g_err, __VA_ARGS__); return 1; } while (0)

/* ------------------------------------------------------------------ model */
typedef struct {
  int nv, nbody, ngeom, nsite, nu, ntendon, npair, npar, obs_dim, frame_skip, task, horizon;
  const int *body_parent, *body_jntadr, *body_jntnum, *body_weld, *body_par, *site_par;
  const double *body_pos, *body_quat, *body_ipos, *body_iquat, *body_mass, *body_inertia, *body_invweight0;
  const int *jnt_type, *jnt_limited, *dof_body, *dof_parent;
  const double *jnt_pos, *jnt_axis, *jnt_range, *jnt_margin, *dof_damping, *dof_armature, *dof_frictionloss,
      *dof_invweight0;
  const int *geom_type, *geom_body;
  const double *geom_pos, *geom_quat, *geom_size, *geom_rbound;
  const int *site_body;
  const double *site_pos, *site_size;
  const int *pair_geom, *pair_condim;
  const double *pair_friction, *pair_margin;
  const int *tendon_dof, *tendon_limited;
  const double *tendon_coef, *tendon_range, *tendon_margin, *tendon_invweight0;
  const int* act_dof;
  const double *act_ctrlrange, *act_gain, *act_bias;
  const int* task_int;
  const double* par0;
  int touch_site;
  double timestep, gravity[3], tolerance, noslip_tolerance, impratio, solref[2], solimp[5], mpr_tolerance,
      meaninertia;
  int iterations, noslip_iterations, mpr_iterations, ls_iterations;
  void* blob;
} OModel;

static const void* blob_find(const unsigned char* b, const char* name, int want_dt, int* count) {
  uint64_t nf;
  memcpy(&nf, b + 8, 8);
  for (uint64_t i = 0; i < nf; i++) {
    const unsigned char* e = b + 16 + 48 * i;
    if (strncmp((const char*)e, name, 32) == 0) {
      uint32_t dt, cnt;
      uint64_t off;
      memcpy(&dt, e + 32, 4);
      memcpy(&cnt, e + 36, 4);
      memcpy(&off, e + 40, 8);
      if ((int)dt != want_dt) return NULL;
      if (count) *count = (int)cnt;
      return b + off;
    }
  }
  return NULL;
}

static int model_from_blob(OModel* m, const void* blob, size_t len) {
  if (len < 16 || memcmp(blob, "ADRB200\x01", 8) != 0) FAIL("bad model blob");
  m->blob = malloc(len);
  memcpy(m->blob, blob, len);
  const unsigned char* b = (const unsigned char*)m->blob;
#define GI(f) if (!(m->f = (const int*)blob_find(b, #f, 0, NULL))) FAIL("blob misses int field %s", #f)
#define GR(f) if (!(m->f = (const double*)blob_find(b, #f, 1, NULL))) FAIL("blob misses real field %s", #f)
  const int* sizes = (const int*)blob_find(b, "sizes", 0, NULL);
  if (!sizes) FAIL("blob misses sizes");
  m->nv = sizes[0]; m->nbody = sizes[1]; m->ngeom = sizes[2]; m->nsite = sizes[3]; m->nu = sizes[4];
  m->ntendon = sizes[5]; m->npair = sizes[6]; m->npar = sizes[7]; m->obs_dim = sizes[8]; m->frame_skip = sizes[9];
  m->task = sizes[10]; m->horizon = sizes[11];
  if (m->nv > MAXV || m->nbody > MAXB || m->ngeom > MAXG || m->nsite > MAXS || m->nu > MAXU || m->ntendon > MAXT ||
      m->npar > MAXPAR)
    FAIL("model exceeds oracle capacity");
  GI(body_parent); GI(body_jntadr); GI(body_jntnum); GI(body_weld); GI(body_par); GI(site_par);
  GR(body_pos); GR(body_quat); GR(body_ipos); GR(body_iquat); GR(body_mass); GR(body_inertia); GR(body_invweight0);
  GI(jnt_type); GI(jnt_limited); GI(dof_body); GI(dof_parent);
  GR(jnt_pos); GR(jnt_axis); GR(jnt_range); GR(jnt_margin); GR(dof_damping); GR(dof_armature);
  GR(dof_frictionloss); GR(dof_invweight0);
  GI(geom_type); GI(geom_body); GR(geom_pos); GR(geom_quat); GR(geom_size); GR(geom_rbound);
  GI(site_body); GR(site_pos); GR(site_size);
  GI(pair_geom); GI(pair_condim); GR(pair_friction); GR(pair_margin);
  GI(tendon_dof); GI(tendon_limited); GR(tendon_coef); GR(tendon_range); GR(tendon_margin); GR(tendon_invweight0);
  GI(act_dof); GR(act_ctrlrange); GR(act_gain); GR(act_bias);
  GI(task_int); GR(par0);
  const int* ts = (const int*)blob_find(b, "touch_site", 0, NULL);
  m->touch_site = ts ? ts[0] : -1;
  const double* o = (const double*)blob_find(b, "opt_real", 1, NULL);
  const int* oi = (const int*)blob_find(b, "opt_int", 0, NULL);
  const double* mi = (const double*)blob_find(b, "meaninertia", 1, NULL);
  if (!o || !oi || !mi) FAIL("blob misses options");
  m->timestep = o[0]; m->gravity[0] = o[1]; m->gravity[1] = o[2]; m->gravity[2] = o[3];
  m->tolerance = o[4]; m->noslip_tolerance = o[5]; m->impratio = o[6];
  m->solref[0] = o[7]; m->solref[1] = o[8];
  for (int i = 0; i < 5; i++) m->solimp[i] = o[9 + i];
  m->mpr_tolerance = o[14];
  m->iterations = oi[0]; m->noslip_iterations = oi[1]; m->mpr_iterations = oi[2]; m->ls_iterations = oi[3];
  m->meaninertia = mi[0];
  return 0;
}

/* ------------------------------------------------------------------ small math */
static inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline void sub3(double* r, const double* a, const double* b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
static inline void add3(double* r, const double* a, const double* b) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
static inline void addscl3(double* r, const double* a, const double* b, double s) { r[0] = a[0] + s * b[0]; r[1] = a[1] + s * b[1]; r[2] = a[2] + s * b[2]; }
static inline void scl3(double* r, const double* a, double s) { r[0] = a[0] * s; r[1] = a[1] * s; r[2] = a[2] * s; }
static inline void copy3(double* r, const double* a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
static inline double norm3(const double* a) { return sqrt(dot3(a, a)); }
static inline double normalize3(double* a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
  a[0] /= n; a[1] /= n; a[2] /= n;
  return n;
}
static void quat_mul(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void quat_norm(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  for (int i = 0; i < 4; i++) q[i] /= n;
}
/* row-major 3x3; columns are the rotated basis vectors */
static void quat2mat(double* R, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}
static inline void mulMV(double* r, const double* R, const double* v) {
  double x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2], y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2],
         z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline void mulMTV(double* r, const double* R, const double* v) {
  double x = R[0] * v[0] + R[3] * v[1] + R[6] * v[2], y = R[1] * v[0] + R[4] * v[1] + R[7] * v[2],
         z = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void mulMM(double* C, const double* A, const double* B) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(C, t, sizeof t);
}

/* ------------------------------------------------------------------ per-env data */
typedef struct {
  double pos[3], frame[9], dist, friction[5], includemargin;
  int dim, geom1, geom2, efc_address;
} OContact;

typedef struct {
  /* persistent state */
  double qpos[MAXV], qvel[MAXV], qacc_warmstart[MAXV], par[MAXPAR];
  int t, episode;
  double ep_return;
  int ep_goal;
  /* finished-episode accumulators */
  double acc_return;
  long acc_any, acc_thr, acc_n;
  /* forward results */
  double ctrl[MAXU];
  double xpos[MAXB][3], xquat[MAXB][4], xmat[MAXB][9], xipos[MAXB][3], ximat[MAXB][9];
  double xanchor[MAXV][3], xaxis[MAXV][3];
  double gpos[MAXG][3], gmat[MAXG][9], spos[MAXS][3], smat[MAXS][9];
  double M[MAXV][MAXV], L[MAXV][MAXV];
  double qfrc_bias[MAXV], qfrc_passive[MAXV], qfrc_actuator[MAXV], qfrc_smooth[MAXV], qacc_smooth[MAXV],
      qfrc_constraint[MAXV], qacc[MAXV];
  int ncon, nefc;
  OContact con[MAXCON];
  double touch;
  int solver_iter;
  double var_size[3];
} OData;

/* solver workspace (per thread) */
typedef struct {
  int type[MAXEFC], id[MAXEFC];
  double J[MAXEFC][MAXV], pos[MAXEFC], margin[MAXEFC], floss[MAXEFC], diagApprox[MAXEFC], R[MAXEFC], D[MAXEFC],
      aref[MAXEFC], vel[MAXEFC], force[MAXEFC], jar[MAXEFC], b[MAXEFC];
  int state[MAXEFC];
  double (*AR)[MAXEFC]; /* nefc x nefc, heap */
} OWork;

typedef struct {
  OModel m;
  int n;
  uint64_t seed;
  int64_t env_off;
  uint32_t flags;
  OData* d;
  float* last_obs;
} OSim;

/* ------------------------------------------------------------------ kinematics */
/* per-env overrides of model fields written by the hammer variations (the model is NOT recompiled: inertia tensor,
   invweight0 and geom_rbound stay stale, hammer_v0.py:115-125 / SURVEY.md App. A.8 Q9) */
static double body_mass_of(const OModel* m, const OData* d, int b) {
  return (m->task_int[8] == VAR_MASS && b == m->task_int[9]) ? d->par[3] : m->body_mass[b];
}
static const double* geom_size_of(const OModel* m, OData* d, int g) {
  if (m->task_int[8] == VAR_SIZE && g == m->task_int[10]) {
    d->var_size[0] = d->par[3]; d->var_size[1] = d->par[4]; d->var_size[2] = m->geom_size[3 * g + 2];
    return d->var_size;
  }
  return m->geom_size + 3 * g;
}

static void o_kinematics(const OModel* m, OData* d) {
  memset(d->xpos[0], 0, sizeof d->xpos[0]);
  d->xquat[0][0] = 1; d->xquat[0][1] = d->xquat[0][2] = d->xquat[0][3] = 0;
  quat2mat(d->xmat[0], d->xquat[0]);
  memset(d->xipos[0], 0, sizeof d->xipos[0]);
  quat2mat(d->ximat[0], d->xquat[0]);
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parent[b];
    double bp[3], bq[4], pos[3], quat[4], R[9], t[3];
    if (m->body_par[2 * b] >= 0) copy3(bp, d->par + m->body_par[2 * b]); else copy3(bp, m->body_pos + 3 * b);
    if (m->body_par[2 * b + 1] >= 0) memcpy(bq, d->par + m->body_par[2 * b + 1], 4 * sizeof(double));
    else memcpy(bq, m->body_quat + 4 * b, 4 * sizeof(double));
    quat_norm(bq);
    mulMV(t, d->xmat[p], bp);
    add3(pos, d->xpos[p], t);
    quat_mul(quat, d->xquat[p], bq);
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k;
      quat2mat(R, quat);
      mulMV(t, R, m->jnt_pos + 3 * j);
      add3(d->xanchor[j], pos, t);
      mulMV(d->xaxis[j], R, m->jnt_axis + 3 * j);
      if (m->jnt_type[j] == J_SLIDE) {
        addscl3(pos, pos, d->xaxis[j], d->qpos[j]);
      } else {
        double h = 0.5 * d->qpos[j], s = sin(h), qj[4] = {cos(h), s * m->jnt_axis[3 * j], s * m->jnt_axis[3 * j + 1], s * m->jnt_axis[3 * j + 2]};
        quat_mul(quat, quat, qj);
        quat2mat(R, quat);
        mulMV(t, R, m->jnt_pos + 3 * j);
        sub3(pos, d->xanchor[j], t);
      }
    }
    quat_norm(quat);
    copy3(d->xpos[b], pos);
    memcpy(d->xquat[b], quat, sizeof quat);
    quat2mat(d->xmat[b], quat);
    mulMV(t, d->xmat[b], m->body_ipos + 3 * b);
    add3(d->xipos[b], pos, t);
    quat2mat(R, m->body_iquat + 4 * b);
    mulMM(d->ximat[b], d->xmat[b], R);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_body[g];
    double t[3], R[9];
    double gp[3] = {m->geom_pos[3 * g], m->geom_pos[3 * g + 1], m->geom_pos[3 * g + 2]};
    if (m->task_int[8] == VAR_POS) {   /* hammer_v0.py:119-122 */
      if (g == m->task_int[10]) gp[0] = d->par[3];
      if (g == m->task_int[11]) gp[0] = d->par[4];
    }
    mulMV(t, d->xmat[b], gp);
    add3(d->gpos[g], d->xpos[b], t);
    quat2mat(R, m->geom_quat + 4 * g);
    mulMM(d->gmat[g], d->xmat[b], R);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_body[s];
    double t[3];
    const double* sp = m->site_par[s] >= 0 ? d->par + m->site_par[s] : m->site_pos + 3 * s;
    mulMV(t, d->xmat[b], sp);
    add3(d->spos[s], d->xpos[b], t);
    memcpy(d->smat[s], d->xmat[b], sizeof d->smat[s]); /* site orientation offsets are not used by the touch test below
                                                         (S_target has none, DAPG_hammer.xml:88) */
  }
}

/* ------------------------------------------------------------------ spatial algebra at the world origin */
/* motion vectors (w, v); force vectors (n, f); inertia {m, h = m*c, I_O (sym 3x3, row-major 9)} */
typedef struct { double m, h[3], I[9]; } SpI;

static void spi_body(SpI* s, double mass, const double* c, const double* Rw, const double* diag) {
  s->m = mass;
  scl3(s->h, c, mass);
  /* I_O = R diag R^T + m(|c|^2 1 - c c^T) */
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double v = 0;
      for (int k = 0; k < 3; k++) v += Rw[3 * i + k] * diag[k] * Rw[3 * j + k];
      v += mass * ((i == j ? dot3(c, c) : 0.0) - c[i] * c[j]);
      s->I[3 * i + j] = v;
    }
}
static void spi_add(SpI* a, const SpI* b) {
  a->m += b->m;
  for (int i = 0; i < 3; i++) a->h[i] += b->h[i];
  for (int i = 0; i < 9; i++) a->I[i] += b->I[i];
}
static void spi_mul(double* out, const SpI* s, const double* mv) { /* out = I * (w,v) */
  double t[3], u[3];
  mulMV(t, s->I, mv);
  cross3(u, s->h, mv + 3);
  add3(out, t, u);
  cross3(u, s->h, mv);
  out[3] = s->m * mv[3] - u[0]; out[4] = s->m * mv[4] - u[1]; out[5] = s->m * mv[5] - u[2];
}
static void dof_axis(const OModel* m, const OData* d, int j, double* S) {
  if (m->jnt_type[j] == J_SLIDE) {
    S[0] = S[1] = S[2] = 0;
    copy3(S + 3, d->xaxis[j]);
  } else {
    copy3(S, d->xaxis[j]);
    cross3(S + 3, d->xanchor[j], d->xaxis[j]);
  }
}
static void cross_motion(double* r, const double* a, const double* b) { /* a x b (motion) */
  double t[3], u[3];
  cross3(r, a, b);
  cross3(t, a, b + 3);
  cross3(u, a + 3, b);
  add3(r + 3, t, u);
}
static void cross_force(double* r, const double* a, const double* f) { /* a x* f */
  double t[3], u[3];
  cross3(t, a, f);
  cross3(u, a + 3, f + 3);
  add3(r, t, u);
  cross3(r + 3, a, f + 3);
}
static inline double dot6(const double* a, const double* b) { return dot3(a, b) + dot3(a + 3, b + 3); }

/* composite-rigid-body mass matrix (+ armature) and dense Cholesky M = L L^T */
static int chol_factor(int n, double A[MAXV][MAXV], double L[MAXV][MAXV]) {
  for (int i = 0; i < n; i++)
    for (int j = 0; j <= i; j++) {
      double s = A[i][j];
      for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
      if (i == j) {
        if (s < MINVAL) s = MINVAL;
        L[i][i] = sqrt(s);
      } else
        L[i][j] = s / L[j][j];
    }
  return 0;
}
static void chol_solve(int n, double L[MAXV][MAXV], double* x) {
  for (int i = 0; i < n; i++) {
    double s = x[i];
    for (int k = 0; k < i; k++) s -= L[i][k] * x[k];
    x[i] = s / L[i][i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[k][i] * x[k];
    x[i] = s / L[i][i];
  }
}

static void o_mass_matrix(const OModel* m, OData* d) {
  SpI crb[MAXB];
  int nv = m->nv;
  for (int b = 0; b < m->nbody; b++) spi_body(&crb[b], body_mass_of(m, d, b), d->xipos[b], d->ximat[b], m->body_inertia + 3 * b);
  for (int b = m->nbody - 1; b > 0; b--) spi_add(&crb[m->body_parent[b]], &crb[b]);
  for (int i = 0; i < nv; i++)
    for (int j = 0; j < nv; j++) d->M[i][j] = 0;
  for (int i = 0; i < nv; i++) {
    double S[6], f[6];
    dof_axis(m, d, i, S);
    spi_mul(f, &crb[m->dof_body[i]], S);
    for (int j = i; j >= 0; j = m->dof_parent[j]) {
      double Sj[6];
      dof_axis(m, d, j, Sj);
      d->M[i][j] = d->M[j][i] = dot6(Sj, f);
    }
    d->M[i][i] += m->dof_armature[i];
  }
  chol_factor(nv, d->M, d->L);
}

/* recursive Newton-Euler bias forces (Coriolis, centrifugal, gravity) */
static void o_bias(const OModel* m, OData* d) {
  double vel[MAXB][6], acc[MAXB][6], frc[MAXB][6];
  memset(vel[0], 0, sizeof vel[0]);
  memset(acc[0], 0, sizeof acc[0]);
  for (int k = 0; k < 3; k++) acc[0][3 + k] = -m->gravity[k];
  memset(frc[0], 0, sizeof frc[0]);
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parent[b];
    double v[6], a[6];
    memcpy(v, vel[p], sizeof v);
    memcpy(a, acc[p], sizeof a);
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k;
      double S[6], vj[6], c[6];
      dof_axis(m, d, j, S);
      for (int i = 0; i < 6; i++) vj[i] = S[i] * d->qvel[j];
      cross_motion(c, v, vj);
      for (int i = 0; i < 6; i++) { a[i] += c[i]; v[i] += vj[i]; }
    }
    memcpy(vel[b], v, sizeof v);
    memcpy(acc[b], a, sizeof a);
    SpI I;
    spi_body(&I, body_mass_of(m, d, b), d->xipos[b], d->ximat[b], m->body_inertia + 3 * b);
    double Ia[6], Iv[6], c[6];
    spi_mul(Ia, &I, a);
    spi_mul(Iv, &I, v);
    cross_force(c, v, Iv);
    for (int i = 0; i < 6; i++) frc[b][i] = Ia[i] + c[i];
  }
  for (int b = m->nbody - 1; b > 0; b--)
    for (int i = 0; i < 6; i++) frc[m->body_parent[b]][i] += frc[b][i];
  for (int j = 0; j < m->nv; j++) {
    double S[6];
    dof_axis(m, d, j, S);
    d->qfrc_bias[j] = dot6(S, frc[m->dof_body[j]]);
  }
}

/* translational / rotational Jacobian of a world point attached to `body` */
static void o_jac(const OModel* m, const OData* d, int body, const double* point, double* jp, double* jr) {
  int nv = m->nv;
  if (jp) memset(jp, 0, 3 * nv * sizeof(double));
  if (jr) memset(jr, 0, 3 * nv * sizeof(double));
  for (int b = body; b > 0; b = m->body_parent[b])
    for (int k = 0; k < m->body_jntnum[b]; k++) {
      int j = m->body_jntadr[b] + k;
      if (m->jnt_type[j] == J_SLIDE) {
        if (jp) for (int i = 0; i < 3; i++) jp[i * nv + j] = d->xaxis[j][i];
      } else {
        double r[3], c[3];
        sub3(r, point, d->xanchor[j]);
        cross3(c, d->xaxis[j], r);
        if (jp) for (int i = 0; i < 3; i++) jp[i * nv + j] = c[i];
        if (jr) for (int i = 0; i < 3; i++) jr[i * nv + j] = d->xaxis[j][i];
      }
    }
}

/* passive + actuation + smooth acceleration */
static void o_smooth(const OModel* m, OData* d) {
  int nv = m->nv;
  for (int j = 0; j < nv; j++) {
    d->qfrc_passive[j] = -m->dof_damping[j] * d->qvel[j];
    d->qfrc_actuator[j] = 0;
  }
  for (int u = 0; u < m->nu; u++) {
    double c = d->ctrl[u], lo = m->act_ctrlrange[2 * u], hi = m->act_ctrlrange[2 * u + 1];
    c = c < lo ? lo : (c > hi ? hi : c);
    int j = m->act_dof[u];
    double force = m->act_gain[u] * c + m->act_bias[3 * u] + m->act_bias[3 * u + 1] * d->qpos[j] + m->act_bias[3 * u + 2] * d->qvel[j];
    d->qfrc_actuator[j] += force;
  }
  for (int j = 0; j < nv; j++) {
    d->qfrc_smooth[j] = d->qfrc_passive[j] - d->qfrc_bias[j] + d->qfrc_actuator[j];
    d->qacc_smooth[j] = d->qfrc_smooth[j];
  }
  chol_solve(nv, d->L, d->qacc_smooth);
}

#include "oracle_collide.inc"
#include "oracle_solver.inc"
#include "oracle_task.inc"
