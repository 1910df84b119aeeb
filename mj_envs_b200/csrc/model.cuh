// model.cuh — device-side model tables, per-warp shared-memory layout and per-env scratch layout.
//
// Layout principles: one warp advances one environment.  Small per-dof vectors, body frames, the
// packed joint-space matrices (M and the Newton Hessian / Cholesky factor) and the per-row constraint scalars live
// in the warp's slice of shared memory; wide or rarely touched arrays (geom frames, spatial velocities, contact
// Jacobians, M^-1) live in a per-env global scratch block that stays L1/L2-resident while the warp works on it.
#pragma once
#include "geom.cuh"

namespace adroit {

constexpr int NVP = 36;        // padded dof count (Adroit scenes: 30..36); a multiple of 4: vectors stay 16-byte aligned
constexpr int NBP = 32;        // bodies (29..31)
constexpr int NGP = 36;        // collidable geoms (26..33)
constexpr int NSP = 32;        // sites (28..32)
#ifndef ADROIT_MAXCON
#define ADROIT_MAXCON 100
#endif
constexpr int MAXCON = ADROIT_MAXCON;    // contacts kept per env = the reference's nconmax (assets/DAPG_assets.xml:4); beyond: counted, adroit_overflow
constexpr int MAXCON_M = 32;   // contact capacity of the shared-memory tables of the finish kernel's common class (more: tables in the context block)
constexpr int CROWS = MAXCON * 4;  // contact-frame Jacobian rows: normal, 2 tangents, torsion
#ifndef ADROIT_MAXSIMPLE
#define ADROIT_MAXSIMPLE 96
#endif
constexpr int MAXSIMPLE = ADROIT_MAXSIMPLE;  // friction-loss + joint-limit + tendon-limit rows
#ifndef ADROIT_NTRI
#define ADROIT_NTRI 720
#endif
constexpr int NTRI = ADROIT_NTRI;      // lower triangle of a 36x36 matrix, rows padded to multiples of 4 (16-byte aligned rows, see tri())
constexpr int NPARP = 8;       // per-env model parameters (board_pos | door_body_pos | target quat | obj pos + target pos)
constexpr int NTF = 24;        // cached task frames (positions the observation / reward read)
constexpr int MAXLVSTEP = 128; // steps of the level-scheduled factorisation program (Adroit: ~50)
constexpr int MAXTENDON = 48;  // fixed tendons (Adroit: 44, all limited; at most two joints each)
#ifndef ADROIT_WPB
#define ADROIT_WPB 4
#endif
constexpr int WPB = ADROIT_WPB;  // warps (= envs) per block
constexpr int MAXSURV = 192;   // candidate pairs surviving the bounding-sphere cull (shared-memory list of the collide kernel)
constexpr int NSURVI = (3 * MAXCON > MAXSURV ? 3 * MAXCON + 4 : MAXSURV);    // the I_SURV scratch region of the context block: max(MAXSURV, 3 * MAXCON tangent rows of the noslip pass), 16-byte multiple

enum { T_HAMMER = 0, T_DOOR = 1, T_PEN = 2, T_RELOCATE = 3 };
enum { J_SLIDE = 2, J_HINGE = 3 };
// hammer variation types (hammer_v0.py:110-129); task_int[8] = type, [9] = Object body, [10] = head geom, [11] = neck geom;
// the drawn values live in the per-env parameters par[3], par[4]
enum { VAR_NONE = 0, VAR_MASS = 1, VAR_POS = 2, VAR_SIZE = 3 };
enum { S_FRICTION = 0, S_JLIMIT = 1, S_TLIMIT = 2 };

// ---- shared memory (per warp), offsets in units of T
enum {
  // persistent / handed-over vectors first, ordered so that every phase kernel's staged set is one contiguous run of the
  // context block (smooth: CTRL..QVEL, solve: WARM..QAS in, QACC..QFC out, finish: QPOS..QFC): one bulk copy each
  V_CTRL = 0, V_QPOS, V_QVEL, V_WARM, V_QFS, V_QAS, V_QACC, V_QFC, V_MA, V_GRAD, V_SEARCH, V_MV, V_T1, V_T2, V_DINV, V_COUNT
};
constexpr int O_PAR = V_COUNT * NVP;
constexpr int O_XPOS = O_PAR + NPARP;
constexpr int O_XQUAT = O_XPOS + NBP * 3;
constexpr int O_XMAT = O_XQUAT + NBP * 4;
constexpr int O_XANC = O_XMAT + NBP * 9;
constexpr int O_XAXIS = O_XANC + NVP * 3;
constexpr int O_CRB = O_XAXIS + NVP * 3;
constexpr int O_M = O_CRB + NBP * 10;
constexpr int O_H = O_M + NTRI;
constexpr int O_SD = O_H + NTRI;
constexpr int O_SAREF = O_SD + MAXSIMPLE;
constexpr int O_SFL = O_SAREF + MAXSIMPLE;
constexpr int O_SFORCE = O_SFL + MAXSIMPLE;
constexpr int O_AC = O_SFORCE + MAXSIMPLE;      // contact-frame acceleration  Jf * qacc   [MAXCON][4]
constexpr int O_AV = O_AC + CROWS;              // contact-frame search dir    Jf * search [MAXCON][4]
constexpr int O_CD = O_AV + CROWS;              // D of every edge of the contact
constexpr int O_CMU = O_CD + MAXCON;            // mu[3] per contact
constexpr int O_CAREFN = O_CMU + MAXCON * 3;    // reference acceleration, normal part
constexpr int O_CBT = O_CAREFN + MAXCON;        // -B*mu_k*vt_k per tangent
constexpr int O_EF = O_CBT + MAXCON * 3;        // edge forces [MAXCON][6]
constexpr int O_W = O_EF + MAXCON * 6;          // per-contact Hessian weights [MAXCON][7]
constexpr int O_TF = O_W + MAXCON * 7;
constexpr int SM_T = ((O_TF + NTF + 3) / 4) * 4;
static_assert(O_M % 4 == 0 && O_H % 4 == 0, "matrix rows must be 16-byte aligned");
// ints
constexpr int I_SINFO = 0;                      // simple rows: kind | id<<2 | (side>0)<<12 | state<<13
constexpr int I_CDIM = I_SINFO + MAXSIMPLE;
constexpr int I_CB1 = I_CDIM + MAXCON;
constexpr int I_CB2 = I_CB1 + MAXCON;
constexpr int I_CMASK = I_CB2 + MAXCON;           // active-edge bit mask of every contact (Newton active-set change detection)
constexpr int I_SURV = I_CMASK + MAXCON;
constexpr int SM_I = I_SURV + NSURVI;
static_assert(NSURVI >= MAXSURV && NSURVI >= 3 * MAXCON && NSURVI % 4 == 0, "I_SURV holds the collide list and the noslip tangent list");

template <typename T> constexpr size_t smem_per_warp() { return SM_T * sizeof(T) + SM_I * sizeof(int); }

// ---- global scratch (per env), offsets in units of T
constexpr int G_GPOS = 0;
constexpr int G_GMAT = G_GPOS + NGP * 3;
constexpr int G_SPOS = G_GMAT + NGP * 9;
constexpr int G_CIN = G_SPOS + NSP * 3;          // single-body inertia about the body origin, world axes [NBP][10]
constexpr int G_VEL = G_CIN + NBP * 10;
constexpr int G_ACC = G_VEL + NBP * 6;
constexpr int G_FRC = G_ACC + NBP * 6;
constexpr int G_CON = G_FRC + NBP * 6;           // contacts [MAXCON][16]: dist, pos3, frame9, margin, -, -
constexpr int G_JR = G_CON + MAXCON * 16;        // contact-frame Jacobian, row-major [CROWS][NVP]
constexpr int G_LINV = G_JR + CROWS * NVP;       // inverse Cholesky factor of M (packed)
constexpr int G_MINV = G_LINV + NTRI;            // M^-1 (packed)
constexpr int G_YT = G_MINV + NTRI;              // M^-1 Jt_k^T for the noslip pass [MAXCON][3][NVP]
constexpr int G_LS = G_YT + MAXCON * 3 * NVP;    // line-search rows of the pyramid edges beyond the register sets (big solve class) [3][6 * MAXCON]: x, y, D
constexpr int MPRQ = 32;                          // MPR pairs per env that go through the block-pooled pass of the collide kernel (more: computed in place)
constexpr int G_MPRRES = G_LS + 18 * MAXCON;     // results of the pooled MPR pass [MPRQ][8]: hit, dist, pos3, nrm3
constexpr int G_DBG = G_MPRRES + 8 * MPRQ;       // debug dump: M packed | qfrc_bias | qfrc_smooth | qacc_smooth | qacc | qfrc_constraint
constexpr int G_DBG_M = G_DBG, G_DBG_BIAS = G_DBG + NTRI, G_DBG_QFS = G_DBG_BIAS + NVP, G_DBG_QAS = G_DBG_QFS + NVP,
              G_DBG_QACC = G_DBG_QAS + NVP, G_DBG_QFC = G_DBG_QACC + NVP, G_DBG_XPOS = G_DBG_QFC + NVP,
              G_DBG_XQUAT = G_DBG_XPOS + NBP * 3, G_DBG_EF = G_DBG_XQUAT + NBP * 4;
constexpr int G_TOTAL = ((G_DBG_EF + MAXCON * 6 + 31) / 32) * 32;
// int scratch per env: ncon, nsimple, newton iterations, noslip sweeps, then contact ints [MAXCON][4]: geom1, geom2, dim, first edge
constexpr int GI_NCON = 0, GI_NSIMPLE = 1, GI_NITER = 2, GI_NSLIP = 3, GI_CON = 4, GI_TOTAL = 4 + MAXCON * 4 + 28;

template <typename T> struct DModel {
  int nv, nbody, ngeom, nsite, nu, ntendon, npair, npar, obs_dim, frame_skip, task, horizon, nlevel, touch_site;
  int iterations, noslip_iterations, mpr_iterations, n_wl, n_tri;
  const int *body_parent, *body_jntadr, *body_jntnum, *body_par, *level_start, *level_body, *child_start, *child_list;
  const unsigned long long* body_dofmask;
  const T *body_pos, *body_quat, *body_ipos, *body_imat, *body_mass, *body_inertia, *body_invweight0;
  const int *jnt_type, *jnt_limited, *dof_body, *dof_parent;
  const T *jnt_pos, *jnt_axis, *jnt_range, *jnt_margin, *dof_damping, *dof_armature, *dof_frictionloss, *dof_invweight0;
  const int *geom_type, *geom_body;
  const T *geom_pos, *geom_mat, *geom_size, *geom_rbound;
  const int *site_body, *site_par;
  const T* site_pos;
  const int *pair_geom, *pair_condim, *pair_word;     // pair_word: geom1 | geom2 << 8 | type1 << 16 | type2 << 24 (one load in the broad phase)
  const T *pair_friction, *pair_margin, *pair_bound;  // pair_bound: rbound1 + rbound2 + margin (plane pairs: rbound2 + margin)
  const int *tendon_dof, *tendon_limited;
  const T *tendon_coef, *tendon_range, *tendon_margin, *tendon_invweight0;
  const int* act_dof;
  // tree-sparse factorisation tables (host-built from dof_parent, see fill_model)
  const int* lv_items;                               // items of the level-scheduled L^T D L program (dynamics.cuh: ltdl_factor), n_lvitems of them
  const int *ld_start, *ld_pair, *wl_pair, *tri_ab;   // [nv+1], [..] tri(i,j)|i<<10|j<<16, [..] k|j<<8, [ntri_real] a|b<<8
  const int *tri_abk;                                // [n_tri] a | b<<8 | lca(a,b)<<16 (255 = none), a >= b
  const int *dof_depth, *anc_tab;                    // [nv] number of proper ancestor dofs, [nv][MAXDEPTH] the chain, nearest first (-1 padded)
  const unsigned char* lca;                          // [nv*nv] deepest common ancestor-or-self dof, 255 = none
  const T *act_ctrlrange, *act_gain, *act_bias;
  const int* render_int;                             // [nrender][4] type, body, par offset of the local pos (-1), is_site
  const T* render_real;                              // [nrender][19] local pos, local mat, size, rgba (render.cuh)
  int task_int[16];
  // fixed-tendon tables by value: kernel-parameter (constant-bank) reads instead of L2 round trips on the row loops' chain
  int n_lvitems;
  int lv_steps[MAXLVSTEP];                           // step s: type (0 pivots, 1 updates, 2 row scaling, 3 end) | first item << 2; it ends where step s+1 starts
  int ld_start_v[NVP + 4];                           // ld_start by value: the pivot loop of ltdl_factor reads it with a uniform index (constant bank)
  int tdof[2 * MAXTENDON];
  T tcoef[2 * MAXTENDON];
  T par0[NPARP];
  T touch_size[3];
  T timestep, gravity[3], tolerance, noslip_tolerance, impratio, solref[2], solimp[5], mpr_tolerance, meaninertia;
  T pen_len, tar_len;
  int any_damping;
  int nrender, light_body;
  T render_cam[8], light_pos[3], light_diffuse[3];   // mjcf.py::_set_render_camera
};

// persistent per-env state handed to the kernels.  qpos / qvel / qacc_warmstart / ctrl / per-env model parameters /
// cached task frames live INSIDE the context block (V_QPOS, V_QVEL, V_WARM, V_CTRL, O_PAR, O_TF regions).
template <typename T> struct DState {
  int n;
  T* ctx;                                       // [n][CTX_T]   (dynamics.cuh)
  int* ctxi;                                    // [n][CTX_I]
  int *t, *episode, *ep_goal;                   // [n]
  T* ep_return;                                 // [n]
  double* acc;                                  // [n][4]: sum return, #any-goal, #goal>thr, #episodes
  unsigned char* reset_flag;                    // [n] set by the step epilogue when AUTO_RESET wants a reset
  unsigned long long* overflow;                 // [2]
  unsigned long long seed;
  long long env_off;
  unsigned flags;
};

// Symmetric / triangular matrices are stored as the lower triangle, row-major, with row i padded to 4*(i/4+1) entries:
// every row starts 16-byte aligned (vector loads in the dot-product loops) and the pad entries are kept at ZERO.
DI int rowoff(int i) { int g = i >> 2; return 4 * (g + 1) * (2 * g + (i & 3)); }
DI int tri(int i, int j) { return rowoff(i) + j; }  // i >= j

template <typename T> DI T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// four warp sums at once: 6 shuffles instead of 20.  Returns sum(v[lane >> 3]) -- lanes 0-7 hold the total of v[0],
// lanes 8-15 of v[1], lanes 16-23 of v[2], lanes 24-31 of v[3].
template <typename T> DI T warp_sum4(T v0, T v1, T v2, T v3, int lane) {
  const unsigned FULL = 0xffffffffu;
  const bool up = lane & 16, hi = lane & 8;
  T a = (up ? v2 : v0) + __shfl_xor_sync(FULL, up ? v0 : v2, 16);
  T b = (up ? v3 : v1) + __shfl_xor_sync(FULL, up ? v1 : v3, 16);
  T c = (hi ? b : a) + __shfl_xor_sync(FULL, hi ? a : b, 8);
  c += __shfl_xor_sync(FULL, c, 4);
  c += __shfl_xor_sync(FULL, c, 2);
  c += __shfl_xor_sync(FULL, c, 1);
  return c;
}
template <typename T> DI T warp_bcast(T v, int src) { return __shfl_sync(0xffffffffu, v, src); }

}  // namespace adroit
