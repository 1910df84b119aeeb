// adroit_b200.cu — kernels and the C ABI (include/adroit_b200.h) of the batched Adroit simulator for sm_100a.
//
// One warp advances one environment.  An env.step() is a short sequence of PHASE KERNELS per physics substep
// (smooth dynamics -> collision -> constraint solve -> noslip/Euler/task epilogue) that hand the per-env context to
// each other through a context block in HBM/L2 (dynamics.cuh: Ctx, MAP).  Each phase kernel stages only the regions it
// works on in shared memory, has its own register budget and occupancy, and is small enough for the instruction
// cache; a step never synchronises with the host.  See DESIGN.md for the data layout and the per-phase roofline; the
// reference call sites each entry point replaces are cited in include/adroit_b200.h.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/adroit_b200.h"
#include "render.cuh"

namespace adroit {

enum { MODE_STEP = 0, MODE_RESET = 1, MODE_FORWARD = 2 };
#ifndef ADROIT_DEFAULT_SUBBATCHES
#define ADROIT_DEFAULT_SUBBATCHES 3   // measured (round 2, tools/gpu_subbatch_sweep.sh): 3 >= 4 at 8192 and 16384 envs, mostly in the end-to-end number
#endif

struct PArgs {
  const float* action;        // smooth: action map (hammer_v0.py:55-59) when non-null
  const uint8_t* mask;        // every phase: only envs with mask[env] != 0 (NULL = all)
  float* obs;                 // finish: observation out (may be NULL)
  float* reward;
  uint8_t* done;
  uint8_t* goal;
  int reset;                  // smooth: reset_model prologue
  int check;                  // smooth: mj_checkPos / mj_checkVel
  int integrate;              // finish: mj_Euler
  int epilogue;               // finish: reward / success / episode counters / auto-reset decision (last substep of a step)
  int write_obs;              // finish: pack the observation
  int debug;                  // export introspection copies (G_DBG_*)
  int env0, env1;             // this launch covers envs [env0, env1): one sub-batch of the handle (run_t)
};

#define VR(V, F) Reg{(V) * NVP, NVP, (F)}
// ---- shared-memory maps of the phase kernels (everything not listed is addressed in the context block directly)
#ifndef ADROIT_WPB_SMOOTH
#define ADROIT_WPB_SMOOTH ADROIT_WPB
#endif
struct MapSmooth {
  static constexpr int W = ADROIT_WPB_SMOOTH;   // warps (= envs) per block of the kernel that uses this map
  static constexpr int NT = 14, NI = 1;
  static constexpr Reg RT[NT] = {VR(V_CTRL, R_IN | R_OUT), VR(V_QPOS, R_IN | R_OUT), VR(V_QVEL, R_IN | R_OUT),
                                 VR(V_QAS, R_ZERO | R_OUT), VR(V_T1, R_ZERO), VR(V_DINV, R_ZERO),
                                 Reg{O_PAR, NPARP, R_IN | R_OUT}, Reg{O_XPOS, NBP * 3, R_OUT}, Reg{O_XQUAT, NBP * 4, R_OUT},
                                 Reg{O_XMAT, NBP * 9, R_OUT}, Reg{O_XANC, NVP * 3, R_OUT}, Reg{O_XAXIS, NVP * 3, R_OUT},
                                 Reg{O_CRB, NBP * 10, 0}, Reg{O_M, NTRI, 0}};   // M: stored early by the kernel itself, then factored in place
  static constexpr Reg RI[NI] = {Reg{I_SURV, 4, 0}};
};
// the geom frames (written by the smooth kernel) are staged: the broad phase walks ~250 candidate pairs per env and
// every test starts with two position reads; the kernel is register-bound (2 blocks/SM), shared memory is free
#ifndef ADROIT_WPB_COLLIDE
#define ADROIT_WPB_COLLIDE ADROIT_WPB
#endif
struct MapCollide {
  static constexpr int W = ADROIT_WPB_COLLIDE;
  static constexpr int EXTRA = W * 16 + W * MPRQ * 2;   // per warp: queue length (16-byte slot) + MPRQ 16-bit pair indices
  static constexpr int NT = 3, NI = 1;
  static constexpr Reg RT[NT] = {Reg{O_W, 4, 0}, Reg{SM_T + G_GPOS, NGP * 3, R_IN}, Reg{SM_T + G_GMAT, NGP * 9, R_IN}};
  static constexpr Reg RI[NI] = {Reg{I_SURV, MAXSURV / 2, 0}};   // 16-bit entries: 3 blocks/SM stay inside the 32 KB shared-memory carve-out (the kernel lives on L1)
};
// The solve kernel comes in two size classes.  Small (MC = MAXCON_S = 16 contacts, shared-memory contact tables, 11 KB per env):
// every env in random play, nearly all under the DAPG policies.  Big (MC = MAXCON = 100, the reference's nconmax): the contact
// tables stay in the context block (L1/L2) -- heavy grasps and pile-ups, rare, and correct up to the reference's capacity
// instead of dropping contacts beyond 32 as round 1 did.  Each class is its own launch over the same env range; a warp whose
// env belongs to the other class exits before staging anything.
#ifndef ADROIT_MAXCON_S
#define ADROIT_MAXCON_S 16
#endif
constexpr int MAXCON_S = ADROIT_MAXCON_S;
#ifndef ADROIT_WPB_SOLVE
#define ADROIT_WPB_SOLVE 5    // 4 blocks x 5 envs per SM: measured best of 4 / 5 / 10 / 20 (profiles/R2c_ab_lockstep.txt)
#endif
#define SOLVE_COMMON_RT                                                                                                          \
  VR(V_WARM, R_IN), VR(V_QFS, R_IN), VR(V_QAS, R_IN), VR(V_QACC, R_ZERO | R_OUT), VR(V_QFC, R_ZERO | R_OUT), VR(V_MA, R_ZERO),    \
      VR(V_GRAD, R_ZERO), VR(V_SEARCH, R_ZERO), VR(V_MV, R_ZERO), VR(V_DINV, R_ZERO),                                            \
      Reg{O_M, NTRI, R_IN},   /* (A/B: reading M from L1/L2 instead costs 30 %) */                                               \
      Reg{O_H, NTRI, 0}, Reg{O_SD, MAXSIMPLE, 0}, Reg{O_SAREF, MAXSIMPLE, R_OUT}, Reg{O_SFL, MAXSIMPLE, R_OUT}, Reg{O_SFORCE, MAXSIMPLE, R_OUT}
// WW = warps per block: ADROIT_WPB_SOLVE for the fp32 small class (block-lockstep Newton, constraint.cuh), WPB otherwise
template <int MC, int WW = WPB> struct MapSolveT {
  static constexpr int W = WW;
  static constexpr int NT = 24, NI = 5;
  static constexpr Reg RT[NT] = {SOLVE_COMMON_RT,
                                 Reg{O_AC, MC * 4, 0}, Reg{O_AV, MC * 4, 0}, Reg{O_CD, MC, 0}, Reg{O_CMU, MC * 3, R_IN},
                                 Reg{O_CAREFN, MC, 0}, Reg{O_CBT, MC * 3, R_OUT}, Reg{O_EF, MC * 6, R_OUT}, Reg{O_W, MC * 7, 0}};
  static constexpr Reg RI[NI] = {Reg{I_SINFO, MAXSIMPLE, R_OUT}, Reg{I_CDIM, MC, R_IN}, Reg{I_CB1, MC, R_IN}, Reg{I_CB2, MC, R_IN},
                                 Reg{I_CMASK, MC, 0}};
};
struct MapSolveBig {   // contact tables in the context block
  static constexpr int W = WPB;
  static constexpr int NT = 16, NI = 1;
  static constexpr Reg RT[NT] = {SOLVE_COMMON_RT};
  static constexpr Reg RI[NI] = {Reg{I_SINFO, MAXSIMPLE, R_OUT}};
};
template <typename T, int MC> constexpr int solve_w() { return (sizeof(T) == 4 && MC < MAXCON) ? ADROIT_WPB_SOLVE : WPB; }
template <typename T, int MC> struct MapSolvePick { using type = MapSolveT<MC, solve_w<T, MC>()>; };
template <typename T> struct MapSolvePick<T, MAXCON> { using type = MapSolveBig; };
template <typename T, int MC> using MapSolveK = typename MapSolvePick<T, MC>::type;
// M^-1 is staged (read-only) for the noslip sweeps; the same shared-memory region then holds the (M + hB) factor.  Two size
// classes like the solve kernel: up to MAXCON_M = 32 contacts the contact tables are staged, beyond they stay in the context block.
// (shared memory bounds the occupancy: qfrc_smooth / qacc_smooth are read from the context block where needed, and the two
// scratch vectors of the noslip pass serve again afterwards -- V_GRAD as the solve vector of the Euler step, V_SEARCH as the
// reciprocal pivots of its factorisation: 7 blocks per SM instead of 6)
#define FINISH_COMMON_RT                                                                                                         \
  VR(V_QPOS, R_IN | R_OUT), VR(V_QVEL, R_IN | R_OUT), VR(V_WARM, R_IN | R_OUT), VR(V_QACC, R_IN | R_OUT), VR(V_QFC, R_IN | R_OUT), \
      VR(V_GRAD, R_ZERO), VR(V_SEARCH, R_ZERO), Reg{SM_T + G_MINV, NTRI, R_IN}, Reg{O_SAREF, MAXSIMPLE, R_IN},                    \
      Reg{O_SFL, MAXSIMPLE, R_IN}, Reg{O_SFORCE, MAXSIMPLE, R_IN | R_OUT}
struct MapFinish {
  static constexpr int W = WPB;
  static constexpr int NT = 16, NI = 3;
  static constexpr Reg RT[NT] = {FINISH_COMMON_RT, Reg{O_AV, MAXCON_M * 3, 0}, Reg{O_CMU, MAXCON_M * 3, R_IN},
                                 Reg{O_CBT, MAXCON_M * 3, R_IN}, Reg{O_EF, MAXCON_M * 6, R_IN | R_OUT}, Reg{O_TF, NTF, R_IN | R_OUT}};
  static constexpr Reg RI[NI] = {Reg{I_SINFO, MAXSIMPLE, R_IN}, Reg{I_CDIM, MAXCON_M, R_IN}, Reg{I_SURV, MAXCON_M * 3, 0}};
};
struct MapFinishBig {
  static constexpr int W = WPB;
  static constexpr int NT = 12, NI = 1;
  static constexpr Reg RT[NT] = {FINISH_COMMON_RT, Reg{O_TF, NTF, R_IN | R_OUT}};
  static constexpr Reg RI[NI] = {Reg{I_SINFO, MAXSIMPLE, R_IN}};
};
template <bool BIG> struct MapFinishPick { using type = MapFinish; };
template <> struct MapFinishPick<true> { using type = MapFinishBig; };
// the auto-reset of a step (reset_kernel): reset_model + the part of mj_forward the observation reads (kinematics, site frames)
struct MapReset {
  static constexpr int W = WPB;
  static constexpr int NT = 12, NI = 1;
  static constexpr Reg RT[NT] = {VR(V_CTRL, R_ZERO | R_OUT), VR(V_QPOS, R_ZERO | R_OUT), VR(V_QVEL, R_ZERO | R_OUT), VR(V_WARM, R_ZERO | R_OUT),
                                 Reg{O_PAR, NPARP, R_IN | R_OUT}, Reg{O_XPOS, NBP * 3, R_ZERO | R_OUT}, Reg{O_XQUAT, NBP * 4, R_ZERO | R_OUT},
                                 Reg{O_XMAT, NBP * 9, R_ZERO | R_OUT}, Reg{O_XANC, NVP * 3, R_ZERO | R_OUT}, Reg{O_XAXIS, NVP * 3, R_ZERO | R_OUT},
                                 Reg{SM_T + G_SPOS, NSP * 3, R_ZERO | R_OUT}, Reg{O_TF, NTF, R_IN | R_OUT}};
  static constexpr Reg RI[NI] = {Reg{I_SURV, 4, 0}};
};
#undef VR

template <typename T> DI bool warp_bad(const T* x, int n, int lane) {
  bool bad = false;
  for (int j = lane; j < n; j += 32) bad |= !(tabs(x[j]) < T(1e10));
  return __any_sync(0xffffffffu, bad);
}
template <class C> DI void zero_state(C& c) {
  using T = typename C::real;
  T *qpos = c.template V<V_QPOS>(), *qvel = c.template V<V_QVEL>(), *warm = c.template V<V_WARM>();
  _Pragma("unroll 1") for (int j = c.lane; j < NVP; j += 32) { qpos[j] = 0; qvel[j] = 0; warm[j] = 0; }
  __syncwarp();
}

template <typename T> __device__ void episode_finish(const DModel<T>& m, const DState<T>& st, int env) {
  if (st.t[env] == 0) return;
  int thr = m.task == T_PEN ? 20 : 25;  // evaluate_success: hammer_v0.py:167-175, pen_v0.py:180-188
  double* a = st.acc + 4 * env;
  a[0] += (double)st.ep_return[env];
  a[1] += st.ep_goal[env] > 0;
  a[2] += st.ep_goal[env] > thr;
  a[3] += 1;
}

// reset_model (hammer_v0.py:106-132, door_v0.py:103-119, pen_v0.py:115-132, relocate_v0.py:85-103): episode bookkeeping,
// the randomised model parameters, qpos = init_qpos (zeros), qvel = 0, ctrl = 0
template <class C, typename T> DI void reset_prologue(C& c, const DModel<T>& m, const DState<T>& st, int env) {
  if (c.lane == 0) {
    episode_finish(m, st, env);
    T par[NPARP];
    draw_params(m, st.seed, (unsigned long long)(st.env_off + env), (unsigned)st.episode[env], par);
    T* P = c.template S<O_PAR>();
    for (int i = 0; i < m.npar; i++) P[i] = par[i];
    st.episode[env] += 1;
    st.t[env] = 0;
    st.ep_return[env] = 0;
    st.ep_goal[env] = 0;
  }
  T* ctrl = c.template V<V_CTRL>();
  _Pragma("unroll 1") for (int j = c.lane; j < NVP; j += 32) ctrl[j] = 0;
  zero_state(c);
}

// common prologue of a phase kernel: env of this warp, context pointers, staging
// programmatic dependent launch: the kernel may have been started while its predecessor in the stream was still draining;
// nothing the predecessor wrote (context block, reset_flag mask, GI_NCON) is read before this point.  A no-op for launches
// without the attribute.  The dependents are released at once: they only occupy slots that this grid no longer fills.
#define PDL_WAIT()                                                                                                        \
  asm volatile("griddepcontrol.wait;" ::: "memory");                                                                      \
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#define PHASE_PROLOGUE(MAP)                                                                                             \
  extern __shared__ __align__(16) unsigned char smem_raw[];                                                              \
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;                                                            \
  const int env = a.env0 + blockIdx.x * MAP::W + warp;                                                                      \
  if (env >= a.env1) return;                      /* no block-wide barrier anywhere in a phase kernel */                 \
  if (a.mask && !a.mask[env]) return;                                                                                    \
  T* sm_ = reinterpret_cast<T*>(smem_raw) + (size_t)warp * map_st<MAP>();                                                \
  int* smi_ = reinterpret_cast<int*>(smem_raw + (size_t)MAP::W * map_st<MAP>() * sizeof(T)) + warp * map_si<MAP>();      \
  Ctx<T, MAP> c(m, sm_, smi_, st.ctx + (size_t)env * CTX_T, st.ctxi + (size_t)env * CTX_I, lane);                        \
  /* one mbarrier per warp (byte count of its bulk stage-in copies) behind the maps, inside the dynamic allocation */   \
  unsigned long long* bar_ = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)MAP::W * map_bytes<T, MAP>()) + 2 * warp; \
  stage_in(c, bar_);                                                                                                     \
  __syncwarp();

// ---------------------------------------------------------------- phase 1: smooth dynamics
// [reset_model prologue] [action map] [mj_checkPos/Vel] kinematics, frames, CRB, RNE, actuation, L^T D L, qacc_smooth, M^-1
template <typename T>
__global__ void __launch_bounds__(MapSmooth::W * 32) smooth_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st,
                                                           const __grid_constant__ PArgs a) {
  PDL_WAIT()
  PHASE_PROLOGUE(MapSmooth)
  const int nv = m.nv;
  if (a.reset) reset_prologue(c, m, st, env);
  if (a.action) {   // action map: hammer_v0.py:55-59 (act_mid / act_rng :49-50)
    T* ctrl = c.template V<V_CTRL>();
    _Pragma("unroll 1") for (int u = lane; u < m.nu; u += 32) {
      T x = tclamp((T)a.action[(size_t)env * m.nu + u], T(-1), T(1));
      T lo = m.act_ctrlrange[2 * u], hi = m.act_ctrlrange[2 * u + 1];
      ctrl[u] = T(0.5) * (lo + hi) + x * T(0.5) * (hi - lo);
    }
    __syncwarp();
  }
  if (a.check && (warp_bad(c.template V<V_QPOS>(), nv, lane) || warp_bad(c.template V<V_QVEL>(), nv, lane))) zero_state(c);  // mj_checkPos / mj_checkVel
  kinematics(c);
  frames_and_inertia(c);
  rne_bias(c);            // (its body-vector scratch lives where M is about to be built)
  mass_matrix(c);
  if (a.debug) {
    const T* M = c.template S<O_M>();
    T* D = c.template G<G_DBG_M>();
    for (int p = lane; p < nv * nv; p += 32) {   // exported as the plain packed lower triangle
      int i = p / nv, j = p - i * nv;
      if (j <= i) D[(i * (i + 1)) / 2 + j] = M[tri(i, j)];
    }
  }
  // M leaves for the context block now (one bulk copy, the solve and finish kernels read it); the same shared-memory
  // region then holds its factor: no second 2.9 KB matrix per env, one more resident block per SM
  store_region_early<O_M, NTRI>(c);
  smooth_dynamics(c, m.noslip_iterations > 0);
  __syncwarp();
  stage_out(c);
}

// ---------------------------------------------------------------- phase 2: collision
#ifndef ADROIT_MPR_SPREAD
#define ADROIT_MPR_SPREAD 3
#endif
#ifndef ADROIT_MPR_SPREAD_MIN
#define ADROIT_MPR_SPREAD_MIN 4    // measured: 2 costs hammer +8 % of its collide kernel, 8 and beyond lose pen's -4.7 %
#endif
#ifndef ADROIT_COLLIDE_BLOCKS
#define ADROIT_COLLIDE_BLOCKS 3   // <= 168 registers: 3 blocks per SM instead of 2 at 202 (collide -11 % hammer, -7.5 % pen; 4 blocks at 128 registers: slower)
#endif
template <typename T>
__global__ void __launch_bounds__(MapCollide::W * 32, ADROIT_COLLIDE_BLOCKS) collide_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st,
                                                            const __grid_constant__ PArgs a) {
  PDL_WAIT()
  // (no early return in this kernel: every warp reaches the two block barriers around the pooled MPR pass)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using MAP = MapCollide;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int env = a.env0 + blockIdx.x * MAP::W + warp;
  const bool alive = env < a.env1 && !(a.mask && !a.mask[env]);
  const int envc = alive ? env : a.env0;             // (pointer arithmetic only for a warp without an env)
  T* sm_ = reinterpret_cast<T*>(smem_raw) + (size_t)warp * map_st<MAP>();
  int* smi_ = reinterpret_cast<int*>(smem_raw + (size_t)MAP::W * map_st<MAP>() * sizeof(T)) + warp * map_si<MAP>();
  unsigned char* tail = smem_raw + (size_t)MAP::W * map_bytes<T, MAP>();
  unsigned long long* bar_ = reinterpret_cast<unsigned long long*>(tail) + 2 * warp;
  int* qn = reinterpret_cast<int*>(tail + MAP::W * 16);                                       // [W] (stride 4 ints)
  unsigned short* qk = reinterpret_cast<unsigned short*>(tail + MAP::W * 32);                 // [W][MPRQ]
  Ctx<T, MAP> c(m, sm_, smi_, st.ctx + (size_t)envc * CTX_T, st.ctxi + (size_t)envc * CTX_I, lane);
  if (lane == 0) qn[4 * warp] = 0;
  int nsurv = 0;
  if (alive) {
    stage_in(c, bar_);
    __syncwarp();
    nsurv = collision_broad(c, st.overflow, qk + warp * MPRQ, qn + 4 * warp);
  }
  __syncthreads();
  {  // pooled MPR: thread e of the block takes the e-th queued pair of the block's envs
    int off[MAP::W + 1];
    off[0] = 0;
#pragma unroll
    for (int w = 0; w < MAP::W; w++) off[w + 1] = off[w] + qn[4 * w];
    // which queued pair a thread takes: 0 = thread e takes pair e (the first warps fill up, the others idle through the pass),
    // 1 = the queue cut into one contiguous chunk per warp, 2 = dealt round-robin over the warps.  A warp's MPR time is set by the
    // divergence among its lanes as much as by its longest pair, so fewer pairs per warp on more warps is the shorter pass.
#if ADROIT_MPR_SPREAD == 1
    const int total_ = off[MAP::W], chunk_ = (total_ + MAP::W - 1) / MAP::W;
    for (int i_ = lane; i_ < chunk_; i_ += 32) {
      const int e = warp * chunk_ + i_;
      if (e >= total_) break;
      int w = 0;
#elif ADROIT_MPR_SPREAD == 2
    for (int e = lane * MAP::W + warp; e < off[MAP::W]; e += MAP::W * 32) {
      int w = 0;
#elif ADROIT_MPR_SPREAD == 3
    // up to a warp's worth of pairs stays on the first warp (the others wait at the barrier without touching the MPR code: hammer,
    // a handful of pairs per block), more is dealt round-robin over the warps (pen: 40-60 pairs per block)
    for (int e = off[MAP::W] <= ADROIT_MPR_SPREAD_MIN ? (int)threadIdx.x : lane * MAP::W + warp; e < off[MAP::W]; e += MAP::W * 32) {
      int w = 0;
#else
    for (int e = threadIdx.x; e < off[MAP::W]; e += MAP::W * 32) {
      int w = 0;
#endif
#pragma unroll
      for (int q = 1; q < MAP::W; q++) w += e >= off[q];
      const int j = e - off[w];
      const T* smw = reinterpret_cast<const T*>(smem_raw) + (size_t)w * map_st<MAP>();
      T* gbw = st.ctx + (size_t)(a.env0 + blockIdx.x * MAP::W + w) * CTX_T;
      collision_mpr_entry<Ctx<T, MAP>>(m, smw + map_soff<MAP>(SM_T + G_GPOS), smw + map_soff<MAP>(SM_T + G_GMAT), gbw, qk[w * MPRQ + j], j);
    }
  }
  __syncthreads();
  if (alive) {
    collision_narrow(c, nsurv, st.overflow);
    if (lane == 0) c.gi[GI_NCON] = c.ncon;
  }
}

// ---------------------------------------------------------------- auto-reset inside env.step()
// The envs the step epilogue flagged (horizon reached / reset-on-done) get reset_model and the part of the following
// sim.forward() that get_obs() reads: kinematics and the site / body frames (hammer_v0.py:126-132: the touch sensor reads 0
// at the reset pose, the hammer lies on the table).  One launch instead of a masked five-launch forward pass: everything
// else a forward computes (M, contacts, qacc) is rebuilt by the first substep of the next step before anything reads it.
template <typename T>
__global__ void __launch_bounds__(WPB * 32) reset_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st,
                                                          const __grid_constant__ PArgs a) {
  PDL_WAIT()
  PHASE_PROLOGUE(MapReset)
  reset_prologue(c, m, st, env);
  kinematics(c);
  frames_and_inertia<decltype(c), false>(c);
  cache_task_frames(c, T(0));
  if (a.obs) write_obs(m, c.template V<V_QPOS>(), c.template V<V_QVEL>(), c.template S<O_TF>(), a.obs + (size_t)env * m.obs_dim, lane);
  __syncwarp();
  stage_out(c);
}

// ---------------------------------------------------------------- phase 3: constraint rows + Newton
// MC = MAXCON_S: envs with ncon <= MAXCON_S; MC = MAXCON: the others (see MapSolveT).  The small class runs 4 blocks of
// 5 envs per SM (<= 102 registers), its warps in block-lockstep through the Newton iteration (constraint.cuh: newton).
#ifndef ADROIT_SOLVE_BLOCKS
#define ADROIT_SOLVE_BLOCKS 4
#endif
template <typename T, int MC>
__global__ void __launch_bounds__(MapSolveK<T, MC>::W * 32, (MC < MAXCON && sizeof(T) == 4) ? ADROIT_SOLVE_BLOCKS : 1)
    solve_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st, const __grid_constant__ PArgs a) {
  PDL_WAIT()
  // class routing: the big class (almost always empty) looks at the contact count BEFORE staging anything; the common class
  // stages first -- its bulk copies are in flight while the count arrives -- and leaves afterwards if the env is not its own
  if (MC == MAXCON) {
    const int env_ = a.env0 + blockIdx.x * MapSolveK<T, MC>::W + (threadIdx.x >> 5);
    if (env_ >= a.env1) return;
    if (st.ctxi[(size_t)env_ * CTX_I + SM_I + GI_NCON] <= MAXCON_S) return;
  }
  using MapK = MapSolveK<T, MC>;
  PHASE_PROLOGUE(MapK)
  c.ncon = c.gi[GI_NCON];
  if (MC < MAXCON && c.ncon > MAXCON_S) return;
  make_constraint(c, st.overflow);
  newton(c);
  if (lane == 0) c.gi[GI_NSIMPLE] = c.nsimple;
  __syncwarp();
  stage_out(c);
}

// ---------------------------------------------------------------- phase 4: noslip, sensors, Euler, task epilogue
// BIG: envs with more than MAXCON_M contacts (contact tables in the context block, see MapFinishBig); a warp whose env belongs
// to the other class exits before staging anything
template <typename T, bool BIG>
__global__ void __launch_bounds__(WPB * 32) finish_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st,
                                                           const __grid_constant__ PArgs a) {
  PDL_WAIT()
  if (BIG) {      // (class routing as in solve_kernel)
    const int env_ = a.env0 + blockIdx.x * WPB + (threadIdx.x >> 5);
    if (env_ >= a.env1) return;
    if (st.ctxi[(size_t)env_ * CTX_I + SM_I + GI_NCON] <= MAXCON_M) return;
  }
  using MapF = typename MapFinishPick<BIG>::type;
  PHASE_PROLOGUE(MapF)
  if (!BIG && c.gi[GI_NCON] > MAXCON_M) return;
  const int nv = m.nv;
  c.ncon = c.gi[GI_NCON];
  c.nsimple = c.gi[GI_NSIMPLE];
  T *qacc = c.template V<V_QACC>(), *qfc = c.template V<V_QFC>();
  if (m.noslip_iterations > 0) {
    noslip(c);
    constraint_force(c);
    T* u = c.template V<V_GRAD>();   // (the noslip pass is done with its copy)
    symv(c.template G<G_MINV>(), qfc, u, nv, lane);
    __syncwarp();
    const T* qas = c.template V<V_QAS>();
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) qacc[j] = qas[j] + u[j];
    __syncwarp();
  }
  // mj_Euler below factors (M + hB) in the shared-memory region that held M^-1: M is fetched from the context block by ONE bulk
  // copy issued here (second phase of the warp's stage-in mbarrier), in flight while the sensor / task-frame work runs
  bool m_prefetched = false;
#if ADROIT_BULK_STAGE
  if (a.integrate && m.any_damping) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the generic-proxy reads of M^-1 are done
    __syncwarp();
    if (lane == 0) {
      mbar_expect_tx(bar_, (unsigned)(NTRI * sizeof(T)));
      bulk_g2s(c.template G<G_MINV>(), c.gb + O_M, (unsigned)(NTRI * sizeof(T)), bar_);
    }
    m_prefetched = true;
  }
#endif
  T touch = touch_sensor(c);
  cache_task_frames(c, touch);
  if (a.debug) {
    const T *qfs = c.template V<V_QFS>(), *qas = c.template V<V_QAS>();
    _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) {
      c.template G<G_DBG_QFS>()[j] = qfs[j];
      c.template G<G_DBG_QAS>()[j] = qas[j];
      c.template G<G_DBG_QACC>()[j] = qacc[j];
      c.template G<G_DBG_QFC>()[j] = qfc[j];
    }
    for (int i = lane; i < 3 * m.nbody; i += 32) c.template G<G_DBG_XPOS>()[i] = c.template S<O_XPOS>()[i];
    for (int i = lane; i < 4 * m.nbody; i += 32) c.template G<G_DBG_XQUAT>()[i] = c.template S<O_XQUAT>()[i];
    for (int i = lane; i < 6 * c.ncon; i += 32) c.template G<G_DBG_EF>()[i] = c.template S<O_EF>()[i];
  }
  __syncwarp();
  T *qpos = c.template V<V_QPOS>(), *qvel = c.template V<V_QVEL>(), *warm = c.template V<V_WARM>();
  if (a.integrate) {
    if (m_prefetched) mbar_wait(bar_, 1);
    // mj_checkAcc: a non-finite acceleration resets the env's data (MuJoCo then re-runs mj_forward on the reset
    // state and integrates it; here the env simply restarts from the reset state on the next substep)
    if (warp_bad(qacc, nv, lane)) {
      zero_state(c);
    } else {
      // mj_Euler: semi-implicit with implicit joint damping, (M + h*diag(damping)) keeps the tree sparsity of M
      T* x = c.template V<V_GRAD>();
      const T h = m.timestep;
      if (m.any_damping) {
        T* H = c.template G<G_MINV>();   // the staged copy of M^-1 (noslip is done with it) doubles as the factor buffer
        const T* M = c.template S<O_M>();
        const T *qfs = c.template V<V_QFS>();
        if (!m_prefetched) {
          for (int p = lane; p < NTRI; p += 32) H[p] = M[p];
          __syncwarp();
        }
        _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) {
          H[tri(j, j)] += h * m.dof_damping[j];
          x[j] = qfs[j] + qfc[j];
        }
        __syncwarp();
        ltdl_factor(m, H, c.template V<V_SEARCH>(), lane);
        ltdl_solve(H, c.template V<V_SEARCH>(), x, nv, lane);
      } else {
        _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) x[j] = qacc[j];
        __syncwarp();
      }
      _Pragma("unroll 1") for (int j = lane; j < nv; j += 32) {
        T v = qvel[j] + h * x[j];
        qvel[j] = v;
        qpos[j] += h * v;
        warm[j] = qacc[j];
      }
      __syncwarp();
    }
  }
  bool reset_now = false;
  if (a.epilogue) {
    T r;
    int g, dn;
    task_reward(m, qpos, qvel, c.template S<O_TF>(), lane, r, g, dn);
    int tnew = st.t[env] + 1;
    bool trunc = (st.flags & ADROIT_AUTO_RESET) && tnew >= m.horizon;
    reset_now = trunc || ((st.flags & ADROIT_AUTO_RESET) && (st.flags & ADROIT_RESET_ON_DONE) && dn);
    if (lane == 0) {
      st.t[env] = tnew;
      st.ep_return[env] += r;
      st.ep_goal[env] += g;
      st.reset_flag[env] = reset_now ? 1 : 0;
      if (a.reward) a.reward[env] = (float)r;
      if (a.goal) a.goal[env] = (uint8_t)g;
      if (a.done) a.done[env] = (uint8_t)(dn || trunc);
    }
    __syncwarp();
  }
  if (a.write_obs && a.obs && !reset_now) write_obs(m, qpos, qvel, c.template S<O_TF>(), a.obs + (size_t)env * m.obs_dim, lane);
  __syncwarp();
  stage_out(c);
}

template <typename T> __global__ void obs_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st, float* obs) {
  int env = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (env >= st.n) return;
  const T* b = st.ctx + (size_t)env * CTX_T;
  write_obs(m, b + V_QPOS * NVP, b + V_QVEL * NVP, b + O_TF, obs + (size_t)env * m.obs_dim, lane);
}

// [n][w] <-> padded [n][stride] conversions between the fp64 host staging buffer and the typed device arrays
template <typename T> __global__ void pack_kernel(T* dst, const double* src, int n, int w, int stride) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * w) return;
  dst[(size_t)(i / w) * stride + i % w] = (T)src[i];
}
template <typename T> __global__ void unpack_kernel(double* dst, const T* src, int n, int w, int stride, int offset) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * w) return;
  dst[i] = (double)src[(size_t)(i / w) * stride + offset + i % w];
}
__global__ void reduce_kernel(const double* acc, const int* t, const int* ep_goal, const float* retf, const double* retd, int n, int thr,
                              int flush, double* out4) {
  __shared__ double sh[4][32];
  double v[4] = {0, 0, 0, 0};
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
    for (int k = 0; k < 4; k++) v[k] += acc[4 * e + k];
    if (flush && t[e] > 0) {
      v[0] += retf ? (double)retf[e] : retd[e];
      v[1] += ep_goal[e] > 0;
      v[2] += ep_goal[e] > thr;
      v[3] += 1;
    }
  }
  for (int k = 0; k < 4; k++) v[k] = warp_sum(v[k]);
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) for (int k = 0; k < 4; k++) sh[k][warp] = v[k];
  __syncthreads();
  if (warp == 0) {
    for (int k = 0; k < 4; k++) {
      double x = lane < (int)(blockDim.x >> 5) ? sh[k][lane] : 0.0;
      x = warp_sum(x);
      if (lane == 0) atomicAdd(out4 + k, x);
    }
  }
}

// ---- adroit_reduce_episode_noalloc: the same statistics with the partial sums in a file-scope array instead of cudaMalloc'ed memory.
// One block per NOALLOC_BLOCK envs writes its four partial sums into its own row of the scratch; a second one-block launch adds the rows in a
// fixed order (the result does not depend on block scheduling, unlike the atomicAdd of reduce_kernel).  Every row a call reads was written by
// that same call, so nothing in the scratch has to start at zero and the previous call's contents are harmless.
constexpr int NOALLOC_BLOCK = 256;
constexpr int NOALLOC_MAX_ENVS = ADROIT_NOALLOC_MAX_ENVS;
constexpr int NOALLOC_MAX_BLOCKS = NOALLOC_MAX_ENVS / NOALLOC_BLOCK;
static_assert(NOALLOC_MAX_ENVS % NOALLOC_BLOCK == 0, "ADROIT_NOALLOC_MAX_ENVS must be a multiple of the block size");
static __device__ __align__(256) double g_noalloc_partials[NOALLOC_MAX_BLOCKS * 4];   // 8 KB per GPU that loads the library, held until exit

__device__ __forceinline__ void block_sum4(double (&v)[4], double (*sh)[32]) {   // result valid in warp 0 (every lane)
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  _Pragma("unroll") for (int k = 0; k < 4; k++) v[k] = warp_sum(v[k]);
  if (lane == 0) { _Pragma("unroll") for (int k = 0; k < 4; k++) sh[k][warp] = v[k]; }
  __syncthreads();
  if (warp == 0)
    _Pragma("unroll") for (int k = 0; k < 4; k++) v[k] = warp_sum(lane < (int)(blockDim.x >> 5) ? sh[k][lane] : 0.0);
}
__global__ void __launch_bounds__(NOALLOC_BLOCK) reduce_partial_kernel(const double* acc, const int* t, const int* ep_goal, const float* retf,
                                                                       const double* retd, int n, int thr, int flush, double* partials) {
  __shared__ double sh[4][32];
  double v[4] = {0, 0, 0, 0};
  int e = blockIdx.x * NOALLOC_BLOCK + threadIdx.x;
  if (e < n) {
    _Pragma("unroll") for (int k = 0; k < 4; k++) v[k] = acc[4 * e + k];
    if (flush && t[e] > 0) {
      v[0] += retf ? (double)retf[e] : retd[e];
      v[1] += ep_goal[e] > 0;
      v[2] += ep_goal[e] > thr;
      v[3] += 1;
    }
  }
  block_sum4(v, sh);
  _Pragma("unroll") for (int k = 0; k < 4; k++) if (threadIdx.x == k) partials[4 * blockIdx.x + k] = v[k];
}
__global__ void __launch_bounds__(NOALLOC_BLOCK) reduce_final_kernel(const double* partials, int nblocks, double* out4) {
  __shared__ double sh[4][32];
  double v[4] = {0, 0, 0, 0};
  for (int b = threadIdx.x; b < nblocks; b += NOALLOC_BLOCK)
    _Pragma("unroll") for (int k = 0; k < 4; k++) v[k] += partials[4 * b + k];
  block_sum4(v, sh);
  _Pragma("unroll") for (int k = 0; k < 4; k++) if (threadIdx.x == k) out4[k] = v[k];
}

}  // namespace adroit


// ================================================================ host side
using namespace adroit;

static thread_local char g_err[512];
#define FAIL(...) do { snprintf(g_err, sizeof g_err, __VA_ARGS__); return 1; } while (0)
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) FAIL("%s failed: %s", #call, cudaGetErrorString(e_)); } while (0)

namespace {

// model blob reader (mjcf.py::Model.to_blob).  Every table entry is checked against the blob length: a truncated or
// corrupt file yields an error from adroit_create, never an out-of-bounds read.
struct Blob {
  const unsigned char* b;
  size_t len;
  void validate() const {
    uint64_t nf;
    memcpy(&nf, b + 8, 8);
    if (nf > (len - 16) / 48) throw std::string("model blob: field table exceeds the blob");
    for (uint64_t i = 0; i < nf; i++) {
      const unsigned char* e = b + 16 + 48 * i;
      uint32_t dt, cnt;
      uint64_t off;
      memcpy(&dt, e + 32, 4); memcpy(&cnt, e + 36, 4); memcpy(&off, e + 40, 8);
      const uint64_t w = dt == 0 ? 4 : 8;
      if (dt > 1 || off > len || (uint64_t)cnt * w > len - off || !memchr(e, 0, 32)) throw std::string("model blob: corrupt field entry");
    }
  }
  const void* find(const char* name, int want_dt, int* count) const {
    uint64_t nf;
    memcpy(&nf, b + 8, 8);
    for (uint64_t i = 0; i < nf; i++) {
      const unsigned char* e = b + 16 + 48 * i;
      if (strncmp((const char*)e, name, 32) == 0) {
        uint32_t dt, cnt;
        uint64_t off;
        memcpy(&dt, e + 32, 4); memcpy(&cnt, e + 36, 4); memcpy(&off, e + 40, 8);
        if ((int)dt != want_dt) return nullptr;
        if (count) *count = (int)cnt;
        return b + off;
      }
    }
    return nullptr;
  }
  std::vector<int> ints(const char* n, size_t min_len = 0) const {
    int c = 0;
    const int* p = (const int*)find(n, 0, &c);
    if (!p) throw std::string("model blob misses int field ") + n;
    if ((size_t)c < min_len) throw std::string("model blob: field too short: ") + n;
    return std::vector<int>(p, p + c);
  }
  std::vector<double> reals(const char* n, size_t min_len = 0) const {
    int c = 0;
    const double* p = (const double*)find(n, 1, &c);
    if (!p) throw std::string("model blob misses real field ") + n;
    if ((size_t)c < min_len) throw std::string("model blob: field too short: ") + n;
    return std::vector<double>(p, p + c);
  }
};

void quat2mat_h(const double* q, double* R) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}

// device arena holding every model table
struct Arena {
  std::vector<unsigned char> host;
  unsigned char* dev = nullptr;
  template <typename U> size_t put(const std::vector<U>& v) {
    size_t off = (host.size() + 15) & ~size_t(15);
    host.resize(off + v.size() * sizeof(U) + 16);
    if (!v.empty()) memcpy(host.data() + off, v.data(), v.size() * sizeof(U));
    return off;
  }
};
template <typename T> std::vector<T> conv(const std::vector<double>& v) { return std::vector<T>(v.begin(), v.end()); }

}  // namespace

struct adroit_sim {
  int n = 0, device = 0;
  bool f64 = false;
  uint32_t flags = 0;
  int nv = 0, nu = 0, obs_dim = 0, npar = 0, horizon = 0, frame_skip = 0, nbody = 0, nsite = 0, task = 0;
  Arena arena;
  DModel<float> mf;
  DModel<double> md;
  DState<float> sf;
  DState<double> sd;
  std::vector<void*> allocs;
  double* stage = nullptr;      // device fp64 staging buffer
  size_t stage_len = 0;
  float *d_action = nullptr, *d_obs = nullptr, *d_reward = nullptr;
  uint8_t *d_done = nullptr, *d_goal = nullptr;
  double* d_out4 = nullptr;
  int64_t launches = 0;
  bool profile = false;
  std::vector<cudaEvent_t> prof_ev[5];   // event pairs per phase (4 = reset_kernel, not reported), pending accumulation
  // sub-batch pipelining (run_t): the envs are split into nsub contiguous ranges, each advanced by its own chain of
  // phase kernels on its own stream, so that the tail of one range's kernel overlaps the other ranges' kernels
  int nsub = 1;
  bool pdl = false;             // phase kernels chained with programmatic dependent launch (ADROIT_PDL=1)
  bool light_reset = true;      // auto-reset of a step through reset_kernel (ADROIT_LIGHT_RESET=0: masked full forward pass)
  std::vector<cudaStream_t> sub_stream;
  std::vector<cudaEvent_t> sub_done;
  cudaEvent_t fork_ev = nullptr;
};

namespace {

template <typename T> void fill_model(DModel<T>& m, Arena& A, const Blob& B) {
  auto sizes = B.ints("sizes", 12);
  m.nv = sizes[0]; m.nbody = sizes[1]; m.ngeom = sizes[2]; m.nsite = sizes[3]; m.nu = sizes[4]; m.ntendon = sizes[5];
  m.npair = sizes[6]; m.npar = sizes[7]; m.obs_dim = sizes[8]; m.frame_skip = sizes[9]; m.task = sizes[10]; m.horizon = sizes[11];
  if (m.nv > 36 || m.nbody > NBP || m.ngeom > NGP || m.nsite > NSP || m.npar > NPARP || m.nu > NVP)
    throw std::string("model exceeds the compiled capacity of the CUDA kernels");
  if (m.npair >= 65536) throw std::string("more candidate pairs than the collide kernel's 16-bit pair list can index");
  auto body_parent = B.ints("body_parent"), jntadr = B.ints("body_jntadr"), jntnum = B.ints("body_jntnum");
  auto depth = B.ints("body_depth");
  int nb = m.nbody, maxd = 0;
  for (int b = 0; b < nb; b++) maxd = depth[b] > maxd ? depth[b] : maxd;
  m.nlevel = maxd + 1;
  std::vector<int> level_start(m.nlevel + 1, 0), level_body, child_start(nb + 1, 0), child_list;
  for (int L = 0; L <= maxd; L++) {
    level_start[L] = (int)level_body.size();
    for (int b = 0; b < nb; b++) if (depth[b] == L) level_body.push_back(b);
  }
  level_start[m.nlevel] = (int)level_body.size();
  for (int p = 0; p < nb; p++) {
    child_start[p] = (int)child_list.size();
    for (int b = 1; b < nb; b++) if (body_parent[b] == p) child_list.push_back(b);
  }
  child_start[nb] = (int)child_list.size();
  std::vector<unsigned long long> dofmask(nb, 0ull);
  for (int b = 1; b < nb; b++) {
    dofmask[b] = dofmask[body_parent[b]];
    for (int k = 0; k < jntnum[b]; k++) dofmask[b] |= 1ull << (jntadr[b] + k);
  }
  auto iquat = B.reals("body_iquat"), gquat = B.reals("geom_quat");
  std::vector<double> imat(9 * nb), gmat(9 * m.ngeom);
  for (int b = 0; b < nb; b++) quat2mat_h(&iquat[4 * b], &imat[9 * b]);
  for (int g = 0; g < m.ngeom; g++) quat2mat_h(&gquat[4 * g], &gmat[9 * g]);
  auto act_dof = B.ints("act_dof");
  for (size_t i = 0; i < act_dof.size(); i++)
    for (size_t j = i + 1; j < act_dof.size(); j++)
      if (act_dof[i] == act_dof[j]) throw std::string("two actuators on one dof are outside the Adroit subset");

  struct Fix { size_t off; const void** slot; };
  std::vector<Fix> fix;
  auto PI = [&](const int*& slot, const std::vector<int>& v) { fix.push_back({A.put(v), (const void**)&slot}); };
  auto PR = [&](const T*& slot, const std::vector<double>& v) { fix.push_back({A.put(conv<T>(v)), (const void**)&slot}); };
  PI(m.body_parent, body_parent); PI(m.body_jntadr, jntadr); PI(m.body_jntnum, jntnum); PI(m.body_par, B.ints("body_par"));
  PI(m.level_start, level_start); PI(m.level_body, level_body); PI(m.child_start, child_start); PI(m.child_list, child_list);
  fix.push_back({A.put(dofmask), (const void**)&m.body_dofmask});
  PR(m.body_pos, B.reals("body_pos")); PR(m.body_quat, B.reals("body_quat")); PR(m.body_ipos, B.reals("body_ipos"));
  PR(m.body_imat, imat); PR(m.body_mass, B.reals("body_mass")); PR(m.body_inertia, B.reals("body_inertia"));
  PR(m.body_invweight0, B.reals("body_invweight0"));
  PI(m.jnt_type, B.ints("jnt_type")); PI(m.jnt_limited, B.ints("jnt_limited")); PI(m.dof_body, B.ints("dof_body"));
  PI(m.dof_parent, B.ints("dof_parent"));
  PR(m.jnt_pos, B.reals("jnt_pos")); PR(m.jnt_axis, B.reals("jnt_axis")); PR(m.jnt_range, B.reals("jnt_range"));
  PR(m.jnt_margin, B.reals("jnt_margin")); PR(m.dof_damping, B.reals("dof_damping")); PR(m.dof_armature, B.reals("dof_armature"));
  PR(m.dof_frictionloss, B.reals("dof_frictionloss")); PR(m.dof_invweight0, B.reals("dof_invweight0"));
  PI(m.geom_type, B.ints("geom_type")); PI(m.geom_body, B.ints("geom_body"));
  PR(m.geom_pos, B.reals("geom_pos")); PR(m.geom_mat, gmat); PR(m.geom_size, B.reals("geom_size")); PR(m.geom_rbound, B.reals("geom_rbound"));
  PI(m.site_body, B.ints("site_body")); PI(m.site_par, B.ints("site_par")); PR(m.site_pos, B.reals("site_pos"));
  PI(m.pair_geom, B.ints("pair_geom")); PI(m.pair_condim, B.ints("pair_condim"));
  {  // packed broad-phase records
    auto pg = B.ints("pair_geom"), gt = B.ints("geom_type");
    auto pm = B.reals("pair_margin"), rb = B.reals("geom_rbound");
    std::vector<int> pw(m.npair);
    std::vector<double> pb(m.npair);
    for (int k = 0; k < m.npair; k++) {
      const int g1 = pg[2 * k], g2 = pg[2 * k + 1];
      if (g1 > 255 || g2 > 255) throw std::string("geom index beyond the packed pair record");
      pw[k] = g1 | (g2 << 8) | (gt[g1] << 16) | (gt[g2] << 24);
      pb[k] = gt[g1] == G_PLANE ? rb[g2] + pm[k] : rb[g1] + rb[g2] + pm[k];
    }
    PI(m.pair_word, pw); PR(m.pair_bound, pb);
  }
  PR(m.pair_friction, B.reals("pair_friction")); PR(m.pair_margin, B.reals("pair_margin"));
  PI(m.tendon_dof, B.ints("tendon_dof")); PI(m.tendon_limited, B.ints("tendon_limited"));
  PR(m.tendon_coef, B.reals("tendon_coef")); PR(m.tendon_range, B.reals("tendon_range")); PR(m.tendon_margin, B.reals("tendon_margin"));
  PR(m.tendon_invweight0, B.reals("tendon_invweight0"));
  {  // tree-sparse factorisation schedules (dynamics.cuh: ltdl_factor / minv_tree)
    auto dof_parent = B.ints("dof_parent");
    int nv = m.nv;
    std::vector<std::vector<int>> anc(nv);          // proper ancestors, nearest first
    std::vector<int> depth(nv, 0);
    for (int k = 0; k < nv; k++) {
      for (int i = dof_parent[k]; i >= 0; i = dof_parent[i]) anc[k].push_back(i);
      depth[k] = (int)anc[k].size();
    }
    std::vector<int> ld_start(nv + 1, 0), ld_pair, wl_pair, tri_ab;
    for (int k = 0; k < nv; k++) {
      ld_start[k] = (int)ld_pair.size();
      for (size_t a = 0; a < anc[k].size(); a++)
        for (size_t b = a; b < anc[k].size(); b++) {                                                    // i >= j
          const int i = anc[k][a], j = anc[k][b], g = i >> 2;
          ld_pair.push_back((4 * (g + 1) * (2 * g + (i & 3)) + j) | (i << 10) | (j << 16));             // model.cuh tri(i, j)
        }
    }
    ld_start[nv] = (int)ld_pair.size();
    int maxd = 0;
    for (int k = 0; k < nv; k++) maxd = depth[k] > maxd ? depth[k] : maxd;
    for (int L = 1; L <= maxd; L++) {               // W rows by depth; each level padded to whole warps
      for (int k = 0; k < nv; k++)
        if (depth[k] == L) for (int j : anc[k]) wl_pair.push_back(k | (j << 8));
      while (wl_pair.size() % 32) wl_pair.push_back(-1);
    }
    m.n_wl = (int)wl_pair.size();
    {  // level-scheduled L^T D L program (dynamics.cuh: ltdl_factor): dofs of equal depth are eliminated together, deepest level
       // first; within a level the (ancestor, ancestor) updates are split into rounds whose targets are all different (dofs
       // of different fingers share the arm / wrist ancestors), so a round is plain load-FMA-store with no atomics and the
       // summation order is fixed
      auto toff = [](int i, int j) { if (i < j) { int t = i; i = j; j = t; } const int g = i >> 2; return 4 * (g + 1) * (2 * g + (i & 3)) + j; };
      std::vector<int> steps, items;
      for (int L = maxd; L >= 0; L--) {
        std::vector<int> piv;
        for (int k = 0; k < nv; k++) if (depth[k] == L) piv.push_back(k);
        if (piv.empty()) continue;
        steps.push_back(0 | ((int)items.size() << 2));
        for (int k : piv) items.push_back(k | (toff(k, k) << 8));
        std::vector<std::vector<int>> rounds;
        std::vector<std::vector<char>> used;
        for (int k : piv)
          for (size_t a = 0; a < anc[k].size(); a++)
            for (size_t b = a; b < anc[k].size(); b++) {
              const int i = anc[k][a], j = anc[k][b], t = toff(i, j);
              size_t r = 0;
              while (r < rounds.size() && used[r][t]) r++;
              if (r == rounds.size()) { rounds.emplace_back(); used.emplace_back(1024, 0); }
              used[r][t] = 1;
              rounds[r].push_back(t | (i << 10) | (j << 16) | (k << 22));
            }
        for (auto& rd : rounds) {
          steps.push_back(1 | ((int)items.size() << 2));
          items.insert(items.end(), rd.begin(), rd.end());
        }
        steps.push_back(2 | ((int)items.size() << 2));
        for (int k : piv) for (int i : anc[k]) items.push_back(toff(k, i) | (k << 10));
      }
      steps.push_back(3 | ((int)items.size() << 2));
      if ((int)steps.size() > MAXLVSTEP) throw std::string("factorisation program longer than model.cuh's MAXLVSTEP");
      for (int i = 0; i < MAXLVSTEP; i++) m.lv_steps[i] = i < (int)steps.size() ? steps[i] : steps.back();
      m.n_lvitems = (int)items.size();
      for (int i = 0; i < 32; i++) items.push_back(0);
      PI(m.lv_items, items);
    }
    std::vector<unsigned char> lca((size_t)nv * nv, 255);
    for (int a = 0; a < nv; a++)
      for (int b = 0; b <= a; b++) {
        tri_ab.push_back(a | (b << 8));
        int x = a, y = b;
        while (x != y && x >= 0 && y >= 0) { if (x > y) x = dof_parent[x]; else y = dof_parent[y]; }
        unsigned char v = (x == y && x >= 0) ? (unsigned char)x : 255;
        lca[(size_t)a * nv + b] = v; lca[(size_t)b * nv + a] = v;
      }
    m.n_tri = (int)tri_ab.size();
    std::vector<int> tri_abk(tri_ab.size());
    for (size_t q = 0; q < tri_ab.size(); q++) {
      int a = tri_ab[q] & 255, b = tri_ab[q] >> 8;
      tri_abk[q] = a | (b << 8) | ((int)lca[(size_t)a * nv + b] << 16);
    }
    PI(m.tri_abk, tri_abk);
    PI(m.ld_start, ld_start); PI(m.ld_pair, ld_pair); PI(m.wl_pair, wl_pair); PI(m.tri_ab, tri_ab);
    for (int i = 0; i < NVP + 4; i++) m.ld_start_v[i] = ld_start[i <= nv ? i : nv];
    if (maxd > MAXDEPTH) throw std::string("dof ancestor chain longer than dynamics.cuh's MAXDEPTH");
    std::vector<int> anc_tab((size_t)nv * MAXDEPTH, -1);
    for (int k = 0; k < nv; k++)
      for (size_t t = 0; t < anc[k].size(); t++) anc_tab[(size_t)k * MAXDEPTH + t] = anc[k][t];
    PI(m.dof_depth, depth); PI(m.anc_tab, anc_tab);
    fix.push_back({A.put(lca), (const void**)&m.lca});
  }
  {  // pixel path tables (render.cuh)
    auto ri = B.ints("render_int");
    m.nrender = (int)ri.size() / 4;
    if (m.nrender > MAXRG) throw std::string("scene has more render primitives than render.cuh's MAXRG");
    PI(m.render_int, ri);
    PR(m.render_real, B.reals("render_real"));
    auto rc = B.reals("render_cam", 8);
    for (int i = 0; i < 8; i++) m.render_cam[i] = (T)rc[i];
    auto lp = B.reals("light_pos"), ld = B.reals("light_diffuse");
    auto lb = B.ints("light_body");
    m.light_body = lb.empty() ? 0 : lb[0];
    for (int i = 0; i < 3; i++) { m.light_pos[i] = lp.size() >= 3 ? (T)lp[i] : T(0); m.light_diffuse[i] = ld.size() >= 3 ? (T)ld[i] : T(0); }
  }
  PI(m.act_dof, act_dof); PR(m.act_ctrlrange, B.reals("act_ctrlrange")); PR(m.act_gain, B.reals("act_gain")); PR(m.act_bias, B.reals("act_bias"));
  // stash offsets; pointers are patched once the arena is on the device
  for (auto& f : fix) *f.slot = (const void*)f.off;

  {
    auto td = B.ints("tendon_dof");
    auto tc = B.reals("tendon_coef");
    if (m.ntendon > MAXTENDON || (int)td.size() < 2 * m.ntendon || (int)tc.size() < 2 * m.ntendon) throw std::string("more fixed tendons than model.cuh's MAXTENDON");
    for (int i = 0; i < 2 * MAXTENDON; i++) { m.tdof[i] = i < 2 * m.ntendon ? td[i] : -1; m.tcoef[i] = i < 2 * m.ntendon ? (T)tc[i] : T(0); }
  }
  auto ti = B.ints("task_int", 16);
  for (int i = 0; i < 16; i++) m.task_int[i] = ti[i];
  auto par0 = B.reals("par0");
  for (int i = 0; i < NPARP; i++) m.par0[i] = i < (int)par0.size() ? (T)par0[i] : T(0);
  m.touch_site = B.ints("touch_site", 1)[0];
  auto ssz = B.reals("site_size");
  for (int i = 0; i < 3; i++) m.touch_size[i] = m.touch_site >= 0 ? (T)ssz[3 * m.touch_site + i] : T(0);
  auto o = B.reals("opt_real", 15);
  auto oi = B.ints("opt_int", 3);
  m.timestep = (T)o[0]; m.gravity[0] = (T)o[1]; m.gravity[1] = (T)o[2]; m.gravity[2] = (T)o[3];
  m.tolerance = (T)o[4]; m.noslip_tolerance = (T)o[5]; m.impratio = (T)o[6]; m.solref[0] = (T)o[7]; m.solref[1] = (T)o[8];
  for (int i = 0; i < 5; i++) m.solimp[i] = (T)o[9 + i];
  m.mpr_tolerance = (T)o[14];
  m.iterations = oi[0]; m.noslip_iterations = oi[1]; m.mpr_iterations = oi[2];
  m.meaninertia = (T)B.reals("meaninertia", 1)[0];
  m.pen_len = m.tar_len = T(1);
  if (m.task == T_PEN) {  // pen_v0.py:57-58: |object_top - object_bottom| (both sites sit on one body)
    auto sp = B.reals("site_pos");
    auto len = [&](int a, int b) {
      double s = 0;
      for (int i = 0; i < 3; i++) s += (sp[3 * a + i] - sp[3 * b + i]) * (sp[3 * a + i] - sp[3 * b + i]);
      return std::sqrt(s);
    };
    m.pen_len = (T)len(ti[4], ti[5]);
    m.tar_len = (T)len(ti[6], ti[7]);
  }
  auto damp = B.reals("dof_damping");
  m.any_damping = 0;
  for (double d : damp) m.any_damping |= d > 0;
}

template <typename T> void patch_model(DModel<T>& m, unsigned char* dev) {
  // every pointer member currently holds an arena offset
  const void** p = (const void**)&m.body_parent;
  const void** end = (const void**)&m.task_int;
  for (; p < end; p++) *p = (const void*)(dev + (size_t)*p);
}

template <typename T> int alloc_state(adroit_sim* s, DState<T>& st) {
  int n = s->n;
  auto A = [&](void** p, size_t bytes) -> int {
    CK(cudaMalloc(p, bytes));
    CK(cudaMemset(*p, 0, bytes));
    s->allocs.push_back(*p);
    return 0;
  };
  st.n = n;
  if (A((void**)&st.ctx, sizeof(T) * (size_t)n * CTX_T) || A((void**)&st.ctxi, sizeof(int) * (size_t)n * CTX_I) ||
      A((void**)&st.t, sizeof(int) * n) || A((void**)&st.episode, sizeof(int) * n) || A((void**)&st.ep_goal, sizeof(int) * n) ||
      A((void**)&st.ep_return, sizeof(T) * n) || A((void**)&st.acc, sizeof(double) * 4 * n) || A((void**)&st.reset_flag, n) ||
      A((void**)&st.overflow, sizeof(unsigned long long) * 2))
    return 1;
  return 0;
}

static int g_extra_smem[5] = {0, 0, 0, 0, 0};   // ADROIT_EXTRA_SMEM_<phase 0..3>: bytes of unused dynamic shared memory added to that phase's launches (occupancy experiments)
// dynamic shared memory of a phase kernel: the WPB map slices + 16 bytes per warp for its stage-in mbarrier
// (+ MAP::EXTRA bytes per block where the map declares them: the collide kernel's MPR queues)
template <class MAP, class = void> struct map_extra { static constexpr size_t value = 0; };
template <class MAP> struct map_extra<MAP, decltype((void)MAP::EXTRA)> { static constexpr size_t value = MAP::EXTRA; };
template <typename T, class MAP> constexpr size_t phase_smem() { return MAP::W * map_bytes<T, MAP>() + MAP::W * 16 + map_extra<MAP>::value; }

template <typename T, class MAP, typename K>
int launch_phase(adroit_sim* s, int phase, K kernel, const DModel<T>& m, const DState<T>& st, const PArgs& a, cudaStream_t stream) {
  size_t smem = phase_smem<T, MAP>() + g_extra_smem[phase];   // (occupancy experiments, tools/gpu_occ_sweep.sh)
  int blocks = (a.env1 - a.env0 + MAP::W - 1) / MAP::W;
  if (blocks <= 0) return 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (s->profile) {
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, stream));
  }
  if (s->pdl) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(MAP::W * 32); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, kernel, m, st, a));
  } else {
    kernel<<<blocks, MAP::W * 32, smem, stream>>>(m, st, a);
    CK(cudaGetLastError());
  }
  if (s->profile) {
    CK(cudaEventRecord(e1, stream));
    s->prof_ev[phase].push_back(e0);
    s->prof_ev[phase].push_back(e1);
  }
  s->launches++;
  return 0;
}

template <typename T> int configure_kernels() {
  for (int p = 0; p < 4; p++) { char nm[32]; snprintf(nm, sizeof nm, "ADROIT_EXTRA_SMEM_%d", p); const char* e = getenv(nm); g_extra_smem[p] = e ? atoi(e) : 0; }
  CK(cudaFuncSetAttribute(smooth_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapSmooth>() + g_extra_smem[0]));
  CK(cudaFuncSetAttribute(collide_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapCollide>() + g_extra_smem[1]));
  CK(cudaFuncSetAttribute(solve_kernel<T, MAXCON>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapSolveK<T, MAXCON>>() + g_extra_smem[2]));
  if (MAXCON_S < MAXCON) CK(cudaFuncSetAttribute(solve_kernel<T, MAXCON_S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapSolveK<T, MAXCON_S>>() + g_extra_smem[2]));
  CK(cudaFuncSetAttribute(finish_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapFinish>() + g_extra_smem[3]));
  CK(cudaFuncSetAttribute(finish_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapFinishBig>() + g_extra_smem[3]));
  CK(cudaFuncSetAttribute(reset_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)phase_smem<T, MapReset>()));
  return 0;
}

// one mj_forward (+ optional Euler / task epilogue) over the envs selected by a.mask: the four phase kernels in order
template <typename T> int forward_pass(adroit_sim* s, const DModel<T>& m, const DState<T>& st, PArgs a, cudaStream_t stream) {
  if (launch_phase<T, MapSmooth>(s, 0, smooth_kernel<T>, m, st, a, stream)) return 1;
  if (launch_phase<T, MapCollide>(s, 1, collide_kernel<T>, m, st, a, stream)) return 1;
  if (MAXCON_S < MAXCON && launch_phase<T, MapSolveK<T, MAXCON_S>>(s, 2, solve_kernel<T, MAXCON_S>, m, st, a, stream)) return 1;
  if (launch_phase<T, MapSolveK<T, MAXCON>>(s, 2, solve_kernel<T, MAXCON>, m, st, a, stream)) return 1;
  if (launch_phase<T, MapFinish>(s, 3, finish_kernel<T, false>, m, st, a, stream)) return 1;
  if (!getenv("ADROIT_NO_BIG_FINISH") && launch_phase<T, MapFinishBig>(s, 3, finish_kernel<T, true>, m, st, a, stream)) return 1;   // (A/B switch)
  return 0;
}

// the launches of one request over the env range [a.env0, a.env1) on one stream
template <typename T>
int run_range(adroit_sim* s, const DModel<T>& m, const DState<T>& st, int mode, PArgs a, const float* action, const uint8_t* mask, float* obs,
              float* reward, uint8_t* done, uint8_t* goal, cudaStream_t stream) {
  if (mode == MODE_FORWARD) {            // sim.forward()
    a.mask = mask;
    return forward_pass(s, m, st, a, stream);
  }
  if (mode == MODE_RESET) {              // env.reset(): reset_model + mj_forward + get_obs
    a.mask = mask; a.reset = 1; a.obs = obs; a.write_obs = obs != nullptr;
    return forward_pass(s, m, st, a, stream);
  }
  // env.step(): action map, frame_skip x mj_step, task epilogue; then reset_model for the envs the epilogue flagged
  for (int k = 0; k < m.frame_skip; k++) {
    PArgs b = a;
    b.action = k == 0 ? action : nullptr;
    b.check = 1;
    b.integrate = 1;
    if (k == m.frame_skip - 1) { b.epilogue = 1; b.write_obs = obs != nullptr; b.obs = obs; b.reward = reward; b.done = done; b.goal = goal; }
    if (forward_pass(s, m, st, b, stream)) return 1;
  }
  if (st.flags & ADROIT_AUTO_RESET) {
    PArgs r = a;
    r.mask = st.reset_flag; r.reset = 1; r.obs = obs; r.write_obs = obs != nullptr;
    if (s->light_reset && !a.debug) return launch_phase<T, MapReset>(s, 4, reset_kernel<T>, m, st, r, stream);
    if (forward_pass(s, m, st, r, stream)) return 1;
  }
  return 0;
}

// Envs are independent, so a request is issued as nsub contiguous sub-batches, each a chain of phase kernels on its own
// stream (forked from / joined to the caller's stream with events: no host sync).  A phase kernel of one sub-batch
// then fills the SMs that the last, partly empty wave of another sub-batch's kernel leaves idle, and kernels bound by
// different resources (registers: collide, shared memory: solve) share an SM.  Per-phase profiling and debug exports
// use one range on the caller's stream.
// Host buffers of adroit_step_host: every sub-batch copies ITS slice of the action in at the head of its chain and its
// slices of the results out at the tail, on its own stream -- the copies of one sub-batch overlap the kernels of the others.
struct HostIO { const float* action; float* obs; float* reward; uint8_t* done; uint8_t* goal; };
static int host_in(adroit_sim* s, const HostIO* h, float* action_dev, int e0, int e1, cudaStream_t st) {
  if (h && h->action) CK(cudaMemcpyAsync(action_dev + (size_t)e0 * s->nu, h->action + (size_t)e0 * s->nu, (size_t)(e1 - e0) * s->nu * sizeof(float), cudaMemcpyHostToDevice, st));
  return 0;
}
static int host_out(adroit_sim* s, const HostIO* h, const float* obs, const float* reward, const uint8_t* done, const uint8_t* goal, int e0, int e1, cudaStream_t st) {
  if (!h) return 0;
  const size_t n = (size_t)(e1 - e0);
  if (h->obs) CK(cudaMemcpyAsync(h->obs + (size_t)e0 * s->obs_dim, obs + (size_t)e0 * s->obs_dim, n * s->obs_dim * sizeof(float), cudaMemcpyDeviceToHost, st));
  if (h->reward) CK(cudaMemcpyAsync(h->reward + e0, reward + e0, n * sizeof(float), cudaMemcpyDeviceToHost, st));
  if (h->done) CK(cudaMemcpyAsync(h->done + e0, done + e0, n, cudaMemcpyDeviceToHost, st));
  if (h->goal) CK(cudaMemcpyAsync(h->goal + e0, goal + e0, n, cudaMemcpyDeviceToHost, st));
  return 0;
}
template <typename T>
int run_t(adroit_sim* s, const DModel<T>& m, const DState<T>& st, int mode, const float* action, const uint8_t* mask, float* obs, float* reward,
          uint8_t* done, uint8_t* goal, int debug, cudaStream_t stream, const HostIO* hio = nullptr) {
  PArgs a;
  memset(&a, 0, sizeof a);
  a.debug = debug;
  int nsub = (s->profile || debug) ? 1 : s->nsub;
  constexpr int EA = WPB * ADROIT_WPB_SOLVE * (ADROIT_WPB_SMOOTH == WPB ? 1 : ADROIT_WPB_SMOOTH) * (ADROIT_WPB_COLLIDE % WPB == 0 ? ADROIT_WPB_COLLIDE / WPB : ADROIT_WPB_COLLIDE);        // sub-batch boundaries fall on block boundaries of every phase kernel
  int per = ((st.n + nsub - 1) / nsub + EA - 1) / EA * EA;
  if (per < 64 * WPB) { per = (64 * WPB + EA - 1) / EA * EA; }          // never less than 64 blocks per launch
  nsub = (st.n + per - 1) / per;
  if (nsub <= 1) {
    a.env0 = 0; a.env1 = st.n;
    if (host_in(s, hio, const_cast<float*>(action), 0, st.n, stream)) return 1;
    if (run_range(s, m, st, mode, a, action, mask, obs, reward, done, goal, stream)) return 1;
    return host_out(s, hio, obs, reward, done, goal, 0, st.n, stream);
  }
  CK(cudaEventRecord(s->fork_ev, stream));
  for (int c = 0; c < nsub; c++) {
    cudaStream_t ss = s->sub_stream[c];
    CK(cudaStreamWaitEvent(ss, s->fork_ev, 0));
    a.env0 = c * per;
    a.env1 = (c + 1) * per < st.n ? (c + 1) * per : st.n;
    if (host_in(s, hio, const_cast<float*>(action), a.env0, a.env1, ss)) return 1;
    if (run_range(s, m, st, mode, a, action, mask, obs, reward, done, goal, ss)) return 1;
    if (host_out(s, hio, obs, reward, done, goal, a.env0, a.env1, ss)) return 1;
    CK(cudaEventRecord(s->sub_done[c], ss));
    CK(cudaStreamWaitEvent(stream, s->sub_done[c], 0));
  }
  return 0;
}

int ensure_stage(adroit_sim* s, size_t n) {
  if (s->stage_len >= n) return 0;
  if (s->stage) cudaFree(s->stage);
  CK(cudaMalloc((void**)&s->stage, n * sizeof(double)));
  s->stage_len = n;
  return 0;
}

int run(adroit_sim* s, int mode, const float* action, const uint8_t* mask, float* obs, float* reward, uint8_t* done, uint8_t* goal,
        int debug, void* stream, const HostIO* hio = nullptr) {
  CK(cudaSetDevice(s->device));
  if (s->f64) return run_t<double>(s, s->md, s->sd, mode, action, mask, obs, reward, done, goal, debug, (cudaStream_t)stream, hio);
  return run_t<float>(s, s->mf, s->sf, mode, action, mask, obs, reward, done, goal, debug, (cudaStream_t)stream, hio);
}

// host fp64 [n][w]  ->  device typed [n][stride]
int upload_field(adroit_sim* s, const double* host, void* dev, int w, int stride, int offset, cudaStream_t st) {
  size_t cnt = (size_t)s->n * w;
  if (ensure_stage(s, cnt)) return 1;
  CK(cudaMemcpyAsync(s->stage, host, cnt * sizeof(double), cudaMemcpyHostToDevice, st));
  int thr = 256, blk = (int)((cnt + thr - 1) / thr);
  if (s->f64) pack_kernel<double><<<blk, thr, 0, st>>>((double*)dev + offset, s->stage, s->n, w, stride);
  else pack_kernel<float><<<blk, thr, 0, st>>>((float*)dev + offset, s->stage, s->n, w, stride);
  CK(cudaGetLastError());
  s->launches++;
  CK(cudaStreamSynchronize(st));
  return 0;
}
int download_field(adroit_sim* s, double* host, const void* dev, int w, int stride, int offset, cudaStream_t st) {
  size_t cnt = (size_t)s->n * w;
  if (ensure_stage(s, cnt)) return 1;
  int thr = 256, blk = (int)((cnt + thr - 1) / thr);
  if (s->f64) unpack_kernel<double><<<blk, thr, 0, st>>>(s->stage, (const double*)dev, s->n, w, stride, offset);
  else unpack_kernel<float><<<blk, thr, 0, st>>>(s->stage, (const float*)dev, s->n, w, stride, offset);
  CK(cudaGetLastError());
  s->launches++;
  CK(cudaMemcpyAsync(host, s->stage, cnt * sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace

#define CTXP (s->f64 ? (void*)s->sd.ctx : (void*)s->sf.ctx)

extern "C" {

const char* adroit_last_error(void) { return g_err; }

int adroit_create(const void* model_blob, size_t blob_len, int n_envs, int device_id, uint64_t seed, int64_t env_id_offset,
                  uint32_t flags, adroit_sim** out) {
  if (!model_blob || blob_len < 16 || memcmp(model_blob, "ADRB200\x01", 8) != 0) FAIL("bad model blob");
  if (n_envs <= 0) FAIL("n_envs must be positive");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) FAIL("no CUDA device: the CUDA path has no CPU fallback");
  if (device_id < 0 || device_id >= ndev) FAIL("device %d out of range (%d devices)", device_id, ndev);
  CK(cudaSetDevice(device_id));
  adroit_sim* s = new adroit_sim();
  s->n = n_envs; s->device = device_id; s->flags = flags; s->f64 = (flags & ADROIT_F64) != 0;
  Blob B{(const unsigned char*)model_blob, blob_len};
  try {
    B.validate();
    if (s->f64) fill_model(s->md, s->arena, B); else fill_model(s->mf, s->arena, B);
  } catch (const std::string& e) {
    adroit_destroy(s);
    FAIL("%s", e.c_str());
  }
  // every failure below releases the handle and what it already owns (adroit_destroy), then reports
#define CKD(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { adroit_destroy(s); FAIL("%s failed: %s", #call, cudaGetErrorString(e_)); } } while (0)
  CKD(cudaMalloc((void**)&s->arena.dev, s->arena.host.size()));
  s->allocs.push_back(s->arena.dev);
  CKD(cudaMemcpy(s->arena.dev, s->arena.host.data(), s->arena.host.size(), cudaMemcpyHostToDevice));
  if (s->f64 ? configure_kernels<double>() : configure_kernels<float>()) { adroit_destroy(s); return 1; }
  {
    const char* ep = getenv("ADROIT_PDL");
    s->pdl = ep && atoi(ep) != 0;
    const char* el = getenv("ADROIT_LIGHT_RESET");
    s->light_reset = !el || atoi(el) != 0;
    const char* e = getenv("ADROIT_SUBBATCHES");
    // default: sub-batches of >= 2048 envs, at most ADROIT_DEFAULT_SUBBATCHES of them (measured: profiles/r3_subbatches.txt)
    int want = e ? atoi(e) : (n_envs / 2048 < ADROIT_DEFAULT_SUBBATCHES ? n_envs / 2048 : ADROIT_DEFAULT_SUBBATCHES);
    s->nsub = want < 1 ? 1 : (want > 16 ? 16 : want);
    CKD(cudaEventCreateWithFlags(&s->fork_ev, cudaEventDisableTiming));
    for (int c = 0; c < s->nsub; c++) {
      cudaStream_t ss;
      cudaEvent_t ev;
      CKD(cudaStreamCreateWithFlags(&ss, cudaStreamNonBlocking));
      CKD(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      s->sub_stream.push_back(ss);
      s->sub_done.push_back(ev);
    }
  }
  int rc;
  if (s->f64) {
    patch_model(s->md, s->arena.dev);
    rc = alloc_state(s, s->sd);
    s->sd.seed = seed; s->sd.env_off = env_id_offset; s->sd.flags = flags;
    s->nv = s->md.nv; s->nu = s->md.nu; s->obs_dim = s->md.obs_dim; s->npar = s->md.npar; s->horizon = s->md.horizon;
    s->frame_skip = s->md.frame_skip; s->nbody = s->md.nbody; s->nsite = s->md.nsite; s->task = s->md.task;
  } else {
    patch_model(s->mf, s->arena.dev);
    rc = alloc_state(s, s->sf);
    s->sf.seed = seed; s->sf.env_off = env_id_offset; s->sf.flags = flags;
    s->nv = s->mf.nv; s->nu = s->mf.nu; s->obs_dim = s->mf.obs_dim; s->npar = s->mf.npar; s->horizon = s->mf.horizon;
    s->frame_skip = s->mf.frame_skip; s->nbody = s->mf.nbody; s->nsite = s->mf.nsite; s->task = s->mf.task;
  }
  if (rc) { adroit_destroy(s); return 1; }
  // model-default parameters, then one forward so that frames / observations are defined before the first reset
  std::vector<double> par((size_t)n_envs * s->npar);
  for (int e = 0; e < n_envs; e++)
    for (int i = 0; i < s->npar; i++) par[(size_t)e * s->npar + i] = s->f64 ? s->md.par0[i] : (double)s->mf.par0[i];
  if (upload_field(s, par.data(), CTXP, s->npar, CTX_T, O_PAR, 0)) { adroit_destroy(s); return 1; }
  if (cudaMalloc((void**)&s->d_out4, 4 * sizeof(double)) != cudaSuccess) { adroit_destroy(s); FAIL("cudaMalloc failed"); }
  s->allocs.push_back(s->d_out4);
  if (run(s, MODE_FORWARD, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr)) { adroit_destroy(s); return 1; }
  if (cudaDeviceSynchronize() != cudaSuccess) { adroit_destroy(s); FAIL("initial forward failed: %s", cudaGetErrorString(cudaGetLastError())); }
#undef CKD
  *out = s;
  return 0;
}

void adroit_destroy(adroit_sim* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  for (auto& v : s->prof_ev) for (cudaEvent_t e : v) cudaEventDestroy(e);
  for (cudaStream_t ss : s->sub_stream) cudaStreamDestroy(ss);
  for (cudaEvent_t e : s->sub_done) cudaEventDestroy(e);
  if (s->fork_ev) cudaEventDestroy(s->fork_ev);
  for (void* p : s->allocs) cudaFree(p);
  if (s->stage) cudaFree(s->stage);
  if (s->d_action) cudaFree(s->d_action);
  if (s->d_obs) cudaFree(s->d_obs);
  if (s->d_reward) cudaFree(s->d_reward);
  if (s->d_done) cudaFree(s->d_done);
  if (s->d_goal) cudaFree(s->d_goal);
  delete s;
}

int adroit_sizes(const adroit_sim* s, int* o) {
  o[0] = s->nv; o[1] = s->nu; o[2] = s->obs_dim; o[3] = s->npar; o[4] = s->horizon; o[5] = s->frame_skip; o[6] = s->nbody; o[7] = s->nsite;
  return 0;
}

int adroit_reset(adroit_sim* s, const uint8_t* mask_dev, float* obs_dev, void* stream) {
  return run(s, MODE_RESET, nullptr, mask_dev, obs_dev, nullptr, nullptr, nullptr, 0, stream);
}

int adroit_step(adroit_sim* s, const float* action_dev, float* obs_dev, float* reward_dev, uint8_t* done_dev, uint8_t* goal_dev, void* stream) {
  if (!action_dev) FAIL("action is NULL");
  return run(s, MODE_STEP, action_dev, nullptr, obs_dev, reward_dev, done_dev, goal_dev, 0, stream);
}

int adroit_step_host(adroit_sim* s, const float* action_host, float* obs_host, float* reward_host, uint8_t* done_host, uint8_t* goal_host,
                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  size_t n = s->n;
  if (!s->d_action) {
    CK(cudaMalloc((void**)&s->d_action, n * s->nu * sizeof(float)));
    CK(cudaMalloc((void**)&s->d_obs, n * s->obs_dim * sizeof(float)));
    CK(cudaMalloc((void**)&s->d_reward, n * sizeof(float)));
    CK(cudaMalloc((void**)&s->d_done, n));
    CK(cudaMalloc((void**)&s->d_goal, n));
  }
  const HostIO hio = {action_host, obs_host, reward_host, done_host, goal_host};   // copies are issued per sub-batch (run_t)
  if (run(s, MODE_STEP, s->d_action, nullptr, s->d_obs, s->d_reward, s->d_done, s->d_goal, 0, stream, &hio)) return 1;
  CK(cudaStreamSynchronize(st));
  return 0;
}

int adroit_get_obs(adroit_sim* s, float* obs_dev, void* stream) {
  CK(cudaSetDevice(s->device));
  int thr = 128, blk = (int)(((size_t)s->n * 32 + thr - 1) / thr);
  if (s->f64) obs_kernel<double><<<blk, thr, 0, (cudaStream_t)stream>>>(s->md, s->sd, obs_dev);
  else obs_kernel<float><<<blk, thr, 0, (cudaStream_t)stream>>>(s->mf, s->sf, obs_dev);
  CK(cudaGetLastError());
  s->launches++;
  return 0;
}

int adroit_get_state(adroit_sim* s, double* qpos, double* qvel, double* par, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  if (qpos && download_field(s, qpos, CTXP, s->nv, CTX_T, V_QPOS * NVP, st)) return 1;
  if (qvel && download_field(s, qvel, CTXP, s->nv, CTX_T, V_QVEL * NVP, st)) return 1;
  if (par && download_field(s, par, CTXP, s->npar, CTX_T, O_PAR, st)) return 1;
  return 0;
}

int adroit_set_state(adroit_sim* s, const double* qpos, const double* qvel, const double* par, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  if (qpos && upload_field(s, qpos, CTXP, s->nv, CTX_T, V_QPOS * NVP, st)) return 1;
  if (qvel && upload_field(s, qvel, CTXP, s->nv, CTX_T, V_QVEL * NVP, st)) return 1;
  if (par && upload_field(s, par, CTXP, s->npar, CTX_T, O_PAR, st)) return 1;
  if (run(s, MODE_FORWARD, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 1, stream)) return 1;
  CK(cudaStreamSynchronize(st));
  return 0;
}

int adroit_set_elapsed_steps(adroit_sim* s, const int32_t* t_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  if (!t_host) FAIL("t is NULL");
  CK(cudaMemcpyAsync(s->f64 ? s->sd.t : s->sf.t, t_host, sizeof(int) * (size_t)s->n, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

int adroit_get_warmstart(adroit_sim* s, double* w, void* stream) {
  CK(cudaSetDevice(s->device));
  return download_field(s, w, CTXP, s->nv, CTX_T, V_WARM * NVP, (cudaStream_t)stream);
}
int adroit_set_warmstart(adroit_sim* s, const double* w, void* stream) {
  CK(cudaSetDevice(s->device));
  return upload_field(s, w, CTXP, s->nv, CTX_T, V_WARM * NVP, (cudaStream_t)stream);
}

int adroit_forward(adroit_sim* s, void* stream) {
  return run(s, MODE_FORWARD, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, stream);
}

int adroit_get_frames(adroit_sim* s, double* bx, double* bq, double* sx, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  // a pure read of the context block: the frames of the LAST forward (one substep behind qpos after a step, exactly
  // what mujoco-py's data.body_xpos / site_xpos hold when the reference's get_env_state reads them, hammer_v0.py:134-142)
  if (bx && download_field(s, bx, CTXP, 3 * s->nbody, CTX_T, O_XPOS, st)) return 1;
  if (bq && download_field(s, bq, CTXP, 4 * s->nbody, CTX_T, O_XQUAT, st)) return 1;
  if (sx && download_field(s, sx, CTXP, 3 * s->nsite, CTX_T, SM_T + G_SPOS, st)) return 1;
  return 0;
}

int adroit_get_solver(adroit_sim* s, double* qacc, int32_t* ncon, int32_t* nefc, double* touch, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  // a pure read, like adroit_get_frames: qacc, the contact / row counts and the touch reading of the LAST forward
  if (qacc && download_field(s, qacc, CTXP, s->nv, CTX_T, V_QACC * NVP, st)) return 1;
  if (touch && download_field(s, touch, CTXP, 1, CTX_T, O_TF + 19, st)) return 1;
  if (ncon || nefc) {
    std::vector<int> gi((size_t)s->n * CTX_I);
    CK(cudaMemcpyAsync(gi.data(), s->f64 ? s->sd.ctxi : s->sf.ctxi, gi.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int e = 0; e < s->n; e++) {
      const int* g = &gi[(size_t)e * CTX_I + SM_I];
      if (ncon) ncon[e] = g[GI_NCON];
      if (nefc) {
        int rows = g[GI_NSIMPLE];
        for (int ci = 0; ci < g[GI_NCON]; ci++) { int dim = g[GI_CON + 4 * ci + 2]; rows += dim == 1 ? 1 : 2 * (dim - 1); }
        nefc[e] = rows;
      }
    }
  }
  return 0;
}

int adroit_reduce_episode(adroit_sim* s, int flush_running, double* out4_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  CK(cudaMemsetAsync(s->d_out4, 0, 4 * sizeof(double), st));
  int thr = s->task == T_PEN ? 20 : 25;
  int blocks = (s->n + 255) / 256;
  if (blocks > 592) blocks = 592;
  if (s->f64) reduce_kernel<<<blocks, 256, 0, st>>>(s->sd.acc, s->sd.t, s->sd.ep_goal, nullptr, s->sd.ep_return, s->n, thr, flush_running, s->d_out4);
  else reduce_kernel<<<blocks, 256, 0, st>>>(s->sf.acc, s->sf.t, s->sf.ep_goal, s->sf.ep_return, nullptr, s->n, thr, flush_running, s->d_out4);
  CK(cudaGetLastError());
  s->launches++;
  CK(cudaMemcpyAsync(out4_host, s->d_out4, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return 0;
}

// Device address of g_noalloc_partials on `device` (which must be current).  In host code the array's name is only a stand-in for the
// symbol, not a device address: it is resolved once per device with cudaGetSymbolAddress (no allocation) and cached.
static int noalloc_scratch(int device, double** out) {
  constexpr int MAXDEV = 64;
  static double* cache[MAXDEV];          // written once per device with the same value by whichever thread gets there first
  if (device < 0 || device >= MAXDEV) FAIL("device id %d outside the no-alloc scratch table", device);
  double* p = __atomic_load_n(&cache[device], __ATOMIC_ACQUIRE);
  if (!p) {
    CK(cudaGetSymbolAddress((void**)&p, g_noalloc_partials));
    if (!p) FAIL("cudaGetSymbolAddress returned a null address for the no-alloc scratch");
    __atomic_store_n(&cache[device], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return 0;
}

int adroit_reduce_episode_noalloc(adroit_sim* s, int flush_running, double* out4_dev, void* stream) {
  if (!s) FAIL("sim is NULL");
  if (!out4_dev) FAIL("out4_dev is NULL");
  if (s->n > NOALLOC_MAX_ENVS)
    FAIL("adroit_reduce_episode_noalloc: %d envs exceed the static scratch (ADROIT_NOALLOC_MAX_ENVS = %d); use adroit_reduce_episode", s->n,
         NOALLOC_MAX_ENVS);
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  double* partials = nullptr;
  if (noalloc_scratch(s->device, &partials)) return 1;
  int thr = s->task == T_PEN ? 20 : 25;
  int blocks = (s->n + NOALLOC_BLOCK - 1) / NOALLOC_BLOCK;
  if (s->f64)
    reduce_partial_kernel<<<blocks, NOALLOC_BLOCK, 0, st>>>(s->sd.acc, s->sd.t, s->sd.ep_goal, nullptr, s->sd.ep_return, s->n, thr, flush_running,
                                                            partials);
  else
    reduce_partial_kernel<<<blocks, NOALLOC_BLOCK, 0, st>>>(s->sf.acc, s->sf.t, s->sf.ep_goal, s->sf.ep_return, nullptr, s->n, thr, flush_running,
                                                            partials);
  CK(cudaGetLastError());
  reduce_final_kernel<<<1, NOALLOC_BLOCK, 0, st>>>(partials, blocks, out4_dev);
  CK(cudaGetLastError());
  s->launches += 2;
  return 0;
}

int adroit_render(adroit_sim* s, uint8_t* rgb_dev, void* stream) {
  if (!rgb_dev) FAIL("rgb is NULL");
  CK(cudaSetDevice(s->device));
  if (s->f64) render_kernel<double><<<s->n, RENDER_THREADS, 0, (cudaStream_t)stream>>>(s->md, s->sd, rgb_dev);
  else render_kernel<float><<<s->n, RENDER_THREADS, 0, (cudaStream_t)stream>>>(s->mf, s->sf, rgb_dev);
  CK(cudaGetLastError());
  s->launches++;
  return 0;
}

int64_t adroit_launch_count(const adroit_sim* s) { return s->launches; }

int adroit_profile(adroit_sim* s, int enable) {
  s->profile = enable != 0;
  return 0;
}
int adroit_phase_times(adroit_sim* s, double* ms4, int64_t* n4, void* stream) {
  CK(cudaSetDevice(s->device));
  CK(cudaStreamSynchronize((cudaStream_t)stream));
  for (int p = 0; p < 5; p++) {
    double acc = 0;
    auto& v = s->prof_ev[p];
    for (size_t i = 0; i + 1 < v.size(); i += 2) {
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, v[i], v[i + 1]));
      acc += ms;
      cudaEventDestroy(v[i]);
      cudaEventDestroy(v[i + 1]);
    }
    if (ms4 && p < 4) ms4[p] = acc;
    if (n4 && p < 4) n4[p] = (int64_t)(v.size() / 2);
    v.clear();
  }
  return 0;
}

int adroit_overflow(adroit_sim* s, int64_t* out2_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaSetDevice(s->device));
  unsigned long long v[2];
  CK(cudaMemcpyAsync(v, s->f64 ? s->sd.overflow : s->sf.overflow, sizeof v, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  out2_host[0] = (int64_t)v[0];
  out2_host[1] = (int64_t)v[1];
  return 0;
}

// debug: raw per-env scratch block (fp64 host buffer [n][count]) starting at T-offset `offset`; layout in csrc/model.cuh
int adroit_debug_scratch(adroit_sim* s, int offset, int count, double* out_host, void* stream) {
  CK(cudaSetDevice(s->device));
  return download_field(s, out_host, CTXP, count, CTX_T, SM_T + offset, (cudaStream_t)stream);
}
int adroit_debug_iscratch(adroit_sim* s, int32_t* out_host, void* stream) {
  CK(cudaSetDevice(s->device));
  CK(cudaMemcpy2DAsync(out_host, GI_TOTAL * sizeof(int), (s->f64 ? s->sd.ctxi : s->sf.ctxi) + SM_I, CTX_I * sizeof(int), GI_TOTAL * sizeof(int),
                       (size_t)s->n, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CK(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}
int adroit_debug_layout(int* out, int max_n) {
  const int v[] = {NVP, NBP, NGP, NSP, MAXCON, CROWS, NTRI, NPARP, NTF, G_TOTAL, GI_TOTAL, G_GPOS, G_GMAT, G_SPOS, G_CON, G_JR, G_MINV,
                   G_DBG_M, G_DBG_BIAS, G_DBG_QFS, G_DBG_QAS, G_DBG_QACC, G_DBG_QFC, G_DBG_XPOS, G_DBG_XQUAT, G_DBG_EF, GI_CON,
                   // dynamic shared memory per block of the fp32 phase kernels (occupancy bookkeeping, tools/occupancy.py)
                   (int)phase_smem<float, MapSmooth>(), (int)phase_smem<float, MapCollide>(), (int)phase_smem<float, MapSolveK<float, MAXCON_S>>(),
                   (int)phase_smem<float, MapSolveK<float, MAXCON>>(), (int)phase_smem<float, MapFinish>(), (int)phase_smem<float, MapReset>(), CTX_T, CTX_I};
  int n = (int)(sizeof v / sizeof v[0]);
  for (int i = 0; i < n && i < max_n; i++) out[i] = v[i];
  return n;
}

}  // extern "C"
