// dynamics.cuh — warp-cooperative smooth dynamics for one environment:
// level-synchronous forward kinematics, composite-rigid-body mass matrix, recursive Newton-Euler bias forces,
// affine actuation, packed Cholesky factorisation / solves and M^-1.
//
// Reference semantics: MuJoCo 2.1 mj_kinematics / mj_crb / mj_rne / mj_fwdActuation / mj_fwdAcceleration as reached
// from the reference through MujocoEnv.do_simulation (hammer_v0.py:60, door_v0.py:61, pen_v0.py:73,
// relocate_v0.py:52) -- third-party code, restated from the published algorithm.
//
// fp32 design note: every spatial quantity of body b is expressed in world axes about the body's OWN origin
// xpos[b]; parent<->child transfers are pure translations by xpos[child]-xpos[parent].  Lever arms therefore stay
// at link scale and the tiny finger inertias (1e-5 kg m^2) are never buried under parallel-axis terms of a far
// away reference point, which a single global origin would do in fp32.
#pragma once
#include "model.cuh"

namespace adroit {

// Per-env context block in global memory: the "S part" (what the single-kernel design kept in shared memory; offsets
// O_*, V_* of model.cuh) followed by the "G part" (offsets G_*).  Every phase kernel names, in a MAP, the regions it
// wants staged in shared memory; everything else is addressed in the context block directly (L1/L2).  The accessors
// resolve at compile time, so a device function works unchanged whichever side a region lives on.
constexpr int CTX_T = SM_T + G_TOTAL;     // T words per env
constexpr int CTX_I = SM_I + GI_TOTAL;    // int words per env
static_assert(CTX_T % 4 == 0 && CTX_I % 4 == 0, "context blocks must stay 16-byte aligned (bulk copies)");
enum { R_IN = 1, R_OUT = 2, R_ZERO = 4 };
struct Reg { int off, len, flags; };
constexpr int pad4(int x) { return (x + 3) / 4 * 4; }
template <class MAP> constexpr int map_soff(int O) {
  int acc = 0;
  for (int i = 0; i < MAP::NT; i++) {
    if (O >= MAP::RT[i].off && O < MAP::RT[i].off + MAP::RT[i].len) return acc + (O - MAP::RT[i].off);
    acc += pad4(MAP::RT[i].len);
  }
  return -1;
}
template <class MAP> constexpr int map_ioff(int O) {
  int acc = 0;
  for (int i = 0; i < MAP::NI; i++) {
    if (O >= MAP::RI[i].off && O < MAP::RI[i].off + MAP::RI[i].len) return acc + (O - MAP::RI[i].off);
    acc += pad4(MAP::RI[i].len);
  }
  return -1;
}
// length of the map's region that starts at context offset O (-1: the region is not staged)
template <class MAP> constexpr int map_rlen(int O) {
  for (int i = 0; i < MAP::NT; i++) if (MAP::RT[i].off == O) return MAP::RT[i].len;
  return -1;
}
template <class MAP> constexpr int map_st() { int a = 0; for (int i = 0; i < MAP::NT; i++) a += pad4(MAP::RT[i].len); return a; }
template <class MAP> constexpr int map_si() { int a = 0; for (int i = 0; i < MAP::NI; i++) a += pad4(MAP::RI[i].len); return a; }
template <typename T, class MAP> constexpr size_t map_bytes() { return map_st<MAP>() * sizeof(T) + map_si<MAP>() * sizeof(int); }

template <typename T_, class MAP_> struct Ctx {
  using real = T_;
  using MAP = MAP_;
  typedef T_ T;
  const DModel<T>& m;
  T* sm;     // this warp's shared-memory slice (T regions of MAP)
  int* smi;  // this warp's shared-memory slice (int regions of MAP)
  T* gb;     // this env's context block (T)
  int* gib;  // this env's context block (int): [SM_I | GI_TOTAL]
  int* gi;   // = gib + SM_I
  int lane;
  int ncon, nsimple, nchanged, nflipped;
  __device__ Ctx(const DModel<T>& mm, T* sm_, int* smi_, T* gb_, int* gib_, int lane_)
      : m(mm), sm(sm_), smi(smi_), gb(gb_), gib(gib_), gi(gib_ + SM_I), lane(lane_), ncon(0), nsimple(0), nchanged(1), nflipped(0) {}
  // contact capacity of this kernel's shared-memory tables (the size class of the solve kernel), MAXCON if they are not staged
  static constexpr int MCAP = map_rlen<MAP>(O_AC) > 0 ? map_rlen<MAP>(O_AC) / 4 : MAXCON;
  template <int O> static constexpr bool in_smem() { return map_soff<MAP>(O) >= 0; }
  template <int O> DI T* S() const {
    constexpr int so = map_soff<MAP>(O);
    if constexpr (so >= 0) return sm + so;
    else return gb + O;
  }
  template <int K> DI T* V() const { return S<K * NVP>(); }
  template <int O> DI T* G() const { return S<SM_T + O>(); }
  template <int O> DI int* I() const {
    constexpr int so = map_ioff<MAP>(O);
    if constexpr (so >= 0) return smi + so;
    else return gib + O;
  }
};

// Address-space hints for the few `SUB_FN` (noinline) helpers: they receive the Ctx BY VALUE (a reference would make the
// caller's Ctx escape to local memory, after which every access in the caller is a generic load with 64-bit address
// arithmetic -- 252 generic LD/ST sites and 90 LDL/STL in the round-1 solve kernel) and state what the pointers are.
template <class CX> DI void assume_spaces(const CX& c) {
  __builtin_assume(__isShared(c.sm));
  __builtin_assume(__isShared(c.smi));
  __builtin_assume(__isGlobal(c.gb));
  __builtin_assume(__isGlobal(c.gib));
  __builtin_assume(__isGlobal(c.gi));
}

// stage the MAP's regions between the context block and shared memory.
//
// Blackwell-native form (ADROIT_BULK_STAGE, default): every region is ONE 1-D bulk copy issued by lane 0 --
// `cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes` on the way in (the warp's own mbarrier counts the
// bytes; the lanes zero the R_ZERO scratch regions meanwhile and then wait on the barrier's phase), and
// `cp.async.bulk.global.shared::cta.bulk_group` on the way out (fire and forget: the warp only waits until the copy engine
// has READ its shared memory, not until the data is in HBM).  Region offsets and lengths are multiples of 16 bytes by
// construction (static_asserts below), the context blocks are 16-byte aligned (CTX_T, CTX_I multiples of 4 words).
// Ampere-style fallback (ADROIT_BULK_STAGE=0, kept for A/B): 16-byte cp.async chunks, lanes over chunks.
#ifndef ADROIT_BULK_STAGE
#define ADROIT_BULK_STAGE 1
#endif
DI unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
DI void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // the copy engine (async proxy) must see the initialised barrier
}
DI void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
DI void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MBAR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MBAR_DONE;\n"
      "bra MBAR_WAIT;\n"
      "MBAR_DONE:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
DI void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)), "l"(gmem_src),
               "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
DI void bulk_s2g(void* gmem_dst, const void* smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
DI void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem_src) : "memory");
}
DI void cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory"); }
template <class MAP, typename T> constexpr unsigned map_in_bytes() {
  unsigned b = 0;
  for (int i = 0; i < MAP::NT; i++) if (MAP::RT[i].flags & R_IN) b += MAP::RT[i].len * sizeof(T);
  for (int i = 0; i < MAP::NI; i++) if (MAP::RI[i].flags & R_IN) b += MAP::RI[i].len * sizeof(int);
  return b;
}
// Runs of regions that are adjacent in the map (hence in shared memory: lengths are multiples of 4 words) AND in the
// context block, with the same direction flag, travel as ONE bulk copy: run_head = the run starts at region i,
// run_len = words from region i to the end of its run.
constexpr bool run_joins(const Reg* R, int n, int i, int flag) {   // does region i continue the run of region i-1?
  return i > 0 && i < n && (R[i].flags & flag) && (R[i - 1].flags & flag) && R[i - 1].off + R[i - 1].len == R[i].off;
}
constexpr int run_len(const Reg* R, int n, int i, int flag) {
  int len = R[i].len;
  for (int k = i + 1; k < n && run_joins(R, n, k, flag); k++) len += R[k].len;
  return len;
}
// lane 0: one bulk copy per run of R_IN regions
template <class CX, int I = 0> DI void stage_in_bulk(CX& c, unsigned long long* bar) {
  using MAP = typename CX::MAP;
  using T = typename CX::real;
  constexpr int E = 16 / sizeof(T);
  if constexpr (I < MAP::NT) {
    constexpr Reg r = MAP::RT[I];
    constexpr int so = map_soff<MAP>(r.off);
    static_assert(r.off % E == 0 && r.len % E == 0, "regions are staged in 16-byte chunks");
    if constexpr ((r.flags & R_IN) != 0 && !run_joins(MAP::RT, MAP::NT, I, R_IN))
      bulk_g2s(c.sm + so, c.gb + r.off, run_len(MAP::RT, MAP::NT, I, R_IN) * sizeof(T), bar);
    stage_in_bulk<CX, I + 1>(c, bar);
  } else if constexpr (I < MAP::NT + MAP::NI) {
    constexpr int K = I - MAP::NT;
    constexpr Reg r = MAP::RI[K];
    constexpr int so = map_ioff<MAP>(r.off);
    static_assert(r.off % 4 == 0 && r.len % 4 == 0, "regions are staged in 16-byte chunks");
    if constexpr ((r.flags & R_IN) != 0 && !run_joins(MAP::RI, MAP::NI, K, R_IN))
      bulk_g2s(c.smi + so, c.gib + r.off, run_len(MAP::RI, MAP::NI, K, R_IN) * sizeof(int), bar);
    stage_in_bulk<CX, I + 1>(c, bar);
  }
}
// all lanes: zero the R_ZERO regions (BULK: only those; otherwise also the cp.async chunks of the R_IN regions)
template <class CX, bool BULK, int I = 0> DI void stage_in_issue(CX& c) {
  using MAP = typename CX::MAP;
  using T = typename CX::real;
  constexpr int E = 16 / sizeof(T);
  if constexpr (I < MAP::NT) {
    constexpr Reg r = MAP::RT[I];
    constexpr int so = map_soff<MAP>(r.off);
    static_assert(r.off % E == 0 && r.len % E == 0, "regions are staged in 16-byte chunks");
    if constexpr ((r.flags & R_IN) != 0) { if constexpr (!BULK) for (int j = c.lane * E; j < r.len; j += 32 * E) cp_async16(c.sm + so + j, c.gb + r.off + j); }
    else if constexpr ((r.flags & R_ZERO) != 0) { for (int j = c.lane; j < r.len; j += 32) c.sm[so + j] = 0; }
    stage_in_issue<CX, BULK, I + 1>(c);
  } else if constexpr (I < MAP::NT + MAP::NI) {
    constexpr Reg r = MAP::RI[I - MAP::NT];
    constexpr int so = map_ioff<MAP>(r.off);
    static_assert(r.off % 4 == 0 && r.len % 4 == 0, "regions are staged in 16-byte chunks");
    if constexpr ((r.flags & R_IN) != 0) { if constexpr (!BULK) for (int j = c.lane * 4; j < r.len; j += 128) cp_async16(c.smi + so + j, c.gib + r.off + j); }
    stage_in_issue<CX, BULK, I + 1>(c);
  }
}
// `bar`: this warp's mbarrier (8 bytes of shared memory, used for one phase)
template <class CX> DI void stage_in(CX& c, unsigned long long* bar) {
#if ADROIT_BULK_STAGE
  using MAP = typename CX::MAP;
  using T = typename CX::real;
  constexpr unsigned bytes = map_in_bytes<MAP, T>();
  if constexpr (bytes > 0) {
    if (c.lane == 0) {
      mbar_init(bar, 1);
      mbar_expect_tx(bar, bytes);
      stage_in_bulk(c, bar);
    }
  }
  stage_in_issue<CX, true>(c);
  __syncwarp();
  if constexpr (bytes > 0) mbar_wait(bar, 0);
#else
  stage_in_issue<CX, false>(c);
  cp_async_wait_all();
#endif
}
template <class CX, int I = 0> DI void stage_out_regions(CX& c) {
  using MAP = typename CX::MAP;
  using T = typename CX::real;
  [[maybe_unused]] constexpr int E = 16 / sizeof(T);
  if constexpr (I < MAP::NT) {
    constexpr Reg r = MAP::RT[I];
    constexpr int so = map_soff<MAP>(r.off);
    if constexpr ((r.flags & R_OUT) != 0) {
#if ADROIT_BULK_STAGE
      if constexpr (!run_joins(MAP::RT, MAP::NT, I, R_OUT)) bulk_s2g(c.gb + r.off, c.sm + so, run_len(MAP::RT, MAP::NT, I, R_OUT) * sizeof(T));
#else
      for (int j = c.lane * E; j < r.len; j += 32 * E) *reinterpret_cast<int4*>(c.gb + r.off + j) = *reinterpret_cast<const int4*>(c.sm + so + j);
#endif
    }
    stage_out_regions<CX, I + 1>(c);
  } else if constexpr (I < MAP::NT + MAP::NI) {
    constexpr Reg r = MAP::RI[I - MAP::NT];
    constexpr int so = map_ioff<MAP>(r.off);
    if constexpr ((r.flags & R_OUT) != 0) {
#if ADROIT_BULK_STAGE
      if constexpr (!run_joins(MAP::RI, MAP::NI, I - MAP::NT, R_OUT))
        bulk_s2g(c.gib + r.off, c.smi + so, run_len(MAP::RI, MAP::NI, I - MAP::NT, R_OUT) * sizeof(int));
#else
      for (int j = c.lane * 4; j < r.len; j += 128) *reinterpret_cast<int4*>(c.gib + r.off + j) = *reinterpret_cast<const int4*>(c.smi + so + j);
#endif
    }
    stage_out_regions<CX, I + 1>(c);
  }
}
// one staged region written back ahead of the epilogue; returns once the copy engine has read the shared memory, so the
// region may be overwritten afterwards
template <int O, int LEN, class CX> DI void store_region_early(CX& c) {
  using T = typename CX::real;
  static_assert(CX::template in_smem<O>(), "store_region_early is for staged regions");
#if ADROIT_BULK_STAGE
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (c.lane == 0) {
    bulk_s2g(c.gb + O, c.template S<O>(), LEN * sizeof(T));
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  __syncwarp();
#else
  constexpr int E = 16 / sizeof(T);
  const T* sp = c.template S<O>();
  for (int j = c.lane * E; j < LEN; j += 32 * E) *reinterpret_cast<int4*>(c.gb + O + j) = *reinterpret_cast<const int4*>(sp + j);
  __syncwarp();
#endif
}
// write the R_OUT regions back (call with the warp converged, after the last shared-memory write)
template <class CX> DI void stage_out(CX& c) {
#if ADROIT_BULK_STAGE
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the lanes' generic-proxy writes -> visible to the copy engine
  __syncwarp();
  if (c.lane == 0) {
    stage_out_regions(c);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory may be released once it has been read
  }
  __syncwarp();
#else
  stage_out_regions(c);
#endif
}

template <typename T> DI void quat_mul(T* r, const T* a, const T* b) {
  T w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  T x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  T y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  T z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <typename T> DI void quat_norm(T* q) {
  T n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < tiny<T>()) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  T inv = T(1) / n;
  q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv;
}
template <typename T> DI void quat2mat(T* R, const T* q) {
  T w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = w * w + x * x - y * y - z * z; R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = w * w - x * x + y * y - z * z; R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = w * w - x * x - y * y + z * z;
}
template <typename T> DI void rot(T* r, const T* R, const T* v) {
  T x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2], y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2], z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> DI void cross3(T* r, const T* a, const T* b) {
  T x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> DI T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> DI void sincos_t(T x, T* s, T* c);
template <> DI void sincos_t<float>(float x, float* s, float* c) { sincosf(x, s, c); }
template <> DI void sincos_t<double>(double x, double* s, double* c) { sincos(x, s, c); }

// ---------------------------------------------------------------- kinematics (level-synchronous over the body tree)
template <class CX> PHASE_FN void kinematics(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  T *xpos = c.template S<O_XPOS>(), *xquat = c.template S<O_XQUAT>(), *xmat = c.template S<O_XMAT>(), *xanc = c.template S<O_XANC>(), *xaxis = c.template S<O_XAXIS>();
  const T *par = c.template S<O_PAR>(), *qpos = c.template V<V_QPOS>();
  if (c.lane == 0) {
    xpos[0] = xpos[1] = xpos[2] = 0;
    xquat[0] = 1; xquat[1] = xquat[2] = xquat[3] = 0;
    for (int i = 0; i < 9; i++) xmat[i] = (i % 4 == 0) ? T(1) : T(0);
  }
  __syncwarp();
  for (int L = 1; L < m.nlevel; L++) {
    int beg = m.level_start[L], end = m.level_start[L + 1];
    for (int idx = beg + c.lane; idx < end; idx += 32) {
      int b = m.level_body[idx], p = m.body_parent[b];
      T bp[3], bq[4], pos[3], quat[4], R[9], t[3];
      int po = m.body_par[2 * b], qo = m.body_par[2 * b + 1];
      for (int i = 0; i < 3; i++) bp[i] = po >= 0 ? par[po + i] : m.body_pos[3 * b + i];
      for (int i = 0; i < 4; i++) bq[i] = qo >= 0 ? par[qo + i] : m.body_quat[4 * b + i];
      quat_norm(bq);
      rot(t, xmat + 9 * p, bp);
      for (int i = 0; i < 3; i++) pos[i] = xpos[3 * p + i] + t[i];
      quat_mul(quat, xquat + 4 * p, bq);
      int ja = m.body_jntadr[b], jn = m.body_jntnum[b];
      for (int k = 0; k < jn; k++) {
        int j = ja + k;
        quat2mat(R, quat);
        rot(t, R, m.jnt_pos + 3 * j);
        for (int i = 0; i < 3; i++) xanc[3 * j + i] = pos[i] + t[i];
        T ax[3];
        rot(ax, R, m.jnt_axis + 3 * j);
        for (int i = 0; i < 3; i++) xaxis[3 * j + i] = ax[i];
        if (m.jnt_type[j] == J_SLIDE) {
          for (int i = 0; i < 3; i++) pos[i] += ax[i] * qpos[j];
        } else {
          T sn, cs;
          sincos_t(T(0.5) * qpos[j], &sn, &cs);
          T qj[4] = {cs, sn * m.jnt_axis[3 * j], sn * m.jnt_axis[3 * j + 1], sn * m.jnt_axis[3 * j + 2]};
          quat_mul(quat, quat, qj);
          quat2mat(R, quat);
          rot(t, R, m.jnt_pos + 3 * j);
          for (int i = 0; i < 3; i++) pos[i] = xanc[3 * j + i] - t[i];
        }
      }
      quat_norm(quat);
      quat2mat(R, quat);
      for (int i = 0; i < 3; i++) xpos[3 * b + i] = pos[i];
      for (int i = 0; i < 4; i++) xquat[4 * b + i] = quat[i];
      for (int i = 0; i < 9; i++) xmat[9 * b + i] = R[i];
    }
    __syncwarp();
  }
}

// geom / site world frames (global scratch) and single-body inertias about the body origin
template <class CX, bool INERTIA = true> PHASE_FN void frames_and_inertia(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const T *xpos = c.template S<O_XPOS>(), *xmat = c.template S<O_XMAT>(), *par = c.template S<O_PAR>();
  for (int g = c.lane; g < m.ngeom; g += 32) {
    int b = m.geom_body[g];
    T t[3];
    T gp[3] = {m.geom_pos[3 * g], m.geom_pos[3 * g + 1], m.geom_pos[3 * g + 2]};
    if (m.task_int[8] == VAR_POS) {   // hammer "pos" variation (hammer_v0.py:119-122): head / neck slide along the handle
      if (g == m.task_int[10]) gp[0] = par[3];
      if (g == m.task_int[11]) gp[0] = par[4];
    }
    rot(t, xmat + 9 * b, gp);
    for (int i = 0; i < 3; i++) c.template G<G_GPOS>()[3 * g + i] = xpos[3 * b + i] + t[i];
    const T *A = xmat + 9 * b, *B = m.geom_mat + 9 * g;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) c.template G<G_GMAT>()[9 * g + 3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  }
  for (int sidx = c.lane; sidx < m.nsite; sidx += 32) {
    int b = m.site_body[sidx], so = m.site_par[sidx];
    T sp[3], t[3];
    for (int i = 0; i < 3; i++) sp[i] = so >= 0 ? par[so + i] : m.site_pos[3 * sidx + i];
    rot(t, xmat + 9 * b, sp);
    for (int i = 0; i < 3; i++) c.template G<G_SPOS>()[3 * sidx + i] = xpos[3 * b + i] + t[i];
  }
  // inertia {m, h[3] = m*c, I[6] = xx,yy,zz,xy,xz,yz} about xpos[b], world axes
  T* crb = c.template S<O_CRB>();
  if constexpr (INERTIA) for (int b = c.lane; b < m.nbody; b += 32) {
    T mass = (m.task_int[8] == VAR_MASS && b == m.task_int[9]) ? par[3] : m.body_mass[b], Rw[9], cv[3];   // hammer_v0.py:115-118
    const T *A = xmat + 9 * b, *B = m.body_imat + 9 * b, *dg = m.body_inertia + 3 * b;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) Rw[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    rot(cv, A, m.body_ipos + 3 * b);
    T cc = dot3(cv, cv);
    T I[6];
    const int ii[6] = {0, 1, 2, 0, 0, 1}, jj[6] = {0, 1, 2, 1, 2, 2};
    for (int e = 0; e < 6; e++) {
      int i = ii[e], j = jj[e];
      T v = Rw[3 * i] * dg[0] * Rw[3 * j] + Rw[3 * i + 1] * dg[1] * Rw[3 * j + 1] + Rw[3 * i + 2] * dg[2] * Rw[3 * j + 2];
      v += mass * ((i == j ? cc : T(0)) - cv[i] * cv[j]);
      I[e] = v;
    }
    T out[10] = {mass, mass * cv[0], mass * cv[1], mass * cv[2], I[0], I[1], I[2], I[3], I[4], I[5]};
    for (int e = 0; e < 10; e++) { crb[10 * b + e] = out[e]; c.template G<G_CIN>()[10 * b + e] = out[e]; }
  }
  __syncwarp();
}

// (n, f) = I * (w, v) for an inertia stored as {m, h, Ixx, Iyy, Izz, Ixy, Ixz, Iyz}
template <typename T> DI void inert_mul(T* out, const T* I, const T* mv) {
  const T* h = I + 1;
  T t[3];
  out[0] = I[4] * mv[0] + I[7] * mv[1] + I[8] * mv[2];
  out[1] = I[7] * mv[0] + I[5] * mv[1] + I[9] * mv[2];
  out[2] = I[8] * mv[0] + I[9] * mv[1] + I[6] * mv[2];
  cross3(t, h, mv + 3);
  out[0] += t[0]; out[1] += t[1]; out[2] += t[2];
  cross3(t, h, mv);
  out[3] = I[0] * mv[3] - t[0]; out[4] = I[0] * mv[4] - t[1]; out[5] = I[0] * mv[5] - t[2];
}
// motion vector of dof j about origin O (world axes)
template <typename T> DI void dof_motion(const DModel<T>& m, const T* xanc, const T* xaxis, int j, const T* O, T* S) {
  const T* a = xaxis + 3 * j;
  if (m.jnt_type[j] == J_SLIDE) {
    S[0] = S[1] = S[2] = 0;
    S[3] = a[0]; S[4] = a[1]; S[5] = a[2];
  } else {
    T r[3] = {xanc[3 * j] - O[0], xanc[3 * j + 1] - O[1], xanc[3 * j + 2] - O[2]};
    S[0] = a[0]; S[1] = a[1]; S[2] = a[2];
    cross3(S + 3, r, a);
  }
}

// composite inertias (children folded into parents, deepest level first) and the joint-space inertia matrix
template <class CX> PHASE_FN void mass_matrix(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  T *crb = c.template S<O_CRB>(), *M = c.template S<O_M>();
  const T *xpos = c.template S<O_XPOS>(), *xanc = c.template S<O_XANC>(), *xaxis = c.template S<O_XAXIS>();
  for (int L = m.nlevel - 1; L >= 1; L--) {
    int beg = m.level_start[L - 1], end = m.level_start[L];
    for (int idx = beg + c.lane; idx < end; idx += 32) {
      int p = m.level_body[idx];
      if (p == 0) continue;  // the world body's composite is never used
      T acc[10];
      for (int e = 0; e < 10; e++) acc[e] = crb[10 * p + e];
      for (int k = m.child_start[p]; k < m.child_start[p + 1]; k++) {
        int ch = m.child_list[k];
        const T* I = crb + 10 * ch;
        T r[3] = {xpos[3 * ch] - xpos[3 * p], xpos[3 * ch + 1] - xpos[3 * p + 1], xpos[3 * ch + 2] - xpos[3 * p + 2]};
        T mass = I[0], hr = I[1] * r[0] + I[2] * r[1] + I[3] * r[2], rr = dot3(r, r);
        T diag = 2 * hr + mass * rr;
        acc[0] += mass;
        acc[1] += I[1] + mass * r[0]; acc[2] += I[2] + mass * r[1]; acc[3] += I[3] + mass * r[2];
        acc[4] += I[4] + diag - (2 * I[1] * r[0] + mass * r[0] * r[0]);
        acc[5] += I[5] + diag - (2 * I[2] * r[1] + mass * r[1] * r[1]);
        acc[6] += I[6] + diag - (2 * I[3] * r[2] + mass * r[2] * r[2]);
        acc[7] += I[7] - (I[1] * r[1] + r[0] * I[2] + mass * r[0] * r[1]);
        acc[8] += I[8] - (I[1] * r[2] + r[0] * I[3] + mass * r[0] * r[2]);
        acc[9] += I[9] - (I[2] * r[2] + r[1] * I[3] + mass * r[1] * r[2]);
      }
      for (int e = 0; e < 10; e++) crb[10 * p + e] = acc[e];
    }
    __syncwarp();
  }
  for (int p = c.lane; p < NTRI; p += 32) M[p] = 0;
  __syncwarp();
  _Pragma("unroll 1") for (int i = c.lane; i < m.nv; i += 32) {
    int b = m.dof_body[i];
    const T* O = xpos + 3 * b;
    T S[6], F[6];
    dof_motion(m, xanc, xaxis, i, O, S);
    inert_mul(F, crb + 10 * b, S);
    M[tri(i, i)] = dot3(S, F) + dot3(S + 3, F + 3) + m.dof_armature[i];
    for (int j = m.dof_parent[i]; j >= 0; j = m.dof_parent[j]) {
      T Sj[6];
      dof_motion(m, xanc, xaxis, j, O, Sj);
      M[tri(i, j)] = dot3(Sj, F) + dot3(Sj + 3, F + 3);
    }
  }
  __syncwarp();
}

// (scratch: the factor region O_H where the kernel stages one, else O_M -- the smooth kernel runs this before it builds M)
template <class CX> constexpr int RNE_SCRATCH() { return CX::template in_smem<O_H>() ? O_H : O_M; }
// recursive Newton-Euler: bias forces (Coriolis + centrifugal + gravity) into V_T1
template <class CX> PHASE_FN void rne_bias(CX& c) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const T *xpos = c.template S<O_XPOS>(), *xanc = c.template S<O_XANC>(), *xaxis = c.template S<O_XAXIS>(), *qvel = c.template V<V_QVEL>();
  // body velocities / accelerations / forces live in the shared-memory region of the (not yet built) factor O_H: the
  // level loops read what the previous level wrote, and from the context block that was an L2 round trip per level
  static_assert(18 * NBP <= NTRI, "rne scratch is aliased onto the matrix region");
  T *vel = c.template S<RNE_SCRATCH<CX>()>(), *acc = vel + 6 * NBP, *frc = vel + 12 * NBP;
  if (c.lane == 0) {
    for (int i = 0; i < 6; i++) { vel[i] = 0; frc[i] = 0; }
    acc[0] = acc[1] = acc[2] = 0;
    acc[3] = -m.gravity[0]; acc[4] = -m.gravity[1]; acc[5] = -m.gravity[2];
  }
  __syncwarp();
  for (int L = 1; L < m.nlevel; L++) {
    int beg = m.level_start[L], end = m.level_start[L + 1];
    for (int idx = beg + c.lane; idx < end; idx += 32) {
      int b = m.level_body[idx], p = m.body_parent[b];
      const T* O = xpos + 3 * b;
      T Ib[10];                                                  // single-body inertia (context block): fetched first
      for (int e = 0; e < 10; e++) Ib[e] = c.template G<G_CIN>()[10 * b + e];
      T r[3] = {O[0] - xpos[3 * p], O[1] - xpos[3 * p + 1], O[2] - xpos[3 * p + 2]};
      T v[6], a[6], t[3];
      for (int i = 0; i < 6; i++) { v[i] = vel[6 * p + i]; a[i] = acc[6 * p + i]; }
      cross3(t, v, r);
      v[3] += t[0]; v[4] += t[1]; v[5] += t[2];
      cross3(t, a, r);
      a[3] += t[0]; a[4] += t[1]; a[5] += t[2];
      int ja = m.body_jntadr[b], jn = m.body_jntnum[b];
      for (int k = 0; k < jn; k++) {
        int j = ja + k;
        T S[6], vj[6], u[3], w[3];
        dof_motion(m, xanc, xaxis, j, O, S);
        for (int i = 0; i < 6; i++) vj[i] = S[i] * qvel[j];
        cross3(t, v, vj);              // w x w'
        cross3(u, v, vj + 3);          // w x v'
        cross3(w, v + 3, vj);          // v x w'
        a[0] += t[0]; a[1] += t[1]; a[2] += t[2];
        a[3] += u[0] + w[0]; a[4] += u[1] + w[1]; a[5] += u[2] + w[2];
        for (int i = 0; i < 6; i++) v[i] += vj[i];
      }
      for (int i = 0; i < 6; i++) { vel[6 * b + i] = v[i]; acc[6 * b + i] = a[i]; }
      const T* I = Ib;
      T Ia[6], Iv[6], f[6], u[3];
      inert_mul(Ia, I, a);
      inert_mul(Iv, I, v);
      cross3(t, v, Iv);          // w x n
      cross3(u, v + 3, Iv + 3);  // v x f
      f[0] = Ia[0] + t[0] + u[0]; f[1] = Ia[1] + t[1] + u[1]; f[2] = Ia[2] + t[2] + u[2];
      cross3(t, v, Iv + 3);      // w x f
      f[3] = Ia[3] + t[0]; f[4] = Ia[4] + t[1]; f[5] = Ia[5] + t[2];
      for (int i = 0; i < 6; i++) frc[6 * b + i] = f[i];
    }
    __syncwarp();
  }
  for (int L = m.nlevel - 1; L >= 2; L--) {
    int beg = m.level_start[L - 1], end = m.level_start[L];
    for (int idx = beg + c.lane; idx < end; idx += 32) {
      int p = m.level_body[idx];
      T f[6];
      for (int i = 0; i < 6; i++) f[i] = frc[6 * p + i];
      for (int k = m.child_start[p]; k < m.child_start[p + 1]; k++) {
        int ch = m.child_list[k];
        T r[3] = {xpos[3 * ch] - xpos[3 * p], xpos[3 * ch + 1] - xpos[3 * p + 1], xpos[3 * ch + 2] - xpos[3 * p + 2]};
        T t[3];
        cross3(t, r, frc + 6 * ch + 3);
        for (int i = 0; i < 3; i++) { f[i] += frc[6 * ch + i] + t[i]; f[3 + i] += frc[6 * ch + 3 + i]; }
      }
      for (int i = 0; i < 6; i++) frc[6 * p + i] = f[i];
    }
    __syncwarp();
  }
  T* bias = c.template V<V_T1>();
  _Pragma("unroll 1") for (int j = c.lane; j < m.nv; j += 32) {
    int b = m.dof_body[j];
    T S[6];
    dof_motion(m, xanc, xaxis, j, xpos + 3 * b, S);
    bias[j] = dot3(S, frc + 6 * b) + dot3(S + 3, frc + 6 * b + 3);
  }
  __syncwarp();
}

// ---------------------------------------------------------------- small linear algebra on the warp
// Storage: lower triangle, row-major, rows padded to multiples of 4 (model.cuh tri()).  n <= 36: lane i owns row i and
// lanes 0..n-33 also own row i+32.  Right-hand sides live in two registers per lane (x0 = x[lane], x1 = x[lane+32])
// and pivots are broadcast with shuffles, so the triangular sweeps issue ~5 instructions per column.
template <typename T> struct Vec4 { T v[4]; };
template <typename T> DI Vec4<T> ld4(const T* p);
template <> DI Vec4<float> ld4<float>(const float* p) {
  float4 t = *reinterpret_cast<const float4*>(p);
  Vec4<float> r; r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w; return r;
}
template <> DI Vec4<double> ld4<double>(const double* p) {
  double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
  Vec4<double> r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; return r;
}
template <typename T> DI T rsqrt_t(T d);
template <> DI float rsqrt_t<float>(float d) {   // bare MUFU.RSQ (callers clamp d >= 1e-30: no denormal fix-up) + one Newton step: ~1 ulp
  float y;                                       // (__frsqrt_rn is the correctly-rounded ~20-instruction sequence, rsqrtf adds range fix-ups)
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(d));
  return y * (1.5f - 0.5f * d * y * y);
}
template <> DI double rsqrt_t<double>(double d) { return 1.0 / sqrt(d); }
template <typename T> DI T rcp_t(T d);
template <> DI float rcp_t<float>(float d) {     // MUFU.RCP + one Newton step (callers clamp d >= 1e-30); ~1 ulp, a third of the IEEE division sequence
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(d));
  return y * (2.0f - d * y);
}
template <> DI double rcp_t<double>(double d) { return 1.0 / d; }

template <typename T> DI void load_vec(const T* x, int n, int lane, T& x0, T& x1) {
  x0 = lane < n ? x[lane] : T(0);
  x1 = lane + 32 < n ? x[lane + 32] : T(0);
}
template <typename T> DI void store_vec(T* x, int n, int lane, T x0, T x1) {
  if (lane < n) x[lane] = x0;
  if (lane + 32 < n) x[lane + 32] = x1;
}

#ifndef ADROIT_BORDERED_SWEEPS
#define ADROIT_BORDERED_SWEEPS 1   // 1: register-row / bordered sweeps below (a third of the instructions of the plain loops, chain = shuffle + FMA per step)
#endif
#if ADROIT_BORDERED_SWEEPS
// Triangular sweeps.  The leading min(n,32) block runs on ONE register per lane (x0 = x[lane]), fully unrolled, with nothing on
// the dependent chain but one shuffle and one FMA per step:
//   forward  (column k of L = entry k of every lane's own row): the lane's row is read ONCE into 32 registers (8 vector loads
//            instead of 32 bank-conflicting scalar ones) and, for SCALE, pre-multiplied by dinv[k], so that the value broadcast in
//            step k is the lane's running, unscaled x[k];
//   backward (row k of L is contiguous: one conflict-free load per step, independent of the chain): every lane carries its
//            entry already multiplied by its own dinv, so that lane k's register IS x[k] when step k broadcasts it.
// Both rely on what the factorisations leave in the padded rows: ZERO on the diagonal (chol_dense / ltdl_factor store it so; the
// pivots live in dinv) and zero pads right of it (never written since the allocation).
// The 1..4 rows beyond lane 31 (n = 33 hammer, 36 relocate) are a border whose entries every lane keeps redundantly in
// registers: each border row costs one coalesced row read and (forward sweep) one warp sum.
// for k = n-1..0: xk = x[k] (*dinv[k] if SCALE);  x[i] -= L[k][i] * xk for i < k      (row k of L is contiguous)
template <typename T, bool SCALE> DI void sweep_rows_desc(const T* L, const T* dinv, T d0, T d1, T& x0, T& x1, int n, int lane) {
  const unsigned FULL = 0xffffffffu;
  (void)dinv;
  if (n > 32) {
    T xb[4], db[4];
#pragma unroll
    for (int q = 0; q < 4; q++) { xb[q] = __shfl_sync(FULL, x1, q); db[q] = __shfl_sync(FULL, d1, q); }
#pragma unroll
    for (int q = 3; q >= 0; q--) {
      if (32 + q < n) {                      // warp-uniform
        const T* row = L + rowoff(32 + q);
        if (SCALE) xb[q] *= db[q];
#pragma unroll
        for (int p = 0; p < q; p++) xb[p] -= row[32 + p] * xb[q];
        x0 -= row[lane] * xb[q];
      }
    }
    x1 = lane == 0 ? xb[0] : lane == 1 ? xb[1] : lane == 2 ? xb[2] : lane == 3 ? xb[3] : T(0);
  }
  // (rows k >= n of the leading block, when n < 32, are zero: their steps change nothing and are not guarded)
  T u = SCALE ? x0 * d0 : x0;
#pragma unroll
  for (int k = 31; k >= 1; k--) {
    T lk = L[rowoff(k) + lane];              // lanes >= k read the zero diagonal / pads or, beyond the row, a later row: discarded below
    if (SCALE) lk *= d0;
    T xk = __shfl_sync(FULL, u, k);
    if (lane < k) u -= lk * xk;
  }
  x0 = u;
}
// for k = 0..n-1: xk = x[k] (*dinv[k] if SCALE);  x[i] -= L[i][k] * xk for i > k      (column k of L: one row per lane)
template <typename T, bool SCALE> DI void sweep_cols_asc(const T* L, const T* dinv, T d0, T d1, T& x0, T& x1, int n, int lane) {
  const unsigned FULL = 0xffffffffu;
  const T* r0 = L + rowoff(lane);
  T a[32];
#pragma unroll
  for (int g = 0; g < 8; g++) {
    Vec4<T> v;
    v.v[0] = v.v[1] = v.v[2] = v.v[3] = T(0);
    if (4 * g <= lane) v = ld4(r0 + 4 * g);  // groups right of the lane's diagonal group belong to later rows
    if (SCALE) {
      Vec4<T> dk = ld4(dinv + 4 * g);        // (broadcast read)
#pragma unroll
      for (int q = 0; q < 4; q++) v.v[q] *= dk.v[q];
    }
#pragma unroll
    for (int q = 0; q < 4; q++) a[4 * g + q] = v.v[q];
  }
#pragma unroll
  for (int k = 0; k < 31; k++) {
    T xk = __shfl_sync(FULL, x0, k);
    x0 -= a[k] * xk;                         // a[k] = 0 for k >= lane
  }
  if (SCALE) x0 *= d0;
  if (n > 32) {
    T xb[4], db[4];
#pragma unroll
    for (int q = 0; q < 4; q++) { xb[q] = __shfl_sync(FULL, x1, q); db[q] = __shfl_sync(FULL, d1, q); }
#pragma unroll
    for (int q = 0; q < 4; q++) {
      if (32 + q < n) {                      // warp-uniform
        const T* row = L + rowoff(32 + q);
        T s = warp_sum(row[lane] * x0);
#pragma unroll
        for (int p = 0; p < q; p++) s += row[32 + p] * xb[p];
        xb[q] -= s;
        if (SCALE) xb[q] *= db[q];
      }
    }
    x1 = lane == 0 ? xb[0] : lane == 1 ? xb[1] : lane == 2 ? xb[2] : lane == 3 ? xb[3] : T(0);
  }
}
#else
// for k = n-1..0: xk = x[k] (*dinv[k] if SCALE);  x[i] -= L[k][i] * xk for i < k      (row k of L is contiguous)
template <typename T, bool SCALE> DI void sweep_rows_desc(const T* L, const T*, T d0, T d1, T& x0, T& x1, int n, int lane) {
#pragma unroll 2
  for (int k = n - 1; k >= 0; k--) {
    T xs = k < 32 ? x0 : x1;
    if (SCALE) xs *= k < 32 ? d0 : d1;
    T xk = __shfl_sync(0xffffffffu, xs, k & 31);
    if (SCALE && lane == (k & 31)) { if (k < 32) x0 = xk; else x1 = xk; }
    const T* row = L + rowoff(k);
    if (lane < k) x0 -= row[lane] * xk;
    if (k > 32 && lane + 32 < k) x1 -= row[lane + 32] * xk;
  }
}
// for k = 0..n-1: xk = x[k] (*dinv[k] if SCALE);  x[i] -= L[i][k] * xk for i > k      (column k of L: one row per lane)
template <typename T, bool SCALE> DI void sweep_cols_asc(const T* L, const T*, T d0, T d1, T& x0, T& x1, int n, int lane) {
  const T *r0 = L + rowoff(lane), *r1 = L + rowoff(lane + 32);
#pragma unroll 2
  for (int k = 0; k < n; k++) {
    T xs = k < 32 ? x0 : x1;
    if (SCALE) xs *= k < 32 ? d0 : d1;
    T xk = __shfl_sync(0xffffffffu, xs, k & 31);
    if (SCALE && lane == (k & 31)) { if (k < 32) x0 = xk; else x1 = xk; }
    if (lane > k && lane < n) x0 -= r0[k] * xk;
    if (lane + 32 > k && lane + 32 < n) x1 -= r1[k] * xk;
  }
}

#endif

// in-place Cholesky A = L L^T; dinv[j] = 1 / L[j][j].  Leading min(n,32) block: right-looking, lane = row, the row held
// in 32 REGISTERS (fully unrolled, compile-time indices): column k is scaled, stored, and broadcast to the warp through
// a 32-word shared-memory buffer (`col`, two of them alternating: one warp barrier per column), then every lane updates
// the rest of its row with register FMAs -- ~1.25 instructions per updated entry instead of a dot product re-read from
// shared memory per entry.  (The kernel is shared-memory-bound at 16 warps/SM, so the registers are free: 128/thread.)
// Entries right of the diagonal hold don't-care values that never reach a stored result.
// Rows 32..n-1 (n <= 36) are "bordered" on afterwards: each is one forward substitution
// against the leading factor plus a few warp-wide dot products, instead of doubling every column step for 1-4 rows.
#ifndef ADROIT_REG_CHOL
#define ADROIT_REG_CHOL 2
#endif
#ifndef ADROIT_CHOL_BORDER_INLOOP
#define ADROIT_CHOL_BORDER_INLOOP 1
#endif
template <typename T, int NBORD> DI void chol_dense_impl(T* A, T* dinv, T* col0, T* col1, int n, int lane) {
  const int nb = n < 32 ? n : 32;
  T* r0 = A + rowoff(lane);
#if ADROIT_REG_CHOL == 2
  // rolled by panels of 4 columns: after each panel the register row is rotated by 4, so that the panel code addresses
  // registers a[0..3] (pivot columns) and a[4..31] (trailing entries) for every panel -- the code footprint of ONE panel
  // (~4 KB instead of ~30 KB fully unrolled: the solve kernel is instruction-cache bound), the executed work of the
  // unrolled version (trailing vector groups beyond column 31 are skipped by warp-uniform predicates).
  T a[32];
#pragma unroll
  for (int g = 0; g < 8; g++) {
    Vec4<T> v = ld4(r0 + 4 * g);
#pragma unroll
    for (int q = 0; q < 4; q++) a[4 * g + q] = v.v[q];
  }
  // Pivot look-ahead: the next pivot A[k+1][k+1] - L[k+1][k]^2 needs nothing but lane k+1's own multiplier, so it is formed and
  // broadcast BEFORE the column goes through shared memory -- the dependent chain of a column is pivot -> rsqrt -> multiplier ->
  // FMA -> shuffle instead of ... -> store -> barrier -> load -> FMA -> shuffle (the kernel is latency-bound, profiles/R2i_*).
  // Same operands, same FMA as the regular update of that entry: bit-identical factor.
  T dnext = __shfl_sync(0xffffffffu, a[0], 0);
#if ADROIT_CHOL_BORDER_INLOOP
  // Rows 32..n-1 ride along: lane j holds A[32+q][j]; at column k lane k's entry times 1/L[k][k] IS L[32+q][k], which is broadcast
  // and folded into the entries right of it with the lanes' own multipliers -- three instructions per column and border row, none of
  // them on the pivot chain, instead of a 32-step forward substitution per border row after the factorisation (a third of the
  // factorisation's latency for hammer's single row, more than half for relocate's four).
  T bq[4];
#pragma unroll
  for (int q = 0; q < 4; q++) bq[q] = (q < NBORD && 32 + q < n) ? A[rowoff(32 + q) + lane] : T(0);
#endif
#pragma unroll 1
  for (int k0 = 0; k0 < nb; k0 += 4) {
#pragma unroll
    for (int c = 0; c < 4; c++) {
      const int k = k0 + c;
      if (NBORD > 0 || k < nb) {                    // warp-uniform; with border rows n > 32 and every column exists: no branch at all
        T d = dnext;
        d = d > tiny<T>() ? d : tiny<T>();
        const T inv = rsqrt_t(d);
        const T l = lane > k ? a[c] * inv : T(0);     // (the diagonal is stored as zero: the pivot lives in dinv, the sweeps rely on it)
        dnext = __shfl_sync(0xffffffffu, a[c + 1] - l * l, (k + 1) & 31);
        T* cb = (c & 1) ? col1 : col0;
        cb[lane] = l;
        if (lane >= k && lane < nb) r0[k] = l;
        if (lane == k) dinv[k] = inv;
#if ADROIT_CHOL_BORDER_INLOOP
#pragma unroll
        for (int q = 0; q < 4; q++) {
          if (q < NBORD) {                          // compile-time: no branch, no reconvergence bookkeeping in the column loop
            const T lb = __shfl_sync(0xffffffffu, bq[q] * inv, k);
            bq[q] = lane == k ? lb : (lane > k ? bq[q] - lb * l : bq[q]);
          }
        }
#endif
        __syncwarp();
#pragma unroll
        for (int g = 0; g < 8; g++) {
          if (4 * g + 3 > c && k0 + 4 * g < 32) {   // second condition warp-uniform
            Vec4<T> cv = ld4(cb + k0 + 4 * g);      // broadcast read of L[k0+4g .. k0+4g+3][k]
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (4 * g + q > c) a[4 * g + q] -= l * cv.v[q];
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 28; j++) a[j] = a[j + 4];
  }
  __syncwarp();
#elif ADROIT_REG_CHOL == 1
  T a[32];
#pragma unroll
  for (int g = 0; g < 8; g++) {
    Vec4<T> v = ld4(r0 + 4 * g);
#pragma unroll
    for (int q = 0; q < 4; q++) a[4 * g + q] = v.v[q];
  }
#pragma unroll
  for (int k = 0; k < 32; k++) {
    if (k < nb) {                                   // warp-uniform
      T d = __shfl_sync(0xffffffffu, a[k], k);
      d = d > tiny<T>() ? d : tiny<T>();
      const T inv = rsqrt_t(d);
      const T l = lane > k ? a[k] * inv : T(0);   // lanes on and above the diagonal sit the column out (diagonal stored as zero)
      T* cb = (k & 1) ? col1 : col0;
      cb[lane] = l;
      if (lane >= k && lane < nb) r0[k] = l;
      if (lane == k) dinv[k] = inv;
      __syncwarp();
#pragma unroll
      for (int g = (k + 1) >> 2; g < 8; g++) {
        Vec4<T> cv = ld4(cb + 4 * g);               // broadcast read of L[4g..4g+3][k]
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (4 * g + q > k) a[4 * g + q] -= l * cv.v[q];
      }
    }
  }
  __syncwarp();
#else
  for (int j = 0; j < nb; j++) {
    const T* rj = A + rowoff(j);
    T s0 = 0;
    int k = 0;
    for (; k + 4 <= j; k += 4) {
      Vec4<T> b = ld4(rj + k), a = ld4(r0 + k);
      s0 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
    }
    if (k < j) {          // last partial chunk: entries at columns >= j are masked out (j is warp-uniform)
      Vec4<T> b = ld4(rj + k), a = ld4(r0 + k);
      s0 += a.v[0] * b.v[0] + (k + 1 < j ? a.v[1] * b.v[1] : T(0)) + (k + 2 < j ? a.v[2] * b.v[2] : T(0));
    }
    s0 = r0[j] - s0;      // lanes with lane < j read a neighbouring row: harmless, never stored
    T d = __shfl_sync(0xffffffffu, s0, j);
    d = d > tiny<T>() ? d : tiny<T>();
    T inv = rsqrt_t(d);
    if (lane == j) { r0[j] = T(0); dinv[j] = inv; }
    else if (lane > j && lane < nb) r0[j] = s0 * inv;
    __syncwarp();
  }
#endif
#if ADROIT_CHOL_BORDER_INLOOP && ADROIT_REG_CHOL == 2
#pragma unroll
  for (int q = 0; q < 4; q++)
    if (q < NBORD && 32 + q < n) A[rowoff(32 + q) + lane] = bq[q];
  __syncwarp();
#endif
  for (int r = 32; r < n; r++) {
    T* rr = A + rowoff(r);
#if !(ADROIT_CHOL_BORDER_INLOOP && ADROIT_REG_CHOL == 2)
    T x = rr[lane], dl = dinv[lane];
#pragma unroll 4
    for (int k = 0; k < 32; k++) {          // L11 y = a
      T yk = __shfl_sync(0xffffffffu, x * dl, k);
      if (lane == k) x = yk;
      if (lane > k) x -= r0[k] * yk;
    }
    rr[lane] = x;
    __syncwarp();
#endif
    for (int q = 32; q <= r; q++) {
      const T* rq = A + rowoff(q);
      T p = rr[lane] * rq[lane];
      if (lane + 32 < q) p += rr[lane + 32] * rq[lane + 32];
      p = warp_sum(p);
      if (q < r) {
        if (lane == 0) rr[q] = (rr[q] - p) * dinv[q];
      } else {
        T d = rr[r] - p;
        d = d > tiny<T>() ? d : tiny<T>();
        T inv = rsqrt_t(d);
        if (lane == 0) { rr[r] = d * inv; dinv[r] = inv; }
      }
      __syncwarp();
    }
  }
}
// (one instance per border-row count the four tasks have: 0 door / pen, 1 hammer, 4 relocate and anything in between)
template <typename T> PHASE_FN void chol_dense(T* A, T* dinv, T* col0, T* col1, int n, int lane) {
#if ADROIT_CHOL_BORDER_INLOOP && ADROIT_REG_CHOL == 2
  if (n <= 32) chol_dense_impl<T, 0>(A, dinv, col0, col1, n, lane);
  else if (n == 33) chol_dense_impl<T, 1>(A, dinv, col0, col1, n, lane);
  else chol_dense_impl<T, 4>(A, dinv, col0, col1, n, lane);
#else
  chol_dense_impl<T, 0>(A, dinv, col0, col1, n, lane);
#endif
}
// x <- (L L^T)^-1 x  (x in shared memory)
template <typename T> DI void chol_solve(const T* L, const T* dinv, T* x, int n, int lane) {
  T x0, x1, d0, d1;
  load_vec(x, n, lane, x0, x1);
  load_vec(dinv, n, lane, d0, d1);
  sweep_cols_asc<T, true>(L, dinv, d0, d1, x0, x1, n, lane);
  sweep_rows_desc<T, true>(L, dinv, d0, d1, x0, x1, n, lane);
  store_vec(x, n, lane, x0, x1);
  __syncwarp();
}

// Tree-sparse factorisation M = L^T D L of a matrix with the sparsity of the joint-space inertia (entry (k,i) non-zero
// only if i is an ancestor dof of k): eliminating leaves first creates no fill-in (Featherstone; MuJoCo's mj_factorM
// uses the same order).  One step per dof, its (ancestor, ancestor) pair updates spread over the lanes; the schedule
// ld_start / ld_pair is built on the host.  On return: strict lower part = L, diagonal = 0, dinv = 1/D.
#ifndef ADROIT_LTDL_LEVEL
#define ADROIT_LTDL_LEVEL 0   // measured equal to the serial form on B200 (profiles/R2c_ab_lockstep.txt): the chain of ~50 barrier-separated steps is as long as 33 pivots
#endif
// Level-scheduled form: the dofs of one depth are eliminated together (the five fingers in parallel, the object's chain
// alongside), deepest level first: 9-12 levels instead of 30-36 serial pivots.  The program (steps in the kernel
// parameters, items in the model arena; built on the host, adroit_b200.cu: fill_model) is, per level: reciprocal pivots ->
// rounds of (ancestor, ancestor) updates with pairwise different targets -> row scaling.  Same products as the serial
// form below, summed in a fixed (different) order.
template <typename T> PHASE_FN void ltdl_factor_lv(const DModel<T>& m, T* A, T* dinv, int lane) {
  const int* items = m.lv_items;
  int s = 0, w0 = m.lv_steps[0], beg = w0 >> 2;
  int nxt = items[beg + lane];                         // (the item array is padded by a warp's worth of words)
#pragma unroll 1
  while ((w0 & 3) != 3) {
    const int type = w0 & 3, w1 = m.lv_steps[s + 1], end = w1 >> 2;
    const int cur = nxt;
    nxt = items[end + lane];                           // first pass of the next step: off the dependent chain
#pragma unroll 1
    for (int p = beg + lane; p < end; p += 32) {
      const int w = p == beg + lane ? cur : items[p];
      if (type == 0) {
        T d = A[w >> 8];
        d = d > tiny<T>() ? d : tiny<T>();
        dinv[w & 255] = rcp_t(d);
      } else if (type == 1) {
        const int k = (w >> 22) & 63, rk = rowoff(k);
        A[w & 1023] -= A[rk + ((w >> 10) & 63)] * A[rk + ((w >> 16) & 63)] * dinv[k];
      } else {
        A[w & 1023] *= dinv[(w >> 10) & 63];
      }
    }
    __syncwarp();
    s++; w0 = w1; beg = end;
  }
}
template <typename T> PHASE_FN void ltdl_factor_serial(const DModel<T>& m, T* A, T* dinv, int lane) {
  const int n = m.nv;
  // schedule words (host-built): packed offset of A[i][j] | i << 10 | j << 16.  The first pass of the NEXT pivot's
  // words is fetched while the current pivot is eliminated: no model-table load on the chain of the 33 serial pivots.
  auto start = [&](int k) { return m.ld_start_v[k]; };   // (by value in the kernel parameters: a uniform constant-bank read)
  int beg = start(n - 1), end = start(n);
  int nxt = beg + lane < end ? m.ld_pair[beg + lane] : 0;
  for (int k = n - 1; k >= 0; k--) {
    const int cb = beg, ce = end, cur = nxt;
    if (k > 0) {
      end = beg;
      beg = start(k - 1);
      nxt = beg + lane < end ? m.ld_pair[beg + lane] : 0;
    }
    const T* rk = A + rowoff(k);
    T dk = rk[k];
    dk = dk > tiny<T>() ? dk : tiny<T>();
    T inv = rcp_t(dk);
    for (int p = cb + lane; p < ce; p += 32) {
      const int w = p == cb + lane ? cur : m.ld_pair[p];
      const int off = w & 1023, i = (w >> 10) & 63, j = (w >> 16) & 63;
      A[off] -= rk[i] * rk[j] * inv;
    }
    __syncwarp();
    // row k: L[k][i] = A[k][i] / D[k] for every i < k (entries outside the ancestor pattern are zero and stay zero)
    for (int i = lane; i < k; i += 32) A[rowoff(k) + i] = rk[i] * inv;
    if (lane == 0) { dinv[k] = inv; A[rowoff(k) + k] = T(0); }   // D lives in dinv; the sweeps want a zero diagonal in the rows
    __syncwarp();
  }
}
template <typename T> PHASE_FN void ltdl_factor(const DModel<T>& m, T* A, T* dinv, int lane) {
#if ADROIT_LTDL_LEVEL
  ltdl_factor_lv(m, A, dinv, lane);
  for (int j = lane; j < m.nv; j += 32) A[tri(j, j)] = T(0);
  __syncwarp();
#else
  ltdl_factor_serial(m, A, dinv, lane);
#endif
}
// x <- (L^T D L)^-1 x  (x in shared memory)
template <typename T> DI void ltdl_solve(const T* L, const T* dinv, T* x, int n, int lane) {
  T x0, x1, d0, d1;
  load_vec(x, n, lane, x0, x1);
  load_vec(dinv, n, lane, d0, d1);
  sweep_rows_desc<T, false>(L, dinv, d0, d1, x0, x1, n, lane);
  x0 *= d0; x1 *= d1;
  sweep_cols_asc<T, false>(L, dinv, d0, d1, x0, x1, n, lane);
  store_vec(x, n, lane, x0, x1);
  __syncwarp();
}
// M^-1 = W D^-1 W^T with W = L^-1 (unit lower, same ancestor pattern as L).
//  1. W, one ROW per lane, in registers: along the ancestor chain j_0 = parent(k), j_1, ... of dof k
//        W[k][j_t] = -L[k][j_t] - sum_{s<t} W[k][j_s] L[j_s][j_t]           (from W L = I)
//     reads only L (shared memory) and the lane's own earlier entries, so the rows need no level ordering and no
//     barrier; the chain table anc_tab is built on the host.  W then replaces L IN PLACE (unit diagonal stored).
//  2. every entry of M^-1 as a short sum over the common ancestors of its two dofs (lca table), W and the parent
//     table read from shared memory: the walk up the chain is a dependent chain of ~25-cycle loads, not L2 round trips.
// `par` is MAXDEPTH.. >= nv ints of shared scratch.  Pad entries of Minv are never written (zero since allocation).
constexpr int MAXDEPTH = 14;   // longest dof ancestor chain (relocate: 6 arm + 2 wrist + 5 finger = 13); checked on the host
template <typename T> PHASE_FN void minv_tree(const DModel<T>& m, T* L, const T* dinv, int* par, T* Minv, int lane) {
  const unsigned FULL = 0xffffffffu;
  const int nv = m.nv;
  // One rolled pass per 32 rows, rows 32.. FIRST: their chains read rows below 32 (still L) and each other (all reads of a pass
  // come before its writes), while the rows below 32 never read a row at or beyond 32 -- so the unrolled chain code exists once.
#pragma unroll 1
  for (int pass = nv > 32 ? 1 : 0; pass >= 0; pass--) {
    const int k = lane + 32 * pass;
    const bool act = k < nv;
    if (act) par[k] = m.dof_parent[k];
    const int d = act ? m.dof_depth[k] : 0;
    const int maxd = __reduce_max_sync(FULL, d);
    const int* an = m.anc_tab + (act ? k : 0) * MAXDEPTH;
    T* rk = L + rowoff(act ? k : 0);
    T w[MAXDEPTH];
    int j[MAXDEPTH], ro[MAXDEPTH];
#pragma unroll
    for (int t = 0; t < MAXDEPTH; t++) {
      j[t] = 0; ro[t] = 0; w[t] = 0;
      if (t < maxd) {                                 // warp-uniform
        j[t] = t < d ? an[t] : 0;
        ro[t] = rowoff(j[t]);
        T acc = rk[j[t]];
#pragma unroll
        for (int q = 0; q < t; q++) acc += w[q] * L[ro[q] + j[t]];
        w[t] = t < d ? -acc : T(0);
      }
    }
    __syncwarp();
    if (act) {
#pragma unroll
      for (int t = 0; t < MAXDEPTH; t++)
        if (t < d) rk[j[t]] = w[t];
      rk[k] = T(1);
    }
    __syncwarp();
  }
  // (a, b, lca(a,b)) packed in one table word, fetched one pass ahead: the model-table load is off the dependent chain
  int nxt = lane < m.n_tri ? m.tri_abk[lane] : 0;
  for (int p = lane; p < m.n_tri; p += 32) {
    const int abk = nxt;
    if (p + 32 < m.n_tri) nxt = m.tri_abk[p + 32];
    const int a = abk & 255, b = (abk >> 8) & 255;     // a >= b
    int k = (abk >> 16) & 255;
    const T *ra = L + rowoff(a), *rb = L + rowoff(b);
    T s = 0;
    while (k != 255) {
      s += ra[k] * rb[k] * dinv[k];
      int pk = par[k];
      k = pk < 0 ? 255 : pk;
    }
    Minv[tri(a, b)] = s;
  }
  __syncwarp();
}

// y = A x for a symmetric matrix (lower storage); x, y in shared memory, y != x
// COMPACT: rolled loops (the solve kernel is instruction-fetch bound: 1.5 KB of code instead of 6 KB)
// SH: A, x and y are all in shared memory (the solve kernel); otherwise A and x are, y may be anywhere
// y = A x for an x that is zero outside a (warp-uniform) index range, found by ballot: a contact tangent of the object against
// the table touches 6 of the 33 dofs, so the row loop runs over 2-3 vector chunks instead of 9 and the transpose loop over 6
// rows instead of 32.  Same products in the same order as symv for the entries that are not skipped (which are exact zeros).
template <typename T> SUB_FN void symv_range(const T* A, const T* x, T* y, int n, int lane) {
  __builtin_assume(__isShared(A));
  __builtin_assume(__isShared(x));
  const int i0 = lane, i1 = lane + 32;
  const T xa = i0 < n ? x[i0] : T(0), xb = i1 < n ? x[i1] : T(0);
  const unsigned m0 = __ballot_sync(0xffffffffu, xa != T(0)), m1 = __ballot_sync(0xffffffffu, xb != T(0));
  T y0 = 0, y1 = 0;
  if (m0 | m1) {
    const int lo = m0 ? __ffs(m0) - 1 : 32 + __ffs(m1) - 1, hi = m1 ? 63 - __clz(m1) : 31 - __clz(m0);
    const int k0 = lo & ~3;
    if (i0 < n) {
      const T* r = A + rowoff(i0);
      const int ke = i0 < hi ? i0 : hi;
      _Pragma("unroll 1") for (int k = k0; k <= ke; k += 4) {
        Vec4<T> a = ld4(r + k), b = ld4(x + k);
        y0 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
      }
    }
    if (i1 < n) {
      const T* r = A + rowoff(i1);
      const int ke = i1 < hi ? i1 : hi;
      _Pragma("unroll 1") for (int k = k0; k <= ke; k += 4) {
        Vec4<T> a = ld4(r + k), b = ld4(x + k);
        y1 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
      }
    }
    _Pragma("unroll 2") for (int j = lo > 1 ? lo : 1; j <= hi; j++) {
      const T* rj = A + rowoff(j);
      T xj = x[j];
      if (i0 < j) y0 += rj[i0] * xj;
      if (j > 32 && i1 < j) y1 += rj[i1] * xj;
    }
  }
  if (i0 < n) y[i0] = y0;
  if (i1 < n) y[i1] = y1;
}
#ifndef ADROIT_FAST_SYMV
#define ADROIT_FAST_SYMV 1   // fully unrolled on registers (x once into 32 registers, compile-time row offsets): ~230 instructions instead of ~540
#endif
template <typename T, bool COMPACT = false, bool SH = false> SUB_FN void symv(const T* A, const T* x, T* y, int n, int lane) {
  __builtin_assume(__isShared(A));
  __builtin_assume(__isShared(x));
  if constexpr (SH) __builtin_assume(__isShared(y));
#if ADROIT_FAST_SYMV
  {
    // Leading 32x32 block: lane i owns y[i].  Relies on zero pads right of the diagonal in the rows of A, on zero rows of A and
    // zero entries of x at and beyond n (all zero since the allocation / the staging: no loop ever writes them).
    T xs[32];
#pragma unroll
    for (int g = 0; g < 8; g++) {
      Vec4<T> v = ld4(x + 4 * g);              // (broadcast read)
#pragma unroll
      for (int q = 0; q < 4; q++) xs[4 * g + q] = v.v[q];
    }
    const T* r0 = A + rowoff(lane);
    T y0 = 0;
#pragma unroll
    for (int g = 0; g < 8; g++) {              // row part: sum_{k <= i} A[i][k] x[k]
      if (4 * g <= lane) {
        Vec4<T> a = ld4(r0 + 4 * g);
        y0 += a.v[0] * xs[4 * g] + a.v[1] * xs[4 * g + 1] + a.v[2] * xs[4 * g + 2] + a.v[3] * xs[4 * g + 3];
      }
    }
#pragma unroll
    for (int j = 1; j < 32; j++) {             // transpose part: row j contributes A[j][i] x[j] to every i < j (row j is contiguous)
      T aj = A[rowoff(j) + lane];
      if (lane < j) y0 += aj * xs[j];
    }
    if (n > 32) {
      const T xl = x[lane];
      Vec4<T> xv = ld4(x + 32);
      T yb[4] = {T(0), T(0), T(0), T(0)};
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (32 + q < n) {                      // warp-uniform
          const T* row = A + rowoff(32 + q);
          const T aq = row[lane];
          y0 += aq * xv.v[q];
          T sq = warp_sum(aq * xl);
#pragma unroll
          for (int p = 0; p < 4; p++) {
            if (p <= q) sq += row[32 + p] * xv.v[p];
            else if (32 + p < n) sq += A[rowoff(32 + p) + 32 + q] * xv.v[p];
          }
          yb[q] = sq;
        }
      }
      const T y1 = lane == 0 ? yb[0] : lane == 1 ? yb[1] : lane == 2 ? yb[2] : yb[3];
      if (lane + 32 < n) y[lane + 32] = y1;
    }
    if (lane < n) y[lane] = y0;
    return;
  }
#endif
  T y0 = 0, y1 = 0;
  const int i0 = lane, i1 = lane + 32;
  if constexpr (COMPACT) {
    if (i0 < n) {
      const T* r = A + rowoff(i0);
      _Pragma("unroll 1") for (int k = 0; k <= i0; k += 4) {
        Vec4<T> a = ld4(r + k), b = ld4(x + k);
        y0 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
      }
    }
    if (i1 < n) {
      const T* r = A + rowoff(i1);
      _Pragma("unroll 1") for (int k = 0; k <= i1; k += 4) {
        Vec4<T> a = ld4(r + k), b = ld4(x + k);
        y1 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
      }
    }
    _Pragma("unroll 2") for (int j = 1; j < n; j++) {
      const T* rj = A + rowoff(j);
      T xj = x[j];
      if (i0 < j) y0 += rj[i0] * xj;
      if (j > 32 && i1 < j) y1 += rj[i1] * xj;
    }
    if (i0 < n) y[i0] = y0;
    if (i1 < n) y[i1] = y1;
    return;
  }
  if (i0 < n) {
    const T* r = A + rowoff(i0);
    for (int k = 0; k <= i0; k += 4) {      // pad entries of the row are zero
      Vec4<T> a = ld4(r + k), b = ld4(x + k);
      y0 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
    }
  }
  if (i1 < n) {
    const T* r = A + rowoff(i1);
    for (int k = 0; k <= i1; k += 4) {
      Vec4<T> a = ld4(r + k), b = ld4(x + k);
      y1 += a.v[0] * b.v[0] + a.v[1] * b.v[1] + a.v[2] * b.v[2] + a.v[3] * b.v[3];
    }
  }
  for (int j = 1; j < n; j++) {             // transpose part: row j contributes A[j][i] x[j] to every i < j
    const T* rj = A + rowoff(j);
    T xj = x[j];
    if (i0 < j) y0 += rj[i0] * xj;
    if (j > 32 && i1 < j) y1 += rj[i1] * xj;
  }
  if (i0 < n) y[i0] = y0;
  if (i1 < n) y[i1] = y1;
}
template <typename T> DI T sym_get(const T* A, int i, int j) { return i >= j ? A[tri(i, j)] : A[tri(j, i)]; }

// passive + actuation + qfrc_smooth + qacc_smooth (+ M^-1 for the noslip pass); M stays in O_M, O_H is scratch
template <class CX> PHASE_FN void smooth_dynamics(CX& c, bool need_minv) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const T *qpos = c.template V<V_QPOS>(), *qvel = c.template V<V_QVEL>(), *ctrl = c.template V<V_CTRL>();
  // shared memory is what bounds this kernel's occupancy (7 blocks/SM with the map as it is): the actuator forces are
  // gathered in the qacc_smooth vector itself, qfrc_smooth goes straight to the context block (V_QFS is not staged here),
  // and the bias vector's slot doubles as the int scratch of minv_tree once the forces are formed
  T *bias = c.template V<V_T1>(), *qfs = c.template V<V_QFS>(), *qas = c.template V<V_QAS>();
  _Pragma("unroll 1") for (int j = c.lane; j < m.nv; j += 32) qas[j] = 0;
  __syncwarp();
  _Pragma("unroll 1") for (int u = c.lane; u < m.nu; u += 32) {
    T cu = tclamp(ctrl[u], m.act_ctrlrange[2 * u], m.act_ctrlrange[2 * u + 1]);
    int j = m.act_dof[u];
    qas[j] = m.act_gain[u] * cu + m.act_bias[3 * u] + m.act_bias[3 * u + 1] * qpos[j] + m.act_bias[3 * u + 2] * qvel[j];
  }
  __syncwarp();
  _Pragma("unroll 1") for (int j = c.lane; j < m.nv; j += 32) {
    T f = -m.dof_damping[j] * qvel[j] - bias[j] + qas[j];
    qfs[j] = f;
    qas[j] = f;
    c.template G<G_DBG_BIAS>()[j] = bias[j];
  }
  T* H = c.template S<O_M>();      // factored in place: M itself has already left for the context block (store_region_early)
  __syncwarp();
  ltdl_factor(m, H, c.template V<V_DINV>(), c.lane);
  ltdl_solve(H, c.template V<V_DINV>(), qas, m.nv, c.lane);
  if (need_minv) minv_tree(m, H, c.template V<V_DINV>(), reinterpret_cast<int*>(bias), c.template G<G_MINV>(), c.lane);   // bias (V_T1): free by now
}

}  // namespace adroit
