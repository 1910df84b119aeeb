// task.cuh — the reference's task layer fused into the step kernel: action map, observation packing, reward,
// goal_achieved / done, reset randomisation (counter-based Philox instead of the reference's unseeded numpy PCG64,
// SURVEY.md App. A.8 Q11).  Paths below are relative to /root/reference/mj_envs_vision/hand_manipulation_suite/.
#pragma once
#include "constraint.cuh"

namespace adroit {

// Philox4x32-10 (Salmon et al., SC'11)
DI void philox4x32(unsigned k0, unsigned k1, unsigned c0, unsigned c1, unsigned c2, unsigned c3, unsigned* out) {
  for (int r = 0; r < 10; r++) {
    unsigned long long p0 = (unsigned long long)0xD2511F53u * c0, p1 = (unsigned long long)0xCD9E8D57u * c2;
    unsigned n0 = (unsigned)(p1 >> 32) ^ c1 ^ k0, n1 = (unsigned)p1, n2 = (unsigned)(p0 >> 32) ^ c3 ^ k1, n3 = (unsigned)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
DI double u01(unsigned x) { return ((double)x + 0.5) * (1.0 / 4294967296.0); }

// reset_model draws (hammer_v0.py:109, door_v0.py:107-109, pen_v0.py:120-122 + utils/quatmath.py:60-76,
// relocate_v0.py:89-93); evaluated in fp64 so that the fp32 and fp64 builds and the CPU oracle draw identical values
template <typename T> DI void draw_params(const DModel<T>& m, unsigned long long seed, unsigned long long genv, unsigned episode, T* par) {
  unsigned r[8];
  philox4x32((unsigned)seed, (unsigned)(seed >> 32), (unsigned)genv, (unsigned)(genv >> 32), episode, 0, r);
  philox4x32((unsigned)seed, (unsigned)(seed >> 32), (unsigned)genv, (unsigned)(genv >> 32), episode, 1, r + 4);
  for (int i = 0; i < m.npar; i++) par[i] = m.par0[i];
  switch (m.task) {
    case T_HAMMER:
      par[2] = (T)(0.1 + u01(r[0]) * (0.25 - 0.1));
      switch (m.task_int[8]) {   // variations, hammer_v0.py:115-125 (drawn after the board height, same order)
        case VAR_MASS: par[3] = (T)(0.05 + u01(r[1]) * (2.5 - 0.05)); break;
        case VAR_POS: {
          double x = -0.24 + u01(r[1]) * (-0.10 + 0.24);
          par[3] = (T)x;
          par[4] = (T)(-0.14 - (-0.24 - x));
        } break;
        case VAR_SIZE:
          par[3] = (T)(0.01 + u01(r[1]) * (0.04 - 0.01));
          par[4] = (T)(0.02 + u01(r[2]) * (0.08 - 0.02));
          break;
        default: break;
      }
      break;
    case T_DOOR:
      par[0] = (T)(-0.3 + u01(r[0]) * 0.1);
      par[1] = (T)(0.25 + u01(r[1]) * 0.1);
      par[2] = (T)(0.252 + u01(r[2]) * (0.35 - 0.252));
      break;
    case T_PEN: {
      double e0 = -1 + 2 * u01(r[0]), e1 = -1 + 2 * u01(r[1]), e2 = 0;
      double ai = e2 / 2, aj = -e1 / 2, ak = e0 / 2;
      double si = sin(ai), sj = sin(aj), sk = sin(ak), ci = cos(ai), cj = cos(aj), ck = cos(ak);
      double cc = ci * ck, cs = ci * sk, sc = si * ck, ss = si * sk;
      par[0] = (T)(cj * cc + sj * ss);
      par[3] = (T)(cj * sc - sj * cs);
      par[2] = (T)(-(cj * ss + sj * cc));
      par[1] = (T)(cj * cs - sj * sc);
    } break;
    default:
      par[0] = (T)(-0.15 + u01(r[0]) * 0.3);
      par[1] = (T)(-0.15 + u01(r[1]) * 0.45);
      par[3] = (T)(-0.2 + u01(r[2]) * 0.4);
      par[4] = (T)(-0.2 + u01(r[3]) * 0.4);
      par[5] = (T)(0.15 + u01(r[4]) * 0.2);
  }
}

// task frames cached after every forward (the reference reads data.site_xpos / body_xpos, which mujoco-py leaves at the
// values of the LAST mj_forward, i.e. one substep behind qpos)
template <class CX> DI void cache_task_frames(CX& c, typename CX::real touch) {
  using T = typename CX::real;
  const DModel<T>& m = c.m;
  const int* ti = m.task_int;
  const T *xpos = c.template S<O_XPOS>(), *xquat = c.template S<O_XQUAT>(), *spos = c.template G<G_SPOS>();
  T* tf = c.template S<O_TF>();
  if (c.lane == 0) {
    auto cp = [&](int o, const T* src) { tf[o] = src[0]; tf[o + 1] = src[1]; tf[o + 2] = src[2]; };
    switch (m.task) {
      case T_HAMMER:
        cp(0, spos + 3 * ti[1]); cp(3, xpos + 3 * ti[2]);
        for (int i = 0; i < 4; i++) tf[6 + i] = xquat[4 * ti[2] + i];
        cp(10, spos + 3 * ti[0]); cp(13, spos + 3 * ti[4]); cp(16, spos + 3 * ti[3]);
        tf[19] = touch;
        break;
      case T_DOOR: cp(0, spos + 3 * ti[1]); cp(3, spos + 3 * ti[2]); break;
      case T_PEN:
        cp(0, xpos + 3 * ti[2]); cp(3, spos + 3 * ti[3]); cp(6, spos + 3 * ti[4]); cp(9, spos + 3 * ti[5]);
        cp(12, spos + 3 * ti[6]); cp(15, spos + 3 * ti[7]);
        break;
      default: cp(0, xpos + 3 * ti[2]); cp(3, spos + 3 * ti[1]); cp(6, spos + 3 * ti[0]);
    }
  }
  __syncwarp();
}

template <typename T> DI T dist3(const T* a, const T* b) {
  T x = a[0] - b[0], y = a[1] - b[1], z = a[2] - b[2];
  return sqrt(x * x + y * y + z * z);
}
template <typename T> DI T clip1(T x) { return x < -1 ? T(-1) : (x > 1 ? T(1) : x); }

// utils/quatmath.py:136-164 (quat2mat) then :79-96 (mat2euler)
template <typename T> DI void quat2euler(const T* q, T* e) {
  T w = q[0], x = q[1], y = q[2], z = q[3], Nq = w * w + x * x + y * y + z * z, M[9];
  if (Nq > T(2.220446049250313e-16)) {
    T s = T(2) / Nq, X = x * s, Y = y * s, Z = z * s;
    T wX = w * X, wY = w * Y, wZ = w * Z, xX = x * X, xY = x * Y, xZ = x * Z, yY = y * Y, yZ = y * Z, zZ = z * Z;
    M[0] = 1 - (yY + zZ); M[1] = xY - wZ; M[2] = xZ + wY;
    M[3] = xY + wZ; M[4] = 1 - (xX + zZ); M[5] = yZ - wX;
    M[6] = xZ - wY; M[7] = yZ + wX; M[8] = 1 - (xX + yY);
  } else {
    for (int i = 0; i < 9; i++) M[i] = (i % 4 == 0) ? T(1) : T(0);
  }
  T cy = sqrt(M[8] * M[8] + M[5] * M[5]);
  if (cy > T(4 * 2.220446049250313e-16)) {
    e[2] = -atan2(M[1], M[0]);
    e[1] = -atan2(-M[2], cy);
    e[0] = -atan2(M[5], M[8]);
  } else {
    e[2] = -atan2(-M[3], M[4]);
    e[1] = -atan2(-M[2], cy);
    e[0] = 0;
  }
}

// get_obs: hammer_v0.py:92-104, door_v0.py:87-101, pen_v0.py:104-113, relocate_v0.py:74-83 (layouts: SURVEY.md App. A.2)
template <typename T> DI void write_obs(const DModel<T>& m, const T* qpos, const T* qvel, const T* tf, float* o, int lane) {
  switch (m.task) {
    case T_HAMMER:
      for (int i = lane; i < 27; i += 32) o[i] = (float)qpos[i];
      if (lane < 6) o[27 + lane] = (float)clip1(qvel[27 + lane]);
      if (lane < 3) { o[33 + lane] = (float)tf[lane]; o[36 + lane] = (float)tf[3 + lane]; o[42 + lane] = (float)tf[10 + lane]; }
      if (lane == 0) {
        T e[3];
        quat2euler(tf + 6, e);
        o[39] = (float)e[0]; o[40] = (float)e[1]; o[41] = (float)e[2];
        o[45] = (float)clip1(tf[19]);
      }
      break;
    case T_DOOR:
      for (int i = lane; i < 27; i += 32) o[i] = (float)qpos[1 + i];
      if (lane == 0) { o[27] = (float)qpos[29]; o[28] = (float)qpos[m.task_int[0]]; o[38] = qpos[m.task_int[0]] > T(1.0) ? 1.0f : -1.0f; }
      if (lane < 3) { o[29 + lane] = (float)tf[lane]; o[32 + lane] = (float)tf[3 + lane]; o[35 + lane] = (float)(tf[lane] - tf[3 + lane]); }
      break;
    case T_PEN:
      for (int i = lane; i < 24; i += 32) o[i] = (float)qpos[i];
      if (lane < 6) o[27 + lane] = (float)qvel[24 + lane];
      if (lane < 3) {
        T oo = (tf[6 + lane] - tf[9 + lane]) / m.pen_len, dd = (tf[12 + lane] - tf[15 + lane]) / m.tar_len;
        o[24 + lane] = (float)tf[lane];
        o[33 + lane] = (float)oo;
        o[36 + lane] = (float)dd;
        o[39 + lane] = (float)(tf[lane] - tf[3 + lane]);
        o[42 + lane] = (float)(oo - dd);
      }
      break;
    default:
      for (int i = lane; i < 30; i += 32) o[i] = (float)qpos[i];
      if (lane < 3) {
        o[30 + lane] = (float)(tf[3 + lane] - tf[lane]);
        o[33 + lane] = (float)(tf[3 + lane] - tf[6 + lane]);
        o[36 + lane] = (float)(tf[lane] - tf[6 + lane]);
      }
  }
}

// reward / goal_achieved / done: hammer_v0.py:62-90, door_v0.py:63-85, pen_v0.py:75-102, relocate_v0.py:54-72
template <typename T> DI void task_reward(const DModel<T>& m, const T* qpos, const T* qvel, const T* tf, int lane, T& reward, int& goal, int& done) {
  T v2 = 0;
  _Pragma("unroll 1") for (int j = lane; j < m.nv; j += 32) v2 += qvel[j] * qvel[j];
  v2 = warp_sum(v2);
  T r = 0;
  done = 0;
  switch (m.task) {
    case T_HAMMER: {
      const T *palm = tf, *obj = tf + 3, *target = tf + 10, *goalp = tf + 13, *tool = tf + 16;
      T dg = dist3(target, goalp);
      r = -T(0.1) * dist3(palm, obj);
      r -= dist3(tool, target);
      r -= 10 * dg;
      r -= T(1e-2) * sqrt(v2);
      if (obj[2] > T(0.04) && tool[2] > T(0.04)) r += 2;
      if (dg < T(0.020)) r += 25;
      if (dg < T(0.010)) r += 75;
      goal = dg < T(0.010);
    } break;
    case T_DOOR: {
      T door = qpos[m.task_int[0]];
      r = -T(0.1) * dist3(tf, tf + 3);
      r += -T(0.1) * (door - T(1.57)) * (door - T(1.57));
      r += -T(1e-5) * v2;
      if (door > T(0.2)) r += 2;
      if (door > T(1.0)) r += 8;
      if (door > T(1.35)) r += 10;
      goal = door >= T(1.35);
    } break;
    case T_PEN: {
      const T *obj = tf, *eps = tf + 3;
      T sim = 0;
      for (int i = 0; i < 3; i++) sim += (tf[6 + i] - tf[9 + i]) / m.pen_len * ((tf[12 + i] - tf[15 + i]) / m.tar_len);
      T dist = dist3(obj, eps);
      r = -dist + sim;
      if (dist < T(0.075) && sim > T(0.9)) r += 10;
      if (dist < T(0.075) && sim > T(0.95)) r += 50;
      if (obj[2] < T(0.075)) { r -= 5; done = 1; }
      goal = dist < T(0.075) && sim > T(0.95);
    } break;
    default: {
      const T *obj = tf, *palm = tf + 3, *tgt = tf + 6;
      T dot = dist3(obj, tgt);
      r = -T(0.1) * dist3(palm, obj);
      if (obj[2] > T(0.04)) { r += 1; r += -T(0.5) * dist3(palm, tgt); r += -T(0.5) * dot; }
      if (dot < T(0.1)) r += 10;
      if (dot < T(0.05)) r += 20;
      goal = dot < T(0.1);
    }
  }
  reward = r;
}

}  // namespace adroit
