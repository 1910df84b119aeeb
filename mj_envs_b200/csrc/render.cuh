// render.cuh — batched 64x64 pixel observations by ray casting.
//
// Stands in for the reference's GL path  HeadlessObserver.render -> sim.render(640, 480) -> vertical flip ->
// CenterCrop(128) (headless_observer.py:34-52)  followed by PlaNet's 128 -> 64 bilinear resize (cv2 INTER_LINEAR at an
// exact factor of 2 = 2x2 box average; train.py:122).  No graphics stack exists on this box: the kernel casts the four
// 128-resolution rays of every output pixel through the reference's free camera (azimuth 90, distance 4.5, elevation
// rule of headless_observer.py:59-67, fovy 45 over a 640x480 frame) against the scene's analytic primitives (plane,
// sphere, capsule, cylinder, box; the STL meshes and texture files are absent, the collision shapes stand in for them),
// shades with MuJoCo's default headlight (ambient 0.1, diffuse 0.4) plus the scene's positional light, alpha-blends
// translucent sites, averages the four samples and writes u8 RGB.  Body frames come from the context block (the frames
// of the last mj_forward, which is what sim.render draws).  One block per environment, primitives staged in shared memory.
#pragma once
#include "task.cuh"

namespace adroit {

constexpr int MAXRG = 48;          // render primitives per scene (Adroit scenes: 32..38)
constexpr int RENDER_THREADS = 256;
constexpr int IMG = 64;            // output resolution
constexpr int CROP = 128, FRAME_W = 640, FRAME_H = 480;

struct RPrim {
  float pos[3], mat[9], size[3], rgba[4];
  float ol[3];      // camera position in the primitive's frame
  float vc[3];      // centre - camera (world)
  float rb2, vv;    // bounding-sphere radius^2, |vc|^2
  int type;
};

// nearest intersection t > tmin of the ray o + t d (primitive frame) with the primitive; local normal in n
DI bool hit_sphere(const float* o, const float* d, float r, float cz, float& t, float* n) {
  float oz = o[2] - cz, b = o[0] * d[0] + o[1] * d[1] + oz * d[2], c = o[0] * o[0] + o[1] * o[1] + oz * oz - r * r, disc = b * b - c;
  if (disc < 0) return false;
  float tt = -b - __fsqrt_rn(disc);
  if (tt <= 1e-4f) return false;
  t = tt;
  float inv = __fdiv_rn(1.0f, r);
  n[0] = (o[0] + tt * d[0]) * inv; n[1] = (o[1] + tt * d[1]) * inv; n[2] = (oz + tt * d[2]) * inv;
  return true;
}
DI bool hit_prim(int type, const float* size, const float* o, const float* d, float& t, float* n) {
  bool found = false;
  t = 3.0e38f;
  if (type == G_PLANE) {
    if (fabsf(d[2]) < 1e-9f) return false;
    float tt = __fdiv_rn(-o[2], d[2]);
    if (tt <= 1e-4f) return false;
    float x = o[0] + tt * d[0], y = o[1] + tt * d[1];
    if ((size[0] > 0 && fabsf(x) > size[0]) || (size[1] > 0 && fabsf(y) > size[1])) return false;
    t = tt; n[0] = 0; n[1] = 0; n[2] = 1;
    return true;
  }
  if (type == G_SPHERE) return hit_sphere(o, d, size[0], 0.0f, t, n);
  if (type == G_BOX) {
    float tn = -3.0e38f, tf = 3.0e38f;
    int ax = 0;
    float sg = 1;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      if (fabsf(d[k]) < 1e-12f) {
        if (fabsf(o[k]) > size[k]) return false;
      } else {
        float inv = __fdiv_rn(1.0f, d[k]), t0 = (-size[k] - o[k]) * inv, t1 = (size[k] - o[k]) * inv, s = -1;
        if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; s = 1; }
        if (t0 > tn) { tn = t0; ax = k; sg = s; }
        tf = fminf(tf, t1);
      }
    }
    if (tn > tf || tn <= 1e-4f) return false;
    t = tn;
    n[0] = n[1] = n[2] = 0;
    n[ax] = sg;
    return true;
  }
  // capsule / cylinder: axis z, radius size[0], half length size[1]
  float r = size[0], h = size[1];
  float a = d[0] * d[0] + d[1] * d[1], b = o[0] * d[0] + o[1] * d[1], c = o[0] * o[0] + o[1] * o[1] - r * r;
  if (a > 1e-12f) {
    float disc = b * b - a * c;
    if (disc < 0) return false;   // misses the infinite cylinder: misses caps of either kind too
    float tt = __fdiv_rn(-b - __fsqrt_rn(disc), a), z = o[2] + tt * d[2];
    if (tt > 1e-4f && fabsf(z) <= h) {
      t = tt; found = true;
      float inv = __fdiv_rn(1.0f, r);
      n[0] = (o[0] + tt * d[0]) * inv; n[1] = (o[1] + tt * d[1]) * inv; n[2] = 0;
    }
  } else if (c > 0) return false;
  if (type == G_CAPSULE) {
#pragma unroll
    for (int s = -1; s <= 1; s += 2) {
      float tt, nn[3];
      if (hit_sphere(o, d, r, s * h, tt, nn) && tt < t && (o[2] + tt * d[2]) * s > h) { t = tt; n[0] = nn[0]; n[1] = nn[1]; n[2] = nn[2]; found = true; }
    }
  } else if (fabsf(d[2]) > 1e-12f) {
#pragma unroll
    for (int s = -1; s <= 1; s += 2) {
      float tt = __fdiv_rn(s * h - o[2], d[2]), x = o[0] + tt * d[0], y = o[1] + tt * d[1];
      if (tt > 1e-4f && tt < t && x * x + y * y <= r * r) { t = tt; n[0] = 0; n[1] = 0; n[2] = (float)s; found = true; }
    }
  }
  return found;
}

DI float prim_rbound(int type, const float* s) {
  if (type == G_SPHERE) return s[0];
  if (type == G_CAPSULE) return s[0] + s[1];
  if (type == G_CYLINDER) return sqrtf(s[0] * s[0] + s[1] * s[1]);
  if (type == G_BOX) return sqrtf(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
  return -1.0f;   // plane: never culled
}

struct RShade { float cam[3], light[3], ldiff; };
DI void shade(const RPrim& P, const RShade& S, const float* d, float t, const float* nl, float* rgb) {
  const float* R = P.mat;
  float n[3] = {R[0] * nl[0] + R[1] * nl[1] + R[2] * nl[2], R[3] * nl[0] + R[4] * nl[1] + R[5] * nl[2], R[6] * nl[0] + R[7] * nl[1] + R[8] * nl[2]};
  float nd = n[0] * d[0] + n[1] * d[1] + n[2] * d[2];
  if (nd > 0) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; nd = -nd; }
  float hp[3] = {S.cam[0] + t * d[0], S.cam[1] + t * d[1], S.cam[2] + t * d[2]};
  float l[3] = {S.light[0] - hp[0], S.light[1] - hp[1], S.light[2] - hp[2]};
  float ll = rsqrtf(l[0] * l[0] + l[1] * l[1] + l[2] * l[2] + 1e-20f);
  float nl_ = fmaxf(0.0f, (n[0] * l[0] + n[1] * l[1] + n[2] * l[2]) * ll);
  float I = 0.1f + 0.4f * (-nd) + S.ldiff * nl_;   // headlight ambient + headlight diffuse (light at the camera) + scene light
  for (int k = 0; k < 3; k++) rgb[k] = fminf(1.0f, P.rgba[k] * I);
}

// rgb [n][64][64][3] u8
template <typename T>
__global__ void __launch_bounds__(RENDER_THREADS) render_kernel(const __grid_constant__ DModel<T> m, const __grid_constant__ DState<T> st,
                                                                 unsigned char* __restrict__ rgb) {
  __shared__ RPrim prim[MAXRG];
  __shared__ unsigned int tile[IMG * IMG * 3 / 4];
  const int env = blockIdx.x, tid = threadIdx.x;
  const T* ctx = st.ctx + (size_t)env * CTX_T;
  const T *xpos = ctx + O_XPOS, *xmat = ctx + O_XMAT, *par = ctx + O_PAR;
  // camera (mjcf.py::_set_render_camera): MuJoCo free camera = lookat, azimuth, elevation, distance
  const float DEG = 0.017453292519943295f;
  const float az = (float)m.render_cam[3] * DEG, el = (float)m.render_cam[4] * DEG, dist = (float)m.render_cam[5];
  const float fwd[3] = {cosf(el) * cosf(az), cosf(el) * sinf(az), sinf(el)};
  const float up[3] = {-sinf(el) * cosf(az), -sinf(el) * sinf(az), cosf(el)};
  const float right[3] = {fwd[1] * up[2] - fwd[2] * up[1], fwd[2] * up[0] - fwd[0] * up[2], fwd[0] * up[1] - fwd[1] * up[0]};
  RShade S;
  for (int k = 0; k < 3; k++) S.cam[k] = (float)m.render_cam[k] - dist * fwd[k];
  {
    int lb = m.light_body;
    for (int k = 0; k < 3; k++)
      S.light[k] = (float)(xpos[3 * lb + k] + xmat[9 * lb + 3 * k] * m.light_pos[0] + xmat[9 * lb + 3 * k + 1] * m.light_pos[1] + xmat[9 * lb + 3 * k + 2] * m.light_pos[2]);
    S.ldiff = (float)m.light_diffuse[0];
  }
  const int nr = m.nrender;
  if (tid < nr) {
    RPrim& P = prim[tid];
    const int* ri = m.render_int + 4 * tid;
    const T* rr = m.render_real + 19 * tid;
    int b = ri[1], po = ri[2];
    float lp[3], sz[3], col[4];
    for (int k = 0; k < 3; k++) { lp[k] = (float)(po >= 0 ? par[po + k] : rr[k]); sz[k] = (float)rr[12 + k]; }
    for (int k = 0; k < 4; k++) col[k] = (float)rr[15 + k];
    // hammer variations are visible too (hammer_v0.py:115-125): head colour / head + neck position / head size
    int vt = m.task_int[8];
    if (vt == VAR_MASS && tid == m.task_int[12]) col[0] = (float)par[3] / 2.5f;
    if (vt == VAR_POS && tid == m.task_int[12]) lp[0] = (float)par[3];
    if (vt == VAR_POS && tid == m.task_int[13]) lp[0] = (float)par[4];
    if (vt == VAR_SIZE && tid == m.task_int[12]) { sz[0] = (float)par[3]; sz[1] = (float)par[4]; }
    const T* X = xmat + 9 * b;
    for (int i = 0; i < 3; i++) {
      P.pos[i] = (float)xpos[3 * b + i] + (float)X[3 * i] * lp[0] + (float)X[3 * i + 1] * lp[1] + (float)X[3 * i + 2] * lp[2];
      for (int j = 0; j < 3; j++)
        P.mat[3 * i + j] = (float)(X[3 * i] * rr[3 + j] + X[3 * i + 1] * rr[6 + j] + X[3 * i + 2] * rr[9 + j]);
    }
    for (int k = 0; k < 3; k++) { P.size[k] = sz[k]; P.vc[k] = P.pos[k] - S.cam[k]; }
    for (int k = 0; k < 4; k++) P.rgba[k] = col[k];
    for (int k = 0; k < 3; k++) P.ol[k] = -(P.mat[k] * P.vc[0] + P.mat[3 + k] * P.vc[1] + P.mat[6 + k] * P.vc[2]);
    P.type = ri[0];
    float rb = prim_rbound(P.type, sz);
    P.rb2 = rb < 0 ? -1.0f : rb * rb;
    P.vv = P.vc[0] * P.vc[0] + P.vc[1] * P.vc[1] + P.vc[2] * P.vc[2];
  }
  __syncthreads();
  const float th = tanf(0.5f * (float)m.render_cam[6] * DEG), aspect = (float)FRAME_W / (float)FRAME_H;
  const int r0 = (FRAME_H - CROP) / 2, c0 = (FRAME_W - CROP) / 2;
  unsigned char* tb = reinterpret_cast<unsigned char*>(tile);
  auto ray_dir = [&](float Rf, float Cf, float* d) {   // unit direction through frame position (row Rf, column Cf), pixel units
    float y = (1.0f - 2.0f * Rf / FRAME_H) * th, x = (2.0f * Cf / FRAME_W - 1.0f) * th * aspect;
    d[0] = fwd[0] + x * right[0] + y * up[0]; d[1] = fwd[1] + x * right[1] + y * up[1]; d[2] = fwd[2] + x * right[2] + y * up[2];
    float inv = rsqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    d[0] *= inv; d[1] *= inv; d[2] *= inv;
  };
  // Each warp walks 8x4-pixel tiles.  Per tile the lanes first test every primitive's bounding sphere against the tile's
  // bounding cone (one primitive per lane, ballot -> two 32-bit masks), then every lane casts the 4 sub-rays of its pixel
  // against the survivors only (typically ground, wall, table and a handful of hand / object shapes instead of ~38).
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int TW = 8, TH = 4, TILES_X = IMG / TW, NTILES = (IMG / TW) * (IMG / TH);
  for (int tl = warp; tl < NTILES; tl += RENDER_THREADS / 32) {
    const int ty0 = (tl / TILES_X) * TH, tx0 = (tl % TILES_X) * TW;
    float dc[3], dk[3];
    ray_dir(r0 + 2.0f * ty0 + TH, c0 + 2.0f * tx0 + TW, dc);          // tile centre
    ray_dir(r0 + 2.0f * ty0, c0 + 2.0f * tx0, dk);                    // a corner (the tile is symmetric about its centre to first order)
    float cost = dc[0] * dk[0] + dc[1] * dk[1] + dc[2] * dk[2];
    ray_dir(r0 + 2.0f * ty0 + 2 * TH, c0 + 2.0f * tx0, dk);
    cost = fminf(cost, dc[0] * dk[0] + dc[1] * dk[1] + dc[2] * dk[2]);
    ray_dir(r0 + 2.0f * ty0, c0 + 2.0f * tx0 + 2 * TW, dk);
    cost = fminf(cost, dc[0] * dk[0] + dc[1] * dk[1] + dc[2] * dk[2]);
    ray_dir(r0 + 2.0f * ty0 + 2 * TH, c0 + 2.0f * tx0 + 2 * TW, dk);
    cost = fminf(cost, dc[0] * dk[0] + dc[1] * dk[1] + dc[2] * dk[2]);
    cost = fmaxf(0.0f, cost - 1e-6f);
    const float sint = __fsqrt_rn(fmaxf(0.0f, 1.0f - cost * cost));
    unsigned keep[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      int g = lane + 32 * q;
      bool k = false;
      if (g < nr) {
        const RPrim& P = prim[g];
        if (P.rb2 < 0 || P.vv <= P.rb2) k = true;
        else {
          float invv = rsqrtf(P.vv), cosA = (P.vc[0] * dc[0] + P.vc[1] * dc[1] + P.vc[2] * dc[2]) * invv;
          float sins = fminf(1.0f, __fsqrt_rn(P.rb2) * invv * 1.0001f + 1e-6f), coss = __fsqrt_rn(fmaxf(0.0f, 1.0f - sins * sins));
          float cossum = cost * coss - sint * sins;                    // cos(tile half-angle + sphere half-angle)
          k = cossum <= 0.0f || cosA >= cossum;
        }
      }
      keep[q] = __ballot_sync(0xffffffffu, k);
    }
    const int I = ty0 + (lane >> 3), J = tx0 + (lane & 7), p = I * IMG + J;
    float acc[3] = {0, 0, 0};
#pragma unroll 1
    for (int sub = 0; sub < 4; sub++) {
      float d[3];
      ray_dir(r0 + 2 * I + (sub >> 1) + 0.5f, c0 + 2 * J + (sub & 1) + 0.5f, d);
      float tbest = 3.0e38f, nbest[3] = {0, 0, 1};
      int ibest = -1;
      float col[3] = {0, 0, 0};                           // background: no skybox texture in these scenes
#pragma unroll 1
      for (int pass = 0; pass < 2; pass++) {              // pass 0: opaque, nearest hit; pass 1: translucent sites blended over it
        if (pass == 1 && ibest >= 0) shade(prim[ibest], S, d, tbest, nbest, col);
#pragma unroll 1
        for (int q = 0; q < 2; q++) {
          unsigned mk = keep[q];
          while (mk) {
            int g = 32 * q + __ffs(mk) - 1;
            mk &= mk - 1;
            const RPrim& P = prim[g];
            if ((P.rgba[3] < 0.999f) != (pass == 1)) continue;
            if (P.rb2 >= 0) {                             // per-ray bounding-sphere test (all rays start at the camera)
              float vd = P.vc[0] * d[0] + P.vc[1] * d[1] + P.vc[2] * d[2];
              if (P.vv - vd * vd > P.rb2 || (vd < 0 && P.vv > P.rb2)) continue;
            }
            float dl[3] = {P.mat[0] * d[0] + P.mat[3] * d[1] + P.mat[6] * d[2], P.mat[1] * d[0] + P.mat[4] * d[1] + P.mat[7] * d[2],
                           P.mat[2] * d[0] + P.mat[5] * d[1] + P.mat[8] * d[2]};
            float t, nl[3];
            if (!hit_prim(P.type, P.size, P.ol, dl, t, nl) || t >= tbest) continue;
            if (pass == 0) { tbest = t; ibest = g; nbest[0] = nl[0]; nbest[1] = nl[1]; nbest[2] = nl[2]; }
            else {
              float c2[3], a = P.rgba[3];
              shade(P, S, d, t, nl, c2);
              for (int k = 0; k < 3; k++) col[k] = a * c2[k] + (1.0f - a) * col[k];
            }
          }
        }
      }
      for (int k = 0; k < 3; k++) acc[k] += col[k];
    }
    for (int k = 0; k < 3; k++) tb[3 * p + k] = (unsigned char)(int)(acc[k] * (255.0f * 0.25f) + 0.5f);
  }
  __syncthreads();
  unsigned int* out = reinterpret_cast<unsigned int*>(rgb + (size_t)env * IMG * IMG * 3);
  for (int w = tid; w < IMG * IMG * 3 / 4; w += RENDER_THREADS) out[w] = tile[w];
}

}  // namespace adroit
