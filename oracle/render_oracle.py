"""CPU restatement (numpy, fp64) of the pixel path of mj_envs_b200/csrc/render.cuh.

TEST INFRASTRUCTURE (only tests/ may import it).  PARITY UNPINNED against the reference's images: the reference renders
through MuJoCo's OpenGL path (hand_manipulation_suite/headless_observer.py:34-52), which cannot run on this box, and it
holds no golden images.  This file restates, independently of the CUDA kernel, the same stand-in:

  camera   free camera of headless_observer.py:20-31,59-67 (azimuth 90, distance 4.5, elevation rule, lookat = median of
           the geom positions [EXT mujoco-py]); constants come from mj_envs_b200/mjcf.py::_set_render_camera
  frame    640x480, fovy 45; vertical flip; CenterCrop(128) (headless_observer.py:38-44); 2x2 box average to 64x64
           (= cv2.resize INTER_LINEAR at factor 2, PlaNet's _images_to_observation, utils/train.py:122)
  scene    analytic primitives (plane / sphere / capsule / cylinder / box) of the render table; meshes and textures are
           absent from the reference tree, so collision shapes stand in; translucent sites are alpha-blended
  shading  MuJoCo default headlight (ambient 0.1, diffuse 0.4) + the scene's positional light (diffuse), no shadows
"""
import numpy as np

G_PLANE, G_SPHERE, G_CAPSULE, G_CYLINDER, G_BOX = 0, 2, 3, 5, 6
IMG, CROP, FW, FH = 64, 128, 640, 480


def _quat2mat(q):
    w, x, y, z = q
    return np.array([[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def _sphere(o, d, r, cz):
    oz = o[2] - cz
    b = o[0] * d[0] + o[1] * d[1] + oz * d[2]
    c = o[0] ** 2 + o[1] ** 2 + oz ** 2 - r * r
    disc = b * b - c
    ok = disc >= 0
    t = -b - np.sqrt(np.where(ok, disc, 0))
    ok &= t > 1e-4
    n = np.stack([(o[0] + t * d[0]) / r, (o[1] + t * d[1]) / r, (oz + t * d[2]) / r])
    return ok, t, n


def _hit(ty, size, o, d):
    """o: (3,) origin in the primitive frame; d: (3, R) directions.  Returns ok (R,), t (R,), normal (3, R)."""
    R = d.shape[1]
    inf = np.full(R, np.inf)
    if ty == G_PLANE:
        with np.errstate(divide="ignore", invalid="ignore"):
            t = -o[2] / d[2]
        ok = (np.abs(d[2]) >= 1e-9) & (t > 1e-4)
        x, y = o[0] + t * d[0], o[1] + t * d[1]
        if size[0] > 0:
            ok &= np.abs(x) <= size[0]
        if size[1] > 0:
            ok &= np.abs(y) <= size[1]
        n = np.zeros((3, R)); n[2] = 1
        return ok, np.where(ok, t, inf), n
    if ty == G_SPHERE:
        ok, t, n = _sphere(o, d, size[0], 0.0)
        return ok, np.where(ok, t, inf), n
    if ty == G_BOX:
        tn, tf = np.full(R, -np.inf), np.full(R, np.inf)
        ax, sg = np.zeros(R, dtype=int), np.ones(R)
        dead = np.zeros(R, dtype=bool)
        for k in range(3):
            par = np.abs(d[k]) < 1e-12
            dead |= par & (abs(o[k]) > size[k])
            with np.errstate(divide="ignore", invalid="ignore"):
                inv = 1.0 / d[k]
                t0, t1 = (-size[k] - o[k]) * inv, (size[k] - o[k]) * inv
            s = np.where(t0 > t1, 1.0, -1.0)
            t0, t1 = np.minimum(t0, t1), np.maximum(t0, t1)
            upd = (~par) & (t0 > tn)
            tn = np.where(upd, t0, tn); ax = np.where(upd, k, ax); sg = np.where(upd, s, sg)
            tf = np.where(par, tf, np.minimum(tf, t1))
        ok = (~dead) & (tn <= tf) & (tn > 1e-4)
        n = np.zeros((3, R)); n[ax, np.arange(R)] = sg
        return ok, np.where(ok, tn, inf), n
    r, h = size[0], size[1]
    a = d[0] ** 2 + d[1] ** 2
    b = o[0] * d[0] + o[1] * d[1]
    c = o[0] ** 2 + o[1] ** 2 - r * r
    side = a > 1e-12
    disc = b * b - a * c
    alive = np.where(side, disc >= 0, c <= 0)
    with np.errstate(divide="ignore", invalid="ignore"):
        tt = (-b - np.sqrt(np.where(disc >= 0, disc, 0))) / a
    z = o[2] + tt * d[2]
    ok = alive & side & (tt > 1e-4) & (np.abs(z) <= h)
    t = np.where(ok, tt, inf)
    n = np.stack([(o[0] + tt * d[0]) / r, (o[1] + tt * d[1]) / r, np.zeros(R)])
    n = np.where(ok, n, 0.0)
    for s in (-1.0, 1.0):
        if ty == G_CAPSULE:
            ok2, t2, n2 = _sphere(o, d, r, s * h)
            ok2 &= alive & (t2 < t) & ((o[2] + t2 * d[2]) * s > h)
        else:
            with np.errstate(divide="ignore", invalid="ignore"):
                t2 = (s * h - o[2]) / d[2]
            x, y = o[0] + t2 * d[0], o[1] + t2 * d[1]
            ok2 = alive & (np.abs(d[2]) > 1e-12) & (t2 > 1e-4) & (t2 < t) & (x * x + y * y <= r * r)
            n2 = np.zeros((3, R)); n2[2] = s
        t = np.where(ok2, t2, t)
        n = np.where(ok2, n2, n)
        ok = ok | ok2
    return ok, t, n


def _shade(P, cam, light, ldiff, d, t, nl):
    n = P["mat"] @ nl
    nd = np.sum(n * d, axis=0)
    flip = nd > 0
    n = np.where(flip, -n, n)
    nd = np.where(flip, -nd, nd)
    hp = cam[:, None] + t * d
    l = light[:, None] - hp
    l = l / np.sqrt(np.sum(l * l, axis=0) + 1e-20)
    nl_ = np.maximum(0.0, np.sum(n * l, axis=0))
    I = 0.1 + 0.4 * (-nd) + ldiff * nl_
    return np.minimum(1.0, P["rgba"][:3, None] * I)


def render(model, xpos, xquat, par=None):
    """One env: body world positions [nbody,3] and quaternions [nbody,4] (frames of the last forward), per-env parameters.
    Returns u8 [64, 64, 3]."""
    A = model.A
    par = A["par0"] if par is None else np.asarray(par)
    ri, rr = A["render_int"], A["render_real"]
    cam_c = A["render_cam"]
    az, el, dist = np.deg2rad(cam_c[3]), np.deg2rad(cam_c[4]), cam_c[5]
    fwd = np.array([np.cos(el) * np.cos(az), np.cos(el) * np.sin(az), np.sin(el)])
    up = np.array([-np.sin(el) * np.cos(az), -np.sin(el) * np.sin(az), np.cos(el)])
    right = np.cross(fwd, up)
    cam = cam_c[:3] - dist * fwd
    xmat = [_quat2mat(q / np.linalg.norm(q)) for q in xquat]
    lb = int(A["light_body"][0])
    light = xpos[lb] + xmat[lb] @ A["light_pos"][0]
    ldiff = A["light_diffuse"][0][0]
    ti = A["task_int"]
    prims = []
    for k, ((ty, b, po, _), r) in enumerate(zip(ri, rr)):
        lp = par[po:po + 3].copy() if po >= 0 else r[:3].copy()
        size, rgba = r[12:15].copy(), r[15:19].copy()
        if ti[8] == 1 and k == ti[12]:
            rgba[0] = par[3] / 2.5
        if ti[8] == 2 and k == ti[12]:
            lp[0] = par[3]
        if ti[8] == 2 and k == ti[13]:
            lp[0] = par[4]
        if ti[8] == 3 and k == ti[12]:
            size[0], size[1] = par[3], par[4]
        mat = xmat[b] @ r[3:12].reshape(3, 3)
        pos = xpos[b] + xmat[b] @ lp
        prims.append(dict(type=int(ty), pos=pos, mat=mat, size=size, rgba=rgba, ol=mat.T @ (cam - pos)))
    th = np.tan(0.5 * np.deg2rad(cam_c[6]))
    r0, c0 = (FH - CROP) // 2, (FW - CROP) // 2
    R_, C_ = np.meshgrid(np.arange(CROP) + r0, np.arange(CROP) + c0, indexing="ij")
    y = (1.0 - 2.0 * (R_.ravel() + 0.5) / FH) * th
    x = (2.0 * (C_.ravel() + 0.5) / FW - 1.0) * th * (FW / FH)
    d = fwd[:, None] + x * right[:, None] + y * up[:, None]
    d = d / np.sqrt(np.sum(d * d, axis=0))
    n_r = d.shape[1]
    tbest = np.full(n_r, np.inf)
    col = np.zeros((3, n_r))
    for P in prims:
        if P["rgba"][3] < 0.999:
            continue
        ok, t, nl = _hit(P["type"], P["size"], P["ol"], P["mat"].T @ d)
        better = ok & (t < tbest)
        if better.any():
            c = _shade(P, cam, light, ldiff, d, np.where(better, t, 0.0), nl)
            col = np.where(better, c, col)
            tbest = np.where(better, t, tbest)
    for P in prims:
        if P["rgba"][3] >= 0.999:
            continue
        ok, t, nl = _hit(P["type"], P["size"], P["ol"], P["mat"].T @ d)
        front = ok & (t < tbest)
        if front.any():
            c = _shade(P, cam, light, ldiff, d, np.where(front, t, 0.0), nl)
            a = P["rgba"][3]
            col = np.where(front, a * c + (1 - a) * col, col)
    img = col.T.reshape(CROP, CROP, 3)
    img = img.reshape(IMG, 2, IMG, 2, 3).mean(axis=(1, 3))
    return np.floor(img * 255.0 + 0.5).astype(np.uint8)
