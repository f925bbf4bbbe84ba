"""TEST INFRASTRUCTURE ONLY: plain numpy restatement of the headless ray-caster (csrc/go1_render.cuh) for single pixels.

There is no reference renderer to compare with -- the reference renders through MuJoCo's OpenGL path (train.py:53-61), which is
unavailable on this box, so pixel parity with the reference is UNPINNED.  This restatement pins the CUDA kernel to an independent
statement of the same geometry and shading: camera `tracking` (go1_torque.xml:84, trackcom), one primary ray per pixel against
the model's collision primitives and the checker floor, headlight + directional light, gradient sky."""
from __future__ import annotations

import numpy as np

SPHERE, CAPSULE, CYLINDER, BOX = 2, 3, 5, 6


def _q2m(q):
    w, x, y, z = q
    return np.array([[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def _axis_rot(u, a):
    c, s = np.cos(a), np.sin(a)
    K = np.array([[0, -u[2], u[1]], [u[2], 0, -u[0]], [-u[1], u[0], 0]])
    return c * np.eye(3) + s * K + (1 - c) * np.outer(u, u)


def world_geoms(model, qpos):
    """Forward kinematics (mj_kinematics) -> list of (type, centre, R, size, body) for the collision geoms, robot COM."""
    m = model
    nb = m.nbody
    R = [np.eye(3)] * nb; p = [np.zeros(3)] * nb
    q = np.asarray(qpos[3:7], float); q = q / np.linalg.norm(q)
    R[1] = _q2m(q); p[1] = np.asarray(qpos[:3], float)
    for b in range(2, nb):
        par = int(m.body_parentid[b]); j = int(m.body_jntadr[b])
        p[b] = p[par] + R[par] @ m.body_pos[b]
        ang = qpos[int(m.jnt_qposadr[j])] - m.qpos0[int(m.jnt_qposadr[j])]
        R[b] = R[par] @ _q2m(m.body_quat[b]) @ _axis_rot(np.asarray(m.jnt_axis[j], float), ang)
    com = sum(m.body_mass[b] * (p[b] + R[b] @ m.body_ipos[b]) for b in range(1, nb)) / sum(m.body_mass[1:])
    prims = []
    for g in range(m.ngeom):
        b = int(m.geom_bodyid[g])
        if b == 0 or int(m.geom_type[g]) not in (SPHERE, CAPSULE, CYLINDER, BOX) or not (m.geom_contype[g] or m.geom_conaffinity[g]):
            continue
        prims.append((int(m.geom_type[g]), p[b] + R[b] @ m.geom_pos[g], R[b] @ _q2m(m.geom_quat[g] / np.linalg.norm(m.geom_quat[g])),
                      np.asarray(m.geom_size[g], float), b, int(m.geom_condim[g])))
    return prims, com, p, R


def _hit(o, d, prim):
    t_, c, Rg, s, _, _ = prim
    z, x = Rg[:, 2], Rg[:, 0]
    oc = o - c
    best, nbest = -1.0, None

    def sphere(cc, r):
        o2 = o - cc
        b = o2 @ d; disc = b * b - (o2 @ o2 - r * r)
        if disc < 0:
            return -1.0, None
        t = -b - np.sqrt(disc)
        return (t, (o2 + t * d) / r) if t > 1e-4 else (-1.0, None)
    if t_ == SPHERE:
        return sphere(c, s[0])
    if t_ in (CAPSULE, CYLINDER):
        r, h = s[0], s[1]
        dz, oz = d @ z, oc @ z
        a, b, cc = 1 - dz * dz, oc @ d - oz * dz, oc @ oc - oz * oz - r * r
        if a > 1e-9 and b * b - a * cc >= 0:
            t = (-b - np.sqrt(b * b - a * cc)) / a
            zt = oz + t * dz
            if t > 1e-4 and abs(zt) <= h:
                best, nbest = t, (oc + t * d - zt * z) / r
        for sg in (-1, 1):
            if t_ == CAPSULE:
                t, n = sphere(c + sg * h * z, r)
                if t > 0 and (best < 0 or t < best):
                    best, nbest = t, n
            elif abs(dz) > 1e-9:
                t = (sg * h - oz) / dz
                if t > 1e-4 and (best < 0 or t < best):
                    qv = oc + t * d
                    if qv @ qv - (qv @ z) ** 2 <= r * r:
                        best, nbest = t, sg * z
        return best, nbest
    y = np.cross(z, x)
    ol = np.array([oc @ x, oc @ y, oc @ z]); dl = np.array([d @ x, d @ y, d @ z])
    t0, t1, ax, sg = -1e30, 1e30, 0, 1.0
    for i in range(3):
        if abs(dl[i]) < 1e-9:
            if abs(ol[i]) > s[i]:
                return -1.0, None
            continue
        ta, tb, sgn = (-s[i] - ol[i]) / dl[i], (s[i] - ol[i]) / dl[i], -1.0
        if ta > tb:
            ta, tb, sgn = tb, ta, 1.0
        if ta > t0:
            t0, ax, sg = ta, i, sgn
        t1 = min(t1, tb)
    if t0 > t1 or t0 <= 1e-4:
        return -1.0, None
    return t0, sg * (x, y, z)[ax]


def pixel(model, qpos, px, py, width=1024, height=768, cam_pos=(0.0, -1.0, 0.8), xaxis=(1, 0, 0), yaxis=(0, 1, 1), fovy=45.0):
    """RGB (uint8 triple) of one pixel; mirrors k_render step by step."""
    prims, com, p, R = world_geoms(model, qpos)
    # trackcom offset: camera position minus the subtree COM of the trunk at qpos0
    _, com0, p0, _ = world_geoms(model, model.qpos0)
    off = np.asarray(cam_pos, float) - (com0 - p0[1])
    x = np.asarray(xaxis, float); x /= np.linalg.norm(x)
    y = np.asarray(yaxis, float); y = y - (x @ y) * x; y /= np.linalg.norm(y)
    z = np.cross(x, y)
    th = np.tan(0.5 * np.deg2rad(fovy))
    u = (2 * (px + 0.5) / width - 1) * th * width / height
    v = (1 - 2 * (py + 0.5) / height) * th
    d = u * x + v * y - z; d /= np.linalg.norm(d)
    o = com + off
    tbest, nb, col = 1e30, None, None
    for pr in prims:
        t, n = _hit(o, d, pr)
        if t > 0 and t < tbest:
            b = pr[4] - 1                                  # kernel body slot: 0 trunk, 1 + 3 leg + link
            shade = 0.62 if b == 0 else ((0.12 if pr[0] == SPHERE else 0.35) if (b - 1) % 3 == 2 else 0.48)
            tbest, nb, col = t, n, np.array([shade, shade, shade * 1.05])
    if d[2] < -1e-6:
        t = -o[2] / d[2]
        if 1e-4 < t < tbest:
            fx, fy = (o[0] + t * d[0]) * 5, (o[1] + t * d[1]) * 5
            chk = (int(np.floor(fx * 2)) + int(np.floor(fy * 2))) & 1
            tbest, nb, col = t, np.array([0, 0, 1.0]), np.array([0.2, 0.3, 0.4]) if chk else np.array([0.1, 0.2, 0.3])
    if nb is None:
        w = min(1.0, max(0.0, 0.5 + d[2]))
        rgb = np.array([0.3, 0.5, 0.7]) * (1 - 0.8 * w)
    else:
        head, sun, fog = max(0.0, -(nb @ d)), max(0.0, nb[2]), np.exp(-0.02 * tbest)
        rgb = col * (0.3 + 0.6 * head + 0.7 * sun) * fog + 0.35 * np.array([0.3, 0.5, 0.7]) * (1 - fog)
    return (np.clip(rgb, 0, 1) * 255 + 0.5).astype(np.uint8)
