// go1_render.cuh -- headless CUDA ray-caster for render_mode="rgb_array" (row f4 of the scope table).
//
// The reference renders through gymnasium's MuJoCo OpenGL renderer (train.py:53-61: camera_name="tracking", 1024 x 768;
// training_callback.py:686-738 appends record_env.render() frames to an mp4).  That path is unavailable on this box; this
// kernel produces the same kind of frame -- an RGB image of the robot from camera `tracking` (go1_torque.xml:84: mode trackcom,
// pos "0 -1 0.8", xyaxes "1 0 0 0 1 1", default fovy 45) -- by casting one ray per pixel against the model's collision
// primitives (spheres, capsules, cylinders, boxes; the visual meshes are not in the repository: trunk.stl is a missing blob) and
// the checker floor (scene_torque.xml:14-21: builtin checker rgb1 .2 .3 .4 / rgb2 .1 .2 .3, texrepeat 5, texuniform), shaded with
// the scene's headlight (diffuse .6, ambient .3) and the directional light from above (scene_torque.xml:20), over the gradient
// skybox (rgb1 .3 .5 .7 -> rgb2 0 0 0).  No graphics API: plain arithmetic into a device buffer.
//
// One block = one 16 x 16 pixel tile of one environment's image; the block first runs forward kinematics for the environment
// (thread g < nflat places geom g in the world from the state row), then every thread traces its pixel.
#pragma once

namespace go1 {

struct RenderParams {
    const DevModel<float>* model;     // render always uses the fp32 view of the model
    const void* sreal;                // state slab (real = float or double)
    int real_is_double, s_stride, n_envs;
    const int* env_ids;               // [n_img] environments to render
    int n_img, width, height;
    float fovy_deg;
    float cam_off[3];                 // camera position relative to the robot's centre of mass (trackcom)
    float cam_x[3], cam_y[3], cam_z[3];   // camera axes in the world (the camera looks along -z)
    unsigned char* rgb;               // [n_img][height][width][3]
};

namespace rt {

struct Prim { int type; float c[3], z[3], x[3], s[3], col[3]; };

__device__ __forceinline__ float dot(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// ray o + t d (|d| = 1) against a sphere; returns the smallest t > tmin or -1
__device__ __forceinline__ float hit_sphere(const float* o, const float* d, const float* c, float r, float* n) {
    const float oc[3] = {o[0] - c[0], o[1] - c[1], o[2] - c[2]};
    const float b = dot(oc, d), cc = dot(oc, oc) - r * r, disc = b * b - cc;
    if (disc < 0.f) return -1.f;
    const float t = -b - sqrtf(disc);
    if (t <= 1e-4f) return -1.f;
    for (int k = 0; k < 3; k++) n[k] = (oc[k] + t * d[k]) / r;
    return t;
}
// finite cylinder (axis z through c, radius r, half length h), optionally with hemispherical ends (capsule)
__device__ __forceinline__ float hit_cyl(const float* o, const float* d, const Prim& p, bool capsule, float* n) {
    const float r = p.s[0], h = p.s[1];
    const float oc[3] = {o[0] - p.c[0], o[1] - p.c[1], o[2] - p.c[2]};
    const float dz = dot(d, p.z), oz = dot(oc, p.z);
    const float a = 1.f - dz * dz, b = dot(oc, d) - oz * dz, c = dot(oc, oc) - oz * oz - r * r;
    float best = -1.f;
    if (a > 1e-9f) {
        const float disc = b * b - a * c;
        if (disc >= 0.f) {
            const float t = (-b - sqrtf(disc)) / a;
            const float zt = oz + t * dz;
            if (t > 1e-4f && fabsf(zt) <= h) {
                best = t;
                for (int k = 0; k < 3; k++) n[k] = (oc[k] + t * d[k] - zt * p.z[k]) / r;
            }
        }
    }
    for (int s = -1; s <= 1; s += 2) {
        if (capsule) {
            const float cc[3] = {p.c[0] + s * h * p.z[0], p.c[1] + s * h * p.z[1], p.c[2] + s * h * p.z[2]};
            float nn[3];
            const float t = hit_sphere(o, d, cc, r, nn);
            if (t > 0.f && (best < 0.f || t < best)) { best = t; n[0] = nn[0]; n[1] = nn[1]; n[2] = nn[2]; }
        } else if (fabsf(dz) > 1e-9f) {      // flat caps
            const float t = (s * h - oz) / dz;
            if (t > 1e-4f && (best < 0.f || t < best)) {
                const float q[3] = {oc[0] + t * d[0], oc[1] + t * d[1], oc[2] + t * d[2]};
                const float qz = dot(q, p.z);
                if (dot(q, q) - qz * qz <= r * r) { best = t; for (int k = 0; k < 3; k++) n[k] = s * p.z[k]; }
            }
        }
    }
    return best;
}
__device__ __forceinline__ float hit_box(const float* o, const float* d, const Prim& p, float* n) {
    float y[3] = {p.z[1] * p.x[2] - p.z[2] * p.x[1], p.z[2] * p.x[0] - p.z[0] * p.x[2], p.z[0] * p.x[1] - p.z[1] * p.x[0]};
    const float oc[3] = {o[0] - p.c[0], o[1] - p.c[1], o[2] - p.c[2]};
    const float ol[3] = {dot(oc, p.x), dot(oc, y), dot(oc, p.z)}, dl[3] = {dot(d, p.x), dot(d, y), dot(d, p.z)};
    float t0 = -1e30f, t1 = 1e30f; int ax = 0; float sg = 1.f;
    for (int i = 0; i < 3; i++) {
        if (fabsf(dl[i]) < 1e-9f) { if (fabsf(ol[i]) > p.s[i]) return -1.f; continue; }
        float ta = (-p.s[i] - ol[i]) / dl[i], tb = (p.s[i] - ol[i]) / dl[i];
        float s = -1.f;
        if (ta > tb) { const float q = ta; ta = tb; tb = q; s = 1.f; }
        if (ta > t0) { t0 = ta; ax = i; sg = s; }
        if (tb < t1) t1 = tb;
    }
    if (t0 > t1 || t0 <= 1e-4f) return -1.f;
    const float* a = ax == 0 ? p.x : (ax == 1 ? y : p.z);
    for (int k = 0; k < 3; k++) n[k] = sg * a[k];
    return t0;
}

}  // namespace rt

__global__ void __launch_bounds__(256) k_render(const __grid_constant__ RenderParams P) {
    using namespace rt;
    __shared__ Prim prim[GFLAT];
    __shared__ float s_com[3];
    __shared__ float s_R[13][9], s_p[13][3];
    const DevModel<float>& m = *P.model;
    const int img = blockIdx.z, env = P.env_ids[img];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    // ---- forward kinematics of this environment (mj_kinematics): thread b < 13 places body b
    auto q = [&](int i) -> float {
        return P.real_is_double ? (float)reinterpret_cast<const double*>(P.sreal)[(size_t)env * P.s_stride + i]
                                : reinterpret_cast<const float*>(P.sreal)[(size_t)env * P.s_stride + i];
    };
    if (tid < 13) {
        float qq[4] = {q(3), q(4), q(5), q(6)};
        const float nq = sqrtf(qq[0] * qq[0] + qq[1] * qq[1] + qq[2] * qq[2] + qq[3] * qq[3]);
        if (nq < 1e-12f) { qq[0] = 1; qq[1] = qq[2] = qq[3] = 0; } else { for (int k = 0; k < 4; k++) qq[k] /= nq; }
        float R[9], p[3] = {q(0), q(1), q(2)};
        quat_to_mat(R, qq);
        if (tid > 0) {
            const int k = (tid - 1) / 3, li = (tid - 1) - 3 * k;
            for (int j = 0; j <= li; j++) {
                float t[3];
                matvec3(t, R, m.body_pos[k][j]);
                p[0] += t[0]; p[1] += t[1]; p[2] += t[2];
                const float* u = m.jnt_axis[k][j];
                float sn, cs;
                sincosf(q(7 + 3 * k + j) - m.qpos0[k][j], &sn, &cs);
                const float oc = 1.f - cs;
                const float Rl[9] = {cs + oc * u[0] * u[0], oc * u[0] * u[1] - sn * u[2], oc * u[0] * u[2] + sn * u[1],
                                     oc * u[0] * u[1] + sn * u[2], cs + oc * u[1] * u[1], oc * u[1] * u[2] - sn * u[0],
                                     oc * u[0] * u[2] - sn * u[1], oc * u[1] * u[2] + sn * u[0], cs + oc * u[2] * u[2]};
                float Rn[9];
                for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Rn[3 * r + c] = R[3 * r] * Rl[c] + R[3 * r + 1] * Rl[3 + c] + R[3 * r + 2] * Rl[6 + c];
                for (int e = 0; e < 9; e++) R[e] = Rn[e];
            }
        }
        for (int e = 0; e < 9; e++) s_R[tid][e] = R[e];
        for (int e = 0; e < 3; e++) s_p[tid][e] = p[e];
    }
    __syncthreads();
    if (tid == 0) {          // subtree COM of the trunk = centre of mass of the robot (what mode="trackcom" follows)
        float c[3] = {0, 0, 0};
        for (int b = 0; b < 13; b++) {
            const float* ip = b == 0 ? m.t_ipos : m.ipos[(b - 1) / 3][(b - 1) % 3];
            const float ms = b == 0 ? m.t_mass : m.mass[(b - 1) / 3][(b - 1) % 3];
            float t[3];
            matvec3(t, s_R[b], ip);
            for (int k = 0; k < 3; k++) c[k] += ms * (s_p[b][k] + t[k]);
        }
        for (int k = 0; k < 3; k++) s_com[k] = c[k] / m.total_mass;
    }
    if (tid < m.nflat) {
        const DevGeom<float>& g = m.flat[tid];
        Prim& p = prim[tid];
        p.type = g.type;
        float t[3];
        matvec3(t, s_R[g.slot], g.pos);
        for (int k = 0; k < 3; k++) { p.c[k] = s_p[g.slot][k] + t[k]; p.s[k] = g.size[k]; }
        matvec3(p.z, s_R[g.slot], g.zax);
        matvec3(p.x, s_R[g.slot], g.xax);
        // colours: trunk light grey, hips / thighs mid grey, calves darker, feet (condim-6 spheres) near black
        const float shade = g.slot == 0 ? 0.62f : ((g.slot - 1) % 3 == 2 ? (g.type == G_SPHERE ? 0.12f : 0.35f) : 0.48f);
        p.col[0] = shade; p.col[1] = shade; p.col[2] = shade * 1.05f;
    }
    __syncthreads();
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= P.width || py >= P.height) return;
    // ---- primary ray (pinhole, vertical field of view fovy; pixel centres)
    const float th = tanf(0.5f * P.fovy_deg * 0.017453292519943295f);
    const float u = (2.f * (px + 0.5f) / P.width - 1.f) * th * (float)P.width / (float)P.height;
    const float v = (1.f - 2.f * (py + 0.5f) / P.height) * th;
    float d[3], o[3];
    for (int k = 0; k < 3; k++) { d[k] = u * P.cam_x[k] + v * P.cam_y[k] - P.cam_z[k]; o[k] = s_com[k] + P.cam_off[k]; }
    const float dn = rsqrtf(dot(d, d));
    d[0] *= dn; d[1] *= dn; d[2] *= dn;
    float tbest = 1e30f, nb[3] = {0, 0, 1}, col[3];
    bool hit = false;
    for (int g = 0; g < m.nflat; g++) {
        const Prim& p = prim[g];
        float n[3];
        float t;
        if (p.type == G_SPHERE) t = hit_sphere(o, d, p.c, p.s[0], n);
        else if (p.type == G_CAPSULE) t = hit_cyl(o, d, p, true, n);
        else if (p.type == G_CYLINDER) t = hit_cyl(o, d, p, false, n);
        else t = hit_box(o, d, p, n);
        if (t > 0.f && t < tbest) { tbest = t; hit = true; nb[0] = n[0]; nb[1] = n[1]; nb[2] = n[2]; col[0] = p.col[0]; col[1] = p.col[1]; col[2] = p.col[2]; }
    }
    // floor plane (geom 0: the model's plane_pos / normal = plane_frame row 0), checker texture
    {
        const float* pn = m.plane_frame;
        const float den = dot(d, pn);
        if (den < -1e-6f) {
            const float op[3] = {o[0] - m.plane_pos[0], o[1] - m.plane_pos[1], o[2] - m.plane_pos[2]};
            const float t = -dot(op, pn) / den;
            if (t > 1e-4f && t < tbest) {
                tbest = t; hit = true; nb[0] = pn[0]; nb[1] = pn[1]; nb[2] = pn[2];
                const float fx = (o[0] + t * d[0]) * 5.f, fy = (o[1] + t * d[1]) * 5.f;       // texrepeat 5 per unit length (texuniform)
                const int chk = ((int)floorf(fx * 2.f) + (int)floorf(fy * 2.f)) & 1;           // builtin checker: 2 x 2 cells per texture
                col[0] = chk ? 0.2f : 0.1f; col[1] = chk ? 0.3f : 0.2f; col[2] = chk ? 0.4f : 0.3f;
            }
        }
    }
    float rgb[3];
    if (hit) {
        const float head = fmaxf(0.f, -dot(nb, d));                 // headlight at the camera
        const float sun = fmaxf(0.f, nb[2]);                        // directional light "0 0 -1" (scene_torque.xml:20), default diffuse 0.7
        // simple distance fog so that the checker fades instead of aliasing at the horizon
        const float fog = __expf(-0.02f * tbest);
        for (int k = 0; k < 3; k++) {
            const float lit = col[k] * (0.3f + 0.6f * head + 0.7f * sun);
            const float sky = k == 0 ? 0.3f : (k == 1 ? 0.5f : 0.7f);
            rgb[k] = lit * fog + 0.35f * sky * (1.f - fog);
        }
    } else {
        const float w = fminf(1.f, fmaxf(0.f, 0.5f + 0.5f * d[2] * 2.f));      // gradient skybox: rgb1 at the horizon band, rgb2 = black above
        rgb[0] = 0.3f * (1.f - w * 0.8f); rgb[1] = 0.5f * (1.f - w * 0.8f); rgb[2] = 0.7f * (1.f - w * 0.8f);
    }
    unsigned char* out = P.rgb + (((size_t)img * P.height + py) * P.width + px) * 3;
    for (int k = 0; k < 3; k++) out[k] = (unsigned char)(fminf(1.f, fmaxf(0.f, rgb[k])) * 255.f + 0.5f);
}

}  // namespace go1
