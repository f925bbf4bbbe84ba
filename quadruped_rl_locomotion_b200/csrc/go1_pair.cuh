// go1_pair.cuh -- robot-robot narrowphase (rows a6/a7 of the scope table: mj_collision for geoms of bodies that are not
// parent and child; go1_torque.xml:9,26 gives every collision geom contype = conaffinity = 1).
//
// One lane = one geom pair.  Same rules as the fp64 oracle (oracle/go1_oracle.c "robot-robot narrowphase"):
//   sphere-sphere, sphere-capsule, capsule-capsule, sphere-box      MuJoCo's analytic functions (engine_collision_primitive.c / _box.c)
//   sphere-cylinder                                                  exact closest feature, one contact (MuJoCo 2.3.0: MPR, tolerance 1e-6)
//   capsule-cylinder, capsule-box                                    centre (box: long patches both ends) of the contact patch of the capsule axis, see seg_patch
//   cylinder-cylinder (left vs right hip motors, coplanar)           rectangle vs rectangle in the common plane, see pp_cyl_cyl (MuJoCo: MPR)
// The normal points from geom1 to geom2, geom1 being the geom of the lower type (mj_collideGeoms' swap).
#pragma once

namespace go1 {

template <typename T> struct GeomW { int type; T c[3], z[3], x[3], s[3]; };            // world centre, z / x axes, size
template <typename T> struct Stage2 { T p[2][3], n[2][3], d[2]; int cnt; };

template <typename T> GO1_DEV void stage2_emit(Stage2<T>& st, const T* pos, const T* n, T dist) {
    if (st.cnt < 2) {
        const int j = st.cnt;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (j == 0) { st.p[0][k] = pos[k]; st.n[0][k] = n[k]; } else { st.p[1][k] = pos[k]; st.n[1][k] = n[k]; }
        }
        if (j == 0) st.d[0] = dist; else st.d[1] = dist;
    }
    st.cnt++;
}

template <typename T> GO1_DEV T normalize3_(T* v) {      // mju_normalize3
    typedef Num<T> N;
    const T n = N::sqrt_(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (n < N::MINVAL) { v[0] = 1; v[1] = 0; v[2] = 0; }
    else { const T inv = T(1) / n; v[0] *= inv; v[1] *= inv; v[2] *= inv; }
    return n;
}

// mju_makeFrame with an empty second axis: f[0..2] unit normal in, f[3..8] tangents out
template <typename T> GO1_DEV void make_frame_(T* f) {
    f[3] = 0; f[4] = 0; f[5] = 0;
    if (f[1] < T(0.5) && f[1] > T(-0.5)) f[4] = 1; else f[5] = 1;
    const T dd = f[0] * f[3] + f[1] * f[4] + f[2] * f[5];
    f[3] -= f[0] * dd; f[4] -= f[1] * dd; f[5] -= f[2] * dd;
    normalize3_(f + 3);
    cross3(f + 6, f, f + 3);
}

// mjraw_SphereSphere; fa / fb: z axes of the two geoms (normal of coincident centres)
template <typename T>
GO1_DEV void pp_sphere_sphere(const T* c1, T r1, const T* c2, T r2, const T* fa, const T* fb, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    const T dif[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
    const T cd2 = dot3(dif, dif), md = margin + r1 + r2;
    if (cd2 > md * md) return;
    const T len = N::sqrt_(cd2), dist = len - r1 - r2;
    T n[3];
    if (len < N::GEOM_EPS) { cross3(n, fa, fb); normalize3_(n); }      // coincident centres (axes that intersect): mjraw_SphereSphere's fallback
    else { const T inv = T(1) / len; n[0] = dif[0] * inv; n[1] = dif[1] * inv; n[2] = dif[2] * inv; }
    if (dist < inc) {
        const T pos[3] = {c1[0] + n[0] * (r1 + T(0.5) * dist), c1[1] + n[1] * (r1 + T(0.5) * dist), c1[2] + n[2] * (r1 + T(0.5) * dist)};
        stage2_emit(st, pos, n, dist);
    }
}

// mjc_SphereBox: sphere (c, r) vs box b
template <typename T>
GO1_DEV void pp_sphere_box(const T* c, T r, const GeomW<T>& b, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    T y[3];
    cross3(y, b.z, b.x);
    const T t[3] = {c[0] - b.c[0], c[1] - b.c[1], c[2] - b.c[2]};
    const T ctr[3] = {dot3(b.x, t), dot3(y, t), dot3(b.z, t)};
    T dv[3];
#pragma unroll
    for (int i = 0; i < 3; i++) dv[i] = N::min_(b.s[i], N::max_(-b.s[i], ctr[i])) - ctr[i];
    const T dist = N::sqrt_(dot3(dv, dv));
    if (dist - r > margin) return;
    T nl[3], pl[3], cd;
    if (dist <= N::MINVAL) {
        T closest = T(1e30); int k = 0;
#pragma unroll
        for (int i = 0; i < 6; i++) {
            const T fd = N::abs_(((i & 1) ? T(1) : T(-1)) * b.s[i >> 1] - ctr[i >> 1]);
            if (fd < closest) { closest = fd; k = i; }
        }
        const T sg = (k & 1) ? T(-1) : T(1);
        nl[0] = (k >> 1) == 0 ? sg : T(0); nl[1] = (k >> 1) == 1 ? sg : T(0); nl[2] = (k >> 1) == 2 ? sg : T(0);
        cd = -closest - r;
#pragma unroll
        for (int i = 0; i < 3; i++) pl[i] = ctr[i] + nl[i] * (r - closest) * T(0.5);
    } else {
        const T inv = T(1) / dist;
        cd = dist - r;
#pragma unroll
        for (int i = 0; i < 3; i++) { nl[i] = dv[i] * inv; pl[i] = ctr[i] + nl[i] * (r + T(0.5) * cd); }
    }
    if (cd < inc) {
        T n[3], pos[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            n[k] = b.x[k] * nl[0] + y[k] * nl[1] + b.z[k] * nl[2];
            pos[k] = b.x[k] * pl[0] + y[k] * pl[1] + b.z[k] * pl[2] + b.c[k];
        }
        stage2_emit(st, pos, n, cd);
    }
}

// sphere (c, r) vs solid cylinder q (radius s[0], half length s[1])
template <typename T>
GO1_DEV void pp_sphere_cyl(const T* c, T r, const GeomW<T>& q, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    const T t[3] = {c[0] - q.c[0], c[1] - q.c[1], c[2] - q.c[2]};
    const T z = dot3(q.z, t);
    const T w[3] = {t[0] - z * q.z[0], t[1] - z * q.z[1], t[2] - z * q.z[2]};
    const T rho = N::sqrt_(dot3(w, w)), R = q.s[0], h = q.s[1];
    T e[3];
    if (rho < N::MINVAL) { e[0] = q.x[0]; e[1] = q.x[1]; e[2] = q.x[2]; }
    else { const T inv = T(1) / rho; e[0] = w[0] * inv; e[1] = w[1] * inv; e[2] = w[2] * inv; }
    const T dz = N::abs_(z) - h, dr = rho - R, sz = z < 0 ? T(-1) : T(1);
    T n[3], pos[3], cd;
    if (dz <= 0 && dr <= 0) {
        T closest;
        if (-dr < -dz) { closest = -dr; n[0] = -e[0]; n[1] = -e[1]; n[2] = -e[2]; }
        else { closest = -dz; n[0] = -sz * q.z[0]; n[1] = -sz * q.z[1]; n[2] = -sz * q.z[2]; }
        if (-closest - r > margin) return;
        cd = -closest - r;
#pragma unroll
        for (int k = 0; k < 3; k++) pos[k] = c[k] + n[k] * (r - closest) * T(0.5);
    } else {
        const T zc = N::min_(h, N::max_(-h, z)), rc = N::min_(rho, R);
        T dv[3];
#pragma unroll
        for (int k = 0; k < 3; k++) dv[k] = q.c[k] + zc * q.z[k] + rc * e[k] - c[k];
        const T dist = N::sqrt_(dot3(dv, dv));
        if (dist - r > margin) return;
        if (dist < N::MINVAL) return;
        const T inv = T(1) / dist;
        cd = dist - r;
#pragma unroll
        for (int k = 0; k < 3; k++) { n[k] = dv[k] * inv; pos[k] = c[k] + n[k] * (r + T(0.5) * cd); }
    }
    if (cd < inc) stage2_emit(st, pos, n, cd);
}

// F(t) = squared distance between the axis point pl + t*al and a box (half sizes sz) / solid cylinder (sz = R, R, h) given in its
// own frame, g(t) = F'(t) / 2 and g'(t).  F is convex along the line, g continuous and non-decreasing.
template <typename T> GO1_DEV void seg_eval(const T* pl, const T* al, const T* sz, bool box, T t, T& F, T& g, T& dg) {
    typedef Num<T> N;
    const T x[3] = {pl[0] + t * al[0], pl[1] + t * al[1], pl[2] + t * al[2]};
    F = 0; g = 0; dg = 0;
    if (box) {
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const T ex = x[i] - N::min_(sz[i], N::max_(-sz[i], x[i]));
            F += ex * ex; g += ex * al[i];
            if (ex != 0) dg += al[i] * al[i];
        }
        return;
    }
    const T ez = x[2] - N::min_(sz[2], N::max_(-sz[2], x[2]));
    F = ez * ez; g = ez * al[2];
    if (ez != 0) dg = al[2] * al[2];
    const T rho = N::sqrt_(x[0] * x[0] + x[1] * x[1]);
    if (rho > sz[0]) {
        const T u = (x[0] * al[0] + x[1] * al[1]) / rho, a2 = al[0] * al[0] + al[1] * al[1];
        F += (rho - sz[0]) * (rho - sz[0]);
        g += (rho - sz[0]) * u;
        dg += u * u + (rho - sz[0]) * (a2 - u * u) / rho;
    }
}

// The part of the capsule axis within `reach` (radius + margin) of the box / cylinder: the contact patch [t1, t2] in axis
// parameters (t in [-1, 1]); returns false when no axis point is that close.  A point of minimum distance is found first (root of
// the monotone g by Newton steps inside a shrinking bracket, bisection when a step leaves it): where the axis runs along a face or
// the cylinder side that point is not unique and rounding picks one -- it only serves as a point INSIDE the patch.  The patch
// ends are roots of the convex F(t) = reach^2 approached by Newton from outside (monotone, no safeguard needed) and are well
// conditioned in every configuration, so contacts placed at the patch centre / ends do not jump.  The oracle runs the same steps.
template <typename T> GO1_DEV bool seg_patch(const T* pl, const T* al, const T* sz, bool box, T reach, T& t1, T& t2) {
    typedef Num<T> N;
    const T eps = sizeof(T) == 4 ? T(1e-6) : T(1e-13);
    const T r2 = reach * reach;
    T Flo, Fhi, glo, ghi, dg, ts, Fs;
    seg_eval(pl, al, sz, box, T(-1), Flo, glo, dg);
    seg_eval(pl, al, sz, box, T(1), Fhi, ghi, dg);
    if (glo >= 0) { ts = -1; Fs = Flo; }
    else if (ghi <= 0) { ts = 1; Fs = Fhi; }
    else {
        T lo = -1, hi = 1;
        ts = lo - glo * (hi - lo) / (ghi - glo);
#pragma unroll 1
        for (int it = 0; it < 16; it++) {
            T gt;
            seg_eval(pl, al, sz, box, ts, Fs, gt, dg);
            if (gt == 0) break;
            if (gt < 0) lo = ts; else hi = ts;
            T tn = dg > 0 ? ts - gt / dg : T(0.5) * (lo + hi);
            if (!(tn > lo && tn < hi)) tn = T(0.5) * (lo + hi);
            const bool done = N::abs_(tn - ts) <= eps;
            ts = tn;
            if (done) break;
        }
        T gt;
        seg_eval(pl, al, sz, box, ts, Fs, gt, dg);
    }
    if (Fs > r2) return false;
    t1 = -1; t2 = 1;
    if (Flo > r2) {
        T t = -1, F = Flo, g = glo;
#pragma unroll 1
        for (int it = 0; it < 12 && F > r2 && g < 0; it++) {
            T tn = t - T(0.5) * (F - r2) / g;
            if (tn > ts) tn = ts;
            const bool done = N::abs_(tn - t) <= eps;
            t = tn;
            if (done) break;
            seg_eval(pl, al, sz, box, t, F, g, dg);
        }
        t1 = t;
    }
    if (Fhi > r2) {
        T t = 1, F = Fhi, g = ghi;
#pragma unroll 1
        for (int it = 0; it < 12 && F > r2 && g > 0; it++) {
            T tn = t - T(0.5) * (F - r2) / g;
            if (tn < ts) tn = ts;
            const bool done = N::abs_(tn - t) <= eps;
            t = tn;
            if (done) break;
            seg_eval(pl, al, sz, box, t, F, g, dg);
        }
        t2 = t;
    }
    return true;
}

// capsule a (geom1) vs box / cylinder q (geom2): sphere-vs-body contacts at the centre of the contact patch; a patch longer than
// half the capsule against a BOX gives two contacts at its ends (mjc_CapsuleBox also returns up to two; a cylinder pair has one, as MPR)
template <typename T>
GO1_DEV void pp_capsule_convex(const GeomW<T>& a, const GeomW<T>& q, bool box, T margin, T inc, Stage2<T>& st) {
    T y[3];
    cross3(y, q.z, q.x);
    const T t[3] = {a.c[0] - q.c[0], a.c[1] - q.c[1], a.c[2] - q.c[2]};
    const T pl[3] = {dot3(q.x, t), dot3(y, t), dot3(q.z, t)};
    const T al[3] = {dot3(q.x, a.z) * a.s[1], dot3(y, a.z) * a.s[1], dot3(q.z, a.z) * a.s[1]};
    const T sz[3] = {q.s[0], box ? q.s[1] : q.s[0], box ? q.s[2] : q.s[1]};
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const T lim = sz[i] + a.s[0] + margin, u = pl[i] - al[i], v = pl[i] + al[i];
        if ((u > lim && v > lim) || (u < -lim && v < -lim)) return;
    }
    T t1, t2;
    if (!seg_patch(pl, al, sz, box, a.s[0] + margin, t1, t2)) return;
    const bool two = box && (t2 - t1 > T(1));
#pragma unroll 1
    for (int e = 0; e < 2; e++) {
        if (e == 1 && !two) break;
        const T tt = two ? (e == 0 ? t1 : t2) : T(0.5) * (t1 + t2);
        const T c[3] = {a.c[0] + a.z[0] * a.s[1] * tt, a.c[1] + a.z[1] * a.s[1] * tt, a.c[2] + a.z[2] * a.s[1] * tt};
        if (box) pp_sphere_box(c, a.s[0], q, margin, inc, st); else pp_sphere_cyl(c, a.s[0], q, margin, inc, st);
    }
}

// mjraw_CapsuleCapsule
template <typename T>
GO1_DEV void pp_capsule_capsule(const GeomW<T>& a, const GeomW<T>& b, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    const T a1[3] = {a.z[0] * a.s[1], a.z[1] * a.s[1], a.z[2] * a.s[1]}, a2[3] = {b.z[0] * b.s[1], b.z[1] * b.s[1], b.z[2] * b.s[1]};
    const T dif[3] = {a.c[0] - b.c[0], a.c[1] - b.c[1], a.c[2] - b.c[2]};
    const T ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2), u = -dot3(a1, dif), v = dot3(a2, dif);
    const T det = ma * mc - mb * mb;
    if (N::abs_(det) >= N::MINVAL) {
        T x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
        if (x1 > 1) { x1 = 1; x2 = (v - mb) / mc; }
        else if (x1 < -1) { x1 = -1; x2 = (v + mb) / mc; }
        if (x2 > 1) { x2 = 1; x1 = (u - mb) / ma; }
        else if (x2 < -1) { x2 = -1; x1 = (u + mb) / ma; }
        x1 = N::min_(T(1), N::max_(T(-1), x1));
        const T v1[3] = {a.c[0] + a1[0] * x1, a.c[1] + a1[1] * x1, a.c[2] + a1[2] * x1};
        const T v2[3] = {b.c[0] + a2[0] * x2, b.c[1] + a2[1] * x2, b.c[2] + a2[2] * x2};
        pp_sphere_sphere(v1, a.s[0], v2, b.s[0], a.z, b.z, margin, inc, st);
        return;
    }
    // parallel axes: up to two contacts.  MuJoCo counts contacts found by the raw test (dist <= margin), the active ones are a subset
    int found = 0;
#pragma unroll 1
    for (int e = 0; e < 4 && found < 2; e++) {
        T x1, x2;
        if (e == 0) { x1 = 1; x2 = N::min_(T(1), N::max_(T(-1), (v - mb) / mc)); }
        else if (e == 1) { x1 = -1; x2 = N::min_(T(1), N::max_(T(-1), (v + mb) / mc)); }
        else if (e == 2) { x2 = 1; x1 = N::min_(T(1), N::max_(T(-1), (u - mb) / ma)); }
        else { x2 = -1; x1 = N::min_(T(1), N::max_(T(-1), (u + mb) / ma)); }
        const T v1[3] = {a.c[0] + a1[0] * x1, a.c[1] + a1[1] * x1, a.c[2] + a1[2] * x1};
        const T v2[3] = {b.c[0] + a2[0] * x2, b.c[1] + a2[1] * x2, b.c[2] + a2[2] * x2};
        const T d3[3] = {v2[0] - v1[0], v2[1] - v1[1], v2[2] - v1[2]};
        const T md = margin + a.s[0] + b.s[0];
        if (dot3(d3, d3) <= md * md) found++;
        pp_sphere_sphere(v1, a.s[0], v2, b.s[0], a.z, b.z, margin, inc, st);
    }
}

// signed distance of the 2-D point (px, py) to the rectangle |x| <= hx, |y| <= hy and its outward unit normal there
template <typename T> GO1_DEV T rect_sdf(T px, T py, T hx, T hy, T& gx, T& gy) {
    typedef Num<T> N;
    const T qx = N::abs_(px) - hx, qy = N::abs_(py) - hy, sx = px < 0 ? T(-1) : T(1), sy = py < 0 ? T(-1) : T(1);
    if (qx <= 0 && qy <= 0) {
        if (qx > qy) { gx = sx; gy = 0; return qx; }
        gx = 0; gy = sy; return qy;
    }
    const T ox = N::max_(qx, T(0)), oy = N::max_(qy, T(0)), d = N::sqrt_(ox * ox + oy * oy), inv = T(1) / d;
    gx = sx * ox * inv; gy = sy * oy * inv;
    return d;
}

// Two solid cylinders with coplanar axes and centres (left / right hip motors: both turn about the trunk's x axis): rectangle vs
// rectangle in that plane, one contact per rectangle whose deepest vertex is within the margin of the other one (oracle:
// cylinderCylinderCoplanar; MuJoCo 2.3.0: MPR, always one contact).  No other cylinder pair of this robot passes the bounding-sphere test.
template <typename T>
GO1_DEV void pp_cyl_cyl(const GeomW<T>& a, const GeomW<T>& b, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    const T t[3] = {b.c[0] - a.c[0], b.c[1] - a.c[1], b.c[2] - a.c[2]};
    T nrm[3], e2[3];
    cross3(nrm, a.z, b.z);
    if (dot3(nrm, nrm) < T(1e-12)) cross3(nrm, a.z, t);
    if (dot3(nrm, nrm) < T(1e-18)) { nrm[0] = a.x[0]; nrm[1] = a.x[1]; nrm[2] = a.x[2]; }
    normalize3_(nrm);
    cross3(e2, nrm, a.z);
    const T cx = dot3(t, a.z), cy = dot3(t, e2), ax = dot3(b.z, a.z), ay = dot3(b.z, e2);
    // the deepest vertex of EACH rectangle in the other one is a contact of its own (no minimum over the two families: they tie
    // in the mirror-symmetric pose and rounding would pick the normal)
    T best = T(1e30), bnx = 0, bny = 0, bpx = 0, bpy = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const T u = (k & 2) ? -b.s[1] : b.s[1], v = (k & 1) ? -b.s[0] : b.s[0];
        const T vx = cx + u * ax - v * ay, vy = cy + u * ay + v * ax;
        T gx, gy;
        const T d = rect_sdf(vx, vy, a.s[1], a.s[0], gx, gy);
        if (d < best) { best = d; bnx = gx; bny = gy; bpx = vx - T(0.5) * d * gx; bpy = vy - T(0.5) * d * gy; }
    }
    if (best <= margin && best < inc) {
        const T n[3] = {bnx * a.z[0] + bny * e2[0], bnx * a.z[1] + bny * e2[1], bnx * a.z[2] + bny * e2[2]};
        const T pos[3] = {a.c[0] + bpx * a.z[0] + bpy * e2[0], a.c[1] + bpx * a.z[1] + bpy * e2[1], a.c[2] + bpx * a.z[2] + bpy * e2[2]};
        stage2_emit(st, pos, n, best);
    }
    best = T(1e30);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const T wx = (k & 2) ? -a.s[1] : a.s[1], wy = (k & 1) ? -a.s[0] : a.s[0];
        const T rx = wx - cx, ry = wy - cy, lx = rx * ax + ry * ay, ly = -rx * ay + ry * ax;
        T gx, gy;
        const T d = rect_sdf(lx, ly, b.s[1], b.s[0], gx, gy);
        if (d < best) {
            const T nx = gx * ax - gy * ay, ny = gx * ay + gy * ax;
            best = d; bnx = -nx; bny = -ny; bpx = wx - T(0.5) * d * nx; bpy = wy - T(0.5) * d * ny;
        }
    }
    if (best <= margin && best < inc) {
        const T n[3] = {bnx * a.z[0] + bny * e2[0], bnx * a.z[1] + bny * e2[1], bnx * a.z[2] + bny * e2[2]};
        const T pos[3] = {a.c[0] + bpx * a.z[0] + bpy * e2[0], a.c[1] + bpx * a.z[1] + bpy * e2[1], a.c[2] + bpx * a.z[2] + bpy * e2[2]};
        stage2_emit(st, pos, n, best);
    }
}

// narrowphase dispatch for an ordered pair (type(a) <= type(b)); returns false for a type pair that is not implemented
template <typename T>
GO1_DEV bool pp_collide(const GeomW<T>& a, const GeomW<T>& b, T margin, T inc, Stage2<T>& st) {
    typedef Num<T> N;
    if (a.type == G_SPHERE) {
        if (b.type == G_SPHERE) pp_sphere_sphere(a.c, a.s[0], b.c, b.s[0], a.z, b.z, margin, inc, st);
        else if (b.type == G_CAPSULE) {
            const T v[3] = {a.c[0] - b.c[0], a.c[1] - b.c[1], a.c[2] - b.c[2]};
            const T x = N::min_(b.s[1], N::max_(-b.s[1], dot3(b.z, v)));
            const T q[3] = {b.c[0] + b.z[0] * x, b.c[1] + b.z[1] * x, b.c[2] + b.z[2] * x};
            pp_sphere_sphere(a.c, a.s[0], q, b.s[0], a.z, b.z, margin, inc, st);
        }
        else if (b.type == G_CYLINDER) pp_sphere_cyl(a.c, a.s[0], b, margin, inc, st);
        else pp_sphere_box(a.c, a.s[0], b, margin, inc, st);
        return true;
    }
    if (a.type == G_CAPSULE) {
        if (b.type == G_CAPSULE) pp_capsule_capsule(a, b, margin, inc, st);
        else pp_capsule_convex(a, b, b.type == G_BOX, margin, inc, st);
        return true;
    }
    if (a.type == G_CYLINDER && b.type == G_CYLINDER) { pp_cyl_cyl(a, b, margin, inc, st); return true; }
    return false;
}

}  // namespace go1
