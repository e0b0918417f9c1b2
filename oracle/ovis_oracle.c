/*
 * oracle/ovis_oracle.c - plain-C restatement of the reference's scene-visibility arithmetic ("port").
 *
 * TEST INFRASTRUCTURE ONLY. This file is the CPU checker for tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py. Nothing under orion_b200/ links, loads or calls
 * it; the product path is the CUDA library and fails loudly without it.
 *
 * Parity status: PINNED. The reference ships no tests or golden vectors for this path (SURVEY.md
 * section 4), so the pin is the reference's own code executed in the build container: every function
 * below is compared bit-for-bit with oracle/_ref/liborion_ref.so (the unmodified reference sources
 * compiled in place, see oracle/Makefile and ref_harness.cc) by tests/test_oracle.py, and against the
 * fixtures under tests/golden/ that tests/golden/make_golden.py generated from that library.
 *
 * All arithmetic is IEEE binary32, round-to-nearest, evaluated in exactly the order GLM 0.9.8.0's
 * scalar templates evaluate it (the reference builds GLM with default precision packed_highp, so no
 * SIMD specialisation is active). Build with -ffp-contract=off and without -march=native/-ffast-math.
 * Reference citations are relative to /root/reference/; "glm/" = engine/3rdparty/glm/repo/glm/ and
 * unqualified module paths (core/, render/, engine/) are under engine/runtime/.
 */
#define _GNU_SOURCE
#include "ovis_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;
typedef struct { float w, x, y, z; } quat;
typedef struct { v4 c[4]; } m4;                 /* c[col], GLM m[col][row] */

const char *orc_name(void) { return "port"; }

/* ------------------------------------------------------------------ GLM vector helpers */

static v3 V3(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static v4 V4(float x, float y, float z, float w) { v4 r = { x, y, z, w }; return r; }
static v3 v3_add(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 v3_sub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 v3_mul(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static v3 v3_scale(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static v4 v4_add(v4 a, v4 b) { return V4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
static v4 v4_sub(v4 a, v4 b) { return V4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
static v4 v4_mul(v4 a, v4 b) { return V4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
static v4 v4_scale(v4 a, float s) { return V4(a.x * s, a.y * s, a.z * s, a.w * s); }

/* glm/detail/func_geometric.inl:53-61: tmp.x + tmp.y + tmp.z, left to right */
static float v3_dot(v3 a, v3 b) { v3 t = v3_mul(a, b); return t.x + t.y + t.z; }
/* glm/detail/func_geometric.inl:73-85 */
static v3 v3_cross(v3 x, v3 y) {
    return V3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
/* glm/detail/func_exponential.inl:128-133 */
static float inversesqrt_f(float x) { return 1.0f / sqrtf(x); }
/* glm/detail/func_geometric.inl:88-97 */
static v3 v3_normalize(v3 v) { return v3_scale(v, inversesqrt_f(v3_dot(v, v))); }
/* glm/detail/func_common.inl:15-28: ternaries, NOT fminf/fmaxf */
static float min_f(float x, float y) { return (x < y) ? x : y; }
static float max_f(float x, float y) { return (x > y) ? x : y; }
static v3 v3_min(v3 a, v3 b) { return V3(min_f(a.x, b.x), min_f(a.y, b.y), min_f(a.z, b.z)); }
static v3 v3_max(v3 a, v3 b) { return V3(max_f(a.x, b.x), max_f(a.y, b.y), max_f(a.z, b.z)); }
/* glm/detail/func_trigonometric.inl:12-17 */
static float radians_f(float deg) { return deg * (float)0.01745329251994329576923690768489; }

/* ------------------------------------------------------------------ GLM quaternion helpers */

static quat Q(float w, float x, float y, float z) { quat r = { w, x, y, z }; return r; }
/* glm/gtc/quaternion.inl:12-20 */
static float q_dot(quat a, quat b) {
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z, tw = a.w * b.w;
    return (tx + ty) + (tz + tw);
}
/* glm/gtc/quaternion.inl:256-270 */
static quat q_mul(quat p, quat q) {
    quat r;
    r.w = p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z;
    r.x = p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y;
    r.y = p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z;
    r.z = p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x;
    return r;
}
/* glm/gtc/quaternion.inl:313-321 */
static v3 q_rotate(quat q, v3 v) {
    v3 u = V3(q.x, q.y, q.z);
    v3 uv = v3_cross(u, v);
    v3 uuv = v3_cross(u, uv);
    return v3_add(v, v3_scale(v3_add(v3_scale(uv, q.w), uuv), 2.0f));
}
/* glm/gtc/quaternion.inl:207-217 (+ compute_quat_div_scalar :50-56) */
static quat q_inverse(quat q) {
    float d = q_dot(q, q);
    return Q(q.w / d, -q.x / d, -q.y / d, -q.z / d);
}
/* glm/gtc/quaternion.inl:378-393 */
static quat q_normalize(quat q) {
    float len = sqrtf(q_dot(q, q));
    if (len <= 0.0f)
        return Q(1.0f, 0.0f, 0.0f, 0.0f);
    float one = 1.0f / len;
    return Q(q.w * one, q.x * one, q.y * one, q.z * one);
}

/* ------------------------------------------------------------------ GLM matrix helpers */

static m4 m4_identity(void) {
    m4 r;
    r.c[0] = V4(1, 0, 0, 0); r.c[1] = V4(0, 1, 0, 0); r.c[2] = V4(0, 0, 1, 0); r.c[3] = V4(0, 0, 0, 1);
    return r;
}
/* glm/detail/type_mat4x4.inl:601-620 */
static m4 m4_mul(m4 a, m4 b) {
    m4 r;
    for (int j = 0; j < 4; j++) {
        v4 bj = b.c[j];
        r.c[j] = v4_add(v4_add(v4_add(v4_scale(a.c[0], bj.x), v4_scale(a.c[1], bj.y)),
                               v4_scale(a.c[2], bj.z)), v4_scale(a.c[3], bj.w));
    }
    return r;
}
/* glm/detail/type_mat4x4.inl:531-542 (mat4 * vec4) */
static v4 m4_mul_v4(m4 m, v4 v) {
    v4 mul0 = v4_scale(m.c[0], v.x), mul1 = v4_scale(m.c[1], v.y);
    v4 add0 = v4_add(mul0, mul1);
    v4 mul2 = v4_scale(m.c[2], v.z), mul3 = v4_scale(m.c[3], v.w);
    v4 add1 = v4_add(mul2, mul3);
    return v4_add(add0, add1);
}
/* glm/gtc/matrix_transform.inl:10-16 */
static m4 m4_translate(m4 m, v3 v) {
    m4 r = m;
    r.c[3] = v4_add(v4_add(v4_add(v4_scale(m.c[0], v.x), v4_scale(m.c[1], v.y)), v4_scale(m.c[2], v.z)), m.c[3]);
    return r;
}
/* glm/gtc/matrix_transform.inl:78-87 */
static m4 m4_scale(m4 m, v3 v) {
    m4 r;
    r.c[0] = v4_scale(m.c[0], v.x); r.c[1] = v4_scale(m.c[1], v.y); r.c[2] = v4_scale(m.c[2], v.z); r.c[3] = m.c[3];
    return r;
}
/* glm/gtc/quaternion.inl:593-624 (mat3_cast widened by mat4(mat3)) */
static m4 m4_cast(quat q) {
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
    float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    m4 r;
    r.c[0] = V4(1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy), 0.0f);
    r.c[1] = V4(2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx), 0.0f);
    r.c[2] = V4(2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy), 0.0f);
    r.c[3] = V4(0.0f, 0.0f, 0.0f, 1.0f);
    return r;
}
/* glm/gtc/matrix_transform.inl:258-278, GLM_DEPTH_ZERO_TO_ONE branch (glm.h:28) */
static m4 m4_perspective(float fovy, float aspect, float zNear, float zFar) {
    float tanHalfFovy = tanf(fovy / 2.0f);
    m4 r;
    memset(&r, 0, sizeof r);
    r.c[0].x = 1.0f / (aspect * tanHalfFovy);
    r.c[1].y = 1.0f / tanHalfFovy;
    r.c[2].w = -1.0f;
    r.c[2].z = zFar / (zNear - zFar);
    r.c[3].z = -(zFar * zNear) / (zFar - zNear);
    return r;
}
/* glm/detail/func_matrix.inl:297-355 */
static m4 m4_inverse(m4 M) {
#define m(c_, r_) (((const float *)&M.c[c_])[r_])
    float Coef00 = m(2,2) * m(3,3) - m(3,2) * m(2,3);
    float Coef02 = m(1,2) * m(3,3) - m(3,2) * m(1,3);
    float Coef03 = m(1,2) * m(2,3) - m(2,2) * m(1,3);
    float Coef04 = m(2,1) * m(3,3) - m(3,1) * m(2,3);
    float Coef06 = m(1,1) * m(3,3) - m(3,1) * m(1,3);
    float Coef07 = m(1,1) * m(2,3) - m(2,1) * m(1,3);
    float Coef08 = m(2,1) * m(3,2) - m(3,1) * m(2,2);
    float Coef10 = m(1,1) * m(3,2) - m(3,1) * m(1,2);
    float Coef11 = m(1,1) * m(2,2) - m(2,1) * m(1,2);
    float Coef12 = m(2,0) * m(3,3) - m(3,0) * m(2,3);
    float Coef14 = m(1,0) * m(3,3) - m(3,0) * m(1,3);
    float Coef15 = m(1,0) * m(2,3) - m(2,0) * m(1,3);
    float Coef16 = m(2,0) * m(3,2) - m(3,0) * m(2,2);
    float Coef18 = m(1,0) * m(3,2) - m(3,0) * m(1,2);
    float Coef19 = m(1,0) * m(2,2) - m(2,0) * m(1,2);
    float Coef20 = m(2,0) * m(3,1) - m(3,0) * m(2,1);
    float Coef22 = m(1,0) * m(3,1) - m(3,0) * m(1,1);
    float Coef23 = m(1,0) * m(2,1) - m(2,0) * m(1,1);

    v4 Fac0 = V4(Coef00, Coef00, Coef02, Coef03), Fac1 = V4(Coef04, Coef04, Coef06, Coef07);
    v4 Fac2 = V4(Coef08, Coef08, Coef10, Coef11), Fac3 = V4(Coef12, Coef12, Coef14, Coef15);
    v4 Fac4 = V4(Coef16, Coef16, Coef18, Coef19), Fac5 = V4(Coef20, Coef20, Coef22, Coef23);
    v4 Vec0 = V4(m(1,0), m(0,0), m(0,0), m(0,0)), Vec1 = V4(m(1,1), m(0,1), m(0,1), m(0,1));
    v4 Vec2 = V4(m(1,2), m(0,2), m(0,2), m(0,2)), Vec3 = V4(m(1,3), m(0,3), m(0,3), m(0,3));

    v4 Inv0 = v4_add(v4_sub(v4_mul(Vec1, Fac0), v4_mul(Vec2, Fac1)), v4_mul(Vec3, Fac2));
    v4 Inv1 = v4_add(v4_sub(v4_mul(Vec0, Fac0), v4_mul(Vec2, Fac3)), v4_mul(Vec3, Fac4));
    v4 Inv2 = v4_add(v4_sub(v4_mul(Vec0, Fac1), v4_mul(Vec1, Fac3)), v4_mul(Vec3, Fac5));
    v4 Inv3 = v4_add(v4_sub(v4_mul(Vec0, Fac2), v4_mul(Vec1, Fac4)), v4_mul(Vec2, Fac5));

    v4 SignA = V4(+1, -1, +1, -1), SignB = V4(-1, +1, -1, +1);
    m4 Inverse;
    Inverse.c[0] = v4_mul(Inv0, SignA); Inverse.c[1] = v4_mul(Inv1, SignB);
    Inverse.c[2] = v4_mul(Inv2, SignA); Inverse.c[3] = v4_mul(Inv3, SignB);

    v4 Row0 = V4(Inverse.c[0].x, Inverse.c[1].x, Inverse.c[2].x, Inverse.c[3].x);
    v4 Dot0 = v4_mul(M.c[0], Row0);
    float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    float OneOverDeterminant = 1.0f / Dot1;
    for (int j = 0; j < 4; j++)
        Inverse.c[j] = v4_scale(Inverse.c[j], OneOverDeterminant);
    return Inverse;
#undef m
}
/* glm/gtc/matrix_transform.inl:521-546 (lookAtRH); only the upper 3x3 is consumed by quat_cast */
static m4 m4_lookat(v3 eye, v3 center, v3 up) {
    v3 f = v3_normalize(v3_sub(center, eye));
    v3 s = v3_normalize(v3_cross(f, up));
    v3 u = v3_cross(s, f);
    m4 r = m4_identity();
    r.c[0].x = s.x; r.c[1].x = s.y; r.c[2].x = s.z;
    r.c[0].y = u.x; r.c[1].y = u.y; r.c[2].y = u.z;
    r.c[0].z = -f.x; r.c[1].z = -f.y; r.c[2].z = -f.z;
    r.c[3].x = -v3_dot(s, eye); r.c[3].y = -v3_dot(u, eye); r.c[3].z = v3_dot(f, eye);
    return r;
}
/* glm/gtc/quaternion.inl:627-688 */
static quat q_cast(m4 M) {
#define m(c_, r_) (((const float *)&M.c[c_])[r_])
    float fourXSquaredMinus1 = m(0,0) - m(1,1) - m(2,2);
    float fourYSquaredMinus1 = m(1,1) - m(0,0) - m(2,2);
    float fourZSquaredMinus1 = m(2,2) - m(0,0) - m(1,1);
    float fourWSquaredMinus1 = m(0,0) + m(1,1) + m(2,2);
    int biggestIndex = 0;
    float fourBiggestSquaredMinus1 = fourWSquaredMinus1;
    if (fourXSquaredMinus1 > fourBiggestSquaredMinus1) { fourBiggestSquaredMinus1 = fourXSquaredMinus1; biggestIndex = 1; }
    if (fourYSquaredMinus1 > fourBiggestSquaredMinus1) { fourBiggestSquaredMinus1 = fourYSquaredMinus1; biggestIndex = 2; }
    if (fourZSquaredMinus1 > fourBiggestSquaredMinus1) { fourBiggestSquaredMinus1 = fourZSquaredMinus1; biggestIndex = 3; }
    float biggestVal = sqrtf(fourBiggestSquaredMinus1 + 1.0f) * 0.5f;
    float mult = 0.25f / biggestVal;
    quat r = Q(0, 0, 0, 0);
    switch (biggestIndex) {
    case 0: r.w = biggestVal; r.x = (m(1,2) - m(2,1)) * mult; r.y = (m(2,0) - m(0,2)) * mult; r.z = (m(0,1) - m(1,0)) * mult; break;
    case 1: r.w = (m(1,2) - m(2,1)) * mult; r.x = biggestVal; r.y = (m(0,1) + m(1,0)) * mult; r.z = (m(2,0) + m(0,2)) * mult; break;
    case 2: r.w = (m(2,0) - m(0,2)) * mult; r.x = (m(0,1) + m(1,0)) * mult; r.y = biggestVal; r.z = (m(1,2) + m(2,1)) * mult; break;
    case 3: r.w = (m(0,1) - m(1,0)) * mult; r.x = (m(2,0) + m(0,2)) * mult; r.y = (m(1,2) + m(2,1)) * mult; r.z = biggestVal; break;
    }
    return r;
#undef m
}

/* ------------------------------------------------------------------ reference math */

/* core/include/math/functions.h:106-111 */
static quat quat_look_at(v3 forward, v3 up) {
    m4 look = m4_lookat(V3(0.0f, 0.0f, 0.0f), forward, up);
    return q_inverse(q_normalize(q_cast(look)));
}
/* core/include/math/functions.h:73-91 */
static quat quat_rotate_between(v3 from, v3 to) {
    float fromToLength = sqrtf(v3_dot(from, from) * v3_dot(to, to));
    float w = fromToLength + v3_dot(from, to);
    v3 n;
    if (w < 1.0e-6f * fromToLength) {
        w = 0.0f;
        n = (fabsf(from.x) > fabsf(from.z)) ? V3(-from.y, from.x, 0.0f) : V3(0.0f, -from.z, from.y);
    } else {
        n = v3_cross(from, to);
    }
    return q_normalize(Q(w, n.x, n.y, n.z));
}
/* core/src/math/transform.cc:26-35 */
static m4 transform_matrix(v3 p, quat q, v3 s) {
    return m4_mul(m4_mul(m4_translate(m4_identity(), p), m4_cast(q)), m4_scale(m4_identity(), s));
}
/* core/src/math/bounding_box.cc:27-40 */
static void aabb_transform(const float *b, m4 M, float *out) {
    v3 c0 = V3(M.c[0].x, M.c[0].y, M.c[0].z), c1 = V3(M.c[1].x, M.c[1].y, M.c[1].z);
    v3 c2 = V3(M.c[2].x, M.c[2].y, M.c[2].z), c3 = V3(M.c[3].x, M.c[3].y, M.c[3].z);
    v3 xa = v3_scale(c0, b[0]), xb = v3_scale(c0, b[3]);
    v3 ya = v3_scale(c1, b[1]), yb = v3_scale(c1, b[4]);
    v3 za = v3_scale(c2, b[2]), zb = v3_scale(c2, b[5]);
    v3 mn = v3_add(v3_add(v3_add(v3_min(xa, xb), v3_min(ya, yb)), v3_min(za, zb)), c3);
    v3 mx = v3_add(v3_add(v3_add(v3_max(xa, xb), v3_max(ya, yb)), v3_max(za, zb)), c3);
    out[0] = mn.x; out[1] = mn.y; out[2] = mn.z; out[3] = mx.x; out[4] = mx.y; out[5] = mx.z;
}

/* core/src/math/intersection.cc:42-57 with bounding_box.h:60-71 and plane.h:70-72 */
static int intersect_aabb(const float *planes, const float *b) {
    for (int i = 0; i < 6; i++) {
        const float *pl = planes + 4 * i;
        float px = (pl[0] >= 0.0f) ? b[3] : b[0];
        float py = (pl[1] >= 0.0f) ? b[4] : b[1];
        float pz = (pl[2] >= 0.0f) ? b[5] : b[2];
        float tx = pl[0] * px, ty = pl[1] * py, tz = pl[2] * pz;
        float d = (tx + ty + tz) - pl[3];
        if (d < 0.0f)
            return 0;
    }
    return 1;
}
/* core/src/math/intersection.cc:28-36 */
static int intersect_sphere(const float *planes, v3 centre, float radius) {
    for (int i = 0; i < 6; i++) {
        const float *pl = planes + 4 * i;
        float tx = pl[0] * centre.x, ty = pl[1] * centre.y, tz = pl[2] * centre.z;
        float d = (tx + ty + tz) - pl[3];
        if (d < -radius)
            return 0;
    }
    return 1;
}

/* ------------------------------------------------------------------ view (render_view.cc:102-135, frustum.cc:27-85) */

static void view_compute(const OrcView *v, m4 *out_view, m4 *out_proj, m4 *out_vp, m4 *out_inv,
                         float planes_out[24], float corners_out[24]) {
    v3 pos = V3(v->position[0], v->position[1], v->position[2]);
    quat ori = Q(v->orientation[0], v->orientation[1], v->orientation[2], v->orientation[3]);

    m4 position = m4_translate(m4_identity(), V3(-pos.x, -pos.y, -pos.z));      /* render_view.cc:109 */
    m4 orientation = m4_cast(q_inverse(ori));                                   /* :110 */
    m4 view = m4_mul(orientation, position);                                    /* :112 */

    float aspect = (float)v->vp_width / (float)v->vp_height;                    /* :84 */
    float fov = radians_f(v->hfov_deg);                                         /* :119 */
    float verticalFOV = 2.0f * atanf(tanf(fov * 0.5f) / aspect);                /* :120 */
    m4 proj = m4_perspective(verticalFOV, aspect, v->znear, v->zfar);           /* :122 */

    m4 vp = m4_mul(proj, view);                                                 /* :128 */
    m4 inv = m4_inverse(vp);                                                    /* :129 */
    if (out_view) *out_view = view;
    if (out_proj) *out_proj = proj;
    if (out_vp) *out_vp = vp;
    if (out_inv) *out_inv = inv;

#define M(c_, r_) (((const float *)&vp.c[c_])[r_])
    float raw[6][4] = {
        {  M(0,3) + M(0,0),  M(1,3) + M(1,0),  M(2,3) + M(2,0), -M(3,3) - M(3,0) },   /* left   frustum.cc:32-35 */
        {  M(0,3) - M(0,0),  M(1,3) - M(1,0),  M(2,3) - M(2,0), -M(3,3) + M(3,0) },   /* right  :37-40 */
        {  M(0,3) - M(0,1),  M(1,3) - M(1,1),  M(2,3) - M(2,1), -M(3,3) + M(3,1) },   /* top    :42-45 */
        {  M(0,3) + M(0,1),  M(1,3) + M(1,1),  M(2,3) + M(2,1), -M(3,3) - M(3,1) },   /* bottom :47-50 */
        {  M(0,3) + M(0,2),  M(1,3) + M(1,2),  M(2,3) + M(2,2), -M(3,3) - M(3,2) },   /* near   :52-55 */
        {  M(0,3) - M(0,2),  M(1,3) - M(1,2),  M(2,3) - M(2,2), -M(3,3) + M(3,2) },   /* far    :57-60 */
    };
#undef M
    for (int i = 0; i < 6; i++) {                                               /* :63-67 */
        float magSquared = (raw[i][0] * raw[i][0]) + (raw[i][1] * raw[i][1]) + (raw[i][2] * raw[i][2]);
        float inv_len = inversesqrt_f(magSquared);
        for (int k = 0; k < 4; k++)
            planes_out[4 * i + k] = raw[i][k] * inv_len;
    }
    if (corners_out) {                                                          /* :70-84 */
        static const float ndc[8][3] = {
            { -1, 1, -1 }, { 1, 1, -1 }, { -1, -1, -1 }, { 1, -1, -1 },
            { -1, 1, 1 }, { 1, 1, 1 }, { -1, -1, 1 }, { 1, -1, 1 } };
        for (int i = 0; i < 8; i++) {
            v4 inverted = m4_mul_v4(inv, V4(ndc[i][0], ndc[i][1], ndc[i][2], 1.0f));
            corners_out[3 * i] = inverted.x / inverted.w;
            corners_out[3 * i + 1] = inverted.y / inverted.w;
            corners_out[3 * i + 2] = inverted.z / inverted.w;
        }
    }
}

void orc_view_planes(const OrcView *view, float out_planes[24]) {
    view_compute(view, 0, 0, 0, 0, out_planes, 0);
}

void orc_view_matrices(const OrcView *view, float out_mats[64], float out_corners[24]) {
    m4 mv, mp, mvp, mi;
    float planes[24];
    view_compute(view, &mv, &mp, &mvp, &mi, planes, out_corners);
    memcpy(out_mats, &mv, 64); memcpy(out_mats + 16, &mp, 64);
    memcpy(out_mats + 32, &mvp, 64); memcpy(out_mats + 48, &mi, 64);
}

/* ------------------------------------------------------------------ transforms */

/* engine/src/entity.cc:383-402 */
void orc_world_trs(int64_t n, const int32_t *parent, const float *local_trs, float *out) {
    for (int64_t i = 0; i < n; i++) {
        const float *l = local_trs + 10 * i;
        v3 position = V3(l[0], l[1], l[2]);
        quat orientation = Q(l[3], l[4], l[5], l[6]);
        v3 scale = V3(l[7], l[8], l[9]);
        if (parent[i] >= 0) {
            const float *p = out + 10 * (int64_t)parent[i];
            v3 pp = V3(p[0], p[1], p[2]);
            quat pq = Q(p[3], p[4], p[5], p[6]);
            v3 ps = V3(p[7], p[8], p[9]);
            position = v3_add(q_rotate(pq, v3_mul(ps, position)), pp);          /* :394 */
            orientation = q_mul(pq, orientation);                               /* :395 */
            scale = v3_mul(ps, scale);                                          /* :396 */
        }
        float *o = out + 10 * i;
        o[0] = position.x; o[1] = position.y; o[2] = position.z;
        o[3] = orientation.w; o[4] = orientation.x; o[5] = orientation.y; o[6] = orientation.z;
        o[7] = scale.x; o[8] = scale.y; o[9] = scale.z;
    }
}

void orc_trs_matrix(int64_t n, const float *t, float *out_mat16) {
    for (int64_t i = 0; i < n; i++) {
        const float *l = t + 10 * i;
        m4 M = transform_matrix(V3(l[0], l[1], l[2]), Q(l[3], l[4], l[5], l[6]), V3(l[7], l[8], l[9]));
        memcpy(out_mat16 + 16 * i, &M, 64);
    }
}

void orc_world_aabb(int64_t n, const float *world_trs, const float *local_aabb, float *out_aabb) {
    for (int64_t i = 0; i < n; i++) {
        const float *l = world_trs + 10 * i;
        m4 M = transform_matrix(V3(l[0], l[1], l[2]), Q(l[3], l[4], l[5], l[6]), V3(l[7], l[8], l[9]));
        aabb_transform(local_aabb + 6 * i, M, out_aabb + 6 * i);                /* render_entity.cc:93 */
    }
}

/* ------------------------------------------------------------------ entity cull */

static int64_t cull_range(int64_t begin, int64_t end, const float *aabb, const float *planes,
                          const uint32_t *flags, uint32_t require, uint32_t *out) {
    int64_t count = 0;
    for (int64_t i = begin; i < end; i++) {                                     /* simple_render_world.cc:38-41 */
        if (intersect_aabb(planes, aabb + 6 * i)) {
            if (!flags || (flags[i] & require) == require)
                out[count++] = (uint32_t)i;
        }
    }
    return count;
}

int64_t orc_cull_aabb(int64_t n, const float *aabb, const float planes[24],
                      const uint32_t *flags, uint32_t require, uint32_t *out_idx) {
    return cull_range(0, n, aabb, planes, flags, require, out_idx);
}

typedef struct {
    int64_t begin, end, count;
    const float *aabb, *planes;
    const uint32_t *flags;
    uint32_t require;
    uint32_t *out;
} cull_job;

static void *cull_thread(void *arg) {
    cull_job *j = (cull_job *)arg;
    j->count = cull_range(j->begin, j->end, j->aabb, j->planes, j->flags, j->require, j->out + j->begin);
    return 0;
}

int64_t orc_cull_aabb_mt(int64_t n, const float *aabb, const float planes[24],
                         const uint32_t *flags, uint32_t require, uint32_t *out_idx, int nthreads) {
    if (nthreads <= 1)
        return orc_cull_aabb(n, aabb, planes, flags, require, out_idx);
    cull_job *jobs = (cull_job *)calloc((size_t)nthreads, sizeof(cull_job));
    pthread_t *tids = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
    for (int t = 0; t < nthreads; t++) {
        cull_job j = { n * t / nthreads, n * (t + 1) / nthreads, 0, aabb, planes, flags, require, out_idx };
        jobs[t] = j;
        pthread_create(&tids[t], 0, cull_thread, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++)
        pthread_join(tids[t], 0);
    int64_t total = jobs[0].count;
    for (int t = 1; t < nthreads; t++) {       /* close the gaps in range order: same list as one loop */
        memmove(out_idx + total, out_idx + jobs[t].begin, (size_t)jobs[t].count * sizeof(uint32_t));
        total += jobs[t].count;
    }
    free(jobs); free(tids);
    return total;
}

/* ------------------------------------------------------------------ lights */

/* render/src/render_light.cc:78-85 (setDirection normalises) */
static v3 light_direction(const OrcLight *l) {
    return v3_normalize(V3(l->direction[0], l->direction[1], l->direction[2]));
}

/* render/src/render_light.cc:227-249 */
void orc_light_spot_aabb(const OrcLight *l, float out_aabb[6]) {
    float radius = l->range * tanf(radians_f(l->cutoff));
    v3 scale = V3(radius, radius, l->range);
    quat orientation = quat_rotate_between(V3(0.0f, 0.0f, -1.0f), light_direction(l));
    m4 M = transform_matrix(V3(l->position[0], l->position[1], l->position[2]), orientation, scale);
    static const float base[6] = { -1.0f, -1.0f, -1.0f, 1.0f, 1.0f, 0.0f };
    aabb_transform(base, M, out_aabb);
}

/* render/src/render_light.cc:187-209; returns 1 when the light is culled */
static int light_culled(const OrcLight *l, const float *planes) {
    v3 colour = V3(l->colour[0], l->colour[1], l->colour[2]);
    if (!l->intensity || !sqrtf(v3_dot(colour, colour)))                        /* :189 */
        return 1;
    switch (l->type) {
    case ORC_LIGHT_AMBIENT:
    case ORC_LIGHT_DIRECTIONAL:
        return 0;
    case ORC_LIGHT_POINT:
        return !intersect_sphere(planes, V3(l->position[0], l->position[1], l->position[2]), l->range);
    case ORC_LIGHT_SPOT: {
        float box[6];
        orc_light_spot_aabb(l, box);
        return !intersect_aabb(planes, box);
    }
    default:
        return 1;
    }
}

int32_t orc_light_cull(int32_t n, const OrcLight *lights, const float planes[24], uint32_t *out_idx) {
    int32_t count = 0;
    for (int32_t i = 0; i < n; i++)                                             /* simple_render_world.cc:44-47 */
        if (!light_culled(&lights[i], planes))
            out_idx[count++] = (uint32_t)i;
    return count;
}

/* render/src/render_light.cc:258-332 */
int32_t orc_light_shadow_views(const OrcLight *l, OrcView out_views[6]) {
    if (!(l->flags & ORC_CASTS_SHADOW))                                         /* :259-260 */
        return 0;
    int32_t count = 0;
    if (l->type == ORC_LIGHT_SPOT) {
        quat o = quat_look_at(light_direction(l), V3(0.0f, 1.0f, 0.0f));        /* :270 */
        OrcView *v = &out_views[0];
        memcpy(v->position, l->position, 12);
        v->orientation[0] = o.w; v->orientation[1] = o.x; v->orientation[2] = o.y; v->orientation[3] = o.z;
        v->hfov_deg = l->cutoff * 2; v->znear = 0.1f; v->zfar = l->range;       /* :273 */
        count = 1;
    } else if (l->type == ORC_LIGHT_POINT) {
        static const float faces[6][6] = {                                      /* :298-308 */
            {  1, 0, 0,  0, -1, 0 }, { -1, 0, 0,  0, -1, 0 }, { 0,  1, 0,  0, 0,  1 },
            { 0, -1, 0,  0, 0, -1 }, { 0, 0,  1,  0, -1, 0 }, { 0, 0, -1,  0, -1, 0 } };
        for (int i = 0; i < 6; i++) {
            quat o = quat_look_at(V3(faces[i][0], faces[i][1], faces[i][2]), V3(faces[i][3], faces[i][4], faces[i][5]));
            OrcView *v = &out_views[i];
            memcpy(v->position, l->position, 12);
            v->orientation[0] = o.w; v->orientation[1] = o.x; v->orientation[2] = o.y; v->orientation[3] = o.z;
            v->hfov_deg = 90.0f; v->znear = 0.1f; v->zfar = l->range;           /* :319 */
        }
        count = 6;
    }
    for (int i = 0; i < count; i++) {                                           /* deferred_render_pipeline.cc:48,264 */
        out_views[i].vp_width = 512;
        out_views[i].vp_height = 512;
    }
    return count;
}

/* ------------------------------------------------------------------ shadow-caster lists */

/* render/src/deferred_render_pipeline.cc:207,247,257-284 */
int32_t orc_shadow_lists(int64_t n, const float *aabb, const uint32_t *flags,
                         int32_t n_lights, const OrcLight *lights, const float main_planes[24],
                         int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                         int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities) {
    uint32_t *visible = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n_lights > 0 ? n_lights : 1));
    uint32_t *scratch = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    int32_t nv = orc_light_cull(n_lights, lights, main_planes, visible);
    int32_t lists = 0;
    int64_t total = 0;
    out_offsets[0] = 0;
    for (int32_t k = 0; k < nv && lists >= 0; k++) {
        const OrcLight *l = &lights[visible[k]];
        OrcView views[6];
        int32_t count = orc_light_shadow_views(l, views);
        for (int32_t i = 0; i < count; i++) {
            float planes[24];
            orc_view_planes(&views[i], planes);
            int64_t c = orc_cull_aabb(n, aabb, planes, flags, ORC_CASTS_SHADOW, scratch);
            if (lists >= cap_lists || total + c > cap_entities) { lists = -1; break; }
            memcpy(out_entities + total, scratch, (size_t)c * sizeof(uint32_t));
            total += c;
            out_list_light[lists] = (int32_t)visible[k];
            out_list_view[lists] = i;
            out_offsets[++lists] = total;
        }
    }
    free(visible); free(scratch);
    return lists;
}

/* ------------------------------------------------------------------ draw lists and uniforms (section 8(f)) */

/* render/src/deferred_render_pipeline.cc:318-330 and :272-283 with render/src/draw_list.cc:35-45 */
void orc_draw_lists(int64_t n_visible, const uint32_t *visible, const uint32_t *pass_counts, const uint32_t *flags,
                    int32_t mode, uint32_t *out_draws0, uint32_t *out_draws1, int64_t out_counts[2]) {
    int64_t c0 = 0, c1 = 0;
    for (int64_t k = 0; k < n_visible; k++) {
        uint32_t e = visible[k];
        uint32_t pc = pass_counts[e];
        uint32_t deferred = pc & 0xffu, basic = (pc >> 8) & 0xffu, caster = (pc >> 16) & 0xffu;
        if (mode == 0) {
            if (deferred > 0) {                                             /* :322-323 */
                for (uint32_t i = 0; i < deferred; i++) { out_draws0[2 * c0] = e; out_draws0[2 * c0 + 1] = i; c0++; }
            } else if (basic > 0) {                                         /* :324-325 */
                for (uint32_t i = 0; i < basic; i++) { out_draws1[2 * c1] = e; out_draws1[2 * c1 + 1] = i; c1++; }
            }                                                               /* else: logWarning, nothing drawn */
        } else {
            if ((flags[e] & ORC_CASTS_SHADOW) && caster > 0)                /* :273-277 */
                for (uint32_t i = 0; i < caster; i++) { out_draws0[2 * c0] = e; out_draws0[2 * c0 + 1] = i; c0++; }
        }
    }
    out_counts[0] = c0;
    out_counts[1] = c1;
}

/* render/src/render_entity.cc:67-69 */
void orc_entity_uniforms(int64_t n, const float *world_trs, float *out20) {
    for (int64_t i = 0; i < n; i++) {
        const float *l = world_trs + 10 * i;
        m4 M = transform_matrix(V3(l[0], l[1], l[2]), Q(l[3], l[4], l[5], l[6]), V3(l[7], l[8], l[9]));
        memcpy(out20 + 20 * i, &M, 64);
        out20[20 * i + 16] = l[0]; out20[20 * i + 17] = l[1]; out20[20 * i + 18] = l[2];
        out20[20 * i + 19] = 0.0f;
    }
}

void orc_entity_uniform_blocks(int64_t n, const float *world_trs, void *out80) {
    orc_entity_uniforms(n, world_trs, (float *)out80);
}

void orc_uniform_sizes(int32_t out[3]) {
    out[0] = 80; out[1] = 288; out[2] = 208;
}

/* render/src/render_view.cc:51-58 (position), :74-90 (viewport), :102-135 (matrices) */
void orc_view_uniforms(const OrcView *view, const int32_t viewport_pos[2], void *out288) {
    float *f = (float *)out288;
    int32_t *i = (int32_t *)out288;
    float planes[24];
    m4 mv, mp, mvp, mi;
    memset(out288, 0, 288);
    view_compute(view, &mv, &mp, &mvp, &mi, planes, 0);
    memcpy(f, &mv, 64); memcpy(f + 16, &mp, 64); memcpy(f + 32, &mvp, 64); memcpy(f + 48, &mi, 64);
    i[64] = viewport_pos[0]; i[65] = viewport_pos[1];                          /* :79 */
    i[66] = view->vp_width; i[67] = view->vp_height;                           /* :80 */
    memcpy(f + 68, view->position, 12);                                        /* :55 */
}

/* render/src/render_light.cc setters :68-158, updateVolumeTransform :212-255, updateShadowViews :258-332 */
void orc_light_uniforms(const OrcLight *l, const float attenuation[3], float shadow_bias, void *out208) {
    float *f = (float *)out208;
    memset(out208, 0, 208);
    v3 pos = V3(l->position[0], l->position[1], l->position[2]);
    v3 dir = light_direction(l);                                               /* :79 */
    memcpy(f, l->position, 12);                                                /* :70 */
    f[3] = l->intensity;                                                       /* :98 */
    f[4] = dir.x; f[5] = dir.y; f[6] = dir.z;                                  /* :80 */
    memcpy(f + 8, l->colour, 12);                                              /* :91 */
    f[11] = l->range;                                                          /* :120, :336 */
    f[48] = shadow_bias;                                                       /* :158 */
    if (l->type == ORC_LIGHT_POINT) {
        m4 M = transform_matrix(pos, Q(1.0f, 0.0f, 0.0f, 0.0f), V3(l->range, l->range, l->range));   /* :222-226 */
        memcpy(f + 12, &M, 64);
        memcpy(f + 45, attenuation, 12);                                       /* :136-138 */
        if (l->flags & ORC_CASTS_SHADOW)
            f[44] = 0.1f;                                                      /* :324 */
    } else if (l->type == ORC_LIGHT_SPOT) {
        f[7] = cosf(radians_f(l->cutoff));                                     /* :110 */
        float radius = l->range * tanf(radians_f(l->cutoff));                  /* :235 */
        quat orientation = quat_rotate_between(V3(0.0f, 0.0f, -1.0f), dir);    /* :239 */
        m4 M = transform_matrix(pos, orientation, V3(radius, radius, l->range));
        memcpy(f + 12, &M, 64);                                                /* :242 */
        memcpy(f + 45, attenuation, 12);
        if (l->flags & ORC_CASTS_SHADOW) {                                     /* :259-260 */
            OrcView sv[6];
            orc_light_shadow_views(l, sv);
            sv[0].vp_width = sv[0].vp_height = 1;                              /* RenderView's initial aspect, render_view.cc:39 */
            m4 mv, mp, mvp, mi;
            float planes[24];
            view_compute(&sv[0], &mv, &mp, &mvp, &mi, planes, 0);
            m4 bias = m4_identity();                                           /* :284-288 */
            bias.c[0] = V4(0.5f, 0.0f, 0.0f, 0.0f);
            bias.c[1] = V4(0.0f, 0.5f, 0.0f, 0.0f);
            bias.c[2] = V4(0.0f, 0.0f, 1.0f, 0.0f);
            bias.c[3] = V4(0.5f, 0.5f, 0.0f, 1.0f);
            m4 shadow_space = m4_mul(m4_mul(bias, mp), mv);                     /* :289-290 */
            memcpy(f + 28, &shadow_space, 64);
        }
    }
}

void orc_quat_rotate(const float q[4], const float v[3], float out[3]) {
    v3 r = q_rotate(Q(q[0], q[1], q[2], q[3]), V3(v[0], v[1], v[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
