// ovis_math.hpp - host-side view / frustum / light-volume set-up for liborion_vis.
//
// O(views + lights) scalar work that stays on the CPU: it calls libm (tanf/atanf/sqrtf) exactly where
// the reference does, and evaluates every expression in the association order of the reference's GLM
// 0.9.8.0 scalar templates, so that the frustum planes handed to the kernels are the reference's planes.
// Compile with -ffp-contract=off (no FMA contraction). Citations are relative to
// /root/reference/engine/runtime/ ("glm/" = /root/reference/engine/3rdparty/glm/repo/glm/).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

#include "orion_vis.h"

namespace ovis {
namespace hm {

struct Vec3 {
    float x, y, z;
};
struct Vec4 {
    float x, y, z, w;
    float operator[](int i) const { return (&x)[i]; }
};
struct Quat {
    float w, x, y, z;
};
struct Mat4 {
    Vec4 col[4];   // col[c][r], GLM m[c][r]
    float at(int c, int r) const { return col[c][r]; }
};

inline Vec3 operator+(Vec3 a, Vec3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline Vec3 operator*(Vec3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
inline Vec4 operator+(Vec4 a, Vec4 b) { return { a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w }; }
inline Vec4 operator-(Vec4 a, Vec4 b) { return { a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w }; }
inline Vec4 operator*(Vec4 a, Vec4 b) { return { a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w }; }
inline Vec4 operator*(Vec4 a, float s) { return { a.x * s, a.y * s, a.z * s, a.w * s }; }

// glm/detail/func_geometric.inl:53-61 - (x + y) + z
inline float dot(Vec3 a, Vec3 b) {
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return tx + ty + tz;
}
// glm/detail/func_geometric.inl:73-85
inline Vec3 cross(Vec3 a, Vec3 b) {
    return { a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y };
}
// glm/detail/func_exponential.inl:128-133 - a true divide of a true sqrt, never an rsqrt
inline float inverseSqrt(float v) { return 1.0f / std::sqrt(v); }
// glm/detail/func_geometric.inl:88-97
inline Vec3 normalize(Vec3 v) { return v * inverseSqrt(dot(v, v)); }
// glm/detail/func_trigonometric.inl:12-17
inline float radians(float degrees) { return degrees * static_cast<float>(0.01745329251994329576923690768489); }
// glm/detail/func_common.inl:15-28
inline float glmMin(float a, float b) { return a < b ? a : b; }
inline float glmMax(float a, float b) { return a > b ? a : b; }

// glm/gtc/quaternion.inl:12-20
inline float dot(Quat a, Quat b) {
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z, tw = a.w * b.w;
    return (tx + ty) + (tz + tw);
}
// glm/gtc/quaternion.inl:207-217, :50-56
inline Quat inverse(Quat q) {
    float d = dot(q, q);
    return { q.w / d, -q.x / d, -q.y / d, -q.z / d };
}
// glm/gtc/quaternion.inl:378-393
inline Quat normalize(Quat q) {
    float len = std::sqrt(dot(q, q));
    if (len <= 0.0f)
        return { 1.0f, 0.0f, 0.0f, 0.0f };
    float inv = 1.0f / len;
    return { q.w * inv, q.x * inv, q.y * inv, q.z * inv };
}

inline Mat4 identity() {
    return { { { 1, 0, 0, 0 }, { 0, 1, 0, 0 }, { 0, 0, 1, 0 }, { 0, 0, 0, 1 } } };
}
// glm/detail/type_mat4x4.inl:601-620 - ((A0*b0 + A1*b1) + A2*b2) + A3*b3 per column
inline Mat4 operator*(const Mat4 &a, const Mat4 &b) {
    Mat4 r;
    for (int j = 0; j < 4; j++)
        r.col[j] = a.col[0] * b.col[j].x + a.col[1] * b.col[j].y + a.col[2] * b.col[j].z + a.col[3] * b.col[j].w;
    return r;
}
// glm/detail/type_mat4x4.inl:531-542 - (m0*x + m1*y) + (m2*z + m3*w)
inline Vec4 operator*(const Mat4 &m, Vec4 v) {
    Vec4 lo = m.col[0] * v.x + m.col[1] * v.y;
    Vec4 hi = m.col[2] * v.z + m.col[3] * v.w;
    return lo + hi;
}
// glm/gtc/matrix_transform.inl:10-16
inline Mat4 translate(const Mat4 &m, Vec3 v) {
    Mat4 r = m;
    r.col[3] = m.col[0] * v.x + m.col[1] * v.y + m.col[2] * v.z + m.col[3];
    return r;
}
// glm/gtc/matrix_transform.inl:78-87
inline Mat4 scale(const Mat4 &m, Vec3 v) {
    return { { m.col[0] * v.x, m.col[1] * v.y, m.col[2] * v.z, m.col[3] } };
}
// glm/gtc/quaternion.inl:593-624
inline Mat4 mat4Cast(Quat q) {
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
    float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    Mat4 r;
    r.col[0] = { 1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy), 0.0f };
    r.col[1] = { 2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx), 0.0f };
    r.col[2] = { 2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy), 0.0f };
    r.col[3] = { 0.0f, 0.0f, 0.0f, 1.0f };
    return r;
}
// glm/gtc/matrix_transform.inl:258-278 with GLM_DEPTH_ZERO_TO_ONE (engine/3rdparty/glm/include/glm.h:28)
inline Mat4 perspectiveRhZo(float fovy, float aspect, float zNear, float zFar) {
    float tanHalfFovy = std::tan(fovy / 2.0f);
    Mat4 r;
    std::memset(&r, 0, sizeof r);
    r.col[0].x = 1.0f / (aspect * tanHalfFovy);
    r.col[1].y = 1.0f / tanHalfFovy;
    r.col[2].w = -1.0f;
    r.col[2].z = zFar / (zNear - zFar);
    r.col[3].z = -(zFar * zNear) / (zFar - zNear);
    return r;
}
// glm/detail/func_matrix.inl:297-355 (cofactor inverse)
inline Mat4 inverse(const Mat4 &m) {
    auto sub = [&](int c0, int r0, int c1, int r1, int c2, int r2, int c3, int r3) {
        return m.at(c0, r0) * m.at(c1, r1) - m.at(c2, r2) * m.at(c3, r3);
    };
    float c00 = sub(2, 2, 3, 3, 3, 2, 2, 3), c02 = sub(1, 2, 3, 3, 3, 2, 1, 3), c03 = sub(1, 2, 2, 3, 2, 2, 1, 3);
    float c04 = sub(2, 1, 3, 3, 3, 1, 2, 3), c06 = sub(1, 1, 3, 3, 3, 1, 1, 3), c07 = sub(1, 1, 2, 3, 2, 1, 1, 3);
    float c08 = sub(2, 1, 3, 2, 3, 1, 2, 2), c10 = sub(1, 1, 3, 2, 3, 1, 1, 2), c11 = sub(1, 1, 2, 2, 2, 1, 1, 2);
    float c12 = sub(2, 0, 3, 3, 3, 0, 2, 3), c14 = sub(1, 0, 3, 3, 3, 0, 1, 3), c15 = sub(1, 0, 2, 3, 2, 0, 1, 3);
    float c16 = sub(2, 0, 3, 2, 3, 0, 2, 2), c18 = sub(1, 0, 3, 2, 3, 0, 1, 2), c19 = sub(1, 0, 2, 2, 2, 0, 1, 2);
    float c20 = sub(2, 0, 3, 1, 3, 0, 2, 1), c22 = sub(1, 0, 3, 1, 3, 0, 1, 1), c23 = sub(1, 0, 2, 1, 2, 0, 1, 1);

    Vec4 f0 = { c00, c00, c02, c03 }, f1 = { c04, c04, c06, c07 }, f2 = { c08, c08, c10, c11 };
    Vec4 f3 = { c12, c12, c14, c15 }, f4 = { c16, c16, c18, c19 }, f5 = { c20, c20, c22, c23 };
    Vec4 v0 = { m.at(1, 0), m.at(0, 0), m.at(0, 0), m.at(0, 0) }, v1 = { m.at(1, 1), m.at(0, 1), m.at(0, 1), m.at(0, 1) };
    Vec4 v2 = { m.at(1, 2), m.at(0, 2), m.at(0, 2), m.at(0, 2) }, v3 = { m.at(1, 3), m.at(0, 3), m.at(0, 3), m.at(0, 3) };

    Vec4 i0 = v1 * f0 - v2 * f1 + v3 * f2;
    Vec4 i1 = v0 * f0 - v2 * f3 + v3 * f4;
    Vec4 i2 = v0 * f1 - v1 * f3 + v3 * f5;
    Vec4 i3 = v0 * f2 - v1 * f4 + v2 * f5;

    const Vec4 signA = { +1, -1, +1, -1 }, signB = { -1, +1, -1, +1 };
    Mat4 inv = { { i0 * signA, i1 * signB, i2 * signA, i3 * signB } };
    Vec4 row0 = { inv.col[0].x, inv.col[1].x, inv.col[2].x, inv.col[3].x };
    Vec4 d = m.col[0] * row0;
    float det = (d.x + d.y) + (d.z + d.w);
    float oneOverDet = 1.0f / det;
    for (auto &c : inv.col)
        c = c * oneOverDet;
    return inv;
}
// glm/gtc/quaternion.inl:627-688 applied to the rotation block of lookAtRH (glm/gtc/matrix_transform.inl:521-546)
inline Quat quatCastLookAt(Vec3 forward, Vec3 up) {
    Vec3 f = normalize(forward - Vec3{ 0.0f, 0.0f, 0.0f });
    Vec3 s = normalize(cross(f, up));
    Vec3 u = cross(s, f);
    // m[c][r]: column c of the look matrix
    float m[3][3] = { { s.x, u.x, -f.x }, { s.y, u.y, -f.y }, { s.z, u.z, -f.z } };
    float fourX = m[0][0] - m[1][1] - m[2][2];
    float fourY = m[1][1] - m[0][0] - m[2][2];
    float fourZ = m[2][2] - m[0][0] - m[1][1];
    float fourW = m[0][0] + m[1][1] + m[2][2];
    int biggest = 0;
    float fourBiggest = fourW;
    if (fourX > fourBiggest) { fourBiggest = fourX; biggest = 1; }
    if (fourY > fourBiggest) { fourBiggest = fourY; biggest = 2; }
    if (fourZ > fourBiggest) { fourBiggest = fourZ; biggest = 3; }
    float biggestVal = std::sqrt(fourBiggest + 1.0f) * 0.5f;
    float mult = 0.25f / biggestVal;
    switch (biggest) {
    case 0: return { biggestVal, (m[1][2] - m[2][1]) * mult, (m[2][0] - m[0][2]) * mult, (m[0][1] - m[1][0]) * mult };
    case 1: return { (m[1][2] - m[2][1]) * mult, biggestVal, (m[0][1] + m[1][0]) * mult, (m[2][0] + m[0][2]) * mult };
    case 2: return { (m[2][0] - m[0][2]) * mult, (m[0][1] + m[1][0]) * mult, biggestVal, (m[1][2] + m[2][1]) * mult };
    default: return { (m[0][1] - m[1][0]) * mult, (m[2][0] + m[0][2]) * mult, (m[1][2] + m[2][1]) * mult, biggestVal };
    }
}

// core/include/math/functions.h:106-111
inline Quat quatLookAt(Vec3 forward, Vec3 up = { 0.0f, 1.0f, 0.0f }) {
    return inverse(normalize(quatCastLookAt(forward, up)));
}
// core/include/math/functions.h:73-91
inline Quat quatRotateBetween(Vec3 from, Vec3 to) {
    float fromToLength = std::sqrt(dot(from, from) * dot(to, to));
    float w = fromToLength + dot(from, to);
    Vec3 n;
    if (w < 1.0e-6f * fromToLength) {
        w = 0.0f;
        n = (std::fabs(from.x) > std::fabs(from.z)) ? Vec3{ -from.y, from.x, 0.0f } : Vec3{ 0.0f, -from.z, from.y };
    } else {
        n = cross(from, to);
    }
    return normalize(Quat{ w, n.x, n.y, n.z });
}
// core/src/math/transform.cc:26-35
inline Mat4 transformMatrix(Vec3 p, Quat q, Vec3 s) {
    return translate(identity(), p) * mat4Cast(q) * scale(identity(), s);
}
// core/src/math/bounding_box.cc:27-40
inline void transformAabb(const float b[6], const Mat4 &m, float out[6]) {
    Vec3 c0 = { m.col[0].x, m.col[0].y, m.col[0].z }, c1 = { m.col[1].x, m.col[1].y, m.col[1].z };
    Vec3 c2 = { m.col[2].x, m.col[2].y, m.col[2].z }, c3 = { m.col[3].x, m.col[3].y, m.col[3].z };
    Vec3 xa = c0 * b[0], xb = c0 * b[3], ya = c1 * b[1], yb = c1 * b[4], za = c2 * b[2], zb = c2 * b[5];
    auto lo = [](Vec3 a, Vec3 b2) { return Vec3{ glmMin(a.x, b2.x), glmMin(a.y, b2.y), glmMin(a.z, b2.z) }; };
    auto hi = [](Vec3 a, Vec3 b2) { return Vec3{ glmMax(a.x, b2.x), glmMax(a.y, b2.y), glmMax(a.z, b2.z) }; };
    Vec3 mn = lo(xa, xb) + lo(ya, yb) + lo(za, zb) + c3;
    Vec3 mx = hi(xa, xb) + hi(ya, yb) + hi(za, zb) + c3;
    out[0] = mn.x; out[1] = mn.y; out[2] = mn.z; out[3] = mx.x; out[4] = mx.y; out[5] = mx.z;
}

struct ViewMatrices {
    Mat4 view, projection, viewProjection, inverseViewProjection;
};

// render/src/render_view.cc:102-135
inline ViewMatrices viewMatrices(const OvisView &v) {
    ViewMatrices r;
    Vec3 p = { v.position[0], v.position[1], v.position[2] };
    Quat q = { v.orientation[0], v.orientation[1], v.orientation[2], v.orientation[3] };
    Mat4 position = translate(identity(), Vec3{ -p.x, -p.y, -p.z });                       // :109
    Mat4 orientation = mat4Cast(inverse(q));                                              // :110
    r.view = orientation * position;                                                      // :112
    float aspect = static_cast<float>(v.vp_width) / static_cast<float>(v.vp_height);      // :84
    float fov = radians(v.hfov_deg);                                                      // :119
    float verticalFov = 2.0f * atanf(tanf(fov * 0.5f) / aspect);                          // :120
    r.projection = perspectiveRhZo(verticalFov, aspect, v.znear, v.zfar);                 // :122
    r.viewProjection = r.projection * r.view;                                             // :128
    r.inverseViewProjection = inverse(r.viewProjection);                                  // :129
    return r;
}

// core/src/math/frustum.cc:27-67 - Gribb/Hartmann extraction with the GL (-1..1) near-plane formula on
// a 0..1-depth projection; that quirk is the reference's behaviour and is reproduced, not corrected.
inline void frustumPlanes(const Mat4 &m, float out[24]) {
    const float r3[4] = { m.at(0, 3), m.at(1, 3), m.at(2, 3), m.at(3, 3) };
    struct { int row; float sign; } spec[6] = { { 0, +1 }, { 0, -1 }, { 1, -1 }, { 1, +1 }, { 2, +1 }, { 2, -1 } };
    for (int i = 0; i < 6; i++) {
        int row = spec[i].row;
        float raw[4];
        if (spec[i].sign > 0) {
            raw[0] = r3[0] + m.at(0, row); raw[1] = r3[1] + m.at(1, row); raw[2] = r3[2] + m.at(2, row);
            raw[3] = -r3[3] - m.at(3, row);
        } else {
            raw[0] = r3[0] - m.at(0, row); raw[1] = r3[1] - m.at(1, row); raw[2] = r3[2] - m.at(2, row);
            raw[3] = -r3[3] + m.at(3, row);
        }
        float magSquared = (raw[0] * raw[0]) + (raw[1] * raw[1]) + (raw[2] * raw[2]);
        float k = inverseSqrt(magSquared);
        for (int c = 0; c < 4; c++)
            out[4 * i + c] = raw[c] * k;
    }
}
// core/src/math/frustum.cc:70-84
inline void frustumCorners(const Mat4 &inv, float out[24]) {
    static const float ndc[8][3] = { { -1, 1, -1 }, { 1, 1, -1 }, { -1, -1, -1 }, { 1, -1, -1 },
                                     { -1, 1, 1 }, { 1, 1, 1 }, { -1, -1, 1 }, { 1, -1, 1 } };
    for (int i = 0; i < 8; i++) {
        Vec4 h = inv * Vec4{ ndc[i][0], ndc[i][1], ndc[i][2], 1.0f };
        out[3 * i] = h.x / h.w; out[3 * i + 1] = h.y / h.w; out[3 * i + 2] = h.z / h.w;
    }
}

inline void viewPlanes(const OvisView &v, float out[24]) { frustumPlanes(viewMatrices(v).viewProjection, out); }

inline Vec3 lightDirection(const OvisLight &l) {                                          // render_light.cc:78-85
    return normalize(Vec3{ l.direction[0], l.direction[1], l.direction[2] });
}
// render/src/render_light.cc:227-249
inline void spotAabb(const OvisLight &l, float out[6]) {
    float radius = l.range * tanf(radians(l.cutoff));
    Quat orientation = quatRotateBetween(Vec3{ 0.0f, 0.0f, -1.0f }, lightDirection(l));
    Mat4 volume = transformMatrix(Vec3{ l.position[0], l.position[1], l.position[2] }, orientation,
                                  Vec3{ radius, radius, l.range });
    const float base[6] = { -1.0f, -1.0f, -1.0f, 1.0f, 1.0f, 0.0f };
    transformAabb(base, volume, out);
}
// render/src/render_light.cc:189 - "!m_intensity || !glm::length(m_colour)"
inline bool lightHasNoEffect(const OvisLight &l) {
    Vec3 c = { l.colour[0], l.colour[1], l.colour[2] };
    return !l.intensity || !std::sqrt(dot(c, c));
}
constexpr float kShadowViewNear = 0.1f;                                                   // render_light.cc:273,319
// render/src/render_light.cc:258-332; viewport square per deferred_render_pipeline.cc:261-265
inline int shadowViews(const OvisLight &l, int resolution, OvisView out[6]) {
    if (!(l.flags & OVIS_CASTS_SHADOW))
        return 0;
    int count = 0;
    auto fill = [&](OvisView &v, Quat o, float fov) {
        std::memcpy(v.position, l.position, sizeof v.position);
        v.orientation[0] = o.w; v.orientation[1] = o.x; v.orientation[2] = o.y; v.orientation[3] = o.z;
        v.hfov_deg = fov; v.znear = kShadowViewNear; v.zfar = l.range;
        v.vp_width = resolution; v.vp_height = resolution;
    };
    if (l.type == OVIS_LIGHT_SPOT) {
        fill(out[0], quatLookAt(lightDirection(l)), l.cutoff * 2);                        // :270-273
        count = 1;
    } else if (l.type == OVIS_LIGHT_POINT) {
        static const Vec3 dir[6] = { { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } };
        static const Vec3 up[6] = { { 0, -1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 }, { 0, -1, 0 }, { 0, -1, 0 } };
        for (int i = 0; i < 6; i++)                                                       // :298-320
            fill(out[i], quatLookAt(dir[i], up[i]), 90.0f);
        count = 6;
    }
    return count;
}

// ViewUniforms (render/include/render_view.h:28-36), 288 bytes: 4 x mat4, ivec2 viewportPosition, ivec2 viewportSize,
// vec3 position (+4 bytes of padding); filled by render_view.cc:55 (position), :79-80 (viewport), :113-133 (matrices).
inline void viewUniforms(const OvisView &v, const int32_t viewportPos[2], void *out288) {
    float *f = static_cast<float *>(out288);
    int32_t *i = static_cast<int32_t *>(out288);
    std::memset(out288, 0, 288);
    ViewMatrices m = viewMatrices(v);
    std::memcpy(f, &m.view, 64);
    std::memcpy(f + 16, &m.projection, 64);
    std::memcpy(f + 32, &m.viewProjection, 64);
    std::memcpy(f + 48, &m.inverseViewProjection, 64);
    i[64] = viewportPos[0]; i[65] = viewportPos[1];
    i[66] = v.vp_width; i[67] = v.vp_height;
    std::memcpy(f + 68, v.position, 12);
}

// LightUniforms (render/include/render_light.h:33-47), 208 bytes, in the state the graphics Light component leaves a
// RenderLight in (graphics/src/light.cc:35-45,58-74,100-108). The block starts zeroed (uniform_buffer.cc:136-137) and
// fields a light type never writes stay zero.
inline void lightUniforms(const OvisLight &l, const float attenuation[3], float shadowBias, void *out208) {
    float *f = static_cast<float *>(out208);
    std::memset(out208, 0, 208);
    Vec3 pos = { l.position[0], l.position[1], l.position[2] };
    Vec3 dir = lightDirection(l);
    std::memcpy(f, l.position, 12);                                                       // render_light.cc:70
    f[3] = l.intensity;                                                                   // :98
    f[4] = dir.x; f[5] = dir.y; f[6] = dir.z;                                             // :80
    std::memcpy(f + 8, l.colour, 12);                                                     // :91
    f[11] = l.range;                                                                      // :120, :336
    f[48] = shadowBias;                                                                   // :158
    if (l.type == OVIS_LIGHT_POINT) {
        Mat4 volume = transformMatrix(pos, Quat{ 1.0f, 0.0f, 0.0f, 0.0f }, Vec3{ l.range, l.range, l.range });   // :222-226
        std::memcpy(f + 12, &volume, 64);
        std::memcpy(f + 45, attenuation, 12);                                             // :136-138
        if (l.flags & OVIS_CASTS_SHADOW)
            f[44] = 0.1f;                                                                 // :324
    } else if (l.type == OVIS_LIGHT_SPOT) {
        f[7] = cosf(radians(l.cutoff));                                                   // :110
        float radius = l.range * tanf(radians(l.cutoff));                                 // :235
        Quat orientation = quatRotateBetween(Vec3{ 0.0f, 0.0f, -1.0f }, dir);             // :239
        Mat4 volume = transformMatrix(pos, orientation, Vec3{ radius, radius, l.range });
        std::memcpy(f + 12, &volume, 64);                                                 // :242
        std::memcpy(f + 45, attenuation, 12);
        OvisView sv[6];
        if (shadowViews(l, 1, sv) == 1) {                                                 // aspect 1: render_view.cc:39
            ViewMatrices m = viewMatrices(sv[0]);
            Mat4 bias;                                                                    // :284-288
            bias.col[0] = { 0.5f, 0.0f, 0.0f, 0.0f };
            bias.col[1] = { 0.0f, 0.5f, 0.0f, 0.0f };
            bias.col[2] = { 0.0f, 0.0f, 1.0f, 0.0f };
            bias.col[3] = { 0.5f, 0.5f, 0.0f, 1.0f };
            Mat4 shadowSpace = bias * m.projection * m.view;                              // :289-290
            std::memcpy(f + 28, &shadowSpace, 64);
        }
    }
}

// Entity::transformed (engine/src/entity.cc:382-411): world TRS of a child from its parent's world TRS. quat * vec3 is
// glm/gtc/quaternion.inl:313-321, quat * quat :256-270 (left-to-right sums, no renormalisation).
struct Trs {
    Vec3 p;
    Quat q;
    Vec3 s;
};
inline Vec3 rotate(Quat q, Vec3 v) {
    Vec3 u = { q.x, q.y, q.z };
    Vec3 uv = cross(u, v);
    Vec3 uuv = cross(u, uv);
    return v + ((uv * q.w) + uuv) * 2.0f;
}
inline Quat mul(Quat p, Quat q) {
    Quat r;
    r.w = p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z;
    r.x = p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y;
    r.y = p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z;
    r.z = p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x;
    return r;
}
inline Trs composeTrs(const Trs &parent, const Trs &local) {
    Trs w;
    Vec3 scaled = { parent.s.x * local.p.x, parent.s.y * local.p.y, parent.s.z * local.p.z };
    w.p = rotate(parent.q, scaled) + parent.p;                                            // entity.cc:394
    w.q = mul(parent.q, local.q);                                                         // :395
    w.s = { parent.s.x * local.s.x, parent.s.y * local.s.y, parent.s.z * local.s.z };     // :396
    return w;
}

}  // namespace hm
}  // namespace ovis
