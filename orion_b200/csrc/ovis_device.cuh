// ovis_device.cuh - device-side arithmetic shared by the kernels.
//
// Everything on a result path is IEEE binary32 round-to-nearest with NO contraction: the library is
// compiled with -fmad=false and the helpers below additionally spell the operations with
// __fmul_rn / __fadd_rn / __fsub_rn (which the compiler never fuses), in the association order of the
// reference's GLM 0.9.8.0 scalar code. Citations are relative to /root/reference/engine/runtime/
// ("glm/" = /root/reference/engine/3rdparty/glm/repo/glm/).
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace ovis {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
// glm/detail/func_common.inl:15-28 - ternaries (NaN and signed-zero behaviour differs from fminf/fmaxf)
__device__ __forceinline__ float glm_min(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float glm_max(float a, float b) { return a > b ? a : b; }

struct DVec3 {
    float x, y, z;
};
struct DVec4 {
    float x, y, z, w;
};
struct DQuat {
    float w, x, y, z;
};
struct DTrs {
    DVec3 p;
    DQuat q;
    DVec3 s;
};

__device__ __forceinline__ DVec4 scale4(DVec4 v, float s) { return { mul(v.x, s), mul(v.y, s), mul(v.z, s), mul(v.w, s) }; }
__device__ __forceinline__ DVec4 add4(DVec4 a, DVec4 b) { return { add(a.x, b.x), add(a.y, b.y), add(a.z, b.z), add(a.w, b.w) }; }

// glm/detail/func_geometric.inl:73-85
__device__ __forceinline__ DVec3 cross3(DVec3 a, DVec3 b) {
    return { sub(mul(a.y, b.z), mul(b.y, a.z)), sub(mul(a.z, b.x), mul(b.z, a.x)), sub(mul(a.x, b.y), mul(b.x, a.y)) };
}

// Entity::transformed, engine/src/entity.cc:394-396 (operators: glm/gtc/quaternion.inl:313-321, :256-270)
__device__ __forceinline__ DTrs compose_trs(const DTrs &parent, const DTrs &local) {
    DTrs w;
    // position = (parentOrientation * (parentScale * position)) + parentPosition
    DVec3 t = { mul(parent.s.x, local.p.x), mul(parent.s.y, local.p.y), mul(parent.s.z, local.p.z) };
    DVec3 u = { parent.q.x, parent.q.y, parent.q.z };
    DVec3 uv = cross3(u, t);
    DVec3 uuv = cross3(u, uv);
    DVec3 r = { add(t.x, mul(add(mul(uv.x, parent.q.w), uuv.x), 2.0f)),
                add(t.y, mul(add(mul(uv.y, parent.q.w), uuv.y), 2.0f)),
                add(t.z, mul(add(mul(uv.z, parent.q.w), uuv.z), 2.0f)) };
    w.p = { add(r.x, parent.p.x), add(r.y, parent.p.y), add(r.z, parent.p.z) };
    // orientation = parentOrientation * orientation (left-to-right sums)
    const DQuat &p = parent.q, &q = local.q;
    w.q.w = sub(sub(sub(mul(p.w, q.w), mul(p.x, q.x)), mul(p.y, q.y)), mul(p.z, q.z));
    w.q.x = sub(add(add(mul(p.w, q.x), mul(p.x, q.w)), mul(p.y, q.z)), mul(p.z, q.y));
    w.q.y = sub(add(add(mul(p.w, q.y), mul(p.y, q.w)), mul(p.z, q.x)), mul(p.x, q.z));
    w.q.z = sub(add(add(mul(p.w, q.z), mul(p.z, q.w)), mul(p.x, q.y)), mul(p.y, q.x));
    // scale = parentScale * scale
    w.s = { mul(parent.s.x, local.s.x), mul(parent.s.y, local.s.y), mul(parent.s.z, local.s.z) };
    return w;
}

struct DMat4 {
    DVec4 c[4];
};

// glm/detail/type_mat4x4.inl:601-620
__device__ __forceinline__ DMat4 mat_mul(const DMat4 &a, const DMat4 &b) {
    DMat4 r;
#pragma unroll
    for (int j = 0; j < 4; j++)
        r.c[j] = add4(add4(add4(scale4(a.c[0], b.c[j].x), scale4(a.c[1], b.c[j].y)), scale4(a.c[2], b.c[j].z)),
                      scale4(a.c[3], b.c[j].w));
    return r;
}

// Transform::matrix, core/src/math/transform.cc:26-35: translate(I,p) * mat4_cast(q) * scale(I,s), evaluated
// as the full 4x4 products GLM performs (so Inf/NaN inputs poison exactly the entries they poison there;
// for finite inputs the compiler folds the multiplications by literal 0/1 that are value-preserving).
__device__ __forceinline__ DMat4 trs_matrix(const DTrs &t) {
    const DVec4 i0 = { 1.0f, 0.0f, 0.0f, 0.0f }, i1 = { 0.0f, 1.0f, 0.0f, 0.0f }, i2 = { 0.0f, 0.0f, 1.0f, 0.0f },
                i3 = { 0.0f, 0.0f, 0.0f, 1.0f };
    DMat4 T;   // glm/gtc/matrix_transform.inl:10-16
    T.c[0] = i0; T.c[1] = i1; T.c[2] = i2;
    T.c[3] = add4(add4(add4(scale4(i0, t.p.x), scale4(i1, t.p.y)), scale4(i2, t.p.z)), i3);
    DMat4 R;   // glm/gtc/quaternion.inl:593-624
    const DQuat &q = t.q;
    float qxx = mul(q.x, q.x), qyy = mul(q.y, q.y), qzz = mul(q.z, q.z);
    float qxz = mul(q.x, q.z), qxy = mul(q.x, q.y), qyz = mul(q.y, q.z);
    float qwx = mul(q.w, q.x), qwy = mul(q.w, q.y), qwz = mul(q.w, q.z);
    R.c[0] = { sub(1.0f, mul(2.0f, add(qyy, qzz))), mul(2.0f, add(qxy, qwz)), mul(2.0f, sub(qxz, qwy)), 0.0f };
    R.c[1] = { mul(2.0f, sub(qxy, qwz)), sub(1.0f, mul(2.0f, add(qxx, qzz))), mul(2.0f, add(qyz, qwx)), 0.0f };
    R.c[2] = { mul(2.0f, add(qxz, qwy)), mul(2.0f, sub(qyz, qwx)), sub(1.0f, mul(2.0f, add(qxx, qyy))), 0.0f };
    R.c[3] = i3;
    DMat4 S;   // glm/gtc/matrix_transform.inl:78-87
    S.c[0] = scale4(i0, t.s.x); S.c[1] = scale4(i1, t.s.y); S.c[2] = scale4(i2, t.s.z); S.c[3] = i3;
    return mat_mul(mat_mul(T, R), S);
}

struct DAabb {
    float mn[3], mx[3];
};

// BoundingBox::transform, core/src/math/bounding_box.cc:27-40
__device__ __forceinline__ DAabb transform_aabb(const DAabb &b, const DMat4 &m) {
    DAabb o;
    const float col0[3] = { m.c[0].x, m.c[0].y, m.c[0].z }, col1[3] = { m.c[1].x, m.c[1].y, m.c[1].z };
    const float col2[3] = { m.c[2].x, m.c[2].y, m.c[2].z }, col3[3] = { m.c[3].x, m.c[3].y, m.c[3].z };
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float xa = mul(col0[k], b.mn[0]), xb = mul(col0[k], b.mx[0]);
        float ya = mul(col1[k], b.mn[1]), yb = mul(col1[k], b.mx[1]);
        float za = mul(col2[k], b.mn[2]), zb = mul(col2[k], b.mx[2]);
        o.mn[k] = add(add(add(glm_min(xa, xb), glm_min(ya, yb)), glm_min(za, zb)), col3[k]);
        o.mx[k] = add(add(add(glm_max(xa, xb), glm_max(ya, yb)), glm_max(za, zb)), col3[k]);
    }
    return o;
}

// One plane of Math::intersect(Frustum, BoundingBox), core/src/math/intersection.cc:49-53 with
// BoundingBox::calcPVertex (core/include/math/bounding_box.h:60-71) and Plane::distanceTo (plane.h:70-72).
// Returns true when the box is OUTSIDE this plane (d < 0; NaN compares false => inside).
__device__ __forceinline__ bool outside_plane(float nx, float ny, float nz, float d,
                                              float mnx, float mny, float mnz, float mxx, float mxy, float mxz) {
    float px = (nx >= 0.0f) ? mxx : mnx;
    float py = (ny >= 0.0f) ? mxy : mny;
    float pz = (nz >= 0.0f) ? mxz : mnz;
    float dist = sub(add(add(mul(nx, px), mul(ny, py)), mul(nz, pz)), d);
    return dist < 0.0f;
}

// ---- the same plane test for the four entities of one float4 group, specialised ---------------------------------------
// (1) The signs of a plane's normal are the same for every entity, so which corner is the P-vertex is decided once per
//     plane (a warp-uniform switch over 8 instantiations) instead of by three selects per entity.
// (2) The three products of two neighbouring entities are formed by one packed instruction each (Blackwell FMUL2:
//     mul.rn.f32x2 rounds both halves exactly like two FMULs). The additions stay scalar: ptxas contracts a packed multiply
//     feeding a packed add into FFMA2 even with -fmad=false, which would change the rounding; scalar add.rn is never fused.
__device__ __forceinline__ void mul2(float a0, float a1, float n, float &r0, float &r1) {
    unsigned long long a, b, d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(n), "f"(n));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r0), "=f"(r1) : "l"(d));
}

// Packed pairs kept packed: two floats in one 64-bit register pair. Only operations whose inputs are not products may be
// packed adds / subtracts (see above: ptxas would contract mul.f32x2 -> add.f32x2 into FFMA2); "sum - d" qualifies.
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long mul2p(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ unsigned long long sub2p(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// Signed distances of four P-vertices (already selected and packed: x01 = {p0.x, p1.x}, x23 = {p2.x, p3.x}, ...) to a plane
// given with every component doubled: qa = (nx, nx, ny, ny), qb = (nz, nz, d, d). ((x + y) + z) - d as the reference sums.
__device__ __forceinline__ float4 plane_dist4_packed(const float4 &qa, const float4 &qb, unsigned long long x01, unsigned long long x23,
                                                     unsigned long long y01, unsigned long long y23, unsigned long long z01,
                                                     unsigned long long z23) {
    const unsigned long long nx = pack2(qa.x, qa.y), ny = pack2(qa.z, qa.w), nz = pack2(qb.x, qb.y), dd = pack2(qb.z, qb.w);
    float x0, x1, x2, x3, y0, y1, y2, y3, z0, z1, z2, z3;
    unpack2(mul2p(x01, nx), x0, x1); unpack2(mul2p(x23, nx), x2, x3);
    unpack2(mul2p(y01, ny), y0, y1); unpack2(mul2p(y23, ny), y2, y3);
    unpack2(mul2p(z01, nz), z0, z1); unpack2(mul2p(z23, nz), z2, z3);
    float4 r;
    unpack2(sub2p(pack2(add(add(x0, y0), z0), add(add(x1, y1), z1)), dd), r.x, r.y);
    unpack2(sub2p(pack2(add(add(x2, y2), z2), add(add(x3, y3), z3)), dd), r.z, r.w);
    return r;
}

template <bool SX, bool SY, bool SZ>
__device__ __forceinline__ uint32_t outside_plane4_signed(float nx, float ny, float nz, float d, const float4 &mnx, const float4 &mny,
                                                          const float4 &mnz, const float4 &mxx, const float4 &mxy, const float4 &mxz) {
    const float4 &px = SX ? mxx : mnx, &py = SY ? mxy : mny, &pz = SZ ? mxz : mnz;
    float x0, x1, x2, x3, y0, y1, y2, y3, z0, z1, z2, z3;
    mul2(px.x, px.y, nx, x0, x1); mul2(px.z, px.w, nx, x2, x3);
    mul2(py.x, py.y, ny, y0, y1); mul2(py.z, py.w, ny, y2, y3);
    mul2(pz.x, pz.y, nz, z0, z1); mul2(pz.z, pz.w, nz, z2, z3);
    uint32_t out = 0;
    if (sub(add(add(x0, y0), z0), d) < 0.0f) out |= 1u;
    if (sub(add(add(x1, y1), z1), d) < 0.0f) out |= 2u;
    if (sub(add(add(x2, y2), z2), d) < 0.0f) out |= 4u;
    if (sub(add(add(x3, y3), z3), d) < 0.0f) out |= 8u;
    return out;
}

// bit j set = entity j of the group is outside the plane (same results as four outside_plane calls)
__device__ __forceinline__ uint32_t outside_plane4(float nx, float ny, float nz, float d, const float4 &mnx, const float4 &mny,
                                                   const float4 &mnz, const float4 &mxx, const float4 &mxy, const float4 &mxz) {
    const int code = (nx >= 0.0f ? 1 : 0) | (ny >= 0.0f ? 2 : 0) | (nz >= 0.0f ? 4 : 0);
    switch (code) {
    case 0: return outside_plane4_signed<false, false, false>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 1: return outside_plane4_signed<true, false, false>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 2: return outside_plane4_signed<false, true, false>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 3: return outside_plane4_signed<true, true, false>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 4: return outside_plane4_signed<false, false, true>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 5: return outside_plane4_signed<true, false, true>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    case 6: return outside_plane4_signed<false, true, true>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    default: return outside_plane4_signed<true, true, true>(nx, ny, nz, d, mnx, mny, mnz, mxx, mxy, mxz);
    }
}

// Signed distances of the four P-vertices of a group to one plane (the value the reference compares with 0).
template <bool SX, bool SY, bool SZ>
__device__ __forceinline__ float4 plane_dist4_signed(float nx, float ny, float nz, float d, const float4 &mnx, const float4 &mny,
                                                     const float4 &mnz, const float4 &mxx, const float4 &mxy, const float4 &mxz) {
    const float4 &px = SX ? mxx : mnx, &py = SY ? mxy : mny, &pz = SZ ? mxz : mnz;
    float x0, x1, x2, x3, y0, y1, y2, y3, z0, z1, z2, z3;
    mul2(px.x, px.y, nx, x0, x1); mul2(px.z, px.w, nx, x2, x3);
    mul2(py.x, py.y, ny, y0, y1); mul2(py.z, py.w, ny, y2, y3);
    mul2(pz.x, pz.y, nz, z0, z1); mul2(pz.z, pz.w, nz, z2, z3);
    return make_float4(sub(add(add(x0, y0), z0), d), sub(add(add(x1, y1), z1), d), sub(add(add(x2, y2), z2), d),
                       sub(add(add(x3, y3), z3), d));
}

// An entity is rejected iff some plane's distance is < 0, i.e. iff the minimum over the planes is < 0 - with fminf's rule
// that a NaN operand is ignored, which is exactly "a NaN distance rejects nothing" (d < 0 is false for NaN). Folding the
// six compares into minima (FMNMX3 takes two new values at a time) leaves one compare per entity.
__device__ __forceinline__ float4 min3_4(const float4 &m, const float4 &a, const float4 &b) {
    return make_float4(fminf(fminf(m.x, a.x), b.x), fminf(fminf(m.y, a.y), b.y), fminf(fminf(m.z, a.z), b.z), fminf(fminf(m.w, a.w), b.w));
}

// plane_dist4_signed chosen by a 3-bit code (bit c = normal component c >= 0): a warp-uniform switch
__device__ __forceinline__ float4 plane_dist4_coded(uint32_t code, float4 q, const float4 &mnx, const float4 &mny, const float4 &mnz,
                                                    const float4 &mxx, const float4 &mxy, const float4 &mxz) {
    switch (code & 7u) {
    case 0: return plane_dist4_signed<false, false, false>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 1: return plane_dist4_signed<true, false, false>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 2: return plane_dist4_signed<false, true, false>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 3: return plane_dist4_signed<true, true, false>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 4: return plane_dist4_signed<false, false, true>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 5: return plane_dist4_signed<true, false, true>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    case 6: return plane_dist4_signed<false, true, true>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    default: return plane_dist4_signed<true, true, true>(q.x, q.y, q.z, q.w, mnx, mny, mnz, mxx, mxy, mxz);
    }
}

// ---- results of one group of 128 entities (4 per lane) ---------------------------------------------------------------
// keep: bit j = entity 4 * lane + j of the group is visible. The 128 bits leave as four bitmap words, assembled with
// three butterfly shuffles (lanes 8k .. 8k+7 own word k).
__device__ __forceinline__ void store_group_bits(uint32_t *survivors, int64_t group, uint32_t keep, int lane) {
    uint32_t v = keep << ((lane & 7) * 4);
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    if ((lane & 7) == 0)
        survivors[group * 4 + (lane >> 3)] = v;
}

// bits of the lane's four entities that exist (ragged tail) and, if MASKED, are set in the alive / caster mask
template <bool MASKED>
__device__ __forceinline__ uint32_t candidate_bits(int64_t e0, int64_t n, uint32_t mask_word) {
    uint32_t k = 0xfu;
    if (MASKED)
        k = (mask_word >> (e0 & 31)) & 0xfu;
    if (e0 + 3 >= n) {
        const int64_t valid = n - e0;
        k &= valid <= 0 ? 0u : ((1u << valid) - 1u);
    }
    return k;
}

__device__ __forceinline__ uint32_t negative_bits(const float4 &lowest) {
    uint32_t out = 0;
    if (lowest.x < 0.0f) out |= 1u;
    if (lowest.y < 0.0f) out |= 2u;
    if (lowest.z < 0.0f) out |= 4u;
    if (lowest.w < 0.0f) out |= 8u;
    return out;
}

}  // namespace ovis
