/*
 * oracle/ovis_raster.c - TEST INFRASTRUCTURE. CPU restatement of the G-buffer pass and the shadow-map passes of the
 * reference's deferred pipeline. Nothing in the product may call this; only tests/ do.
 *
 * Restated from the reference, statement for statement, in float arithmetic:
 *   engine/shaders/lit_vtx.glsl:41-58            vtxNormal = vec3(entity.transform * vec4(attribNormal, 0.0));
 *                                                gl_Position = view.viewProjection * entity.transform * vec4(attribPosition, 1.0)
 *   engine/shaders/shadow_vtx.glsl:25            the same gl_Position
 *   engine/shaders/lit_deferred_frag.glsl:45-74  the variant without TEXTURED / NORMALMAP: bufferA = (normalize(normal) + 1) / 2,
 *                                                bufferB = diffuseColour, bufferC = specularColour / shininess or (0, 0, 0, 1)
 *   engine/shaders/shadow_frag.glsl              (empty: depth only)
 * with GLM's operation order (detail/type_mat4x4.inl: mat4 * mat4 left to right over the columns, mat4 * vec4 as
 * (m0 x + m1 y) + (m2 z + m3 w); func_geometric.inl: normalize = v * inversesqrt(dot(v, v))), and the pipeline state of
 * render/src/deferred_render_pipeline.cc:335-370, 384-412 with gpu/include/state.h:105-145: clear depth 1 and colours 0,
 * depth test LessOrEqual with depth write, back faces culled, counter-clockwise front faces, primitives in list order.
 *
 * PARITY UNPINNED by reference code for the fixed-function part: the reference hands rasterisation to OpenGL / Vulkan,
 * which is unavailable on this box, and those APIs leave sub-pixel snapping and interpolation precision to the
 * implementation. What this file fixes as the definition (the same rules the CUDA path documents):
 *   - window coordinates with the origin at the bottom left, pixel centres at (x + 0.5, y + 0.5), row 0 = bottom row;
 *   - coverage by 2-D homogeneous edge functions in float (rows of the adjugate of [X; Y; w] with X = (x_c + w_c) W / 2,
 *     Y = (y_c + w_c) H / 2), a centre exactly on an edge belongs to the triangle for which that edge is a top or left one;
 *     no geometric clipping: a triangle that crosses the near plane is rasterised as it is;
 *   - with all three vertices in front of the eye (w > 0) only the centres inside the bounding box of the projected
 *     vertices, widened by a pixel, are tested (setup() below: the box is part of the definition, not an optimisation -
 *     the edge functions of a triangle far smaller than a pixel are cancellation noise away from it);
 *   - depth = (sum E_i z_i) / (sum E_i w_i) with the clip-space z and w of the vertices (perspective-correct z / w on a
 *     0..1 depth range), fragments with depth outside [0, 1] or a non-positive w are discarded (near / far clip);
 *   - attributes interpolated with the weights E_i / (E_0 + E_1 + E_2);
 *   - equal depths: the later primitive wins (LessOrEqual in submission order).
 * The CUDA path is compared with this file value for value in tests/test_raster_gpu.py.
 */
#include "ovis_oracle.h"

#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static void mat4_mul(const float *A, const float *B, float *R) {
    for (int j = 0; j < 4; j++)
        for (int r = 0; r < 4; r++)
            R[4 * j + r] = ((A[r] * B[4 * j] + A[4 + r] * B[4 * j + 1]) + A[8 + r] * B[4 * j + 2]) + A[12 + r] * B[4 * j + 3];
}

static void mat4_vec4(const float *M, float x, float y, float z, float w, float out[4]) {
    for (int r = 0; r < 4; r++)
        out[r] = (M[r] * x + M[4 + r] * y) + (M[8 + r] * z + M[12 + r] * w);
}

/* the two shader stages on their own, for comparison with the reference's shader source (oracle/glsl_harness.cc) */
void orc_raster_vertex(const float *transform, const float *view_projection, const float *position, const float *normal,
                       float *out_clip, float *out_normal) {
    float mvp[16], nrm[4];
    mat4_mul(view_projection, transform, mvp);
    mat4_vec4(mvp, position[0], position[1], position[2], 1.0f, out_clip);
    mat4_vec4(transform, normal[0], normal[1], normal[2], 0.0f, nrm);
    out_normal[0] = nrm[0]; out_normal[1] = nrm[1]; out_normal[2] = nrm[2];
}

void orc_raster_fragment(const float *n, const float *mat, float *a, float *b, float *c) {
    const float inv = 1.0f / sqrtf((n[0] * n[0] + n[1] * n[1]) + n[2] * n[2]);
    a[0] = (n[0] * inv + 1.0f) / 2.0f; a[1] = (n[1] * inv + 1.0f) / 2.0f; a[2] = (n[2] * inv + 1.0f) / 2.0f; a[3] = 1.0f;
    b[0] = mat[0]; b[1] = mat[1]; b[2] = mat[2]; b[3] = 1.0f;
    if (mat[7] != 0.0f) {
        c[0] = mat[3]; c[1] = mat[4]; c[2] = mat[5]; c[3] = mat[6];
    } else {
        c[0] = c[1] = c[2] = 0.0f; c[3] = 1.0f;
    }
}

/* vtxBitangent = cross(vtxNormal, vtxTangent) * attribTangent.w (lit_vtx.glsl:55); GLM cross:
 * (a.y b.z - b.y a.z, a.z b.x - b.z a.x, a.x b.y - b.x a.y) */
void orc_raster_bitangent(const float *n, const float *t, float w, float *out) {
    out[0] = (n[1] * t[2] - t[1] * n[2]) * w;
    out[1] = (n[2] * t[0] - t[2] * n[0]) * w;
    out[2] = (n[0] * t[1] - t[0] * n[1]) * w;
}

typedef struct {
    float a[3], b[3], c[3], z[3], w[3], n[9];
    float uv[6], tg[9], bt[9];                   /* vtxTexcoord, vtxTangent, vtxBitangent (lit_vtx.glsl:41-55) */
    int draw, x0, x1, y0, y1;
} tri_t;

static void setup_tan(const float *mvp, const float *model, const float *v0, const float *v1, const float *v2, const float *t0,
                      const float *t1, const float *t2, int width, int height, tri_t *t);

static void setup(const float *mvp, const float *model, const float *v0, const float *v1, const float *v2, int width, int height, tri_t *t) {
    static const float zero[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
    setup_tan(mvp, model, v0, v1, v2, zero, zero, zero, width, height, t);
}

static void setup_tan(const float *mvp, const float *model, const float *v0, const float *v1, const float *v2, const float *t0,
                      const float *t1, const float *t2, int width, int height, tri_t *t) {
    const float *vtx[3] = { v0, v1, v2 };
    const float *tan4[3] = { t0, t1, t2 };
    float X[3], Y[3];
    const float half_w = 0.5f * (float)width, half_h = 0.5f * (float)height;
    for (int i = 0; i < 3; i++) {
        float clip[4], nrm[4];
        mat4_vec4(mvp, vtx[i][0], vtx[i][1], vtx[i][2], 1.0f, clip);            /* = orc_raster_vertex with the product hoisted */
        mat4_vec4(model, vtx[i][3], vtx[i][4], vtx[i][5], 0.0f, nrm);
        X[i] = (clip[0] + clip[3]) * half_w;
        Y[i] = (clip[1] + clip[3]) * half_h;
        t->z[i] = clip[2];
        t->w[i] = clip[3];
        t->n[3 * i] = nrm[0]; t->n[3 * i + 1] = nrm[1]; t->n[3 * i + 2] = nrm[2];
        {
            float tw[4];
            mat4_vec4(model, tan4[i][0], tan4[i][1], tan4[i][2], 0.0f, tw);
            t->uv[2 * i] = vtx[i][6]; t->uv[2 * i + 1] = vtx[i][7];
            t->tg[3 * i] = tw[0]; t->tg[3 * i + 1] = tw[1]; t->tg[3 * i + 2] = tw[2];
            orc_raster_bitangent(nrm, tw, tan4[i][3], t->bt + 3 * i);
        }
    }
    for (int i = 0; i < 3; i++) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
        t->a[i] = Y[i1] * t->w[i2] - Y[i2] * t->w[i1];
        t->b[i] = X[i2] * t->w[i1] - X[i1] * t->w[i2];
        t->c[i] = X[i1] * Y[i2] - X[i2] * Y[i1];
    }
    const float det = (X[0] * t->a[0] + Y[0] * t->b[0]) + t->w[0] * t->c[0];
    t->draw = det > 0.0f;
    if (t->z[0] < 0.0f && t->z[1] < 0.0f && t->z[2] < 0.0f) t->draw = 0;
    if (t->z[0] > t->w[0] && t->z[1] > t->w[1] && t->z[2] > t->w[2]) t->draw = 0;
    /* The pixels a triangle may touch are part of the definition: with all three vertices in front of the eye, only the
     * centres inside the box of the projected vertices (one pixel of slack for the rounding of the divisions) are tested.
     * Without the box a triangle far smaller than a pixel, whose edge functions are all cancellation noise, lights up
     * centres nowhere near it. A triangle with a vertex at or behind the eye has no finite box: every centre is tested. */
    t->x0 = 0; t->x1 = width - 1; t->y0 = 0; t->y1 = height - 1;
    if (t->draw && t->w[0] > 0.0f && t->w[1] > 0.0f && t->w[2] > 0.0f) {
        const float px[3] = { X[0] / t->w[0], X[1] / t->w[1], X[2] / t->w[2] };
        const float py[3] = { Y[0] / t->w[0], Y[1] / t->w[1], Y[2] / t->w[2] };
        const float lx = fminf(fminf(px[0], px[1]), px[2]), hx = fmaxf(fmaxf(px[0], px[1]), px[2]);
        const float ly = fminf(fminf(py[0], py[1]), py[2]), hy = fmaxf(fmaxf(py[0], py[1]), py[2]);
        if (!(hx >= 0.0f) || !(lx <= (float)width) || !(hy >= 0.0f) || !(ly <= (float)height)) {
            t->draw = 0;
        } else {
            /* (clamped as floats first: a float beyond int's range converts to anything in C; CUDA's conversion saturates) */
            const int bx0 = (int)fmaxf(floorf(lx - 0.5f), 0.0f), bx1 = (int)fminf(ceilf(hx - 0.5f), (float)(width - 1));
            const int by0 = (int)fmaxf(floorf(ly - 0.5f), 0.0f), by1 = (int)fminf(ceilf(hy - 0.5f), (float)(height - 1));
            if (bx0 > t->x0) t->x0 = bx0;
            if (bx1 < t->x1) t->x1 = bx1;
            if (by0 > t->y0) t->y0 = by0;
            if (by1 < t->y1) t->y1 = by1;
        }
    }
}

static int sample(const tri_t *t, int px, int py, float e[3], float *depth) {
    const float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
    int inside = 1;
    for (int i = 0; i < 3; i++) {
        e[i] = (t->a[i] * fx + t->b[i] * fy) + t->c[i];
        inside = inside && (e[i] > 0.0f || (e[i] == 0.0f && (t->a[i] > 0.0f || (t->a[i] == 0.0f && t->b[i] > 0.0f))));
    }
    const float zn = (e[0] * t->z[0] + e[1] * t->z[1]) + e[2] * t->z[2];
    const float wn = (e[0] * t->w[0] + e[1] * t->w[1]) + e[2] * t->w[2];
    *depth = zn / wn;
    return inside && wn > 0.0f && *depth >= 0.0f && *depth <= 1.0f;
}

/* instances: per drawn entity, in draw order: matrix[16 n] (EntityUniforms::transform, column-major), mesh[3 n] =
 * (first vertex, first index, triangles), material[8 n] = (diffuse rgb, specular rgb, shininess, flags as a float).
 * vertices: 8 floats each (position, normal, texcoord). vp: viewProjection, column-major. Outputs row-major, row 0 at the
 * bottom; out_a / out_b / out_c may be NULL for a depth-only pass (a shadow map). */
void orc_raster(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                const uint32_t *indices, const float *vp, int32_t width, int32_t height, float *out_a, float *out_b, float *out_c,
                float *out_depth) {
    const int64_t pixels = (int64_t)width * height;
    for (int64_t p = 0; p < pixels; p++) {
        out_depth[p] = 1.0f;
        if (out_a)
            for (int c = 0; c < 4; c++)
                out_a[4 * p + c] = out_b[4 * p + c] = out_c[4 * p + c] = 0.0f;
    }
    for (int64_t k = 0; k < n_instances; k++) {
        const float *M = matrix + 16 * k;
        float mvp[16];
        mat4_mul(vp, M, mvp);
        const int64_t first_vertex = mesh[3 * k], first_index = mesh[3 * k + 1], n_tri = mesh[3 * k + 2];
        const float *mat = material + 8 * k;
        for (int64_t tix = 0; tix < n_tri; tix++) {
            tri_t t;
            const uint32_t *ix = indices + first_index + 3 * tix;
            setup(mvp, M, vertices + 8 * (first_vertex + ix[0]), vertices + 8 * (first_vertex + ix[1]), vertices + 8 * (first_vertex + ix[2]),
                  width, height, &t);
            if (!t.draw)
                continue;
            for (int py = t.y0; py <= t.y1; py++)
                for (int px = t.x0; px <= t.x1; px++) {
                    float e[3], depth;
                    if (!sample(&t, px, py, e, &depth))
                        continue;
                    const int64_t p = (int64_t)py * width + px;
                    if (!(depth <= out_depth[p]))                    /* LessOrEqual: a later primitive of equal depth passes */
                        continue;
                    out_depth[p] = depth;
                    if (!out_a)
                        continue;
                    const float sum = (e[0] + e[1]) + e[2];
                    const float b0 = e[0] / sum, b1 = e[1] / sum, b2 = e[2] / sum;
                    float n[3];
                    for (int c = 0; c < 3; c++)
                        n[c] = (b0 * t.n[c] + b1 * t.n[3 + c]) + b2 * t.n[6 + c];
                    orc_raster_fragment(n, mat, out_a + 4 * p, out_b + 4 * p, out_c + 4 * p);
                }
        }
    }
}

/* DeferredRenderPipeline::renderBasic (render/src/deferred_render_pipeline.cc:506-522) with engine/shaders/unlit_vtx.glsl
 * (the same gl_Position as lit_vtx.glsl) and unlit_frag.glsl without TEXTURED (fragColour = vec4(diffuseColour, 1.0)):
 * the instances are drawn into rgba / depth as they stand (what the light pass and the G-buffer pass left), depth test
 * LessOrEqual with write, no blending. */
void orc_raster_basic(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                      const uint32_t *indices, const float *vp, int32_t width, int32_t height, float *rgba, float *depth_buffer) {
    for (int64_t k = 0; k < n_instances; k++) {
        const float *M = matrix + 16 * k;
        float mvp[16];
        mat4_mul(vp, M, mvp);
        const int64_t first_vertex = mesh[3 * k], first_index = mesh[3 * k + 1], n_tri = mesh[3 * k + 2];
        const float *mat = material + 8 * k;
        for (int64_t tix = 0; tix < n_tri; tix++) {
            tri_t t;
            const uint32_t *ix = indices + first_index + 3 * tix;
            setup(mvp, M, vertices + 8 * (first_vertex + ix[0]), vertices + 8 * (first_vertex + ix[1]), vertices + 8 * (first_vertex + ix[2]),
                  width, height, &t);
            if (!t.draw)
                continue;
            for (int py = t.y0; py <= t.y1; py++)
                for (int px = t.x0; px <= t.x1; px++) {
                    float e[3], depth;
                    if (!sample(&t, px, py, e, &depth))
                        continue;
                    const int64_t p = (int64_t)py * width + px;
                    if (!(depth <= depth_buffer[p]))
                        continue;
                    depth_buffer[p] = depth;
                    rgba[4 * p] = mat[0]; rgba[4 * p + 1] = mat[1]; rgba[4 * p + 2] = mat[2]; rgba[4 * p + 3] = 1.0f;
                }
        }
    }
}

/* ---- textures: the sampler this library defines (include/orion_vis.h, ovis_raster_add_texture) ---------------------------
 * texels decoded once: c / 255, colour channels of an sRGB texture through the transfer function of the GL specification;
 * texture(sampler2D, uv) = bilinear on the base level, texel centres at (i + 0.5) / size, repeat or clamp addressing. */
typedef struct {
    const uint8_t *rgba;
    int32_t width, height, srgb, wrap;
} orc_texture;

static int tex_coord(int i, int size, int wrap) {
    if (wrap) {
        i %= size;
        return i < 0 ? i + size : i;
    }
    return i < 0 ? 0 : (i > size - 1 ? size - 1 : i);
}

static void texel(const orc_texture *t, const float *linear, const float *decoded, int i, int j, float out[4]) {
    const uint8_t *p = t->rgba + 4 * ((int64_t)j * t->width + i);
    const float *table = t->srgb ? decoded : linear;
    out[0] = table[p[0]]; out[1] = table[p[1]]; out[2] = table[p[2]]; out[3] = linear[p[3]];
}

void orc_texture_sample(const void *texture, float u, float v, float out[4]) {
    const orc_texture *t = (const orc_texture *)texture;
    static float linear[256], decoded[256];
    static int tables_ready = 0;
    if (!tables_ready) {
        for (int i = 0; i < 256; i++) {
            linear[i] = (float)i / 255.0f;
            decoded[i] = linear[i] <= 0.04045f ? linear[i] / 12.92f : powf((linear[i] + 0.055f) / 1.055f, 2.4f);
        }
        tables_ready = 1;
    }
    float x = u * (float)t->width - 0.5f, y = v * (float)t->height - 0.5f;
    if (!(fabsf(x) < 1.0e9f)) x = 0.0f;
    if (!(fabsf(y) < 1.0e9f)) y = 0.0f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float fx = x - fx0, fy = y - fy0;
    const int i0 = tex_coord((int)fx0, t->width, t->wrap), i1 = tex_coord((int)fx0 + 1, t->width, t->wrap);
    const int j0 = tex_coord((int)fy0, t->height, t->wrap), j1 = tex_coord((int)fy0 + 1, t->height, t->wrap);
    float t00[4], t10[4], t01[4], t11[4];
    texel(t, linear, decoded, i0, j0, t00); texel(t, linear, decoded, i1, j0, t10);
    texel(t, linear, decoded, i0, j1, t01); texel(t, linear, decoded, i1, j1, t11);
    const float gx = 1.0f - fx, gy = 1.0f - fy;
    for (int c = 0; c < 4; c++)
        out[c] = (t00[c] * gx + t10[c] * fx) * gy + (t01[c] * gx + t11[c] * fx) * fy;
}

static void normalize3(float v[3]) {
    const float inv = 1.0f / sqrtf((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
    v[0] = v[0] * inv; v[1] = v[1] * inv; v[2] = v[2] * inv;
}

/* lit_deferred_frag.glsl with whatever of SPECULAR / TEXTURED / NORMALMAP the material's flags name. material: 10 floats =
 * diffuse rgb, specular rgb, shininess, flags, diffuse texture, normal texture (the last three as numbers). */
void orc_raster_fragment_textured(const float *normal, const float *uv, const float *tangent, const float *bitangent, const float *mat,
                                  const void *textures, float *a, float *b, float *c) {
    const orc_texture *tex = (const orc_texture *)textures;
    const int flags = (int)mat[7];
    float n[3] = { normal[0], normal[1], normal[2] };
    float m8[8] = { mat[0], mat[1], mat[2], mat[3], mat[4], mat[5], mat[6], (flags & 1) ? 1.0f : 0.0f };
    if (flags & 2) {
        float d[4];
        orc_texture_sample(tex + (int)mat[8], uv[0], uv[1], d);
        m8[0] = d[0]; m8[1] = d[1]; m8[2] = d[2];
    }
    if (flags & 4) {
        float s4[4];
        orc_texture_sample(tex + (int)mat[9], uv[0], uv[1], s4);
        float sn[3] = { s4[0] * 2.0f - 1.0f, s4[1] * 2.0f - 1.0f, s4[2] * 2.0f - 1.0f };
        normalize3(sn);
        float T[3] = { tangent[0], tangent[1], tangent[2] }, B[3] = { bitangent[0], bitangent[1], bitangent[2] }, N[3] = { n[0], n[1], n[2] };
        normalize3(T); normalize3(B); normalize3(N);
        for (int k = 0; k < 3; k++)
            n[k] = (T[k] * sn[0] + B[k] * sn[1]) + N[k] * sn[2];
    }
    orc_raster_fragment(n, m8, a, b, c);
}

/* orc_raster with textures: tangents = 4 floats per vertex (parallel to vertices), material = 10 floats per instance,
 * textures = n_textures orc_texture records. */
void orc_raster_textured(int64_t n_instances, const float *matrix, const int64_t *mesh, const float *material, const float *vertices,
                         const float *tangents, const uint32_t *indices, const void *textures, const float *vp, int32_t width,
                         int32_t height, float *out_a, float *out_b, float *out_c, float *out_depth) {
    const int64_t pixels = (int64_t)width * height;
    for (int64_t p = 0; p < pixels; p++) {
        out_depth[p] = 1.0f;
        for (int c = 0; c < 4; c++)
            out_a[4 * p + c] = out_b[4 * p + c] = out_c[4 * p + c] = 0.0f;
    }
    for (int64_t k = 0; k < n_instances; k++) {
        const float *M = matrix + 16 * k;
        float mvp[16];
        mat4_mul(vp, M, mvp);
        const int64_t first_vertex = mesh[3 * k], first_index = mesh[3 * k + 1], n_tri = mesh[3 * k + 2];
        const float *mat = material + 10 * k;
        for (int64_t tix = 0; tix < n_tri; tix++) {
            tri_t t;
            const uint32_t *ix = indices + first_index + 3 * tix;
            const int64_t v0 = first_vertex + ix[0], v1 = first_vertex + ix[1], v2 = first_vertex + ix[2];
            setup_tan(mvp, M, vertices + 8 * v0, vertices + 8 * v1, vertices + 8 * v2, tangents + 4 * v0, tangents + 4 * v1, tangents + 4 * v2,
                      width, height, &t);
            if (!t.draw)
                continue;
            for (int py = t.y0; py <= t.y1; py++)
                for (int px = t.x0; px <= t.x1; px++) {
                    float e[3], depth;
                    if (!sample(&t, px, py, e, &depth))
                        continue;
                    const int64_t p = (int64_t)py * width + px;
                    if (!(depth <= out_depth[p]))
                        continue;
                    out_depth[p] = depth;
                    const float sum = (e[0] + e[1]) + e[2];
                    const float b0 = e[0] / sum, b1 = e[1] / sum, b2 = e[2] / sum;
                    float n[3], T[3], B[3], uv[2];
                    for (int c = 0; c < 3; c++) {
                        n[c] = (b0 * t.n[c] + b1 * t.n[3 + c]) + b2 * t.n[6 + c];
                        T[c] = (b0 * t.tg[c] + b1 * t.tg[3 + c]) + b2 * t.tg[6 + c];
                        B[c] = (b0 * t.bt[c] + b1 * t.bt[3 + c]) + b2 * t.bt[6 + c];
                    }
                    uv[0] = (b0 * t.uv[0] + b1 * t.uv[2]) + b2 * t.uv[4];
                    uv[1] = (b0 * t.uv[1] + b1 * t.uv[3]) + b2 * t.uv[5];
                    orc_raster_fragment_textured(n, uv, T, B, mat, textures, out_a + 4 * p, out_b + 4 * p, out_c + 4 * p);
                }
        }
    }
}
