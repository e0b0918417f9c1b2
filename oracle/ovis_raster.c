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

typedef struct {
    float a[3], b[3], c[3], z[3], w[3], n[9];
    int draw, x0, x1, y0, y1;
} tri_t;

static void setup(const float *mvp, const float *model, const float *v0, const float *v1, const float *v2, int width, int height, tri_t *t) {
    const float *vtx[3] = { v0, v1, v2 };
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
    /* every pixel is visited here; the bounding box only has to be one the CUDA path's box is inside of */
    t->x0 = 0; t->x1 = width - 1; t->y0 = 0; t->y1 = height - 1;
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
            for (int py = 0; py < height; py++)
                for (int px = 0; px < width; px++) {
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
