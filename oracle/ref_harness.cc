/*
 * oracle/ref_harness.cc - C entry points over the reference's OWN compiled classes.
 *
 * TEST INFRASTRUCTURE ONLY. Linked with the unmodified reference sources named in oracle/Makefile
 * into oracle/_ref/liborion_ref.so. Every function here is a thin driver: the arithmetic that
 * decides a result is executed by the reference's code (and its vendored GLM 0.9.8.0), not by
 * anything written in this file. The two places where reference code cannot be linked and its
 * control flow is restated with the same GLM operators are marked RESTATED.
 *
 *   tier A (oref_world_*): SimpleRenderWorld + RenderEntity + RenderLight + RenderView as shipped.
 *   tier B (orc_*):        the flat API of ovis_oracle.h driven through Math::intersect,
 *                          BoundingBox::transform, Transform::matrix, RenderView, RenderLight.
 */

#include "core/math/bounding_box.h"
#include "core/math/frustum.h"
#include "core/math/intersection.h"
#include "core/math/sphere.h"
#include "core/math/transform.h"

#include "render/render_entity.h"
#include "render/render_light.h"
#include "render/render_view.h"
#include "render/simple_render_world.h"

#include "render_core/geometry.h"

#include "ovis_oracle.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>
#include <unordered_map>
#include <vector>

void oref_shim_init();

namespace {

static_assert(sizeof(Plane) == 16, "Plane is one vec4");
static_assert(sizeof(BoundingBox) == 24, "BoundingBox is two vec3");

inline glm::vec3 v3(const float *p) { return glm::vec3(p[0], p[1], p[2]); }
inline glm::quat qt(const float *p) { return glm::quat(p[0], p[1], p[2], p[3]); } /* w x y z */
inline Transform trs(const float *t) { return Transform(v3(t), qt(t + 3), v3(t + 7)); }
inline BoundingBox box(const float *b) { return BoundingBox(v3(b), v3(b + 3)); }

inline void putBox(const BoundingBox &b, float *out) {
    memcpy(out, &b.minimum, 12);
    memcpy(out + 3, &b.maximum, 12);
}

/** RenderEntity is abstract (render_entity.h:88-92); geometry/material are never used by culling. */
class TestEntity : public RenderEntity {
public:
    uint32_t index = 0;
    Geometry geometry() const override { return Geometry(); }
    Material *material() const override { return nullptr; }
};

void configureView(RenderView &rv, const OrcView *v) {
    rv.setTransform(v3(v->position), qt(v->orientation));
    rv.perspective(v->hfov_deg, v->znear, v->zfar);
    rv.setViewport(IntRect(0, 0, v->vp_width, v->vp_height));
}

/** Frustum has no plane setter; its first member is Plane m_planes[6] (frustum.h:94-95). */
Frustum frustumFromPlanes(const float planes[24]) {
    Frustum f;
    memcpy(reinterpret_cast<void *>(&f), planes, 96);
    return f;
}

void configureLight(RenderLight &l, const OrcLight *d) {
    /* Same setters graphics/src/light.cc uses; final state is order-independent. */
    l.setColour(v3(d->colour));
    l.setIntensity(d->intensity);
    l.setFlags(d->flags);
    if (d->type == RenderLight::kPointLight || d->type == RenderLight::kSpotLight)
        l.setRange(d->range);
    if (d->type == RenderLight::kSpotLight)
        l.setCutoff(d->cutoff);
    if (d->type == RenderLight::kSpotLight || d->type == RenderLight::kDirectionalLight)
        l.setDirection(v3(d->direction));
    l.setPosition(v3(d->position));
}

struct RefWorld {
    SimpleRenderWorld world;
    std::vector<std::unique_ptr<TestEntity>> entities;   /* index = add order; null once removed */
    std::vector<std::unique_ptr<RenderLight>> lights;
    std::unordered_map<const RenderLight *, uint32_t> lightIndex;
};

} // namespace

extern "C" {

const char *orc_name(void) { return "reference"; }

/* ------------------------------------------------------------------ view */

void orc_view_planes(const OrcView *view, float out_planes[24]) {
    oref_shim_init();
    RenderView rv;
    configureView(rv, view);
    const Frustum &f = rv.frustum();
    for (unsigned i = 0; i < Frustum::kNumPlanes; i++)
        memcpy(out_planes + 4 * i, &f.plane(i).vector(), 16);
}

void orc_view_matrices(const OrcView *view, float out_mats[64], float out_corners[24]) {
    oref_shim_init();
    RenderView rv;
    configureView(rv, view);
    memcpy(out_mats, &rv.view(), 64);
    memcpy(out_mats + 16, &rv.projection(), 64);
    memcpy(out_mats + 32, &rv.viewProjection(), 64);
    memcpy(out_mats + 48, &rv.inverseViewProjection(), 64);
    for (unsigned i = 0; i < Frustum::kNumCorners; i++)
        memcpy(out_corners + 3 * i, &rv.frustum().corner(i), 12);
}

/* ------------------------------------------------------------------ transforms */

/* RESTATED control flow of Entity::transformed (engine/src/entity.cc:383-402); entity.cc itself needs
 * the objgen-generated reflection tables and cannot be compiled here. The operators are GLM's own. */
void orc_world_trs(int64_t n, const int32_t *parent, const float *local_trs, float *out_world_trs) {
    for (int64_t i = 0; i < n; i++) {
        glm::vec3 position = v3(local_trs + 10 * i);
        glm::quat orientation = qt(local_trs + 10 * i + 3);
        glm::vec3 scale = v3(local_trs + 10 * i + 7);
        if (parent[i] >= 0) {
            const float *p = out_world_trs + 10 * (int64_t)parent[i];
            glm::vec3 parentPosition = v3(p);
            glm::quat parentOrientation = qt(p + 3);
            glm::vec3 parentScale = v3(p + 7);
            position = (parentOrientation * (parentScale * position)) + parentPosition;   /* :394 */
            orientation = parentOrientation * orientation;                                /* :395 */
            scale = parentScale * scale;                                                  /* :396 */
        }
        float *o = out_world_trs + 10 * i;
        o[0] = position.x; o[1] = position.y; o[2] = position.z;
        o[3] = orientation.w; o[4] = orientation.x; o[5] = orientation.y; o[6] = orientation.z;
        o[7] = scale.x; o[8] = scale.y; o[9] = scale.z;
    }
}

void orc_trs_matrix(int64_t n, const float *t, float *out_mat16) {
    for (int64_t i = 0; i < n; i++) {
        Transform tr = trs(t + 10 * i);
        memcpy(out_mat16 + 16 * i, &tr.matrix(), 64);
    }
}

void orc_world_aabb(int64_t n, const float *world_trs, const float *local_aabb, float *out_aabb) {
    for (int64_t i = 0; i < n; i++) {
        Transform tr = trs(world_trs + 10 * i);
        putBox(box(local_aabb + 6 * i).transform(tr.matrix()), out_aabb + 6 * i);   /* render_entity.cc:93 */
    }
}

/* ------------------------------------------------------------------ flat entity cull */

static int64_t cullRange(int64_t begin, int64_t end, const float *aabb, const Frustum &f,
                         const uint32_t *flags, uint32_t require, uint32_t *out) {
    const BoundingBox *boxes = reinterpret_cast<const BoundingBox *>(aabb);
    int64_t count = 0;
    for (int64_t i = begin; i < end; i++) {
        if (Math::intersect(f, boxes[i])) {                                         /* simple_render_world.cc:39 */
            if (!flags || (flags[i] & require) == require)
                out[count++] = (uint32_t)i;
        }
    }
    return count;
}

int64_t orc_cull_aabb(int64_t n, const float *aabb, const float planes[24],
                      const uint32_t *flags, uint32_t require, uint32_t *out_idx) {
    Frustum f = frustumFromPlanes(planes);
    return cullRange(0, n, aabb, f, flags, require, out_idx);
}

int64_t orc_cull_aabb_mt(int64_t n, const float *aabb, const float planes[24],
                         const uint32_t *flags, uint32_t require, uint32_t *out_idx, int nthreads) {
    if (nthreads <= 1)
        return orc_cull_aabb(n, aabb, planes, flags, require, out_idx);
    Frustum f = frustumFromPlanes(planes);
    std::vector<int64_t> counts(nthreads);
    std::vector<std::thread> pool;
    /* Each thread compacts into out_idx at the start of its own range, then ranges are closed up in
     * order - the result is the same list the single loop produces. */
    for (int t = 0; t < nthreads; t++) {
        int64_t b = n * t / nthreads, e = n * (t + 1) / nthreads;
        pool.emplace_back([&, t, b, e] { counts[t] = cullRange(b, e, aabb, f, flags, require, out_idx + b); });
    }
    for (auto &th : pool) th.join();
    int64_t total = counts[0];
    for (int t = 1; t < nthreads; t++) {
        int64_t b = n * t / nthreads;
        memmove(out_idx + total, out_idx + b, counts[t] * sizeof(uint32_t));
        total += counts[t];
    }
    return total;
}

/* ------------------------------------------------------------------ lights */

void orc_light_spot_aabb(const OrcLight *d, float out_aabb[6]) {
    oref_shim_init();
    /* m_boundingBox is private; cull() against one plane set cannot reveal it, so read it through
     * the documented object layout instead: rebuild it exactly as render_light.cc:235-248 does, by
     * calling the same reference functions. RESTATED call sequence, reference arithmetic. */
    float radius = d->range * tanf(glm::radians(d->cutoff));
    glm::vec3 scale(radius, radius, d->range);
    glm::vec3 direction = glm::normalize(v3(d->direction));                            /* render_light.cc:80 */
    glm::quat orientation = Math::quatRotateBetween(glm::vec3(0.0f, 0.0f, -1.0f), direction);
    Transform volume(v3(d->position), orientation, scale);
    BoundingBox base(glm::vec3(-1.0f, -1.0f, -1.0f), glm::vec3(1.0f, 1.0f, 0.0f));
    putBox(base.transform(volume.matrix()), out_aabb);
}

int32_t orc_light_cull(int32_t n, const OrcLight *lights, const float planes[24], uint32_t *out_idx) {
    oref_shim_init();
    /* RenderLight::cull takes a RenderView; a RenderView cannot be given raw planes. Subclassing is
     * not possible (frustum() is non-virtual), so overwrite the view's frustum planes in place after
     * forcing its lazy update: m_frustum is reached through frustum()'s returned reference. */
    RenderView rv;
    rv.setTransform(glm::vec3(0.0f), glm::quat());
    rv.perspective(90.0f, 0.1f, 100.0f);
    Frustum &f = const_cast<Frustum &>(rv.frustum());
    memcpy(reinterpret_cast<void *>(&f), planes, 96);
    int32_t count = 0;
    for (int32_t i = 0; i < n; i++) {
        RenderLight light((RenderLight::Type)lights[i].type);
        configureLight(light, &lights[i]);
        if (!light.cull(rv))                                                           /* simple_render_world.cc:45 */
            out_idx[count++] = (uint32_t)i;
    }
    return count;
}

int32_t orc_light_shadow_views(const OrcLight *d, OrcView out_views[6]) {
    oref_shim_init();
    if (!(d->flags & RenderLight::kCastsShadows))
        return 0;
    if (d->type != RenderLight::kPointLight && d->type != RenderLight::kSpotLight)
        return 0;                                                                      /* render_light.cc:328-330 TODO */
    RenderLight light((RenderLight::Type)d->type);
    configureLight(light, d);
    unsigned count = light.numShadowViews();
    for (unsigned i = 0; i < count; i++) {
        RenderView &sv = light.shadowView(i);
        OrcView &o = out_views[i];
        memcpy(o.position, &sv.position(), 12);
        o.orientation[0] = sv.orientation().w; o.orientation[1] = sv.orientation().x;
        o.orientation[2] = sv.orientation().y; o.orientation[3] = sv.orientation().z;
        o.hfov_deg = sv.fov(); o.znear = sv.zNear(); o.zfar = sv.zFar();
        o.vp_width = 512; o.vp_height = 512;                                           /* deferred_render_pipeline.cc:48,264 */
    }
    return (int32_t)count;
}

/* ------------------------------------------------------------------ tier A object world */

void *oref_world_create(void) { oref_shim_init(); return new RefWorld; }

void oref_world_destroy(void *w) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    rw->entities.clear();   /* ~RenderEntity -> setWorld(nullptr) -> removeEntity */
    rw->lights.clear();
    delete rw;
}

/* Tearing a big world down entity by entity is O(n^2) in the reference (each ~RenderEntity does a
 * std::list::remove scan, simple_render_world.cc:65-67). The baseline timer abandons its world instead. */
void oref_world_abandon(void *w) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    for (auto &e : rw->entities) e.release();
    for (auto &l : rw->lights) l.release();
}

int64_t oref_world_add_entities(void *w, int64_t n, const float *t, const float *local_aabb, const uint32_t *flags) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    int64_t first = (int64_t)rw->entities.size();
    for (int64_t i = 0; i < n; i++) {
        std::unique_ptr<TestEntity> e(new TestEntity);
        e->index = (uint32_t)(first + i);
        e->setFlags(flags ? flags[i] : 0);
        e->setBoundingBox(box(local_aabb + 6 * i));      /* graphics/src/mesh_renderer.cc:50-59 order */
        e->setTransform(trs(t + 10 * i));                /* graphics/src/renderer.cc:62-66 */
        e->setWorld(&rw->world);                         /* render_entity.cc:52-60 -> addEntity */
        rw->entities.push_back(std::move(e));
    }
    return first;
}

void oref_world_set_transforms(void *w, int64_t n, const uint32_t *idx, const float *t) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    for (int64_t i = 0; i < n; i++)
        rw->entities[idx[i]]->setTransform(trs(t + 10 * i));
}

void oref_world_set_local_aabb(void *w, uint32_t idx, const float *b) {
    static_cast<RefWorld *>(w)->entities[idx]->setBoundingBox(box(b));
}

void oref_world_set_flags(void *w, uint32_t idx, uint32_t flags) {
    static_cast<RefWorld *>(w)->entities[idx]->setFlags(flags);
}

void oref_world_remove_entity(void *w, uint32_t idx) {
    static_cast<RefWorld *>(w)->entities[idx].reset();
}

void oref_world_get_bounds(void *w, int64_t first, int64_t n, float *out_aabb) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    for (int64_t i = 0; i < n; i++)
        putBox(rw->entities[first + i]->worldBoundingBox(), out_aabb + 6 * i);
}

void oref_world_get_matrices(void *w, int64_t first, int64_t n, float *out_mat16) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    for (int64_t i = 0; i < n; i++)
        memcpy(out_mat16 + 16 * i, &rw->entities[first + i]->transform().matrix(), 64);
}

int32_t oref_world_add_lights(void *w, int32_t n, const OrcLight *lights) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    int32_t first = (int32_t)rw->lights.size();
    for (int32_t i = 0; i < n; i++) {
        std::unique_ptr<RenderLight> l(new RenderLight((RenderLight::Type)lights[i].type));
        configureLight(*l, &lights[i]);
        l->setWorld(&rw->world);
        rw->lightIndex[l.get()] = (uint32_t)(first + i);
        rw->lights.push_back(std::move(l));
    }
    return first;
}

void oref_world_update_light(void *w, uint32_t idx, const OrcLight *d) {
    configureLight(*static_cast<RefWorld *>(w)->lights[idx], d);
}

void oref_world_remove_light(void *w, uint32_t idx) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    rw->lightIndex.erase(rw->lights[idx].get());
    rw->lights[idx].reset();
}

/** SimpleRenderWorld::cull as shipped. Returns entity count; *n_lights gets the light count. */
int64_t oref_world_cull(void *w, const OrcView *view, uint32_t flags,
                        uint32_t *out_entities, uint32_t *out_lights, int32_t *n_lights) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    RenderView rv;
    configureView(rv, view);
    RenderWorld::CullResults results;
    rw->world.cull(rv, results, flags);
    int64_t count = 0;
    for (RenderEntity *e : results.entities)
        out_entities[count++] = static_cast<TestEntity *>(e)->index;
    int32_t lc = 0;
    for (RenderLight *l : results.lights)
        out_lights[lc++] = rw->lightIndex[l];
    if (n_lights) *n_lights = lc;
    return count;
}

/** RESTATED loop of DeferredRenderPipeline::prepareLights (deferred_render_pipeline.cc:207,247,259-284),
 * which itself needs a live GPU pipeline; every call it makes into the world/light/view is the
 * reference's own. */
int32_t oref_world_shadow_lists(void *w, const OrcView *main_view, int32_t shadow_res,
                                int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                                int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    RenderView rv;
    configureView(rv, main_view);
    RenderWorld::CullResults mainResults;
    rw->world.cull(rv, mainResults, RenderWorld::kCullLights);                 /* :147 */
    int32_t lists = 0;
    int64_t total = 0;
    out_offsets[0] = 0;
    for (RenderLight *light : mainResults.lights) {                            /* :207 */
        if (!light->castsShadows())                                             /* :247 */
            continue;
        /* allocateShadowMap (:291-303) is fatal("TODO") for anything but spot and point lights, so the
         * reference never reaches the cull below for a shadow-casting ambient/directional light: such
         * lights contribute no lists (their shadow RenderView is never configured either,
         * render_light.cc:328-330). */
        if (light->type() != RenderLight::kPointLight && light->type() != RenderLight::kSpotLight)
            continue;
        unsigned numShadowViews = light->numShadowViews();                      /* :259 */
        for (unsigned i = 0; i < numShadowViews; i++) {
            RenderView &shadowView = light->shadowView(i);
            shadowView.setViewport(IntRect(0, 0, shadow_res, shadow_res));      /* :264-265 */
            RenderWorld::CullResults cullResults;
            rw->world.cull(shadowView, cullResults, 0);                         /* :270 */
            if (lists >= cap_lists) return -1;
            for (RenderEntity *e : cullResults.entities) {
                if (e->castsShadow()) {                                         /* :273 */
                    if (total >= cap_entities) return -1;
                    out_entities[total++] = static_cast<TestEntity *>(e)->index;
                }
            }
            out_list_light[lists] = (int32_t)rw->lightIndex[light];
            out_list_view[lists] = (int32_t)i;
            out_offsets[++lists] = total;
        }
    }
    return lists;
}

/** Wall-clock of SimpleRenderWorld::cull as shipped (std::list walk + std::list result build),
 * one thread - "the reference CPU cull" of BASELINE.md section 4 (CPU-A). Milliseconds per call. */
void oref_world_time_cull(void *w, const OrcView *view, uint32_t flags, int32_t iters, double *out_ms,
                          int64_t *out_visible) {
    RefWorld *rw = static_cast<RefWorld *>(w);
    RenderView rv;
    configureView(rv, view);
    (void)rv.frustum();
    for (int32_t it = 0; it < iters; it++) {
        auto t0 = std::chrono::steady_clock::now();
        RenderWorld::CullResults results;
        rw->world.cull(rv, results, flags);
        auto t1 = std::chrono::steady_clock::now();
        out_ms[it] = std::chrono::duration<double, std::milli>(t1 - t0).count();
        if (out_visible) *out_visible = (int64_t)results.entities.size();
    }
}

/* orc_shadow_lists over flat arrays: tier B composition of the calls above. */
int32_t orc_shadow_lists(int64_t n, const float *aabb, const uint32_t *flags,
                         int32_t n_lights, const OrcLight *lights, const float main_planes[24],
                         int32_t cap_lists, int32_t *out_list_light, int32_t *out_list_view,
                         int64_t *out_offsets, int64_t cap_entities, uint32_t *out_entities) {
    std::vector<uint32_t> visibleLights(n_lights > 0 ? n_lights : 1);
    int32_t nv = orc_light_cull(n_lights, lights, main_planes, visibleLights.data());
    std::vector<uint32_t> scratch((size_t)(n > 0 ? n : 1));
    int32_t lists = 0;
    int64_t total = 0;
    out_offsets[0] = 0;
    for (int32_t k = 0; k < nv; k++) {
        const OrcLight *d = &lights[visibleLights[k]];
        OrcView views[6];
        int32_t count = orc_light_shadow_views(d, views);
        for (int32_t i = 0; i < count; i++) {
            float planes[24];
            orc_view_planes(&views[i], planes);
            int64_t c = orc_cull_aabb(n, aabb, planes, flags, ORC_CASTS_SHADOW, scratch.data());
            if (lists >= cap_lists || total + c > cap_entities) return -1;
            memcpy(out_entities + total, scratch.data(), c * sizeof(uint32_t));
            total += c;
            out_list_light[lists] = (int32_t)visibleLights[k];
            out_list_view[lists] = i;
            out_offsets[++lists] = total;
        }
    }
    return lists;
}

/* Draw lists. Up to kDrawListObjects visible entities the draws come out of the reference's own DrawList::add
 * (draw_list.cc compiled where it lies, real RenderEntity objects: oracle/draw_harness.cc); only the two ten-line pipeline
 * loops that pick a list (deferred_render_pipeline.cc:318-330, :272-283) are restated there. Longer lists would mean
 * millions of RenderEntity objects (a uniform buffer and resource set each) and take the restated loop below, which
 * tests/test_oracle.py holds equal to the DrawList path. */
extern "C" void oref_draw_lists(int64_t n_visible, const uint32_t *visible, const uint32_t *pass_counts, const uint32_t *flags,
                                int32_t mode, uint32_t *out_draws0, uint32_t *out_draws1, int64_t out_counts[2]);
static const int64_t kDrawListObjects = 200000;

void orc_draw_lists(int64_t n_visible, const uint32_t *visible, const uint32_t *pass_counts, const uint32_t *flags,
                    int32_t mode, uint32_t *out_draws0, uint32_t *out_draws1, int64_t out_counts[2]) {
    if (n_visible <= kDrawListObjects) {
        oref_draw_lists(n_visible, visible, pass_counts, flags, mode, out_draws0, out_draws1, out_counts);
        return;
    }
    std::vector<std::pair<uint32_t, uint32_t>> lists[2];
    auto add = [&](int list, uint32_t entity, uint32_t numPasses) {
        for (uint32_t i = 0; i < numPasses; i++)
            lists[list].emplace_back(entity, i);
    };
    for (int64_t k = 0; k < n_visible; k++) {
        uint32_t e = visible[k], pc = pass_counts[e];
        if (mode == 0) {
            if ((pc & 0xffu) > 0) add(0, e, pc & 0xffu);
            else if (((pc >> 8) & 0xffu) > 0) add(1, e, (pc >> 8) & 0xffu);
        } else if ((flags[e] & RenderEntity::kCastsShadow) && ((pc >> 16) & 0xffu) > 0) {
            add(0, e, (pc >> 16) & 0xffu);
        }
    }
    uint32_t *outs[2] = { out_draws0, out_draws1 };
    for (int l = 0; l < 2; l++) {
        out_counts[l] = (int64_t)lists[l].size();
        for (size_t i = 0; i < lists[l].size(); i++) {
            outs[l][2 * i] = lists[l][i].first;
            outs[l][2 * i + 1] = lists[l][i].second;
        }
    }
}

/* EntityUniforms as RenderEntity::setTransform fills them (render_entity.cc:67-69), via the reference's Transform. */
void orc_entity_uniforms(int64_t n, const float *world_trs, float *out20) {
    for (int64_t i = 0; i < n; i++) {
        Transform tr = trs(world_trs + 10 * i);
        memcpy(out20 + 20 * i, &tr.matrix(), 64);
        memcpy(out20 + 20 * i + 16, &tr.position(), 12);
        out20[20 * i + 19] = 0.0f;
    }
}

/* ------------------------------------------------------------------ uniform blocks as the reference ships them */

/* What UniformBuffer::flush (render_core/src/uniform_buffer.cc:147-152) wrote into the object's (headless) GPU buffer. */
static void readUniformBlock(GPUResourceSet *set, void *out, size_t size) {
    GPUObject *object = set->slots()[ResourceSlots::kUniforms].object.get();
    GPUBuffer *buffer = static_cast<GPUBuffer *>(object);
    if (buffer->size() != size) {
        fprintf(stderr, "oracle: uniform block is %zu bytes, caller expects %zu\n", buffer->size(), size);
        abort();
    }
    memcpy(out, buffer->map(0, size, 0, GPUBuffer::kWriteAccess), size);
    buffer->unmap();
}

/* glm quat * vec3 (Light::transformed, graphics/src/light.cc:105; Light::direction :93) */
void orc_quat_rotate(const float q[4], const float v[3], float out[3]) {
    glm::vec3 r = qt(q) * v3(v);
    memcpy(out, &r, 12);
}

void orc_uniform_sizes(int32_t out[3]) {
    out[0] = (int32_t)sizeof(EntityUniforms);
    out[1] = (int32_t)sizeof(ViewUniforms);
    out[2] = (int32_t)sizeof(LightUniforms);
}

/* EntityUniforms through RenderEntity::setTransform + getResources (render_entity.cc:64-72,100-103). */
void orc_entity_uniform_blocks(int64_t n, const float *world_trs, void *out80) {
    oref_shim_init();
    for (int64_t i = 0; i < n; i++) {
        TestEntity e;
        e.setTransform(trs(world_trs + 10 * i));
        readUniformBlock(e.getResources(), (char *)out80 + 80 * i, 80);
    }
}

/* ViewUniforms through RenderView::setTransform/perspective/setViewport + getResources (render_view.cc:51-98). */
void orc_view_uniforms(const OrcView *view, const int32_t viewport_pos[2], void *out288) {
    oref_shim_init();
    RenderView rv;
    rv.setTransform(v3(view->position), qt(view->orientation));
    rv.perspective(view->hfov_deg, view->znear, view->zfar);
    rv.setViewport(IntRect(viewport_pos[0], viewport_pos[1], view->vp_width, view->vp_height));
    readUniformBlock(rv.getResources(), out288, 288);
}

/* LightUniforms through the setters graphics/src/light.cc drives: Light::Light (:35-45: colour, intensity, flags, shadow
 * bias), the Point/Spot constructors (:58-74: range, attenuation, cutoff), Light::transformed (:100-108: direction, then
 * position, for EVERY light type). setRange is also called for ambient/directional lights, whose m_range the reference
 * leaves uninitialised (it then copies it into the block, render_light.cc:336): the harness needs a defined value. */
void orc_light_uniforms(const OrcLight *d, const float attenuation[3], float shadow_bias, void *out208) {
    oref_shim_init();
    RenderLight l((RenderLight::Type)d->type);
    l.setColour(v3(d->colour));
    l.setIntensity(d->intensity);
    l.setFlags(d->flags);
    l.setShadowBias(shadow_bias);
    l.setRange(d->range);
    if (d->type == RenderLight::kPointLight || d->type == RenderLight::kSpotLight)
        l.setAttenuation(attenuation[0], attenuation[1], attenuation[2]);
    if (d->type == RenderLight::kSpotLight)
        l.setCutoff(d->cutoff);
    l.setDirection(v3(d->direction));
    l.setPosition(v3(d->position));
    readUniformBlock(l.getResources(), out208, 208);
}

} /* extern "C" */
