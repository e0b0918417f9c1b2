/*
 * oracle/adapter_check.cc - end-to-end drop-in check of adapter/CudaRenderWorld against SimpleRenderWorld.
 *
 * TEST INFRASTRUCTURE ONLY (built into oracle/_ref/adapter_check together with the reference's own sources, the
 * headless shim and the adapter; linked against liborion_vis.so). Two identical scenes of the reference's
 * RenderEntity / RenderLight objects are attached, one to the reference's SimpleRenderWorld and one to
 * CudaRenderWorld; after every mutation (transform changes, box changes, removals, light changes) both worlds are
 * culled through the RenderWorld interface and the CullResults lists must name the same objects in the same order,
 * for the main view and for every shadow view of every visible light (the prepareLights loop).
 * Exit code 0 = identical everywhere. Needs a CUDA device.
 */
#include "render/cuda_render_world.h"
#include "render/render_entity.h"
#include "render/render_light.h"
#include "render/render_view.h"
#include "render/simple_render_world.h"

#include "render_core/geometry.h"

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <random>
#include <vector>

void oref_shim_init();

/* core/src/error.cc needs SDL; the harness supplies the one symbol the adapter's error path uses. */
void __fatal(const char *file, int line, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "fatal (%s:%d): ", file, line);
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
    exit(3);
}

namespace {

class TestEntity : public RenderEntity {
public:
    size_t index = 0;
    Geometry geometry() const override { return Geometry(); }
    Material *material() const override { return nullptr; }
};

struct Scene {
    std::unique_ptr<RenderWorld> world;
    std::vector<std::unique_ptr<TestEntity>> entities;
    std::vector<std::unique_ptr<RenderLight>> lights;
};

std::mt19937 rng(1234);
float uni(float a, float b) { return std::uniform_real_distribution<float>(a, b)(rng); }

Transform randomTransform(float extent) {
    glm::quat q = glm::normalize(glm::quat(uni(-1, 1), uni(-1, 1), uni(-1, 1), uni(-1, 1)));
    return Transform(glm::vec3(uni(-extent, extent), uni(-extent, extent), uni(-extent, extent)), q,
                     glm::vec3(uni(-2, 2), uni(0.2f, 2), uni(0.2f, 2)));
}

BoundingBox randomBox() {
    glm::vec3 h(uni(0.05f, 2), uni(0.05f, 2), uni(0.05f, 2)), c(uni(-1, 1), uni(-1, 1), uni(-1, 1));
    return BoundingBox(c - h, c + h);
}

int failures = 0;

template <typename T, typename F>
void compareLists(const char *what, const std::list<T *> &a, const std::list<T *> &b, F indexOf) {
    bool same = a.size() == b.size();
    auto ia = a.begin();
    auto ib = b.begin();
    for (; same && ia != a.end(); ++ia, ++ib)
        same = indexOf(*ia) == indexOf(*ib);
    if (!same) {
        fprintf(stderr, "MISMATCH in %s: reference %zu items, adapter %zu items\n", what, a.size(), b.size());
        failures++;
    }
}

void compareWorlds(Scene &ref, Scene &gpu, RenderView &view, const char *stage) {
    RenderWorld::CullResults ra, rb;
    ref.world->cull(view, ra, RenderWorld::kCullLights);
    gpu.world->cull(view, rb, RenderWorld::kCullLights);
    auto entityIndex = [](RenderEntity *e) { return static_cast<TestEntity *>(e)->index; };
    compareLists(stage, ra.entities, rb.entities, entityIndex);
    /* lights: compare by position in the scene's light vector */
    auto lightIndexIn = [](Scene &s) {
        return [&s](RenderLight *l) {
            for (size_t i = 0; i < s.lights.size(); i++)
                if (s.lights[i].get() == l) return i;
            return (size_t)-1;
        };
    };
    {
        std::vector<size_t> la, lb;
        auto fa = lightIndexIn(ref), fb = lightIndexIn(gpu);
        for (RenderLight *l : ra.lights) la.push_back(fa(l));
        for (RenderLight *l : rb.lights) lb.push_back(fb(l));
        if (la != lb) {
            fprintf(stderr, "MISMATCH in %s lights: %zu vs %zu\n", stage, la.size(), lb.size());
            failures++;
        }
    }
    /* the prepareLights loop (deferred_render_pipeline.cc:207,247,259-284) on both worlds */
    auto ita = ra.lights.begin();
    auto itb = rb.lights.begin();
    size_t shadowViews = 0;
    for (; ita != ra.lights.end() && itb != rb.lights.end(); ++ita, ++itb) {
        RenderLight *la = *ita, *lb = *itb;
        if (!la->castsShadows() || (la->type() != RenderLight::kPointLight && la->type() != RenderLight::kSpotLight))
            continue;
        for (unsigned i = 0; i < la->numShadowViews(); i++) {
            la->shadowView(i).setViewport(IntRect(0, 0, 512, 512));
            lb->shadowView(i).setViewport(IntRect(0, 0, 512, 512));
            RenderWorld::CullResults sa, sb;
            ref.world->cull(la->shadowView(i), sa, 0);
            gpu.world->cull(lb->shadowView(i), sb, 0);
            compareLists("shadow view", sa.entities, sb.entities, entityIndex);
            shadowViews++;
        }
    }
    printf("%-28s main view: %zu entities, %zu lights, %zu shadow views compared\n", stage, ra.entities.size(),
           ra.lights.size(), shadowViews);
}

/* One frame as DeferredRenderPipeline drives a RenderWorld: the main cull with lights, then one cull per shadow view of
 * every visible shadow-casting light (deferred_render_pipeline.cc:147, 257-284). Returns the entity-view tests done. */
size_t frame(Scene &s, RenderView &view, size_t *shadowViews) {
    RenderWorld::CullResults main;
    s.world->cull(view, main, RenderWorld::kCullLights);
    size_t views = 1, listed = main.entities.size();
    for (RenderLight *l : main.lights) {
        if (!l->castsShadows() || (l->type() != RenderLight::kPointLight && l->type() != RenderLight::kSpotLight))
            continue;
        for (unsigned i = 0; i < l->numShadowViews(); i++) {
            l->shadowView(i).setViewport(IntRect(0, 0, 512, 512));
            RenderWorld::CullResults shadow;
            s.world->cull(l->shadowView(i), shadow, 0);
            listed += shadow.entities.size();
            views++;
        }
    }
    *shadowViews = views - 1;
    return listed;
}

double timeFrames(Scene &s, RenderView &view, int frames, size_t *shadowViews, size_t *listed) {
    frame(s, view, shadowViews);                                /* warm-up */
    const auto t0 = std::chrono::steady_clock::now();
    for (int f = 0; f < frames; f++)
        *listed = frame(s, view, shadowViews);
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / frames;
}

}  // namespace

int main(int argc, char **argv) {
    const size_t numEntities = argc > 1 ? strtoul(argv[1], nullptr, 10) : 20000;
    /* a second argument reseeds the scenario and widens it: light ranges down to centimetres (at or below the shadow
     * views' near plane the reference's face frusta open outwards, render_light.cc:319) */
    const bool wild = argc > 2;
    if (wild)
        rng.seed((unsigned)strtoul(argv[2], nullptr, 10));
    oref_shim_init();
    Scene ref, gpu;
    ref.world.reset(new SimpleRenderWorld);
    gpu.world.reset(new CudaRenderWorld(0));

    for (size_t i = 0; i < numEntities; i++) {
        Transform t = randomTransform(60.0f);
        BoundingBox box = randomBox();
        uint32_t flags = (rng() & 1) ? RenderEntity::kCastsShadow : 0;
        for (Scene *s : { &ref, &gpu }) {
            std::unique_ptr<TestEntity> e(new TestEntity);
            e->index = i;
            e->setFlags(flags);
            e->setBoundingBox(box);
            e->setTransform(t);
            e->setWorld(s->world.get());
            s->entities.push_back(std::move(e));
        }
    }
    const RenderLight::Type types[] = { RenderLight::kAmbientLight, RenderLight::kDirectionalLight, RenderLight::kPointLight,
                                        RenderLight::kSpotLight, RenderLight::kPointLight, RenderLight::kSpotLight };
    for (size_t i = 0; i < 24; i++) {
        RenderLight::Type type = types[i % 6];
        glm::vec3 pos(uni(-40, 40), uni(-40, 40), uni(-40, 40)), dir(uni(-1, 1), uni(-1, 1), uni(-1, 1)), col(uni(0, 1), uni(0, 1), uni(0, 1));
        float intensity = (i == 7) ? 0.0f : uni(0.1f, 2), range = (wild && rng() % 3 == 0) ? uni(0.02f, 0.3f) : uni(5, 60), cutoff = uni(5, 45);
        uint32_t flags = (i % 3) ? RenderLight::kCastsShadows : 0;
        for (Scene *s : { &ref, &gpu }) {
            std::unique_ptr<RenderLight> l(new RenderLight(type));
            l->setColour(col);
            l->setIntensity(intensity);
            l->setRange(range);
            l->setCutoff(cutoff);
            l->setDirection(dir);
            l->setPosition(pos);
            l->setFlags(flags);
            l->setWorld(s->world.get());
            s->lights.push_back(std::move(l));
        }
    }

    RenderView view;
    view.setTransform(glm::vec3(0, 0, 90), glm::quat());
    view.perspective(75.0f, 0.1f, 400.0f);
    view.setViewport(IntRect(0, 0, 1920, 1080));
    compareWorlds(ref, gpu, view, "initial");

    for (size_t k = 0; k < numEntities / 10; k++) {            /* move a tenth of the entities */
        size_t i = rng() % numEntities;
        Transform t = randomTransform(60.0f);
        ref.entities[i]->setTransform(t);
        gpu.entities[i]->setTransform(t);
    }
    compareWorlds(ref, gpu, view, "after transforms");

    for (size_t k = 0; k < numEntities / 20; k++) {            /* new local boxes */
        size_t i = rng() % numEntities;
        BoundingBox box = randomBox();
        ref.entities[i]->setBoundingBox(box);
        gpu.entities[i]->setBoundingBox(box);
    }
    compareWorlds(ref, gpu, view, "after box changes");

    for (size_t k = 0; k < numEntities / 10; k++) {            /* destroy entities: survivors keep their order */
        size_t i = rng() % numEntities;
        ref.entities[i].reset();
        gpu.entities[i].reset();
    }
    compareWorlds(ref, gpu, view, "after removals");

    for (size_t k = 0; k < numEntities / 2; k++) {             /* heavy despawn: the adapter reclaims the slots, ids change inside */
        size_t i = rng() % numEntities;
        ref.entities[i].reset();
        gpu.entities[i].reset();
    }
    compareWorlds(ref, gpu, view, "after heavy removals");
    for (size_t i = 0; i < numEntities; i += 7) {              /* ... and spawning goes on afterwards */
        if (ref.entities[i])
            continue;
        Transform t = randomTransform(60.0f);
        BoundingBox box = randomBox();
        for (Scene *s : { &ref, &gpu }) {
            std::unique_ptr<TestEntity> e(new TestEntity);
            e->index = i;
            e->setBoundingBox(box);
            e->setTransform(t);
            e->setWorld(s->world.get());
            s->entities[i] = std::move(e);
        }
    }
    compareWorlds(ref, gpu, view, "after respawns");

    for (size_t i = 0; i < ref.lights.size(); i += 3) {        /* setters that do NOT notify the world */
        float range = (wild && rng() % 3 == 0) ? uni(0.02f, 0.3f) : uni(5, 80);
        glm::vec3 col = (i == 6) ? glm::vec3(0.0f) : glm::vec3(1.0f);
        for (Scene *s : { &ref, &gpu }) {
            s->lights[i]->setRange(range);
            s->lights[i]->setColour(col);
        }
    }
    ref.lights[4].reset();
    gpu.lights[4].reset();
    view.setTransform(glm::vec3(30, 10, 70), glm::normalize(glm::quat(0.95f, 0.05f, 0.3f, 0.0f)));
    compareWorlds(ref, gpu, view, "after light changes");

    /* The batched shadow-view path must have been the one that answered (and agreed, above). */
    CudaRenderWorld *cuda = static_cast<CudaRenderWorld *>(gpu.world.get());
    printf("shadow views served from %zu batched culls: %zu\n", cuda->shadowBatches(), cuda->shadowViewsServed());
    if (cuda->shadowViewsServed() == 0 || cuda->shadowBatches() == 0) {
        fprintf(stderr, "the adapter never used the batched shadow-view cull\n");
        failures++;
    }
    /* Frame time through the RenderWorld interface, both worlds, same scene. */
    {
        size_t sv = 0, listedRef = 0, listedGpu = 0;
        const int frames = numEntities > 200000 ? 3 : 20;
        const double refUs = timeFrames(ref, view, frames, &sv, &listedRef);
        const double gpuUs = timeFrames(gpu, view, frames, &sv, &listedGpu);
        printf("FRAME_US entities=%zu shadow_views=%zu listed=%zu reference=%.1f adapter=%.1f speedup=%.2f\n", numEntities, sv, listedRef,
               refUs, gpuUs, refUs / gpuUs);
        if (listedRef != listedGpu) {
            fprintf(stderr, "timed frames disagree: %zu vs %zu listed entities\n", listedRef, listedGpu);
            failures++;
        }
    }

    /* Leak the scenes instead of tearing them down one std::list::remove at a time (O(n^2)). */
    for (Scene *s : { &ref, &gpu }) {
        for (auto &e : s->entities) e.release();
        for (auto &l : s->lights) l.release();
        s->world.release();
    }
    printf("%s\n", failures ? "ADAPTER MISMATCH" : "ADAPTER OK");
    return failures ? 1 : 0;
}
