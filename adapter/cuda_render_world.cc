/*
 * cuda_render_world.cc - see cuda_render_world.h. Install as engine/runtime/render/src/cuda_render_world.cc.
 */
#include "render/cuda_render_world.h"

#include "render/render_entity.h"
#include "render/render_light.h"
#include "render/render_view.h"

#include "orion_vis.h"

#include <cstring>

/** The reference's cull() cannot fail; unrecoverable conditions go through fatal() like everywhere in the engine. */
#define OVIS_CHECK(call) \
    do { \
        if ((call) != OVIS_OK) \
            fatal("liborion_vis: %s", ovis_last_error()); \
    } while (0)

static OvisLight describeLight(const RenderLight *light) {
    OvisLight d;
    memset(&d, 0, sizeof(d));
    d.type = static_cast<int32_t>(light->type());
    d.flags = light->flags();
    memcpy(d.position, &light->position(), sizeof(d.position));
    memcpy(d.direction, &light->direction(), sizeof(d.direction));
    memcpy(d.colour, &light->colour(), sizeof(d.colour));
    d.intensity = light->intensity();
    d.range = light->range();
    d.cutoff = light->cutoff();
    /* Only the fields RenderLight::cull (render_light.cc:187-209) reads for this type may be uninitialised
     * garbage in the RenderLight; give the others defined values. */
    if (d.type != RenderLight::kSpotLight) {
        d.cutoff = 0.0f;
        d.direction[0] = 0.0f; d.direction[1] = 0.0f; d.direction[2] = -1.0f;
    }
    if (d.type != RenderLight::kSpotLight && d.type != RenderLight::kPointLight)
        d.range = 0.0f;
    return d;
}

static OvisView describeView(const RenderView &view) {
    OvisView v;
    memset(&v, 0, sizeof(v));
    memcpy(v.position, &view.position(), sizeof(v.position));
    const glm::quat &q = view.orientation();
    v.orientation[0] = q.w; v.orientation[1] = q.x; v.orientation[2] = q.y; v.orientation[3] = q.z;
    v.hfov_deg = view.fov();
    v.znear = view.zNear();
    v.zfar = view.zFar();
    v.vp_width = view.viewport().width;
    v.vp_height = view.viewport().height;
    return v;
}

CudaRenderWorld::CudaRenderWorld(int device) :
    m_world                (nullptr),
    m_entityResult         (nullptr),
    m_entityResultCapacity (0),
    m_shadowCacheValid     (false)
{
    OVIS_CHECK(ovis_world_create(device, &m_world));
}

CudaRenderWorld::~CudaRenderWorld() {
    if (m_entityResult)
        ovis_host_free(m_entityResult);
    ovis_world_destroy(m_world);
}

void CudaRenderWorld::addEntity(RenderEntity *entity) {
    /* Ids are slots handed out in order, so the id is known now and the upload can wait: a level load that activates a
     * million entities becomes ONE ovis_entity_add at the next cull instead of a million calls with two launches each. */
    const uint32_t id = static_cast<uint32_t>(m_entities.size());
    const BoundingBox &box = entity->boundingBox();
    const float local[6] = { box.minimum.x, box.minimum.y, box.minimum.z, box.maximum.x, box.maximum.y, box.maximum.z };
    m_newLocal.insert(m_newLocal.end(), local, local + 6);
    m_newFlags.push_back(entity->flags());
    m_entities.push_back(entity);
    m_isPending.push_back(false);
    m_entityIds[entity] = id;
    m_shadowCacheValid = false;
    updateEntity(entity);
}

/** Hand the entities added since the last flush to the library, in one call. */
void CudaRenderWorld::flushAdds() const {
    if (m_newFlags.empty())
        return;
    uint32_t first = 0;
    OVIS_CHECK(ovis_entity_add(m_world, static_cast<int64_t>(m_newFlags.size()), m_newLocal.data(), m_newFlags.data(), &first));
    if (first + m_newFlags.size() != m_entities.size())
        fatal("liborion_vis: entity ids out of step (%u + %zu != %zu)", first, m_newFlags.size(), m_entities.size());
    m_newLocal.clear();
    m_newFlags.clear();
}

void CudaRenderWorld::updateEntity(RenderEntity *entity) {
    auto it = m_entityIds.find(entity);
    if (it == m_entityIds.end())
        return;
    if (!m_isPending[it->second]) {
        m_isPending[it->second] = true;
        m_pending.push_back(it->second);
    }
    m_shadowCacheValid = false;
}

void CudaRenderWorld::removeEntity(RenderEntity *entity) {
    auto it = m_entityIds.find(entity);
    if (it == m_entityIds.end())
        return;
    const uint32_t id = it->second;
    flushAdds();
    OVIS_CHECK(ovis_entity_remove(m_world, 1, &id));
    m_entities[id] = nullptr;
    m_entityIds.erase(it);
    m_removed++;
    m_shadowCacheValid = false;
}

/** SimpleRenderWorld's std::list shrinks when an entity leaves (simple_render_world.cc:65-67). Here a removed entity keeps
 *  its slot until more than a quarter of the slots are dead; then the library closes the gaps (order preserved) and the
 *  pointer <-> id maps are rewritten from the remap it returns. */
void CudaRenderWorld::compactIfWorthIt() const {
    if (m_removed < 1024 || m_removed * 4 < m_entities.size())
        return;
    std::vector<uint32_t> remap(m_entities.size());
    int64_t oldSlots = 0, newSlots = 0;
    OVIS_CHECK(ovis_world_compact(m_world, static_cast<int64_t>(remap.size()), remap.data(), &oldSlots, &newSlots));
    std::vector<RenderEntity *> entities(static_cast<size_t>(newSlots), nullptr);
    for (size_t old = 0; old < m_entities.size(); old++) {
        if (remap[old] == 0xffffffffu)
            continue;
        entities[remap[old]] = m_entities[old];
        m_entityIds[m_entities[old]] = remap[old];
    }
    m_entities.swap(entities);
    std::vector<uint32_t> pending;
    for (uint32_t id : m_pending)
        if (remap[id] != 0xffffffffu)
            pending.push_back(remap[id]);
    m_pending.swap(pending);
    m_isPending.assign(m_entities.size(), false);
    for (uint32_t id : m_pending)
        m_isPending[id] = true;
    m_removed = 0;
}

void CudaRenderWorld::addLight(RenderLight *light) {
    uint32_t id = 0;
    OvisLight d = describeLight(light);
    OVIS_CHECK(ovis_light_add(m_world, 1, &d, &id));
    if (m_lights.size() <= id) {
        m_lights.resize(id + 1, nullptr);
        m_lightSent.resize(id + 1);
    }
    m_lights[id] = light;
    m_lightIds[light] = id;
    m_lightSent[id] = d;
    m_shadowCacheValid = false;
}

void CudaRenderWorld::updateLight(RenderLight *light) {
    /* Nothing to send here: most RenderLight setters never call this (render_light.cc:67-142), so cull() re-reads
     * every light's parameters anyway (syncLights). */
    m_shadowCacheValid = false;
}

void CudaRenderWorld::removeLight(RenderLight *light) {
    auto it = m_lightIds.find(light);
    if (it == m_lightIds.end())
        return;
    OVIS_CHECK(ovis_light_remove(m_world, it->second));
    m_lights[it->second] = nullptr;
    m_lightIds.erase(it);
    m_shadowCacheValid = false;
}

/** Push every world bounding box that changed since the last cull, in one batch. */
void CudaRenderWorld::flush() const {
    flushAdds();
    compactIfWorthIt();
    m_ids.clear();
    m_boxes.clear();
    for (uint32_t id : m_pending) {
        m_isPending[id] = false;
        const RenderEntity *entity = m_entities[id];
        if (!entity)
            continue;
        const BoundingBox &box = entity->worldBoundingBox();
        const float world[6] = { box.minimum.x, box.minimum.y, box.minimum.z, box.maximum.x, box.maximum.y, box.maximum.z };
        m_ids.push_back(id);
        m_boxes.insert(m_boxes.end(), world, world + 6);
    }
    m_pending.clear();
    if (!m_ids.empty())
        OVIS_CHECK(ovis_entity_set_world_aabbs_by_id(m_world, static_cast<int64_t>(m_ids.size()), m_ids.data(), m_boxes.data()));
}

/** Re-read every light (the reference reads them in place at each cull); only the changed ones cross the ABI. */
void CudaRenderWorld::syncLights() const {
    for (size_t id = 0; id < m_lights.size(); id++) {
        if (!m_lights[id])
            continue;
        const OvisLight d = describeLight(m_lights[id]);
        if (memcmp(&d, &m_lightSent[id], sizeof(d)) != 0) {
            OVIS_CHECK(ovis_light_update(m_world, static_cast<uint32_t>(id), &d));
            m_lightSent[id] = d;
            m_shadowCacheValid = false;
        }
    }
}

/** Answer a flags == 0 cull from the prefetched shadow-view lists, prefetching them first if this is the first shadow view
 *  since the last kCullLights cull. Returns false when the view is not a shadow view of a light that cull reported. */
bool CudaRenderWorld::serveShadowView(const OvisView &v, CullResults &outResults) const {
    if (m_visibleShadowLights.empty() || v.vp_width != v.vp_height || v.vp_width <= 0)
        return false;
    if (!m_shadowCacheValid || (m_cachedViews.size() && m_cachedViews[0].vp_width != v.vp_width)) {
        /* Is it one of their shadow views at all? (RenderLight::updateShadowViews, restated by ovis_light_shadow_views) */
        m_cachedViews.clear();
        bool mine = false;
        for (uint32_t id : m_visibleShadowLights) {
            OvisView views[6];
            int32_t count = 0;
            OVIS_CHECK(ovis_light_shadow_views(&m_lightSent[id], v.vp_width, views, &count));
            for (int32_t i = 0; i < count; i++) {
                mine = mine || memcmp(&views[i], &v, sizeof(v)) == 0;
                m_cachedViews.push_back(views[i]);
            }
        }
        if (!mine) {
            m_cachedViews.clear();
            return false;
        }
        const int32_t capLists = static_cast<int32_t>(m_cachedViews.size());
        m_cachedListLight.resize(capLists);
        m_cachedListView.resize(capLists);
        m_cachedOffsets.resize(capLists + 1);
        if (m_cachedEntities.size() < 1024)
            m_cachedEntities.resize(1024);
        for (;;) {
            int32_t numLists = 0;
            int64_t numEntities = 0;
            const int rc = ovis_cull_shadow_views(m_world, static_cast<int32_t>(m_visibleShadowLights.size()), m_visibleShadowLights.data(),
                                                  v.vp_width, 0, capLists, m_cachedListLight.data(), m_cachedListView.data(),
                                                  m_cachedOffsets.data(), &numLists, static_cast<int64_t>(m_cachedEntities.size()),
                                                  m_cachedEntities.data(), &numEntities);
            if (rc == OVIS_ERR_CAPACITY && numLists <= capLists) {
                m_cachedEntities.resize(static_cast<size_t>(numEntities + numEntities / 4 + 1024));
                continue;
            }
            OVIS_CHECK(rc);
            break;
        }
        m_shadowCacheValid = true;
        m_batches++;
    }
    for (size_t k = 0; k < m_cachedViews.size(); k++) {
        if (memcmp(&m_cachedViews[k], &v, sizeof(v)) != 0)
            continue;
        for (int64_t i = m_cachedOffsets[k]; i < m_cachedOffsets[k + 1]; i++)
            outResults.entities.emplace_back(m_entities[m_cachedEntities[i]]);
        m_served++;
        return true;
    }
    return false;
}

void CudaRenderWorld::cull(RenderView &view, CullResults &outResults, uint32_t flags) const {
    flush();
    const OvisView v = describeView(view);
    if (flags & kCullLights) {
        syncLights();
    } else if (serveShadowView(v, outResults)) {
        return;                                              /* one of this frame's shadow views: already culled */
    }

    const size_t needed = m_entities.size() ? m_entities.size() : 1;
    if (needed > m_entityResultCapacity) {
        if (m_entityResult)
            OVIS_CHECK(ovis_host_free(m_entityResult));
        m_entityResult = nullptr;
        m_entityResultCapacity = needed + needed / 2 + 1024;
        void *ptr = nullptr;
        OVIS_CHECK(ovis_host_alloc(static_cast<int64_t>(m_entityResultCapacity * sizeof(uint32_t)), &ptr));
        m_entityResult = static_cast<uint32_t *>(ptr);
    }
    m_lightResult.resize(m_lights.size() ? m_lights.size() : 1);
    int64_t numEntities = 0;
    int32_t numLights = 0;
    OVIS_CHECK(ovis_cull(m_world, &v, (flags & kCullLights) ? OVIS_CULL_LIGHTS : 0,
                         m_entityResult, static_cast<int64_t>(m_entityResultCapacity), &numEntities,
                         m_lightResult.data(), static_cast<int32_t>(m_lightResult.size()), &numLights));

    /* Appended, never cleared, in insertion order - as simple_render_world.cc:38-47 does. */
    for (int64_t i = 0; i < numEntities; i++)
        outResults.entities.emplace_back(m_entities[m_entityResult[i]]);
    for (int32_t i = 0; i < numLights; i++)
        outResults.lights.emplace_back(m_lights[m_lightResult[i]]);

    if (flags & kCullLights) {
        /* The lights whose shadow views the pipeline will ask for next (deferred_render_pipeline.cc:257-260). */
        m_visibleShadowLights.clear();
        for (int32_t i = 0; i < numLights; i++)
            if (m_lightSent[m_lightResult[i]].flags & OVIS_CASTS_SHADOW)
                m_visibleShadowLights.push_back(m_lightResult[i]);
        m_shadowCacheValid = false;
    }
}
