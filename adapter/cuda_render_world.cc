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

CudaRenderWorld::CudaRenderWorld(int device) :
    m_world                (nullptr),
    m_entityResult         (nullptr),
    m_entityResultCapacity (0)
{
    OVIS_CHECK(ovis_world_create(device, &m_world));
}

CudaRenderWorld::~CudaRenderWorld() {
    if (m_entityResult)
        ovis_host_free(m_entityResult);
    ovis_world_destroy(m_world);
}

void CudaRenderWorld::addEntity(RenderEntity *entity) {
    uint32_t id = 0;
    const BoundingBox &box = entity->boundingBox();
    const float local[6] = { box.minimum.x, box.minimum.y, box.minimum.z, box.maximum.x, box.maximum.y, box.maximum.z };
    const uint32_t flags = entity->flags();
    OVIS_CHECK(ovis_entity_add(m_world, 1, local, &flags, &id));
    if (m_entities.size() <= id) {
        m_entities.resize(id + 1, nullptr);
        m_isPending.resize(id + 1, false);
    }
    m_entities[id] = entity;
    m_entityIds[entity] = id;
    updateEntity(entity);
}

void CudaRenderWorld::updateEntity(RenderEntity *entity) {
    auto it = m_entityIds.find(entity);
    if (it == m_entityIds.end())
        return;
    if (!m_isPending[it->second]) {
        m_isPending[it->second] = true;
        m_pending.push_back(it->second);
    }
}

void CudaRenderWorld::removeEntity(RenderEntity *entity) {
    auto it = m_entityIds.find(entity);
    if (it == m_entityIds.end())
        return;
    const uint32_t id = it->second;
    OVIS_CHECK(ovis_entity_remove(m_world, 1, &id));
    m_entities[id] = nullptr;
    m_entityIds.erase(it);
}

void CudaRenderWorld::addLight(RenderLight *light) {
    uint32_t id = 0;
    OvisLight d = describeLight(light);
    OVIS_CHECK(ovis_light_add(m_world, 1, &d, &id));
    if (m_lights.size() <= id)
        m_lights.resize(id + 1, nullptr);
    m_lights[id] = light;
    m_lightIds[light] = id;
}

void CudaRenderWorld::updateLight(RenderLight *light) {
    /* Nothing to do: most RenderLight setters never call this (render_light.cc:67-142), so cull() re-reads
     * every light's parameters anyway. */
}

void CudaRenderWorld::removeLight(RenderLight *light) {
    auto it = m_lightIds.find(light);
    if (it == m_lightIds.end())
        return;
    OVIS_CHECK(ovis_light_remove(m_world, it->second));
    m_lights[it->second] = nullptr;
    m_lightIds.erase(it);
}

/** Push every world bounding box that changed since the last cull, in one batch. */
void CudaRenderWorld::flush() const {
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

void CudaRenderWorld::cull(RenderView &view, CullResults &outResults, uint32_t flags) const {
    flush();

    if (flags & kCullLights) {
        for (size_t id = 0; id < m_lights.size(); id++) {
            if (m_lights[id]) {
                OvisLight d = describeLight(m_lights[id]);
                OVIS_CHECK(ovis_light_update(m_world, static_cast<uint32_t>(id), &d));
            }
        }
    }

    OvisView v;
    memcpy(v.position, &view.position(), sizeof(v.position));
    const glm::quat &q = view.orientation();
    v.orientation[0] = q.w; v.orientation[1] = q.x; v.orientation[2] = q.y; v.orientation[3] = q.z;
    v.hfov_deg = view.fov();
    v.znear = view.zNear();
    v.zfar = view.zFar();
    v.vp_width = view.viewport().width;
    v.vp_height = view.viewport().height;

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
}
