/*
 * cuda_render_world.h - RenderWorld implementation backed by liborion_vis (B200).
 *
 * Drop-in for SimpleRenderWorld (engine/runtime/render/include/simple_render_world.h:29-53): same base class,
 * same call contract, results in the same (insertion) order. Install as render/include/cuda_render_world.h and
 * create it at engine/runtime/graphics/src/graphics_system.cc:35. All visibility arithmetic runs on the GPU
 * through the C ABI of include/orion_vis.h; this class is bookkeeping only (pointer <-> id maps, pending updates).
 */
#pragma once

#include "render/render_world.h"

#include <unordered_map>
#include <vector>

struct OvisWorld;

class CudaRenderWorld : public RenderWorld {
public:
    explicit CudaRenderWorld(int device = 0);
    ~CudaRenderWorld();

    void cull(RenderView &view, CullResults &outResults, uint32_t flags) const override;

    void addEntity(RenderEntity *entity) override;
    void updateEntity(RenderEntity *entity) override;
    void removeEntity(RenderEntity *entity) override;

    void addLight(RenderLight *light) override;
    void updateLight(RenderLight *light) override;
    void removeLight(RenderLight *light) override;
private:
    void flush() const;
private:
    OvisWorld *m_world;

    /** Entity/light by id (null once removed) and id by pointer. Ids are never reused. */
    std::vector<RenderEntity *> m_entities;
    std::unordered_map<const RenderEntity *, uint32_t> m_entityIds;
    std::vector<RenderLight *> m_lights;
    std::unordered_map<const RenderLight *, uint32_t> m_lightIds;

    /** Entities whose world bounding box changed since the last cull (cull() is const in the interface). */
    mutable std::vector<uint32_t> m_pending;
    mutable std::vector<bool> m_isPending;

    /** Scratch reused between culls. */
    mutable std::vector<uint32_t> m_lightResult, m_ids;
    /** Visible entity ids land here: page-locked (ovis_host_alloc), so the cull's gather kernel writes them itself. */
    mutable uint32_t *m_entityResult;
    mutable size_t m_entityResultCapacity;
    mutable std::vector<float> m_boxes;
};
