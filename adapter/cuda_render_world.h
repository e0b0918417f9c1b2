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

#include "orion_vis.h"


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

    /** Diagnostics: cull() calls answered from a prefetched batch / batches made (oracle/adapter_check.cc). */
    size_t shadowViewsServed() const { return m_served; }
    size_t shadowBatches() const { return m_batches; }
private:
    void flush() const;
    void flushAdds() const;
    void compactIfWorthIt() const;
    void syncLights() const;
    bool serveShadowView(const OvisView &view, CullResults &outResults) const;
private:
    OvisWorld *m_world;

    /** Entity/light by id (null once removed) and id by pointer. Entity ids change only in compactIfWorthIt(). */
    mutable std::vector<RenderEntity *> m_entities;
    mutable std::unordered_map<const RenderEntity *, uint32_t> m_entityIds;
    mutable size_t m_removed = 0;
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
    /** Entities added since the last cull: local boxes and flags, uploaded in one call (flushAdds). */
    mutable std::vector<float> m_newLocal;
    mutable std::vector<uint32_t> m_newFlags;

    /** What the library was last told about each light: RenderLight's setters mostly do not call updateLight()
     *  (render_light.cc:67-142), so every light is re-read at each kCullLights cull - and compared here, on the host,
     *  so that only lights that really changed cross the C ABI. */
    mutable std::vector<OvisLight> m_lightSent;

    /** One frame's shadow-view culls, prefetched in a single batched pass.
     *  DeferredRenderPipeline::prepareLights (deferred_render_pipeline.cc:257-284) calls cull(shadowView, results, 0) once per
     *  shadow view of every visible shadow-casting light: 1 + 6 x lights full world scans and host round trips. The first
     *  such call after a kCullLights cull culls ALL shadow views of the lights that cull reported (ovis_cull_shadow_views:
     *  one pass over the bounds, point-light faces through the light grid) and keeps the lists; that call and the
     *  following ones are answered from them as long as nothing was mutated and the view is one of the prefetched ones. */
    mutable bool m_shadowCacheValid;
    mutable size_t m_served = 0, m_batches = 0;
    mutable std::vector<uint32_t> m_visibleShadowLights;
    mutable std::vector<OvisView> m_cachedViews;
    mutable std::vector<int64_t> m_cachedOffsets;
    mutable std::vector<uint32_t> m_cachedEntities;
    mutable std::vector<int32_t> m_cachedListLight, m_cachedListView;
};
