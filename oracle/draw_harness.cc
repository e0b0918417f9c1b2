/* draw_harness.cc - TEST INFRASTRUCTURE: draw-list construction pinned by the reference's own DrawList::add.
 *
 * render/src/draw_list.cc is compiled where it lies (oracle/Makefile) against oracle/draw_stub/render_core/{shader,
 * material}.h - the real Shader / Material are assets with generated reflection and GPU programs and cannot exist
 * headless; DrawList::add only uses numPasses(type) and getPass(type, i). Real RenderEntity objects (the reference's
 * class, render_entity.cc compiled verbatim) carry a stub material each. What stays RESTATED are the two ten-line loops
 * of the pipeline that decide which list an entity goes to (deferred_render_pipeline.cc:318-330 prepareEntities, :272-283
 * the shadow-caster loop of prepareLights): the pipeline class needs a live GPUManager and compiled shaders. The draws
 * themselves - one per pass, in order - come out of the reference's DrawList. */
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#define private public                  /* DrawList keeps its draws private (std::deque<Draw> m_draws): read them back */
#include "render/draw_list.h"
#undef private
#include "render/render_entity.h"

#include "render_core/geometry.h"
#include "render_core/material.h"
#include "render_core/shader.h"

void oref_shim_init();                  /* ref_shim.cc: the headless GPUManager + RenderResources every RenderEntity needs */

namespace {

const std::string kDeferredPassType("Deferred");          /* deferred_render_pipeline.cc:32 */
const std::string kShadowCasterPassType("ShadowCaster");  /* :33 */
const std::string kBasicPassType("Basic");                /* Pass::kBasicType, render_core/src/pass.cc */

class DrawEntity : public RenderEntity {
public:
    DrawEntity(uint32_t id, uint32_t passCounts, bool castsShadow) : id(id), m_material(&m_shader) {
        m_shader.setPasses(kDeferredPassType, passCounts & 0xffu);
        m_shader.setPasses(kBasicPassType, (passCounts >> 8) & 0xffu);
        m_shader.setPasses(kShadowCasterPassType, (passCounts >> 16) & 0xffu);
        setFlags(castsShadow ? (uint32_t)kCastsShadow : 0u);
    }
    Geometry geometry() const override { return Geometry(); }
    Material *material() const override { return const_cast<Material *>(&m_material); }
    uint32_t id;
private:
    Shader m_shader;
    Material m_material;
};

void emit(const DrawList &list, uint32_t *out, int64_t *count) {
    *count = (int64_t)list.m_draws.size();
    size_t i = 0;
    for (const auto &draw : list.m_draws) {
        out[2 * i] = static_cast<const DrawEntity *>(draw.entity)->id;
        out[2 * i + 1] = (uint32_t)draw.pass->index;
        i++;
    }
}

}  // namespace

extern "C" void oref_draw_lists(int64_t n_visible, const uint32_t *visible, const uint32_t *pass_counts, const uint32_t *flags,
                                int32_t mode, uint32_t *out_draws0, uint32_t *out_draws1, int64_t out_counts[2]) {
    oref_shim_init();
    std::vector<std::unique_ptr<DrawEntity>> entities;
    entities.reserve((size_t)n_visible);
    for (int64_t k = 0; k < n_visible; k++)
        entities.emplace_back(new DrawEntity(visible[k], pass_counts[visible[k]], (flags[visible[k]] & RenderEntity::kCastsShadow) != 0));
    DrawList first, second;
    for (auto &owned : entities) {
        RenderEntity *entity = owned.get();
        Shader *shader = entity->material()->shader();
        if (mode == 0) {                                    /* prepareEntities, deferred_render_pipeline.cc:318-330 */
            if (shader->numPasses(kDeferredPassType) > 0)
                first.add(entity, kDeferredPassType);
            else if (shader->numPasses(kBasicPassType) > 0)
                second.add(entity, kBasicPassType);
        } else if (entity->castsShadow()) {                 /* prepareLights, :272-283 */
            if (shader->numPasses(kShadowCasterPassType) > 0)
                first.add(entity, kShadowCasterPassType);
        }
    }
    emit(first, out_draws0, &out_counts[0]);
    emit(second, out_draws1, &out_counts[1]);
}
