/*
 * oracle/ref_shim.cc - headless stand-ins so the reference's render classes link with no graphics API.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/README.md). Compiled together with the reference's own,
 * unmodified sources (listed in oracle/Makefile) into oracle/_ref/liborion_ref.so.
 *
 * Why it is needed: RenderEntity / RenderView / RenderLight each create a uniform buffer and a
 * resource set in their constructors (engine/runtime/render/src/render_entity.cc:41-42,
 * render_view.cc:41-42, render_light.cc:45-46) through the abstract GPUManager
 * (engine/runtime/gpu/include/gpu_manager.h:53-244) and through g_renderResources
 * (engine/runtime/render_core/src/render_resources.cc:31). That bookkeeping has no effect on any
 * visibility result. Here the GPU buffer is plain host memory and every draw/texture entry point is
 * a stub; no graphics library is loaded.
 */

#include "gpu/gpu_manager.h"
#include "render_core/defs.h"
#include "render_core/render_resources.h"

#include <cstdlib>
#include <vector>

namespace {

/** GPUBuffer whose storage is host memory (interface: gpu/include/buffer.h:40-121). */
class HostBuffer : public GPUBuffer {
public:
    explicit HostBuffer(const GPUBufferDesc &desc) : GPUBuffer(desc), m_data(desc.size) {}
    void *map(size_t offset, size_t, uint32_t, uint32_t) override { return m_data.data() + offset; }
    void unmap() override {}
private:
    std::vector<char> m_data;
};

[[noreturn]] void unsupported(const char *what) {
    fprintf(stderr, "oracle shim: GPUManager::%s has no headless implementation\n", what);
    abort();
}

/** GPUManager with only the object-creation calls the render classes' constructors use. */
class HeadlessGPUManager : public GPUManager {
public:
    GPUBufferPtr createBuffer(const GPUBufferDesc &desc) override { return new HostBuffer(desc); }

    GPUPipelinePtr createPipeline(GPUPipelineDesc &&) override { unsupported("createPipeline"); }
    GPUQueryPoolPtr createQueryPool(const GPUQueryPoolDesc &) override { unsupported("createQueryPool"); }
    GPUTexturePtr createTexture(const GPUTextureDesc &) override { unsupported("createTexture"); }
    GPUTexturePtr createTextureView(const GPUTextureViewDesc &) override { unsupported("createTextureView"); }
    GPUProgramPtr createProgram(GPUProgramDesc &&) override { unsupported("createProgram"); }
    void endFrame() override {}
    void blit(const GPUTextureImageRef &, const GPUTextureImageRef &, glm::ivec2, glm::ivec2, glm::ivec2) override {
        unsupported("blit");
    }
    GPUCommandList *beginRenderPass(const GPURenderPassInstanceDesc &) override { unsupported("beginRenderPass"); }
    void submitRenderPass(GPUCommandList *) override { unsupported("submitRenderPass"); }
    void endQuery(GPUQueryPool *, uint32_t) override {}
};

} // namespace

/** Same global the reference defines at render_core/src/render_resources.cc:31. */
GlobalResource<RenderResources> g_renderResources;

/**
 * Headless RenderResources: only the three resource-set layouts the render classes ask for
 * (behaviour of render_core/src/render_resources.cc:67-86). The vertex layout and the quad /
 * sphere / cone geometry (:35-65, :88-93) need real GPU objects and are never touched by culling.
 */
RenderResources::RenderResources() {
    GPUResourceSetLayoutDesc desc;

    desc.slots.resize(ResourceSlots::kNumEntityResources);
    desc.slots[ResourceSlots::kUniforms].type = GPUResourceType::kUniformBuffer;
    m_entityResourceSetLayout = g_gpuManager->createResourceSetLayout(std::move(desc));

    desc.slots.resize(ResourceSlots::kNumViewResources);
    desc.slots[ResourceSlots::kUniforms].type = GPUResourceType::kUniformBuffer;
    m_viewResourceSetLayout = g_gpuManager->createResourceSetLayout(std::move(desc));

    desc.slots.resize(ResourceSlots::kNumLightResources);
    desc.slots[ResourceSlots::kUniforms].type = GPUResourceType::kUniformBuffer;
    desc.slots[ResourceSlots::kShadowMap].type = GPUResourceType::kTexture;
    m_lightResourceSetLayout = g_gpuManager->createResourceSetLayout(std::move(desc));
}

RenderResources::~RenderResources() {}

/** Called once by the harness before any render object is constructed. */
void oref_shim_init() {
    if (!g_gpuManager) {
        g_gpuManager = new HeadlessGPUManager;
        g_renderResources.init();
    }
}
