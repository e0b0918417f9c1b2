/* TEST INFRASTRUCTURE - stand-in for the reference's render_core/shader.h, seen ONLY by the two translation units that
 * pin draw-list construction: the reference's own render/src/draw_list.cc (compiled where it lies) and
 * oracle/draw_harness.cc. The real Shader is an Asset with objgen-generated reflection and GPU programs, neither of
 * which exists headless; DrawList::add (draw_list.cc:35-45) only asks it for numPasses(type) and getPass(type, i), so
 * this class answers exactly those two questions from a table of pass counts. Not a copy of the reference's header. */
#pragma once

#include <cstddef>
#include <map>
#include <string>
#include <vector>

#include "gpu/command_list.h"

class Pass {
public:
    explicit Pass(size_t index) : index(index) {}
    void setDrawState(GPUCommandList *, const ShaderKeywordSet &) const {}
    size_t index;                       /* position inside its pass type: what a draw's pass pointer resolves to */
};

class Shader {
public:
    void setPasses(const std::string &type, size_t count) {
        std::vector<Pass> &passes = m_passes[type];
        passes.clear();
        for (size_t i = 0; i < count; i++)
            passes.emplace_back(i);
    }
    size_t numPasses(const std::string &type) const {
        auto it = m_passes.find(type);
        return it != m_passes.end() ? it->second.size() : 0;
    }
    const Pass *getPass(const std::string &type, size_t index) const {
        auto it = m_passes.find(type);
        return it != m_passes.end() && index < it->second.size() ? &it->second[index] : nullptr;
    }
private:
    std::map<std::string, std::vector<Pass>> m_passes;
};
