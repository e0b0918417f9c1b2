/* TEST INFRASTRUCTURE - stand-in for the reference's render_core/material.h (see shader.h beside it): DrawList only
 * follows material()->shader(). */
#pragma once

#include "render_core/shader.h"

class Material {
public:
    explicit Material(Shader *shader) : m_shader(shader) {}
    Shader *shader() const { return m_shader; }
    void setDrawState(GPUCommandList *) const {}
private:
    Shader *m_shader;
};
