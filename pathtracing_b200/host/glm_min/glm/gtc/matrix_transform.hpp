// translate / rotate / scale with glm's semantics (result = m * T): see glm_min/glm/glm.hpp.
#pragma once
#include "../glm.hpp"

namespace glm {

inline mat4 translate(const mat4& m, const vec3& v){
    mat4 t(1.0f);
    t[3] = vec4(v, 1.0f);
    return m * t;
}
inline mat4 scale(const mat4& m, const vec3& v){
    mat4 s(1.0f);
    s[0].x = v.x; s[1].y = v.y; s[2].z = v.z;
    return m * s;
}
inline mat4 rotate(const mat4& m, float angle, const vec3& axisIn){
    const float c = std::cos(angle), s = std::sin(angle);
    const vec3 a = normalize(axisIn);
    const vec3 t = a * (1.0f - c);
    mat4 r(1.0f);
    r[0] = vec4(c + t.x * a.x, t.x * a.y + s * a.z, t.x * a.z - s * a.y, 0.0f);
    r[1] = vec4(t.y * a.x - s * a.z, c + t.y * a.y, t.y * a.z + s * a.x, 0.0f);
    r[2] = vec4(t.z * a.x + s * a.y, t.z * a.y - s * a.x, c + t.z * a.z, 0.0f);
    return m * r;
}

} // namespace glm
