// TEST STAND-IN for the slice of glm the application's Camera exposes: vec3, column-major mat4,
// mat4 * mat4, lookAt (right-handed), perspective (right-handed, -1..1 depth), unary minus.
#pragma once
#include <cmath>
namespace glm {
struct vec3 {
    float x, y, z;
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline vec3 normalize(vec3 a) { const float s = 1.0f / std::sqrt(dot(a, a)); return {a.x * s, a.y * s, a.z * s}; }
struct mat4 {
    float c[4][4];   // c[column][row]
    float* operator[](int i) { return c[i]; }
    const float* operator[](int i) const { return c[i]; }
};
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; j++)
        for (int i = 0; i < 4; i++) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2] + a.c[3][i] * b.c[j][3];
    return r;
}
inline mat4 operator-(const mat4& a) {
    mat4 r;
    for (int j = 0; j < 4; j++) for (int i = 0; i < 4; i++) r.c[j][i] = -a.c[j][i];
    return r;
}
inline mat4 lookAt(vec3 eye, vec3 center, vec3 up) {
    const vec3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
    mat4 r = {{{s.x, u.x, -f.x, 0}, {s.y, u.y, -f.y, 0}, {s.z, u.z, -f.z, 0}, {-dot(s, eye), -dot(u, eye), dot(f, eye), 1}}};
    return r;
}
inline mat4 perspective(float fovy, float aspect, float zn, float zf) {
    const float t = std::tan(fovy / 2.0f);
    mat4 r = {};
    r.c[0][0] = 1.0f / (aspect * t); r.c[1][1] = 1.0f / t; r.c[2][2] = -(zf + zn) / (zf - zn); r.c[2][3] = -1.0f;
    r.c[3][2] = -(2.0f * zf * zn) / (zf - zn);
    return r;
}
inline float radians(float d) { return d * 0.01745329251994329576923690768489f; }
}  // namespace glm
