// Camera maths of the application layer: src/Camera.cpp:9-86 (Fibonacci spheres, preview camera, view / projection) in
// float32 and in the operation order of the glm 0.9.9.9 functions the reference calls (lookAtRH, perspectiveRH_NO,
// mat4 * mat4, angleAxis -> mat4_cast, inverse), so matrices agree with the application's to the last bit.
// Pinned by tests/golden/cameras.json and tests/golden/app_project.json (generated from the reference's own sources).
#include "app_common.h"

#include <cmath>
#include <cstring>

namespace gsbapp {

namespace {
struct V3 { float x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(V3 a, V3 b) { const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z; return tx + ty + tz; }   // glm compute_dot<vec3>
inline V3 cross(V3 x, V3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
inline V3 normalize(V3 v) { const float s = 1.0f / std::sqrt(dot(v, v)); return {v.x * s, v.y * s, v.z * s}; }      // v * inversesqrt(dot)
}  // namespace

Mat4 look_at_rh(const float eye[3], const float center[3], const float up[3]) {
    const V3 e{eye[0], eye[1], eye[2]}, c{center[0], center[1], center[2]}, u0{up[0], up[1], up[2]};
    const V3 f = normalize(sub(c, e));
    const V3 s = normalize(cross(f, u0));
    const V3 u = cross(s, f);
    Mat4 r;
    memset(&r, 0, sizeof(r));
    r.m[0][0] = s.x; r.m[1][0] = s.y; r.m[2][0] = s.z;
    r.m[0][1] = u.x; r.m[1][1] = u.y; r.m[2][1] = u.z;
    r.m[0][2] = -f.x; r.m[1][2] = -f.y; r.m[2][2] = -f.z;
    r.m[3][0] = -dot(s, e);
    r.m[3][1] = -dot(u, e);
    r.m[3][2] = dot(f, e);
    r.m[3][3] = 1.0f;
    return r;
}

Mat4 perspective_rh_no(float fovy, float aspect, float z_near, float z_far) {
    const float t = std::tan(fovy / 2.0f);
    Mat4 r;
    memset(&r, 0, sizeof(r));
    r.m[0][0] = 1.0f / (aspect * t);
    r.m[1][1] = 1.0f / t;
    r.m[2][2] = -(z_far + z_near) / (z_far - z_near);
    r.m[2][3] = -1.0f;
    r.m[3][2] = -(2.0f * z_far * z_near) / (z_far - z_near);
    return r;
}

Mat4 mat_mul(const Mat4& a, const Mat4& b) {          // glm operator*(mat4, mat4): column c = a0*b[c][0] + a1*b[c][1] + a2*b[c][2] + a3*b[c][3]
    Mat4 r;
    for (int c = 0; c < 4; c++)
        for (int k = 0; k < 4; k++)
            r.m[c][k] = a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1] + a.m[2][k] * b.m[c][2] + a.m[3][k] * b.m[c][3];
    return r;
}

Mat4 angle_axis_mat(float angle, float ax, float ay, float az) {     // (glm::mat4) glm::angleAxis(angle, axis)
    const float s = std::sin(angle * 0.5f);
    const float w = std::cos(angle * 0.5f);
    const float x = ax * s, y = ay * s, z = az * s;
    const float qxx = x * x, qyy = y * y, qzz = z * z, qxz = x * z, qxy = x * y, qyz = y * z, qwx = w * x, qwy = w * y, qwz = w * z;
    Mat4 r;
    memset(&r, 0, sizeof(r));
    r.m[0][0] = 1.0f - 2.0f * (qyy + qzz); r.m[0][1] = 2.0f * (qxy + qwz); r.m[0][2] = 2.0f * (qxz - qwy);
    r.m[1][0] = 2.0f * (qxy - qwz); r.m[1][1] = 1.0f - 2.0f * (qxx + qzz); r.m[1][2] = 2.0f * (qyz + qwx);
    r.m[2][0] = 2.0f * (qxz + qwy); r.m[2][1] = 2.0f * (qyz - qwx); r.m[2][2] = 1.0f - 2.0f * (qxx + qyy);
    r.m[3][3] = 1.0f;
    return r;
}

Mat4 inverse(const Mat4& mm) {                        // glm detail::compute_inverse<4,4> (func_matrix.inl)
    const float (*m)[4] = mm.m;
    const float c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3], c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3], c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3], c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3], c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2], c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2], c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3], c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3], c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2], c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2], c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1], c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1], c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    const float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    const float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    const float v0[4] = {m[1][0], m[0][0], m[0][0], m[0][0]}, v1[4] = {m[1][1], m[0][1], m[0][1], m[0][1]};
    const float v2[4] = {m[1][2], m[0][2], m[0][2], m[0][2]}, v3[4] = {m[1][3], m[0][3], m[0][3], m[0][3]};
    const float sa[4] = {1.0f, -1.0f, 1.0f, -1.0f}, sb[4] = {-1.0f, 1.0f, -1.0f, 1.0f};
    Mat4 inv;
    for (int k = 0; k < 4; k++) {
        inv.m[0][k] = (v1[k] * f0[k] - v2[k] * f1[k] + v3[k] * f2[k]) * sa[k];
        inv.m[1][k] = (v0[k] * f0[k] - v2[k] * f3[k] + v3[k] * f4[k]) * sb[k];
        inv.m[2][k] = (v0[k] * f1[k] - v1[k] * f3[k] + v3[k] * f5[k]) * sa[k];
        inv.m[3][k] = (v0[k] * f2[k] - v1[k] * f4[k] + v2[k] * f5[k]) * sb[k];
    }
    const float d0 = m[0][0] * inv.m[0][0], d1 = m[0][1] * inv.m[1][0], d2 = m[0][2] * inv.m[2][0], d3 = m[0][3] * inv.m[3][0];
    const float det = (d0 + d1) + (d2 + d3);
    const float ood = 1.0f / det;
    for (int c = 0; c < 4; c++)
        for (int k = 0; k < 4; k++) inv.m[c][k] *= ood;
    return inv;
}

Mat4 camera_view(const gsb_camera_desc& c) {          // -glm::lookAt(location, target, (0,1,0)), CA:79-82
    const float up[3] = {0.0f, 1.0f, 0.0f};
    Mat4 v = look_at_rh(c.location, c.target, up);
    for (int a = 0; a < 4; a++)
        for (int b = 0; b < 4; b++) v.m[a][b] = -v.m[a][b];
    return v;
}

Mat4 camera_projection(const gsb_camera_desc& c, float aspect) {     // CA:84-86
    return perspective_rh_no(radians(c.fovDegY), aspect, 0.1f, 100.0f);
}

namespace {

// getFibonacciSphere, CA:9-27
void fibonacci_sphere(int count, float distance, std::vector<V3>& out) {
    static const float goldenRatio = (1.0f + sqrtf(5.0f)) / 2.0f;
    static const float angleStep = 2.0f * (float)3.1415926535897932384626433832795 * goldenRatio;
    for (int i = 0; i < count; i++) {
        const float t = (float)i / (float)count;
        const float angle1 = acosf(1.0f - 2.0f * t);
        const float angle2 = angleStep * (float)i;
        out.push_back({sinf(angle1) * cosf(angle2) * distance, sinf(angle1) * sinf(angle2) * distance, cosf(angle1) * distance});
    }
}

void sphere_cameras(const gsb_camera_sphere& s, std::vector<gsb_camera_desc>& out) {     // CA:39-46
    const Mat4 rot = mat_mul(angle_axis_mat(radians(s.rotX), 0.0f, 1.0f, 0.0f), angle_axis_mat(radians(s.rotY), 1.0f, 0.0f, 0.0f));
    std::vector<V3> locs;
    fibonacci_sphere(s.count, s.distance, locs);
    for (const V3& l : locs) {
        float v[4];
        for (int k = 0; k < 4; k++) {                 // glm mat4 * vec4: Mov0*m[0] + Mov1*m[1] then + (Mov2*m[2] + Mov3*m[3])
            const float a0 = rot.m[0][k] * l.x, a1 = rot.m[1][k] * l.y, a2 = rot.m[2][k] * l.z, a3 = rot.m[3][k] * 1.0f;
            v[k] = (a0 + a1) + (a2 + a3);
        }
        gsb_camera_desc c;
        c.location[0] = v[0] / v[3]; c.location[1] = v[1] / v[3]; c.location[2] = v[2] / v[3];
        c.target[0] = c.target[1] = c.target[2] = 0.0f;
        c.fovDegY = s.fovDeg;
        out.push_back(c);
    }
}

}  // namespace
}  // namespace gsbapp

using namespace gsbapp;

extern "C" {

int gsb_cameras_count(const gsb_project_data* d) { return d ? d->sphere1.count + d->sphere2.count : 0; }

int gsb_cameras_get(const gsb_project_data* d, gsb_camera_desc* out, int cap) {
    if (!d || (!out && cap > 0)) return fail("gsb_cameras_get: invalid argument");
    std::vector<gsb_camera_desc> cams;
    sphere_cameras(d->sphere1, cams);
    sphere_cameras(d->sphere2, cams);
    if ((int)cams.size() > cap) return fail("gsb_cameras_get: output array too small");
    for (size_t i = 0; i < cams.size(); i++) out[i] = cams[i];
    return GSB_OK;
}

int gsb_camera_preview(const gsb_project_data* d, gsb_camera_desc* out) {     // Camera::getPreviewCamera, CA:60-76
    if (!d || !out) return fail("gsb_camera_preview: invalid argument");
    if (d->previewTruth) {
        std::vector<gsb_camera_desc> cams;
        sphere_cameras(d->sphere1, cams);
        sphere_cameras(d->sphere2, cams);
        if (d->previewTruthIndex < 0 || d->previewTruthIndex >= (int)cams.size())
            return fail("vector::_M_range_check: previewTruthIndex out of range");      // std::vector::at throws in the reference
        *out = cams[(size_t)d->previewTruthIndex];
        return GSB_OK;
    }
    const float degRotOrbit = d->previewFreeOrbit ? d->previewTimer * d->previewFreeOrbitSpeed : 0.0f;
    const Mat4 rot = mat_mul(angle_axis_mat(radians(d->previewFreeRotY) + degRotOrbit, 0.0f, 1.0f, 0.0f),
                             angle_axis_mat(radians(d->previewFreeRotX), 1.0f, 0.0f, 0.0f));
    const float l[4] = {0.0f, 0.0f, -d->previewFreeDistance, 1.0f};
    float v[4];
    for (int k = 0; k < 4; k++) {
        const float a0 = rot.m[0][k] * l[0], a1 = rot.m[1][k] * l[1], a2 = rot.m[2][k] * l[2], a3 = rot.m[3][k] * l[3];
        v[k] = (a0 + a1) + (a2 + a3);
    }
    out->location[0] = v[0] / v[3]; out->location[1] = v[1] / v[3]; out->location[2] = v[2] / v[3];
    out->target[0] = out->target[1] = out->target[2] = 0.0f;
    out->fovDegY = d->previewFreeFovDeg;
    return GSB_OK;
}

int gsb_camera_constants(const gsb_camera_desc* c, int width, int height, int square_fov, gsb_camera* out) {
    if (!c || !out || width <= 0 || height <= 0) return fail("gsb_camera_constants: invalid argument");
    const Mat4 view = camera_view(*c);
    const Mat4 projview = mat_mul(camera_projection(*c, (float)width / (float)height), view);
    memcpy(out->view, view.m, sizeof(out->view));
    memcpy(out->projview, projview.m, sizeof(out->projview));
    for (int k = 0; k < 3; k++) out->location[k] = c->location[k];
    out->tan_fovy = std::tan(radians(c->fovDegY) * 0.5f);
    out->tan_fovx = square_fov ? out->tan_fovy : std::tan(radians((float)width * c->fovDegY / (float)height) * 0.5f);
    return GSB_OK;
}

int gsb_camera_inverse_projview(const gsb_camera_desc* c, float aspect, float out16[16]) {
    if (!c || !out16) return fail("gsb_camera_inverse_projview: invalid argument");
    const Mat4 inv = inverse(mat_mul(camera_projection(*c, aspect), camera_view(*c)));
    memcpy(out16, inv.m, 16 * sizeof(float));
    return GSB_OK;
}

}  // extern "C"
