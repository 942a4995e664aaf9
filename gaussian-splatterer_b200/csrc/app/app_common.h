// Shared declarations of the headless application layer (include/gsb_app.h).  Host C++ only.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../../include/gsb_app.h"

int gsb_fail_msg(const char* msg);          // defined in rasterizer_api.cu: records the message, returns GSB_ERR_ARG

namespace gsbapp {

inline int fail(const std::string& msg) { return gsb_fail_msg(msg.c_str()); }

// ---- Project with its three strings (src/Project.h) -----------------------------------------------------------------
struct Project {
    gsb_project_data d;
    std::string perspective, pathModel, pathTextureDiffuse;
    Project();                                   // Project{} member initialisers, src/Project.h:10-62
};

// ---- float32 camera maths in glm 0.9.9.9's operation order (camera.cpp) ------------------------------------------------
struct Mat4 { float m[4][4]; };                  // m[col][row], 16 contiguous floats like glm::value_ptr
Mat4 look_at_rh(const float eye[3], const float center[3], const float up[3]);
Mat4 perspective_rh_no(float fovy, float aspect, float z_near, float z_far);
Mat4 mat_mul(const Mat4& a, const Mat4& b);
Mat4 angle_axis_mat(float angle_rad, float ax, float ay, float az);
Mat4 inverse(const Mat4& m);
Mat4 camera_view(const gsb_camera_desc& c);                      // -lookAt, src/Camera.cpp:79-82
Mat4 camera_projection(const gsb_camera_desc& c, float aspect);  // src/Camera.cpp:84-86
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }   // glm::radians<float>

// ---- ModelSplatsHost (src/ModelSplatsHost.{h,cpp}) ----------------------------------------------------------------------
struct ModelHost {
    int capacity = 0, shDegree = 0, shCoeffs = 0, count = 0;
    std::vector<float> locations, shs, scales, opacities, rotations;     // sized for `capacity` like the reference's new[]
    ModelHost(int capacity, int shDegree, int shCoeffs);
    void push_back(const float loc[3], const float* sh, int n_sh, const float scale[3], float opacity, const float rot[4]);
    void copy(int to, int from);
};

// OBJ subset the reference reads (v / vt / f with 3 or 4 corners, src/rtx/RtxHost.cpp:107-180, src/ui/UiFrame.cpp:185-229)
struct ObjMesh {
    std::vector<float> vertices;        // 3 per vertex
    std::vector<int> triangles;         // 3 per triangle, zero-based
    std::vector<float> texcoords;       // 2 per corner, 3 corners per triangle ((0,0) where a corner has no vt)
};
void parse_obj(const std::string& path, ObjMesh& out, bool want_texcoords);   // throws std::runtime_error like the reference

}  // namespace gsbapp

// opaque handles of the C ABI
struct gsb_project { gsbapp::Project p; };
struct gsb_model_host { gsbapp::ModelHost m; explicit gsb_model_host(int c, int d, int k) : m(c, d, k) {} };
