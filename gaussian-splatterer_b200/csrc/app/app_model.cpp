// ModelSplatsHost (src/ModelSplatsHost.{h,cpp}), the .gobj text format (UiFrame::saveSplats / loadSplats,
// src/ui/UiFrame.cpp:333-356,372-450), the three field initialisers (UiFrame::initFieldGrid / Mono / Model,
// src/ui/UiFrame.cpp:137-264) and the OBJ subset both the tracer and initFieldModel read
// (src/rtx/RtxHost.cpp:107-180, src/ui/UiFrame.cpp:185-229).  Host C++, no GUI.
#include "app_common.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <fstream>
#include <optional>
#include <sstream>
#include <stdexcept>

namespace gsbapp {

ModelHost::ModelHost(int capacity_, int shDegree_, int shCoeffs_) : capacity(capacity_), shDegree(shDegree_), shCoeffs(shCoeffs_) {
    // the reference news capacity-sized arrays (MH:9-14); vectors grow lazily here, the capacity rule is kept in push_back
}

void ModelHost::push_back(const float loc[3], const float* sh, int n_sh, const float scale[3], float opacity, const float rot[4]) {
    if (count >= capacity) throw std::runtime_error("Model ran out of capacity!");                 // MH:73
    if (n_sh < shCoeffs * 3) throw std::out_of_range("vector::_M_range_check: sh has fewer than 3 * shCoeffs entries");   // sh.at(i), MH:77
    locations.insert(locations.end(), loc, loc + 3);
    shs.insert(shs.end(), sh, sh + shCoeffs * 3);
    scales.insert(scales.end(), scale, scale + 3);
    opacities.push_back(opacity);
    rotations.insert(rotations.end(), rot, rot + 4);
    count++;
}

void ModelHost::copy(int to, int from) {
    if (to < 0 || to >= count || from < 0 || from >= count)
        throw std::runtime_error("Can't copy splat in model, incorrect bounds and/or no capacity!");   // MH:87-89
    memmove(&locations[(size_t)to * 3], &locations[(size_t)from * 3], 3 * sizeof(float));
    memmove(&shs[(size_t)to * 3 * shCoeffs], &shs[(size_t)from * 3 * shCoeffs], (size_t)3 * shCoeffs * sizeof(float));
    memmove(&scales[(size_t)to * 3], &scales[(size_t)from * 3], 3 * sizeof(float));
    opacities[(size_t)to] = opacities[(size_t)from];
    memmove(&rotations[(size_t)to * 4], &rotations[(size_t)from * 4], 4 * sizeof(float));
}

// One corner "v/vt/vn": digits accumulated exactly like the reference (no sign, no validation), RH:137-150 / UF:208-221.
static void parse_corner(const std::string& s, int out[3]) {
    out[0] = out[1] = out[2] = 0;
    size_t k = 0;
    for (int part = 0; part < 3; part++) {
        while (k < s.length() && s[k] != '/') out[part] = out[part] * 10 + (s[k++] - '0');
        k++;
    }
}

void parse_obj(const std::string& path, ObjMesh& out, bool want_texcoords) {
    std::vector<float> ref_uv;            // textureCoordsRefVec
    std::vector<int> corner_uv;           // textureCoordsVec (1-based vt index per corner, 0 = none)
    std::ifstream ifs(path);
    std::string line;
    while (getline(ifs, line)) {
        std::istringstream iss(line);
        std::string prefix;
        iss >> prefix;
        if (prefix == "v") {
            float x = 0, y = 0, z = 0;
            iss >> x >> y >> z;
            out.vertices.insert(out.vertices.end(), {x, y, z});
        } else if (prefix == "vt") {
            float u = 0, v = 0;
            iss >> u >> v;
            ref_uv.insert(ref_uv.end(), {u, v});
        } else if (prefix == "f") {
            std::vector<std::array<int, 3>> fv;
            std::string tok;
            while (iss >> tok) {
                std::array<int, 3> c;
                parse_corner(tok, c.data());
                fv.push_back(c);
            }
            auto tri = [&](int a, int b, int c) {
                out.triangles.insert(out.triangles.end(), {fv[a][0] - 1, fv[b][0] - 1, fv[c][0] - 1});
                corner_uv.insert(corner_uv.end(), {fv[a][1], fv[b][1], fv[c][1]});
            };
            if (fv.size() == 4) { tri(0, 1, 2); tri(0, 2, 3); }
            else if (fv.size() == 3) tri(0, 1, 2);
            else throw std::runtime_error("Unexpected vertex count in face list!" + std::to_string(fv.size()));
        }
    }
    if (!want_texcoords) return;
    const size_t nt = out.triangles.size() / 3;
    for (size_t i = 0; i < nt; i++) {                                   // RH:166-178
        const int a = corner_uv[i * 3], b = corner_uv[i * 3 + 1], c = corner_uv[i * 3 + 2];
        if (a > 0 && b > 0 && c > 0) {
            for (int idx : {a, b, c}) {
                if ((size_t)idx * 2 > ref_uv.size()) throw std::runtime_error("OBJ texture coordinate index out of range");
                out.texcoords.push_back(ref_uv[(size_t)(idx - 1) * 2]);
                out.texcoords.push_back(ref_uv[(size_t)(idx - 1) * 2 + 1]);
            }
        } else out.texcoords.insert(out.texcoords.end(), 6, 0.0f);
    }
}

}  // namespace gsbapp

using namespace gsbapp;

namespace {

const int SPLATS_LIMIT = 1000000, SPLATS_SH_DEGREE = 1, SPLATS_SH_COEF = 4;      // src/Config.h:17-20

template <typename F>
int guarded(F&& f) {
    try {
        f();
    } catch (const std::exception& e) {
        return fail(e.what());
    }
    return GSB_OK;
}

// glm::angleAxis(angle, axis) in memory order (w, x, y, z): the axis is NOT normalised (ext/quaternion_trigonometric.inl:30-36)
void angle_axis_quat(float angle, const float axis[3], float out[4]) {
    const float s = std::sin(angle * 0.5f);
    out[0] = std::cos(angle * 0.5f);
    out[1] = axis[0] * s; out[2] = axis[1] * s; out[3] = axis[2] * s;
}

}  // namespace

extern "C" {

int gsb_model_host_create(gsb_model_host** out, int capacity, int shDegree, int shCoeffs) {
    if (!out || capacity < 0 || shCoeffs < 0) return fail("gsb_model_host_create: invalid argument");
    *out = new gsb_model_host(capacity, shDegree, shCoeffs);
    return GSB_OK;
}

int gsb_model_host_from_arrays(gsb_model_host** out, const float* locations, size_t n_locations, const float* shs, size_t n_shs,
                               const float* scales, size_t n_scales, const float* opacities, size_t n_opacities,
                               const float* rotations, size_t n_rotations) {
    if (!out) return fail("gsb_model_host_from_arrays: invalid argument");
    // MH:31-63
    size_t capacity = 1000000;
    while (capacity < n_locations / 3) capacity *= 10;
    const int count = (int)n_locations / 3;
    if (count == 0) return fail("Inconsistent feature dimensions supplied when creating a host model!");   // the reference divides by zero here
    const int shDegree = (((int)n_shs / (3 * count)) - 1) / 3;
    const int shCoeffs = ((int)n_shs / (3 * count));
    if (n_locations != (size_t)count * 3 || n_shs != (size_t)count * 3 * shCoeffs || n_scales != (size_t)count * 3 ||
        n_opacities != (size_t)count || n_rotations != (size_t)count * 4)
        return fail("Inconsistent feature dimensions supplied when creating a host model!");
    gsb_model_host* h = new gsb_model_host((int)capacity, shDegree, shCoeffs);
    h->m.count = count;
    h->m.locations.assign(locations, locations + n_locations);
    h->m.shs.assign(shs, shs + n_shs);
    h->m.scales.assign(scales, scales + n_scales);
    h->m.opacities.assign(opacities, opacities + n_opacities);
    h->m.rotations.assign(rotations, rotations + n_rotations);
    *out = h;
    return GSB_OK;
}

int gsb_model_host_destroy(gsb_model_host* m) { delete m; return GSB_OK; }

int gsb_model_host_get(gsb_model_host* h, gsb_model_host_view* out) {
    if (!h || !out) return fail("gsb_model_host_get: invalid argument");
    ModelHost& m = h->m;
    out->capacity = m.capacity; out->shDegree = m.shDegree; out->shCoeffs = m.shCoeffs; out->count = m.count;
    out->locations = m.locations.data(); out->shs = m.shs.data(); out->scales = m.scales.data();
    out->opacities = m.opacities.data(); out->rotations = m.rotations.data();
    return GSB_OK;
}

int gsb_model_host_push_back(gsb_model_host* h, const float location[3], const float* sh, int n_sh, const float scale[3],
                             float opacity, const float rotation_wxyz[4]) {
    if (!h || !location || (!sh && n_sh > 0) || !scale || !rotation_wxyz) return fail("gsb_model_host_push_back: invalid argument");
    return guarded([&] { h->m.push_back(location, sh, n_sh, scale, opacity, rotation_wxyz); });
}

int gsb_model_host_copy(gsb_model_host* h, int indexTo, int indexFrom) {
    if (!h) return fail("gsb_model_host_copy: invalid argument");
    return guarded([&] { h->m.copy(indexTo, indexFrom); });
}

// UiFrame::saveSplats, UF:333-356: `file << float` = ostream default formatting (%g, 6 significant digits).
int gsb_gobj_save(gsb_model_host* h, const char* path) {
    if (!h || !path) return fail("gsb_gobj_save: invalid argument");
    const ModelHost& model = h->m;
    std::ofstream file(path);
    if (!file) return fail(std::string("Failed to write splats file at \"") + path + "\"!");
    for (int i = 0; i < model.count; i++) {
        file << "v " << model.locations[(size_t)i * 3] << " " << model.locations[(size_t)i * 3 + 1] << " " << model.locations[(size_t)i * 3 + 2] << "\n";
        file << "sh";
        for (int f = 0; f < model.shCoeffs * 3; f++) file << " " << model.shs[(size_t)i * 3 * model.shCoeffs + f];
        file << "\n";
        file << "s " << model.scales[(size_t)i * 3] << " " << model.scales[(size_t)i * 3 + 1] << " " << model.scales[(size_t)i * 3 + 2] << "\n";
        file << "a " << model.opacities[(size_t)i] << "\n";
        file << "r " << model.rotations[(size_t)i * 4] << " " << model.rotations[(size_t)i * 4 + 1] << " " << model.rotations[(size_t)i * 4 + 2] << " "
             << model.rotations[(size_t)i * 4 + 3] << "\n";
    }
    return GSB_OK;
}

// UiFrame::loadSplats, UF:372-450.
int gsb_gobj_load(gsb_model_host** out, const char* path) {
    if (!out || !path) return fail("gsb_gobj_load: invalid argument");
    std::ifstream file(path);
    if (!file) return fail(std::string("Failed to load splats file at \"") + path + "\"!");          // UF:374 dialog text
    std::optional<int> shCoeffs;
    std::vector<float> locations, shs, scales, opacities, rotations;
    std::string line;
    try {
        while (std::getline(file, line)) {
            std::istringstream iss(line);
            std::string prefix;
            iss >> prefix;
            auto read_n = [&](std::vector<float>& dst, int n) {
                for (int f = 0; f < n; f++) { float x = 0.0f; iss >> x; dst.push_back(x); }      // a failed read pushes 0, as operator>> does since C++11
            };
            if (prefix == "v") read_n(locations, 3);
            else if (prefix == "sh") {
                int n = 0;
                float x;
                while (iss >> x) { shs.push_back(x); n++; }
                if (!shCoeffs) shCoeffs = n;
                else if (shCoeffs != n) throw std::runtime_error("Inconsistent SH degree!");
            } else if (prefix == "s") read_n(scales, 3);
            else if (prefix == "a") read_n(opacities, 1);
            else if (prefix == "r") read_n(rotations, 4);
        }
    } catch (const std::exception& e) {
        return fail(e.what());
    }
    return gsb_model_host_from_arrays(out, locations.data(), locations.size(), shs.data(), shs.size(), scales.data(), scales.size(),
                                      opacities.data(), opacities.size(), rotations.data(), rotations.size());
}

int gsb_field_grid(gsb_model_host** out) {                              // UF:137-160
    if (!out) return fail("gsb_field_grid: invalid argument");
    gsb_model_host* h = new gsb_model_host(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    static const float dim = 4.0f, step = 0.5f;
    const std::vector<float> shs((size_t)3 * h->m.shCoeffs, 0.0f);
    const float axis[3] = {0.0f, 1.0f, 0.0f};
    float rot[4];
    angle_axis_quat(0.0f, axis, rot);
    const float scale[3] = {step * 0.1f, step * 0.1f, step * 0.1f};
    for (float x = -dim; x <= dim; x += step)
        for (float y = -dim; y <= dim; y += step)
            for (float z = -dim; z <= dim; z += step) {
                const float loc[3] = {x, y, z};
                h->m.push_back(loc, shs.data(), (int)shs.size(), scale, 1.0f, rot);
            }
    *out = h;
    return GSB_OK;
}

int gsb_field_mono(gsb_model_host** out) {                              // UF:162-176
    if (!out) return fail("gsb_field_mono: invalid argument");
    gsb_model_host* h = new gsb_model_host(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    const std::vector<float> shs((size_t)3 * h->m.shCoeffs, 0.0f);
    const float axis[3] = {0.0f, 1.0f, 0.0f}, loc[3] = {0.0f, 0.0f, 0.0f}, scale[3] = {0.3f, 0.3f, 0.3f};
    float rot[4];
    angle_axis_quat(0.0f, axis, rot);
    h->m.push_back(loc, shs.data(), (int)shs.size(), scale, 1.0f, rot);
    *out = h;
    return GSB_OK;
}

int gsb_field_model(gsb_model_host** out, const char* obj_path) {       // UF:178-264
    if (!out || !obj_path) return fail("gsb_field_model: invalid argument");
    {
        std::ifstream probe(obj_path);
        if (!probe) return fail("Failed to load the model file for splat field generation!");    // UF:180 dialog text
    }
    ObjMesh mesh;
    gsb_model_host* h = nullptr;
    const int rc = guarded([&] {
        parse_obj(obj_path, mesh, false);
        h = new gsb_model_host(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
        const std::vector<float> shs((size_t)3 * h->m.shCoeffs, 0.0f);
        const size_t nv = mesh.vertices.size() / 3;
        for (size_t t = 0; t + 2 < mesh.triangles.size(); t += 3) {
            const float* v[3];
            for (int k = 0; k < 3; k++) {
                const int idx = mesh.triangles[t + k];
                if (idx < 0 || (size_t)idx >= nv) throw std::runtime_error("OBJ face references a vertex that does not exist");
                v[k] = &mesh.vertices[(size_t)idx * 3];
            }
            // glm vec3 arithmetic in the reference's order (UF:238-254)
            float loc[3], e1[3], e2[3];
            for (int k = 0; k < 3; k++) {
                loc[k] = ((v[0][k] + v[1][k]) + v[2][k]) / 3.0f;
                e1[k] = v[1][k] - v[0][k];
                e2[k] = v[2][k] - v[0][k];
            }
            auto dot3 = [](const float* a, const float* b) { const float x = a[0] * b[0], y = a[1] * b[1], z = a[2] * b[2]; return x + y + z; };
            float scale[3] = {std::sqrt(dot3(e1, e1)), std::sqrt(dot3(e2, e2)), 0.005f};
            for (float& s : scale) s *= 0.2f;
            float n[3] = {e1[1] * e2[2] - e2[1] * e1[2], e1[2] * e2[0] - e2[2] * e1[0], e1[0] * e2[1] - e2[0] * e1[1]};
            const float inv = 1.0f / std::sqrt(dot3(n, n));
            for (float& c : n) c *= inv;
            const float up[3] = {0.0f, 0.0f, 1.0f};
            const float axis[3] = {up[1] * n[2] - n[1] * up[2], up[2] * n[0] - n[2] * up[0], up[0] * n[1] - n[0] * up[1]};
            const float angle = std::acos(dot3(up, n));
            float rot[4];
            angle_axis_quat(angle, axis, rot);
            h->m.push_back(loc, shs.data(), (int)shs.size(), scale, 1.0f, rot);
        }
    });
    if (rc != GSB_OK) { delete h; return rc; }
    *out = h;
    return GSB_OK;
}

}  // extern "C"
