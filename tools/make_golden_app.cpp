// Generates tests/golden/app_project.json from the reference's OWN src/Project.h (nlohmann macros), src/Camera.cpp and
// glm, compiled in place (nothing copied): run tools/make_golden_app.sh where /root/reference exists.
//
//   "default_dump"   nlohmann dump of Project{} exactly as UiFrame::saveSettings writes it (src/ui/UiFrame.cpp:323-331)
//   "new_dump"       the same after UiFrame::initProject (src/ui/UiFrame.cpp:130-135)
//   "parse"          [{in, out | error}]: from_json on hand-written inputs (missing keys, int<->float, bools, wrong types)
//   "preview"        Camera::getPreviewCamera for a few projects (src/Camera.cpp:60-76)
//   "inverse"        glm::inverse(proj * view) as RtxHost::render builds it (src/rtx/RtxHost.cpp:212)
//   "field_model"    the per-triangle splat of UiFrame::initFieldModel (src/ui/UiFrame.cpp:233-254) for a few triangles
#include <cstdio>
#include <string>
#include <vector>
#include "Camera.h"
#include "Project.h"
#include <diff-gaussian-rasterization/third_party/glm/glm/gtc/type_ptr.hpp>
#include <diff-gaussian-rasterization/third_party/glm/glm/gtc/quaternion.hpp>

using nlohmann::json;

static json floats(const float* p, int n) {
    json a = json::array();
    for (int i = 0; i < n; i++) a.push_back((double)p[i]);
    return a;
}

int main(int argc, char** argv) {
    json out;
    {
        Project p;
        json j; nlohmann::to_json(j, p);
        out["default_dump"] = j.dump();
        p.sphere2.count = 0; p.sphere2.fovDeg = 30.0f;
        json k; nlohmann::to_json(k, p);
        out["new_dump"] = k.dump();
        p.perspective = "layout2|name=\"a\\b\";caption=3. Train Splats|\n\ttab";
        p.pathModel = "C:\\models\\caf\xc3\xa9.obj"; p.pathTextureDiffuse = "/tmp/tex.png";
        p.lrLocation = 1.25e-7f; p.lrSh = 123456.0f; p.paramScaleMax = 1e15f; p.paramCullSize = 1e16f; p.lrScale = 0.1f;
        p.previewTimer = 1234.5678f; p.iterations = 98765; p.previewTruth = true; p.previewFreeOrbit = false;
        p.sphere1.rotX = 359.99f; p.sphere1.rotY = -0.0f; p.paramSplitScale = 1e-5f; p.paramCloneDistance = 0.0001f;
        json m; nlohmann::to_json(m, p);
        out["modified_dump"] = m.dump();
    }
    const char* inputs[] = {
        "{}",
        "{\"iterations\": 7, \"lrSh\": 1, \"rtSamples\": 12.9, \"sphere1\": {\"count\": 3}}",
        "{\"previewTruth\": true, \"intervalCapture\": true, \"lrOpacity\": false, \"renderResX\": -5}",
        "{\"sphere2\": {\"count\": 4, \"distance\": 2.5, \"fovDeg\": 20, \"rotX\": 1e2, \"rotY\": -3.5e-1}, \"unknown\": [1, {\"a\": null}]}",
        "{\"pathModel\": \"a\\u00e9\\n\\\"q\\\"\\\\\", \"perspective\": \"\"}",
        "{\"lrLocation\": \"fast\"}",
        "{\"previewTruth\": 1}",
        "{\"sphere1\": 5}",
        "{\"pathModel\": 3}",
        "[1, 2]",
        "{\"iterations\": 1e3, \"intervalDensify\": 4294967296, \"lrRotation\": 1e-50}",
        "{\"iterations\": null}",
        "{\"iterations\": 3,}",
    };
    out["parse"] = json::array();
    for (const char* in : inputs) {
        json e; e["in"] = in;
        try {
            json j = json::parse(in);
            Project p;
            p.iterations = 555; p.lrSh = 9.0f; p.pathModel = "keep?";       // WITH_DEFAULT resets to Project{} defaults
            nlohmann::from_json(j, p);
            json k; nlohmann::to_json(k, p);
            e["out"] = k.dump();
        } catch (const std::exception& ex) {
            e["error"] = ex.what();
        }
        out["parse"].push_back(e);
    }
    out["preview"] = json::array();
    for (int c = 0; c < 4; c++) {
        Project p;
        p.sphere2.count = 3; p.sphere2.fovDeg = 25.0f; p.sphere1.rotX = 40.0f;
        if (c == 1) { p.previewTimer = 12.75f; p.previewFreeOrbitSpeed = 0.3f; p.previewFreeRotX = -15.0f; p.previewFreeRotY = 200.0f; p.previewFreeDistance = 6.5f; p.previewFreeFovDeg = 35.0f; }
        if (c == 2) { p.previewFreeOrbit = false; p.previewTimer = 99.0f; p.previewFreeRotY = 33.0f; }
        if (c == 3) { p.previewTruth = true; p.previewTruthIndex = 17; }
        Camera cam = Camera::getPreviewCamera(p);
        json j; nlohmann::to_json(j, p);
        json e; e["project"] = j.dump(); e["location"] = floats(&cam.location[0], 3); e["target"] = floats(&cam.target[0], 3);
        e["fov"] = (double)cam.fovDegY;
        out["preview"].push_back(e);
    }
    out["inverse"] = json::array();
    {
        Project p; p.sphere1.count = 5; p.sphere1.rotX = 20.0f; p.sphere1.rotY = 75.0f; p.sphere2.count = 2; p.sphere2.fovDeg = 20.0f;
        const float aspects[] = {1.0f, 1920.0f / 1080.0f, 0.5f};
        for (const Camera& cam : Camera::getCameras(p))
            for (float a : aspects) {
                glm::mat4 inv = glm::inverse(cam.getProjection(a) * cam.getView());
                json e; e["location"] = floats(&cam.location[0], 3); e["fov"] = (double)cam.fovDegY; e["aspect"] = (double)a;
                e["inverse"] = floats(glm::value_ptr(inv), 16);
                out["inverse"].push_back(e);
            }
    }
    out["field_model"] = json::array();
    {
        const float tris[][9] = {{0, 0, 0, 1, 0, 0, 0, 1, 0}, {0, 0, 0, 0, 1, 0, 1, 0, 0}, {0.5f, -1.25f, 2, 1.5f, 0.75f, -0.5f, -2, 0.25f, 1},
                                 {1, 2, 3, 1.001f, 2, 3, 1, 2.5f, 3.75f}, {-3, 0.1f, 0.2f, -3, 1.1f, 0.2f, -3, 0.1f, 1.2f}};
        for (auto& t : tris) {
            glm::vec3 v0(t[0], t[1], t[2]), v1(t[3], t[4], t[5]), v2(t[6], t[7], t[8]);
            glm::vec3 location = (v0 + v1 + v2) / 3.0f;
            glm::vec3 scale(glm::length(v1 - v0), glm::length(v2 - v0), 0.005f);
            scale *= 0.2f;
            glm::vec3 splatUp = glm::vec3(0.0f, 0.0f, 1.0f);
            glm::vec3 triNormal = glm::normalize(glm::cross(v1 - v0, v2 - v0));
            glm::vec3 axis = glm::cross(splatUp, triNormal);
            float angle = glm::acos(glm::dot(splatUp, triNormal));
            glm::quat q = glm::angleAxis(angle, axis);
            float rot[4]; memcpy(rot, &q, 16);                       // memory order as ModelSplatsHost::pushBack copies it
            json e; e["tri"] = floats(t, 9); e["location"] = floats(&location[0], 3); e["scale"] = floats(&scale[0], 3);
            e["rotation"] = floats(rot, 4);
            out["field_model"].push_back(e);
        }
        glm::quat q0 = glm::angleAxis(0.0f, glm::vec3(0.0f, 1.0f, 0.0f));    // initFieldGrid / initFieldMono rotation
        float rot[4]; memcpy(rot, &q0, 16);
        out["field_identity_rotation"] = floats(rot, 4);
    }
    FILE* f = fopen(argv[1], "w");
    std::string s = out.dump(1);
    fwrite(s.data(), 1, s.size(), f);
    fclose(f);
    return 0;
}
