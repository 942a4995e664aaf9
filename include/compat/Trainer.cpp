// compat/Trainer.cpp -- implementation of the reference's Trainer interface (src/Trainer.cuh) on the
// C ABI of libgsb200.so.  Compile this file in the application INSTEAD of src/Trainer.cu; it includes
// the application's own Camera.h / Project.h / ModelSplatsDevice.h / ModelSplatsHost.h / Config.h
// (resolved through the application's include path) and needs no nvcc: it is plain host C++.
//
// Behaviour kept from src/Trainer.cu: the temporary 0-splat model of the constructor (:112), the
// exception text when train() runs without truths (:253), iterations++ at the start of train (:255),
// white passes and black passes of every camera with S = 2 x cameras (:311-419), square-image
// tan_fov in train (:355-356) and the aspect-scaled angle in render (:196-197), `delete model; model =
// new ...` on densify (:540-541).  `model` may be replaced from outside between calls
// (src/ui/UiFrame.cpp:157-158): it is re-bound on every call.

#include "Trainer.cuh"

#include <cmath>
#include <stdexcept>

#include <cuda_runtime.h>

#include "Config.h"
#include "Camera.h"
#include "Project.h"
#include "ModelSplatsDevice.h"
#include "ModelSplatsHost.h"
#ifndef GSB_COMPAT_NO_RTX
#include "rtx/RtxHost.h"
#endif

#include "../gsb.h"

namespace {

void check(int code) {
    if (code != GSB_OK) throw std::runtime_error(gsb_last_error());
}

float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }   // glm::radians

// view = -lookAt, projview = perspective * view (src/Trainer.cu:321-326); glm::mat4 is 16 contiguous
// column-major floats.
template <typename CameraT>
gsb_camera camera_constants(const CameraT& camera, float aspect, float tan_fovx, float tan_fovy) {
    gsb_camera c;
    const auto view = camera.getView();
    const auto projview = camera.getProjection(aspect) * view;
    const float* v = &view[0][0];
    const float* pv = &projview[0][0];
    for (int i = 0; i < 16; i++) { c.view[i] = v[i]; c.projview[i] = pv[i]; }
    for (int i = 0; i < 3; i++) c.location[i] = camera.location[i];
    c.tan_fovx = tan_fovx;
    c.tan_fovy = tan_fovy;
    return c;
}

gsb_project_params project_params(const Project& p) {
    gsb_project_params q;
    q.lrLocation = p.lrLocation; q.lrSh = p.lrSh; q.lrScale = p.lrScale; q.lrOpacity = p.lrOpacity; q.lrRotation = p.lrRotation;
    q.paramScaleMax = p.paramScaleMax;
    q.paramCullOpacity = p.paramCullOpacity; q.paramCullSize = p.paramCullSize; q.paramDensifyVariance = p.paramDensifyVariance;
    q.paramSplitSize = p.paramSplitSize; q.paramSplitDistance = p.paramSplitDistance; q.paramSplitScale = p.paramSplitScale;
    q.paramCloneDistance = p.paramCloneDistance;
    return q;
}

void bind(gsb_trainer* t, ModelSplatsDevice* m) {
    check(gsb_trainer_bind_model(t, m->count, m->capacity, m->shDegree, m->shCoeffs, m->devLocations, m->devShs, m->devScales,
                                 m->devOpacities, m->devRotations));
}

}  // namespace

Trainer::Trainer() {
    check(gsb_trainer_create(&impl, RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y, nullptr));
    // Temporary model so callers never see a null pointer; replaced on project load (src/Trainer.cu:111-112).
    model = new ModelSplatsDevice(ModelSplatsHost(0, 0, 1));
}

Trainer::~Trainer() {
    gsb_trainer_destroy(impl);
    delete model;
    for (uint32_t* frameBuffer : truthFrameBuffersW) cudaFree(frameBuffer);
    for (uint32_t* frameBuffer : truthFrameBuffersB) cudaFree(frameBuffer);
}

void Trainer::render(uint32_t* frameBuffer, int sizeX, int sizeY, float splatScale, const Camera& camera) {
    bind(impl, model);
    const gsb_camera c = camera_constants(camera, (float)sizeX / (float)sizeY,
                                          std::tan(radians((float)sizeX * camera.fovDegY / (float)sizeY) * 0.5f),
                                          std::tan(radians(camera.fovDegY) * 0.5f));
    check(gsb_trainer_render(impl, frameBuffer, sizeX, sizeY, splatScale, &c));   // ends synchronised, like :215
}

#ifndef GSB_COMPAT_NO_RTX
void Trainer::captureTruths(const Project& project, RtxHost& rtx) {
    const std::vector<Camera> cameras = Camera::getCameras(project);
    if (lastTruthCount != (int)cameras.size()) {
        lastTruthCount = (int)cameras.size();
        for (uint32_t* frameBuffer : truthFrameBuffersW) cudaFree(frameBuffer);
        for (uint32_t* frameBuffer : truthFrameBuffersB) cudaFree(frameBuffer);
        truthFrameBuffersW.clear();
        truthFrameBuffersB.clear();
        for (size_t i = 0; i < cameras.size(); i++) {
            uint32_t *w = nullptr, *b = nullptr;
            cudaMalloc(&w, (size_t)RENDER_RESOLUTION_X * RENDER_RESOLUTION_Y * sizeof(uint32_t));
            cudaMalloc(&b, (size_t)RENDER_RESOLUTION_X * RENDER_RESOLUTION_Y * sizeof(uint32_t));
            truthFrameBuffersW.push_back(w);
            truthFrameBuffersB.push_back(b);
        }
    }
    truthCameras = cameras;
    for (size_t i = 0; i < cameras.size(); i++) {
        rtx.render(truthFrameBuffersW[i], {RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y}, cameras[i], {1.0f, 1.0f, 1.0f}, project.rtSamples, {});
        rtx.render(truthFrameBuffersB[i], {RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y}, cameras[i], {0.0f, 0.0f, 0.0f}, project.rtSamples, {});
    }
}
#endif

void Trainer::train(Project& project, bool densify) {
    if (truthFrameBuffersW.empty()) throw std::runtime_error("Can't run training iteration, no truth data available!");

    project.iterations++;

    bind(impl, model);
    std::vector<gsb_camera> cams;
    cams.reserve(truthCameras.size());
    for (const Camera& camera : truthCameras) {
        const float t = std::tan(radians(camera.fovDegY) * 0.5f);    // square training image: tan_fovx = tan_fovy
        cams.push_back(camera_constants(camera, (float)RENDER_RESOLUTION_X / (float)RENDER_RESOLUTION_Y, t, t));
    }
    const int n = (int)truthCameras.size();
    check(gsb_trainer_set_truths(impl, n, truthFrameBuffersW.data(), truthFrameBuffersB.data(), cams.data(), n));

    const gsb_project_params params = project_params(project);
    int newCount = 0;
    check(gsb_trainer_train(impl, &params, densify ? 1 : 0, &newCount));
    if (!densify) return;

    // The library planned the cull / split / clone on the device; hand it freshly allocated arrays of
    // the new size and swap the model, as the reference does after its CPU pass (src/Trainer.cu:540-541).
    ModelSplatsDevice* next = new ModelSplatsDevice(ModelSplatsHost(0, model->shDegree, model->shCoeffs));
    cudaFree(next->devLocations); cudaFree(next->devShs); cudaFree(next->devScales);
    cudaFree(next->devOpacities); cudaFree(next->devRotations);
    next->capacity = model->capacity;
    next->count = newCount;
    const size_t c = (size_t)newCount;
    cudaMalloc(&next->devLocations, c * 3 * sizeof(float));
    cudaMalloc(&next->devShs, c * 3 * next->shCoeffs * sizeof(float));
    cudaMalloc(&next->devScales, c * 3 * sizeof(float));
    cudaMalloc(&next->devOpacities, c * sizeof(float));
    cudaMalloc(&next->devRotations, c * 4 * sizeof(float));
    check(gsb_trainer_densify_commit(impl, next->devLocations, next->devShs, next->devScales, next->devOpacities,
                                     next->devRotations));
    delete model;
    model = next;
}
