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

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>

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

void cuda_check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

}  // namespace

// ---- GSB_COMPAT_DEVICES: one process, G devices ---------------------------------------------------------------
// Rank 0 is the application's own device and `model`.  Rank r >= 1 owns a replica of the five parameter arrays on
// device r (re-broadcast whenever the application replaces `model`; afterwards every replica applies the same
// all-reduced means with the same deterministic kernels, so they stay bit-identical), reads its cameras' truth
// frames straight out of device 0's memory over NVLink (peer access; copied per iteration where peer access is
// unavailable) and is driven by its own host thread for the duration of a train() call.
struct Trainer::Multi {
    struct Replica {
        int device = 0;
        gsb_trainer* impl = nullptr;
        gsb_comm* comm = nullptr;
        float *loc = nullptr, *shs = nullptr, *scales = nullptr, *opac = nullptr, *rot = nullptr;
        int count = 0;
        bool peer = false;
        std::vector<uint32_t*> ownW, ownB;     // frame copies when peer access is unavailable
        std::string error;
    };
    int home = 0;                              // the application's device
    std::vector<Replica> ranks;                // ranks[0] mirrors `impl` / `model`
    std::vector<gsb_comm*> comms;
    const void* syncedModel = nullptr;         // identity of the model the replicas were broadcast from
    float* syncedLoc = nullptr;
    int syncedCount = -1;

    static void freeArrays(Replica& r) {
        cudaFree(r.loc); cudaFree(r.shs); cudaFree(r.scales); cudaFree(r.opac); cudaFree(r.rot);
        r.loc = r.shs = r.scales = r.opac = r.rot = nullptr;
    }
    static void allocArrays(Replica& r, int count, int shCoeffs) {
        const size_t c = (size_t)(count > 0 ? count : 1);
        cuda_check(cudaMalloc(&r.loc, c * 3 * sizeof(float)), "replica alloc");
        cuda_check(cudaMalloc(&r.shs, c * 3 * shCoeffs * sizeof(float)), "replica alloc");
        cuda_check(cudaMalloc(&r.scales, c * 3 * sizeof(float)), "replica alloc");
        cuda_check(cudaMalloc(&r.opac, c * sizeof(float)), "replica alloc");
        cuda_check(cudaMalloc(&r.rot, c * 4 * sizeof(float)), "replica alloc");
        r.count = count;
    }
};

Trainer::Trainer() {
    check(gsb_trainer_create(&impl, RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y, nullptr));
    // Temporary model so callers never see a null pointer; replaced on project load (src/Trainer.cu:111-112).
    model = new ModelSplatsDevice(ModelSplatsHost(0, 0, 1));

    const char* env = std::getenv("GSB_COMPAT_DEVICES");
    const int want = env ? std::atoi(env) : 1;
    if (want > 1) {
        int have = 0, home = 0;
        cuda_check(cudaGetDeviceCount(&have), "cudaGetDeviceCount");
        cuda_check(cudaGetDevice(&home), "cudaGetDevice");
        if (want > have) throw std::runtime_error("GSB_COMPAT_DEVICES asks for more GPUs than are visible");
        multi = new Multi();
        multi->home = home;
        std::vector<int> devs;
        devs.push_back(home);
        for (int d = 0; d < have && (int)devs.size() < want; d++) if (d != home) devs.push_back(d);
        multi->comms.resize((size_t)want);
        check(gsb_comm_init_all(multi->comms.data(), want, devs.data()));
        multi->ranks.resize((size_t)want);
        for (int r = 0; r < want; r++) {
            Multi::Replica& R = multi->ranks[(size_t)r];
            R.device = devs[(size_t)r];
            R.comm = multi->comms[(size_t)r];
            if (r == 0) { R.impl = impl; R.peer = true; continue; }
            cuda_check(cudaSetDevice(R.device), "cudaSetDevice");
            int can = 0;
            cudaDeviceCanAccessPeer(&can, R.device, home);
            if (can) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(home, 0);
                R.peer = (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled);
                cudaGetLastError();
            }
            check(gsb_trainer_create(&R.impl, RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y, nullptr));
            check(gsb_trainer_set_comm(R.impl, R.comm));
        }
        cuda_check(cudaSetDevice(home), "cudaSetDevice");
        check(gsb_trainer_set_comm(impl, multi->comms[0]));
    }
}

Trainer::~Trainer() {
    if (multi) {
        for (size_t r = 1; r < multi->ranks.size(); r++) {
            Multi::Replica& R = multi->ranks[r];
            cudaSetDevice(R.device);
            gsb_trainer_destroy(R.impl);
            Multi::freeArrays(R);
            for (uint32_t* f : R.ownW) cudaFree(f);
            for (uint32_t* f : R.ownB) cudaFree(f);
        }
        cudaSetDevice(multi->home);
        gsb_trainer_set_comm(impl, nullptr);
        for (gsb_comm* c : multi->comms) gsb_comm_destroy(c);
        delete multi;
    }
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

    if (multi) { trainMulti(project, densify); return; }

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

void Trainer::trainMulti(Project& project, bool densify) {
    Multi& m = *multi;
    const int G = (int)m.ranks.size();
    const int n = (int)truthCameras.size();
    const int n1 = std::min(n, std::max(0, (int)project.sphere1.count));
    const int M = model->shCoeffs;

    // replicas follow `model`: broadcast when the application replaced it (src/ui/UiFrame.cpp:157-158)
    if (m.syncedModel != model || m.syncedLoc != model->devLocations || m.syncedCount != model->count) {
        cuda_check(cudaDeviceSynchronize(), "sync before replica broadcast");
        for (int r = 1; r < G; r++) {
            Multi::Replica& R = m.ranks[(size_t)r];
            cuda_check(cudaSetDevice(R.device), "cudaSetDevice");
            Multi::freeArrays(R);
            Multi::allocArrays(R, model->count, M);
            const size_t c = (size_t)model->count;
            cuda_check(cudaMemcpyPeer(R.loc, R.device, model->devLocations, m.home, c * 3 * sizeof(float)), "peer copy");
            cuda_check(cudaMemcpyPeer(R.shs, R.device, model->devShs, m.home, c * 3 * M * sizeof(float)), "peer copy");
            cuda_check(cudaMemcpyPeer(R.scales, R.device, model->devScales, m.home, c * 3 * sizeof(float)), "peer copy");
            cuda_check(cudaMemcpyPeer(R.opac, R.device, model->devOpacities, m.home, c * sizeof(float)), "peer copy");
            cuda_check(cudaMemcpyPeer(R.rot, R.device, model->devRotations, m.home, c * 4 * sizeof(float)), "peer copy");
        }
        cuda_check(cudaSetDevice(m.home), "cudaSetDevice");
    }

    for (int r = 0; r < G; r++)   // every rank joins the all-reduce, so every rank needs at least one camera
        if ((n1 > r ? (n1 - r + G - 1) / G : 0) + (n - n1 > r ? (n - n1 - r + G - 1) / G : 0) == 0)
            throw std::runtime_error("GSB_COMPAT_DEVICES exceeds the number of truth cameras");

    const gsb_project_params params = project_params(project);
    std::vector<int> newCount((size_t)G, 0);
    std::vector<float*> next((size_t)G * 5, nullptr);

    auto rankBody = [&](int r) {
        Multi::Replica& R = m.ranks[(size_t)r];
        try {
            cuda_check(cudaSetDevice(R.device), "cudaSetDevice");
            // camera partition of SURVEY.md 8(e): rank r owns cameras r, r+G, ... of each sphere
            std::vector<gsb_camera> cams;
            std::vector<const uint32_t*> fw, fb;
            std::vector<int> mine;
            for (int i = r; i < n1; i += G) mine.push_back(i);
            for (int i = n1 + r; i < n; i += G) mine.push_back(i);
            for (size_t k = 0; k < mine.size(); k++) {
                const Camera& camera = truthCameras[(size_t)mine[k]];
                const float t = std::tan(radians(camera.fovDegY) * 0.5f);
                cams.push_back(camera_constants(camera, (float)RENDER_RESOLUTION_X / (float)RENDER_RESOLUTION_Y, t, t));
                const uint32_t* w = truthFrameBuffersW[(size_t)mine[k]];
                const uint32_t* b = truthFrameBuffersB[(size_t)mine[k]];
                if (!R.peer) {
                    const size_t bytes = (size_t)RENDER_RESOLUTION_X * RENDER_RESOLUTION_Y * sizeof(uint32_t);
                    while (R.ownW.size() <= k) {
                        uint32_t *a = nullptr, *c = nullptr;
                        cuda_check(cudaMalloc(&a, bytes), "frame copy alloc");
                        cuda_check(cudaMalloc(&c, bytes), "frame copy alloc");
                        R.ownW.push_back(a); R.ownB.push_back(c);
                    }
                    cuda_check(cudaMemcpyPeer(R.ownW[k], R.device, w, m.home, bytes), "frame copy");
                    cuda_check(cudaMemcpyPeer(R.ownB[k], R.device, b, m.home, bytes), "frame copy");
                    w = R.ownW[k]; b = R.ownB[k];
                }
                fw.push_back(w); fb.push_back(b);
            }
            if (r == 0) bind(R.impl, model);
            else check(gsb_trainer_bind_model(R.impl, R.count, model->capacity, model->shDegree, M, R.loc, R.shs, R.scales, R.opac, R.rot));
            check(gsb_trainer_set_truths(R.impl, (int)cams.size(), fw.data(), fb.data(), cams.data(), n));
            check(gsb_trainer_train(R.impl, &params, densify ? 1 : 0, &newCount[(size_t)r]));
            if (densify) {
                const size_t c = (size_t)(newCount[(size_t)r] > 0 ? newCount[(size_t)r] : 1);
                float** nx = &next[(size_t)r * 5];
                cuda_check(cudaMalloc(&nx[0], c * 3 * sizeof(float)), "densify alloc");
                cuda_check(cudaMalloc(&nx[1], c * 3 * M * sizeof(float)), "densify alloc");
                cuda_check(cudaMalloc(&nx[2], c * 3 * sizeof(float)), "densify alloc");
                cuda_check(cudaMalloc(&nx[3], c * sizeof(float)), "densify alloc");
                cuda_check(cudaMalloc(&nx[4], c * 4 * sizeof(float)), "densify alloc");
                check(gsb_trainer_densify_commit(R.impl, nx[0], nx[1], nx[2], nx[3], nx[4]));
            }
        } catch (const std::exception& e) {
            R.error = e.what();
        }
    };

    std::vector<std::thread> workers;
    for (int r = 1; r < G; r++) workers.emplace_back(rankBody, r);
    rankBody(0);
    for (std::thread& t : workers) t.join();
    cuda_check(cudaSetDevice(m.home), "cudaSetDevice");
    for (int r = 0; r < G; r++)
        if (!m.ranks[(size_t)r].error.empty()) {
            const std::string e = m.ranks[(size_t)r].error;
            m.ranks[(size_t)r].error.clear();
            throw std::runtime_error("rank " + std::to_string(r) + ": " + e);
        }

    if (densify) {
        for (int r = 1; r < G; r++)
            if (newCount[(size_t)r] != newCount[0]) throw std::runtime_error("replicas diverged: densify counts differ across devices");
        ModelSplatsDevice* nextModel = new ModelSplatsDevice(ModelSplatsHost(0, model->shDegree, model->shCoeffs));
        cudaFree(nextModel->devLocations); cudaFree(nextModel->devShs); cudaFree(nextModel->devScales);
        cudaFree(nextModel->devOpacities); cudaFree(nextModel->devRotations);
        nextModel->capacity = model->capacity;
        nextModel->count = newCount[0];
        nextModel->devLocations = next[0]; nextModel->devShs = next[1]; nextModel->devScales = next[2];
        nextModel->devOpacities = next[3]; nextModel->devRotations = next[4];
        delete model;
        model = nextModel;
        for (int r = 1; r < G; r++) {
            Multi::Replica& R = m.ranks[(size_t)r];
            cuda_check(cudaSetDevice(R.device), "cudaSetDevice");
            Multi::freeArrays(R);
            float** nx = &next[(size_t)r * 5];
            R.loc = nx[0]; R.shs = nx[1]; R.scales = nx[2]; R.opac = nx[3]; R.rot = nx[4];
            R.count = newCount[(size_t)r];
        }
        cuda_check(cudaSetDevice(m.home), "cudaSetDevice");
    }
    m.syncedModel = model; m.syncedLoc = model->devLocations; m.syncedCount = model->count;
}
