// C++ caller written the way the reference application uses its own classes (src/ui/UiFrame.cpp:137-160,
// 266-298; src/Trainer.cu:148-216), compiled against include/compat/ and linked with libgsb200.so.
// It proves the drop-in boundary: `Trainer`, `CudaRasterizer::Rasterizer` and the model containers
// are used through the reference's signatures only.  Plain host C++ (g++), no nvcc.
//
// Prints one JSON line; exit code 0 only if training reduced the image error and densify grew the model.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>

#include <cuda_runtime.h>

#include "Config.h"
#include "Camera.h"
#include "Project.h"
#include "ModelSplatsDevice.h"
#include "ModelSplatsHost.h"
#include "Trainer.cuh"
#include "rasterizer.h"

int Camera::getCamerasCount(const Project& project) { return project.sphere1.count + project.sphere2.count; }
std::vector<Camera> Camera::getCameras(const Project& project) {
    std::vector<Camera> out;
    for (int i = 0; i < project.sphere1.count; i++) {
        const float a = 6.2831853f * (float)i / (float)project.sphere1.count;
        out.push_back(Camera({project.sphere1.distance * std::sin(a), 2.0f, project.sphere1.distance * std::cos(a)}, {0, 0, 0},
                             project.sphere1.fovDeg));
    }
    return out;
}

static uint32_t lcg_state = 12345u;
static float rnd() { lcg_state = lcg_state * 1664525u + 1013904223u; return (float)(lcg_state >> 8) * (1.0f / 16777216.0f); }

static void fill(ModelSplatsHost& h, int n, float spread) {
    for (int i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++) h.locations[3 * i + k] = (rnd() * 2.0f - 1.0f) * spread;
        for (int k = 0; k < 3; k++) h.scales[3 * i + k] = 0.05f + 0.15f * rnd();
        h.opacities[i] = 0.3f + 0.7f * rnd();
        float q[4], n2 = 0.f;
        for (int k = 0; k < 4; k++) { q[k] = rnd() * 2.0f - 1.0f; n2 += q[k] * q[k]; }
        for (int k = 0; k < 4; k++) h.rotations[4 * i + k] = q[k] / std::sqrt(n2);
        for (int k = 0; k < 12; k++) h.shs[12 * i + k] = k < 3 ? (rnd() - 0.5f) / 0.2820948f : 0.1f * (rnd() - 0.5f);
    }
    h.count = n;
}

static double image_error(Trainer& trainer, const std::vector<Camera>& cams, std::vector<uint32_t>& scratch, uint32_t* devFb) {
    const int N = RENDER_RESOLUTION_X * RENDER_RESOLUTION_Y;
    std::vector<uint32_t> truth(N);
    double err = 0.0;
    for (size_t i = 0; i < cams.size(); i++) {
        trainer.render(devFb, RENDER_RESOLUTION_X, RENDER_RESOLUTION_Y, 1.0f, cams[i]);
        cudaMemcpy(scratch.data(), devFb, N * sizeof(uint32_t), cudaMemcpyDeviceToHost);
        cudaMemcpy(truth.data(), trainer.truthFrameBuffersB[i], N * sizeof(uint32_t), cudaMemcpyDeviceToHost);
        for (int p = 0; p < N; p++)
            for (int c = 0; c < 3; c++) {
                const double d = (double)((scratch[p] >> (8 * c)) & 0xFF) - (double)((truth[p] >> (8 * c)) & 0xFF);
                err += d * d;
            }
    }
    return err / ((double)cams.size() * N * 3);
}

int main() {
    const int W = RENDER_RESOLUTION_X, H = RENDER_RESOLUTION_Y, N = W * H;
    Project project;
    project.sphere1.count = 4;
    project.sphere2.count = 0;
    const std::vector<Camera> cams = Camera::getCameras(project);

    Trainer trainer;

    // "truth capture": a teacher model rendered on black through Trainer::render and on white through
    // CudaRasterizer::Rasterizer::forward called the way src/Trainer.cu:175-201 calls it
    ModelSplatsHost teacher(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    fill(teacher, 2000, 2.0f);
    delete trainer.model;
    trainer.model = new ModelSplatsDevice(teacher);
    float *devBg, *devView, *devProjView, *devCamLoc, *devColor;
    cudaMalloc(&devBg, 12); cudaMalloc(&devView, 64); cudaMalloc(&devProjView, 64); cudaMalloc(&devCamLoc, 12);
    cudaMalloc(&devColor, (size_t)N * 3 * sizeof(float));
    const float white[3] = {1.f, 1.f, 1.f};
    cudaMemcpy(devBg, white, 12, cudaMemcpyHostToDevice);
    int rendered = 0;
    for (const Camera& camera : cams) {
        uint32_t *fbW, *fbB;
        cudaMalloc(&fbW, N * sizeof(uint32_t)); cudaMalloc(&fbB, N * sizeof(uint32_t));
        trainer.render(fbB, W, H, 1.0f, camera);
        const glm::mat4 view = camera.getView();
        const glm::mat4 projView = camera.getProjection((float)W / (float)H) * view;
        cudaMemcpy(devView, &view[0][0], 64, cudaMemcpyHostToDevice);
        cudaMemcpy(devProjView, &projView[0][0], 64, cudaMemcpyHostToDevice);
        cudaMemcpy(devCamLoc, &camera.location[0], 12, cudaMemcpyHostToDevice);
        char *geomBuffer = nullptr, *binningBuffer = nullptr, *imgBuffer = nullptr;
        const float t = std::tan(glm::radians(camera.fovDegY) * 0.5f);
        rendered = CudaRasterizer::Rasterizer::forward(
            [&](size_t n) { cudaMalloc(&geomBuffer, n); return geomBuffer; },
            [&](size_t n) { cudaMalloc(&binningBuffer, n); return binningBuffer; },
            [&](size_t n) { cudaMalloc(&imgBuffer, n); return imgBuffer; },
            trainer.model->count, trainer.model->shDegree, trainer.model->shCoeffs, devBg, W, H, trainer.model->devLocations,
            trainer.model->devShs, nullptr, trainer.model->devOpacities, trainer.model->devScales, 1.0f,
            trainer.model->devRotations, nullptr, devView, devProjView, devCamLoc, t, t, false, devColor, nullptr, true);
        gsb_image_float_to_int(devColor, fbW, W, H, nullptr);
        cudaDeviceSynchronize();
        cudaFree(geomBuffer); cudaFree(binningBuffer); cudaFree(imgBuffer);
        trainer.truthFrameBuffersW.push_back(fbW);
        trainer.truthFrameBuffersB.push_back(fbB);
        trainer.truthCameras.push_back(camera);
    }

    // train() with no truths must throw the reference's message: check on a second trainer
    bool threw = false;
    {
        Trainer empty;
        try { empty.train(project, false); } catch (const std::runtime_error&) { threw = true; }
    }

    // the student: what UiFrame::initField* does -- delete trainer->model, assign a new one
    ModelSplatsHost student(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    fill(student, 3000, 2.0f);
    delete trainer.model;
    trainer.model = new ModelSplatsDevice(student);

    std::vector<uint32_t> scratch(N);
    uint32_t* devFb;
    cudaMalloc(&devFb, N * sizeof(uint32_t));
    const double err0 = image_error(trainer, cams, scratch, devFb);
    project.lrLocation = 0.02f; project.lrSh = 0.1f; project.lrScale = 0.002f; project.lrOpacity = 0.05f; project.lrRotation = 0.002f;
    for (int it = 0; it < 60; it++) trainer.train(project, false);
    const double err1 = image_error(trainer, cams, scratch, devFb);
    const int count0 = trainer.model->count;
    project.paramDensifyVariance = -1.0f;      // every surviving splat splits or clones
    trainer.train(project, true);
    const int count1 = trainer.model->count;
    ModelSplatsHost back(*trainer.model);      // device -> host copy constructor, as saveSplats does
    bool finite = true;
    for (int i = 0; i < back.count * 3; i++) finite = finite && std::isfinite(back.scales[i]);
    trainer.train(project, false);             // the iteration after a densify re-binds the new arrays

    printf("{\"rendered_last_truth\": %d, \"err0\": %.4f, \"err1\": %.4f, \"count0\": %d, \"count1\": %d, \"iterations\": %d, "
           "\"threw_without_truths\": %s, \"finite\": %s}\n",
           rendered, err0, err1, count0, count1, project.iterations, threw ? "true" : "false", finite ? "true" : "false");
    const bool ok = rendered > 0 && err1 < 0.8 * err0 && count1 > count0 && threw && finite && project.iterations == 62;
    return ok ? 0 : 1;
}
