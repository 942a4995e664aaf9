// C++ caller written the way the reference application uses its own classes (src/ui/UiFrame.cpp:137-160,
// 266-298; src/Trainer.cu:148-216), compiled against include/compat/ and linked with libgsb200.so.
// It proves the drop-in boundary: `Trainer`, `CudaRasterizer::Rasterizer` and the model containers
// are used through the reference's signatures only.  Plain host C++ (g++), no nvcc.
//
// Two builds (tests/compat/Makefile):
//   GSB_DEMO_REAL_APP  the application headers are the REFERENCE'S OWN src/Camera.h, src/Camera.cpp, src/Project.h,
//                      src/Config.h, src/ModelSplatsDevice.h, src/ModelSplatsHost.h, read in place from the reference
//                      tree (with its in-tree glm and nlohmann); only the bodies of ModelSplats*.cpp are restated
//                      (app_impl/ModelSplatsImpl.cpp) because the originals include ui/UiPanelViewInput.h (wxWidgets)
//   otherwise          the hand-written look-alikes in tests/compat/app/ (where the reference tree is absent)
//
//   compat_demo               single GPU: training must reduce the image error, densify must grow the model
//   compat_demo --devices G   the same job twice, once on one GPU and once with GSB_COMPAT_DEVICES=G (camera-sharded,
//                             native NCCL all-reduce inside libgsb200.so): parameters after k iterations incl. a densify
//                             must agree ("G in {1,2,4,8} produce parameters equal to G = 1")
// Prints one JSON line; exit code 0 on success.
#include <cmath>
#include <cstdint>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <vector>

#include <cuda_runtime.h>

#include "Config.h"
#include "Camera.h"
#include "Project.h"
#include "ModelSplatsDevice.h"
#include "ModelSplatsHost.h"
#include "Trainer.cuh"
#include "rasterizer.h"

#ifndef GSB_DEMO_REAL_APP
int Camera::getCamerasCount(const Project& project) { return project.sphere1.count + project.sphere2.count; }
std::vector<Camera> Camera::getCameras(const Project& project) {
    std::vector<Camera> out;
    for (int s = 0; s < 2; s++) {
        const auto& sphere = s == 0 ? project.sphere1 : project.sphere2;
        for (int i = 0; i < sphere.count; i++) {
            const float a = 6.2831853f * ((float)i + 0.37f * (float)s) / (float)sphere.count;
            out.push_back(Camera({sphere.distance * std::sin(a), 2.0f + (float)s, sphere.distance * std::cos(a)}, {0, 0, 0}, sphere.fovDeg));
        }
    }
    return out;
}
#endif

static const char* g_stage = "start";   // reported when an exception escapes
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

// "truth capture": a teacher model rendered on black through Trainer::render and on white through
// CudaRasterizer::Rasterizer::forward called the way src/Trainer.cu:175-201 calls it
static int capture_truths(Trainer& trainer, const std::vector<Camera>& cams, int teacherSplats) {
    const int W = RENDER_RESOLUTION_X, H = RENDER_RESOLUTION_Y, N = W * H;
    ModelSplatsHost teacher(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    lcg_state = 12345u;
    fill(teacher, teacherSplats, 2.0f);
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
    cudaFree(devBg); cudaFree(devView); cudaFree(devProjView); cudaFree(devCamLoc); cudaFree(devColor);
    return rendered;
}

static void set_student(Trainer& trainer, int splats) {
    // what UiFrame::initField* does -- delete trainer->model, assign a new one
    ModelSplatsHost student(SPLATS_LIMIT, SPLATS_SH_DEGREE, SPLATS_SH_COEF);
    lcg_state = 777u;
    fill(student, splats, 2.0f);
    delete trainer.model;
    trainer.model = new ModelSplatsDevice(student);
}

static void set_rates(Project& project) {
    // the gradients are sums over pixels (src/Trainer.cu:33-44 has no 1/N): rates tuned at 256x256 scale with the pixel count
    const float k = (256.0f * 256.0f) / ((float)RENDER_RESOLUTION_X * (float)RENDER_RESOLUTION_Y);
    project.lrLocation = 0.02f * k; project.lrSh = 0.1f * k; project.lrScale = 0.002f * k; project.lrOpacity = 0.05f * k;
    project.lrRotation = 0.002f * k;
}

static int single_device_demo() {
    const int N = RENDER_RESOLUTION_X * RENDER_RESOLUTION_Y;
    Project project;
    project.sphere1.count = 4;
    project.sphere2.count = 0;
    const std::vector<Camera> cams = Camera::getCameras(project);

    g_stage = "construct";
    Trainer trainer;
    g_stage = "capture_truths";
    const int rendered = capture_truths(trainer, cams, 2000);
    g_stage = "train without truths";

    // train() with no truths must throw the reference's message: check on a second trainer
    bool threw = false;
    {
        Trainer empty;
        try { empty.train(project, false); } catch (const std::runtime_error&) { threw = true; }
    }

    g_stage = "student";
    set_student(trainer, 3000);
    std::vector<uint32_t> scratch(N);
    uint32_t* devFb;
    cudaMalloc(&devFb, N * sizeof(uint32_t));
    const double err0 = image_error(trainer, cams, scratch, devFb);
    set_rates(project);
    g_stage = "train x60";
    for (int it = 0; it < 60; it++) trainer.train(project, false);
    g_stage = "after training";
    const double err1 = image_error(trainer, cams, scratch, devFb);
    const int count0 = trainer.model->count;
    project.paramDensifyVariance = -1.0f;      // every surviving splat splits or clones
    trainer.train(project, true);
    const int count1 = trainer.model->count;
    ModelSplatsHost back(*trainer.model);      // device -> host copy constructor, as saveSplats does
    bool finite = true;
    for (int i = 0; i < back.count * 3; i++) finite = finite && std::isfinite(back.scales[i]);
    trainer.train(project, false);             // the iteration after a densify re-binds the new arrays

#ifdef GSB_DEMO_REAL_APP
    const char* app = "reference headers (src/Camera.h, Camera.cpp, Project.h, Config.h, ModelSplats*.h)";
#else
    const char* app = "stand-in headers (tests/compat/app)";
#endif
    printf("{\"rendered_last_truth\": %d, \"err0\": %.4f, \"err1\": %.4f, \"count0\": %d, \"count1\": %d, \"iterations\": %d, "
           "\"threw_without_truths\": %s, \"finite\": %s, \"resolution\": %d, \"app\": \"%s\"}\n",
           rendered, err0, err1, count0, count1, project.iterations, threw ? "true" : "false", finite ? "true" : "false",
           RENDER_RESOLUTION_X, app);
    const bool ok = rendered > 0 && err1 < 0.8 * err0 && count1 > count0 && threw && finite && project.iterations == 62;
    return ok ? 0 : 1;
}

struct RunResult {
    int count = 0;
    std::vector<float> loc, shs, scales, opac, rot;
};

// k iterations of the same job (6 + 6 cameras, densify on iteration 3 of 5) on `devices` GPUs
static int g_devices_max = 1;
static RunResult run_job(int devices) {
    if (devices > 1) setenv("GSB_COMPAT_DEVICES", std::to_string(devices).c_str(), 1);
    else unsetenv("GSB_COMPAT_DEVICES");
    Project project;
    project.sphere1.count = std::max(6, g_devices_max);      // every rank needs a camera: both runs use the larger set
    project.sphere2.count = std::max(6, g_devices_max);
    project.sphere2.fovDeg = 30.0f;
    const std::vector<Camera> cams = Camera::getCameras(project);
    Trainer trainer;
    capture_truths(trainer, cams, 2000);
    set_student(trainer, 3000);
    set_rates(project);
    // every surviving splat densifies (no near-tie on the variance threshold), large ones split, small ones clone, faint
    // ones are culled: all three branches of src/Trainer.cu:448-499 on the all-reduced means
    project.paramDensifyVariance = -1.0f;
    project.paramCullOpacity = 0.35f;
    project.paramSplitSize = 0.2f;
    for (int it = 0; it < 5; it++) trainer.train(project, it == 2);
    ModelSplatsHost back(*trainer.model);
    RunResult r;
    r.count = back.count;
    r.loc.assign(back.locations, back.locations + 3 * back.count);
    r.shs.assign(back.shs, back.shs + 3 * back.shCoeffs * back.count);
    r.scales.assign(back.scales, back.scales + 3 * back.count);
    r.opac.assign(back.opacities, back.opacities + back.count);
    r.rot.assign(back.rotations, back.rotations + 4 * back.count);
    return r;
}

static double rel_diff(const std::vector<float>& a, const std::vector<float>& b) {
    if (a.size() != b.size()) return 1e30;
    double num = 0.0, den = 0.0;
    for (size_t i = 0; i < a.size(); i++) {
        const double d = (double)a[i] - (double)b[i];
        num += d * d; den += (double)b[i] * (double)b[i];
    }
    return std::sqrt(num / (den > 0.0 ? den : 1.0));
}

static int multi_device_demo(int devices) {
    int have = 0;
    cudaGetDeviceCount(&have);
    if (have < devices) {
        printf("{\"skipped\": \"needs %d GPUs, %d visible\"}\n", devices, have);
        return 77;
    }
    g_devices_max = devices;
    const RunResult one = run_job(1);
    const RunResult many = run_job(devices);
    const double d[5] = {rel_diff(many.loc, one.loc), rel_diff(many.shs, one.shs), rel_diff(many.scales, one.scales),
                         rel_diff(many.opac, one.opac), rel_diff(many.rot, one.rot)};
    double worst = 0.0;
    for (double x : d) worst = std::max(worst, x);
    printf("{\"devices\": %d, \"count_1\": %d, \"count_G\": %d, \"rel_loc\": %.3e, \"rel_sh\": %.3e, \"rel_scale\": %.3e, "
           "\"rel_opacity\": %.3e, \"rel_rot\": %.3e}\n", devices, one.count, many.count, d[0], d[1], d[2], d[3], d[4]);
    return (one.count == many.count && one.count != 3000 && worst <= 1e-5) ? 0 : 1;
}

int main(int argc, char** argv) {
    int devices = 1;
    for (int i = 1; i + 1 < argc; i++)
        if (!strcmp(argv[i], "--devices")) devices = std::atoi(argv[i + 1]);
    try {
        return devices > 1 ? multi_device_demo(devices) : single_device_demo();
    } catch (const std::exception& e) {
        fprintf(stderr, "compat_demo: %s (stage: %s)\n", e.what(), g_stage);
        return 2;
    }
}
