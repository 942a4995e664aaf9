// The headless UiFrame: Project + Trainer (+ its ModelSplatsDevice) + RtxHost and the idle-loop scheduler
// (reference src/ui/UiFrame.cpp:126-298,323-450, src/ui/tools/UiPanelToolsTruth.cpp:185-197,
// src/ui/tools/UiPanelToolsView.cpp:227-259, src/Trainer.cu:218-250).  Host C++ over the C ABI of gsb.h / gsb_app.h:
// no window system, no GL; frames live in device buffers and leave as PNG files.
#include "app_common.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include <cuda_runtime.h>

namespace {

using gsbapp::fail;

#define APP_CUDA_OK(expr)                                                                       \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) return fail(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)
#define APP_OK(expr)                  \
    do {                              \
        const int rc_ = (expr);       \
        if (rc_ != GSB_OK) return rc_; \
    } while (0)

const float AUTO_TRAIN_BUDGET = 100.0f;           // src/Config.h:10

// ModelSplatsDevice (src/ModelSplatsDevice.h:8-28): five device arrays sized for `count`
struct ModelDevice {
    int capacity = 0, shDegree = 0, shCoeffs = 1, count = 0;
    float* dev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};      // locations, shs, scales, opacities, rotations
    size_t width(int k) const { return k == 0 ? 3 : k == 1 ? (size_t)3 * shCoeffs : k == 2 ? 3 : k == 3 ? 1 : 4; }
    void release() { for (float*& p : dev) { cudaFree(p); p = nullptr; } count = 0; }
    int allocate(int n) {
        for (int k = 0; k < 5; k++) APP_CUDA_OK(cudaMalloc(&dev[k], std::max<size_t>((size_t)n * width(k), 1) * sizeof(float)));
        count = n;
        return GSB_OK;
    }
};

}  // namespace

struct gsb_app {
    int W = 0, H = 0;
    gsb_project* project = nullptr;
    gsb_rtx* rtx = nullptr;
    gsb_trainer* trainer = nullptr;
    ModelDevice model;
    std::vector<uint32_t*> truthW, truthB;         // Trainer::truthFrameBuffersW / B
    std::vector<gsb_camera_desc> truthCameras;     // Trainer::truthCameras
    int lastTruthCount = -1;
    bool autoTraining = false;
    float autoTrainingBudget = 0.0f;
};

namespace {

int upload_model(gsb_app* a, const gsbapp::ModelHost& h) {         // ModelSplatsDevice(const ModelSplatsHost&)
    a->model.release();
    a->model.capacity = h.capacity; a->model.shDegree = h.shDegree; a->model.shCoeffs = h.shCoeffs;
    APP_OK(a->model.allocate(h.count));
    const std::vector<float>* src[5] = {&h.locations, &h.shs, &h.scales, &h.opacities, &h.rotations};
    for (int k = 0; k < 5; k++) {
        const size_t n = (size_t)h.count * a->model.width(k);
        if (n) APP_CUDA_OK(cudaMemcpy(a->model.dev[k], src[k]->data(), n * sizeof(float), cudaMemcpyHostToDevice));
    }
    return GSB_OK;
}

int bind(gsb_app* a) {
    ModelDevice& m = a->model;
    return gsb_trainer_bind_model(a->trainer, m.count, m.capacity, m.shDegree, m.shCoeffs, m.dev[0], m.dev[1], m.dev[2], m.dev[3], m.dev[4]);
}

void free_truths(gsb_app* a) {
    for (uint32_t* p : a->truthW) cudaFree(p);
    for (uint32_t* p : a->truthB) cudaFree(p);
    a->truthW.clear(); a->truthB.clear(); a->truthCameras.clear();
}

int alloc_truths(gsb_app* a, int n) {
    free_truths(a);
    for (int i = 0; i < n; i++) {
        uint32_t *w = nullptr, *b = nullptr;
        APP_CUDA_OK(cudaMalloc(&w, (size_t)a->W * a->H * sizeof(uint32_t)));
        a->truthW.push_back(w);
        APP_CUDA_OK(cudaMalloc(&b, (size_t)a->W * a->H * sizeof(uint32_t)));
        a->truthB.push_back(b);
    }
    a->truthCameras.resize((size_t)n);
    return GSB_OK;
}

int set_field(gsb_app* a, gsb_model_host* h) {
    const int rc = upload_model(a, h->m);
    gsb_model_host_destroy(h);
    if (rc == GSB_OK) gsb_project_fields(a->project)->iterations = 0;
    return rc;
}

}  // namespace

extern "C" {

int gsb_app_create(gsb_app** out, int train_width, int train_height) {
    if (!out || train_width <= 0 || train_height <= 0) return fail("gsb_app_create: invalid argument");
    gsb_app* a = new gsb_app();
    a->W = train_width; a->H = train_height;
    int rc = gsb_project_create(&a->project);
    if (rc == GSB_OK) rc = gsb_rtx_create(&a->rtx);
    if (rc == GSB_OK) rc = gsb_trainer_create(&a->trainer, train_width, train_height, nullptr);
    if (rc == GSB_OK) rc = gsb_app_new_project(a);
    if (rc != GSB_OK) { gsb_app_destroy(a); return rc; }
    *out = a;
    return GSB_OK;
}

int gsb_app_destroy(gsb_app* a) {
    if (!a) return GSB_OK;
    gsb_trainer_destroy(a->trainer);
    gsb_rtx_destroy(a->rtx);
    gsb_project_destroy(a->project);
    a->model.release();
    free_truths(a);
    delete a;
    return GSB_OK;
}

gsb_project* gsb_app_project(gsb_app* a) { return a ? a->project : nullptr; }
gsb_rtx* gsb_app_rtx(gsb_app* a) { return a ? a->rtx : nullptr; }
gsb_trainer* gsb_app_trainer(gsb_app* a) { return a ? a->trainer : nullptr; }

int gsb_app_new_project(gsb_app* a) {
    if (!a) return fail("gsb_app_new_project: invalid argument");
    APP_OK(gsb_project_init_new(a->project));
    APP_OK(gsb_app_init_field(a, 0));
    return gsb_rtx_reset(a->rtx);                  // refreshProject, UF:300-301
}

int gsb_app_init_field(gsb_app* a, int kind) {
    if (!a) return fail("gsb_app_init_field: invalid argument");
    gsb_model_host* h = nullptr;
    if (kind == 0) APP_OK(gsb_field_grid(&h));
    else if (kind == 1) APP_OK(gsb_field_mono(&h));
    else if (kind == 2) APP_OK(gsb_field_model(&h, gsb_project_get_string(a->project, "pathModel")));
    else return fail("gsb_app_init_field: kind must be 0 (grid), 1 (mono) or 2 (model)");
    return set_field(a, h);
}

int gsb_app_set_model(gsb_app* a, gsb_model_host* m) {
    if (!a || !m) return fail("gsb_app_set_model: invalid argument");
    return upload_model(a, m->m);
}

int gsb_app_get_model(gsb_app* a, gsb_model_host** out) {
    if (!a || !out) return fail("gsb_app_get_model: invalid argument");
    const ModelDevice& m = a->model;
    gsb_model_host* h = new gsb_model_host(m.capacity, m.shDegree, m.shCoeffs);
    std::vector<float>* dst[5] = {&h->m.locations, &h->m.shs, &h->m.scales, &h->m.opacities, &h->m.rotations};
    h->m.count = m.count;
    for (int k = 0; k < 5; k++) {
        dst[k]->resize((size_t)m.count * m.width(k));
        if (!dst[k]->empty()) {
            const cudaError_t e = cudaMemcpy(dst[k]->data(), m.dev[k], dst[k]->size() * sizeof(float), cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { delete h; return fail(std::string("cudaMemcpy: ") + cudaGetErrorString(e)); }
        }
    }
    *out = h;
    return GSB_OK;
}

int gsb_app_model_count(gsb_app* a) { return a ? a->model.count : 0; }

int gsb_app_model_device(gsb_app* a, int* count, int* capacity, int* shDegree, int* shCoeffs, float** dev5) {
    if (!a) return fail("gsb_app_model_device: invalid argument");
    if (count) *count = a->model.count;
    if (capacity) *capacity = a->model.capacity;
    if (shDegree) *shDegree = a->model.shDegree;
    if (shCoeffs) *shCoeffs = a->model.shCoeffs;
    if (dev5) for (int k = 0; k < 5; k++) dev5[k] = a->model.dev[k];
    return GSB_OK;
}

int gsb_app_load_input(gsb_app* a) {               // UiPanelToolsInput::refreshProject, :39-40
    if (!a) return fail("gsb_app_load_input: invalid argument");
    APP_OK(gsb_rtx_load_model(a->rtx, gsb_project_get_string(a->project, "pathModel")));
    return gsb_rtx_load_texture_diffuse(a->rtx, gsb_project_get_string(a->project, "pathTextureDiffuse"));
}

int gsb_app_random_rotate(gsb_app* a) {            // UiPanelToolsTruth::onButtonRandomRotate, :192-197
    if (!a) return fail("gsb_app_random_rotate: invalid argument");
    gsb_project_data* d = gsb_project_fields(a->project);
    d->sphere1.rotX = ((float)rand() / (float)RAND_MAX) * 360.0f;
    d->sphere1.rotY = ((float)rand() / (float)RAND_MAX) * 360.0f;
    d->sphere2.rotX = ((float)rand() / (float)RAND_MAX) * 360.0f;
    d->sphere2.rotY = ((float)rand() / (float)RAND_MAX) * 360.0f;
    return GSB_OK;
}

int gsb_app_capture_truths(gsb_app* a) {           // Trainer::captureTruths, src/Trainer.cu:218-250
    if (!a) return fail("gsb_app_capture_truths: invalid argument");
    const gsb_project_data* d = gsb_project_fields(a->project);
    const int n = gsb_cameras_count(d);
    if (a->lastTruthCount != n) {                  // the number of truth frames changed: re-allocate
        a->lastTruthCount = n;
        APP_OK(alloc_truths(a, n));
    }
    std::vector<gsb_camera_desc> cams((size_t)std::max(n, 1));
    APP_OK(gsb_cameras_get(d, cams.data(), n));
    for (int i = 0; i < n; i++) {
        // white and black frame of one camera from one trace (the reference launches the tracer twice, :230-236)
        APP_OK(gsb_rtx_render_pair(a->rtx, a->truthW[(size_t)i], a->truthB[(size_t)i], a->W, a->H, &cams[(size_t)i], d->rtSamples, nullptr));
        a->truthCameras[(size_t)i] = cams[(size_t)i];
    }
    return GSB_OK;
}

int gsb_app_truth_count(gsb_app* a) { return a ? (int)a->truthW.size() : 0; }

int gsb_app_truth_frames(gsb_app* a, int index, uint32_t** dev_white, uint32_t** dev_black, gsb_camera_desc* camera) {
    if (!a || index < 0 || index >= (int)a->truthW.size()) return fail("gsb_app_truth_frames: index out of range");
    if (dev_white) *dev_white = a->truthW[(size_t)index];
    if (dev_black) *dev_black = a->truthB[(size_t)index];
    if (camera) *camera = a->truthCameras[(size_t)index];
    return GSB_OK;
}

int gsb_app_set_truths(gsb_app* a, int n, const uint32_t* const* dev_white, const uint32_t* const* dev_black,
                       const gsb_camera_desc* cameras) {
    if (!a || n < 0 || (n > 0 && (!dev_white || !dev_black || !cameras))) return fail("gsb_app_set_truths: invalid argument");
    if ((int)a->truthW.size() != n) APP_OK(alloc_truths(a, n));
    a->lastTruthCount = n;
    const size_t bytes = (size_t)a->W * a->H * sizeof(uint32_t);
    for (int i = 0; i < n; i++) {
        APP_CUDA_OK(cudaMemcpy(a->truthW[(size_t)i], dev_white[i], bytes, cudaMemcpyDeviceToDevice));
        APP_CUDA_OK(cudaMemcpy(a->truthB[(size_t)i], dev_black[i], bytes, cudaMemcpyDeviceToDevice));
        a->truthCameras[(size_t)i] = cameras[i];
    }
    return GSB_OK;
}

int gsb_app_train(gsb_app* a, int densify) {       // Trainer::train, src/Trainer.cu:252-543
    if (!a) return fail("gsb_app_train: invalid argument");
    if (a->truthW.empty()) { fail("Can't run training iteration, no truth data available!"); return GSB_ERR_STATE; }
    gsb_project_data* d = gsb_project_fields(a->project);
    d->iterations++;
    APP_OK(bind(a));
    const int n = (int)a->truthCameras.size();
    std::vector<gsb_camera> cams((size_t)n);
    for (int i = 0; i < n; i++) APP_OK(gsb_camera_constants(&a->truthCameras[(size_t)i], a->W, a->H, 1, &cams[(size_t)i]));   // :355-356
    APP_OK(gsb_trainer_set_truths(a->trainer, n, a->truthW.data(), a->truthB.data(), cams.data(), n));
    gsb_project_params params;
    APP_OK(gsb_project_train_params(d, &params));
    int new_count = 0;
    APP_OK(gsb_trainer_train(a->trainer, &params, densify ? 1 : 0, &new_count));
    if (!densify) return GSB_OK;
    ModelDevice next;                               // `delete model; model = new ModelSplatsDevice(modelHost)`, :540-541
    next.capacity = a->model.capacity; next.shDegree = a->model.shDegree; next.shCoeffs = a->model.shCoeffs;
    APP_OK(next.allocate(new_count));
    const int rc = gsb_trainer_densify_commit(a->trainer, next.dev[0], next.dev[1], next.dev[2], next.dev[3], next.dev[4]);
    if (rc != GSB_OK) { next.release(); return rc; }
    a->model.release();
    a->model = next;
    return GSB_OK;
}

int gsb_app_set_auto_training(gsb_app* a, int on) {
    if (!a) return fail("gsb_app_set_auto_training: invalid argument");
    a->autoTraining = on != 0;
    return GSB_OK;
}

int gsb_app_update(gsb_app* a, float delta, int* ran) {             // UiFrame::update, UF:266-298
    if (!a) return fail("gsb_app_update: invalid argument");
    if (ran) *ran = 0;
    gsb_project_data* d = gsb_project_fields(a->project);
    d->previewTimer += delta;
    if (!a->autoTraining) return GSB_OK;
    a->autoTrainingBudget = std::min(1.0f, a->autoTrainingBudget + delta * AUTO_TRAIN_BUDGET);
    if (!(a->autoTrainingBudget >= 1.0f)) return GSB_OK;
    a->autoTrainingBudget = 0.0f;
    const bool capture = d->intervalCapture > 0 && d->iterations % d->intervalCapture == 0;
    const bool densify = d->intervalDensify > 0 && d->iterations % d->intervalDensify == 0;
    int flags = 0;
    if (capture) {
        APP_OK(gsb_app_random_rotate(a));
        APP_OK(gsb_app_capture_truths(a));
        flags |= 2;
    }
    APP_OK(gsb_app_train(a, densify ? 1 : 0));
    flags |= 1 | (densify ? 4 : 0);
    if (ran) *ran = flags;
    return GSB_OK;
}

int gsb_app_save_settings(gsb_app* a, const char* path) { return a ? gsb_project_save(a->project, path) : fail("gsb_app_save_settings: invalid argument"); }

int gsb_app_load_settings(gsb_app* a, const char* path) {
    return a ? gsb_project_load(a->project, path) : fail("gsb_app_load_settings: invalid argument");
}

int gsb_app_save_splats(gsb_app* a, const char* path) {
    if (!a || !path) return fail("gsb_app_save_splats: invalid argument");
    gsb_model_host* h = nullptr;
    APP_OK(gsb_app_get_model(a, &h));
    const int rc = gsb_gobj_save(h, path);
    gsb_model_host_destroy(h);
    return rc;
}

int gsb_app_load_splats(gsb_app* a, const char* path) {
    if (!a || !path) return fail("gsb_app_load_splats: invalid argument");
    gsb_model_host* h = nullptr;
    APP_OK(gsb_gobj_load(&h, path));
    const int rc = upload_model(a, h->m);           // loadSplats keeps project.iterations (UF:447-449)
    gsb_model_host_destroy(h);
    return rc;
}

int gsb_app_render_splats(gsb_app* a, uint32_t* dev_frame_buffer, int width, int height, float splat_scale,
                          const gsb_camera_desc* camera) {
    if (!a || !dev_frame_buffer || !camera) return fail("gsb_app_render_splats: invalid argument");
    APP_OK(bind(a));
    gsb_camera c;
    APP_OK(gsb_camera_constants(camera, width, height, 0, &c));      // Trainer::render's aspect-scaled fov, src/Trainer.cu:196-197
    return gsb_trainer_render(a->trainer, dev_frame_buffer, width, height, splat_scale, &c);
}

static int frame_to_png(const uint32_t* dev_fb, int w, int h, const char* path) {
    std::vector<uint32_t> host((size_t)w * h);
    APP_CUDA_OK(cudaMemcpy(host.data(), dev_fb, host.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return gsb_png_write(path, host.data(), w, h, 1);                  // stbi_flip_vertically_on_write(true), TV:236,254
}

int gsb_app_render_splats_png(gsb_app* a, const char* path) {        // onButtonRenderSplats, TV:244-259
    if (!a || !path) return fail("gsb_app_render_splats_png: invalid argument");
    const gsb_project_data* d = gsb_project_fields(a->project);
    if (d->renderResX <= 0 || d->renderResY <= 0) return fail("gsb_app_render_splats_png: bad render resolution");
    gsb_camera_desc cam;
    APP_OK(gsb_camera_preview(d, &cam));
    uint32_t* fb = nullptr;
    APP_CUDA_OK(cudaMalloc(&fb, (size_t)d->renderResX * d->renderResY * sizeof(uint32_t)));
    int rc = gsb_app_render_splats(a, fb, d->renderResX, d->renderResY, d->previewSplatScale, &cam);
    if (rc == GSB_OK) rc = frame_to_png(fb, d->renderResX, d->renderResY, path);
    cudaFree(fb);
    return rc;
}

int gsb_app_render_rtx_png(gsb_app* a, const char* path) {           // onButtonRenderRtx, TV:227-242
    if (!a || !path) return fail("gsb_app_render_rtx_png: invalid argument");
    const gsb_project_data* d = gsb_project_fields(a->project);
    if (d->renderResX <= 0 || d->renderResY <= 0) return fail("gsb_app_render_rtx_png: bad render resolution");
    gsb_camera_desc cam;
    APP_OK(gsb_camera_preview(d, &cam));
    uint32_t* fb = nullptr;
    APP_CUDA_OK(cudaMalloc(&fb, (size_t)d->renderResX * d->renderResY * sizeof(uint32_t)));
    const float black[3] = {0.0f, 0.0f, 0.0f};
    int rc = gsb_rtx_render(a->rtx, fb, d->renderResX, d->renderResY, &cam, black, d->rtSamples, nullptr, 0, nullptr);
    if (rc == GSB_OK) rc = frame_to_png(fb, d->renderResX, d->renderResY, path);
    cudaFree(fb);
    return rc;
}

}  // extern "C"
