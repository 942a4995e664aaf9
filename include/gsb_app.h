/* gsb_app.h -- C ABI of the headless application layer around the training hot path.
 *
 * The reference is a wxWidgets application; what sits either side of Trainer::train is re-built here without any GUI
 * or graphics API, in C++ inside libgsb200.so, behind plain-C entry points:
 *
 *   gsb_project_*      Project + settings.json             src/Project.h:6-73, src/ui/UiFrame.cpp:323-331,358-370   (PJ, UF)
 *   gsb_cameras_*      Camera::getCameras / preview        src/Camera.cpp:29-86, src/Camera.h:8-26                  (CA)
 *   gsb_model_host_*   ModelSplatsHost                      src/ModelSplatsHost.{h,cpp}                              (MH)
 *   gsb_gobj_*         .gobj text format                    src/ui/UiFrame.cpp:333-356,372-450                       (UF)
 *   gsb_field_*        grid / mono / per-triangle fields    src/ui/UiFrame.cpp:137-264                               (UF)
 *   gsb_png_*          screenshot PNG (vertical flip)       src/ui/tools/UiPanelToolsView.cpp:227-259, lib/stb       (TV)
 *   gsb_rtx_*          truth ray caster (RtxHost/RtxDevice) src/rtx/RtxHost.{h,cpp}, src/rtx/RtxDevice.{cu,cuh}      (RH, RD)
 *   gsb_app_*          the UiFrame::update scheduler        src/ui/UiFrame.cpp:266-298, src/ui/tools/UiPanelToolsTruth.cpp:185-197
 *
 * Conventions as in gsb.h: every function returns GSB_OK or an error code and sets gsb_last_error(); pointers are host
 * pointers unless named dev_*; no torch / C++ types cross the boundary.
 */
#ifndef GSB_APP_H
#define GSB_APP_H

#include <stddef.h>
#include <stdint.h>

#include "gsb.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------ Project (PJ:6-73) */
typedef struct gsb_camera_sphere {          /* Project::CameraSphere PJ:14-22 */
    int count;
    float distance, fovDeg, rotX, rotY;
} gsb_camera_sphere;

/* Every non-string member of Project in declaration order (PJ:14-62); bools are ints. */
typedef struct gsb_project_data {
    gsb_camera_sphere sphere1, sphere2;
    int rtSamples;
    float lrLocation, lrSh, lrScale, lrOpacity, lrRotation;
    float paramScaleMax;
    float paramCullOpacity, paramCullSize, paramDensifyVariance, paramSplitSize, paramSplitDistance, paramSplitScale,
        paramCloneDistance;
    int iterations, intervalCapture, intervalDensify;
    float previewTimer;
    int previewRtSamples;
    float previewSplatScale;
    int previewTruth, previewTruthIndex;
    int previewFreeOrbit;
    float previewFreeOrbitSpeed, previewFreeDistance, previewFreeFovDeg, previewFreeRotX, previewFreeRotY;
    int renderResX, renderResY;
} gsb_project_data;

typedef struct gsb_project gsb_project;     /* gsb_project_data + perspective, pathModel, pathTextureDiffuse */

int gsb_project_create(gsb_project** out);                       /* Project{} defaults (PJ) */
int gsb_project_init_new(gsb_project* p);                        /* UiFrame::initProject UF:130-135: sphere2 = {0 cameras, fov 30} */
int gsb_project_destroy(gsb_project* p);
gsb_project_data* gsb_project_fields(gsb_project* p);            /* live pointer, valid until destroy */
const char* gsb_project_get_string(gsb_project* p, const char* name);           /* "perspective" | "pathModel" | "pathTextureDiffuse" */
int gsb_project_set_string(gsb_project* p, const char* name, const char* value);
/* nlohmann to_json + dump(): compact, keys sorted, floats widened to double and printed shortest-round-trip (UF:323-331).
 * Writes at most cap bytes including the terminator; *needed = strlen + 1. */
int gsb_project_to_json(gsb_project* p, char* buf, size_t cap, size_t* needed);
/* nlohmann from_json of NLOHMANN_DEFINE_TYPE_INTRUSIVE_WITH_DEFAULT: absent keys become Project{} defaults, wrong types
 * are errors (the messages of json.exception.type_error.302/306, parse_error.101), unknown keys are ignored (UF:358-370). */
int gsb_project_from_json(gsb_project* p, const char* text);
int gsb_project_save(gsb_project* p, const char* path);          /* UiFrame::saveSettings (perspective is whatever was set) */
int gsb_project_load(gsb_project* p, const char* path);          /* UiFrame::loadSettings: missing file is an error */
int gsb_project_train_params(const gsb_project_data* d, gsb_project_params* out);    /* the 13 floats Trainer::train reads */

/* ------------------------------------------------------------------------------------------------- Camera (CA:29-86) */
typedef struct gsb_camera_desc {            /* Camera CA(h):16-21 */
    float location[3];
    float target[3];
    float fovDegY;
} gsb_camera_desc;

int gsb_cameras_count(const gsb_project_data* d);                                  /* Camera::getCamerasCount */
int gsb_cameras_get(const gsb_project_data* d, gsb_camera_desc* out, int cap);     /* Camera::getCameras, sphere1 then sphere2 */
int gsb_camera_preview(const gsb_project_data* d, gsb_camera_desc* out);           /* Camera::getPreviewCamera */
/* view = -lookAt, projview = perspective(fovY, w/h, 0.1, 100) * view; square_fov != 0 is Trainer::train's
 * tan_fovx = tan_fovy (src/Trainer.cu:355-356), else Trainer::render's aspect-scaled angle (src/Trainer.cu:196-197). */
int gsb_camera_constants(const gsb_camera_desc* c, int width, int height, int square_fov, gsb_camera* out);
int gsb_camera_inverse_projview(const gsb_camera_desc* c, float aspect, float out16[16]);   /* glm::inverse, RH:212 */

/* ----------------------------------------------------------------------------------- ModelSplatsHost (MH), .gobj, fields */
typedef struct gsb_model_host gsb_model_host;

typedef struct gsb_model_host_view {        /* the public members of ModelSplatsHost MH(h):11-21 */
    int capacity, shDegree, shCoeffs, count;
    float *locations, *shs, *scales, *opacities, *rotations;
} gsb_model_host_view;

int gsb_model_host_create(gsb_model_host** out, int capacity, int shDegree, int shCoeffs);           /* MH:6-15 */
/* The five-vector constructor MH:31-63: capacity = 1e6 * 10^k >= count, shCoeffs / shDegree derived from the SH vector,
 * "Inconsistent feature dimensions supplied when creating a host model!" on mismatch. n_* are float counts. */
int gsb_model_host_from_arrays(gsb_model_host** out, const float* locations, size_t n_locations, const float* shs, size_t n_shs,
                               const float* scales, size_t n_scales, const float* opacities, size_t n_opacities,
                               const float* rotations, size_t n_rotations);
int gsb_model_host_destroy(gsb_model_host* m);
int gsb_model_host_get(gsb_model_host* m, gsb_model_host_view* out);
int gsb_model_host_push_back(gsb_model_host* m, const float location[3], const float* sh, int n_sh, const float scale[3],
                             float opacity, const float rotation_wxyz[4]);                           /* MH:72-84 */
int gsb_model_host_copy(gsb_model_host* m, int indexTo, int indexFrom);                              /* MH:86-98 */
int gsb_gobj_save(gsb_model_host* m, const char* path);          /* UiFrame::saveSplats UF:333-356 (ostream default precision 6) */
int gsb_gobj_load(gsb_model_host** out, const char* path);       /* UiFrame::loadSplats UF:372-450 ("Inconsistent SH degree!") */
int gsb_field_grid(gsb_model_host** out);                        /* UiFrame::initFieldGrid UF:137-160: 17^3 splats, capacity 1e6 */
int gsb_field_mono(gsb_model_host** out);                        /* UiFrame::initFieldMono UF:162-176 */
int gsb_field_model(gsb_model_host** out, const char* obj_path); /* UiFrame::initFieldModel UF:178-264: one splat per triangle */

/* --------------------------------------------------------------------------------------------------------- PNG (TV, stb) */
/* stbi_flip_vertically_on_write(flip) + stbi_write_png(path, w, h, 4, rgba, w*4) (TV:236-238,254-256). host memory. */
int gsb_png_write(const char* path, const uint32_t* rgba, int width, int height, int flip_vertically);
/* stbi_load(path, &w, &h, nullptr, STBI_rgb_alpha) for PNG files (RH:15-17): 8/16-bit grey, grey+alpha, RGB, RGBA, palette,
 * non-interlaced.  *rgba_out is malloc'ed; release with gsb_free. */
int gsb_png_read(const char* path, uint8_t** rgba_out, int* width, int* height);
void gsb_free(void* p);

/* ------------------------------------------------------------------------------- truth ray caster (RH, RD): SURVEY 8 f1 */
/* RtxHost + RtxDevice without OptiX/OWL: a CUDA path tracer over a BVH of the triangle mesh, one thread per pixel,
 * `samples` jittered primary rays per pixel, <= 50 bounces, alpha-tested lambertian scattering, sky = min(1, 1 + dir.y),
 * the background colour where no surface reflected, RGBA8 packing int(c * 256) (RD:8-158, owl_device.h:79-90).
 * Per-pixel random numbers come from owl's LCG<4> seeded with (pixel.x, pixel.y): the reference leaves the generator
 * state uninitialised (RD:57 `PerRayData prd;`, owl random.h:52-55), so its sequence is undefined; this is the seeding owl's
 * own samples use.  The hardware triangle test of OptiX is replaced by Moeller-Trumbore (no culling). */
typedef struct gsb_rtx gsb_rtx;

int gsb_rtx_create(gsb_rtx** out);                               /* RtxHost::RtxHost: no model, blank 8x8 grey texture */
int gsb_rtx_destroy(gsb_rtx* r);
int gsb_rtx_reset(gsb_rtx* r);                                   /* RtxHost::reset RH:72-75 */
int gsb_rtx_load_model(gsb_rtx* r, const char* obj_path);        /* RtxHost::loadModel RH:107-182 (missing path: no-op like :108) */
/* The same from memory (procedural meshes): vertices[3*nv], triangles[3*nt] zero-based, texcoords[6*nt] per corner or NULL. */
int gsb_rtx_set_mesh(gsb_rtx* r, const float* vertices, int nv, const int* triangles, int nt, const float* texcoords);
int gsb_rtx_load_texture_diffuse(gsb_rtx* r, const char* png_path);                  /* RtxHost::loadTextureDiffuse RH:184-189 */
int gsb_rtx_set_texture_diffuse(gsb_rtx* r, const uint8_t* rgba, int width, int height);
int gsb_rtx_model_present(gsb_rtx* r);
/* RtxHost::render RH:191-221.  dev_frame_buffer: width*height RGBA8 on the device.  splat_cameras: n x 3 host floats (the
 * camera indicator orbs of the preview, RD:33-47) or NULL.  Without a model the frame is cleared to 0 (RH:220). */
int gsb_rtx_render(gsb_rtx* r, uint32_t* dev_frame_buffer, int width, int height, const gsb_camera_desc* camera,
                   const float background[3], int samples, const float* splat_cameras, int n_splat_cameras, void* stream);
/* Both truth frames of one camera (white and black background, src/Trainer.cu:230-237) from ONE trace: with a seeded
 * generator the two renders follow identical paths and differ only where no surface reflected. */
int gsb_rtx_render_pair(gsb_rtx* r, uint32_t* dev_frame_white, uint32_t* dev_frame_black, int width, int height,
                        const gsb_camera_desc* camera, int samples, void* stream);
/* rays traced (primary + bounce segments) by the last render call, for the Mray/s figure */
int gsb_rtx_last_ray_count(gsb_rtx* r, unsigned long long* out);

/* ------------------------------------------------------------- headless UiFrame: the training driver */
/* The application object without its windows: one Project, one Trainer (with its ModelSplatsDevice), one RtxHost, and
 * the idle-loop scheduler of UiFrame::update (UF:266-298).  The training resolution is RENDER_RESOLUTION_X/Y of
 * src/Config.h:13-14 in the reference (1024 x 1024); here it is a constructor argument. */
typedef struct gsb_app gsb_app;

int gsb_app_create(gsb_app** out, int train_width, int train_height);   /* UiFrame ctor: initProject + initFieldGrid (UF:126-160) */
int gsb_app_destroy(gsb_app* a);
gsb_project* gsb_app_project(gsb_app* a);
gsb_rtx* gsb_app_rtx(gsb_app* a);
gsb_trainer* gsb_app_trainer(gsb_app* a);
int gsb_app_new_project(gsb_app* a);                             /* FILE_NEW_PROJECT: initProject, initFieldGrid, rtx reset (UF:452-458,300-301) */
int gsb_app_init_field(gsb_app* a, int kind);                    /* 0 grid, 1 mono, 2 from project.pathModel; iterations = 0 (UF:137-264) */
int gsb_app_set_model(gsb_app* a, gsb_model_host* m);            /* delete trainer->model; model = new ModelSplatsDevice(host) (UF:157-158) */
int gsb_app_get_model(gsb_app* a, gsb_model_host** out);         /* ModelSplatsHost(*trainer->model) (MH:17-29) */
int gsb_app_model_count(gsb_app* a);
/* the five device arrays of trainer->model (src/ModelSplatsDevice.h:15-19) for callers that render or inspect them */
int gsb_app_model_device(gsb_app* a, int* count, int* capacity, int* shDegree, int* shCoeffs, float** dev5);
int gsb_app_load_input(gsb_app* a);                              /* rtx.loadModel(project.pathModel) + loadTextureDiffuse(project.pathTextureDiffuse) */
int gsb_app_random_rotate(gsb_app* a);                           /* UiPanelToolsTruth::onButtonRandomRotate: four rand() / RAND_MAX * 360 */
int gsb_app_capture_truths(gsb_app* a);                          /* Trainer::captureTruths (src/Trainer.cu:218-250) through the ray caster */
int gsb_app_truth_count(gsb_app* a);
int gsb_app_truth_frames(gsb_app* a, int index, uint32_t** dev_white, uint32_t** dev_black, gsb_camera_desc* camera);
/* Caller-supplied truth frames instead of the ray caster (device RGBA8 at the training resolution, copied). */
int gsb_app_set_truths(gsb_app* a, int n, const uint32_t* const* dev_white, const uint32_t* const* dev_black,
                       const gsb_camera_desc* cameras);
int gsb_app_train(gsb_app* a, int densify);                      /* Trainer::train(project, densify), the manual buttons (UiPanelToolsTrain.cpp:166-169) */
int gsb_app_set_auto_training(gsb_app* a, int on);               /* autoTraining flag (UF(h)) ; budget restarts at 0 */
/* UiFrame::update with `delta_seconds` in place of the wall clock: previewTimer += delta; if auto-training, the budget
 * grows by delta * AUTO_TRAIN_BUDGET (100 / s) capped at 1, and a full budget buys one iteration: capture (random
 * rotation + captureTruths) when iterations % intervalCapture == 0, densify when iterations % intervalDensify == 0,
 * then Trainer::train.  *ran (optional): 0 none, 1 trained, |2 captured, |4 densified. */
int gsb_app_update(gsb_app* a, float delta_seconds, int* ran);
int gsb_app_save_settings(gsb_app* a, const char* path);         /* UF:323-331 */
int gsb_app_load_settings(gsb_app* a, const char* path);         /* UF:358-370 */
int gsb_app_save_splats(gsb_app* a, const char* path);           /* UF:333-356 */
int gsb_app_load_splats(gsb_app* a, const char* path);           /* UF:372-450 */
/* Trainer::render at (width, height) from `camera` -> RGBA8 in caller device memory (the preview panel's frame buffer,
 * src/ui/UiPanelViewOutput.cpp:47-55; the GL blit of FboRenderer is unavailable on this box). */
int gsb_app_render_splats(gsb_app* a, uint32_t* dev_frame_buffer, int width, int height, float splat_scale,
                          const gsb_camera_desc* camera);
/* onButtonRenderSplats / onButtonRenderRtx (TV:227-259): project.renderResX x renderResY from Camera::getPreviewCamera,
 * previewSplatScale / rtSamples, written as a vertically flipped PNG. */
int gsb_app_render_splats_png(gsb_app* a, const char* path);
int gsb_app_render_rtx_png(gsb_app* a, const char* path);

#ifdef __cplusplus
}
#endif

#endif /* GSB_APP_H */
