// oracle/ref_shim.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// extern "C" shim around the UNMODIFIED reference rasterizer, compiled from the
// sources where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libgsref.so.  It exposes
//   * CudaRasterizer::Rasterizer::{forward,backward,markVisible}
//       (lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:24-84)
//   * the private Geometry/Binning/Image state arrays, re-derived with the
//       reference's own fromChunk (rasterizer_impl.cu:155-194), so parity tests can
//       diff radii / keys / ranges / n_contrib bit for bit
//   * a harness that replays Trainer::train's call sequence (src/Trainer.cu:252-435)
//       around that rasterizer.  src/Trainer.cu cannot be compiled stand-alone (it
//       pulls wxWidgets / GL / OptiX through src/OpenGLIncludes.h), so its four
//       kernels are extracted verbatim AT BUILD TIME (sed line ranges, see Makefile)
//       into oracle/_ref/trainer_kernels.inc, and the CPU densify block into
//       oracle/_ref/densify_body.inc.  Nothing from the reference is stored in git.
//
// This file itself contains no reference code: only calls and a restated call order.

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <vector>
#include <unordered_set>
#include <stdexcept>
#include <functional>
#include <cuda.h>
#include <cuda_runtime.h>

#include "rasterizer.h"        // reference header (via -I)
#include "rasterizer_impl.h"   // reference header: GeometryState / BinningState / ImageState

#define GLM_FORCE_CUDA
#include <glm/glm.hpp>
#include <glm/gtc/quaternion.hpp>
#include <glm/gtc/type_ptr.hpp>

// The four Trainer kernels, verbatim from src/Trainer.cu:18-101 (generated file).
#include "trainer_kernels.inc"

namespace {

struct Buffers {
    char* geom = nullptr; size_t geom_sz = 0;
    char* bin = nullptr;  size_t bin_sz = 0;
    char* img = nullptr;  size_t img_sz = 0;
};

// "as shipped": cudaMalloc inside the callback every call (src/Trainer.cu:335-337).
// "pooled": grow-only buffers, the best case for the reference.
char* grab(char*& ptr, size_t& cap, size_t n, bool pooled) {
    if (pooled) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            cudaMalloc(&ptr, want);
            cap = want;
        }
        return ptr;
    }
    if (ptr) { cudaFree(ptr); ptr = nullptr; }
    cudaMalloc(&ptr, n);
    cap = n;
    return ptr;
}

void release(Buffers& b) {
    if (b.geom) cudaFree(b.geom);
    if (b.bin) cudaFree(b.bin);
    if (b.img) cudaFree(b.img);
    b = Buffers();
}

Buffers g_buf;          // buffers of the stand-alone forward/backward entry points
char g_err[512] = "";

}  // namespace

extern "C" {

const char* gsref_last_error() { return g_err; }

// Rasterizer::forward (rasterizer.h:33-56).  Returns num_rendered, or -1 on exception.
int gsref_forward(int pooled, int P, int D, int M, const float* background, int width, int height,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* opacities, const float* scales, float scale_modifier,
                  const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                  const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy,
                  int prefiltered, float* out_color, int* radii, int debug) {
    try {
        Buffers& b = g_buf;
        bool pl = pooled != 0;
        return CudaRasterizer::Rasterizer::forward(
            [&](size_t n) { return grab(b.geom, b.geom_sz, n, pl); },
            [&](size_t n) { return grab(b.bin, b.bin_sz, n, pl); },
            [&](size_t n) { return grab(b.img, b.img_sz, n, pl); },
            P, D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx,
            tan_fovy, prefiltered != 0, out_color, radii, debug != 0);
    } catch (const std::exception& e) {
        snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

// Rasterizer::backward (rasterizer.h:58-83) on the buffers of the last gsref_forward.
int gsref_backward(int P, int D, int M, int R, const float* background, int width, int height,
                   const float* means3D, const float* shs, const float* colors_precomp,
                   const float* scales, float scale_modifier, const float* rotations,
                   const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                   const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                   const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                   float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                   float* dL_dscale, float* dL_drot, int debug) {
    try {
        CudaRasterizer::Rasterizer::backward(
            P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
            tan_fovy, radii, g_buf.geom, g_buf.bin, g_buf.img, dL_dpix, dL_dmean2D, dL_dconic,
            dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
        return 0;
    } catch (const std::exception& e) {
        snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

void gsref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

void gsref_release() { release(g_buf); }

// Device pointers into the private state of the last gsref_forward, in this order:
//  0 depths  1 clamped  2 internal_radii  3 means2D  4 cov3D  5 conic_opacity  6 rgb
//  7 point_offsets  8 tiles_touched  9 point_list_keys_unsorted  10 point_list_keys
//  11 point_list_unsorted  12 point_list  13 ranges  14 n_contrib  15 accum_alpha
void gsref_state(int P, int width, int height, int R, void** out) {
    char* g = g_buf.geom;
    char* b = g_buf.bin;
    char* i = g_buf.img;
    CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(g, P);
    CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(i, (size_t)width * height);
    out[0] = gs.depths; out[1] = gs.clamped; out[2] = gs.internal_radii; out[3] = gs.means2D;
    out[4] = gs.cov3D; out[5] = gs.conic_opacity; out[6] = gs.rgb; out[7] = gs.point_offsets;
    out[8] = gs.tiles_touched;
    if (b != nullptr && R > 0) {
        CudaRasterizer::BinningState bs = CudaRasterizer::BinningState::fromChunk(b, R);
        out[9] = bs.point_list_keys_unsorted; out[10] = bs.point_list_keys;
        out[11] = bs.point_list_unsorted; out[12] = bs.point_list;
    } else {
        out[9] = out[10] = out[11] = out[12] = nullptr;
    }
    out[13] = is.ranges; out[14] = is.n_contrib; out[15] = is.accum_alpha;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Trainer::train replay (src/Trainer.cu:252-435).  Same order of mallocs, memsets, copies,
// kernel launches and frees; `as_shipped` keeps debug=true and the per-view cudaMalloc/cudaFree,
// the best case drops both (debug=false, pooled workspaces) and nothing else.
// ---------------------------------------------------------------------------------------------
namespace {

struct RefProject {   // the fields of src/Project.h:26-39 that train() reads
    float lrLocation, lrSh, lrScale, lrOpacity, lrRotation;
    float paramScaleMax;
    float paramCullOpacity, paramCullSize, paramDensifyVariance;
    float paramSplitSize, paramSplitDistance, paramSplitScale, paramCloneDistance;
};

// Minimal stand-ins with the member names the extracted densify block uses
// (src/ModelSplatsHost.h:11-37, src/ModelSplatsDevice.h:8-19).
class ModelSplatsDevice;
class ModelSplatsHost {
public:
    int capacity, shDegree, shCoeffs, count = 0;
    float *locations, *shs, *scales, *opacities, *rotations;
    ModelSplatsHost(int capacity_, int shDegree_, int shCoeffs_)
        : capacity(capacity_), shDegree(shDegree_), shCoeffs(shCoeffs_) {
        locations = new float[(size_t)capacity * 3];
        shs = new float[(size_t)capacity * 3 * shCoeffs];
        scales = new float[(size_t)capacity * 3];
        opacities = new float[(size_t)capacity];
        rotations = new float[(size_t)capacity * 4];
    }
    explicit ModelSplatsHost(const ModelSplatsDevice& d);
    ~ModelSplatsHost() {
        delete[] locations; delete[] shs; delete[] scales; delete[] opacities; delete[] rotations;
    }
    void copy(int to, int from) {
        if (to < 0 || to >= count || from < 0 || from >= count) throw std::runtime_error("copy bounds");
        memcpy(&locations[to * 3], &locations[from * 3], 3 * sizeof(float));
        for (int i = 0; i < shCoeffs * 3; i++) shs[to * 3 * shCoeffs + i] = shs[from * 3 * shCoeffs + i];
        memcpy(&scales[to * 3], &scales[from * 3], 3 * sizeof(float));
        opacities[to] = opacities[from];
        memcpy(&rotations[to * 4], &rotations[from * 4], 4 * sizeof(float));
    }
};

class ModelSplatsDevice {
public:
    int capacity, shDegree, shCoeffs, count = 0;
    float *devLocations = nullptr, *devShs = nullptr, *devScales = nullptr, *devOpacities = nullptr,
          *devRotations = nullptr;
    explicit ModelSplatsDevice(const ModelSplatsHost& h)
        : capacity(h.capacity), shDegree(h.shDegree), shCoeffs(h.shCoeffs) {
        count = h.count;
        cudaMalloc(&devLocations, (size_t)count * 3 * sizeof(float));
        cudaMalloc(&devShs, (size_t)count * 3 * shCoeffs * sizeof(float));
        cudaMalloc(&devScales, (size_t)count * 3 * sizeof(float));
        cudaMalloc(&devOpacities, (size_t)count * sizeof(float));
        cudaMalloc(&devRotations, (size_t)count * 4 * sizeof(float));
        cudaMemcpy(devLocations, h.locations, (size_t)count * 3 * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devShs, h.shs, (size_t)count * 3 * shCoeffs * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devScales, h.scales, (size_t)count * 3 * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devOpacities, h.opacities, (size_t)count * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(devRotations, h.rotations, (size_t)count * 4 * sizeof(float), cudaMemcpyHostToDevice);
    }
    ~ModelSplatsDevice() {
        cudaFree(devLocations); cudaFree(devShs); cudaFree(devScales); cudaFree(devOpacities);
        cudaFree(devRotations);
    }
};

ModelSplatsHost::ModelSplatsHost(const ModelSplatsDevice& d)
    : ModelSplatsHost(d.capacity, d.shDegree, d.shCoeffs) {
    count = d.count;
    cudaMemcpy(locations, d.devLocations, (size_t)count * 3 * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(shs, d.devShs, (size_t)count * 3 * shCoeffs * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(scales, d.devScales, (size_t)count * 3 * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(opacities, d.devOpacities, (size_t)count * sizeof(float), cudaMemcpyDeviceToHost);
    cudaMemcpy(rotations, d.devRotations, (size_t)count * 4 * sizeof(float), cudaMemcpyDeviceToHost);
}

struct RefTrainer {
    int W, H;
    int lastCount = -1;
    float *devBackground, *devMatView, *devMatProjView, *devCameraLocation, *devRasterized;
    float *devVarLocations = nullptr, *devAvgGradLocations = nullptr, *devAvgGradShs = nullptr,
          *devAvgGradScales = nullptr, *devAvgGradOpacities = nullptr, *devAvgGradRotations = nullptr;
    float* devLossPixels = nullptr;
    float *devGradLocations = nullptr, *devGradShs = nullptr, *devGradScales = nullptr,
          *devGradOpacities = nullptr, *devGradRotations = nullptr;
    float *devGradMean2D = nullptr, *devGradConic = nullptr, *devGradColor = nullptr,
          *devGradCov3D = nullptr;
    ModelSplatsDevice* model = nullptr;
    std::vector<const uint32_t*> truthW, truthB;
    std::vector<float> view, projview, campos, tanx, tany;   // per camera
    Buffers pool;
    long long rendered_total = 0;
};

void free_splat_buffers(RefTrainer* t) {
    cudaFree(t->devVarLocations); cudaFree(t->devAvgGradLocations); cudaFree(t->devAvgGradShs);
    cudaFree(t->devAvgGradScales); cudaFree(t->devAvgGradOpacities); cudaFree(t->devAvgGradRotations);
    cudaFree(t->devLossPixels);
    cudaFree(t->devGradLocations); cudaFree(t->devGradShs); cudaFree(t->devGradScales);
    cudaFree(t->devGradOpacities); cudaFree(t->devGradRotations);
    cudaFree(t->devGradMean2D); cudaFree(t->devGradConic); cudaFree(t->devGradColor);
    cudaFree(t->devGradCov3D);
}

}  // namespace

extern "C" {

void* gsref_trainer_create(int W, int H) {
    RefTrainer* t = new RefTrainer();
    t->W = W; t->H = H;
    cudaMalloc(&t->devBackground, 3 * sizeof(float));
    cudaMalloc(&t->devMatView, 16 * sizeof(float));
    cudaMalloc(&t->devMatProjView, 16 * sizeof(float));
    cudaMalloc(&t->devCameraLocation, 3 * sizeof(float));
    cudaMalloc(&t->devRasterized, (size_t)W * H * 3 * sizeof(float));
    return t;
}

void gsref_trainer_destroy(void* h) {
    RefTrainer* t = (RefTrainer*)h;
    cudaFree(t->devBackground); cudaFree(t->devMatView); cudaFree(t->devMatProjView);
    cudaFree(t->devCameraLocation); cudaFree(t->devRasterized);
    free_splat_buffers(t);
    delete t->model;
    release(t->pool);
    delete t;
}

void gsref_trainer_set_model(void* h, int count, int capacity, int D, int M, const float* loc,
                             const float* shs, const float* scales, const float* opac,
                             const float* rot) {
    RefTrainer* t = (RefTrainer*)h;
    ModelSplatsHost host(capacity, D, M);
    host.count = count;
    memcpy(host.locations, loc, (size_t)count * 3 * sizeof(float));
    memcpy(host.shs, shs, (size_t)count * 3 * M * sizeof(float));
    memcpy(host.scales, scales, (size_t)count * 3 * sizeof(float));
    memcpy(host.opacities, opac, (size_t)count * sizeof(float));
    memcpy(host.rotations, rot, (size_t)count * 4 * sizeof(float));
    delete t->model;
    t->model = new ModelSplatsDevice(host);
}

int gsref_trainer_count(void* h) { return ((RefTrainer*)h)->model->count; }

void gsref_trainer_get_model(void* h, float* loc, float* shs, float* scales, float* opac, float* rot) {
    RefTrainer* t = (RefTrainer*)h;
    ModelSplatsHost host(*t->model);
    memcpy(loc, host.locations, (size_t)host.count * 3 * sizeof(float));
    memcpy(shs, host.shs, (size_t)host.count * 3 * host.shCoeffs * sizeof(float));
    memcpy(scales, host.scales, (size_t)host.count * 3 * sizeof(float));
    memcpy(opac, host.opacities, (size_t)host.count * sizeof(float));
    memcpy(rot, host.rotations, (size_t)host.count * 4 * sizeof(float));
}

// Device pointers: 0 var 1 avgLoc 2 avgSh 3 avgScale 4 avgOpacity 5 avgRot 6 rasterized 7 loss
void gsref_trainer_buffers(void* h, void** out) {
    RefTrainer* t = (RefTrainer*)h;
    out[0] = t->devVarLocations; out[1] = t->devAvgGradLocations; out[2] = t->devAvgGradShs;
    out[3] = t->devAvgGradScales; out[4] = t->devAvgGradOpacities; out[5] = t->devAvgGradRotations;
    out[6] = t->devRasterized; out[7] = t->devLossPixels;
}

// truth frames stay owned by the caller (device uint32 RGBA, src/Trainer.cuh:52-53).
void gsref_trainer_set_truths(void* h, int n, const uint32_t* const* devW, const uint32_t* const* devB,
                              const float* view, const float* projview, const float* campos,
                              const float* tanx, const float* tany) {
    RefTrainer* t = (RefTrainer*)h;
    t->truthW.assign(devW, devW + n);
    t->truthB.assign(devB, devB + n);
    t->view.assign(view, view + 16 * n);
    t->projview.assign(projview, projview + 16 * n);
    t->campos.assign(campos, campos + 3 * n);
    t->tanx.assign(tanx, tanx + n);
    t->tany.assign(tany, tany + n);
}

long long gsref_trainer_rendered_total(void* h) { return ((RefTrainer*)h)->rendered_total; }

// One iteration without densify.  `apply` = 0 stops before applyGradients so tests can read
// the averaged gradients.  Returns 0, or -1 on exception.
int gsref_trainer_train(void* h, const float* p13, int as_shipped, int apply) {
    RefTrainer* t = (RefTrainer*)h;
    RefProject project;
    memcpy(&project, p13, sizeof(RefProject));
    ModelSplatsDevice* model = t->model;
    const int RX = t->W, RY = t->H;
    const bool dbg = as_shipped != 0;
    try {
        if (t->truthW.empty()) throw std::runtime_error("Can't run training iteration, no truth data available!");
        if (model->count != t->lastCount) {
            t->lastCount = model->count;
            free_splat_buffers(t);
            size_t c = model->count;
            cudaMalloc(&t->devVarLocations, c * sizeof(float));
            cudaMalloc(&t->devAvgGradLocations, c * 3 * sizeof(float));
            cudaMalloc(&t->devAvgGradShs, c * 3 * model->shCoeffs * sizeof(float));
            cudaMalloc(&t->devAvgGradScales, c * 3 * sizeof(float));
            cudaMalloc(&t->devAvgGradOpacities, c * sizeof(float));
            cudaMalloc(&t->devAvgGradRotations, c * 4 * sizeof(float));
            cudaMalloc(&t->devLossPixels, (size_t)RX * RY * 3 * sizeof(float));
            cudaMalloc(&t->devGradLocations, c * 3 * sizeof(float));
            cudaMalloc(&t->devGradShs, c * 3 * model->shCoeffs * sizeof(float));
            cudaMalloc(&t->devGradScales, c * 3 * sizeof(float));
            cudaMalloc(&t->devGradOpacities, c * sizeof(float));
            cudaMalloc(&t->devGradRotations, c * 4 * sizeof(float));
            cudaMalloc(&t->devGradMean2D, c * 3 * sizeof(float));
            cudaMalloc(&t->devGradConic, c * 4 * sizeof(float));
            cudaMalloc(&t->devGradColor, c * 3 * sizeof(float));
            cudaMalloc(&t->devGradCov3D, c * 6 * sizeof(float));
        }
        size_t c = model->count;
        cudaMemset(t->devVarLocations, 0, c * sizeof(float));
        cudaMemset(t->devAvgGradLocations, 0, c * 3 * sizeof(float));
        cudaMemset(t->devAvgGradShs, 0, c * 3 * model->shCoeffs * sizeof(float));
        cudaMemset(t->devAvgGradScales, 0, c * 3 * sizeof(float));
        cudaMemset(t->devAvgGradOpacities, 0, c * sizeof(float));
        cudaMemset(t->devAvgGradRotations, 0, c * 4 * sizeof(float));

        const size_t n = t->truthW.size();
        for (size_t i = 0; i < n * 2; i++) {
            bool white = i < n;
            const uint32_t* truth = white ? t->truthW[i] : t->truthB[i - n];
            size_t cam = i % n;
            std::vector<float> background;
            for (int k = 0; k < 3; k++) background.emplace_back(white ? 1.0f : 0.0f);
            cudaMemcpy(t->devBackground, background.data(), 3 * sizeof(float), cudaMemcpyHostToDevice);
            cudaMemcpy(t->devMatView, &t->view[16 * cam], 16 * sizeof(float), cudaMemcpyHostToDevice);
            cudaMemcpy(t->devMatProjView, &t->projview[16 * cam], 16 * sizeof(float), cudaMemcpyHostToDevice);
            cudaMemcpy(t->devCameraLocation, &t->campos[3 * cam], 3 * sizeof(float), cudaMemcpyHostToDevice);

            Buffers local;
            Buffers& b = as_shipped ? local : t->pool;
            const bool pooled = !as_shipped;
            int countRendered = CudaRasterizer::Rasterizer::forward(
                [&](size_t N) { return grab(b.geom, b.geom_sz, N, pooled); },
                [&](size_t N) { return grab(b.bin, b.bin_sz, N, pooled); },
                [&](size_t N) { return grab(b.img, b.img_sz, N, pooled); },
                model->count, model->shDegree, model->shCoeffs, t->devBackground, RX, RY,
                model->devLocations, model->devShs, nullptr, model->devOpacities, model->devScales,
                1.0f, model->devRotations, nullptr, t->devMatView, t->devMatProjView,
                t->devCameraLocation, t->tanx[cam], t->tany[cam], false, t->devRasterized, nullptr, dbg);
            t->rendered_total += countRendered;

            imageIntToLoss<<<256, 256>>>(truth, t->devRasterized, t->devLossPixels, 256 * 256, RX, RY);

            cudaMemset(t->devGradLocations, 0, c * 3 * sizeof(float));
            cudaMemset(t->devGradShs, 0, c * 3 * model->shCoeffs * sizeof(float));
            cudaMemset(t->devGradScales, 0, c * 3 * sizeof(float));
            cudaMemset(t->devGradOpacities, 0, c * sizeof(float));
            cudaMemset(t->devGradRotations, 0, c * 4 * sizeof(float));
            cudaMemset(t->devGradMean2D, 0, c * 3 * sizeof(float));
            cudaMemset(t->devGradConic, 0, c * 4 * sizeof(float));
            cudaMemset(t->devGradColor, 0, c * 3 * sizeof(float));
            cudaMemset(t->devGradCov3D, 0, c * 6 * sizeof(float));

            CudaRasterizer::Rasterizer::backward(
                model->count, model->shDegree, model->shCoeffs, countRendered, t->devBackground, RX, RY,
                model->devLocations, model->devShs, nullptr, model->devScales, 1.0f,
                model->devRotations, nullptr, t->devMatView, t->devMatProjView, t->devCameraLocation,
                t->tanx[cam], t->tany[cam], nullptr, b.geom, b.bin, b.img, t->devLossPixels,
                t->devGradMean2D, t->devGradConic, t->devGradOpacities, t->devGradColor,
                t->devGradLocations, t->devGradCov3D, t->devGradShs, t->devGradScales,
                t->devGradRotations, dbg);

            accumulateGradients<<<256, 256>>>(
                t->devVarLocations, t->devAvgGradLocations, t->devAvgGradShs, t->devAvgGradScales,
                t->devAvgGradOpacities, t->devAvgGradRotations, t->devGradLocations, t->devGradShs,
                t->devGradScales, t->devGradOpacities, t->devGradRotations, (float)n * 2.0f,
                model->shCoeffs, 256 * 256, model->count);

            if (as_shipped) release(local);   // cudaFree x3 (src/Trainer.cu:422-424)
        }
        if (apply)
            applyGradients<<<256, 256>>>(
                model->devLocations, model->devShs, model->devScales, model->devOpacities,
                model->devRotations, t->devAvgGradLocations, t->devAvgGradShs, t->devAvgGradScales,
                t->devAvgGradOpacities, t->devAvgGradRotations, project.lrLocation, project.lrSh,
                project.lrScale, project.lrOpacity, project.lrRotation, project.paramScaleMax,
                model->shCoeffs, 256 * 256, model->count);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e));
        return 0;
    } catch (const std::exception& e) {
        snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

// The CPU densify block of Trainer::train (src/Trainer.cu:437-542), verbatim (generated file),
// run on the averaged buffers left by the last gsref_trainer_train.
int gsref_trainer_densify(void* h, const float* p13) {
    RefTrainer* t = (RefTrainer*)h;
    RefProject project;
    memcpy(&project, p13, sizeof(RefProject));
    ModelSplatsDevice*& model = t->model;
    float* devVarLocations = t->devVarLocations;
    float* devAvgGradLocations = t->devAvgGradLocations;
    const bool densify = true;
    try {
#include "densify_body.inc"
        return 0;
    } catch (const std::exception& e) {
        snprintf(g_err, sizeof(g_err), "%s", e.what());
        return -1;
    }
}

// Trainer::render's tail (src/Trainer.cu:204): CHW float -> RGBA8.
void gsref_image_float_to_int(const float* src, uint32_t* fb, int w, int h) {
    imageFloatToInt<<<256, 256>>>(src, fb, 256 * 256, w, h);
}

void gsref_image_int_to_loss(const uint32_t* truth, const float* rasterized, float* loss, int w, int h) {
    imageIntToLoss<<<256, 256>>>(truth, rasterized, loss, 256 * 256, w, h);
}

}  // extern "C"
