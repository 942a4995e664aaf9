// compat/rasterizer.h -- CudaRasterizer::Rasterizer with the reference's call signatures
// (lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:24-84), implemented as inline
// forwarders to the C ABI of libgsb200.so (include/gsb.h).
//
// Put this directory ahead of the reference's cuda_rasterizer/ on the include path (or install it as
// the `CudaRasterizer` target's public include dir) and link libgsb200.so: callers such as
// src/Trainer.cu:8,175,334,378 and rasterize_points.cu:83,170,214 compile unchanged.
//
// Differences a caller can observe: none in results (see DESIGN.md for tolerances).  Errors: the
// reference throws std::runtime_error only when `debug` is set (auxiliary.h:166-173) and otherwise
// ignores CUDA errors; here every failure of the underlying call throws std::runtime_error carrying
// gsb_last_error().  Work is issued on the legacy default stream, like the reference.
#ifndef GSB_COMPAT_CUDA_RASTERIZER_H_INCLUDED
#define GSB_COMPAT_CUDA_RASTERIZER_H_INCLUDED
#ifndef CUDA_RASTERIZER_H_INCLUDED
#define CUDA_RASTERIZER_H_INCLUDED   // keep the reference's own header from being pulled in as well
#endif

#include <cstddef>
#include <functional>
#include <stdexcept>
#include <vector>

#include "../gsb.h"

namespace CudaRasterizer {

class Rasterizer {
    using Alloc = std::function<char*(size_t)>;
    // gsb_alloc_fn trampoline: `user` points at the caller's std::function
    static char* call_alloc(size_t bytes, void* user) { return (*static_cast<Alloc*>(user))(bytes); }
    static void check(int code) {
        if (code != GSB_OK) throw std::runtime_error(gsb_last_error());
    }

public:
    // rasterizer.h:24-29
    static void markVisible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
        check(gsb_rasterizer_mark_visible(P, means3D, viewmatrix, projmatrix, reinterpret_cast<unsigned char*>(present), nullptr));
    }

    // rasterizer.h:31-56; returns num_rendered
    static int forward(Alloc geometryBuffer, Alloc binningBuffer, Alloc imageBuffer, const int P, int D, int M,
                       const float* background, const int width, int height, const float* means3D, const float* shs,
                       const float* colors_precomp, const float* opacities, const float* scales,
                       const float scale_modifier, const float* rotations, const float* cov3D_precomp,
                       const float* viewmatrix, const float* projmatrix, const float* cam_pos, const float tan_fovx,
                       float tan_fovy, const bool prefiltered, float* out_color, int* radii = nullptr,
                       bool debug = false) {
        int rendered = 0;
        check(gsb_rasterizer_forward(&call_alloc, &geometryBuffer, &call_alloc, &binningBuffer, &call_alloc, &imageBuffer, P,
                                     D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales,
                                     scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx,
                                     tan_fovy, prefiltered ? 1 : 0, out_color, radii, debug ? 1 : 0, nullptr, &rendered));
        return rendered;
    }

    // rasterizer.h:58-83
    static void backward(const int P, int D, int M, int R, const float* background, const int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp, const float* scales,
                         const float scale_modifier, const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* campos, const float tan_fovx,
                         float tan_fovy, const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
                         const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                         float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, bool debug) {
        check(gsb_rasterizer_backward(P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
                                      scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
                                      tan_fovy, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_dmean2D,
                                      dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
                                      debug ? 1 : 0, nullptr));
    }
};

}  // namespace CudaRasterizer

#endif
