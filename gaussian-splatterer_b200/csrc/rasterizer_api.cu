// C ABI of the rasterizer level: gsb_rasterizer_forward / backward / mark_visible.
// Orchestration counterpart of CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (reference cuda_rasterizer/rasterizer_impl.cu:141-153,198-434); the workspace carving below
// plays the role of Geometry/Binning/ImageState::fromChunk (:155-194) with this library's own
// private layout inside the three caller-allocated blobs.

#include "workspace.cuh"
#include "trainer_kernels.cuh"
#include "../../include/gsb.h"

#include <cstdio>
#include <cstring>
#include <mutex>

namespace {
thread_local char g_err[512] = "";
}  // namespace

int gsb_fail(cudaError_t e, const char* what, const char* file, int line) {
    snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
    return GSB_ERR_CUDA;
}
int gsb_fail_msg(const char* msg) {
    snprintf(g_err, sizeof(g_err), "%s", msg);
    return GSB_ERR_ARG;
}

// getHigherMsb (rasterizer_impl.cu:35-50): smallest b with n < 2^b ... as the reference computes it.
int gsb_tile_bits(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return (int)msb;
}

namespace {

__global__ void fill_background_kernel(float* out, const float* bg, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = bg[0]; out[i + n] = bg[1]; out[i + 2 * (size_t)n] = bg[2];
}

int check(cudaStream_t s, bool debug) {
    if (debug) {
        GSB_CUDA_OK(cudaStreamSynchronize(s));
    }
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

}  // namespace

extern "C" {

const char* gsb_last_error(void) { return g_err; }
int gsb_version(void) { return 100; }


int gsb_rasterizer_forward(gsb_alloc_fn geometry_alloc, void* geometry_user, gsb_alloc_fn binning_alloc,
                           void* binning_user, gsb_alloc_fn image_alloc, void* image_user, int P, int D, int M,
                           const float* background, int width, int height, const float* means3D, const float* shs,
                           const float* colors_precomp, const float* opacities, const float* scales,
                           float scale_modifier, const float* rotations, const float* cov3D_precomp,
                           const float* viewmatrix, const float* projmatrix, const float* cam_pos, float tan_fovx,
                           float tan_fovy, int prefiltered, float* out_color, int* radii, int debug, void* stream,
                           int* num_rendered) {
    cudaStream_t s = (cudaStream_t)stream;
    if (num_rendered) *num_rendered = 0;
    if (P < 0 || width <= 0 || height <= 0 || !geometry_alloc || !binning_alloc || !image_alloc)
        return gsb_fail_msg("gsb_rasterizer_forward: invalid argument");
    if (colors_precomp == nullptr && shs == nullptr)
        return gsb_fail_msg("gsb_rasterizer_forward: neither shs nor colors_precomp given");
    const bool dbg = debug != 0;
    const int grid_x = (width + GSB_TILE - 1) / GSB_TILE, grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    const int tiles = grid_x * grid_y;

    GeomWs g;
    char* gptr = geometry_alloc(GeomWs::required((size_t)P), geometry_user);
    if (!gptr) { gsb_fail_msg("geometry allocator returned NULL"); return GSB_ERR_ALLOC; }
    g.carve(gptr, (size_t)P);
    ImageWs im;
    char* iptr = image_alloc(ImageWs::required((size_t)width * height, (size_t)tiles), image_user);
    if (!iptr) { gsb_fail_msg("image allocator returned NULL"); return GSB_ERR_ALLOC; }
    im.carve(iptr, (size_t)width * height, (size_t)tiles);

    GSB_CUDA_OK(cudaMemsetAsync(g.ctrl, 0, sizeof(GsbCtrl), s));
    GSB_CUDA_OK(cudaMemsetAsync(im.tile_max, 0, im.zero_bytes((size_t)tiles), s));   // tile_max, tile_count, look-back words

    PreprocessArgs pa;
    pa.P = P; pa.D = D; pa.M = M;
    pa.means3D = means3D; pa.scales = scales; pa.scale_modifier = scale_modifier; pa.rotations = rotations;
    pa.opacities = opacities; pa.shs = shs; pa.cov3D_precomp = cov3D_precomp; pa.colors_precomp = colors_precomp;
    pa.viewmatrix = viewmatrix; pa.projmatrix = projmatrix; pa.cam_pos = cam_pos;
    pa.W = width; pa.H = height; pa.tan_fovx = tan_fovx; pa.tan_fovy = tan_fovy;
    pa.focal_y = height / (2.0f * tan_fovy);
    pa.focal_x = width / (2.0f * tan_fovx);
    pa.grid_x = grid_x; pa.grid_y = grid_y; pa.prefiltered = prefiltered;
    pa.radii = g.radii; pa.radii_user = radii; pa.depths = g.depths; pa.tiles_touched = g.tiles_touched;
    pa.recs = g.recs; pa.cov3D = g.cov3D; pa.clamped = g.clamped; pa.tile_count = im.tile_count; pa.ctrl = g.ctrl;
    launch_preprocess_fwd(pa, s);
    if (int e = check(s, dbg)) return e;

    uint32_t epoch = 0;
    BinningArgs ba;
    ba.P = P; ba.r_cap = 0xFFFFFFFFu; ba.grid_x = grid_x; ba.grid_y = grid_y;
    ba.radii = g.radii; ba.tiles_touched = g.tiles_touched; ba.depths = g.depths; ba.recs = g.recs;
    ba.tile_count = im.tile_count; ba.ranges = im.ranges; ba.tile_max = im.tile_max; ba.big_list = im.big_list;
    ba.tile_order = im.tile_order;
    ba.pairs = nullptr; ba.point_list = nullptr; ba.pairs2 = nullptr; ba.chunk_list = nullptr;
    ba.lookback = im.lookback; ba.epoch = &epoch; ba.ctrl = g.ctrl; ba.num_rendered_host = nullptr;
    launch_tile_scan(ba, s);                       // ranges of every tile, R
    if (int e = check(s, dbg)) return e;

    // The one forced device->host crossing of this API (rasterizer_impl.cu:281): R sizes the
    // binning workspace the caller allocates.
    uint32_t R = 0;
    if (P > 0) {
        GSB_CUDA_OK(cudaMemcpyAsync(&R, &g.ctrl->num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        GSB_CUDA_OK(cudaStreamSynchronize(s));
    }
    if (R > 0x7FFFFFFFu) return gsb_fail_msg("gsb_rasterizer_forward: duplicate count exceeds int range");
    BinWs b;
    char* bptr = binning_alloc(BinWs::required((size_t)R), binning_user);
    if (!bptr) { gsb_fail_msg("binning allocator returned NULL"); return GSB_ERR_ALLOC; }
    b.carve(bptr, (size_t)R);
    if (R > 0) {
        ba.r_cap = R;
        ba.pairs = b.pairs; ba.point_list = b.point_list; ba.pairs2 = b.pairs2; ba.chunk_list = b.chunk_list;
        launch_queue_long_lists(ba, s);      // the scan ran before this arena existed
        launch_emit(ba, s);
        launch_tile_sort(ba, s);
        if (int e = check(s, dbg)) return e;
    }

    RenderFwdArgs ra;
    ra.W = width; ra.H = height; ra.grid_x = grid_x; ra.grid_y = grid_y;
    ra.ranges = im.ranges; ra.point_list = b.point_list; ra.recs = g.recs; ra.bg = background;
    ra.out_color = out_color; ra.final_T = im.final_T; ra.n_contrib = im.n_contrib; ra.tile_max = im.tile_max;
    ra.frame = nullptr; ra.tile_order = im.tile_order; ra.ctrl = g.ctrl;
    launch_render_fwd(ra, s);
    if (int e = check(s, dbg)) return e;
    if (num_rendered) *num_rendered = (int)R;
    return GSB_OK;
}

int gsb_rasterizer_backward(int P, int D, int M, int R, const float* background, int width, int height,
                            const float* means3D, const float* shs, const float* colors_precomp, const float* scales,
                            float scale_modifier, const float* rotations, const float* cov3D_precomp,
                            const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                            float tan_fovy, const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
                            const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                            float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                            float* dL_drot, int debug, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (P <= 0) return GSB_OK;
    if (!geom_buffer || !image_buffer || !dL_dpix) return gsb_fail_msg("gsb_rasterizer_backward: invalid argument");
    const bool dbg = debug != 0;
    const int grid_x = (width + GSB_TILE - 1) / GSB_TILE, grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    const int tiles = grid_x * grid_y;
    GeomWs g; g.carve(geom_buffer, (size_t)P);
    ImageWs im; im.carve(image_buffer, (size_t)width * height, (size_t)tiles);
    BinWs b; b.carve(binning_buffer, (size_t)(R > 0 ? R : 0));
    if (radii == nullptr) radii = g.radii;

    GSB_CUDA_OK(cudaMemsetAsync(g.grads, 0, (size_t)P * sizeof(GradRec), s));
    if (R > 0) {
        RenderBwdArgs rb;
        rb.W = width; rb.H = height; rb.grid_x = grid_x; rb.grid_y = grid_y;
        rb.ranges = im.ranges; rb.point_list = b.point_list; rb.recs = g.recs; rb.bg = background;
        rb.final_T = im.final_T; rb.n_contrib = im.n_contrib; rb.tile_max = im.tile_max;
        rb.dL_dpix = dL_dpix; rb.truth = nullptr; rb.rendered = nullptr; rb.grads = g.grads; rb.truth2 = nullptr; rb.grads2 = nullptr; rb.loss_out = nullptr;
        rb.tile_order = im.tile_order; rb.ctrl = g.ctrl;
        launch_render_bwd(rb, s);
        if (int e = check(s, dbg)) return e;
    }
    PreprocessBwdArgs pb;
    memset(&pb, 0, sizeof(pb));
    pb.P = P; pb.D = D; pb.M = M;
    pb.means3D = means3D; pb.scales = scales; pb.scale_modifier = scale_modifier; pb.rotations = rotations; pb.shs = shs;
    pb.cov3D = cov3D_precomp ? cov3D_precomp : g.cov3D;
    pb.cov3D_is_precomp = (scales == nullptr) ? 1 : 0;
    pb.colors_precomp = (shs == nullptr) ? 1 : 0;
    pb.viewmatrix = viewmatrix; pb.projmatrix = projmatrix; pb.cam_pos = campos;
    pb.tan_fovx = tan_fovx; pb.tan_fovy = tan_fovy;
    pb.focal_y = height / (2.0f * tan_fovy);
    pb.focal_x = width / (2.0f * tan_fovx);
    pb.radii = radii; pb.clamped = g.clamped; pb.grads = g.grads; pb.recs = g.recs; pb.W = width; pb.H = height;
    pb.dL_dmean2D = dL_dmean2D; pb.dL_dconic = dL_dconic; pb.dL_dopacity = dL_dopacity; pb.dL_dcolor = dL_dcolor;
    pb.dL_dmean3D = dL_dmean3D; pb.dL_dcov3D = dL_dcov3D; pb.dL_dsh = dL_dsh; pb.dL_dscale = dL_dscale; pb.dL_drot = dL_drot;
    pb.avg = nullptr; pb.samples = 1.f;
    launch_preprocess_bwd(pb, s);
    return check(s, dbg);
}

int gsb_rasterizer_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                                unsigned char* present, void* stream) {
    (void)projmatrix;   // the reference's test only uses the view-space depth (auxiliary.h:152-154)
    launch_mark_visible(P, means3D, viewmatrix, reinterpret_cast<bool*>(present), (cudaStream_t)stream);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

int gsb_rasterizer_state(int P, int width, int height, int R, char* geom_buffer, char* binning_buffer,
                         char* image_buffer, gsb_state_view* out) {
    if (!out || !geom_buffer || !image_buffer) return gsb_fail_msg("gsb_rasterizer_state: invalid argument");
    const int grid_x = (width + GSB_TILE - 1) / GSB_TILE, grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    const int tiles = grid_x * grid_y;
    GeomWs g; g.carve(geom_buffer, (size_t)P);
    ImageWs im; im.carve(image_buffer, (size_t)width * height, (size_t)tiles);
    memset(out, 0, sizeof(*out));
    out->radii = g.radii; out->depths = g.depths; out->tiles_touched = g.tiles_touched;
    out->recs = reinterpret_cast<const float*>(g.recs); out->cov3D = g.cov3D; out->clamped = g.clamped;
    if (binning_buffer && R > 0) {
        BinWs b; b.carve(binning_buffer, (size_t)R);
        out->point_list = b.point_list;
    }
    out->ranges = reinterpret_cast<const uint32_t*>(im.ranges); out->n_contrib = im.n_contrib; out->final_T = im.final_T;
    out->tile_max = im.tile_max; out->counters = &g.ctrl->num_rendered;
    return GSB_OK;
}

int gsb_rasterizer_sorted_keys(int P, int width, int height, int R, char* geom_buffer, char* binning_buffer,
                               char* image_buffer, uint64_t* dev_keys_out, void* stream) {
    if (R <= 0) return GSB_OK;
    if (!geom_buffer || !binning_buffer || !image_buffer || !dev_keys_out) return gsb_fail_msg("gsb_rasterizer_sorted_keys: invalid argument");
    const int grid_x = (width + GSB_TILE - 1) / GSB_TILE, grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    const int tiles = grid_x * grid_y;
    GeomWs g; g.carve(geom_buffer, (size_t)P);
    BinWs b; b.carve(binning_buffer, (size_t)R);
    ImageWs im; im.carve(image_buffer, (size_t)width * height, (size_t)tiles);
    launch_sorted_keys(im.ranges, b.point_list, g.depths, (uint32_t)tiles, dev_keys_out, (cudaStream_t)stream);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

// The order in which the tile grids (tile sort, render forward / backward) take their tiles after a forward pass: CTA
// number t works on tile host_order_out[t] (heaviest list-length class first, common.cuh gsb_lpt_tile).  For tests.
int gsb_rasterizer_tile_order(int P, int width, int height, char* geom_buffer, char* image_buffer, uint32_t* host_order_out,
                              void* stream) {
    if (!geom_buffer || !image_buffer || !host_order_out) return gsb_fail_msg("gsb_rasterizer_tile_order: invalid argument");
    cudaStream_t s = (cudaStream_t)stream;
    const int grid_x = (width + GSB_TILE - 1) / GSB_TILE, grid_y = (height + GSB_TILE - 1) / GSB_TILE;
    const size_t tiles = (size_t)grid_x * grid_y;
    GeomWs g; g.carve(geom_buffer, (size_t)P);
    ImageWs im; im.carve(image_buffer, (size_t)width * height, tiles);
    GsbCtrl ctrl;
    GSB_CUDA_OK(cudaMemcpyAsync(&ctrl, g.ctrl, sizeof(ctrl), cudaMemcpyDeviceToHost, s));
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    size_t at = 0;
    for (int b = GSB_LPT_BUCKETS - 1; b >= 0; b--) {
        const size_t n = ctrl.bucket_n[b];
        if (at + n > tiles) return gsb_fail_msg("gsb_rasterizer_tile_order: class counts exceed the tile count");
        if (n) GSB_CUDA_OK(cudaMemcpyAsync(host_order_out + at, im.tile_order + (size_t)b * tiles, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        at += n;
    }
    GSB_CUDA_OK(cudaStreamSynchronize(s));
    if (at != tiles) return gsb_fail_msg("gsb_rasterizer_tile_order: class counts do not add up to the tile count");
    return GSB_OK;
}

int gsb_measure_fp32_peak(double* out2, void* stream) {
    if (!out2) return gsb_fail_msg("gsb_measure_fp32_peak: null argument");
    return gsb_measure_fp32_peak_impl(out2, (cudaStream_t)stream);
}

int gsb_image_float_to_int(const float* dev_src, uint32_t* dev_fb, int w, int h, void* stream) {
    launch_image_float_to_int(dev_src, dev_fb, w, h, (cudaStream_t)stream);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

int gsb_image_int_to_loss(const uint32_t* dev_truth, const float* dev_rasterized, float* dev_loss, int w, int h,
                          void* stream) {
    launch_image_int_to_loss(dev_truth, dev_rasterized, dev_loss, w, h, (cudaStream_t)stream);
    GSB_CUDA_OK(cudaGetLastError());
    return GSB_OK;
}

}  // extern "C"
