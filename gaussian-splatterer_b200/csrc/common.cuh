// Shared declarations for the sm_100a kernels of the training hot path.
// Path boundary: CudaRasterizer::Rasterizer::{forward,backward}
//   (reference lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:24-84) and
//   Trainer::train (reference src/Trainer.cu:252-543).
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <cstddef>

#define GSB_TILE 16                 // tile edge in pixels: fixed by key/range parity (config.h:15-16)
#define GSB_TILE_PIX (GSB_TILE * GSB_TILE)

// One 48-byte record per splat and view: everything the render kernels gather per list entry.
// Three float4 so a record moves as three 16-byte cp.async (LDGSTS) transfers.
struct __align__(16) SplatRec {
    float4 a;   // mean2D.x, mean2D.y, conic.x, conic.y
    float4 b;   // conic.z, opacity, r, g
    float4 c;   // b, cull threshold on power, cull half-extent x, cull half-extent y
};

// Per-splat scratch written by the render backward (one vector red per float4) and consumed
// (then cleared) by the preprocess backward.
// With h = G * dL/dalpha and d = mean2D - pixel, summed over all blended (pixel, splat) pairs:
struct __align__(16) GradRec {
    float4 a;   // H0 = sum h, Hx = sum h dx, Hy = sum h dy, Hxx = sum h dx dx
    float4 b;   // Hxy = sum h dx dy, Hyy = sum h dy dy, dL/dr, dL/dg
    float4 c;   // dL/db, unused x3
};

// Fused white+black camera pass (Trainer): the same nine sums of both passes, interleaved as (white, black) pairs
// so the render backward adds a pair with one 64-bit RED; v[r] follows the float order of GradRec (r < 9).
struct __align__(16) GradRec2 {
    float2 v[12];   // v[0..5] = H0, Hx, Hy, Hxx, Hxy, Hyy; v[6..8] = dL/dr, dL/dg, dL/db; v[9..11] unused
};

// Device-resident control block, cleared by one (32-byte) memset per view.
struct GsbCtrl {
    uint32_t num_rendered;          // R: total duplicates (device copy; host copy is optional)
    uint32_t num_visible;           // splats with radius > 0
    uint32_t overflow;              // 1 if R exceeded the binning capacity
    uint32_t ticket;                // dynamic scan-tile ids of tile_scan_kernel
    uint32_t num_big;               // tiles queued for tile_merge_kernel
    uint32_t num_chunks;            // 1024-key chunks of those tiles queued for chunk_sort_kernel
    uint32_t pad[2];
#ifndef GSB_LPT_FINE
#define GSB_LPT_FINE 0              // 1: four classes per octave (32 classes) instead of two (16): A/B knob
#endif
    uint32_t bucket_n[GSB_LPT_FINE ? 32 : 16];   // tiles per list-length class (gsb_lpt_bucket): the heavy-first tile order of the tile grids
};
constexpr int GSB_LPT_BUCKETS = GSB_LPT_FINE ? 32 : 16;

#define GSB_CUDA_OK(expr)                                                              \
    do {                                                                               \
        cudaError_t _e = (expr);                                                       \
        if (_e != cudaSuccess) return gsb_fail(_e, #expr, __FILE__, __LINE__);         \
    } while (0)

int gsb_fail(cudaError_t e, const char* what, const char* file, int line);
int gsb_fail_msg(const char* msg);

static inline int gsb_div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

// Warp-cooperative walk over the tile rectangles of the warp's 32 splats (device code, full warp converged).
// A per-thread loop over its own rectangle runs with ~5 of 32 lanes active (rectangle sizes differ), and every RED /
// ATOM instruction costs its ~40 LSU cycles whatever the number of active lanes.  Here the (splat, tile) pairs of all
// lanes are flattened with a warp prefix sum: lane l takes pairs l, l + 32, ... (two per round), finds the owning lane by
// a 5-step binary search over the prefix (shuffles) and gets the rectangle from that lane.
//   cnt = tiles of this lane's rectangle (0 if the splat is invisible), rx0 / ry0 / w = its origin and width in tiles.
//   f(on[2], src_lane[2], tile_x[2], tile_y[2]) is called by ALL lanes once per round with this lane's two pairs
//   (on[u] false = no pair), so a caller can shuffle from src_lane and issue two atomics back to back (emit_kernel).
#ifdef __CUDACC__
template <typename F>
__device__ __forceinline__ void gsb_warp_walk_tiles(uint32_t rx0, uint32_t ry0, uint32_t w, uint32_t cnt, F&& f) {
    const int lane = threadIdx.x & 31;
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    const uint32_t excl = inc - cnt;
    for (uint32_t base = 0; base < total; base += 64) {
        int src[2];
        uint32_t tx[2], ty[2];
        bool on[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const uint32_t it = base + 32 * u + lane;
            int sl = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {       // largest lane whose exclusive prefix is <= it
                const int cand = sl + step;
                const uint32_t e = __shfl_sync(0xffffffffu, excl, cand);
                if (e <= it) sl = cand;
            }
            const uint32_t k = it - __shfl_sync(0xffffffffu, excl, sl);
            const uint32_t sx = __shfl_sync(0xffffffffu, rx0, sl), sy = __shfl_sync(0xffffffffu, ry0, sl);
            const uint32_t sw = max(__shfl_sync(0xffffffffu, w, sl), 1u);
            const uint32_t ky = (uint32_t)__fdividef((float)k + 0.5f, (float)sw);   // k, sw < 2^20: exact floor(k / sw)
            src[u] = sl; tx[u] = sx + (k - ky * sw); ty[u] = sy + ky; on[u] = it < total;
        }
        f(on, src, tx, ty);
    }
}
#endif

// Heavy-first order of the tile grids.  The render kernels and the tile sort run one CTA per tile and a tile's list is walked
// serially, so a long list that starts late IS the tail of the kernel (ncu: SMs active 90 % of the forward's duration in
// raster order).  The tile scan therefore files every tile under a list-length class (two per octave from 32 entries up)
// and CTA number t takes the t-th tile counting from the heaviest class down -- longest-processing-time-first, to the
// resolution of the classes.  Every tile appears exactly once (empty ones in class 0).
#ifdef __CUDACC__
__device__ __forceinline__ int gsb_lpt_bucket(uint32_t n) {
    if (n < 32u) return 0;
    const int e = 31 - __clz((int)n);                       // >= 5
#if GSB_LPT_FINE
    return min(GSB_LPT_BUCKETS - 1, 4 * (e - 5) + (int)((n >> (e - 2)) & 3u) + 1);
#else
    return min(GSB_LPT_BUCKETS - 1, 2 * (e - 5) + (int)((n >> (e - 1)) & 1u) + 1);
#endif
}
__device__ __forceinline__ uint32_t gsb_lpt_tile(const GsbCtrl* ctrl, const uint32_t* order, uint32_t tiles, uint32_t t) {
    if (order == nullptr) return t;
#pragma unroll 1
    for (int b = GSB_LPT_BUCKETS - 1; b > 0; b--) {
        const uint32_t n = ctrl->bucket_n[b];
        if (t < n) return order[(size_t)b * tiles + t];
        t -= n;
    }
    return order[t];
}
#endif

// NVTX ranges around the host-side stages: compiled in with `make EXTRA=-DGSB_NVTX`.
#ifdef GSB_NVTX
#include <nvtx3/nvToolsExt.h>
struct GsbNvtxRange {
    explicit GsbNvtxRange(const char* name) { nvtxRangePushA(name); }
    ~GsbNvtxRange() { nvtxRangePop(); }
};
#define GSB_NVTX_CAT2(a, b) a##b
#define GSB_NVTX_CAT(a, b) GSB_NVTX_CAT2(a, b)
#define GSB_NVTX(name) GsbNvtxRange GSB_NVTX_CAT(gsb_nvtx_, __LINE__)(name)
#else
#define GSB_NVTX(name) do { } while (0)
#endif

// ------------------------------------------------------------------------------------------
// Kernel launchers (defined in the .cu files of this directory).  All take a stream and never
// synchronise; sizes that live on the device are passed as pointers into GsbCtrl.
// ------------------------------------------------------------------------------------------

struct PreprocessArgs {
    int P, D, M;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* opacities;
    const float* shs;
    const float* cov3D_precomp;
    const float* colors_precomp;
    const float* viewmatrix;        // device, 16 floats column-major
    const float* projmatrix;        // device, 16 floats column-major
    const float* cam_pos;           // device, 3 floats
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y;
    int prefiltered;
    // outputs
    int* radii;                     // [P]
    int* radii_user;                // [P] or null (caller's optional array)
    float* depths;                  // [P]
    uint32_t* tiles_touched;        // [P]
    SplatRec* recs;                 // [P]
    float* cov3D;                   // [P][6]
    uint8_t* clamped;               // [P] bit c set if channel c was clamped at 0
    uint32_t* tile_count;           // [tiles] += 1 for every tile of a visible splat's rectangle (zero on entry)
    GsbCtrl* ctrl;                  // num_visible += visible splats
};
void launch_preprocess_fwd(const PreprocessArgs& a, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s);

// binning.cu
struct BinningArgs {
    int P;
    uint32_t r_cap;                 // capacity of the duplicate arrays
    int grid_x, grid_y;
    const int* radii;
    const uint32_t* tiles_touched;
    const float* depths;
    const SplatRec* recs;
    uint32_t* tile_count;           // [tiles] per-tile duplicate counts from preprocess; cursors after the scan; zero after the sort
    uint2* ranges;                  // [tiles] written for every tile by the scan
    uint32_t* tile_max;             // [tiles] zeroed by the sort (render forward takes atomicMax into it)
    uint32_t* big_list;             // [tiles] queue of the tiles whose lists one warp cannot sort in registers
    uint32_t* tile_order;           // [GSB_LPT_BUCKETS][tiles] tiles by list-length class, written by the scan (heavy-first order of the tile grids)
    uint2* chunk_list;              // [R / 512 + 2] (tile, chunk) queue of their 1024-key chunks
    unsigned long long* pairs2;     // [R] second buffer of the merge rounds
    unsigned long long* pairs;      // [r_cap] depth_bits << 32 | splat, bucketed by tile, sorted in place for long lists
    uint32_t* point_list;           // [r_cap] splat ids sorted by (tile, depth, splat)
    unsigned long long* lookback;   // [scan tiles] chained-scan status words
    uint32_t* epoch;                // host counter, bumped per look-back launch
    GsbCtrl* ctrl;
    uint32_t* num_rendered_host;    // pinned host mirror of R, or null
};
// The stages, in order (preprocess has filled tile_count).
void launch_tile_scan(const BinningArgs& a, cudaStream_t s);
void launch_emit(const BinningArgs& a, cudaStream_t s);
void launch_queue_long_lists(const BinningArgs& a, cudaStream_t s);   // only when the tile scan ran without a chunk queue
void launch_tile_sort(const BinningArgs& a, cudaStream_t s);
void launch_sorted_keys(const uint2* ranges, const uint32_t* point_list, const float* depths, uint32_t tiles, uint64_t* out,
                        cudaStream_t s);

// render_fwd.cu
struct RenderFwdArgs {
    int W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const SplatRec* recs;
    const float* bg;                // device, 3 floats
    float* out_color;               // [3][H][W]
    float* final_T;                 // [H*W]
    uint32_t* n_contrib;            // [H*W]
    uint32_t* tile_max;             // [tiles] max n_contrib in tile
    uint32_t* frame;                // preview mode (Trainer::render): RGBA8 [H*W] written instead of everything above
    const uint32_t* tile_order;     // [GSB_LPT_BUCKETS][tiles] + ctrl: CTA -> tile, longest lists first (null: raster order)
    const GsbCtrl* ctrl;
    float expf_k1, expf_k252;       // set by the launcher: two constants of expf2_pinned (render_common.cuh) as kernel parameters
};
void launch_render_fwd(const RenderFwdArgs& a, cudaStream_t s);

// render_bwd.cu
// The fused loss is one double atomic per warp.  Into a single address they serialise in one L2 slice: 32 640 of them per
// launch at 1080p cost 14 us per launch (0.9 ms of a 65 ms step, tools/e2e_bisect.py); spread by tile over this many
// addresses (the host adds the slots up after the copy) they disappear under the kernel.
#ifndef GSB_LOSS_SLOTS_N
#define GSB_LOSS_SLOTS_N 128
#endif
constexpr int GSB_LOSS_SLOTS = GSB_LOSS_SLOTS_N;   // a power of two

struct RenderBwdArgs {
    int W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const SplatRec* recs;
    const float* bg;
    const float* final_T;
    const uint32_t* n_contrib;
    const uint32_t* tile_max;
    const float* dL_dpix;           // [3][H][W], or null when the loss is fused:
    const uint32_t* truth;          //   RGBA8 truth frame and
    const float* rendered;          //   the forward's out_color (dL = truth/255 - rendered)
    GradRec* grads;                 // [P], accumulated with red.global.add
    // dual mode (non-null truth2): `truth` is the white-background frame, `truth2` the black one,
    // `rendered` the black-background image; both passes accumulate into grads2 (grads is not touched)
    const uint32_t* truth2;
    GradRec2* grads2;
    double* loss_out;               // optional: [GSB_LOSS_SLOTS] partial sums of squared residuals, += (fused-loss modes only)
    const uint32_t* tile_order;     // [GSB_LPT_BUCKETS][tiles] + ctrl: CTA -> tile, longest lists first (null: raster order)
    const GsbCtrl* ctrl;
    float expf_k1, expf_k252;       // set by the launcher (see RenderFwdArgs)
};
void launch_render_bwd(const RenderBwdArgs& a, cudaStream_t s);

// measure.cu -- measurement only: (pixel, splat) pair counts of one view as the reference's
// algorithm defines them, for the FLOP side of the roofline.
// out[6] += { E_eval_f, E_exp_f, E_blend_f, E_eval_b, E_exp_b, E_blend_b }
void launch_count_pairs(const RenderFwdArgs& a, unsigned long long* out, cudaStream_t s);
int gsb_measure_fp32_peak_impl(double* out2, cudaStream_t s);
void launch_bwd_skeleton(const RenderBwdArgs& a, cudaStream_t s);   // A/B lower bound, see measure.cu
void launch_reduction_stats(const RenderBwdArgs& a, unsigned long long* out40, cudaStream_t s);   // measurement, see measure.cu
void launch_staging_ab(const RenderFwdArgs& a, const void* tensor_map, int mode, float* sink, cudaStream_t s);   // A/B, see measure.cu

// preprocess_bwd.cu
struct PreprocessBwdArgs {
    int P, D, M;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* shs;
    const float* cov3D;             // precomputed or the forward's
    int cov3D_is_precomp;
    int colors_precomp;             // nonzero: no SH backward
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    const int* radii;
    const uint8_t* clamped;
    GradRec* grads;                 // consumed and cleared
    GradRec2* grads2;               // Trainer mode only: both passes of a fused camera pass (replaces grads), or null
    const SplatRec* recs;           // the forward's records (conic, opacity)
    int W, H;
    // mode A (Rasterizer::backward): caller arrays, reference semantics
    float* dL_dmean2D;              // [P][3] +=
    float* dL_dconic;               // [P][4] +=
    float* dL_dopacity;             // [P]    +=
    float* dL_dcolor;               // [P][3] +=
    float* dL_dmean3D;              // [P][3] overwritten for visible splats
    float* dL_dcov3D;               // [P][6]
    float* dL_dsh;                  // [P][M][3]
    float* dL_dscale;               // [P][3]
    float* dL_drot;                 // [P][4]
    // mode B (Trainer): running means, layout [var P | loc 3P | sh 3MP | scale 3P | opacity P | rot 4P]
    float* avg;                     // null selects mode A
    float samples;                  // S = 2 * number of cameras
};
void launch_preprocess_bwd(const PreprocessBwdArgs& a, cudaStream_t s);
