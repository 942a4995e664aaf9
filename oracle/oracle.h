/* oracle.h -- TEST INFRASTRUCTURE ONLY.  C interface of the scalar CPU restatement (oracle.cpp) of
 * the reference's hot path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * reference legs may load liboracle.so; the product never does.
 *
 * Parity status: PINNED against the reference's own CUDA build (oracle/_ref, compiled from
 * /root/reference) through the golden vectors in tests/golden/ (generated on a B200 by
 * tools/make_golden.py).  The reference ships no CPU path, tests or fixtures of its own.
 */
#ifndef GS_ORACLE_H
#define GS_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_view oracle_view;

/* Rasterizer::forward restated (rasterizer_impl.cu:198-336).  All pointers are HOST memory.
 * colors_precomp / cov3D_precomp may be NULL.  out_color is [3][H][W]. */
oracle_view* oracle_forward(int P, int D, int M, const float* background, int width, int height,
                            const float* means3D, const float* shs, const float* colors_precomp,
                            const float* opacities, const float* scales, float scale_modifier,
                            const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy,
                            float* out_color);
void oracle_view_free(oracle_view* v);
int oracle_view_num_rendered(const oracle_view* v);
/* field ids: 0 radii(i32 P) 1 depths(f32 P) 2 tiles_touched(u32 P) 3 means2D(f32 2P) 4 cov3D(f32 6P)
 * 5 conic_opacity(f32 4P) 6 rgb(f32 3P) 7 clamped(u8 3P) 8 point_offsets(u32 P) 9 keys sorted(u64 R)
 * 10 point_list(u32 R) 11 ranges(u32 2*tiles) 12 n_contrib(u32 HW) 13 final_T(f32 HW)
 * 14 keys unsorted(u64 R) 15 point_list unsorted(u32 R) */
const void* oracle_view_field(const oracle_view* v, int field);
/* pair counters of the last forward: evaluated (power computed), reached exp, blended */
void oracle_view_pair_counts(const oracle_view* v, long long* out3);

/* Rasterizer::backward restated (rasterizer_impl.cu:340-434, backward.cu).  Gradient arrays follow
 * the reference: mean2D[3P] conic[4P] opacity[P] color[3P] are accumulated into, the rest overwritten
 * for visible splats.  Per-splat sums are accumulated in double (deterministic). */
void oracle_backward(const oracle_view* v, const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                     float* dL_dsh, float* dL_dscale, float* dL_drot);

void oracle_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present);

/* Trainer kernels restated (src/Trainer.cu:19-101). */
void oracle_image_float_to_int(const float* src, uint32_t* fb, int w, int h);
void oracle_image_int_to_loss(const uint32_t* truth, const float* rasterized, float* loss, int w, int h);
void oracle_accumulate(float* var, float* avgLoc, float* avgSh, float* avgScale, float* avgOp, float* avgRot,
                       const float* loc, const float* sh, const float* scale, const float* op, const float* rot,
                       float samples, int shCoeffs, int n);
void oracle_apply(float* loc, float* sh, float* scale, float* op, float* rot, const float* gLoc, const float* gSh,
                  const float* gScale, const float* gOp, const float* gRot, float lr, float lrSh, float lrScale,
                  float lrOpacity, float lrRotation, float maxScale, int shCoeffs, int count);

/* Densify block restated (src/Trainer.cu:437-542) with a canonical order: splits by ascending index,
 * then clones by ascending index, then stable compaction.  Arrays must have room for `capacity`
 * splats; returns the new count.  p13 = gsb_project_params order. */
int oracle_densify(int count, int capacity, int shCoeffs, float* loc, float* sh, float* scale, float* op, float* rot,
                   const float* var, const float* avgLoc, const float* p13);

#ifdef __cplusplus
}
#endif
#endif
