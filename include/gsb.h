/* gsb.h -- C ABI of the B200-native Gaussian-Splatterer training hot path (libgsb200.so).
 *
 * Plain pointers and sizes only; every pointer named dev* / documented "device" is CUDA device
 * memory of the current device.  All functions return 0 on success or a non-zero error code and
 * leave a message retrievable with gsb_last_error().  Nothing here falls back to the CPU: without
 * a CUDA device every compute entry point fails.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference
 * repository osreboot/Gaussian-Splatterer @ 5ca294e):
 *   R  = lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h
 *   RI = lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu
 *   T  = src/Trainer.cu, TH = src/Trainer.cuh, P = src/Project.h
 * The C++ headers in include/compat/ wrap this ABI in the reference's own class signatures.
 */
#ifndef GSB_H_INCLUDED
#define GSB_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_OK 0
#define GSB_ERR_CUDA 1        /* a CUDA call failed (message holds cudaGetErrorString) */
#define GSB_ERR_ARG 2         /* invalid argument */
#define GSB_ERR_STATE 3       /* call order / missing data (e.g. train() without truth frames) */
#define GSB_ERR_ALLOC 4       /* a caller-supplied allocator returned NULL */

const char* gsb_last_error(void);
int gsb_version(void);

/* Workspace allocator: the callee asks for `bytes`, the caller returns device memory it owns.
 * Replaces the three std::function<char*(size_t)> arguments of Rasterizer::forward (R:34-36). */
typedef char* (*gsb_alloc_fn)(size_t bytes, void* user);

/* ---- rasterizer level ------------------------------------------------------------------- */

/* Rasterizer::forward (R:33-56, RI:198-336).  Same arguments in the same order; the three
 * std::function allocators become (fn, user) pairs, bool becomes int, the return value moves to
 * *num_rendered and `stream` (a cudaStream_t, NULL = legacy default stream) is added.
 * One host synchronisation is unavoidable: the binning workspace is sized from R (RI:281-285). */
int gsb_rasterizer_forward(gsb_alloc_fn geometry_alloc, void* geometry_user,
                           gsb_alloc_fn binning_alloc, void* binning_user,
                           gsb_alloc_fn image_alloc, void* image_user,
                           int P, int D, int M,
                           const float* background, int width, int height,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* opacities, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp,
                           const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                           float tan_fovx, float tan_fovy, int prefiltered,
                           float* out_color, int* radii, int debug, void* stream,
                           int* num_rendered);

/* Rasterizer::backward (R:58-83, RI:340-434).  Gradient outputs follow the reference exactly:
 * dL_dmean2D / dL_dconic / dL_dopacity / dL_dcolor are accumulated into (caller pre-zeroes them,
 * T:365-375); dL_dmean3D / dL_dcov3D / dL_dsh / dL_dscale / dL_drot are overwritten for visible
 * splats only. */
int gsb_rasterizer_backward(int P, int D, int M, int R,
                            const float* background, int width, int height,
                            const float* means3D, const float* shs, const float* colors_precomp,
                            const float* scales, float scale_modifier, const float* rotations,
                            const float* cov3D_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* campos,
                            float tan_fovx, float tan_fovy, const int* radii,
                            char* geom_buffer, char* binning_buffer, char* image_buffer,
                            const float* dL_dpix,
                            float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                            float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                            float* dL_dscale, float* dL_drot, int debug, void* stream);

/* Rasterizer::markVisible (R:24-29, RI:141-153). */
int gsb_rasterizer_mark_visible(int P, const float* means3D, const float* viewmatrix,
                                const float* projmatrix, unsigned char* present, void* stream);


/* Read-only view into the private workspace layout after gsb_rasterizer_forward, for parity tests
 * (counterpart of Geometry/Binning/ImageState::fromChunk, RI:155-194).  Device pointers. */
typedef struct gsb_state_view {
    const int* radii;                 /* [P] */
    const float* depths;              /* [P] (valid where radii > 0) */
    const uint32_t* tiles_touched;    /* [P] */
    const float* recs;                /* [P][12]: x y conic.x conic.y / conic.z opacity r g / b cull-threshold cull-ex cull-ey */
    const float* cov3D;               /* [P][6] */
    const unsigned char* clamped;     /* [P] bit c = channel c clamped */
    const uint32_t* point_list;       /* [R] splat ids sorted by (tile, depth) */
    const uint32_t* ranges;           /* [tiles][2] */
    const uint32_t* n_contrib;        /* [H*W] */
    const float* final_T;             /* [H*W] */
    const uint32_t* tile_max;         /* [tiles] */
    const uint32_t* counters;         /* [3]: R, visible, overflow */
} gsb_state_view;
int gsb_rasterizer_state(int P, int width, int height, int R, char* geom_buffer,
                         char* binning_buffer, char* image_buffer, gsb_state_view* out);
/* Materialise the reference's 64-bit sort keys (tile << 32 | depth bits, RI:102-104) of the
 * sorted list into dev_keys_out[R]. */
int gsb_rasterizer_sorted_keys(int P, int width, int height, int R, char* geom_buffer,
                               char* binning_buffer, char* image_buffer, uint64_t* dev_keys_out,
                               void* stream);
/* Test hook: the order in which the one-CTA-per-tile grids (tile sort, render forward / backward) take their tiles after a
 * forward pass -- CTA number t works on tile host_order_out[t], longest lists first; a permutation of 0 .. tiles-1. */
int gsb_rasterizer_tile_order(int P, int width, int height, char* geom_buffer, char* image_buffer,
                              uint32_t* host_order_out, void* stream);

/* ---- trainer level ---------------------------------------------------------------------- */

/* The Project fields Trainer::train reads (P:26-39), in declaration order. */
typedef struct gsb_project_params {
    float lrLocation, lrSh, lrScale, lrOpacity, lrRotation;
    float paramScaleMax;
    float paramCullOpacity, paramCullSize, paramDensifyVariance;
    float paramSplitSize, paramSplitDistance, paramSplitScale, paramCloneDistance;
} gsb_project_params;

/* What train() derives from a Camera per view (T:321-326,355-356): view = -lookAt,
 * projview = perspective * view (both column-major), camera location, tan(fov/2). */
typedef struct gsb_camera {
    float view[16];
    float projview[16];
    float location[3];
    float tan_fovx, tan_fovy;
} gsb_camera;

typedef struct gsb_trainer gsb_trainer;

/* Trainer::Trainer / ~Trainer (T:103-146).  Resolution is a runtime parameter here (the
 * reference fixes 1024x1024 at compile time, src/Config.h:13-14). */
int gsb_trainer_create(gsb_trainer** out, int width, int height, void* stream);
int gsb_trainer_destroy(gsb_trainer* t);

/* Bind the five ModelSplatsDevice arrays (src/ModelSplatsDevice.h:8-19).  They stay the canonical
 * parameter storage, owned by the caller; call again whenever `model` is replaced. */
int gsb_trainer_bind_model(gsb_trainer* t, int count, int capacity, int shDegree, int shCoeffs,
                           float* devLocations, float* devShs, float* devScales,
                           float* devOpacities, float* devRotations);

/* truthFrameBuffersW / truthFrameBuffersB / truthCameras (TH:52-54): n device RGBA8 frames per
 * background, owned by the caller.  `total_cameras` is the camera count over ALL ranks (S = 2 *
 * total_cameras scales the running means, T:419); pass n on a single GPU. */
int gsb_trainer_set_truths(gsb_trainer* t, int n, const uint32_t* const* devW,
                           const uint32_t* const* devB, const gsb_camera* cameras,
                           int total_cameras);

/* Optional caller-owned buffer for the 24*count running means
 * [var | loc 3 | sh 3*shCoeffs | scale 3 | opacity | rot 4] so the caller can all-reduce it. */
int gsb_trainer_set_grad_buffer(gsb_trainer* t, float* dev_avg);
size_t gsb_trainer_grad_floats(const gsb_trainer* t);
float* gsb_trainer_grad_buffer(gsb_trainer* t);

/* Exchange step between accumulate and apply on multi-GPU runs: sum `count` floats at `dev`
 * across ranks, in place, ordered after `stream`. */
typedef int (*gsb_allreduce_fn)(float* dev, size_t count, void* stream, void* user);
int gsb_trainer_set_allreduce(gsb_trainer* t, gsb_allreduce_fn fn, void* user);

/* Native exchange step: NCCL communicators owned by the library, so a C++ caller of this ABI
 * (or of include/compat/Trainer.cuh) can shard the truth cameras over several GPUs without any Python.  NCCL is bound
 * at run time (libnccl.so.2); single-GPU callers never load it.
 *   one process per GPU:  rank 0 calls gsb_comm_get_unique_id, ships the 128 bytes to the other ranks out of band
 *                         (file, socket, MPI, torch.distributed), every rank calls gsb_comm_init on ITS current device
 *   one process, G GPUs:  gsb_comm_init_all fills out[0..ndev) (devices = NULL means 0..ndev-1); each communicator is
 *                         then driven from its own host thread with its device current
 * gsb_trainer_set_comm makes gsb_trainer_train sum the running means over the communicator (on the trainer's stream)
 * between accumulate and apply; it takes precedence over the gsb_trainer_set_allreduce hook.  Every rank must call
 * train() the same number of times.  NCCL returns bit-identical sums on all ranks and apply / densify are
 * deterministic, so the replicas stay bit-identical without a parameter broadcast. */
#define GSB_COMM_UNIQUE_ID_BYTES 128
typedef struct gsb_comm gsb_comm;
int gsb_comm_nccl_version(int* version);
int gsb_comm_get_unique_id(void* id_out);
int gsb_comm_init(gsb_comm** out, const void* nccl_unique_id, int rank, int world);
int gsb_comm_init_all(gsb_comm** out, int ndev, const int* devices);
int gsb_comm_destroy(gsb_comm* c);
int gsb_comm_rank(const gsb_comm* c);
int gsb_comm_world(const gsb_comm* c);
int gsb_comm_device(const gsb_comm* c);
int gsb_comm_allreduce_sum(gsb_comm* c, float* dev, size_t count, void* stream);
int gsb_comm_broadcast(gsb_comm* c, void* dev, size_t bytes, int root, void* stream);
int gsb_trainer_set_comm(gsb_trainer* t, gsb_comm* comm);

/* Peer windows: the exchange as the library's own kernel over NVLink peer memory instead of ncclAllReduce.
 * gsb_comm_window_create (COLLECTIVE over the communicator) allocates `floats` floats on this rank, ships the
 * allocation to every other rank (cudaIpc handles between processes, raw pointers + peer access between the devices of
 * one process) and maps theirs.  gsb_window_allreduce_sum then runs a two-shot kernel on `stream`: rank r reads slice r
 * of all G windows, adds the G values in rank order and stores the sum into all G windows, between two system-scope
 * flag barriers; when it returns (in stream order) the local window holds the global sums and may be overwritten.
 * Each element is summed by one rank and broadcast as bits: replicas are bit-identical by construction.  If any rank
 * could not map any peer (no P2P / IPC, ranks on different hosts, GSB_PEER_EXCHANGE=0) every rank reports
 * gsb_window_peer_mapped() == 0 and gsb_window_allreduce_sum forwards to ncclAllReduce on the same buffer.
 * gsb_comm_window_destroy is COLLECTIVE too and must precede gsb_comm_destroy.  A trainer with a communicator keeps
 * its running means in such a window (gsb_trainer_exchange_path: 0 none, 1 ncclAllReduce, 2 peer kernel) unless the
 * caller supplied its own buffer with gsb_trainer_set_grad_buffer. */
typedef struct gsb_window gsb_window;
int gsb_comm_window_create(gsb_comm* c, size_t floats, gsb_window** out);
int gsb_comm_window_destroy(gsb_window* w);
float* gsb_window_data(gsb_window* w);
size_t gsb_window_floats(const gsb_window* w);
int gsb_window_peer_mapped(const gsb_window* w);
int gsb_window_allreduce_sum(gsb_window* w, size_t count, void* stream);
int gsb_trainer_exchange_path(const gsb_trainer* t);

/* Trainer::train (T:252-543) split at the point where ranks exchange gradients:
 *   accumulate = zero the means, then forward + loss + backward + accumulate of every local view
 *   apply      = applyGradients (T:428-431)
 *   train      = accumulate, all-reduce hook (if set), apply, optional densify plan
 * With densify != 0 train() stops after gsb_trainer_densify_plan; the caller allocates arrays for
 * *new_count splats and finishes with gsb_trainer_densify_commit (the reference reallocates the
 * model the same way, T:540-541). */
int gsb_trainer_accumulate(gsb_trainer* t);
int gsb_trainer_apply(gsb_trainer* t, const gsb_project_params* p);
int gsb_trainer_train(gsb_trainer* t, const gsb_project_params* p, int densify, int* new_count);
int gsb_trainer_densify_plan(gsb_trainer* t, const gsb_project_params* p, int* new_count);
int gsb_trainer_densify_commit(gsb_trainer* t, float* devLocations, float* devShs, float* devScales,
                               float* devOpacities, float* devRotations);

/* Trainer::render (T:148-216): black background, splatScale as scale_modifier, RGBA8 out. */
int gsb_trainer_render(gsb_trainer* t, uint32_t* dev_frame_buffer, int width, int height,
                       float splat_scale, const gsb_camera* camera);

/* Monitoring / measurement (no reference counterpart). */
typedef struct gsb_trainer_stats {
    long long rendered_total;      /* sum of R over the view passes of the last accumulate */
    long long visible_total;       /* sum of visible splats */
    int views;                     /* view passes of the last accumulate */
    int retries;                   /* accumulate re-runs caused by binning-capacity growth */
    long long kernel_launches;     /* kernels launched by this trainer since creation */
    double sum_sq_residual;        /* sum over views of sum(residual^2), if enabled */
    long long epoch_wraps;         /* look-back epoch counter restarts (every ~2^30 look-back launches) */
    double exchange_ms, apply_ms;  /* with stage timing on: all-reduce and applyGradients, summed over timed_iterations */
    long long timed_iterations;
} gsb_trainer_stats;
int gsb_trainer_get_stats(gsb_trainer* t, gsb_trainer_stats* out);
int gsb_trainer_enable_loss(gsb_trainer* t, int on);
/* Sum of squared residuals of the latest iteration (iterations_back = 0) or of the one before it (1).  The scalar is
 * copied to pinned host memory asynchronously at the end of every iteration; this call waits for that copy only, so
 * reading iteration k-1 while iteration k runs does not drain the device. */
int gsb_trainer_get_loss(gsb_trainer* t, int iterations_back, double* out);
/* White and black pass of a camera (T:311-314) share preprocess, sorts, the forward blend and the
 * alpha / T recomputation of the backward walk (default on; same results up to the summation order
 * of the running means).  0 = one full pass per (camera, background) in the reference's order. */
int gsb_trainer_set_fuse_backgrounds(gsb_trainer* t, int on);

/* Measurement: run training view pass `pass_index` (0..2n-1, white passes first) once with CUDA
 * events between its stages and return the milliseconds of each, in this order:
 *   0 clear, 1 preprocess (+ per-tile counts), 2 tile scan (ranges, R), 3 emit, 4 per-tile sort,
 *   5 render forward, 6 render backward (+fused loss), 7 preprocess backward (+accumulate).
 * counters (optional) receives R and the visible-splat count of that pass.  Needs one prior
 * gsb_trainer_accumulate (arenas sized).  The running means are perturbed; call it outside training. */
#define GSB_NUM_STAGES 8
int gsb_trainer_profile_pass(gsb_trainer* t, int pass_index, float* stage_ms, unsigned int* counters);

/* View pass i+1's forward chain (preprocess, sorts, forward blend) runs on a second internal stream
 * underneath view pass i's backward chain (default on; same results: the two chains share no writable
 * state).  0 = every kernel on the trainer's stream, one at a time. */
int gsb_trainer_set_overlap(gsb_trainer* t, int on);

/* Measurement: with timing on, every view pass of gsb_trainer_accumulate / gsb_trainer_train is
 * bracketed stage by stage with CUDA events on the trainer's stream; get_stage_times returns the
 * milliseconds summed per stage (order of gsb_trainer_profile_pass) over `passes` view-pass
 * launches since timing was switched on (a fused white+black camera pass counts once). */
int gsb_trainer_set_stage_timing(gsb_trainer* t, int on);
int gsb_trainer_get_stage_times(gsb_trainer* t, double* stage_ms, long long* passes);
/* Measurement: (pixel, splat) pair counts of local camera `camera_index` under the reference's
 * control flow: out8 = { E_eval_f, E_exp_f, E_blend_f, E_eval_b, E_exp_b, E_blend_b, R, visible }
 * (forward.cu:310-357, backward.cu:476-503).  Needs one prior gsb_trainer_accumulate. */
int gsb_trainer_count_pairs(gsb_trainer* t, int camera_index, unsigned long long* out8);

/* Measurement: FP32 issue peak of the current device from an FFMA-only kernel (8 independent chains per thread,
 * 8 x 256 threads per SM): out2[0] = scalar FFMA TFLOP/s, out2[1] = packed fma.rn.f32x2 TFLOP/s.  The measured
 * denominator of the render kernels' roofline (MEASURED_PEAKS.json has no FP32 SIMT figure). */
int gsb_measure_fp32_peak(double* out2, void* stream);

/* Measurement (A/B): time of the shipped fused render backward (out_ms2[0]) and of a skeleton of
 * the "lanes = list entries" mapping (out_ms2[1], a lower bound: no forward checkpoints, results not gradients) on the
 * state of local camera `camera_index`, best of `reps`.  Needs one prior gsb_trainer_accumulate. */
int gsb_trainer_measure_bwd_skeleton(gsb_trainer* t, int camera_index, int reps, float* out_ms2);
/* Measurement (A/B): staging of the render kernels' 48-byte records over the tile lists of one
 * training camera, out_ms3 = {16-byte cp.async (what ships), TMA row gather (cp.async.bulk.tensor.2d tile::gather4)
 * issued by eight lanes, the same issued by one lane}; -1 for a mode whose copies did not complete. */
int gsb_trainer_measure_staging(gsb_trainer* t, int camera_index, int reps, float* out_ms3);
/* Measurement: statistics of the render backward's per-(8x8 block, list entry) reductions of one training camera --
 * out40[1..32] histogram of the lanes holding a blended pixel, [33] reductions through the shared-memory panel, [34] / [35] /
 * [36] of those with the blended pixels inside one half of the block's rows / columns / either, [37] blended pairs,
 * [38] (block, entry) pairs the bounding-box test lets through. */
int gsb_trainer_reduction_stats(gsb_trainer* t, int camera_index, unsigned long long* out40);

/* Testing: preset the look-back epoch counter (30-bit tags in the chained-scan status words) so the wrap
 * path can be exercised without running 2^30 launches. */
int gsb_trainer_debug_set_epoch(gsb_trainer* t, unsigned int epoch);

/* Image helpers: imageFloatToInt / imageIntToLoss (T:19-44). */
int gsb_image_float_to_int(const float* dev_src, uint32_t* dev_fb, int w, int h, void* stream);
int gsb_image_int_to_loss(const uint32_t* dev_truth, const float* dev_rasterized, float* dev_loss,
                          int w, int h, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSB_H_INCLUDED */
