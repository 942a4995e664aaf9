// Launchers of the trainer-side kernels (trainer_kernels.cu).
#pragma once
#include "common.cuh"

enum : uint8_t { DENS_NONE = 0, DENS_CULL = 1, DENS_KEEP = 2, DENS_SPLIT = 3, DENS_CLONE = 4 };

struct ApplyArgs {
    float *locations, *shs, *scales, *opacities, *rotations;   // ModelSplatsDevice arrays
    const float* avg;                                           // packed running means (see PreprocessBwdArgs)
    float lrLocation, lrSh, lrScale, lrOpacity, lrRotation, maxScale;
    int M, count;
};

struct DensifyArgs {
    const float *locations, *shs, *scales, *opacities, *rotations;
    const float* avg;               // var at [0,P), mean location gradient at [P,4P)
    int M, count;
    float cullOpacity, cullSize, densifyVariance, splitSize, splitDistance, splitScale, cloneDistance;
    uint8_t* flags;                 // [count]
    uint32_t* keep_idx;             // [count] destination of a survivor
    uint32_t* aux_idx;              // [count] rank among splits (or among clones)
    uint32_t* totals;               // [3] survivors, splits, clones
    unsigned long long* lookback;   // [tiles][4]
    uint32_t epoch;
    uint32_t* ticket;
};

struct DensifyOut {
    float *locations, *shs, *scales, *opacities, *rotations;    // new arrays, sized for the new count
    uint32_t n_keep, n_split_done, n_clone_done;
};

void launch_image_float_to_int(const float* src, uint32_t* fb, int w, int h, cudaStream_t s);
void launch_image_int_to_loss(const uint32_t* truth, const float* rasterized, float* loss, int w, int h, cudaStream_t s);
void launch_apply_gradients(const ApplyArgs& a, cudaStream_t s);
void launch_densify_classify_scan(const DensifyArgs& a, cudaStream_t s);
void launch_densify_scatter(const DensifyArgs& a, const DensifyOut& o, cudaStream_t s);
