// Trainer-side kernels: image conversions, the SGD update and the GPU densify / split / clone /
// cull.  Replaces imageFloatToInt, imageIntToLoss, applyGradients (reference
// src/Trainer.cu:19-44,81-101) and the CPU densify block (src/Trainer.cu:437-542).

#include "trainer_kernels.cuh"

namespace {

__device__ __forceinline__ unsigned long long lb_pack(uint32_t epoch, uint32_t flag, uint32_t value) {
    return ((unsigned long long)((epoch << 2) | flag) << 32) | value;
}

__global__ void __launch_bounds__(256) image_float_to_int_kernel(const float* __restrict__ src, uint32_t* __restrict__ fb,
                                                                 int n) {
    // int(c * 256) clamped to [0, 255], alpha = 0xFF (src/Trainer.cu:24-27)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int r = min(255, max(0, (int)(src[i] * 256.0f)));
        const int g = min(255, max(0, (int)(src[i + n] * 256.0f)));
        const int b = min(255, max(0, (int)(src[i + 2 * (size_t)n] * 256.0f)));
        fb[i] = (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | 0xFF000000u;
    }
}

__global__ void __launch_bounds__(256) image_int_to_loss_kernel(const uint32_t* __restrict__ truth,
                                                                const float* __restrict__ rasterized,
                                                                float* __restrict__ loss, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t t = truth[i];
        loss[i] = (float)(t & 0xFF) / 255.0f - rasterized[i];
        loss[i + n] = (float)((t >> 8) & 0xFF) / 255.0f - rasterized[i + n];
        loss[i + 2 * (size_t)n] = (float)((t >> 16) & 0xFF) / 255.0f - rasterized[i + 2 * (size_t)n];
    }
}

// theta += lr_theta * mean gradient, clamps as the reference (src/Trainer.cu:85-99).  One thread
// per parameter float so every access is coalesced; `avg` uses the trainer's packed layout.
__global__ void __launch_bounds__(256) apply_gradients_kernel(ApplyArgs a) {
    const size_t P = (size_t)a.count;
    const size_t nsh = 3 * (size_t)a.M * P;
    const size_t total = P * 11 + nsh;
    const float* avg_loc = a.avg + P;
    const float* avg_sh = a.avg + 4 * P;
    const float* avg_scale = avg_sh + nsh;
    const float* avg_op = avg_scale + 3 * P;
    const float* avg_rot = avg_op + P;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        if (e < 3 * P) {
            a.locations[e] = fmaf(avg_loc[e], a.lrLocation, a.locations[e]);
        } else if (e < 3 * P + nsh) {
            const size_t i = e - 3 * P;
            a.shs[i] = fmaf(avg_sh[i], a.lrSh, a.shs[i]);
        } else if (e < 6 * P + nsh) {
            const size_t i = e - 3 * P - nsh;
            a.scales[i] = fminf(a.maxScale, fmaxf(0.0f, fmaf(avg_scale[i], a.lrScale, a.scales[i])));
        } else if (e < 7 * P + nsh) {
            const size_t i = e - 6 * P - nsh;
            a.opacities[i] = fminf(1.0f, fmaxf(0.0f, fmaf(avg_op[i], a.lrOpacity, a.opacities[i])));
        } else {
            const size_t i = e - 7 * P - nsh;
            a.rotations[i] = fmaf(avg_rot[i], a.lrRotation, a.rotations[i]);
        }
    }
}

// ---- densify -------------------------------------------------------------------------------
// classify (src/Trainer.cu:448-456) + three exclusive scans in one chained-scan kernel.
constexpr int DS_IPT = 4;
constexpr int DS_TILE = 256 * DS_IPT;

__device__ __forceinline__ uint32_t chained_excl(unsigned long long* lb, uint32_t slot, uint32_t tile, uint32_t epoch,
                                                 uint32_t total) {
    unsigned long long* me = lb + (size_t)tile * 4 + slot;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(me), "l"(lb_pack(epoch, tile == 0 ? 2u : 1u, total)) : "memory");
    uint32_t excl = 0;
    for (int t = (int)tile - 1; t >= 0; --t) {
        unsigned long long v;
        uint32_t tag;
        do {
            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(lb + (size_t)t * 4 + slot) : "memory");
            tag = (uint32_t)(v >> 32);
        } while ((tag >> 2) != epoch || (tag & 3u) == 0u);
        excl += (uint32_t)v;
        if ((tag & 3u) == 2u) break;
    }
    if (tile != 0)
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(me), "l"(lb_pack(epoch, 2u, excl + total)) : "memory");
    return excl;
}

__global__ void __launch_bounds__(256) densify_classify_scan_kernel(DensifyArgs a) {
    __shared__ uint32_t s_w[3][8];
    __shared__ uint32_t s_excl[3];
    __shared__ uint32_t s_tile;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(a.ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t P = (uint32_t)a.count;
    if (tile * DS_TILE >= P) return;
    const uint32_t base = tile * DS_TILE + tid * DS_IPT;
    const float* var = a.avg;
    const float* gl = a.avg + (size_t)P;
    uint8_t f[DS_IPT];
    uint32_t c[3] = {0, 0, 0};   // keep, split, clone
#pragma unroll
    for (int k = 0; k < DS_IPT; k++) {
        const uint32_t i = base + k;
        f[k] = DENS_NONE;
        if (i < P) {
            const float sx = a.scales[3 * i], sy = a.scales[3 * i + 1], sz = a.scales[3 * i + 2];
            const float mag = sqrtf(sx * sx + sy * sy + sz * sz);
            if (a.opacities[i] <= a.cullOpacity || mag < a.cullSize) {
                f[k] = DENS_CULL;
            } else {
                const float gx = gl[3 * i], gy = gl[3 * i + 1], gz = gl[3 * i + 2];
                f[k] = DENS_KEEP;
                if (var[i] - sqrtf(gx * gx + gy * gy + gz * gz) > a.densifyVariance)
                    f[k] = (mag > a.splitSize) ? DENS_SPLIT : DENS_CLONE;
            }
            a.flags[i] = f[k];
        }
        c[0] += (f[k] >= DENS_KEEP);
        c[1] += (f[k] == DENS_SPLIT);
        c[2] += (f[k] == DENS_CLONE);
    }
    uint32_t inc[3], tot[3], wp[3];
#pragma unroll
    for (int q = 0; q < 3; q++) {
        inc[q] = c[q];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc[q], o);
            if (lane >= o) inc[q] += t;
        }
        if (lane == 31) s_w[q][warp] = inc[q];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 3; q++) {
        wp[q] = 0; tot[q] = 0;
        for (int w = 0; w < 8; w++) {
            if (w < warp) wp[q] += s_w[q][w];
            tot[q] += s_w[q][w];
        }
    }
    if (tid < 3) s_excl[tid] = chained_excl(a.lookback, tid, tile, a.epoch, tid == 0 ? tot[0] : (tid == 1 ? tot[1] : tot[2]));
    __syncthreads();
    uint32_t run[3];
#pragma unroll
    for (int q = 0; q < 3; q++) run[q] = s_excl[q] + wp[q] + inc[q] - c[q];
#pragma unroll
    for (int k = 0; k < DS_IPT; k++) {
        const uint32_t i = base + k;
        if (i < P) {
            a.keep_idx[i] = run[0];
            a.aux_idx[i] = (f[k] == DENS_SPLIT) ? run[1] : run[2];
        }
        run[0] += (f[k] >= DENS_KEEP);
        run[1] += (f[k] == DENS_SPLIT);
        run[2] += (f[k] == DENS_CLONE);
    }
    if ((tile + 1) * DS_TILE >= P && tid == 255) {
        a.totals[0] = run[0]; a.totals[1] = run[1]; a.totals[2] = run[2];
    }
}

// R(q) column-major exactly as glm::mat3_cast (quaternion stored w,x,y,z, not normalised):
// returns R(q) * v.
__device__ __forceinline__ float3 quat_rotate(const float* q, float3 v) {
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    const float c00 = 1.f - 2.f * (y * y + z * z), c01 = 2.f * (x * y + w * z), c02 = 2.f * (x * z - w * y);
    const float c10 = 2.f * (x * y - w * z), c11 = 1.f - 2.f * (x * x + z * z), c12 = 2.f * (y * z + w * x);
    const float c20 = 2.f * (x * z + w * y), c21 = 2.f * (y * z - w * x), c22 = 1.f - 2.f * (x * x + y * y);
    return make_float3(c00 * v.x + c10 * v.y + c20 * v.z, c01 * v.x + c11 * v.y + c21 * v.z, c02 * v.x + c12 * v.y + c22 * v.z);
}

// Canonical output order: survivors in index order, then split twins (ascending source index),
// then clones (ascending source index).  The reference appends in unordered_set iteration order
// and compacts afterwards; as a multiset of splats the result is the same.
__global__ void __launch_bounds__(256) densify_scatter_kernel(DensifyArgs a, DensifyOut o) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (uint32_t)a.count) return;
    const uint8_t f = a.flags[i];
    if (f == DENS_CULL || f == DENS_NONE) return;
    const int M3 = 3 * a.M;
    const uint32_t dst = a.keep_idx[i];
    float loc[3] = {a.locations[3 * i], a.locations[3 * i + 1], a.locations[3 * i + 2]};
    float sc[3] = {a.scales[3 * i], a.scales[3 * i + 1], a.scales[3 * i + 2]};
    const float q[4] = {a.rotations[4 * i], a.rotations[4 * i + 1], a.rotations[4 * i + 2], a.rotations[4 * i + 3]};
    const float op = a.opacities[i];

    int twin = -1;
    float tloc[3] = {loc[0], loc[1], loc[2]};
    if (f == DENS_SPLIT && a.aux_idx[i] < o.n_split_done) {
        // largest scale axis (src/Trainer.cu:467-474), offset = R(q) (s_k e_k)
        float3 axis = make_float3(0.f, 0.f, 0.f);
        if (sc[0] > sc[1] && sc[0] > sc[2]) axis.x = sc[0];
        else if (sc[1] > sc[2]) axis.y = sc[1];
        else axis.z = sc[2];
        const float3 v = quat_rotate(q, axis);
        const float h = a.splitDistance * 0.5f;
        tloc[0] = loc[0] - v.x * h; tloc[1] = loc[1] - v.y * h; tloc[2] = loc[2] - v.z * h;
        loc[0] = loc[0] + v.x * h; loc[1] = loc[1] + v.y * h; loc[2] = loc[2] + v.z * h;
        sc[0] *= a.splitScale; sc[1] *= a.splitScale; sc[2] *= a.splitScale;
        twin = (int)(o.n_keep + a.aux_idx[i]);
    } else if (f == DENS_CLONE && a.aux_idx[i] < o.n_clone_done) {
        // clone offset = (R(q) s) * normalize(mean location gradient) * cloneDistance, component-wise
        const float* gl = a.avg + (size_t)a.count + 3 * (size_t)i;
        const float inv = rsqrtf(gl[0] * gl[0] + gl[1] * gl[1] + gl[2] * gl[2]);   // inf/NaN for a zero gradient, as glm::normalize
        const float3 v = quat_rotate(q, make_float3(sc[0], sc[1], sc[2]));
        tloc[0] = loc[0] + v.x * (gl[0] * inv) * a.cloneDistance;
        tloc[1] = loc[1] + v.y * (gl[1] * inv) * a.cloneDistance;
        tloc[2] = loc[2] + v.z * (gl[2] * inv) * a.cloneDistance;
        twin = (int)(o.n_keep + o.n_split_done + a.aux_idx[i]);
    }
    for (int pass = 0; pass < 2; pass++) {
        const int d = pass == 0 ? (int)dst : twin;
        if (d < 0) break;
        const float* L = pass == 0 ? loc : tloc;
        o.locations[3 * d] = L[0]; o.locations[3 * d + 1] = L[1]; o.locations[3 * d + 2] = L[2];
        o.scales[3 * d] = sc[0]; o.scales[3 * d + 1] = sc[1]; o.scales[3 * d + 2] = sc[2];
        o.opacities[d] = op;
        o.rotations[4 * d] = q[0]; o.rotations[4 * d + 1] = q[1]; o.rotations[4 * d + 2] = q[2]; o.rotations[4 * d + 3] = q[3];
        for (int k = 0; k < M3; k++) o.shs[(size_t)d * M3 + k] = a.shs[(size_t)i * M3 + k];
    }
}

}  // namespace

void launch_image_float_to_int(const float* src, uint32_t* fb, int w, int h, cudaStream_t s) {
    const int n = w * h;
    if (n <= 0) return;
    image_float_to_int_kernel<<<min(gsb_div_up(n, 256), 148 * 8), 256, 0, s>>>(src, fb, n);
}

void launch_image_int_to_loss(const uint32_t* truth, const float* rasterized, float* loss, int w, int h, cudaStream_t s) {
    const int n = w * h;
    if (n <= 0) return;
    image_int_to_loss_kernel<<<min(gsb_div_up(n, 256), 148 * 8), 256, 0, s>>>(truth, rasterized, loss, n);
}

void launch_apply_gradients(const ApplyArgs& a, cudaStream_t s) {
    if (a.count <= 0) return;
    const size_t total = (size_t)a.count * (11 + 3 * (size_t)a.M);
    apply_gradients_kernel<<<(int)min((size_t)148 * 16, (total + 255) / 256), 256, 0, s>>>(a);
}

void launch_densify_classify_scan(const DensifyArgs& a, cudaStream_t s) {
    if (a.count <= 0) return;
    densify_classify_scan_kernel<<<gsb_div_up(a.count, DS_TILE), 256, 0, s>>>(a);
}

void launch_densify_scatter(const DensifyArgs& a, const DensifyOut& o, cudaStream_t s) {
    if (a.count <= 0) return;
    densify_scatter_kernel<<<gsb_div_up(a.count, 256), 256, 0, s>>>(a, o);
}
