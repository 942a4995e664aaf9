// Forward preprocess: per-splat frustum cull, 3D covariance, EWA projection to a 2D conic,
// screen-space radius / tile rectangle, SH -> RGB.
// Replaces preprocessCUDA<3> (reference cuda_rasterizer/forward.cu:155-256) and its helpers
// computeCov3D (:118-152), computeCov2D (:74-113), computeColorFromSH (:20-71), in_frustum /
// getRect / ndc2Pix (auxiliary.h:41-56,139-164), plus checkFrustum (rasterizer_impl.cu:54-66).
//
// Everything that decides an integer downstream (depth bits in the sort key, radius, tile
// rectangle, tiles_touched) is written with explicit round-to-nearest intrinsics in the exact
// operation order the reference build executes (mul / fma pairing taken from the sm_100 SASS of
// the reference's preprocessCUDA, see DESIGN.md "bit-exact discretisation"), so the compiler can
// neither contract nor re-associate it and keys / ranges / counts match bit for bit.

#include "common.cuh"
#include <cstdio>

namespace {

__device__ __forceinline__ float mul_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// ((m_a*x + m_b*y) + m_c*z) + m_d with the reference's pairing: fma(z, m_c, fma(x, m_a, y*m_b)) + m_d
// (transformPoint4x3 / transformPoint4x4, auxiliary.h:58-78).
__device__ __forceinline__ float affine_row(float x, float y, float z, float ma, float mb, float mc, float md) {
    float r = mul_(y, mb);
    r = fma_(x, ma, r);
    r = fma_(z, mc, r);
    return add_(r, md);
}

// 3-term dot in the reference's order: fma(c0, c1, fma(a0, a1, b0*b1))
__device__ __forceinline__ float dot3_bac(float a0, float a1, float b0, float b1, float c0, float c1) {
    return fma_(c0, c1, fma_(a0, a1, mul_(b0, b1)));
}

// Sigma = (S R)^T (S R) for an UNNORMALISED quaternion stored (w,x,y,z); upper triangle out.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx_, float sy_, float sz_, float mod, float r,
                                                     float x, float y, float z, float* c) {
    const float xz = mul_(x, z);
    const float rx = mul_(r, x);
    const float A = fma_(r, y, xz);          // xz + ry
    const float B = fma_(-r, y, xz);         // xz - ry
    const float rz = mul_(r, z);
    const float Cm = fma_(y, z, -rx);        // yz - rx
    const float yy = mul_(y, y);
    const float Dp = fma_(y, z, rx);         // yz + rx
    const float zz = mul_(z, z);
    const float E = fma_(x, y, -rz);         // xy - rz
    const float F = fma_(x, y, rz);          // xy + rz
    const float xxyy = fma_(x, x, yy);
    const float yyzz = add_(yy, zz);
    const float xxzz = fma_(x, x, zz);
    const float R00 = add_(-add_(yyzz, yyzz), 1.0f);
    const float R11 = add_(-add_(xxzz, xxzz), 1.0f);
    const float R22 = add_(-add_(xxyy, xxyy), 1.0f);
    const float sx = mul_(sx_, mod), sy = mul_(sy_, mod), sz = mul_(sz_, mod);
    // M = S * R (columns of glm's R literal; only one product survives per entry)
    const float m00 = mul_(sx, R00), m01 = mul_(sx, add_(F, F)), m02 = mul_(sx, add_(B, B));
    const float m10 = mul_(sy, add_(E, E)), m11 = mul_(sy, R11), m12 = mul_(sy, add_(Dp, Dp));
    const float m20 = mul_(sz, add_(A, A)), m21 = mul_(sz, add_(Cm, Cm)), m22 = mul_(sz, R22);
    c[0] = dot3_bac(m00, m00, m10, m10, m20, m20);
    c[1] = dot3_bac(m00, m01, m10, m11, m20, m21);
    c[2] = dot3_bac(m00, m02, m10, m12, m20, m22);
    c[3] = dot3_bac(m01, m01, m11, m11, m21, m21);
    c[4] = dot3_bac(m01, m02, m11, m12, m21, m22);
    c[5] = dot3_bac(m02, m02, m12, m12, m22, m22);
}

__device__ const float kSH_C0 = 0.28209479177387814f;
__device__ const float kSH_C1 = 0.4886025119029199f;
__device__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                    -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                    0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                    -0.5900435899266435f};

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 ld3(const float* p) { return {p[0], p[1], p[2]}; }
__device__ __forceinline__ F3 operator*(float s, F3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

// SH -> RGB, degree <= 3, +0.5, clamp at 0 with per-channel flags (forward.cu:20-71).
__device__ F3 sh_to_rgb(int deg, int M, float px, float py, float pz, const float* cam, const float* sh, uint8_t& flags) {
    float dx = px - cam[0], dy = py - cam[1], dz = pz - cam[2];
    float len = sqrtf(fma_(dz, dz, fma_(dx, dx, mul_(dy, dy))));
    float x = dx / len, y = dy / len, z = dz / len;
    F3 res = kSH_C0 * ld3(sh);
    if (deg > 0) {
        res = res - (kSH_C1 * y) * ld3(sh + 3) + (kSH_C1 * z) * ld3(sh + 6) - (kSH_C1 * x) * ld3(sh + 9);
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            res = res + (kSH_C2[0] * xy) * ld3(sh + 12) + (kSH_C2[1] * yz) * ld3(sh + 15) +
                  (kSH_C2[2] * (2.0f * zz - xx - yy)) * ld3(sh + 18) + (kSH_C2[3] * xz) * ld3(sh + 21) +
                  (kSH_C2[4] * (xx - yy)) * ld3(sh + 24);
            if (deg > 2) {
                res = res + (kSH_C3[0] * y * (3.0f * xx - yy)) * ld3(sh + 27) + (kSH_C3[1] * xy * z) * ld3(sh + 30) +
                      (kSH_C3[2] * y * (4.0f * zz - xx - yy)) * ld3(sh + 33) +
                      (kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy)) * ld3(sh + 36) +
                      (kSH_C3[4] * x * (4.0f * zz - xx - yy)) * ld3(sh + 39) +
                      (kSH_C3[5] * z * (xx - yy)) * ld3(sh + 42) + (kSH_C3[6] * x * (xx - 3.0f * yy)) * ld3(sh + 45);
            }
        }
    }
    res.x += 0.5f; res.y += 0.5f; res.z += 0.5f;
    flags = (uint8_t)((res.x < 0.f ? 1 : 0) | (res.y < 0.f ? 2 : 0) | (res.z < 0.f ? 4 : 0));
    return {fmaxf(res.x, 0.f), fmaxf(res.y, 0.f), fmaxf(res.z, 0.f)};
}

// The same for the trainer's record shape (M = 4: degree <= 1, twelve floats = three 16-byte words per splat, loaded as
// such): identical expression tree, operands from registers.
__device__ __forceinline__ F3 sh_to_rgb_m4(int deg, float px, float py, float pz, const float* cam, const float4 s0, const float4 s1,
                                           const float4 s2, uint8_t& flags) {
    float dx = px - cam[0], dy = py - cam[1], dz = pz - cam[2];
    float len = sqrtf(fma_(dz, dz, fma_(dx, dx, mul_(dy, dy))));
    float x = dx / len, y = dy / len, z = dz / len;
    F3 res = kSH_C0 * F3{s0.x, s0.y, s0.z};
    if (deg > 0) res = res - (kSH_C1 * y) * F3{s0.w, s1.x, s1.y} + (kSH_C1 * z) * F3{s1.z, s1.w, s2.x} - (kSH_C1 * x) * F3{s2.y, s2.z, s2.w};
    res.x += 0.5f; res.y += 0.5f; res.z += 0.5f;
    flags = (uint8_t)((res.x < 0.f ? 1 : 0) | (res.y < 0.f ? 2 : 0) | (res.z < 0.f ? 4 : 0));
    return {fmaxf(res.x, 0.f), fmaxf(res.y, 0.f), fmaxf(res.z, 0.f)};
}

// VEC: rotations, the 12-float SH record and the covariance are moved as 16- / 8-byte words (the arrays are contiguous
// per splat and the caller's bases are suitably aligned: checked by the launcher); means and scales (stride 3) stay scalar.
#ifndef GSB_PP_THREADS
#define GSB_PP_THREADS 256   // threads per block (A/B knob)
#endif
template <bool VEC>
__global__ void __launch_bounds__(GSB_PP_THREADS) preprocess_fwd_kernel(const PreprocessArgs a) {
    // every lane stays to the end (the tile counts are added warp-cooperatively); lanes beyond P shadow splat P-1 and
    // write nothing
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = gid < a.P;
    const int idx = valid ? gid : a.P - 1;

    // Culled unless proven visible (forward.cu:189-190).
    int radius = 0;
    uint32_t tiles = 0;

    const float px = a.means3D[3 * idx], py = a.means3D[3 * idx + 1], pz = a.means3D[3 * idx + 2];
    const float* V = a.viewmatrix;
    const float* Pm = a.projmatrix;

    // p_view.z, the sort depth (in_frustum, auxiliary.h:152).  Only a z cull; NaN passes.
    const float depth = affine_row(px, py, pz, __ldg(V + 2), __ldg(V + 6), __ldg(V + 10), __ldg(V + 14));
    bool alive = valid && !(depth <= 0.2f);
    if (valid && !alive && a.prefiltered) {
        printf("Point is filtered although prefiltered is set. This shouldn't happen!");
        __trap();
    }

    float cov[6];
    float ix = 0.f, iy = 0.f, conx = 0.f, cony = 0.f, conz = 0.f, covx = 0.f, covy = 0.f;
    bool pos_def = true;
    uint32_t rect_x0 = 0, rect_y0 = 0, rect_w = 0;
    if (alive) {
        const float hx = affine_row(px, py, pz, __ldg(Pm + 0), __ldg(Pm + 4), __ldg(Pm + 8), __ldg(Pm + 12));
        const float hy = affine_row(px, py, pz, __ldg(Pm + 1), __ldg(Pm + 5), __ldg(Pm + 9), __ldg(Pm + 13));
        const float hw = affine_row(px, py, pz, __ldg(Pm + 3), __ldg(Pm + 7), __ldg(Pm + 11), __ldg(Pm + 15));
        const float pw = __frcp_rn(add_(hw, 0.0000001f));
        const float ndcx = mul_(hx, pw), ndcy = mul_(hy, pw);

        if (a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov[k] = a.cov3D_precomp[6 * idx + k];
        } else {
            if constexpr (VEC) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
                cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.scale_modifier,
                                     q.x, q.y, q.z, q.w, cov);
                float2* c2 = reinterpret_cast<float2*>(a.cov3D + 6 * (size_t)idx);
                c2[0] = make_float2(cov[0], cov[1]); c2[1] = make_float2(cov[2], cov[3]); c2[2] = make_float2(cov[4], cov[5]);
            } else {
                cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.scale_modifier,
                                     a.rotations[4 * idx], a.rotations[4 * idx + 1], a.rotations[4 * idx + 2],
                                     a.rotations[4 * idx + 3], cov);
#pragma unroll
                for (int k = 0; k < 6; k++) a.cov3D[6 * idx + k] = cov[k];
            }
        }

        // EWA projection (computeCov2D).  t = view * p; x/z and y/z clamped to 1.3 tan(fov/2).
        const float v0 = __ldg(V + 0), v1 = __ldg(V + 1), v2 = __ldg(V + 2);
        const float v4 = __ldg(V + 4), v5 = __ldg(V + 5), v6 = __ldg(V + 6);
        const float v8 = __ldg(V + 8), v9 = __ldg(V + 9), v10 = __ldg(V + 10);
        const float tz = depth;
        const float tx = affine_row(px, py, pz, v0, v4, v8, __ldg(V + 12));
        const float ty = affine_row(px, py, pz, v1, v5, v9, __ldg(V + 13));
        const float limx = mul_(a.tan_fovx, 1.3f), limy = mul_(a.tan_fovy, 1.3f);
        const float clx = fminf(fmaxf(__fdiv_rn(tx, tz), -limx), limx);
        const float cly = fminf(fmaxf(__fdiv_rn(ty, tz), -limy), limy);
        const float tz2 = mul_(tz, tz);
        const float J00 = __fdiv_rn(a.focal_x, tz);
        const float J02 = __fdiv_rn(mul_(mul_(tz, -clx), a.focal_x), tz2);
        const float J11 = __fdiv_rn(a.focal_y, tz);
        const float J12 = __fdiv_rn(mul_(mul_(tz, -cly), a.focal_y), tz2);
        // T = W * J: two non-trivial rows
        const float a0 = fma_(v2, J02, mul_(v0, J00)), a1 = fma_(v6, J02, mul_(v4, J00)), a2 = fma_(v10, J02, mul_(v8, J00));
        const float b0 = fma_(v2, J12, mul_(v1, J11)), b1 = fma_(v6, J12, mul_(v5, J11)), b2 = fma_(v10, J12, mul_(v9, J11));
        // Vrk * rows
        const float ua0 = dot3_bac(a0, cov[0], a1, cov[1], a2, cov[2]);
        const float ua1 = dot3_bac(a0, cov[1], a1, cov[3], a2, cov[4]);
        const float ua2 = dot3_bac(a0, cov[2], a1, cov[4], a2, cov[5]);
        const float ub0 = dot3_bac(b0, cov[0], b1, cov[1], b2, cov[2]);
        const float ub1 = dot3_bac(b0, cov[1], b1, cov[3], b2, cov[4]);
        const float ub2 = dot3_bac(b0, cov[2], b1, cov[4], b2, cov[5]);
        const float cxx = add_(dot3_bac(a0, ua0, a1, ua1, a2, ua2), 0.3f);   // + low-pass
        const float cxy = dot3_bac(a0, ub0, a1, ub1, a2, ub2);
        const float cyy = add_(dot3_bac(b0, ub0, b1, ub1, b2, ub2), 0.3f);

        const float det = fma_(cxx, cyy, -mul_(cxy, cxy));
        alive = (det != 0.0f);
        if (alive) {
            const float det_inv = __frcp_rn(det);
            conx = mul_(cyy, det_inv);
            cony = mul_(cxy, -det_inv);
            conz = mul_(cxx, det_inv);
            covx = cxx; covy = cyy;
            pos_def = det > 0.0f && cxx > 0.0f && cyy > 0.0f;
            const float mid = mul_(add_(cxx, cyy), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(fma_(mid, mid, -det), 0.1f));
            const float lam = fmaxf(add_(mid, sq), sub_(mid, sq));
            const int r = (int)ceilf(mul_(__fsqrt_rn(lam), 3.0f));
            // ndc2Pix is evaluated in double (auxiliary.h:41-44)
            ix = __double2float_rn(__dmul_rn(__fma_rn(__dadd_rn((double)ndcx, 1.0), (double)a.W, -1.0), 0.5));
            iy = __double2float_rn(__dmul_rn(__fma_rn(__dadd_rn((double)ndcy, 1.0), (double)a.H, -1.0), 0.5));
            // getRect (auxiliary.h:46-56)
            const float rf = (float)r;
            const int x0 = (int)mul_(sub_(ix, rf), 0.0625f), y0 = (int)mul_(sub_(iy, rf), 0.0625f);
            const int x1 = (int)mul_(add_(add_(add_(ix, rf), 16.0f), -1.0f), 0.0625f);
            const int y1 = (int)mul_(add_(add_(add_(iy, rf), 16.0f), -1.0f), 0.0625f);
            const uint32_t gx = (uint32_t)a.grid_x, gy = (uint32_t)a.grid_y;
            const uint32_t rx0 = min(gx, (uint32_t)max(0, x0)), ry0 = min(gy, (uint32_t)max(0, y0));
            const uint32_t rx1 = min(gx, (uint32_t)max(0, x1)), ry1 = min(gy, (uint32_t)max(0, y1));
            const uint32_t t = (rx1 - rx0) * (ry1 - ry0);
            alive = (t != 0);
            if (alive) { radius = r; tiles = t; rect_x0 = rx0; rect_y0 = ry0; rect_w = rx1 - rx0; }
        }
    }

    if (valid) {
        a.radii[idx] = radius;
        if (a.radii_user) a.radii_user[idx] = radius;
        a.tiles_touched[idx] = tiles;
    }
    {
        const unsigned vis = __ballot_sync(0xffffffffu, alive);
        if (vis && (threadIdx.x & 31) == 0) atomicAdd(&a.ctrl->num_visible, (uint32_t)__popc(vis));
        // first half of the binning (binning.cu): how many splats cover each tile, one RED per (splat, tile) pair
        if (vis) {
            const uint32_t gxw = (uint32_t)a.grid_x;
            uint32_t* const tc = a.tile_count;
            gsb_warp_walk_tiles(rect_x0, rect_y0, rect_w, tiles, [&](const bool* on, const int*, const uint32_t* tx, const uint32_t* ty) {
#pragma unroll
                for (int u = 0; u < 2; u++)
                    if (on[u]) atomicAdd(&tc[ty[u] * gxw + tx[u]], 1u);
            });
        }
    }
    if (!alive) return;

    F3 rgb;
    uint8_t flags = 0;
    if (a.colors_precomp == nullptr) {
        if constexpr (VEC) {
            const float4* sp = reinterpret_cast<const float4*>(a.shs) + 3 * (size_t)idx;
            rgb = sh_to_rgb_m4(a.D, px, py, pz, a.cam_pos, __ldg(sp), __ldg(sp + 1), __ldg(sp + 2), flags);
        } else {
            rgb = sh_to_rgb(a.D, a.M, px, py, pz, a.cam_pos, a.shs + (size_t)idx * a.M * 3, flags);
        }
    } else {
        rgb = ld3(a.colors_precomp + 3 * (size_t)idx);
    }
    a.clamped[idx] = flags;
    a.depths[idx] = depth;
    // Conservative culling data for the render kernels.  A pair can only blend if
    // opacity * exp(power) >= 1/255, i.e. power >= -log(255 * opacity); `lim` sits 0.05 below that
    // (far more than the rounding error of power and expf), and (ex, ey) is the half-extent of the
    // bounding box of { power >= lim - 0.05 }, inflated.  Neither value enters any blended result.
    const float opac = a.opacities[idx];
    float lim = __int_as_float(0x7f800000), ex = -1e30f, ey = -1e30f;   // opacity < 1/255: never blends
    if (!(opac < 1.0f / 255.0f)) {
        lim = -logf(opac * 255.0f) - 0.05f;
        const float k = -2.0f * (lim - 0.05f);
        ex = sqrtf(fmaxf(k * covx, 0.f)) * 1.001f + 0.01f;
        ey = sqrtf(fmaxf(k * covy, 0.f)) * 1.001f + 0.01f;
        // NaN covariance, or a conic that is not positive definite (a caller's cov3D_precomp can produce det < 0:
        // { power >= lim } is then unbounded): never cull by the box, only by the exact threshold on power
        if (!(ex == ex) || !(ey == ey) || !pos_def) { ex = 1e30f; ey = 1e30f; }
    }
    SplatRec rec;
    rec.a = make_float4(ix, iy, conx, cony);
    rec.b = make_float4(conz, opac, rgb.x, rgb.y);
    rec.c = make_float4(rgb.z, lim, ex, ey);
    a.recs[idx] = rec;
}

__global__ void mark_visible_kernel(int P, const float* means3D, const float* V, bool* present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float depth = affine_row(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2], __ldg(V + 2),
                                   __ldg(V + 6), __ldg(V + 10), __ldg(V + 14));
    present[idx] = !(depth <= 0.2f);
}

}  // namespace

void launch_preprocess_fwd(const PreprocessArgs& a, cudaStream_t s) {
    if (a.P <= 0) return;
    auto al = [](const void* p, uintptr_t n) { return reinterpret_cast<uintptr_t>(p) % n == 0; };
    const bool vec = a.cov3D_precomp == nullptr && a.colors_precomp == nullptr && a.M == 4 && a.D <= 1 && al(a.rotations, 16) &&
                     al(a.shs, 16) && al(a.cov3D, 8);
    if (vec) preprocess_fwd_kernel<true><<<gsb_div_up(a.P, GSB_PP_THREADS), GSB_PP_THREADS, 0, s>>>(a);
    else preprocess_fwd_kernel<false><<<gsb_div_up(a.P, GSB_PP_THREADS), GSB_PP_THREADS, 0, s>>>(a);
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, bool* present, cudaStream_t s) {
    if (P <= 0) return;
    mark_visible_kernel<<<gsb_div_up(P, 256), 256, 0, s>>>(P, means3D, viewmatrix, present);
}
