// Backward of the per-splat preprocess, one fused kernel:
//   conic gradient -> 2D covariance -> 3D covariance + mean   (computeCov2DCUDA,
//       reference cuda_rasterizer/backward.cu:144-274)
//   screen-space mean gradient -> 3D mean, SH backward, covariance -> scale / rotation
//       (preprocessCUDA backward :346-396, computeColorFromSH :20-139, computeCov3D :278-341)
// and, in Trainer mode, the running-mean accumulation of accumulateGradients
// (reference src/Trainer.cu:47-77) so the 23 per-view gradient floats never round-trip HBM.
// It consumes the packed GradRec scratch written by the render backward and clears it, which
// replaces the nine per-view cudaMemsets of src/Trainer.cu:366-375.
//
// Quirks kept: the quaternion is used unnormalised and its gradient is not projected
// (backward.cu:281,340); clamped SH channels get zero colour gradient; the x/z, y/z clamp zeroes
// the matching mean gradient term.

#include "common.cuh"
#include <cmath>

namespace {

__device__ const float kC0 = 0.28209479177387814f;
__device__ const float kC1 = 0.4886025119029199f;
__device__ const float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 ld3(const float* p) { return {p[0], p[1], p[2]}; }
__device__ __forceinline__ V3 operator*(float s, V3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// Derivatives of the colour w.r.t. the (normalised) view direction, plus the unnormalised direction.
struct ShDirs { V3 d0, dx, dy, dz; };

// Writes dL/dsh (M x 3) through `emit(coeff, value)` for the colour gradient dRGB and returns what the
// view-direction term of dL/dmean needs (dir_grad below); everything here is linear in dRGB.
template <typename Emit>
__device__ __forceinline__ ShDirs sh_backward(int deg, float px, float py, float pz, const float* cam, const float* sh, V3 dRGB, Emit emit) {
    const V3 d0 = {px - cam[0], py - cam[1], pz - cam[2]};
    const float len = sqrtf(d0.x * d0.x + d0.y * d0.y + d0.z * d0.z);
    const float x = d0.x / len, y = d0.y / len, z = d0.z / len;
    V3 dx = {0, 0, 0}, dy = {0, 0, 0}, dz = {0, 0, 0};
    emit(0, kC0 * dRGB);
    if (deg > 0) {
        emit(1, (-kC1 * y) * dRGB);
        emit(2, (kC1 * z) * dRGB);
        emit(3, (-kC1 * x) * dRGB);
        dx = (-kC1) * ld3(sh + 9);
        dy = (-kC1) * ld3(sh + 3);
        dz = kC1 * ld3(sh + 6);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            emit(4, (kC2[0] * xy) * dRGB);
            emit(5, (kC2[1] * yz) * dRGB);
            emit(6, (kC2[2] * (2.f * zz - xx - yy)) * dRGB);
            emit(7, (kC2[3] * xz) * dRGB);
            emit(8, (kC2[4] * (xx - yy)) * dRGB);
            const V3 s4 = ld3(sh + 12), s5 = ld3(sh + 15), s6 = ld3(sh + 18), s7 = ld3(sh + 21), s8 = ld3(sh + 24);
            dx = dx + (kC2[0] * y) * s4 + (kC2[2] * 2.f * -x) * s6 + (kC2[3] * z) * s7 + (kC2[4] * 2.f * x) * s8;
            dy = dy + (kC2[0] * x) * s4 + (kC2[1] * z) * s5 + (kC2[2] * 2.f * -y) * s6 + (kC2[4] * 2.f * -y) * s8;
            dz = dz + (kC2[1] * y) * s5 + (kC2[2] * 2.f * 2.f * z) * s6 + (kC2[3] * x) * s7;
            if (deg > 2) {
                emit(9, (kC3[0] * y * (3.f * xx - yy)) * dRGB);
                emit(10, (kC3[1] * xy * z) * dRGB);
                emit(11, (kC3[2] * y * (4.f * zz - xx - yy)) * dRGB);
                emit(12, (kC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dRGB);
                emit(13, (kC3[4] * x * (4.f * zz - xx - yy)) * dRGB);
                emit(14, (kC3[5] * z * (xx - yy)) * dRGB);
                emit(15, (kC3[6] * x * (xx - 3.f * yy)) * dRGB);
                const V3 s9 = ld3(sh + 27), s10 = ld3(sh + 30), s11 = ld3(sh + 33), s12 = ld3(sh + 36),
                         s13 = ld3(sh + 39), s14 = ld3(sh + 42), s15 = ld3(sh + 45);
                dx = dx + (kC3[0] * 3.f * 2.f * xy) * s9 + (kC3[1] * yz) * s10 + (kC3[2] * -2.f * xy) * s11 +
                     (kC3[3] * -3.f * 2.f * xz) * s12 + (kC3[4] * (-3.f * xx + 4.f * zz - yy)) * s13 +
                     (kC3[5] * 2.f * xz) * s14 + (kC3[6] * 3.f * (xx - yy)) * s15;
                dy = dy + (kC3[0] * 3.f * (xx - yy)) * s9 + (kC3[1] * xz) * s10 +
                     (kC3[2] * (-3.f * yy + 4.f * zz - xx)) * s11 + (kC3[3] * -3.f * 2.f * yz) * s12 +
                     (kC3[4] * -2.f * xy) * s13 + (kC3[5] * -2.f * yz) * s14 + (kC3[6] * -3.f * 2.f * xy) * s15;
                dz = dz + (kC3[1] * xy) * s10 + (kC3[2] * 4.f * 2.f * yz) * s11 +
                     (kC3[3] * 3.f * (2.f * zz - xx - yy)) * s12 + (kC3[4] * 4.f * 2.f * xz) * s13 +
                     (kC3[5] * (xx - yy)) * s14;
            }
        }
    }
    return {d0, dx, dy, dz};
}

// dL/dmean through dir = d0 / |d0|  (dnormvdv, auxiliary.h:108-119) for one colour gradient
__device__ __forceinline__ V3 dir_grad(const ShDirs& q, V3 dRGB) {
    const V3 d0 = q.d0;
    const V3 dd = {dot(q.dx, dRGB), dot(q.dy, dRGB), dot(q.dz, dRGB)};
    const float sum2 = d0.x * d0.x + d0.y * d0.y + d0.z * d0.z;
    const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    V3 g;
    g.x = ((sum2 - d0.x * d0.x) * dd.x - d0.y * d0.x * dd.y - d0.z * d0.x * dd.z) * inv32;
    g.y = (-d0.x * d0.y * dd.x + (sum2 - d0.y * d0.y) * dd.y - d0.z * d0.y * dd.z) * inv32;
    g.z = (-d0.x * d0.z * dd.x - d0.y * d0.z * dd.y + (sum2 - d0.z * d0.z) * dd.z) * inv32;
    return g;
}

// NS = number of GradRec sets consumed in one pass over the splat: 1 everywhere except the Trainer's
// fused white+black camera pass (NS = 2: a.grads2 holds the (white, black) pairs of both passes, GradRec2).  Every gradient
// is linear in a set's moments, so the per-splat geometry (T = W J, Vrk T, the conic Jacobian, the
// projection Jacobian, the SH direction derivatives) is computed once; only dL/dmean3D is finished
// per set, because varLocations sums the per-pass norms (src/Trainer.cu:52-54).
// M4 = the application's shape (SH degree 1, 4 coefficients, src/Config.h:19-20): the 24 running-mean
// floats of the splat are loaded up front into registers (independent loads in flight together, hidden
// behind the arithmetic) and stored once at the end, instead of 24 dependent read-modify-writes.
// GSB_PB_RED: the Trainer shape adds grad/S to the running means with fire-and-forget REDs (vector REDs for the 16-byte
// aligned SH and rotation records) instead of loading the 24 floats up front and storing them back: no load latency on
// the critical path and 24 registers fewer.  Every element is touched by exactly one thread per launch, so the sums
// (and their order across launches of one stream) are those of the load / add / store form.
#ifndef GSB_PB_RED
#define GSB_PB_RED 1
#endif
__device__ __forceinline__ void pb_red(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void pb_red4(float* p, float x, float y, float z, float w) {
    asm volatile("red.global.v4.f32.add [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
#ifndef GSB_PB_THREADS
#define GSB_PB_THREADS 128   // 95 registers: five 128-thread blocks per SM against two of 256
#define GSB_PB_BLOCKS 1
#endif
// POW2S: the sample count S is a power of two (2 x cameras is, for every preset of the reference): x / S (what
// accumulateGradients computes, src/Trainer.cu:54-75) is then bit for bit x * (1 / S), and the 27 IEEE divisions of the
// accumulate epilogue become multiplications.
template <int NS, bool M4, bool POW2S = false>
__global__ void __launch_bounds__(GSB_PB_THREADS, GSB_PB_BLOCKS) preprocess_bwd_kernel(const PreprocessBwdArgs a) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P || !(a.radii[idx] > 0)) return;
    const bool modeA = (a.avg == nullptr);
    const size_t P = (size_t)a.P;
    const int M = M4 ? 4 : a.M;
    float* avg_loc = modeA ? nullptr : a.avg + P;
    float* avg_sh = modeA ? nullptr : a.avg + 4 * P;
    float* avg_scale = modeA ? nullptr : a.avg + 4 * P + 3 * (size_t)M * P;
    float* avg_op = modeA ? nullptr : avg_scale + 3 * P;
    float* avg_rot = modeA ? nullptr : avg_op + P;
    float o_var = 0.f, o_loc[3] = {0.f, 0.f, 0.f}, o_op = 0.f, o_scale[3] = {0.f, 0.f, 0.f}, o_rot[4] = {0.f, 0.f, 0.f, 0.f};
    float o_sh[12];
    constexpr bool RED = M4 && (GSB_PB_RED != 0);   // o_* start at zero and are added with REDs at the end
    if (RED) {
#pragma unroll
        for (int k = 0; k < 12; k++) o_sh[k] = 0.f;
    } else if (!modeA) {
        o_var = a.avg[idx];
        o_op = avg_op[idx];
#pragma unroll
        for (int k = 0; k < 3; k++) { o_loc[k] = avg_loc[3 * idx + k]; o_scale[k] = avg_scale[3 * idx + k]; }
#pragma unroll
        for (int k = 0; k < 4; k++) o_rot[k] = avg_rot[4 * idx + k];
        if (M4) {
            const float4* p = reinterpret_cast<const float4*>(avg_sh + 12 * (size_t)idx);   // 48-byte records: 16-byte aligned
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const float4 t = p[k];
                o_sh[4 * k] = t.x; o_sh[4 * k + 1] = t.y; o_sh[4 * k + 2] = t.z; o_sh[4 * k + 3] = t.w;
            }
        }
    }

    // The render backward accumulates the moments H0, Hx, Hy, Hxx, Hxy, Hyy of h = G dL/dalpha (see
    // GradRec); the gradients of backward.cu:536-554 are linear in them with per-splat coefficients.
    float gm2x[NS], gm2y[NS], gcx[NS], gcy[NS], gcw[NS], gop[NS];
    V3 gcol[NS];
    {
        const SplatRec& r = a.recs[idx];
        const float conx = r.a.z, cony = r.a.w, conz = r.b.x, opac = r.b.y;
        float H[NS][9];
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (NS == 2) {
            // fused camera pass: (white, black) pairs of the nine sums in one GradRec2; take and clear
            float4* gp = reinterpret_cast<float4*>(a.grads2 + idx);
            float4 q[5];
#pragma unroll
            for (int i = 0; i < 5; i++) { q[i] = gp[i]; gp[i] = z; }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                H[0][2 * i] = q[i].x; H[1][2 * i] = q[i].y; H[0][2 * i + 1] = q[i].z; H[1][2 * i + 1] = q[i].w;
            }
            H[0][8] = q[4].x; H[1][8] = q[4].y;
        } else {
            GradRec* gp = a.grads + idx;
            const GradRec g = *gp;
            gp->a = z; gp->b = z; gp->c = z;                      // take and clear the per-view scratch
            H[0][0] = g.a.x; H[0][1] = g.a.y; H[0][2] = g.a.z; H[0][3] = g.a.w; H[0][4] = g.b.x; H[0][5] = g.b.y;
            H[0][6] = g.b.z; H[0][7] = g.b.w; H[0][8] = g.c.x;
        }
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const float H0 = H[s][0], Hx = H[s][1], Hy = H[s][2], Hxx = H[s][3], Hxy = H[s][4], Hyy = H[s][5];
            gop[s] = H0;
            gm2x[s] = -(0.5f * a.W) * opac * (conx * Hx + cony * Hy);
            gm2y[s] = -(0.5f * a.H) * opac * (conz * Hy + cony * Hx);
            gcx[s] = -0.5f * opac * Hxx;
            gcy[s] = -0.5f * opac * Hxy;
            gcw[s] = -0.5f * opac * Hyy;
            gcol[s] = {H[s][6], H[s][7], H[s][8]};
        }
    }
    if (modeA) {
        // Rasterizer::backward accumulates into caller arrays (atomicAdd semantics) and the
        // per-splat stage then reads the totals.
        gm2x[0] = (a.dL_dmean2D[3 * idx] += gm2x[0]);
        gm2y[0] = (a.dL_dmean2D[3 * idx + 1] += gm2y[0]);
        gcx[0] = (a.dL_dconic[4 * idx] += gcx[0]);
        gcy[0] = (a.dL_dconic[4 * idx + 1] += gcy[0]);
        gcw[0] = (a.dL_dconic[4 * idx + 3] += gcw[0]);
        gop[0] = (a.dL_dopacity[idx] += gop[0]);
        gcol[0].x = (a.dL_dcolor[3 * idx] += gcol[0].x);
        gcol[0].y = (a.dL_dcolor[3 * idx + 1] += gcol[0].y);
        gcol[0].z = (a.dL_dcolor[3 * idx + 2] += gcol[0].z);
    }

    const float mx = a.means3D[3 * idx], my = a.means3D[3 * idx + 1], mz = a.means3D[3 * idx + 2];
    const float* V = a.viewmatrix;
    const float* Pm = a.projmatrix;
    const float v0 = __ldg(V + 0), v1 = __ldg(V + 1), v2 = __ldg(V + 2);
    const float v4 = __ldg(V + 4), v5 = __ldg(V + 5), v6 = __ldg(V + 6);
    const float v8 = __ldg(V + 8), v9 = __ldg(V + 9), v10 = __ldg(V + 10);

    // ---- per-splat geometry of the covariance path (computeCov2DCUDA) ----
    float c3[6];
    if constexpr (M4) {   // the trainer shape moves covariance, SH record and quaternion as 8- / 16-byte words (launcher checks alignment)
        const float2* c2 = reinterpret_cast<const float2*>(a.cov3D + 6 * (size_t)idx);
        const float2 u0 = c2[0], u1 = c2[1], u2 = c2[2];
        c3[0] = u0.x; c3[1] = u0.y; c3[2] = u1.x; c3[3] = u1.y; c3[4] = u2.x; c3[5] = u2.y;
    } else {
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = a.cov3D[6 * idx + k];
    }
    float tx = v0 * mx + v4 * my + v8 * mz + __ldg(V + 12);
    float ty = v1 * mx + v5 * my + v9 * mz + __ldg(V + 13);
    const float tz = v2 * mx + v6 * my + v10 * mz + __ldg(V + 14);
    const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float hx = a.focal_x, hy = a.focal_y;
    const float J00 = hx / tz, J02 = -(hx * tx) / (tz * tz), J11 = hy / tz, J12 = -(hy * ty) / (tz * tz);
    const float a0 = v0 * J00 + v2 * J02, a1 = v4 * J00 + v6 * J02, a2 = v8 * J00 + v10 * J02;   // T col 0
    const float b0 = v1 * J11 + v2 * J12, b1 = v5 * J11 + v6 * J12, b2 = v9 * J11 + v10 * J12;   // T col 1
    const float ua0 = a0 * c3[0] + a1 * c3[1] + a2 * c3[2];
    const float ua1 = a0 * c3[1] + a1 * c3[3] + a2 * c3[4];
    const float ua2 = a0 * c3[2] + a1 * c3[4] + a2 * c3[5];
    const float ub0 = b0 * c3[0] + b1 * c3[1] + b2 * c3[2];
    const float ub1 = b0 * c3[1] + b1 * c3[3] + b2 * c3[4];
    const float ub2 = b0 * c3[2] + b1 * c3[4] + b2 * c3[5];
    const float ca = a0 * ua0 + a1 * ua1 + a2 * ua2 + 0.3f;
    const float cb = a0 * ub0 + a1 * ub1 + a2 * ub2;
    const float cc = b0 * ub0 + b1 * ub1 + b2 * ub2 + 0.3f;
    const float denom = ca * cc - cb * cb;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    // projection Jacobian of the screen-space mean
    const float p0 = __ldg(Pm + 0), p1 = __ldg(Pm + 1), p3 = __ldg(Pm + 3);
    const float p4 = __ldg(Pm + 4), p5 = __ldg(Pm + 5), p7 = __ldg(Pm + 7);
    const float p8 = __ldg(Pm + 8), p9 = __ldg(Pm + 9), p11 = __ldg(Pm + 11);
    const float hw = p3 * mx + p7 * my + p11 * mz + __ldg(Pm + 15);
    const float m_w = 1.0f / (hw + 0.0000001f);
    const float mul1 = (p0 * mx + p4 * my + p8 * mz + __ldg(Pm + 12)) * m_w * m_w;
    const float mul2 = (p1 * mx + p5 * my + p9 * mz + __ldg(Pm + 13)) * m_w * m_w;

    // ---- per set: conic -> cov2D, and the mean gradient of that pass ----
    float dLa[NS], dLb[NS], dLc[NS];
    V3 gmean[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        dLa[s] = dLb[s] = dLc[s] = 0.f;
        if (denom2inv != 0.f) {
            dLa[s] = denom2inv * (-cc * cc * gcx[s] + 2.f * cb * cc * gcy[s] + (denom - ca * cc) * gcw[s]);
            dLc[s] = denom2inv * (-ca * ca * gcw[s] + 2.f * ca * cb * gcy[s] + (denom - ca * cc) * gcx[s]);
            dLb[s] = denom2inv * 2.f * (cb * cc * gcx[s] - (denom + 2.f * cb * cb) * gcy[s] + ca * cb * gcw[s]);
        }
        const float dT00 = 2.f * ua0 * dLa[s] + ub0 * dLb[s], dT01 = 2.f * ua1 * dLa[s] + ub1 * dLb[s], dT02 = 2.f * ua2 * dLa[s] + ub2 * dLb[s];
        const float dT10 = 2.f * ub0 * dLc[s] + ua0 * dLb[s], dT11 = 2.f * ub1 * dLc[s] + ua1 * dLb[s], dT12 = 2.f * ub2 * dLc[s] + ua2 * dLb[s];
        const float dJ00 = v0 * dT00 + v4 * dT01 + v8 * dT02;
        const float dJ02 = v2 * dT00 + v6 * dT01 + v10 * dT02;
        const float dJ11 = v1 * dT10 + v5 * dT11 + v9 * dT12;
        const float dJ12 = v2 * dT10 + v6 * dT11 + v10 * dT12;
        const float dtx = x_grad_mul * -hx * itz2 * dJ02;
        const float dty = y_grad_mul * -hy * itz2 * dJ12;
        const float dtz = -hx * itz2 * dJ00 - hy * itz2 * dJ11 + (2.f * hx * tx) * itz3 * dJ02 + (2.f * hy * ty) * itz3 * dJ12;
        gmean[s] = {v0 * dtx + v1 * dty + v2 * dtz, v4 * dtx + v5 * dty + v6 * dtz, v8 * dtx + v9 * dty + v10 * dtz};
        // screen-space mean gradient through the perspective divide
        gmean[s].x += (p0 * m_w - p3 * mul1) * gm2x[s] + (p1 * m_w - p3 * mul2) * gm2y[s];
        gmean[s].y += (p4 * m_w - p7 * mul1) * gm2x[s] + (p5 * m_w - p7 * mul2) * gm2y[s];
        gmean[s].z += (p8 * m_w - p11 * mul1) * gm2x[s] + (p9 * m_w - p11 * mul2) * gm2y[s];
    }
    // sums over the sets: everything downstream except the var term only needs these
    float sLa = dLa[0], sLb = dLb[0], sLc = dLc[0], sop = gop[0];
#pragma unroll
    for (int s = 1; s < NS; s++) { sLa += dLa[s]; sLb += dLb[s]; sLc += dLc[s]; sop += gop[s]; }
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0.f) {
        dcov[0] = a0 * a0 * sLa + a0 * b0 * sLb + b0 * b0 * sLc;
        dcov[3] = a1 * a1 * sLa + a1 * b1 * sLb + b1 * b1 * sLc;
        dcov[5] = a2 * a2 * sLa + a2 * b2 * sLb + b2 * b2 * sLc;
        dcov[1] = 2.f * a0 * a1 * sLa + (a0 * b1 + a1 * b0) * sLb + 2.f * b0 * b1 * sLc;
        dcov[2] = 2.f * a0 * a2 * sLa + (a0 * b2 + a2 * b0) * sLb + 2.f * b0 * b2 * sLc;
        dcov[4] = 2.f * a2 * a1 * sLa + (a1 * b2 + a2 * b1) * sLb + 2.f * b1 * b2 * sLc;
    }

    const float S = a.samples;
    const float invS = 1.0f / S;
    auto divS = [&](float x) -> float { return POW2S ? x * invS : x / S; };

    // ---- SH backward ----
    if (!a.colors_precomp) {
        const uint8_t cl = a.clamped[idx];
        V3 dRGB[NS];
#pragma unroll
        for (int s = 0; s < NS; s++)
            dRGB[s] = {(cl & 1) ? 0.f : gcol[s].x, (cl & 2) ? 0.f : gcol[s].y, (cl & 4) ? 0.f : gcol[s].z};
        V3 dsum = dRGB[0];
#pragma unroll
        for (int s = 1; s < NS; s++) dsum = dsum + dRGB[s];
        const float* sh = a.shs + (size_t)idx * M * 3;
        ShDirs dirs;
        if (modeA) {
            float* out = a.dL_dsh + (size_t)idx * M * 3;
            dirs = sh_backward(a.D, mx, my, mz, a.cam_pos, sh, dsum, [&](int c, V3 v) {
                out[3 * c] = v.x; out[3 * c + 1] = v.y; out[3 * c + 2] = v.z;
            });
        } else if (M4) {
            float shl[12];
            {
                const float4* sp = reinterpret_cast<const float4*>(sh);
                const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1), s2 = __ldg(sp + 2);
                shl[0] = s0.x; shl[1] = s0.y; shl[2] = s0.z; shl[3] = s0.w; shl[4] = s1.x; shl[5] = s1.y; shl[6] = s1.z; shl[7] = s1.w;
                shl[8] = s2.x; shl[9] = s2.y; shl[10] = s2.z; shl[11] = s2.w;
            }
            dirs = sh_backward(1, mx, my, mz, a.cam_pos, shl, dsum, [&](int c, V3 v) {
                o_sh[3 * c] += divS(v.x); o_sh[3 * c + 1] += divS(v.y); o_sh[3 * c + 2] += divS(v.z);
            });
            float4* p = reinterpret_cast<float4*>(avg_sh + 12 * (size_t)idx);
#pragma unroll
            for (int k = 0; k < 3; k++) {
                if (RED) pb_red4(reinterpret_cast<float*>(p + k), o_sh[4 * k], o_sh[4 * k + 1], o_sh[4 * k + 2], o_sh[4 * k + 3]);
                else p[k] = make_float4(o_sh[4 * k], o_sh[4 * k + 1], o_sh[4 * k + 2], o_sh[4 * k + 3]);
            }
        } else {
            float* out = avg_sh + (size_t)idx * M * 3;
            dirs = sh_backward(a.D, mx, my, mz, a.cam_pos, sh, dsum, [&](int c, V3 v) {
                out[3 * c] += divS(v.x); out[3 * c + 1] += divS(v.y); out[3 * c + 2] += divS(v.z);
            });
        }
#pragma unroll
        for (int s = 0; s < NS; s++) gmean[s] = gmean[s] + dir_grad(dirs, dRGB[s]);
    }

    if (modeA) {
        a.dL_dmean3D[3 * idx] = gmean[0].x; a.dL_dmean3D[3 * idx + 1] = gmean[0].y; a.dL_dmean3D[3 * idx + 2] = gmean[0].z;
#pragma unroll
        for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * idx + k] = dcov[k];
    } else {
        float var = 0.f;
        V3 gs_ = {0.f, 0.f, 0.f};
#pragma unroll
        for (int s = 0; s < NS; s++) {
            var += divS(sqrtf(gmean[s].x * gmean[s].x + gmean[s].y * gmean[s].y + gmean[s].z * gmean[s].z));
            gs_.x += divS(gmean[s].x); gs_.y += divS(gmean[s].y); gs_.z += divS(gmean[s].z);
        }
        if (RED) {
            pb_red(a.avg + idx, var);
            pb_red(avg_loc + 3 * idx, gs_.x); pb_red(avg_loc + 3 * idx + 1, gs_.y); pb_red(avg_loc + 3 * idx + 2, gs_.z);
            pb_red(avg_op + idx, sop * invS);
        } else {
            a.avg[idx] = o_var + var;
            avg_loc[3 * idx] = o_loc[0] + gs_.x; avg_loc[3 * idx + 1] = o_loc[1] + gs_.y; avg_loc[3 * idx + 2] = o_loc[2] + gs_.z;
            avg_op[idx] = o_op + sop * invS;
        }
    }

    // ---- cov3D -> scale, rotation ----
    if (!a.cov3D_is_precomp) {
        float r, x, y, z;
        if constexpr (M4) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
            r = q.x; x = q.y; y = q.z; z = q.w;
        } else {
            r = a.rotations[4 * idx]; x = a.rotations[4 * idx + 1]; y = a.rotations[4 * idx + 2]; z = a.rotations[4 * idx + 3];
        }
        const float s[3] = {a.scale_modifier * a.scales[3 * idx], a.scale_modifier * a.scales[3 * idx + 1],
                            a.scale_modifier * a.scales[3 * idx + 2]};
        // Rg[c][r]: columns of the reference's R literal
        const float Rg[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        const float Dg[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
        // dMt[c][r] = 2 * sum_k (s_c Rg[k][c]) * Dg[r][k]
        float dMt[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++)
                dMt[c][rr] = 2.0f * (s[c] * Rg[0][c] * Dg[rr][0] + s[c] * Rg[1][c] * Dg[rr][1] + s[c] * Rg[2][c] * Dg[rr][2]);
        float gs[3];
#pragma unroll
        for (int c = 0; c < 3; c++) gs[c] = Rg[0][c] * dMt[c][0] + Rg[1][c] * dMt[c][1] + Rg[2][c] * dMt[c][2];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) dMt[c][rr] *= s[c];
        float gq[4];
        gq[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
        gq[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
        gq[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
        gq[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
        if (modeA) {
#pragma unroll
            for (int c = 0; c < 3; c++) a.dL_dscale[3 * idx + c] = gs[c];
#pragma unroll
            for (int c = 0; c < 4; c++) a.dL_drot[4 * idx + c] = gq[c];
        } else if (RED) {
#pragma unroll
            for (int c = 0; c < 3; c++) pb_red(avg_scale + 3 * idx + c, divS(gs[c]));
            pb_red4(avg_rot + 4 * idx, divS(gq[0]), divS(gq[1]), divS(gq[2]), divS(gq[3]));
        } else {
#pragma unroll
            for (int c = 0; c < 3; c++) avg_scale[3 * idx + c] = o_scale[c] + divS(gs[c]);
#pragma unroll
            for (int c = 0; c < 4; c++) avg_rot[4 * idx + c] = o_rot[c] + divS(gq[c]);
        }
    }
}

}  // namespace

void launch_preprocess_bwd(const PreprocessBwdArgs& a, cudaStream_t s) {
    if (a.P <= 0) return;
    const int grid = gsb_div_up(a.P, GSB_PB_THREADS);
    const bool trainer = a.avg != nullptr;
    // the fast shape needs 16-byte aligned 12-float SH records in the running-mean buffer
    const bool m4 = trainer && a.M == 4 && a.D == 1 && !a.colors_precomp &&
                    (reinterpret_cast<uintptr_t>(a.avg) % 16 == 0) && (reinterpret_cast<uintptr_t>(a.shs) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(a.rotations) % 16 == 0) && (reinterpret_cast<uintptr_t>(a.cov3D) % 8 == 0) &&
                    !a.cov3D_is_precomp;
    int ex = 0;
    const bool pow2 = trainer && a.samples > 0.f && frexpf(a.samples, &ex) == 0.5f;   // S = 2^(ex - 1)
    if (trainer && a.grads2 != nullptr) {
        if (m4 && pow2) preprocess_bwd_kernel<2, true, true><<<grid, GSB_PB_THREADS, 0, s>>>(a);
        else if (m4) preprocess_bwd_kernel<2, true><<<grid, GSB_PB_THREADS, 0, s>>>(a);
        else preprocess_bwd_kernel<2, false><<<grid, GSB_PB_THREADS, 0, s>>>(a);
    } else if (m4 && pow2) {
        preprocess_bwd_kernel<1, true, true><<<grid, GSB_PB_THREADS, 0, s>>>(a);
    } else if (m4) {
        preprocess_bwd_kernel<1, true><<<grid, GSB_PB_THREADS, 0, s>>>(a);
    } else {
        preprocess_bwd_kernel<1, false><<<grid, GSB_PB_THREADS, 0, s>>>(a);
    }
}
