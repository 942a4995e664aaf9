// Measurement only (never on the training path): counts the (pixel, splat) pairs one view costs
// under the REFERENCE's control flow, so bench.py can turn a measured kernel time into achieved
// FLOP/s with the per-pair figures of SURVEY.md section 8(d):
//   forward  (reference forward.cu:310-357): a pixel evaluates `power` for every list entry until it
//            is done (T would fall below 1e-4) or the list ends          -> E_eval_f
//            entries with power <= 0 reach the exp                        -> E_exp_f
//            entries with alpha >= 1/255 before the stop are blended      -> E_blend_f
//   backward (reference backward.cu:476-503): a pixel evaluates the entries below its n_contrib
//                                                                         -> E_eval_b, E_exp_b, E_blend_b
// One thread per pixel, no staging: this replays the blend decisions with the same pinned arithmetic
// as render_fwd.cu, so the counts are those of the images actually produced.

#include <cuda.h>
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) count_pairs_kernel(const RenderFwdArgs a, unsigned long long* out) {
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int px = blockIdx.x * GSB_TILE + (threadIdx.x & 15), py = blockIdx.y * GSB_TILE + (threadIdx.x >> 4);
    const bool inside = px < a.W && py < a.H;
    const uint2 range = a.ranges[tile];
    const int len = (int)(range.y - range.x);
    const float pxf = (float)px, pyf = (float)py;
    unsigned int ev = 0, ex = 0, bl = 0;          // forward counters
    unsigned int ev_b = 0, ex_b = 0;              // backward: values of ev / ex at the last blend
    float T = 1.0f;
    if (inside) {
        for (int j = 0; j < len; j++) {
            const SplatRec& r = a.recs[a.point_list[range.x + j]];
            const float4 ra = r.a;
            const float conz = r.b.x, opac = r.b.y;
            ev++;
            const float dx = __fsub_rn(ra.x, pxf), dy = __fsub_rn(ra.y, pyf);
            const float ay = __fmul_rn(dy, __fmul_rn(dy, conz));
            const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, ra.z), ay), -0.5f, -__fmul_rn(dy, __fmul_rn(dx, ra.w)));
            if (power > 0.0f) continue;
            ex++;
            const float alpha = fminf(0.99f, __fmul_rn(opac, expf(power)));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
            if (test_T < 0.0001f) break;
            T = test_T;
            bl++;
            ev_b = ev; ex_b = ex;
        }
    }
    unsigned int v[5] = {ev, ex, bl, ev_b, ex_b};
#pragma unroll
    for (int q = 0; q < 5; q++) v[q] = __reduce_add_sync(0xffffffffu, v[q]);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out + 0, (unsigned long long)v[0]);
        atomicAdd(out + 1, (unsigned long long)v[1]);
        atomicAdd(out + 2, (unsigned long long)v[2]);
        atomicAdd(out + 3, (unsigned long long)v[3]);
        atomicAdd(out + 4, (unsigned long long)v[4]);
        atomicAdd(out + 5, (unsigned long long)v[2]);
    }
}

// ---- A/B skeleton: render backward with lanes = list entries (Mallick et al. 2024) ---------------
// A LOWER BOUND on the time of a per-splat backward at this workload, measured instead of argued: every warp takes
// buckets of 32 consecutive list entries of its tile, each lane holds one splat's record and its nine (white, black)
// sums in registers, and the 256 pixels of the tile stream through the lanes in list order (pixel p is at lane l at
// step p + l) with the per-pixel running state (T and the two residual-weighted colour prefixes, 3 floats) handed from
// lane to lane by shfl_up.  Per-pixel constants (two residuals, the two final colour-behind scalars, n_contrib) sit in
// shared memory.  What a real kernel would need on top is deliberately left out: the forward's per-bucket checkpoints
// (lane 0 starts every pixel from final_T here), exact equality of the blend decisions, and the fused loss.  Its
// results are therefore NOT gradients; only its run time is used (gsb_trainer_measure_bwd_skeleton).
__global__ void __launch_bounds__(128) render_bwd_lanes_are_splats_skeleton(const RenderBwdArgs a) {
    __shared__ float4 s_pix[256][3];   // {dLw.rgb, SDw}, {dLb.rgb, SDb}, {n_contrib, T_fin, -, -}
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int len = (int)a.tile_max[tile];
    if (len == 0) return;
    const uint2 range = a.ranges[tile];
    const uint32_t* list = a.point_list + range.x;
    const size_t plane = (size_t)a.W * a.H;
    for (int p = tid; p < 256; p += 128) {
        const int x = blockIdx.x * 16 + (p & 15), y = blockIdx.y * 16 + (p >> 4);
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f), b = w, c = w;
        if (x < a.W && y < a.H) {
            const size_t pid = (size_t)y * a.W + x;
            const float cr = a.rendered[pid], cg = a.rendered[plane + pid], cb = a.rendered[2 * plane + pid];
            const float Tf = a.final_T[pid];
            const uint32_t tw = a.truth[pid], tb = a.truth2[pid];
            w = make_float4((float)(tw & 0xFF) / 255.0f - (cr + Tf), (float)((tw >> 8) & 0xFF) / 255.0f - (cg + Tf),
                            (float)((tw >> 16) & 0xFF) / 255.0f - (cb + Tf), 0.f);
            b = make_float4((float)(tb & 0xFF) / 255.0f - cr, (float)((tb >> 8) & 0xFF) / 255.0f - cg,
                            (float)((tb >> 16) & 0xFF) / 255.0f - cb, 0.f);
            w.w = w.x * (cr + Tf) + w.y * (cg + Tf) + w.z * (cb + Tf);     // residual . final colour (white)
            b.w = b.x * cr + b.y * cg + b.z * cb;
            c = make_float4(__uint_as_float(a.n_contrib[pid]), Tf, 0.f, 0.f);
        }
        s_pix[p][0] = w; s_pix[p][1] = b; s_pix[p][2] = c;
    }
    __syncthreads();
    const float ox = (float)(blockIdx.x * 16), oy = (float)(blockIdx.y * 16);
    const int buckets = (len + 31) / 32;
    for (int bk = warp; bk < buckets; bk += 4) {
        const int idx = bk * 32 + lane;
        const bool valid = idx < len;
        float4 ra = make_float4(0.f, 0.f, 0.f, 0.f), rb = ra;
        float2 rc = make_float2(0.f, 0.f);
        uint32_t splat = 0;
        if (valid) {
            splat = list[idx];
            const SplatRec* g = a.recs + splat;
            ra = g->a; rb = g->b; rc = make_float2(g->c.x, g->c.y);
        }
        float2 sA = make_float2(0.f, 0.f), sB = sA, sC = sA, sD = sA, sE = sA, sF = sA, sr = sA, sg = sA, sb = sA;
        float T = 1.f, Dw = 0.f, Db = 0.f;
        bool any = false;
        for (int step = 0; step < 256 + 31; step++) {
            // hand-off: the state of pixel p leaves lane l-1 and enters lane l
            float Tin = __shfl_up_sync(0xffffffffu, T, 1), Dwin = __shfl_up_sync(0xffffffffu, Dw, 1), Dbin = __shfl_up_sync(0xffffffffu, Db, 1);
            const int p = step - lane;
            if (lane == 0) { Tin = (p < 256) ? s_pix[p & 255][2].y : 1.f; Dwin = 0.f; Dbin = 0.f; }   // (the forward's checkpoint in a real kernel)
            T = Tin; Dw = Dwin; Db = Dbin;
            if (p < 0 || p >= 256 || !valid) continue;
            const float4 pc = s_pix[p][2];
            if (idx >= (int)__float_as_uint(pc.x)) continue;
            const float dx = ra.x - (ox + (float)(p & 15)), dy = ra.y - (oy + (float)(p >> 4));
            const float power = -0.5f * (ra.z * dx * dx + rb.x * dy * dy) - ra.w * dx * dy;
            if (power > 0.f || power < rc.y) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, rb.y * G);
            if (alpha < 1.0f / 255.0f) continue;
            any = true;
            const float4 w = s_pix[p][0], b = s_pix[p][1];
            const float wgt = alpha * T;
            const float2 cd = make_float2(rb.z * w.x + rb.w * w.y + rc.x * w.z, rb.z * b.x + rb.w * b.y + rc.x * b.z);
            Dw = fmaf(wgt, cd.x, Dw); Db = fmaf(wgt, cd.y, Db);
            const float inv = 1.f / (1.f - alpha);
            const float2 h = make_float2(G * (T * cd.x - (w.w - Dw) * inv), G * (T * cd.y - (b.w - Db) * inv));
            T *= (1.f - alpha);
            sA.x += h.x; sA.y += h.y;
            sB.x = fmaf(h.x, dx, sB.x); sB.y = fmaf(h.y, dx, sB.y);
            sC.x = fmaf(h.x, dy, sC.x); sC.y = fmaf(h.y, dy, sC.y);
            sD.x = fmaf(h.x * dx, dx, sD.x); sD.y = fmaf(h.y * dx, dx, sD.y);
            sE.x = fmaf(h.x * dx, dy, sE.x); sE.y = fmaf(h.y * dx, dy, sE.y);
            sF.x = fmaf(h.x * dy, dy, sF.x); sF.y = fmaf(h.y * dy, dy, sF.y);
            sr.x = fmaf(wgt, w.x, sr.x); sr.y = fmaf(wgt, b.x, sr.y);
            sg.x = fmaf(wgt, w.y, sg.x); sg.y = fmaf(wgt, b.y, sg.y);
            sb.x = fmaf(wgt, w.z, sb.x); sb.y = fmaf(wgt, b.z, sb.y);
        }
        if (any) {   // one 64-bit RED per pair, issued by the owning lane: no cross-lane reduction at all
            float* g = reinterpret_cast<float*>(a.grads2 + splat);
            const float2 v[9] = {sA, sB, sC, sD, sE, sF, sr, sg, sb};
#pragma unroll
            for (int q = 0; q < 9; q++)
                asm volatile("red.global.v2.f32.add [%0], {%1, %2};" ::"l"(g + 2 * q), "f"(v[q].x), "f"(v[q].y) : "memory");
        }
    }
}

// ---- A/B: staging of the render kernels' records, cp.async (LDGSTS) against the TMA row gather ------------------------
// VERDICT r01 item 8b.  sm_100's cp.async.bulk.tensor.2d...tile::gather4 fetches four rows of a 2-D tensor by index: with
// the record array described as [P][12 floats] and a box of {12, 1} one instruction brings four 48-byte records into
// 192 contiguous bytes of shared memory (tools/tma_gather4_probe.cu found the tensor-map shape; the destination must be
// 128-byte aligned, so every group of four records takes a 256-byte slot).  This kernel does only what the render
// kernels' staging does -- every warp of a tile walks the tile's whole list in chunks of 32 entries, one chunk ahead,
// double-buffered -- and consumes a chunk by reading one word of every record, so the time is the staging's:
//   MODE 0: three 16-byte cp.async per lane and chunk (what ships)
//   MODE 1: lanes 0-7 issue one gather4 each (the instruction takes its coordinates from uniform registers: the
//           compiler serialises the eight lanes), completion through an mbarrier per buffer
//   MODE 2: lane 0 issues the eight gather4 of a chunk
template <int MODE>
__global__ void __launch_bounds__(128) staging_ab_kernel(const RenderFwdArgs a, const __grid_constant__ CUtensorMap tmap, float* sink) {
    __shared__ __align__(128) float4 s_rec[4][2][8][16];        // [warp][buffer][group of four records][256 B slot]
    __shared__ uint32_t s_id[4][2][32];
    __shared__ __align__(8) unsigned long long s_bar[4][2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const uint2 range = a.ranges[tile];
    const int len = (int)(range.y - range.x);
    const int chunks = (len + 31) / 32;
    const uint32_t* list = a.point_list + range.x;
    if (MODE != 0) {
        if (lane < 2) {
            const uint32_t b = (uint32_t)__cvta_generic_to_shared(&s_bar[warp][lane]);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
    }
    auto issue = [&](int c) {
        const int buf = c & 1;
        const int q = c * 32 + lane;
        const uint32_t id = q < len ? list[q] : 0u;               // (rows past the end of the list fetch record 0: never consumed)
        if (MODE == 0) {
            if (q < len) {
                const SplatRec* g = a.recs + id;
                float4* d = &s_rec[warp][buf][lane >> 2][(lane & 3) * 3];
                const uint32_t da = (uint32_t)__cvta_generic_to_shared(d);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(&g->a) : "memory");
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da + 16), "l"(&g->b) : "memory");
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da + 32), "l"(&g->c) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        } else {
            s_id[warp][buf][lane] = id;
            __syncwarp();
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar[warp][buf]);
            if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(8 * 192) : "memory");
            __syncwarp();
            if (MODE == 1) {
                if (lane < 8) {
                    const uint4 r = *reinterpret_cast<const uint4*>(&s_id[warp][buf][4 * lane]);
                    const uint32_t da = (uint32_t)__cvta_generic_to_shared(&s_rec[warp][buf][lane][0]);
                    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                                 ::"r"(da), "l"(&tmap), "r"(0), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(bar) : "memory");
                }
            } else if (lane == 0) {
#pragma unroll 1
                for (int g = 0; g < 8; g++) {
                    const uint4 r = *reinterpret_cast<const uint4*>(&s_id[warp][buf][4 * g]);
                    const uint32_t da = (uint32_t)__cvta_generic_to_shared(&s_rec[warp][buf][g][0]);
                    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                                 ::"r"(da), "l"(&tmap), "r"(0), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(bar) : "memory");
                }
            }
        }
    };
    float acc = 0.f;
    if (chunks > 0) issue(0);
    for (int c = 0; c < chunks; c++) {
        const int buf = c & 1;
        if (MODE == 0) {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        } else {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar[warp][buf]);
            const uint32_t parity = (uint32_t)((c >> 1) & 1);
            uint32_t done = 0;
            for (int it = 0; it < (1 << 22) && !done; it++)       // bounded: a copy that never lands must not hang the box
                asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }"
                             : "=r"(done) : "r"(bar), "r"(parity) : "memory");
            if (!done) { if (lane == 0) sink[1] = -1.0f; return; }
        }
        __syncwarp();
        if (c + 1 < chunks) issue(c + 1);
        if (c * 32 + lane < len) acc += s_rec[warp][buf][lane >> 2][(lane & 3) * 3 + 1].y;   // opacity of this lane's record
    }
    if (MODE == 0) asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (acc == 1.2345e30f) sink[0] = acc;
}

// ---- statistics of the render backward's reductions ---------------------------------------------------------------------
// For every (8x8 block, list entry) with at least one blended pixel (= one reduction of render_bwd_kernel): how many of
// the warp's 32 lanes hold a blended pixel (lane = column x row pair (r, r + 4), the backward's mapping), and whether the
// blended pixels stay inside the top or bottom four rows / the left or right four columns of the block.
//   out[0..32]  histogram of the number of lanes with a blended pixel (index 0 unused)
//   out[33]     reductions that go through the shared-memory panel (more than GSB_BWD_DIRECT = 4 lanes)
//   out[34]     ... of those, blended pixels confined to rows 0-3 or rows 4-7
//   out[35]     ... of those, confined to columns 0-3 or columns 4-7
//   out[36]     ... of those, either
//   out[37]     blended (pixel, entry) pairs
//   out[38]     (warp, entry) pairs the bounding-box test lets through
__global__ void __launch_bounds__(128) reduction_stats_kernel(const RenderBwdArgs a, unsigned long long* out) {
    __shared__ unsigned int s_h[40];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 40) s_h[tid] = 0u;
    __syncthreads();
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int start = (int)a.tile_max[tile];
    const uint2 range = a.ranges[tile];
    const int bx0 = blockIdx.x * GSB_TILE + (warp & 1) * 8, by0 = blockIdx.y * GSB_TILE + (warp >> 1) * 8;
    const int pix_x = bx0 + (lane & 7);
    const float pxf = (float)pix_x, bcx = (float)bx0 + 3.5f, bcy = (float)by0 + 3.5f;
    float pyf[2];
    int nc[2];
    for (int k = 0; k < 2; k++) {
        const int py = by0 + (lane >> 3) + 4 * k;
        pyf[k] = (float)py;
        nc[k] = (pix_x < a.W && py < a.H) ? (int)a.n_contrib[(size_t)py * a.W + pix_x] : 0;
    }
    for (int j = 0; j < start; j++) {
        const SplatRec& r = a.recs[a.point_list[range.x + j]];
        const float4 ra = r.a, rb = r.b, rc = r.c;
        if (!((fabsf(ra.x - bcx) <= rc.z + 3.5f) && (fabsf(ra.y - bcy) <= rc.w + 3.5f))) continue;   // uniform over the warp
        bool hit[2];
        for (int k = 0; k < 2; k++) {
            const float dx = __fsub_rn(ra.x, pxf), dy = __fsub_rn(ra.y, pyf[k]);
            const float ay = __fmul_rn(dy, __fmul_rn(dy, rb.x));
            const float power = __fmaf_rn(__fmaf_rn(dx, __fmul_rn(dx, ra.z), ay), -0.5f, -__fmul_rn(dy, __fmul_rn(dx, ra.w)));
            hit[k] = j < nc[k] && !(power > 0.0f) && !(fminf(0.99f, __fmul_rn(rb.y, expf(power))) < 1.0f / 255.0f);
        }
        const unsigned int b0 = __ballot_sync(0xffffffffu, hit[0]), b1 = __ballot_sync(0xffffffffu, hit[1]);
        if (lane == 0) {
            atomicAdd(&s_h[38], 1u);
            const unsigned int any = b0 | b1;
            if (any) {
                const int lanes = __popc(any);
                atomicAdd(&s_h[lanes], 1u);
                atomicAdd(&s_h[37], (unsigned int)(__popc(b0) + __popc(b1)));
                if (lanes > 4) {
                    unsigned int cols = (any | (any >> 8) | (any >> 16) | (any >> 24)) & 0xFFu;
                    unsigned int rows = 0;
                    for (int q = 0; q < 4; q++) {
                        if ((b0 >> (8 * q)) & 0xFFu) rows |= 1u << q;
                        if ((b1 >> (8 * q)) & 0xFFu) rows |= 1u << (q + 4);
                    }
                    const bool tb = (rows & 0xF0u) == 0u || (rows & 0x0Fu) == 0u;
                    const bool lr = (cols & 0xF0u) == 0u || (cols & 0x0Fu) == 0u;
                    atomicAdd(&s_h[33], 1u);
                    if (tb) atomicAdd(&s_h[34], 1u);
                    if (lr) atomicAdd(&s_h[35], 1u);
                    if (tb || lr) atomicAdd(&s_h[36], 1u);
                }
            }
        }
    }
    __syncthreads();
    if (tid < 40 && s_h[tid]) atomicAdd(out + tid, (unsigned long long)s_h[tid]);
}

// FP32 issue peak, measured: every thread runs 8 independent FFMA chains (scalar) or 8 independent packed FFMA2
// chains (fma.rn.f32x2: two FMAs per issue slot) long enough for the launch overhead to vanish.  The render kernels'
// roofline denominator is the scalar figure; the packed one shows what the f32x2 path could add on top.
template <bool PACKED>
__global__ void __launch_bounds__(256) ffma_peak_kernel(float* out, float a, float b, int iters) {
    if constexpr (PACKED) {
        unsigned long long x[8];
        unsigned long long a2, b2;
        asm("mov.b64 %0, {%1, %1};" : "=l"(a2) : "f"(a));
        asm("mov.b64 %0, {%1, %1};" : "=l"(b2) : "f"(b));
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float v = (float)(threadIdx.x + k) * 1e-3f;
            asm("mov.b64 %0, {%1, %1};" : "=l"(x[k]) : "f"(v));
        }
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int k = 0; k < 8; k++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[k]) : "l"(a2), "l"(b2));
        }
        unsigned long long acc = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) acc ^= x[k];
        if (acc == 0x123456789abcdef0ull) out[0] = 1.0f;
    } else {
        float x[8];
#pragma unroll
        for (int k = 0; k < 8; k++) x[k] = (float)(threadIdx.x + k) * 1e-3f;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int k = 0; k < 8; k++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[k]) : "f"(a), "f"(b));
        }
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) acc += x[k];
        if (acc == 12345.678f) out[0] = acc;
    }
}

}  // namespace

// out2[0] = scalar FFMA TFLOP/s, out2[1] = packed FFMA2 TFLOP/s (2 FLOP per FMA), best of `reps` timed launches each.
int gsb_measure_fp32_peak_impl(double* out2, cudaStream_t s) {
    int dev = 0, sms = 0;
    GSB_CUDA_OK(cudaGetDevice(&dev));
    GSB_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    float* d_out = nullptr;
    GSB_CUDA_OK(cudaMalloc(&d_out, sizeof(float)));
    cudaEvent_t e0, e1;
    GSB_CUDA_OK(cudaEventCreate(&e0));
    GSB_CUDA_OK(cudaEventCreate(&e1));
    const int blocks = sms * 8, threads = 256, iters = 4096, reps = 5;
    for (int mode = 0; mode < 2; mode++) {
        double best = 0.0;
        for (int r = 0; r < reps + 1; r++) {                      // first launch is the warm-up
            GSB_CUDA_OK(cudaEventRecord(e0, s));
            if (mode == 0) ffma_peak_kernel<false><<<blocks, threads, 0, s>>>(d_out, 0.999f, 1e-3f, iters);
            else ffma_peak_kernel<true><<<blocks, threads, 0, s>>>(d_out, 0.999f, 1e-3f, iters);
            GSB_CUDA_OK(cudaEventRecord(e1, s));
            GSB_CUDA_OK(cudaEventSynchronize(e1));
            float ms = 0.f;
            GSB_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
            const double fma = (double)blocks * threads * (double)iters * 32.0 * (mode == 0 ? 1.0 : 2.0);
            const double tf = 2.0 * fma / (ms * 1e-3) / 1e12;
            if (r > 0 && tf > best) best = tf;
        }
        out2[mode] = best;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d_out);
    GSB_CUDA_OK(cudaGetLastError());
    return 0;
}

void launch_bwd_skeleton(const RenderBwdArgs& a, cudaStream_t s) {
    render_bwd_lanes_are_splats_skeleton<<<dim3(a.grid_x, a.grid_y), 128, 0, s>>>(a);
}

// opacity check value of mode `mode` is not needed: the three modes are timed only.  Returns the launch error, if any.
void launch_staging_ab(const RenderFwdArgs& a, const void* tensor_map, int mode, float* sink, cudaStream_t s) {
    const CUtensorMap tm = *static_cast<const CUtensorMap*>(tensor_map);
    const dim3 grid(a.grid_x, a.grid_y);
    if (mode == 0) staging_ab_kernel<0><<<grid, 128, 0, s>>>(a, tm, sink);
    else if (mode == 1) staging_ab_kernel<1><<<grid, 128, 0, s>>>(a, tm, sink);
    else staging_ab_kernel<2><<<grid, 128, 0, s>>>(a, tm, sink);
}

void launch_reduction_stats(const RenderBwdArgs& a, unsigned long long* out40, cudaStream_t s) {
    reduction_stats_kernel<<<dim3(a.grid_x, a.grid_y), 128, 0, s>>>(a, out40);
}

void launch_count_pairs(const RenderFwdArgs& a, unsigned long long* out, cudaStream_t s) {
    count_pairs_kernel<<<dim3(a.grid_x, a.grid_y), 256, 0, s>>>(a, out);
}
