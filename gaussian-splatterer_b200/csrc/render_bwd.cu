// Backward tile renderer: per-pixel gradients of the alpha blend, reduced per (warp, splat) before
// they touch global memory.  Replaces renderCUDA<3> backward (reference
// cuda_rasterizer/backward.cu:399-557), which issues 9 global atomicAdds per blended
// (pixel, splat) pair, and -- when the loss is fused -- imageIntToLoss (src/Trainer.cu:33-44).
//
// Same tiling, staging and survivor masks as the forward (render_fwd.cu), walked back to front and
// starting at the tile's largest n_contrib instead of the end of the list; a warp additionally drops
// entries beyond the largest n_contrib of its own 64 pixels.
//
// What is accumulated per (pixel, splat) pair is not the reference's nine gradients but the nine
// sums they are linear in: with h = G * dL/dalpha and d = mean2D - pixel,
//     H0 = sum h, Hx = sum h dx, Hy = sum h dy, Hxx = sum h dx dx, Hxy = sum h dx dy, Hyy = sum h dy dy
// and the three colour sums.  dL/dopacity = H0, dL/dmean2D = -(W/2, H/2) * opacity * (conic * (Hx,Hy)),
// dL/dconic = -opacity/2 * (Hxx, Hxy, Hyy): per-splat constants, applied once per splat by the
// preprocess backward (preprocess_bwd.cu) instead of once per pair here.
//
// For every surviving list entry the nine partial sums are added over the lane's two pixels in
// registers and reduced across the warp (skipped when no lane blended the entry) by a transposed
// pass through a per-warp shared-memory panel; the lanes that end up holding a total issue one
// red.global.add each into the packed scratch (one RED instruction per warp and entry).
// On the measured scenes 1.2 warps blend a given (tile, splat) duplicate, so a further cross-warp
// stage would save almost nothing.
//
// The colour behind an entry is not tracked as the reference's normalised three-channel recurrence
// (accum_rec / last_alpha / last_color, backward.cu:489-500) but as SD = dL . (un-normalised colour
// behind, background included), one FMA per residual: T_i (c_i - accum_rec_i) . dL equals
// T_i (c_i . dL) - SD_i / (1 - alpha_i).
//
// DUAL mode serves Trainer::train's white and black pass of one camera (src/Trainer.cu:311-314) in
// a single walk: both passes blend the same splats with the same alpha and T (the background only
// enters through out = C + T * bg, forward.cu:372), so power / exp / T are computed once and every
// per-residual quantity is a (white, black) pair: the arithmetic is issued as packed FP32 pairs
// (FFMA2 / FMUL2 / FADD2), the panel holds float2 and the scratch is GradRec2 (interleaved pairs),
// so one 64-bit RED per reader lane adds both passes' sums.  Round-1 ncu history
// (profiles/r01_render_bwd_source.txt): shuffle butterflies 48 % of the warp instructions ->
// transposed scalar panel 30 % -> packed panel 22 %; 622 M -> 421 M warp instructions per launch.

#include "common.cuh"
#include "render_common.cuh"

namespace {

#ifndef GSB_BWD_VOTE2
#define GSB_BWD_VOTE2 0   // second vote after the alpha test (exits in < 1 % of the cases; without it such an entry adds zeros)
#endif
#ifndef GSB_BWD_DIRECT
#define GSB_BWD_DIRECT 4   // measured: 4 - 6 best (-1.5 %), 12 loses 8 % (same-address REDs serialise in L2)
#endif
#ifndef GSB_BWD_CONV
#define GSB_BWD_CONV 1   // fused mode: straight-line evaluation of both pixels of a lane
#endif
#ifndef GSB_BWD_CTAS
#define GSB_BWD_CTAS 7   // minimum resident CTAs per SM handed to the register allocator: at 7 the fused kernel takes 64 registers (62 at
                         // 8) -- eight CTAs still fit -- and runs 1.5 % faster; 6 and 9 lose (profiles/r02_ab_variants.txt)
#endif

#ifndef GSB_BWD_REDUX
#define GSB_BWD_REDUX 1   // fused mode: eight panel rows read by 8 x 4 lanes, the ninth pair (H0) summed with REDUX in fixed point
#endif
constexpr int PANEL8_ROW = 320;  // GSB_BWD_REDUX: bytes per panel row: 32 lanes x 8 B + 64 (consecutive rows land on opposite 64-byte halves of the banks)

constexpr int RED_STRIDE = 36;   // floats per row of the reduction panel: 32 lanes + 4 pad (rows stay 16-byte aligned)
constexpr int RED_STRIDE2 = 38;  // dual mode: float2 (white, black) per lane: 32 lanes + 4 zero slots (18 chunks of 16 B) + 2 pad; the row
                                 // stride of 304 B = 48 mod 128 keeps the three readers of consecutive rows (48 contiguous bytes per row
                                 // and load) on different banks

// shared-memory accesses through a 32-bit address computed once (kept opaque so it is not re-derived per entry)
__device__ __forceinline__ void sts64(uint32_t addr, float2 v) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
// one RED for a (white, black) pair: REDG.E.ADD.F32x2
__device__ __forceinline__ void red_add2(float* p, float2 v) {
    asm volatile("red.global.v2.f32.add [%0], {%1, %2};" ::"l"(p), "f"(v.x), "f"(v.y) : "memory");
}
// 1 / x for x in [0.01, 1]: the bare MUFU.RCP (no denormal guard needed in that range)
__device__ __forceinline__ float rcp_fast(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

constexpr int WB = 32;           // list entries per warp-private chunk (one per lane)

struct BwdSmem {
    SplatRec rec[RENDER_WARPS][2][WB];
    uint32_t id[RENDER_WARPS][2][WB];
};

// MODE 0: dL/dpixel given (Rasterizer::backward).  MODE 1: residual of one truth frame fused.
// MODE 2: white and black truth frames of one camera fused (two residuals, two GradRec sets).
template <int MODE>
__global__ void __launch_bounds__(RENDER_THREADS, GSB_BWD_CTAS) render_bwd_kernel(const RenderBwdArgs a) {
    constexpr int NS = (MODE == 2) ? 2 : 1;   // residual sets
    __shared__ BwdSmem sm;
    // reduction panel of a warp: NS == 1: [9][RED_STRIDE] floats; NS == 2: [9][RED_STRIDE2] (white, black) pairs
    __shared__ __align__(16) float s_red[RENDER_WARPS][NS == 2 ? 9 * 2 * RED_STRIDE2 : 9 * RED_STRIDE];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* const red = s_red[warp];
    // dual mode: lanes r and r + 9 (r < 9) each sum one half of row r of the panel and feed pair r of the splat's
    // GradRec2; the addresses are computed once and kept opaque so they are not re-derived per entry
    // lane = 3 * row + third (< 27) reads chunks third, third + 3, ..., third + 15 of its row (six 128-bit loads in flight
    // together: one shared-memory round trip; slots 32..35 of a row stay zero); the lanes with third == 0 collect the
    // other two partial pairs with two shuffles and issue the RED for pair `row` of the GradRec2
#if GSB_BWD_REDUX
    // fused mode, GSB_BWD_REDUX: the panel holds pairs 1..8 (Hx .. b) as eight rows of 32 (white, black) pairs; lane = 4 * row +
    // quarter reads chunks quarter, quarter + 4, + 8, + 12 of its row (four 128-bit loads: every quarter-warp covers two rows =
    // all 32 banks once), two xor-shuffles add the four partial pairs, the lanes with quarter == 0 issue the RED of pair
    // row + 1; pair 0 (H0) never touches shared memory (see below) and is added by lane 1
    const int rrow = lane >> 2, third = lane & 3;
    uint32_t col_addr = (uint32_t)__cvta_generic_to_shared(red) + lane * 8;
    uint32_t row_addr = (uint32_t)__cvta_generic_to_shared(red) + rrow * PANEL8_ROW + third * 16;
    float* gbase = third == 0 ? reinterpret_cast<float*>(a.grads2) + (rrow + 1) * 2
                              : (lane == 1 ? reinterpret_cast<float*>(a.grads2) : nullptr);
    if (NS == 2) asm volatile("" : "+r"(col_addr), "+r"(row_addr), "+l"(gbase));
#else
    const int rrow = lane / 3, third = lane - 3 * rrow;
    uint32_t col_addr = (uint32_t)__cvta_generic_to_shared(red) + lane * 8;
    uint32_t row_addr = lane < 27 ? (uint32_t)__cvta_generic_to_shared(red) + rrow * RED_STRIDE2 * 8 + third * 16 : 0u;
    float* gbase = (lane < 27 && third == 0) ? reinterpret_cast<float*>(a.grads2) + rrow * 2 : nullptr;
    if (NS == 2) {
        for (int r = 0; r < 9; r++)
            if (lane < 6) sts64((uint32_t)__cvta_generic_to_shared(red) + (r * RED_STRIDE2 + 32 + lane) * 8, make_float2(0.f, 0.f));
        __syncwarp();
        asm volatile("" : "+r"(col_addr), "+r"(row_addr), "+l"(gbase));
    }
#endif
#if GSB_EXPF_PACKED
    const ExpfConsts expk = {a.expf_k1, a.expf_k252};
#endif
    // CTA -> tile, longest lists first (common.cuh gsb_lpt_tile)
    const int tile = (int)gsb_lpt_tile(a.ctrl, a.tile_order, (uint32_t)(a.grid_x * a.grid_y), blockIdx.y * a.grid_x + blockIdx.x);
    const int tile_y = tile / a.grid_x, tile_x = tile - tile_y * a.grid_x;
    const int start = (int)a.tile_max[tile];       // entries at or beyond this index were blended by no pixel
    if (start == 0 && (MODE == 0 || a.loss_out == nullptr)) return;
    const uint2 range = a.ranges[tile];
    const uint32_t* list = a.point_list + range.x;

    const int bx0 = tile_x * GSB_TILE + (warp & 1) * 8, by0 = tile_y * GSB_TILE + (warp >> 1) * 8;
    const int pix_x = bx0 + (lane & 7);
    const float pxf = (float)pix_x;
    const float bcx = (float)bx0 + 3.5f, bcy = (float)by0 + 3.5f;
    const size_t plane = (size_t)a.W * a.H;

    // SD[s][k] = dL_s . (un-normalised colour of everything behind the current entry, background included):
    // the reference's T_i (c_i - accum_rec_i) . dL - T_final / (1 - alpha_i) bg . dL (backward.cu:489-508) is
    // T_i (c_i . dL) - SD_i / (1 - alpha_i), and SD_{i-1} = SD_i + alpha_i T_i (c_i . dL) needs one FMA per set
    // where the normalised colour behind needs a three-channel recurrence.
    float pyf[2], T[2], SD[NS][2];
    float dLr[NS][2], dLg[NS][2], dLb[NS][2];
    int ncontrib[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int py = by0 + (lane >> 3) + 4 * k;
        pyf[k] = (float)py;
        const bool inside = pix_x < a.W && py < a.H;
        const size_t pid = (size_t)py * a.W + pix_x;
        const float Tfin = inside ? a.final_T[pid] : 0.f;
        T[k] = Tfin;
        ncontrib[k] = inside ? (int)a.n_contrib[pid] : 0;
#pragma unroll
        for (int s = 0; s < NS; s++) dLr[s][k] = dLg[s][k] = dLb[s][k] = 0.f;
        float bgdot = 0.f;                      // set 0 only (set 1 of the dual mode is the black pass)
        if (inside) {
            if (MODE == 0) {
                dLr[0][k] = a.dL_dpix[pid];
                dLg[0][k] = a.dL_dpix[plane + pid];
                dLb[0][k] = a.dL_dpix[2 * plane + pid];
                bgdot = __ldg(a.bg) * dLr[0][k] + __ldg(a.bg + 1) * dLg[0][k] + __ldg(a.bg + 2) * dLb[0][k];
            } else if (MODE == 1) {
                // residual = truth/255 - rendered (imageIntToLoss, src/Trainer.cu:38-42)
                const uint32_t t = a.truth[pid];
                dLr[0][k] = (float)(t & 0xFF) / 255.0f - a.rendered[pid];
                dLg[0][k] = (float)((t >> 8) & 0xFF) / 255.0f - a.rendered[plane + pid];
                dLb[0][k] = (float)((t >> 16) & 0xFF) / 255.0f - a.rendered[2 * plane + pid];
                bgdot = __ldg(a.bg) * dLr[0][k] + __ldg(a.bg + 1) * dLg[0][k] + __ldg(a.bg + 2) * dLb[0][k];
            } else {
                // `rendered` is the black-background image C; the white one is fma(1, T, C) = C + T
                const float cr = a.rendered[pid], cg = a.rendered[plane + pid], cb = a.rendered[2 * plane + pid];
                const uint32_t tw = a.truth[pid], tb = a.truth2[pid];
                dLr[0][k] = (float)(tw & 0xFF) / 255.0f - __fadd_rn(cr, Tfin);
                dLg[0][k] = (float)((tw >> 8) & 0xFF) / 255.0f - __fadd_rn(cg, Tfin);
                dLb[0][k] = (float)((tw >> 16) & 0xFF) / 255.0f - __fadd_rn(cb, Tfin);
                dLr[NS - 1][k] = (float)(tb & 0xFF) / 255.0f - cr;
                dLg[NS - 1][k] = (float)((tb >> 8) & 0xFF) / 255.0f - cg;
                dLb[NS - 1][k] = (float)((tb >> 16) & 0xFF) / 255.0f - cb;
                bgdot = dLr[0][k] + dLg[0][k] + dLb[0][k];   // white background (1,1,1)
            }
        }
#pragma unroll
        for (int s = 0; s < NS; s++) SD[s][k] = 0.f;
        SD[0][k] = Tfin * bgdot;                // the background is the last, fully opaque layer
    }
    if (MODE != 0 && a.loss_out != nullptr) {
        // monitoring only: sum of squared residuals of the pass(es) this launch serves
        float sq = 0.f;
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int s = 0; s < NS; s++)
                sq += dLr[s][k] * dLr[s][k] + dLg[s][k] * dLg[s][k] + dLb[s][k] * dLb[s][k];
        // the warp's 64 pixels in float (384 squares: 1e-6 relative at worst, the value is for monitoring), one double
        // atomic per warp into the tile's slot
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if (lane == 0) atomicAdd(a.loss_out + (tile & (GSB_LOSS_SLOTS - 1)), (double)sq);
        if (start == 0) return;
    }
    asm volatile("" : "+f"(pyf[0]), "+f"(pyf[1]));   // keep the pixel rows in registers (not re-derived per entry)
    const int warp_max = __reduce_max_sync(0xffffffffu, max(ncontrib[0], ncontrib[1]));

    // Every warp walks the list on its own: no block-wide barrier, a warp whose 64 pixels blended
    // nothing at or beyond `wstart` never looks at those entries, and warps of one tile drift apart
    // freely.  Chunk c holds list positions wstart-1-32c ... (one entry per lane), back to front.
    const int wstart = min(start, warp_max);
    if (wstart == 0) return;
    const int chunks = (wstart + WB - 1) / WB;
    auto fetch_id = [&](int c) -> uint32_t {
        const int q = wstart - 1 - (c * WB + lane);
        return q >= 0 ? list[q] : 0xFFFFFFFFu;
    };
    uint32_t next_id = fetch_id(0);
    auto issue = [&](int buf) {
        if (next_id != 0xFFFFFFFFu) {
            const SplatRec* g = a.recs + next_id;
            SplatRec* d = &sm.rec[warp][buf][lane];
            cp_async16(&d->a, &g->a);
            cp_async16(&d->b, &g->b);
            cp_async16(&d->c, &g->c);
            sm.id[warp][buf][lane] = next_id;
        }
        cp_async_commit();
    };
    issue(0);
    next_id = fetch_id(1);

    for (int i = 0; i < chunks; i++) {
        cp_async_wait_all();
        __syncwarp();   // chunk i visible to the whole warp, chunk i-1 fully consumed
        if (i + 1 < chunks) {
            issue((i + 1) & 1);
            next_id = fetch_id(i + 2);
        }
        const SplatRec* buf = sm.rec[warp][i & 1];
        const uint32_t* ids = sm.id[warp][i & 1];
        const int cnt = min(WB, wstart - i * WB);
        const int idx0 = wstart - 1 - i * WB;         // list position of entry 0 of this chunk

        bool keep = false;
        if (lane < cnt) {
            const float2 xy = *reinterpret_cast<const float2*>(&buf[lane].a);
            const float2 ext = *reinterpret_cast<const float2*>(&buf[lane].c.z);
            keep = (fabsf(xy.x - bcx) <= ext.x + 3.5f) && (fabsf(xy.y - bcy) <= ext.y + 3.5f);
        }
        {
            uint32_t bits = __ballot_sync(0xffffffffu, keep);
            while (bits) {
                const int j = __ffs(bits) - 1;
                bits &= bits - 1;
                const int idx = idx0 - j;
                const float4 ra = buf[j].a;   // x, y, conic.x, conic.y
                const float4 rb = buf[j].b;   // conic.z, opacity, r, g
                const float2 rc = *reinterpret_cast<const float2*>(&buf[j].c);   // b, power threshold
                const uint32_t splat = ids[j];
                const float dx = __fsub_rn(ra.x, pxf);
                const float bx = __fmul_rn(dx, ra.z);
                const float cxy = __fmul_rn(dx, ra.w);
                // per lane and set: A = sum h, B = sum h dy, C = sum h dy^2 over the lane's two pixels (they
                // share dx, so the dx factors of the six moments are applied once after the pixel loop) and
                // the three colour sums.  NS == 2: every one of them is a (white, black) pair.
                float mA[NS], mB[NS], mC[NS], vr[NS], vg[NS], vb[NS];
                bool lane_hit = false;
#if GSB_BWD_CONV
                if constexpr (NS == 2) {
                    // Both pixels of the lane in one straight-line pass (no divergent per-pixel blocks: the two
                    // dependency chains interleave).  A pixel that does not blend the entry runs with alpha = G = 0,
                    // which leaves T and SD bit for bit unchanged (1 / (1 - 0) = 1, w = 0) and contributes zero sums.
                    float pw[2], dyv[2];
                    bool pass[2];
                    {
                        // power of both pixels with packed pairs, in the forward's evaluation order (bit-identical)
                        const float2 dy2 = add2(dup2(ra.y), make_float2(-pyf[0], -pyf[1]));
                        const float2 ay2 = mul2(dy2, mul2(dy2, dup2(rb.x)));
                        const float2 pw2 = fma2(fma2(dup2(dx), dup2(bx), ay2), dup2(-0.5f), mul2(dy2, dup2(-cxy)));
                        pw[0] = pw2.x; pw[1] = pw2.y; dyv[0] = dy2.x; dyv[1] = dy2.y;
#pragma unroll
                        for (int k = 0; k < 2; k++) pass[k] = (idx < ncontrib[k]) && !(pw[k] < rc.y) && !(pw[k] > 0.0f);
                    }
                    if (!__any_sync(0xffffffffu, pass[0] || pass[1])) continue;
                    float G[2], alpha[2];
                    bool hit = false;
#if GSB_EXPF_PACKED
                    const float2 g2 = expf2_pinned(make_float2(pw[0], pw[1]), expk);
                    const float2 og2 = mul2(dup2(rb.y), g2);
                    const float gv[2] = {g2.x, g2.y}, ogv[2] = {og2.x, og2.y};
#endif
#pragma unroll
                    for (int k = 0; k < 2; k++) {
#if GSB_EXPF_PACKED
                        const float g = gv[k];
                        const float al = fminf(0.99f, ogv[k]);
#else
                        const float g = expf(pw[k]);
                        const float al = fminf(0.99f, __fmul_rn(rb.y, g));
#endif
                        const bool h = pass[k] && !(al < 1.0f / 255.0f);
                        G[k] = h ? g : 0.f;
                        alpha[k] = h ? al : 0.f;
                        hit = hit || h;
                    }
                    lane_hit = hit;
#if GSB_BWD_VOTE2
                    if (!__any_sync(0xffffffffu, hit)) continue;
#endif
                    float2 sA, sB, sC, sr, sg, sb;
#pragma unroll
                    for (int k = 0; k < 2; k++) {
                        const float inv = rcp_fast(1.f - alpha[k]);
                        T[k] *= inv;
                        const float w = alpha[k] * T[k];
                        const float GT = G[k] * T[k], Ginv = G[k] * inv;
                        const float2 w2 = dup2(w), dy2 = dup2(dyv[k]);
                        const float2 dr = make_float2(dLr[0][k], dLr[1][k]), dg = make_float2(dLg[0][k], dLg[1][k]),
                                     db = make_float2(dLb[0][k], dLb[1][k]);
                        const float2 cd = fma2(dup2(rc.x), db, fma2(dup2(rb.w), dg, mul2(dup2(rb.z), dr)));
                        const float2 sd = make_float2(SD[0][k], SD[1][k]);
                        const float2 h = fma2(dup2(GT), cd, mul2(dup2(-Ginv), sd));   // h = G * dL/dalpha
                        const float2 v = fma2(w2, cd, sd); SD[0][k] = v.x; SD[1][k] = v.y;
                        const float2 hy = mul2(h, dy2);
                        if (k == 0) {
                            sr = mul2(w2, dr); sg = mul2(w2, dg); sb = mul2(w2, db);
                            sA = h; sB = hy; sC = mul2(hy, dy2);
                        } else {
                            sr = fma2(w2, dr, sr); sg = fma2(w2, dg, sg); sb = fma2(w2, db, sb);
                            sA = add2(sA, h); sB = add2(sB, hy); sC = fma2(hy, dy2, sC);
                        }
                    }
                    mA[0] = sA.x; mA[1] = sA.y; mB[0] = sB.x; mB[1] = sB.y; mC[0] = sC.x; mC[1] = sC.y;
                    vr[0] = sr.x; vr[1] = sr.y; vg[0] = sg.x; vg[1] = sg.y; vb[0] = sb.x; vb[1] = sb.y;
                } else
#endif
                {
#pragma unroll
                for (int s = 0; s < NS; s++) mA[s] = mB[s] = mC[s] = vr[s] = vg[s] = vb[s] = 0.f;
                bool hit = false;
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    if (idx >= ncontrib[k]) continue;
                    const float dy = __fsub_rn(ra.y, pyf[k]);
                    const float ay = __fmul_rn(dy, __fmul_rn(dy, rb.x));
                    const float power = __fmaf_rn(__fmaf_rn(dx, bx, ay), -0.5f, -__fmul_rn(dy, cxy));
                    if (power < rc.y) continue;
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, __fmul_rn(rb.y, G));
                    if (alpha < 1.0f / 255.0f) continue;
                    hit = true;
                    const float inv = rcp_fast(1.f - alpha);      // alpha <= 0.99
                    T[k] *= inv;                                   // T in front of this entry
                    const float w = alpha * T[k];
                    const float GT = G * T[k], Ginv = G * inv;
                    if constexpr (NS == 2) {
                        const float2 w2 = dup2(w), dy2 = dup2(dy);
                        const float2 dr = make_float2(dLr[0][k], dLr[1][k]), dg = make_float2(dLg[0][k], dLg[1][k]),
                                     db = make_float2(dLb[0][k], dLb[1][k]);
                        float2 v;
                        v = fma2(w2, dr, make_float2(vr[0], vr[1])); vr[0] = v.x; vr[1] = v.y;
                        v = fma2(w2, dg, make_float2(vg[0], vg[1])); vg[0] = v.x; vg[1] = v.y;
                        v = fma2(w2, db, make_float2(vb[0], vb[1])); vb[0] = v.x; vb[1] = v.y;
                        const float2 cd = fma2(dup2(rc.x), db, fma2(dup2(rb.w), dg, mul2(dup2(rb.z), dr)));
                        const float2 sd = make_float2(SD[0][k], SD[1][k]);
                        const float2 h = fma2(dup2(GT), cd, mul2(dup2(-Ginv), sd));   // h = G * dL/dalpha
                        v = fma2(w2, cd, sd); SD[0][k] = v.x; SD[1][k] = v.y;
                        const float2 hy = mul2(h, dy2);
                        v = add2(make_float2(mA[0], mA[1]), h); mA[0] = v.x; mA[1] = v.y;
                        v = add2(make_float2(mB[0], mB[1]), hy); mB[0] = v.x; mB[1] = v.y;
                        v = fma2(hy, dy2, make_float2(mC[0], mC[1])); mC[0] = v.x; mC[1] = v.y;
                    } else {
                        vr[0] = fmaf(w, dLr[0][k], vr[0]);
                        vg[0] = fmaf(w, dLg[0][k], vg[0]);
                        vb[0] = fmaf(w, dLb[0][k], vb[0]);
                        const float cd = fmaf(rc.x, dLb[0][k], fmaf(rb.w, dLg[0][k], rb.z * dLr[0][k]));
                        const float h = fmaf(GT, cd, -(Ginv * SD[0][k]));              // h = G * dL/dalpha
                        SD[0][k] = fmaf(w, cd, SD[0][k]);
                        const float hy = h * dy;
                        mA[0] += h;
                        mB[0] += hy;
                        mC[0] = fmaf(hy, dy, mC[0]);
                    }
                }
                if (!__any_sync(0xffffffffu, hit)) continue;
                }
#if GSB_BWD_DIRECT > 0
                if constexpr (NS == 2) {
                    // entries at most GSB_BWD_DIRECT lanes blended: those lanes add their nine pairs themselves (nine 64-bit REDs,
                    // no shared-memory round trip: the shared-memory data pipe bounds this kernel)
                    const uint32_t hits = __ballot_sync(0xffffffffu, lane_hit);
                    if (__popc(hits) <= GSB_BWD_DIRECT) {
                        if (lane_hit) {
                            float* g = reinterpret_cast<float*>(a.grads2 + splat);
                            const float2 dx2 = dup2(dx);
                            const float2 A2 = make_float2(mA[0], mA[1]), B2 = make_float2(mB[0], mB[1]);
                            const float2 ax = mul2(A2, dx2);
                            red_add2(g + 0, A2);
                            red_add2(g + 2, ax);
                            red_add2(g + 4, B2);
                            red_add2(g + 6, mul2(ax, dx2));
                            red_add2(g + 8, mul2(B2, dx2));
                            red_add2(g + 10, make_float2(mC[0], mC[1]));
                            red_add2(g + 12, make_float2(vr[0], vr[1]));
                            red_add2(g + 14, make_float2(vg[0], vg[1]));
                            red_add2(g + 16, make_float2(vb[0], vb[1]));
                        }
                        continue;
                    }
                }
#endif
                if constexpr (NS == 2) {
                    // Transposed reduction through shared memory: every lane parks its nine (white, black) pairs in
                    // column `lane` of the warp's [9][RED_STRIDE2] panel with 64-bit stores; lanes 3r, 3r + 1, 3r + 2
                    // add up a third of row r each (six 128-bit loads, a tree of packed adds), lane 3r collects the
                    // three partial pairs with two shuffles and issues one 64-bit RED for pair r of the splat's
                    // GradRec2 -- nine distinct addresses, one RED instruction per warp and entry.
                    const float2 dx2 = dup2(dx);
                    const float2 A2 = make_float2(mA[0], mA[1]), B2 = make_float2(mB[0], mB[1]);
                    const float2 ax = mul2(A2, dx2);
#if GSB_BWD_REDUX
                    // Eight of the nine pairs go through the panel: 8 x 32 pairs = 128 chunks of 16 B = four 128-bit loads
                    // for each of 32 reader lanes, no padding, every quarter-warp of a load on all 32 banks once
                    // (16 + 16 wavefronts against 18 + 24 for nine rows read by 27 lanes).
                    sts64(col_addr + 0 * PANEL8_ROW, ax);                                        // Hx
                    sts64(col_addr + 1 * PANEL8_ROW, B2);                                        // Hy
                    sts64(col_addr + 2 * PANEL8_ROW, mul2(ax, dx2));                             // Hxx
                    sts64(col_addr + 3 * PANEL8_ROW, mul2(B2, dx2));                             // Hxy
                    sts64(col_addr + 4 * PANEL8_ROW, make_float2(mC[0], mC[1]));                 // Hyy
                    sts64(col_addr + 5 * PANEL8_ROW, make_float2(vr[0], vr[1]));
                    sts64(col_addr + 6 * PANEL8_ROW, make_float2(vg[0], vg[1]));
                    sts64(col_addr + 7 * PANEL8_ROW, make_float2(vb[0], vb[1]));
                    __syncwarp();
                    // The ninth pair, H0 = sum h, is added across the warp by the integer REDUX unit instead (no shared-memory
                    // wavefront at all): the lanes' values are scaled by a common power of two chosen from the largest
                    // magnitude in the warp (REDUX.MAX on the float bits) so that each fits 25 bits, rounded to integers,
                    // summed exactly, and scaled back.  Every addend is represented to 2^-24 of the largest one -- the
                    // resolution a float sum of these values has anyway.  Non-finite or huge values (exponent >= 127) take a
                    // shuffle butterfly instead.
                    float2 h0;
                    {
                        const uint32_t mb = __float_as_uint(fmaxf(fabsf(A2.x), fabsf(A2.y)));
                        const uint32_t wm = __reduce_max_sync(0xffffffffu, mb);
                        if (wm < 0x7F000000u) {
                            const int E = max((int)(wm >> 23), 25);                       // biased exponent of the largest magnitude
                            const float scale = __int_as_float((278 - E) << 23);          // 2^(24 - (E - 127)): |x| * scale < 2^25
                            const float inv = __int_as_float((E - 24) << 23);             // 1 / scale
                            const int qw = __float2int_rn(A2.x * scale), qb = __float2int_rn(A2.y * scale);
                            const int sw = __reduce_add_sync(0xffffffffu, qw), sb = __reduce_add_sync(0xffffffffu, qb);
                            h0 = make_float2((float)sw * inv, (float)sb * inv);
                        } else {
                            h0 = A2;
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) {
                                h0.x += __shfl_xor_sync(0xffffffffu, h0.x, o);
                                h0.y += __shfl_xor_sync(0xffffffffu, h0.y, o);
                            }
                        }
                    }
                    {
                        float2 p[4];
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const float4 t = lds128(row_addr + c * 64);
                            p[c] = add2(make_float2(t.x, t.y), make_float2(t.z, t.w));
                        }
                        float2 part = add2(add2(p[0], p[1]), add2(p[2], p[3]));
                        part = add2(part, make_float2(__shfl_xor_sync(0xffffffffu, part.x, 1), __shfl_xor_sync(0xffffffffu, part.y, 1)));
                        part = add2(part, make_float2(__shfl_xor_sync(0xffffffffu, part.x, 2), __shfl_xor_sync(0xffffffffu, part.y, 2)));
                        if (gbase != nullptr) red_add2(gbase + (size_t)splat * (sizeof(GradRec2) / 4), lane == 1 ? h0 : part);
                    }
#else
                    sts64(col_addr + 0 * RED_STRIDE2 * 8, A2);                                   // H0
                    sts64(col_addr + 1 * RED_STRIDE2 * 8, ax);                                   // Hx
                    sts64(col_addr + 2 * RED_STRIDE2 * 8, B2);                                   // Hy
                    sts64(col_addr + 3 * RED_STRIDE2 * 8, mul2(ax, dx2));                        // Hxx
                    sts64(col_addr + 4 * RED_STRIDE2 * 8, mul2(B2, dx2));                        // Hxy
                    sts64(col_addr + 5 * RED_STRIDE2 * 8, make_float2(mC[0], mC[1]));            // Hyy
                    sts64(col_addr + 6 * RED_STRIDE2 * 8, make_float2(vr[0], vr[1]));
                    sts64(col_addr + 7 * RED_STRIDE2 * 8, make_float2(vg[0], vg[1]));
                    sts64(col_addr + 8 * RED_STRIDE2 * 8, make_float2(vb[0], vb[1]));
                    __syncwarp();
                    {
                        float2 part = make_float2(0.f, 0.f);
                        if (row_addr != 0u) {
                            float2 p[6];
#pragma unroll
                            for (int c = 0; c < 6; c++) {
                                const float4 t = lds128(row_addr + c * 48);
                                p[c] = add2(make_float2(t.x, t.y), make_float2(t.z, t.w));
                            }
                            part = add2(add2(add2(p[0], p[1]), add2(p[2], p[3])), add2(p[4], p[5]));
                        }
                        const float2 d1 = make_float2(__shfl_down_sync(0xffffffffu, part.x, 1), __shfl_down_sync(0xffffffffu, part.y, 1));
                        const float2 d2 = make_float2(__shfl_down_sync(0xffffffffu, part.x, 2), __shfl_down_sync(0xffffffffu, part.y, 2));
                        if (gbase != nullptr) red_add2(gbase + (size_t)splat * (sizeof(GradRec2) / 4), add2(part, add2(d1, d2)));
                    }
#endif
                } else {
                    // one set: [9][RED_STRIDE] floats (conflict-free: bank = 4*row + lane), lane r adds up row r
                    // with eight 128-bit loads and issues the one RED for GradRec float r
                    float* col = red + lane;
                    const float ax = mA[0] * dx;
                    col[0 * RED_STRIDE] = mA[0];          // H0
                    col[1 * RED_STRIDE] = ax;             // Hx
                    col[2 * RED_STRIDE] = mB[0];          // Hy
                    col[3 * RED_STRIDE] = ax * dx;        // Hxx
                    col[4 * RED_STRIDE] = mB[0] * dx;     // Hxy
                    col[5 * RED_STRIDE] = mC[0];          // Hyy
                    col[6 * RED_STRIDE] = vr[0];
                    col[7 * RED_STRIDE] = vg[0];
                    col[8 * RED_STRIDE] = vb[0];
                    __syncwarp();
                    if (lane < 9) {
                        const float4* row = reinterpret_cast<const float4*>(red + lane * RED_STRIDE);
                        float4 acc = row[0];
#pragma unroll
                        for (int c = 1; c < 8; c++) {
                            const float4 t = row[c];
                            acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                        }
                        const float sum = (acc.x + acc.y) + (acc.z + acc.w);
                        atomicAdd(reinterpret_cast<float*>(a.grads + splat) + lane, sum);
                    }
                }
                __syncwarp();
            }
        }
    }
    cp_async_wait_all();
}

}  // namespace

void launch_render_bwd(const RenderBwdArgs& a, cudaStream_t s) {
    const dim3 grid(a.grid_x, a.grid_y);
    RenderBwdArgs b = a;
    b.expf_k1 = GSB_EXPF_K1; b.expf_k252 = GSB_EXPF_K252;
    if (b.dL_dpix != nullptr) render_bwd_kernel<0><<<grid, RENDER_THREADS, 0, s>>>(b);
    else if (b.truth2 == nullptr) render_bwd_kernel<1><<<grid, RENDER_THREADS, 0, s>>>(b);
    else render_bwd_kernel<2><<<grid, RENDER_THREADS, 0, s>>>(b);
}
