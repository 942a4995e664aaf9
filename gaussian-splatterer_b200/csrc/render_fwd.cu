// Forward tile renderer: front-to-back alpha blending of each tile's depth-ordered splat list.
// Replaces renderCUDA<3> forward (reference cuda_rasterizer/forward.cu:261-374).
//
// One 128-thread CTA per 16x16 tile.  Each warp owns an 8x8 pixel sub-block (two pixels per lane,
// same column, four rows apart) so a staged record read from shared memory serves two pixels, and
// walks the tile's list independently of the other three warps (no barrier inside the loop).
// Per chunk of 32 list entries (one per lane):
//   * records (48 B, see SplatRec) are gathered into the warp's double-buffered shared-memory ring with
//     16-byte cp.async copies issued one chunk ahead; the list indices of the chunk after that are
//     prefetched into registers, so neither the dependent gather nor the index load is exposed;
//   * the warp builds a survivor mask with one ballot: an entry survives if the bounding
//     box of { opacity * exp(power) >= 1/255 } (stored, conservatively inflated, in the record) meets
//     the warp's sub-block.  The blend loop then visits survivors only (ffs over the mask), in list
//     order.  Culled pairs are exactly those the reference would `continue` over, so nothing changes
//     in any pixel's T / colour / n_contrib;
//   * a survivor is first tested against the record's threshold on `power`, skipping exp for pairs that
//     cannot reach alpha 1/255.
// The arithmetic of a blended pair (power, expf, alpha, the 1/255 and 1e-4 tests, the colour FMA
// order) is pinned with explicit rn intrinsics to what the reference build executes, so every
// blend / skip / stop decision and n_contrib are identical and images agree bit for bit.

#include "common.cuh"
#include "render_common.cuh"

namespace {

constexpr int WB = 32;   // list entries per warp-private chunk (one per lane)

// 1: both pixels of a lane are evaluated in one straight-line pass with packed FP32 pairs (FFMA2 / FMUL2 / FADD2, IEEE
// round-to-nearest like the scalar intrinsics, so every bit of T / colour / n_contrib is unchanged); a pixel that does
// not blend the entry runs with alpha = 0, which leaves T (T * 1) and the colour (C + T * 0) bit for bit unchanged.
#ifndef GSB_FWD_CA
#define GSB_FWD_CA 0      // 1: record staging through L1 (see cp_async16): the stage alone gains 0.7 %, the pipelined step nothing
#endif
#ifndef GSB_FWD_CTAS
#define GSB_FWD_CTAS 8    // minimum resident CTAs per SM handed to the register allocator: 55 registers, nine CTAs resident -- 0.292 ms against
                          // 0.294 (no bound: 45 registers, ten CTAs), 0.299 (12), 0.301 (9), 0.309 (10: 44 registers); same-box A/B
#endif
#ifndef GSB_FWD_CONV
#define GSB_FWD_CONV 1
#endif

// PREVIEW: Trainer::render's shape -- the pixel goes out as RGBA8 (imageFloatToInt, src/Trainer.cu:21-31, applied to the
// same float colour) and nothing a backward pass would need (final_T, n_contrib, the float planes, tile_max) is written.
template <bool PREVIEW>
__global__ void __launch_bounds__(RENDER_THREADS, GSB_FWD_CTAS) render_fwd_kernel(const RenderFwdArgs a) {
    __shared__ SplatRec s_rec[RENDER_WARPS][2][WB];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // CTA -> tile, longest lists first (common.cuh gsb_lpt_tile)
    const int tile = (int)gsb_lpt_tile(a.ctrl, a.tile_order, (uint32_t)(a.grid_x * a.grid_y), blockIdx.y * a.grid_x + blockIdx.x);
    const int tile_y = tile / a.grid_x, tile_x = tile - tile_y * a.grid_x;
    // warp sub-block origin and this lane's two pixels
    const int bx0 = tile_x * GSB_TILE + (warp & 1) * 8, by0 = tile_y * GSB_TILE + (warp >> 1) * 8;
    const int pix_x = bx0 + (lane & 7);
    const int pix_y[2] = {by0 + (lane >> 3), by0 + (lane >> 3) + 4};
    const float pxf = (float)pix_x;
    const float pyf[2] = {(float)pix_y[0], (float)pix_y[1]};
    const float2 npy2 = make_float2(-pyf[0], -pyf[1]);
    const float bcx = (float)bx0 + 3.5f, bcy = (float)by0 + 3.5f;   // sub-block centre
    bool done[2];
    float T[2] = {1.0f, 1.0f}, Cr[2] = {0.f, 0.f}, Cg[2] = {0.f, 0.f}, Cb[2] = {0.f, 0.f};
    uint32_t last[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 2; k++) done[k] = !(pix_x < a.W && pix_y[k] < a.H);

#if GSB_EXPF_PACKED
    // (folded with a value the compiler cannot see through -- W is never negative -- so that it keeps both in registers
    // instead of re-loading the parameters in front of every survivor's first FMA)
    const int opaque0 = a.W >> 31;
    const ExpfConsts expk = {__int_as_float(__float_as_int(a.expf_k1) ^ opaque0), __int_as_float(__float_as_int(a.expf_k252) ^ opaque0)};
#endif
    const uint2 range = a.ranges[tile];
    const int len = (int)(range.y - range.x);
    const int chunks = (len + WB - 1) / WB;
    const uint32_t* list = a.point_list + range.x;

    // Every warp walks the tile's list on its own (no block-wide barrier in the loop): chunk c holds
    // list positions 32c ... 32c+31, one entry per lane, gathered one chunk ahead.
    uint32_t next_id = (lane < len) ? list[lane] : 0xFFFFFFFFu;
    auto issue = [&](int buf) {
        if (next_id != 0xFFFFFFFFu) {
            const SplatRec* g = a.recs + next_id;
            SplatRec* d = &s_rec[warp][buf][lane];
            cp_async16<GSB_FWD_CA != 0>(&d->a, &g->a);
            cp_async16<GSB_FWD_CA != 0>(&d->b, &g->b);
            cp_async16<GSB_FWD_CA != 0>(&d->c, &g->c);
        }
        cp_async_commit();
    };
    if (chunks > 0) {
        issue(0);
        next_id = (WB + lane < len) ? list[WB + lane] : 0xFFFFFFFFu;
    }

    for (int i = 0; i < chunks; i++) {
        // the warp stops as soon as its 64 pixels are finished (the reference votes per block, forward.cu:306-311;
        // a finished pixel ignores every later entry either way)
        if (__all_sync(0xffffffffu, done[0] && done[1])) break;
        cp_async_wait_all();
        __syncwarp();   // chunk i visible to the whole warp, chunk i-1 fully consumed
        if (i + 1 < chunks) {
            issue((i + 1) & 1);
            const int q = (i + 2) * WB + lane;
            next_id = (q < len) ? list[q] : 0xFFFFFFFFu;
        }
        const SplatRec* buf = s_rec[warp][i & 1];
        const int cnt = min(WB, len - i * WB);

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
                const float4 ra = buf[j].a;   // x, y, conic.x, conic.y
                const float4 rb = buf[j].b;   // conic.z, opacity, r, g
                const float2 rc = *reinterpret_cast<const float2*>(&buf[j].c);   // b, power threshold
                const float dx = __fsub_rn(ra.x, pxf);
                const float bx = __fmul_rn(dx, ra.z);
                const float cxy = __fmul_rn(dx, ra.w);
#if GSB_FWD_CONV
                // power = -0.5 (cx dx^2 + cz dy^2) - cy dx dy for both pixels, in the reference's evaluation order
                const float2 dy2 = add2(dup2(ra.y), npy2);                      // ra.y - py
                const float2 ay2 = mul2(dy2, mul2(dy2, dup2(rb.x)));
                const float2 pw2 = fma2(fma2(dup2(dx), dup2(bx), ay2), dup2(-0.5f), mul2(dy2, dup2(-cxy)));
                const bool pass0 = !done[0] && !(pw2.x < rc.y) && !(pw2.x > 0.0f);
                const bool pass1 = !done[1] && !(pw2.y < rc.y) && !(pw2.y > 0.0f);
                if (!__any_sync(0xffffffffu, pass0 || pass1)) continue;
#if GSB_EXPF_PACKED
                const float2 og2 = mul2(dup2(rb.y), expf2_pinned(pw2, expk));          // opacity * expf(power), both pixels
                const float a0 = fminf(0.99f, og2.x), a1 = fminf(0.99f, og2.y);
#else
                const float a0 = fminf(0.99f, __fmul_rn(rb.y, expf(pw2.x)));
                const float a1 = fminf(0.99f, __fmul_rn(rb.y, expf(pw2.y)));
#endif
                const bool hit0 = pass0 && !(a0 < 1.0f / 255.0f), hit1 = pass1 && !(a1 < 1.0f / 255.0f);
                // (no second vote: a power that passes the stored threshold fails the alpha test in well under 1 % of the cases)
                const float2 T2 = make_float2(T[0], T[1]);
                const float2 test2 = mul2(T2, add2(dup2(1.0f), make_float2(hit0 ? -a0 : -0.0f, hit1 ? -a1 : -0.0f)));
                const bool stop0 = hit0 && test2.x < 0.0001f, stop1 = hit1 && test2.y < 0.0001f;
                const bool do0 = hit0 && !stop0, do1 = hit1 && !stop1;
                const float2 al2 = make_float2(do0 ? a0 : 0.0f, do1 ? a1 : 0.0f);
                float2 c2;
                c2 = fma2(T2, mul2(al2, dup2(rb.z)), make_float2(Cr[0], Cr[1])); Cr[0] = c2.x; Cr[1] = c2.y;
                c2 = fma2(T2, mul2(al2, dup2(rb.w)), make_float2(Cg[0], Cg[1])); Cg[0] = c2.x; Cg[1] = c2.y;
                c2 = fma2(T2, mul2(al2, dup2(rc.x)), make_float2(Cb[0], Cb[1])); Cb[0] = c2.x; Cb[1] = c2.y;
                const uint32_t pos = (uint32_t)(i * WB + j + 1);
                if (do0) { T[0] = test2.x; last[0] = pos; }
                if (do1) { T[1] = test2.y; last[1] = pos; }
                done[0] = done[0] || stop0;
                done[1] = done[1] || stop1;
#else
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    if (done[k]) continue;
                    const float dy = __fsub_rn(ra.y, pyf[k]);
                    // power = -0.5 (cx dx^2 + cz dy^2) - cy dx dy, in the reference's evaluation order
                    const float ay = __fmul_rn(dy, __fmul_rn(dy, rb.x));
                    const float power = __fmaf_rn(__fmaf_rn(dx, bx, ay), -0.5f, -__fmul_rn(dy, cxy));
                    if (power < rc.y) continue;          // cannot reach alpha 1/255
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, __fmul_rn(rb.y, expf(power)));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = __fmul_rn(T[k], __fsub_rn(1.0f, alpha));
                    if (test_T < 0.0001f) {
                        done[k] = true;
                        continue;
                    }
                    Cr[k] = __fmaf_rn(T[k], __fmul_rn(alpha, rb.z), Cr[k]);
                    Cg[k] = __fmaf_rn(T[k], __fmul_rn(alpha, rb.w), Cg[k]);
                    Cb[k] = __fmaf_rn(T[k], __fmul_rn(alpha, rc.x), Cb[k]);
                    T[k] = test_T;
                    last[k] = (uint32_t)(i * WB + j + 1);
                }
#endif
            }
        }
    }
    cp_async_wait_all();

    const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
    const size_t plane = (size_t)a.W * a.H;
    uint32_t mx = 0;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        if (pix_x < a.W && pix_y[k] < a.H) {
            const size_t pid = (size_t)pix_y[k] * a.W + pix_x;
            if constexpr (PREVIEW) {
                // int(c * 256) clamped to [0, 255], alpha = 0xFF
                const int r = min(255, max(0, (int)(__fmaf_rn(bg0, T[k], Cr[k]) * 256.0f)));
                const int g = min(255, max(0, (int)(__fmaf_rn(bg1, T[k], Cg[k]) * 256.0f)));
                const int b = min(255, max(0, (int)(__fmaf_rn(bg2, T[k], Cb[k]) * 256.0f)));
                a.frame[pid] = (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | 0xFF000000u;
                continue;
            }
            a.final_T[pid] = T[k];
            a.n_contrib[pid] = last[k];
            a.out_color[pid] = __fmaf_rn(bg0, T[k], Cr[k]);
            a.out_color[plane + pid] = __fmaf_rn(bg1, T[k], Cg[k]);
            a.out_color[2 * plane + pid] = __fmaf_rn(bg2, T[k], Cb[k]);
            mx = max(mx, last[k]);
        }
    }
    // largest n_contrib of the tile: one atomicMax per warp into tile_max (zeroed by tile_sort_kernel / the caller), so
    // a finished warp retires without waiting at a block barrier for the slowest warp of its tile
    if constexpr (PREVIEW) return;
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && mx) atomicMax(&a.tile_max[tile], mx);
}

}  // namespace

void launch_render_fwd(const RenderFwdArgs& a, cudaStream_t s) {
    RenderFwdArgs b = a;
    b.expf_k1 = GSB_EXPF_K1; b.expf_k252 = GSB_EXPF_K252;
    if (b.frame != nullptr) render_fwd_kernel<true><<<dim3(b.grid_x, b.grid_y), RENDER_THREADS, 0, s>>>(b);
    else render_fwd_kernel<false><<<dim3(b.grid_x, b.grid_y), RENDER_THREADS, 0, s>>>(b);
}
