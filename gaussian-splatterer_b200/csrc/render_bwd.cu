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
// red.global.add each into the packed GradRec scratch (one RED instruction per warp and entry).
// On the measured scenes 1.2 warps blend a given (tile, splat) duplicate, so a further cross-warp
// stage would save almost nothing.  (Round-1 ncu: the earlier shuffle butterfly was 48 % of the
// kernel's warp instructions, profiles/r01_render_bwd_source.txt.)
//
// DUAL mode serves Trainer::train's white and black pass of one camera (src/Trainer.cu:311-314) in
// a single walk: both passes blend the same splats with the same alpha and T (the background only
// enters through out = C + T * bg, forward.cu:372), so power / exp / T / colour-behind are computed
// once and only the two residuals' sums are kept apart (two GradRec sets).

#include "common.cuh"
#include "render_common.cuh"

namespace {

constexpr int RED_STRIDE = 36;   // floats per row of the reduction panel: 32 lanes + 4 pad (rows stay 16-byte aligned)

constexpr int WB = 32;           // list entries per warp-private chunk (one per lane)

struct BwdSmem {
    SplatRec rec[RENDER_WARPS][2][WB];
    uint32_t id[RENDER_WARPS][2][WB];
};

// MODE 0: dL/dpixel given (Rasterizer::backward).  MODE 1: residual of one truth frame fused.
// MODE 2: white and black truth frames of one camera fused (two residuals, two GradRec sets).
template <int MODE>
__global__ void __launch_bounds__(RENDER_THREADS, 6) render_bwd_kernel(const RenderBwdArgs a) {
    constexpr int NS = (MODE == 2) ? 2 : 1;   // residual sets
    __shared__ BwdSmem sm;
    __shared__ __align__(16) float s_red[RENDER_WARPS][9 * NS * RED_STRIDE];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* const red = s_red[warp];
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int start = (int)a.tile_max[tile];       // entries at or beyond this index were blended by no pixel
    if (start == 0 && (MODE == 0 || a.loss_out == nullptr)) return;
    const uint2 range = a.ranges[tile];
    const uint32_t* list = a.point_list + range.x;

    const int bx0 = blockIdx.x * GSB_TILE + (warp & 1) * 8, by0 = blockIdx.y * GSB_TILE + (warp >> 1) * 8;
    const int pix_x = bx0 + (lane & 7);
    const float pxf = (float)pix_x;
    const float bcx = (float)bx0 + 3.5f, bcy = (float)by0 + 3.5f;
    const size_t plane = (size_t)a.W * a.H;

    float pyf[2], T[2], Tfin[2], last_alpha[2];
    float lcr[2], lcg[2], lcb[2], arr[2], arg[2], arb[2];
    float dLr[NS][2], dLg[NS][2], dLb[NS][2], bgdot[2];   // bgdot: set 0 only (set 1 is the black pass)
    int ncontrib[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int py = by0 + (lane >> 3) + 4 * k;
        pyf[k] = (float)py;
        const bool inside = pix_x < a.W && py < a.H;
        const size_t pid = (size_t)py * a.W + pix_x;
        Tfin[k] = inside ? a.final_T[pid] : 0.f;
        T[k] = Tfin[k];
        ncontrib[k] = inside ? (int)a.n_contrib[pid] : 0;
#pragma unroll
        for (int s = 0; s < NS; s++) dLr[s][k] = dLg[s][k] = dLb[s][k] = 0.f;
        bgdot[k] = 0.f;
        if (inside) {
            if (MODE == 0) {
                dLr[0][k] = a.dL_dpix[pid];
                dLg[0][k] = a.dL_dpix[plane + pid];
                dLb[0][k] = a.dL_dpix[2 * plane + pid];
                bgdot[k] = __ldg(a.bg) * dLr[0][k] + __ldg(a.bg + 1) * dLg[0][k] + __ldg(a.bg + 2) * dLb[0][k];
            } else if (MODE == 1) {
                // residual = truth/255 - rendered (imageIntToLoss, src/Trainer.cu:38-42)
                const uint32_t t = a.truth[pid];
                dLr[0][k] = (float)(t & 0xFF) / 255.0f - a.rendered[pid];
                dLg[0][k] = (float)((t >> 8) & 0xFF) / 255.0f - a.rendered[plane + pid];
                dLb[0][k] = (float)((t >> 16) & 0xFF) / 255.0f - a.rendered[2 * plane + pid];
                bgdot[k] = __ldg(a.bg) * dLr[0][k] + __ldg(a.bg + 1) * dLg[0][k] + __ldg(a.bg + 2) * dLb[0][k];
            } else {
                // `rendered` is the black-background image C; the white one is fma(1, T, C) = C + T
                const float cr = a.rendered[pid], cg = a.rendered[plane + pid], cb = a.rendered[2 * plane + pid];
                const uint32_t tw = a.truth[pid], tb = a.truth2[pid];
                dLr[0][k] = (float)(tw & 0xFF) / 255.0f - __fadd_rn(cr, Tfin[k]);
                dLg[0][k] = (float)((tw >> 8) & 0xFF) / 255.0f - __fadd_rn(cg, Tfin[k]);
                dLb[0][k] = (float)((tw >> 16) & 0xFF) / 255.0f - __fadd_rn(cb, Tfin[k]);
                dLr[NS - 1][k] = (float)(tb & 0xFF) / 255.0f - cr;
                dLg[NS - 1][k] = (float)((tb >> 8) & 0xFF) / 255.0f - cg;
                dLb[NS - 1][k] = (float)((tb >> 16) & 0xFF) / 255.0f - cb;
                bgdot[k] = dLr[0][k] + dLg[0][k] + dLb[0][k];   // white background (1,1,1)
            }
        }
        last_alpha[k] = 0.f;
        lcr[k] = lcg[k] = lcb[k] = 0.f;
        arr[k] = arg[k] = arb[k] = 0.f;
    }
    if (MODE != 0 && a.loss_out != nullptr) {
        // monitoring only: sum of squared residuals of the pass(es) this launch serves
        float sq = 0.f;
#pragma unroll
        for (int k = 0; k < 2; k++)
#pragma unroll
            for (int s = 0; s < NS; s++)
                sq += dLr[s][k] * dLr[s][k] + dLg[s][k] * dLg[s][k] + dLb[s][k] * dLb[s][k];
        double acc = (double)sq;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) atomicAdd(a.loss_out, acc);
        if (start == 0) return;
    }
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
                const float dx = __fsub_rn(ra.x, pxf);
                const float bx = __fmul_rn(dx, ra.z);
                const float cxy = __fmul_rn(dx, ra.w);
                // per lane and set: A = sum h, B = sum h dy, C = sum h dy^2 over the lane's two pixels (they
                // share dx, so the dx factors of the six moments are applied once after the pixel loop) and
                // the three colour sums
                float mA[NS], mB[NS], mC[NS], vr[NS], vg[NS], vb[NS];
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
                    const float inv = __fdividef(1.0f, 1.f - alpha);
                    T[k] *= inv;
                    const float w = alpha * T[k];
                    // colour behind this splat, reconstructed back to front
                    arr[k] = fmaf(last_alpha[k], lcr[k] - arr[k], arr[k]);
                    arg[k] = fmaf(last_alpha[k], lcg[k] - arg[k], arg[k]);
                    arb[k] = fmaf(last_alpha[k], lcb[k] - arb[k], arb[k]);
                    lcr[k] = rb.z; lcg[k] = rb.w; lcb[k] = rc.x;
                    last_alpha[k] = alpha;
                    const float er = rb.z - arr[k], eg = rb.w - arg[k], eb = rc.x - arb[k];
                    const float GT = G * T[k];
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        vr[s] = fmaf(w, dLr[s][k], vr[s]);
                        vg[s] = fmaf(w, dLg[s][k], vg[s]);
                        vb[s] = fmaf(w, dLb[s][k], vb[s]);
                        // h = G * dL/dalpha; the background term (alpha also scales how much of bg shows
                        // through) exists for a non-black background only
                        float h = GT * (er * dLr[s][k] + eg * dLg[s][k] + eb * dLb[s][k]);
                        if (s == 0) h = fmaf(-(G * Tfin[k] * inv), bgdot[k], h);
                        const float hy = h * dy;
                        mA[s] += h;
                        mB[s] += hy;
                        mC[s] = fmaf(hy, dy, mC[s]);
                    }
                }
                if (!__any_sync(0xffffffffu, hit)) continue;
                // Transposed reduction through shared memory: every lane parks its 9*NS partial sums in
                // column `lane` of the warp's [9*NS][RED_STRIDE] panel (conflict-free: bank = 4*row + lane),
                // then lane r adds up row r with eight 128-bit loads and issues the one RED for GradRec
                // float r.  ~60 warp instructions per (warp, entry) against ~95 for two shuffle butterflies.
#pragma unroll
                for (int s = 0; s < NS; s++) {
                    float* col = red + s * 9 * RED_STRIDE + lane;
                    const float ax = mA[s] * dx;
                    col[0 * RED_STRIDE] = mA[s];          // H0
                    col[1 * RED_STRIDE] = ax;             // Hx
                    col[2 * RED_STRIDE] = mB[s];          // Hy
                    col[3 * RED_STRIDE] = ax * dx;        // Hxx
                    col[4 * RED_STRIDE] = mB[s] * dx;     // Hxy
                    col[5 * RED_STRIDE] = mC[s];          // Hyy
                    col[6 * RED_STRIDE] = vr[s];
                    col[7 * RED_STRIDE] = vg[s];
                    col[8 * RED_STRIDE] = vb[s];
                }
                __syncwarp();
                if (lane < 9 * NS) {
                    const float4* row = reinterpret_cast<const float4*>(red + lane * RED_STRIDE);
                    float4 acc = row[0];
#pragma unroll
                    for (int c = 1; c < 8; c++) {
                        const float4 t = row[c];
                        acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                    }
                    const float sum = (acc.x + acc.y) + (acc.z + acc.w);
                    GradRec* g = ((NS == 1 || lane < 9) ? a.grads : a.grads2) + ids[j];
                    atomicAdd(reinterpret_cast<float*>(g) + ((NS == 1 || lane < 9) ? lane : lane - 9), sum);
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
    if (a.dL_dpix != nullptr) render_bwd_kernel<0><<<grid, RENDER_THREADS, 0, s>>>(a);
    else if (a.truth2 == nullptr) render_bwd_kernel<1><<<grid, RENDER_THREADS, 0, s>>>(a);
    else render_bwd_kernel<2><<<grid, RENDER_THREADS, 0, s>>>(a);
}
