// Backward tile renderer: per-pixel gradients of the alpha blend, reduced per (tile, splat) before
// they touch global memory.  Replaces renderCUDA<3> backward (reference
// cuda_rasterizer/backward.cu:399-557), which issues 9 global atomicAdds per blended
// (pixel, splat) pair, and -- when the loss is fused -- imageIntToLoss (src/Trainer.cu:33-44).
//
// Same tiling, staging and survivor masks as the forward (render_fwd.cu), walked back to front and
// starting at the tile's largest n_contrib instead of the end of the list; a warp additionally drops
// entries beyond the largest n_contrib of its own 64 pixels.  For every surviving list entry the 9
// partial gradients are
//   1. summed over the lane's two pixels in registers,
//   2. reduced across the warp with a reduce-scatter butterfly (14 shuffles for 9 values instead of
//      45) -- skipped when no lane of the warp blended the entry,
//   3. parked in a per-warp shared-memory slot,
// and after each batch one thread per entry adds the warps' slots and issues three reductions
// (red.global.add.v4.f32 x2 + one scalar) into the packed GradRec scratch: 3 L2 transactions per
// (tile, splat) instead of 9 per (pixel, splat).

#include "common.cuh"
#include "render_common.cuh"

namespace {

__device__ __forceinline__ void red_add_v4(float4* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

struct BwdSmem {
    SplatRec rec[2][RB];
    uint32_t id[2][RB];
    float acc[RENDER_WARPS][RB][12];
    uint32_t mask[RENDER_WARPS][RB / 32];
};

template <bool FUSED_LOSS>
__global__ void __launch_bounds__(RENDER_THREADS) render_bwd_kernel(const RenderBwdArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem& sm = *reinterpret_cast<BwdSmem*>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int start = (int)a.tile_max[tile];       // entries at or beyond this index were blended by no pixel
    if (start == 0) return;
    const uint2 range = a.ranges[tile];
    const uint32_t* list = a.point_list + range.x;
    const int rounds = (start + RB - 1) / RB;

    const int bx0 = blockIdx.x * GSB_TILE + (warp & 1) * 8, by0 = blockIdx.y * GSB_TILE + (warp >> 1) * 8;
    const int pix_x = bx0 + (lane & 7);
    const float pxf = (float)pix_x;
    const float bcx = (float)bx0 + 3.5f, bcy = (float)by0 + 3.5f;
    const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
    const size_t plane = (size_t)a.W * a.H;
    const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;

    float pyf[2], T[2], Tfin[2], bgdot[2], last_alpha[2];
    float dLr[2], dLg[2], dLb[2], lcr[2], lcg[2], lcb[2], arr[2], arg[2], arb[2];
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
        dLr[k] = dLg[k] = dLb[k] = 0.f;
        if (inside) {
            if (FUSED_LOSS) {
                // residual = truth/255 - rendered (imageIntToLoss, src/Trainer.cu:38-42)
                const uint32_t t = a.truth[pid];
                dLr[k] = (float)(t & 0xFF) / 255.0f - a.rendered[pid];
                dLg[k] = (float)((t >> 8) & 0xFF) / 255.0f - a.rendered[plane + pid];
                dLb[k] = (float)((t >> 16) & 0xFF) / 255.0f - a.rendered[2 * plane + pid];
            } else {
                dLr[k] = a.dL_dpix[pid];
                dLg[k] = a.dL_dpix[plane + pid];
                dLb[k] = a.dL_dpix[2 * plane + pid];
            }
        }
        bgdot[k] = bg0 * dLr[k] + bg1 * dLg[k] + bg2 * dLb[k];
        last_alpha[k] = 0.f;
        lcr[k] = lcg[k] = lcb[k] = 0.f;
        arr[k] = arg[k] = arb[k] = 0.f;
    }
    const int warp_max = __reduce_max_sync(0xffffffffu, max(ncontrib[0], ncontrib[1]));

    // back to front: entry p of batch b is list position start - 1 - (b * RB + p)
    auto fetch_id = [&](int batch) -> uint32_t {
        const int q = start - 1 - (batch * RB + tid);
        return q >= 0 ? list[q] : 0xFFFFFFFFu;
    };
    uint32_t next_id = fetch_id(0);
    auto issue = [&](int buf) {
        if (next_id != 0xFFFFFFFFu) {
            const SplatRec* g = a.recs + next_id;
            cp_async16(&sm.rec[buf][tid].a, &g->a);
            cp_async16(&sm.rec[buf][tid].b, &g->b);
            cp_async16(&sm.rec[buf][tid].c, &g->c);
            sm.id[buf][tid] = next_id;
        }
        cp_async_commit();
    };
    issue(0);
    next_id = fetch_id(1);

    for (int i = 0; i < rounds; i++) {
        cp_async_wait_all();
        __syncthreads();   // batch i visible; reduction of batch i-1 finished (acc / mask reusable)
        if (i + 1 < rounds) {
            issue((i + 1) & 1);
            next_id = fetch_id(i + 2);
        }
        const SplatRec* buf = sm.rec[i & 1];
        const int cnt = min(RB, start - i * RB);
        const int idx0 = start - 1 - i * RB;          // list position of entry 0 of this batch

        uint32_t mask[RB / 32], touched[RB / 32];
#pragma unroll
        for (int m = 0; m < RB / 32; m++) {
            const int e = m * 32 + lane;
            bool hit = false;
            if (e < cnt && idx0 - e < warp_max) {
                const float2 xy = *reinterpret_cast<const float2*>(&buf[e].a);
                const float2 ext = *reinterpret_cast<const float2*>(&buf[e].c.z);
                hit = (fabsf(xy.x - bcx) <= ext.x + 3.5f) && (fabsf(xy.y - bcy) <= ext.y + 3.5f);
            }
            mask[m] = __ballot_sync(0xffffffffu, hit);
            touched[m] = 0;
        }
#pragma unroll
        for (int m = 0; m < RB / 32; m++) {
            uint32_t bits = mask[m];
            while (bits) {
                const int b = __ffs(bits) - 1;
                const int j = m * 32 + b;
                bits &= bits - 1;
                const int idx = idx0 - j;
                const float4 ra = buf[j].a;   // x, y, conic.x, conic.y
                const float4 rb = buf[j].b;   // conic.z, opacity, r, g
                const float2 rc = *reinterpret_cast<const float2*>(&buf[j].c);   // b, power threshold
                const float dx = __fsub_rn(ra.x, pxf);
                const float bx = __fmul_rn(dx, ra.z);
                const float cxy = __fmul_rn(dx, ra.w);
                float v[9];
#pragma unroll
                for (int q = 0; q < 9; q++) v[q] = 0.f;
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
                    const float om = 1.f - alpha;
                    T[k] = T[k] / om;
                    const float w = alpha * T[k];
                    // colour behind this splat, reconstructed back to front
                    arr[k] = last_alpha[k] * lcr[k] + (1.f - last_alpha[k]) * arr[k];
                    arg[k] = last_alpha[k] * lcg[k] + (1.f - last_alpha[k]) * arg[k];
                    arb[k] = last_alpha[k] * lcb[k] + (1.f - last_alpha[k]) * arb[k];
                    lcr[k] = rb.z; lcg[k] = rb.w; lcb[k] = rc.x;
                    float dL_dalpha = (rb.z - arr[k]) * dLr[k] + (rb.w - arg[k]) * dLg[k] + (rc.x - arb[k]) * dLb[k];
                    v[6] += w * dLr[k];
                    v[7] += w * dLg[k];
                    v[8] += w * dLb[k];
                    dL_dalpha *= T[k];
                    last_alpha[k] = alpha;
                    dL_dalpha += (-Tfin[k] / om) * bgdot[k];
                    const float dL_dG = rb.y * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * ra.z - gdy * ra.w;
                    const float dG_ddely = -gdy * rb.x - gdx * ra.w;
                    v[0] += dL_dG * dG_ddelx * ddelx_dx;
                    v[1] += dL_dG * dG_ddely * ddely_dy;
                    v[2] += -0.5f * gdx * dx * dL_dG;
                    v[3] += -0.5f * gdx * dy * dL_dG;
                    v[4] += -0.5f * gdy * dy * dL_dG;
                    v[5] += G * dL_dalpha;
                }
                if (!__any_sync(0xffffffffu, hit)) continue;
                touched[m] |= 1u << b;
                // reduce-scatter butterfly over v[0..7]: afterwards lanes 4q..4q+3 hold the warp sum of v[q]
                float u[4], t2[2], s1;
                {
                    const bool hi = lane & 16;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float send = hi ? v[q] : v[q + 4];
                        const float keep = hi ? v[q + 4] : v[q];
                        u[q] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
                }
                {
                    const bool hi = lane & 8;
#pragma unroll
                    for (int q = 0; q < 2; q++) {
                        const float send = hi ? u[q] : u[q + 2];
                        const float keep = hi ? u[q + 2] : u[q];
                        t2[q] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                    }
                }
                {
                    const bool hi = lane & 4;
                    const float send = hi ? t2[0] : t2[1];
                    const float keep = hi ? t2[1] : t2[0];
                    s1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                }
                s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
                s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
                float s8 = v[8];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s8 += __shfl_xor_sync(0xffffffffu, s8, o);
                // lane bits (16,8,4) select the value index: 4*b16 + 2*b8 + b4
                if ((lane & 3) == 0) sm.acc[warp][j][lane >> 2] = s1;
                if (lane == 1) sm.acc[warp][j][8] = s8;
            }
        }
#pragma unroll
        for (int m = 0; m < RB / 32; m++)
            if (lane == m) sm.mask[warp][m] = touched[m];   // warp-uniform values
        __syncthreads();
        // one thread per entry: add the warps' slots, 3 reductions into the packed scratch
        if (tid < cnt) {
            const int e = tid;
            float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1v = s0;
            float s2 = 0.f;
            bool any = false;
#pragma unroll
            for (int w = 0; w < RENDER_WARPS; w++) {
                if ((sm.mask[w][e >> 5] >> (e & 31)) & 1u) {
                    const float4 p0 = *reinterpret_cast<const float4*>(&sm.acc[w][e][0]);
                    const float4 p1 = *reinterpret_cast<const float4*>(&sm.acc[w][e][4]);
                    s0.x += p0.x; s0.y += p0.y; s0.z += p0.z; s0.w += p0.w;
                    s1v.x += p1.x; s1v.y += p1.y; s1v.z += p1.z; s1v.w += p1.w;
                    s2 += sm.acc[w][e][8];
                    any = true;
                }
            }
            if (any) {
                GradRec* g = a.grads + sm.id[i & 1][e];
                red_add_v4(&g->a, s0);
                red_add_v4(&g->b, s1v);
                atomicAdd(&g->c.x, s2);
            }
        }
    }
    cp_async_wait_all();
}

template <bool FUSED>
void launch_one(const RenderBwdArgs& a, cudaStream_t s) {
    static bool configured = false;
    const size_t bytes = sizeof(BwdSmem);
    if (!configured) {
        cudaFuncSetAttribute(render_bwd_kernel<FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        configured = true;
    }
    render_bwd_kernel<FUSED><<<dim3(a.grid_x, a.grid_y), RENDER_THREADS, bytes, s>>>(a);
}

}  // namespace

void launch_render_bwd(const RenderBwdArgs& a, cudaStream_t s) {
    if (a.dL_dpix == nullptr) launch_one<true>(a, s);
    else launch_one<false>(a, s);
}
