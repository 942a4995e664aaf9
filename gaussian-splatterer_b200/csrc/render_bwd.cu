// Backward tile renderer: per-pixel gradients of the alpha blend, reduced per (tile, splat) before
// they touch global memory.  Replaces renderCUDA<3> backward (reference
// cuda_rasterizer/backward.cu:399-557), which issues 9 global atomicAdds per blended
// (pixel, splat) pair, and -- when the loss is fused -- imageIntToLoss (src/Trainer.cu:33-44).
//
// Same tiling and staging as the forward (one CTA per tile, PPT pixels per thread, cp.async
// double-buffered 48-byte records, walked back to front).  The walk starts at the tile's largest
// n_contrib instead of the end of the list.  For every list entry the 9 partial gradients are
//   1. summed over the thread's own PPT pixels in registers,
//   2. reduced across the warp with a reduce-scatter butterfly (14 shuffles for 9 values instead of
//      45) -- skipped entirely when no lane of the warp blended the entry,
//   3. parked in a per-warp shared-memory slot,
// and after each batch one thread per entry adds the warps' slots and issues three vector
// reductions (red.global.add.v4.f32 x2 + one scalar) into the packed GradRec scratch: 3 L2
// transactions per (tile, splat) instead of 9 per (pixel, splat).

#include "common.cuh"

namespace {

constexpr int RB = 128;   // list entries per staged batch

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void red_add_v4(float4* addr, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

template <int PPT>
struct BwdSmem {
    static constexpr int NT = GSB_TILE_PIX / PPT;
    static constexpr int NW = NT / 32;
    SplatRec rec[2][RB];
    float acc[NW][RB][12];
    uint32_t mask[NW][RB / 32];
};

template <int PPT, bool FUSED_LOSS>
__global__ void __launch_bounds__(GSB_TILE_PIX / PPT) render_bwd_kernel(const RenderBwdArgs a) {
    constexpr int NT = GSB_TILE_PIX / PPT;
    constexpr int NW = NT / 32;
    constexpr int ROWS = GSB_TILE / PPT;
    constexpr int EPT = (RB + NT - 1) / NT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmem<PPT>& sm = *reinterpret_cast<BwdSmem<PPT>*>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & (GSB_TILE - 1), ty = tid >> 4;
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int start = (int)a.tile_max[tile];       // entries at or beyond this index were blended by no pixel
    if (start == 0) return;
    const uint2 range = a.ranges[tile];
    const uint32_t* list = a.point_list + range.x;
    const int rounds = (start + RB - 1) / RB;

    const int pix_x = blockIdx.x * GSB_TILE + tx;
    const float pxf = (float)pix_x;
    const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
    const size_t plane = (size_t)a.W * a.H;
    const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;

    float pyf[PPT], T[PPT], Tfin[PPT], bgdot[PPT], last_alpha[PPT];
    float dLr[PPT], dLg[PPT], dLb[PPT], lcr[PPT], lcg[PPT], lcb[PPT], arr[PPT], arg[PPT], arb[PPT];
    int ncontrib[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const int py = blockIdx.y * GSB_TILE + ty + k * ROWS;
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

    uint32_t ids[EPT];
    auto load_ids = [&](int batch) {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
            const int p = e * NT + tid;
            const int q = start - 1 - (batch * RB + p);     // back to front
            ids[e] = (p < RB && q >= 0) ? list[q] : 0xFFFFFFFFu;
        }
    };
    auto issue = [&](int buf) {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
            const int p = e * NT + tid;
            if (p < RB && ids[e] != 0xFFFFFFFFu) {
                const SplatRec* g = a.recs + ids[e];
                cp_async16(&sm.rec[buf][p].a, &g->a);
                cp_async16(&sm.rec[buf][p].b, &g->b);
                cp_async16(&sm.rec[buf][p].c, &g->c);
            }
        }
        cp_async_commit();
    };
    load_ids(0);
    issue(0);
    load_ids(1);

    for (int i = 0; i < rounds; i++) {
        cp_async_wait_all();
        __syncthreads();   // batch i visible; reduction of batch i-1 finished (acc / mask reusable)
        if (i + 1 < rounds) {
            issue((i + 1) & 1);
            load_ids(i + 2);
        }
        const SplatRec* buf = sm.rec[i & 1];
        const int cnt = min(RB, start - i * RB);
        uint32_t wmask[RB / 32];
#pragma unroll
        for (int m = 0; m < RB / 32; m++) wmask[m] = 0;

        for (int j = 0; j < cnt; j++) {
            const int idx = start - 1 - (i * RB + j);     // position in the tile's list
            const float4 ra = buf[j].a;   // x, y, conic.x, conic.y
            const float4 rb = buf[j].b;   // conic.z, opacity, r, g
            const float blue = buf[j].c.x;
            const float dx = __fsub_rn(ra.x, pxf);
            const float bx = __fmul_rn(dx, ra.z);
            const float cxy = __fmul_rn(dx, ra.w);
            float v[9];
#pragma unroll
            for (int q = 0; q < 9; q++) v[q] = 0.f;
            bool hit = false;
#pragma unroll
            for (int k = 0; k < PPT; k++) {
                if (idx >= ncontrib[k]) continue;
                const float dy = __fsub_rn(ra.y, pyf[k]);
                const float ay = __fmul_rn(dy, __fmul_rn(dy, rb.x));
                const float power = __fmaf_rn(__fmaf_rn(dx, bx, ay), -0.5f, -__fmul_rn(dy, cxy));
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
                lcr[k] = rb.z; lcg[k] = rb.w; lcb[k] = blue;
                float dL_dalpha = (rb.z - arr[k]) * dLr[k] + (rb.w - arg[k]) * dLg[k] + (blue - arb[k]) * dLb[k];
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
            wmask[j >> 5] |= 1u << (j & 31);
            // reduce-scatter butterfly over v[0..7]: afterwards lane 4q holds the warp sum of v[q]
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
            if ((lane & 3) == 0) sm.acc[warp][j][((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)] = s1;
            if (lane == 1) sm.acc[warp][j][8] = s8;
        }
        if (lane < RB / 32) sm.mask[warp][lane] = wmask[lane];   // wmask is warp-uniform
        __syncthreads();
        // one thread per entry: add the warps' slots, 3 vector reds into the packed scratch
        for (int e = tid; e < cnt; e += NT) {
            float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1v = s0;
            float s2 = 0.f;
            bool any = false;
#pragma unroll
            for (int w = 0; w < NW; w++) {
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
                GradRec* g = a.grads + __float_as_int(buf[e].c.z);
                red_add_v4(&g->a, s0);
                red_add_v4(&g->b, s1v);
                atomicAdd(&g->c.x, s2);
            }
        }
    }
    cp_async_wait_all();
}

template <int PPT, bool FUSED>
void launch_one(const RenderBwdArgs& a, cudaStream_t s) {
    static bool configured = false;
    const size_t bytes = sizeof(BwdSmem<PPT>);
    if (!configured) {
        cudaFuncSetAttribute(render_bwd_kernel<PPT, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        configured = true;
    }
    render_bwd_kernel<PPT, FUSED><<<dim3(a.grid_x, a.grid_y), GSB_TILE_PIX / PPT, bytes, s>>>(a);
}

}  // namespace

void launch_render_bwd(const RenderBwdArgs& a, cudaStream_t s) {
    const bool fused = a.dL_dpix == nullptr;
    switch (a.ppt) {
        case 1: fused ? launch_one<1, true>(a, s) : launch_one<1, false>(a, s); break;
        case 4: fused ? launch_one<4, true>(a, s) : launch_one<4, false>(a, s); break;
        default: fused ? launch_one<2, true>(a, s) : launch_one<2, false>(a, s); break;
    }
}
