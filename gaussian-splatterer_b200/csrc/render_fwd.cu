// Forward tile renderer: front-to-back alpha blending of each tile's depth-ordered splat list.
// Replaces renderCUDA<3> forward (reference cuda_rasterizer/forward.cu:261-374).
//
// One CTA per 16x16 tile; each thread owns PPT pixels of one column so a staged splat record is
// read from shared memory once and reused PPT times.  Records (48 B, see SplatRec) are gathered
// into a double-buffered shared-memory ring with 16-byte cp.async copies issued one batch ahead,
// and the list indices for the batch after that are prefetched into registers, so neither the
// dependent gather nor the index load sits on the critical path of the blend loop.
// The per-pair arithmetic (power, exp, alpha, the 1/255 and 1e-4 tests, the colour FMA order) is
// pinned with explicit rn intrinsics to what the reference build executes, so every blend /
// skip / stop decision and therefore n_contrib is identical and images agree to float rounding.

#include "common.cuh"

namespace {

constexpr int RB = 128;   // list entries per staged batch

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int PPT>
__global__ void __launch_bounds__(GSB_TILE_PIX / PPT) render_fwd_kernel(const RenderFwdArgs a) {
    constexpr int NT = GSB_TILE_PIX / PPT;      // threads per CTA
    constexpr int ROWS = GSB_TILE / PPT;        // thread rows
    constexpr int EPT = (RB + NT - 1) / NT;     // staged entries per thread per batch
    __shared__ SplatRec s_rec[2][RB];
    __shared__ uint32_t s_max;

    const int tid = threadIdx.x;
    const int tx = tid & (GSB_TILE - 1), ty = tid >> 4;
    const int tile = blockIdx.y * a.grid_x + blockIdx.x;
    const int pix_x = blockIdx.x * GSB_TILE + tx;
    const float pxf = (float)pix_x;
    int pix_y[PPT];
    float pyf[PPT];
    bool done[PPT];
    float T[PPT], Cr[PPT], Cg[PPT], Cb[PPT];
    uint32_t last[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        pix_y[k] = blockIdx.y * GSB_TILE + ty + k * ROWS;
        pyf[k] = (float)pix_y[k];
        done[k] = !(pix_x < a.W && pix_y[k] < a.H);
        T[k] = 1.0f; Cr[k] = Cg[k] = Cb[k] = 0.f;
        last[k] = 0;
    }
    if (tid == 0) s_max = 0;

    const uint2 range = a.ranges[tile];
    const int len = (int)(range.y - range.x);
    const int rounds = (len + RB - 1) / RB;
    const uint32_t* list = a.point_list + range.x;

    // prologue: ids of batch 0 -> gather batch 0; ids of batch 1 -> registers
    uint32_t ids[EPT];
#pragma unroll
    for (int e = 0; e < EPT; e++) {
        const int p = e * NT + tid;
        ids[e] = (p < RB && p < len) ? list[p] : 0xFFFFFFFFu;
    }
    auto issue = [&](int buf) {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
            const int p = e * NT + tid;
            if (p < RB && ids[e] != 0xFFFFFFFFu) {
                const SplatRec* g = a.recs + ids[e];
                cp_async16(&s_rec[buf][p].a, &g->a);
                cp_async16(&s_rec[buf][p].b, &g->b);
                cp_async16(&s_rec[buf][p].c, &g->c);
            }
        }
        cp_async_commit();
    };
    auto load_ids = [&](int batch) {
#pragma unroll
        for (int e = 0; e < EPT; e++) {
            const int p = e * NT + tid;
            const int q = batch * RB + p;
            ids[e] = (p < RB && q < len) ? list[q] : 0xFFFFFFFFu;
        }
    };
    if (rounds > 0) {
        issue(0);
        load_ids(1);
    }

    for (int i = 0; i < rounds; i++) {
        bool all = true;
#pragma unroll
        for (int k = 0; k < PPT; k++) all = all && done[k];
        cp_async_wait_all();
        // batch i is visible, batch i-1 fully consumed; vote like the reference's __syncthreads_count
        if (__syncthreads_and(all)) break;
        if (i + 1 < rounds) {
            issue((i + 1) & 1);
            load_ids(i + 2);
        }
        const SplatRec* buf = s_rec[i & 1];
        const int cnt = min(RB, len - i * RB);
        for (int j = 0; j < cnt; j++) {
            const float4 ra = buf[j].a;   // x, y, conic.x, conic.y
            const float4 rb = buf[j].b;   // conic.z, opacity, r, g
            const float dx = __fsub_rn(ra.x, pxf);
            const float bx = __fmul_rn(dx, ra.z);
            const float cxy = __fmul_rn(dx, ra.w);
#pragma unroll
            for (int k = 0; k < PPT; k++) {
                if (done[k]) continue;
                const float dy = __fsub_rn(ra.y, pyf[k]);
                // power = -0.5 (cx dx^2 + cz dy^2) - cy dx dy, in the reference's evaluation order
                const float ay = __fmul_rn(dy, __fmul_rn(dy, rb.x));
                const float power = __fmaf_rn(__fmaf_rn(dx, bx, ay), -0.5f, -__fmul_rn(dy, cxy));
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, __fmul_rn(rb.y, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = __fmul_rn(T[k], __fsub_rn(1.0f, alpha));
                if (test_T < 0.0001f) {
                    done[k] = true;
                    continue;
                }
                const float blue = buf[j].c.x;
                Cr[k] = __fmaf_rn(T[k], __fmul_rn(alpha, rb.z), Cr[k]);
                Cg[k] = __fmaf_rn(T[k], __fmul_rn(alpha, rb.w), Cg[k]);
                Cb[k] = __fmaf_rn(T[k], __fmul_rn(alpha, blue), Cb[k]);
                T[k] = test_T;
                last[k] = (uint32_t)(i * RB + j + 1);
            }
        }
    }
    cp_async_wait_all();

    const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
    const size_t plane = (size_t)a.W * a.H;
    uint32_t mx = 0;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        if (pix_x < a.W && pix_y[k] < a.H) {
            const size_t pid = (size_t)pix_y[k] * a.W + pix_x;
            a.final_T[pid] = T[k];
            a.n_contrib[pid] = last[k];
            a.out_color[pid] = __fmaf_rn(bg0, T[k], Cr[k]);
            a.out_color[plane + pid] = __fmaf_rn(bg1, T[k], Cg[k]);
            a.out_color[2 * plane + pid] = __fmaf_rn(bg2, T[k], Cb[k]);
            mx = max(mx, last[k]);
        }
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    __syncthreads();
    if ((tid & 31) == 0 && mx) atomicMax(&s_max, mx);
    __syncthreads();
    if (tid == 0) a.tile_max[tile] = s_max;
}

}  // namespace

void launch_render_fwd(const RenderFwdArgs& a, cudaStream_t s) {
    const dim3 grid(a.grid_x, a.grid_y);
    switch (a.ppt) {
        case 1: render_fwd_kernel<1><<<grid, 256, 0, s>>>(a); break;
        case 4: render_fwd_kernel<4><<<grid, 64, 0, s>>>(a); break;
        default: render_fwd_kernel<2><<<grid, 128, 0, s>>>(a); break;
    }
}
