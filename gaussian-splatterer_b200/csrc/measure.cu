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

}  // namespace

void launch_count_pairs(const RenderFwdArgs& a, unsigned long long* out, cudaStream_t s) {
    count_pairs_kernel<<<dim3(a.grid_x, a.grid_y), 256, 0, s>>>(a, out);
}
