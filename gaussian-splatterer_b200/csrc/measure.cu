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

void launch_count_pairs(const RenderFwdArgs& a, unsigned long long* out, cudaStream_t s) {
    count_pairs_kernel<<<dim3(a.grid_x, a.grid_y), 256, 0, s>>>(a, out);
}
