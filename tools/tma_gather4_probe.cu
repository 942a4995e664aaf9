// Probe for cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4 (PTX ISA 8.6, sm_100) on the record array of the render
// kernels: [P] rows of 12 floats (48 B).  There is no documentation of the tensor-map shape the mode wants in this
// image, so the shapes are tried one per process (a wrong one faults the context):
//     tma_gather4_probe <variant>      variant 0: dims {12, P}, box {16, 1} (rows padded to 64 B by out-of-bounds fill)
//                                      variant 1: dims {12, P}, box {12, 1}
//                                      variant 2: dims {12, P}, box {16, 4}
//                                      variant 3: dims {12, P}, box {12, 4}
//                                      variant 4 / 5: box {12, 1}, destination 192 B / 48 B into the buffer (64- / 16-byte aligned)
// One warp gathers rows {5, 1000, 3, 77777} and prints what landed in shared memory.  The wait is bounded: a copy that
// never completes is reported, not waited for.
// build: nvcc -std=c++17 -O2 -gencode arch=compute_100a,code=sm_100a tools/tma_gather4_probe.cu -o tools/tma_gather4_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("{\"error\": \"%s at %s:%d\"}\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

__global__ void probe_kernel(const __grid_constant__ CUtensorMap tmap, int r0, int r1, int r2, int r3, uint32_t tx_bytes,
                             float* out, int* status, uint32_t dst_off) {
    __shared__ __align__(128) float dst[4 * 16 + 64];
    __shared__ __align__(8) uint64_t bar;
    const int lane = threadIdx.x;
    for (int i = lane; i < 4 * 16 + 64; i += 32) dst[i] = -1.0f;
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst_a = (uint32_t)__cvta_generic_to_shared(dst) + dst_off;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    if (lane == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(tx_bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
            ::"r"(dst_a), "l"(&tmap), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar_a) : "memory");
    }
    int ok = 0;
    for (int it = 0; it < 200000 && !ok; it++) {
        uint32_t p;
        asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }"
                     : "=r"(p) : "r"(bar_a), "r"(0) : "memory");
        ok = (int)p;
    }
    __syncwarp();
    if (lane == 0) *status = ok;
    for (int i = lane; i < 4 * 16 + 64; i += 32) out[i] = dst[i];
}

int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    const int P = 100000;
    std::vector<float> h((size_t)P * 12 + 16);
    for (size_t i = 0; i < h.size(); i++) h[i] = (float)(i / 12) + 0.01f * (float)(i % 12);   // row + 0.01 * column
    float* d = nullptr;
    CK(cudaMalloc(&d, h.size() * sizeof(float)));
    CK(cudaMemcpy(d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));

    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn) { printf("{\"error\": \"no cuTensorMapEncodeTiled\"}\n"); return 2; }
    const cuuint32_t boxc = (variant == 0 || variant == 2) ? 16 : 12, boxr = (variant == 2 || variant == 3) ? 4 : 1;
    const uint32_t dst_off = variant == 4 ? 192u : (variant == 5 ? 48u : 0u);
    CUtensorMap tm;
    const cuuint64_t dims[2] = {12, (cuuint64_t)P};
    const cuuint64_t strides[1] = {48};
    const cuuint32_t box[2] = {boxc, boxr};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = ((EncodeFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("{\"variant\": %d, \"encode_error\": %d}\n", variant, (int)r); return 1; }
    float* out = nullptr;
    int* st = nullptr;
    CK(cudaMalloc(&out, 128 * sizeof(float)));
    CK(cudaMalloc(&st, sizeof(int)));
    CK(cudaMemset(st, 0xFF, sizeof(int)));
    probe_kernel<<<1, 32>>>(tm, 5, 1000, 3, 77777, 4 * boxc * 4, out, st, dst_off);
    const cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("{\"variant\": %d, \"box\": [%u, %u], \"kernel_error\": \"%s\"}\n", variant, boxc, boxr, cudaGetErrorString(e)); return 1; }
    float ho[128];
    int hs = 0;
    CK(cudaMemcpy(ho, out, sizeof(ho), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&hs, st, sizeof(int), cudaMemcpyDeviceToHost));
    printf("{\"variant\": %d, \"box\": [%u, %u], \"dst_offset\": %u, \"completed\": %d, \"smem\": [", variant, boxc, boxr, dst_off, hs);
    for (int i = 0; i < 128; i++) printf("%s%.2f", i ? ", " : "", ho[i]);
    printf("]}\n");
    return 0;
}
