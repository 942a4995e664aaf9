// Binning: produce, per 16x16 tile, the list of splats that touch it, ordered by depth.
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs +
// identifyTileRanges (reference cuda_rasterizer/rasterizer_impl.cu:70-138,277-318).
//
// The reference sorts R (tile<<32 | depth_bits, splat) pairs with a stable 64-bit LSD sort.
// Because that sort is stable and the key is (tile, depth), the same order is obtained by
//   1. a stable sort of the P splats by depth bits            (4 x 8-bit passes over P items)
//   2. emitting the duplicates in that order                   (tile id, splat)
//   3. a stable sort of the R duplicates by tile id only       (ceil(tile_bits / 8) passes)
// which moves 8-byte pairs through 2 passes over R instead of 12-byte pairs through 6.
// Ties (same tile, same depth) stay in splat-index order exactly as with the reference's sort.
//
// Both sorts use one hand-written onesweep pass kernel: a single read of the input per digit,
// chained-scan decoupled look-back across tiles for the digit offsets, warp match_any ranking for
// stability.  Tile ids are handed out by an atomic ticket so a tile only ever waits on tiles that
// already started (forward progress without co-residency assumptions).  Look-back words carry an
// epoch so the status array never needs clearing between passes.

#include "common.cuh"

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
#ifndef GSB_RS_IPT
#define GSB_RS_IPT 16
#endif
constexpr int RS_IPT = GSB_RS_IPT;
constexpr int RS_TILE = RS_THREADS * RS_IPT;   // items per tile

constexpr uint32_t LB_AGG = 1u, LB_INCL = 2u;

__device__ __forceinline__ unsigned long long lb_pack(uint32_t epoch, uint32_t flag, uint32_t value) {
    return ((unsigned long long)((epoch << 2) | flag) << 32) | value;
}
__device__ __forceinline__ unsigned long long lb_load(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void lb_store(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Chained look-back for one counter: publish `total`, return the sum over all earlier tiles.
__device__ __forceinline__ uint32_t lookback_exclusive(unsigned long long* lb, uint32_t stride, uint32_t slot,
                                                       uint32_t tile, uint32_t epoch, uint32_t total) {
    lb_store(lb + (size_t)tile * stride + slot, lb_pack(epoch, tile == 0 ? LB_INCL : LB_AGG, total));
    uint32_t excl = 0;
    for (int t = (int)tile - 1; t >= 0; --t) {
        unsigned long long v;
        uint32_t tag;
        do {
            v = lb_load(lb + (size_t)t * stride + slot);
            tag = (uint32_t)(v >> 32);
        } while ((tag >> 2) != epoch || (tag & 3u) == 0u);
        excl += (uint32_t)v;
        if ((tag & 3u) == LB_INCL) break;
    }
    if (tile != 0) lb_store(lb + (size_t)tile * stride + slot, lb_pack(epoch, LB_INCL, excl + total));
    return excl;
}

// Digit histograms of up to 4 passes in one read of the keys.
template <int NPASS>
__global__ void __launch_bounds__(256) radix_hist_kernel(const uint32_t* __restrict__ keys, const uint32_t* n_ptr,
                                                         uint32_t n_imm, uint32_t n_cap, int shift0,
                                                         uint32_t* __restrict__ g_hist) {
    __shared__ uint32_t s_hist[NPASS][256];
    const uint32_t n = n_ptr ? min(*n_ptr, n_cap) : n_imm;
    for (int p = 0; p < NPASS; p++) s_hist[p][threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t k = keys[i];
#pragma unroll
        for (int p = 0; p < NPASS; p++) atomicAdd(&s_hist[p][(k >> (shift0 + 8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int p = 0; p < NPASS; p++) {
        const uint32_t c = s_hist[p][threadIdx.x];
        if (c) atomicAdd(&g_hist[p * 256 + threadIdx.x], c);
    }
}

// One stable 8-bit LSD pass over (key, value) pairs.
__global__ void __launch_bounds__(RS_THREADS) onesweep_pass_kernel(
    const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, const uint32_t* n_ptr, uint32_t n_imm, uint32_t n_cap, int shift,
    const uint32_t* __restrict__ g_hist, unsigned long long* lookback, uint32_t epoch, uint32_t* ticket) {
    __shared__ uint32_t s_cnt[RS_WARPS][256];
    __shared__ uint32_t s_base[256];
    __shared__ uint32_t s_scan[RS_WARPS];
    __shared__ uint32_t s_tile;

    const uint32_t n = n_ptr ? min(*n_ptr, n_cap) : n_imm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) s_cnt[w][tid] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    if ((unsigned long long)tile * RS_TILE >= n) return;

    // Exclusive prefix of the global digit histogram (each block redoes this 256-wide scan).
    uint32_t gx;
    {
        const uint32_t h = g_hist[tid];
        uint32_t inc = h;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_scan[warp] = inc;
        __syncthreads();
        uint32_t wp = 0;
        for (int w = 0; w < warp; w++) wp += s_scan[w];
        gx = wp + inc - h;
    }

    // Warp-contiguous blocked-by-round arrangement keeps memory order == rank order.
    const uint32_t base = tile * RS_TILE + warp * (32 * RS_IPT) + lane;
    uint32_t key[RS_IPT], val[RS_IPT], rank[RS_IPT];
#pragma unroll
    for (int r = 0; r < RS_IPT; r++) {
        const uint32_t i = base + r * 32;
        key[r] = (i < n) ? keys_in[i] : 0xFFFFFFFFu;
        val[r] = (i < n) ? vals_in[i] : 0u;
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_IPT; r++) {
        const uint32_t d = (key[r] >> shift) & 255u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader) {
            old = s_cnt[warp][d];
            s_cnt[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = old + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();

    // Digit `tid`: exclusive scan over warps, then the cross-tile look-back.
    {
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t c = s_cnt[w][tid];
            s_cnt[w][tid] = sum;
            sum += c;
        }
        const uint32_t excl = lookback_exclusive(lookback, 256, tid, tile, epoch, sum);
        s_base[tid] = gx + excl;
    }
    __syncthreads();

#pragma unroll
    for (int r = 0; r < RS_IPT; r++) {
        const uint32_t i = base + r * 32;
        if (i < n) {
            const uint32_t d = (key[r] >> shift) & 255u;
            const uint32_t pos = s_base[d] + s_cnt[warp][d] + rank[r];
            keys_out[pos] = key[r];
            vals_out[pos] = val[r];
        }
    }
}

// Exclusive scan of tiles_touched taken in depth order; total -> ctrl->num_rendered.
constexpr int SC_IPT = 8;
constexpr int SC_TILE = 256 * SC_IPT;

__global__ void __launch_bounds__(256) scan_offsets_kernel(const uint32_t* __restrict__ order,
                                                           const uint32_t* __restrict__ tiles_touched, uint32_t P,
                                                           uint32_t* __restrict__ offsets, unsigned long long* lookback,
                                                           uint32_t epoch, uint32_t* ticket, GsbCtrl* ctrl,
                                                           uint32_t r_cap, uint32_t* num_rendered_host) {
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_tile, s_excl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile * SC_TILE >= P) return;
    const uint32_t base = tile * SC_TILE + tid * SC_IPT;
    uint32_t v[SC_IPT], sum = 0, vis = 0;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        const uint32_t i = base + k;
        v[k] = (i < P) ? tiles_touched[order[i]] : 0u;
        sum += v[k];
        vis += (v[k] != 0u);
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    vis = __reduce_add_sync(0xffffffffu, vis);
    if (lane == 0 && vis) atomicAdd(&ctrl->num_visible, vis);
    __syncthreads();
    uint32_t wp = 0, total = 0;
    for (int w = 0; w < 8; w++) {
        if (w < warp) wp += s_warp[w];
        total += s_warp[w];
    }
    if (tid == 0) s_excl = lookback_exclusive(lookback, 256, 0, tile, epoch, total);
    __syncthreads();
    uint32_t run = s_excl + wp + inc - sum;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        const uint32_t i = base + k;
        if (i < P) offsets[i] = run;
        run += v[k];
    }
    if ((tile + 1) * SC_TILE >= P && tid == 255) {   // last tile, last thread: run == R
        ctrl->num_rendered = run;
        if (run > r_cap) ctrl->overflow = 1u;
        if (num_rendered_host) *num_rendered_host = run;
    }
}

// Duplicate emission in depth order (duplicateWithKeys, rasterizer_impl.cu:70-111): the rectangle
// is recomputed from the stored pixel centre and integer radius exactly as the reference does.
__global__ void __launch_bounds__(256) emit_kernel(const uint32_t* __restrict__ order,
                                                   const uint32_t* __restrict__ offsets,
                                                   const uint32_t* __restrict__ tiles_touched,
                                                   const int* __restrict__ radii, const SplatRec* __restrict__ recs,
                                                   uint32_t P, int grid_x, int grid_y, uint32_t r_cap,
                                                   uint32_t* __restrict__ out_keys, uint32_t* __restrict__ out_vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t s = order[i];
    if (tiles_touched[s] == 0u) return;
    uint32_t off = offsets[i];
    const float4 ra = recs[s].a;
    const float rf = (float)radii[s];
    const int x0 = (int)__fmul_rn(__fsub_rn(ra.x, rf), 0.0625f), y0 = (int)__fmul_rn(__fsub_rn(ra.y, rf), 0.0625f);
    const int x1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.x, rf), 16.0f), -1.0f), 0.0625f);
    const int y1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.y, rf), 16.0f), -1.0f), 0.0625f);
    const uint32_t gx = (uint32_t)grid_x, gy = (uint32_t)grid_y;
    const uint32_t rx0 = min(gx, (uint32_t)max(0, x0)), ry0 = min(gy, (uint32_t)max(0, y0));
    const uint32_t rx1 = min(gx, (uint32_t)max(0, x1)), ry1 = min(gy, (uint32_t)max(0, y1));
    for (uint32_t y = ry0; y < ry1; y++)
        for (uint32_t x = rx0; x < rx1; x++) {
            if (off < r_cap) {
                out_keys[off] = y * gx + x;
                out_vals[off] = s;
            }
            off++;
        }
}

// identifyTileRanges (rasterizer_impl.cu:116-138) on the sorted tile ids.
__global__ void __launch_bounds__(256) tile_ranges_kernel(const uint32_t* __restrict__ keys, const uint32_t* n_ptr,
                                                          uint32_t n_cap, uint2* __restrict__ ranges) {
    const uint32_t n = min(*n_ptr, n_cap);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t cur = keys[i];
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = keys[i - 1];
            if (cur != prev) {
                ranges[prev].y = i;
                ranges[cur].x = i;
            }
        }
        if (i == n - 1) ranges[cur].y = n;
    }
}

}  // namespace

void launch_depth_sort(const BinningArgs& a, cudaStream_t s) {
    const uint32_t P = (uint32_t)a.P;
    if (P == 0) return;
    // stable sort of splats by depth bits: 4 passes, result back in pkeys[0]/pvals[0]
    radix_hist_kernel<4><<<min(gsb_div_up(P, 256 * 8), 592), 256, 0, s>>>(a.pkeys[0], nullptr, P, P, 0, &a.ctrl->hist[0][0]);
    const int tiles = gsb_div_up(P, RS_TILE);
    for (int p = 0; p < 4; p++) {
        const int src = p & 1, dst = src ^ 1;
        onesweep_pass_kernel<<<tiles, RS_THREADS, 0, s>>>(a.pkeys[src], a.pvals[src], a.pkeys[dst], a.pvals[dst], nullptr, P, P,
                                                          8 * p, &a.ctrl->hist[p][0], a.lookback, ++(*a.epoch),
                                                          &a.ctrl->tickets[p]);
    }
}

void launch_scan_offsets(const BinningArgs& a, cudaStream_t s) {
    const uint32_t P = (uint32_t)a.P;
    if (P == 0) return;
    // offsets of every splat's duplicates, in depth order
    scan_offsets_kernel<<<gsb_div_up(P, SC_TILE), 256, 0, s>>>(a.pvals[0], a.tiles_touched, P, a.offsets, a.lookback,
                                                               ++(*a.epoch), &a.ctrl->tickets[4], a.ctrl, a.r_cap,
                                                               a.num_rendered_host);
}

void launch_emit(const BinningArgs& a, cudaStream_t s) {
    const uint32_t P = (uint32_t)a.P;
    if (P == 0 || a.r_cap == 0) return;
    emit_kernel<<<gsb_div_up(P, 256), 256, 0, s>>>(a.pvals[0], a.offsets, a.tiles_touched, a.radii, a.recs, P, a.grid_x,
                                                   a.grid_y, a.r_cap, a.rkeys[0], a.rvals[0]);
}

int gsb_tile_sort_passes(int tile_bits) { return max(1, (tile_bits + 7) / 8); }

int launch_tile_sort(const BinningArgs& a, cudaStream_t s) {
    const int npass = gsb_tile_sort_passes(a.tile_bits);
    if (a.P == 0 || a.r_cap == 0) return npass & 1;
    // stable sort of the duplicates by tile id
    const int grid_cap = gsb_div_up(a.r_cap, RS_TILE);
    const int hist_grid = min(max(grid_cap / 2, 1), 592);
    switch (npass) {
        case 1: radix_hist_kernel<1><<<hist_grid, 256, 0, s>>>(a.rkeys[0], &a.ctrl->num_rendered, 0, a.r_cap, 0, &a.ctrl->hist[4][0]); break;
        case 2: radix_hist_kernel<2><<<hist_grid, 256, 0, s>>>(a.rkeys[0], &a.ctrl->num_rendered, 0, a.r_cap, 0, &a.ctrl->hist[4][0]); break;
        case 3: radix_hist_kernel<3><<<hist_grid, 256, 0, s>>>(a.rkeys[0], &a.ctrl->num_rendered, 0, a.r_cap, 0, &a.ctrl->hist[4][0]); break;
        default: radix_hist_kernel<4><<<hist_grid, 256, 0, s>>>(a.rkeys[0], &a.ctrl->num_rendered, 0, a.r_cap, 0, &a.ctrl->hist[4][0]); break;
    }
    for (int p = 0; p < npass; p++) {
        const int src = p & 1, dst = src ^ 1;
        onesweep_pass_kernel<<<grid_cap, RS_THREADS, 0, s>>>(a.rkeys[src], a.rvals[src], a.rkeys[dst], a.rvals[dst],
                                                             &a.ctrl->num_rendered, 0, a.r_cap, 8 * p, &a.ctrl->hist[4 + p][0],
                                                             a.lookback, ++(*a.epoch), &a.ctrl->tickets[5 + p]);
    }
    return npass & 1;
}

void launch_tile_ranges(const BinningArgs& a, int fin, cudaStream_t s) {
    if (a.P == 0 || a.r_cap == 0) return;
    const int grid_cap = gsb_div_up(a.r_cap, RS_TILE);
    tile_ranges_kernel<<<min(max(grid_cap * 4, 1), 2368), 256, 0, s>>>(a.rkeys[fin], &a.ctrl->num_rendered, a.r_cap, a.ranges);
}

void launch_binning_depth_and_scan(const BinningArgs& a, cudaStream_t s) {
    launch_depth_sort(a, s);
    launch_scan_offsets(a, s);
}

int launch_binning_emit_and_sort(const BinningArgs& a, cudaStream_t s) {
    launch_emit(a, s);
    const int fin = launch_tile_sort(a, s);
    launch_tile_ranges(a, fin, s);
    return fin;
}
