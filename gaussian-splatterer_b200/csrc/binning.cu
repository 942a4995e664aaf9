// Binning: produce, per 16x16 tile, the list of splats that touch it, ordered by (depth, splat index).
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs +
// identifyTileRanges (reference cuda_rasterizer/rasterizer_impl.cu:70-138,277-318).
//
// The reference sorts R (tile << 32 | depth_bits, splat) pairs with one stable 64-bit LSD radix sort over all R
// duplicates (6 onesweep passes on sm_100).  The key's high half is the tile, so the result is nothing but one
// independent sorted list per tile; within a tile the order is ascending depth bits, ties (equal depth) in emission
// order = ascending splat index (duplicateWithKeys emits in index order and the sort is stable).  That order is
// reproduced without any global sort:
//   1. preprocess counts, per tile, the splats whose rectangle covers it           (tile_count, one RED per duplicate)
//   2. tile_scan_kernel: exclusive scan over the tiles -> ranges (= identifyTileRanges' output) and per-tile cursors
//   3. emit_kernel: every duplicate claims a slot in its tile's range (atomic cursor) and stores depth_bits << 32 | splat
//   4. tile_sort_kernel: one CTA per tile sorts its own list by that 64-bit key (bitonic network in shared memory;
//      lists longer than the shared-memory capacity fall back to a hybrid shared/global network) and writes the
//      splat ids.  The key is unique within a tile, so the unordered claims of step 3 cannot show in the result.
// Round-1's decomposed radix sort (4 passes over P + 2 passes over R, fourteen dependent launches, 0.27 ms per view
// at C4) is replaced by three kernels that move 20 bytes per duplicate.

#include "common.cuh"

namespace {

constexpr uint32_t LB_AGG = 1u, LB_INCL = 2u;

// Look-back status word: (30-bit epoch << 2 | flag) in the high half, the value in the low half.  The host keeps the
// epoch below 2^30 (trainer_api.cu epoch_guard), so a stale word of an earlier launch can never match.
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

// Chained look-back for one counter: publish `total`, return the sum over all earlier scan tiles.  Scan-tile ids are
// handed out by an atomic ticket, so a tile only ever waits on tiles that already started.
__device__ __forceinline__ uint32_t lookback_exclusive(unsigned long long* lb, uint32_t tile, uint32_t epoch, uint32_t total) {
    lb_store(lb + tile, lb_pack(epoch, tile == 0 ? LB_INCL : LB_AGG, total));
    uint32_t excl = 0;
    for (int t = (int)tile - 1; t >= 0; --t) {
        unsigned long long v;
        uint32_t tag;
        do {
            v = lb_load(lb + t);
            tag = (uint32_t)(v >> 32);
        } while ((tag >> 2) != epoch || (tag & 3u) == 0u);
        excl += (uint32_t)v;
        if ((tag & 3u) == LB_INCL) break;
    }
    if (tile != 0) lb_store(lb + tile, lb_pack(epoch, LB_INCL, excl + total));
    return excl;
}

// Exclusive scan of the per-tile duplicate counts.  ranges[t] = [start, start + count) (empty tiles keep (0, 0) like
// the reference, whose identifyTileRanges never writes them); tile_count[t] becomes the tile's write cursor for
// emit_kernel; the total R goes to ctrl->num_rendered and, optionally, to pinned host memory.  If R exceeds the
// capacity of the duplicate arrays the ranges are clipped to it, so every later kernel stays in bounds, and
// ctrl->overflow is raised (the caller grows the arena and repeats the pass).
constexpr int SC_IPT = 8;
constexpr int SC_TILE = 256 * SC_IPT;

__global__ void __launch_bounds__(256) tile_scan_kernel(uint32_t* __restrict__ tile_count, uint32_t tiles, uint2* __restrict__ ranges,
                                                        unsigned long long* lookback, uint32_t epoch, GsbCtrl* ctrl,
                                                        uint32_t r_cap, uint32_t* num_rendered_host) {
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_tile, s_excl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&ctrl->ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile * SC_TILE >= tiles) return;
    const uint32_t base = tile * SC_TILE + tid * SC_IPT;
    uint32_t v[SC_IPT], sum = 0;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        const uint32_t i = base + k;
        v[k] = (i < tiles) ? tile_count[i] : 0u;
        sum += v[k];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wp = 0, total = 0;
    for (int w = 0; w < 8; w++) {
        if (w < warp) wp += s_warp[w];
        total += s_warp[w];
    }
    if (tid == 0) s_excl = lookback_exclusive(lookback, tile, epoch, total);
    __syncthreads();
    uint32_t run = s_excl + wp + inc - sum;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        const uint32_t i = base + k;
        if (i < tiles) {
            ranges[i] = v[k] ? make_uint2(min(run, r_cap), min(run + v[k], r_cap)) : make_uint2(0u, 0u);
            tile_count[i] = run;
        }
        run += v[k];
    }
    if ((tile + 1) * SC_TILE >= tiles && tid == 255) {   // last scan tile, last thread: run == R
        ctrl->num_rendered = run;
        if (run > r_cap) ctrl->overflow = 1u;
        if (num_rendered_host) *num_rendered_host = run;
    }
}

// Duplicate emission (duplicateWithKeys, rasterizer_impl.cu:70-111): the rectangle is recomputed from the stored
// pixel centre and integer radius exactly as the reference does; each duplicate claims the next free slot of its tile.
__global__ void __launch_bounds__(256) emit_kernel(const uint32_t* __restrict__ tiles_touched, const int* __restrict__ radii,
                                                   const float* __restrict__ depths, const SplatRec* __restrict__ recs,
                                                   uint32_t P, int grid_x, int grid_y, uint32_t r_cap,
                                                   uint32_t* __restrict__ cursor, unsigned long long* __restrict__ pairs) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P) return;
    if (tiles_touched[s] == 0u) return;
    const float4 ra = recs[s].a;
    const float rf = (float)radii[s];
    const unsigned long long key = ((unsigned long long)__float_as_uint(depths[s]) << 32) | s;
    const int x0 = (int)__fmul_rn(__fsub_rn(ra.x, rf), 0.0625f), y0 = (int)__fmul_rn(__fsub_rn(ra.y, rf), 0.0625f);
    const int x1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.x, rf), 16.0f), -1.0f), 0.0625f);
    const int y1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.y, rf), 16.0f), -1.0f), 0.0625f);
    const uint32_t gx = (uint32_t)grid_x, gy = (uint32_t)grid_y;
    const uint32_t rx0 = min(gx, (uint32_t)max(0, x0)), ry0 = min(gy, (uint32_t)max(0, y0));
    const uint32_t rx1 = min(gx, (uint32_t)max(0, x1)), ry1 = min(gy, (uint32_t)max(0, y1));
    for (uint32_t y = ry0; y < ry1; y++)
        for (uint32_t x = rx0; x < rx1; x++) {
            const uint32_t pos = atomicAdd(&cursor[y * gx + x], 1u);
            if (pos < r_cap) pairs[pos] = key;
        }
}

// ---- per-tile sort ---------------------------------------------------------------------------------------------
// Bitonic network in its "always ascending" form: the first sub-stage of merge level k compares element o of a
// k-block with element k-1-o (mirror), the remaining sub-stages compare i with i + j.  Every comparison puts the
// smaller key at the lower index, so virtual +inf padding above the list length never moves and a compare-exchange
// whose upper index is beyond the list is simply skipped.
constexpr int TS_THREADS = 256;
constexpr int TS_CAP = 4096;          // keys per shared-memory chunk (32 KB)

__device__ __forceinline__ void ce(unsigned long long& a, unsigned long long& b) {
    const unsigned long long lo = min(a, b), hi = max(a, b);
    a = lo; b = hi;
}

// sub-stages j = j0, j0/2, ..., 1 of the plain (i, i + j) form on m keys in shared memory (m a power of two)
__device__ __forceinline__ void smem_plain_stages(unsigned long long* s, int m, int j0) {
    for (int j = j0; j > 0; j >>= 1) {
        for (int i = threadIdx.x; i < (m >> 1); i += TS_THREADS) {
            const int lo = 2 * i - (i & (j - 1));
            ce(s[lo], s[lo + j]);
        }
        __syncthreads();
    }
}

// full sort of m keys in shared memory (m a power of two, padded with ~0)
__device__ __forceinline__ void smem_sort(unsigned long long* s, int m) {
    for (int k = 2; k <= m; k <<= 1) {
        const int half = k >> 1;
        for (int i = threadIdx.x; i < (m >> 1); i += TS_THREADS) {
            const int b = i / half, o = i - b * half;
            const int lo = b * k + o;
            ce(s[lo], s[b * k + k - 1 - o]);
        }
        __syncthreads();
        smem_plain_stages(s, m, half >> 1);
    }
}

__global__ void __launch_bounds__(TS_THREADS) tile_sort_kernel(const uint2* __restrict__ ranges, unsigned long long* __restrict__ pairs,
                                                               uint32_t* __restrict__ point_list, uint32_t* __restrict__ tile_count,
                                                               uint32_t* __restrict__ tile_max) {
    __shared__ unsigned long long s[TS_CAP];
    const uint32_t tile = blockIdx.x;
    const int tid = threadIdx.x;
    // housekeeping for the kernels around this one: the count / cursor of this tile starts the next view pass at zero,
    // and the render forward takes the maximum of n_contrib over the tile with atomicMax
    if (tid == 0) { tile_count[tile] = 0u; tile_max[tile] = 0u; }
    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    if (n <= 0) return;
    unsigned long long* keys = pairs + range.x;
    uint32_t* out = point_list + range.x;
    if (n == 1) {
        if (tid == 0) out[0] = (uint32_t)keys[0];
        return;
    }
    int m = 2;
    while (m < n) m <<= 1;
    if (m <= TS_CAP) {
        for (int i = tid; i < m; i += TS_THREADS) s[i] = (i < n) ? keys[i] : ~0ull;
        __syncthreads();
        smem_sort(s, m);
        for (int i = tid; i < n; i += TS_THREADS) out[i] = (uint32_t)s[i];
        return;
    }
    // Long list: sort TS_CAP-key chunks in shared memory, then merge levels above TS_CAP with their wide sub-stages
    // (partner at least TS_CAP away) done in place in global memory and the narrow ones per chunk in shared memory.
    const int chunks = m / TS_CAP;
    for (int c = 0; c < chunks; c++) {
        const int base = c * TS_CAP;
        if (base >= n) break;
        for (int i = tid; i < TS_CAP; i += TS_THREADS) s[i] = (base + i < n) ? keys[base + i] : ~0ull;
        __syncthreads();
        smem_sort(s, TS_CAP);
        for (int i = tid; i < TS_CAP; i += TS_THREADS)
            if (base + i < n) keys[base + i] = s[i];
        __syncthreads();
    }
    for (int k = 2 * TS_CAP; k <= m; k <<= 1) {
        const int half = k >> 1;
        for (int i = tid; i < (m >> 1); i += TS_THREADS) {          // mirror sub-stage
            const int b = i / half, o = i - b * half;
            const int lo = b * k + o, hi = b * k + k - 1 - o;
            if (hi < n) {
                unsigned long long a = keys[lo], z = keys[hi];
                if (z < a) { keys[lo] = z; keys[hi] = a; }
            }
        }
        __syncthreads();
        for (int j = half >> 1; j >= TS_CAP; j >>= 1) {             // wide plain sub-stages
            for (int i = tid; i < (m >> 1); i += TS_THREADS) {
                const int lo = 2 * i - (i & (j - 1)), hi = lo + j;
                if (hi < n) {
                    unsigned long long a = keys[lo], z = keys[hi];
                    if (z < a) { keys[lo] = z; keys[hi] = a; }
                }
            }
            __syncthreads();
        }
        for (int c = 0; c < chunks; c++) {                          // narrow sub-stages, chunk by chunk
            const int base = c * TS_CAP;
            if (base >= n) break;
            for (int i = tid; i < TS_CAP; i += TS_THREADS) s[i] = (base + i < n) ? keys[base + i] : ~0ull;
            __syncthreads();
            smem_plain_stages(s, TS_CAP, TS_CAP >> 1);
            for (int i = tid; i < TS_CAP; i += TS_THREADS)
                if (base + i < n) keys[base + i] = s[i];
            __syncthreads();
        }
    }
    for (int i = tid; i < n; i += TS_THREADS) out[i] = (uint32_t)keys[i];
}

// The reference's 64-bit keys of the sorted list, for the bit-exact parity tests: tile << 32 | depth bits.
__global__ void __launch_bounds__(256) sorted_keys_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                                          const float* __restrict__ depths, uint64_t* __restrict__ out) {
    const uint32_t tile = blockIdx.x;
    const uint2 range = ranges[tile];
    for (uint32_t i = range.x + threadIdx.x; i < range.y; i += blockDim.x)
        out[i] = ((uint64_t)tile << 32) | (uint64_t)__float_as_uint(depths[point_list[i]]);
}

}  // namespace

void launch_tile_scan(const BinningArgs& a, cudaStream_t s) {
    const uint32_t tiles = (uint32_t)(a.grid_x * a.grid_y);
    tile_scan_kernel<<<gsb_div_up(tiles, SC_TILE), 256, 0, s>>>(a.tile_count, tiles, a.ranges, a.lookback, ++(*a.epoch), a.ctrl, a.r_cap,
                                                                a.num_rendered_host);
}

void launch_emit(const BinningArgs& a, cudaStream_t s) {
    const uint32_t P = (uint32_t)a.P;
    if (P == 0 || a.r_cap == 0) return;
    emit_kernel<<<gsb_div_up(P, 256), 256, 0, s>>>(a.tiles_touched, a.radii, a.depths, a.recs, P, a.grid_x, a.grid_y, a.r_cap,
                                                   a.tile_count, a.pairs);
}

void launch_tile_sort(const BinningArgs& a, cudaStream_t s) {
    const uint32_t tiles = (uint32_t)(a.grid_x * a.grid_y);
    tile_sort_kernel<<<tiles, TS_THREADS, 0, s>>>(a.ranges, a.pairs, a.point_list, a.tile_count, a.tile_max);
}

void launch_sorted_keys(const uint2* ranges, const uint32_t* point_list, const float* depths, uint32_t tiles, uint64_t* out,
                        cudaStream_t s) {
    sorted_keys_kernel<<<tiles, 256, 0, s>>>(ranges, point_list, depths, out);
}
