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
//   4. tile_sort_kernel: every tile's list is sorted on its own by that 64-bit key -- by one warp in registers (bitonic
//      network over 32 lanes x up to 32 keys, shuffles between lanes) up to 1024 entries, by the whole CTA in shared /
//      global memory beyond -- and the splat ids are written out.  The key is unique within a tile, so the unordered
//      claims of step 3 cannot show in the result.
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
constexpr uint32_t SORT_WARP_MAX = 1024;   // longest list (or chunk) one warp sorts in registers (32 keys per lane)
constexpr int SC_IPT = 8;
constexpr int SC_TILE = 256 * SC_IPT;

__global__ void __launch_bounds__(256) tile_scan_kernel(uint32_t* __restrict__ tile_count, uint32_t tiles, uint2* __restrict__ ranges,
                                                        unsigned long long* lookback, uint32_t epoch, GsbCtrl* ctrl,
                                                        uint32_t r_cap, uint32_t* num_rendered_host,
                                                        uint32_t* __restrict__ big_list, uint2* __restrict__ chunk_list,
                                                        uint32_t* __restrict__ tile_order) {
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_tile, s_excl;
    __shared__ uint32_t s_bn[GSB_LPT_BUCKETS], s_bbase[GSB_LPT_BUCKETS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&ctrl->ticket, 1u);
    if (tid < GSB_LPT_BUCKETS) s_bn[tid] = 0u;
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
    // heavy-first tile order (common.cuh gsb_lpt_tile): rank inside the CTA by shared-memory atomics, one global atomic per
    // (CTA, class) for the base
    uint32_t lrank[SC_IPT];
    if (tile_order) {
#pragma unroll
        for (int k = 0; k < SC_IPT; k++)
            if (base + k < tiles) lrank[k] = atomicAdd(&s_bn[gsb_lpt_bucket(v[k])], 1u);
    }
    if (tid == 0) s_excl = lookback_exclusive(lookback, tile, epoch, total);
    __syncthreads();
    if (tile_order && tid < GSB_LPT_BUCKETS && s_bn[tid]) s_bbase[tid] = atomicAdd(&ctrl->bucket_n[tid], s_bn[tid]);
    __syncthreads();
    if (tile_order) {
#pragma unroll
        for (int k = 0; k < SC_IPT; k++)
            if (base + k < tiles) {
                const int b = gsb_lpt_bucket(v[k]);
                tile_order[(size_t)b * tiles + s_bbase[b] + lrank[k]] = base + k;
            }
    }
    uint32_t run = s_excl + wp + inc - sum;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        const uint32_t i = base + k;
        if (i < tiles) {
            const uint32_t lo = min(run, r_cap), hi = min(run + v[k], r_cap);
            ranges[i] = v[k] ? make_uint2(lo, hi) : make_uint2(0u, 0u);
            tile_count[i] = run;
            if (chunk_list && hi - lo > SORT_WARP_MAX) {
                // a list too long for one warp's registers: queue the tile for the merge kernel and its 1024-key
                // chunks for the chunk-sort kernel (sum of chunks <= R / 1024 + number of such tiles <= R / 512)
                const uint32_t chunks = (hi - lo + SORT_WARP_MAX - 1) / SORT_WARP_MAX;
                big_list[atomicAdd(&ctrl->num_big, 1u)] = i;
                const uint32_t at = atomicAdd(&ctrl->num_chunks, chunks);
                for (uint32_t c = 0; c < chunks; c++) chunk_list[at + c] = make_uint2(i, c);
            }
        }
        run += v[k];
    }
    if ((tile + 1) * SC_TILE >= tiles && tid == 255) {   // last scan tile, last thread: run == R
        ctrl->num_rendered = run;
        if (run > r_cap) ctrl->overflow = 1u;
        if (num_rendered_host) *num_rendered_host = run;
    }
}

// The same queueing as a kernel of its own, for callers whose duplicate arena (and with it the chunk queue) is only
// allocated once R is known (gsb_rasterizer_forward: the scan runs before the binning allocation).
__global__ void __launch_bounds__(256) queue_long_lists_kernel(const uint2* __restrict__ ranges, uint32_t tiles, uint32_t* __restrict__ big_list,
                                                               uint2* __restrict__ chunk_list, GsbCtrl* ctrl) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tiles) return;
    const uint2 r = ranges[i];
    if (r.y - r.x > SORT_WARP_MAX) {
        const uint32_t chunks = (r.y - r.x + SORT_WARP_MAX - 1) / SORT_WARP_MAX;
        big_list[atomicAdd(&ctrl->num_big, 1u)] = i;
        const uint32_t at = atomicAdd(&ctrl->num_chunks, chunks);
        for (uint32_t c = 0; c < chunks; c++) chunk_list[at + c] = make_uint2(i, c);
    }
}

// Duplicate emission (duplicateWithKeys, rasterizer_impl.cu:70-111): the rectangle is recomputed from the stored
// pixel centre and integer radius exactly as the reference does; each duplicate claims the next free slot of its tile.
// The (splat, tile) pairs of a warp's 32 splats are walked cooperatively (gsb_warp_walk_tiles): all lanes busy, two
// independent ATOMs in flight per lane and round.
__global__ void __launch_bounds__(256) emit_kernel(const uint32_t* __restrict__ tiles_touched, const int* __restrict__ radii,
                                                   const float* __restrict__ depths, const SplatRec* __restrict__ recs,
                                                   uint32_t P, int grid_x, int grid_y, uint32_t r_cap,
                                                   uint32_t* __restrict__ cursor, unsigned long long* __restrict__ pairs) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t cnt = 0, rx0 = 0, ry0 = 0, w = 0, dbits = 0;
    if (s < P && tiles_touched[s] != 0u) {
        const float4 ra = recs[s].a;
        const float rf = (float)radii[s];
        dbits = __float_as_uint(depths[s]);
        const int x0 = (int)__fmul_rn(__fsub_rn(ra.x, rf), 0.0625f), y0 = (int)__fmul_rn(__fsub_rn(ra.y, rf), 0.0625f);
        const int x1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.x, rf), 16.0f), -1.0f), 0.0625f);
        const int y1 = (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(ra.y, rf), 16.0f), -1.0f), 0.0625f);
        const uint32_t gx = (uint32_t)grid_x, gy = (uint32_t)grid_y;
        rx0 = min(gx, (uint32_t)max(0, x0)); ry0 = min(gy, (uint32_t)max(0, y0));
        const uint32_t rx1 = min(gx, (uint32_t)max(0, x1)), ry1 = min(gy, (uint32_t)max(0, y1));
        w = rx1 - rx0;
        cnt = w * (ry1 - ry0);
    }
    if (!__any_sync(0xffffffffu, cnt != 0u)) return;
    const uint32_t warp_first = s - (threadIdx.x & 31);
    const uint32_t gxw = (uint32_t)grid_x;
    gsb_warp_walk_tiles(rx0, ry0, w, cnt, [&](const bool* on, const int* src, const uint32_t* tx, const uint32_t* ty) {
        uint32_t pos[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
        unsigned long long key[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const uint32_t d = __shfl_sync(0xffffffffu, dbits, src[u]);
            key[u] = ((unsigned long long)d << 32) | (unsigned long long)(warp_first + (uint32_t)src[u]);
            if (on[u]) pos[u] = atomicAdd(&cursor[ty[u] * gxw + tx[u]], 1u);
        }
#pragma unroll
        for (int u = 0; u < 2; u++)
            if (pos[u] < r_cap) pairs[pos[u]] = key[u];
    });
}

// ---- per-tile sort ---------------------------------------------------------------------------------------------
// Lists of up to 1024 entries (all but a few tiles at C4) are sorted by ONE WARP entirely in registers: lane l holds
// elements l*E .. l*E+E-1 of a 32*E-element bitonic network, sub-stages whose partner sits in the same lane are
// compare-exchanges between registers, the others exchange with lane l ^ (j / E) by shuffle.  No shared memory, no
// barrier.  The list is loaded striped (coalesced; the network sorts any input order) and padded with ~0.
// Longer lists (every tile at C5) are cut into 1024-key chunks, each sorted the same way by its own warp (in place, as
// keys), and then merged: one CTA per list, log2(chunks) rounds of pairwise merge-path merges between `pairs` and a
// second buffer, the last round writing the splat ids.
constexpr int TS_WARP_MAX = (int)SORT_WARP_MAX;
#ifndef GSB_TS_CTAS
#define GSB_TS_CTAS 20   // one-warp CTAs per SM the tile sort is cut for: 95 registers and no spill at 20 (0.0575 ms) against 80 + spills at 24 (0.0615), 32: 0.066
#endif
constexpr int MG_THREADS = 256;       // merge CTA
constexpr int MG_ITEMS = 8;           // outputs per thread per block of MG_THREADS * MG_ITEMS = 2048 merged keys

// The network is written so that its code stays small: the levels that live inside one lane (k <= E) are unrolled (their
// compare directions depend on the register index), every later level k = 2E .. 32E runs through ONE rolled body -- a
// runtime loop over the cross-lane distances (shfl_xor takes the lane mask from a register) followed by the unrolled
// in-lane merge-down, whose direction is the lane's for all of them.  A fully unrolled network is 6.7 k instructions for
// E = 32 alone and 26.7 k for the six list classes together (427 KB of SASS): with every warp of an SM at a different
// place in a different class, ncu showed 75 % of the stall samples on "no instruction" (instruction-cache misses) and
// 0.04 instructions per clock and warp.
// One copy of each list class in the kernel (__noinline__, the store form is a runtime flag): the tile path and the
// chunk path share the E = 32 body.
template <int E>
__device__ __noinline__ void warp_sort_tile(const unsigned long long* __restrict__ keys, int n, void* __restrict__ out, bool ids) {
    const int lane = threadIdx.x & 31;
    unsigned long long v[E];
#pragma unroll
    for (int r = 0; r < E; r++) {
        const int i = r * 32 + lane;
        v[r] = (i < n) ? keys[i] : ~0ull;
    }
    // element index of register r of lane l: i = l * E + r.  Blocks of k elements alternate ascending / descending; the
    // last level (k = 32 E) is ascending everywhere.
#pragma unroll
    for (int k = 2; k <= E; k <<= 1) {
        const bool asc_lane = ((lane * E) & k) == 0;      // only k == E depends on the lane
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int r = 0; r < E; r++) {
                if ((r & j) == 0) {
                    const bool asc = (k >= E) ? asc_lane : ((r & k) == 0);
                    const unsigned long long a = v[r], b = v[r + j];
                    if ((b < a) == asc) { v[r] = b; v[r + j] = a; }
                }
            }
        }
    }
#pragma unroll 1
    for (int k = 2 * E; k <= 32 * E; k <<= 1) {
        const bool asc_lane = ((lane * E) & k) == 0;
#pragma unroll 1
        for (int lj = k / (2 * E); lj > 0; lj >>= 1) {
            const bool keep_min = (((lane & lj) == 0) == asc_lane);
#pragma unroll
            for (int r = 0; r < E; r++) {
                const unsigned long long o = __shfl_xor_sync(0xffffffffu, v[r], lj);
                if ((o < v[r]) == keep_min) v[r] = o;     // keep_min: min(v, o); else max(v, o) (keys are distinct or both ~0)
            }
        }
#pragma unroll
        for (int j = E >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int r = 0; r < E; r++) {
                if ((r & j) == 0) {
                    const unsigned long long a = v[r], b = v[r + j];
                    if ((b < a) == asc_lane) { v[r] = b; v[r + j] = a; }
                }
            }
        }
    }
    // (in place is safe: every key of the list is in a register by now)
#pragma unroll
    for (int r = 0; r < E; r++) {
        const int i = lane * E + r;
        if (i < n) {
            if (ids) static_cast<uint32_t*>(out)[i] = (uint32_t)v[r];
            else static_cast<unsigned long long*>(out)[i] = v[r];
        }
    }
}

__device__ __forceinline__ void warp_sort_dispatch(const unsigned long long* keys, int n, void* out, bool ids) {
    if (n <= 32) warp_sort_tile<1>(keys, n, out, ids);
    else if (n <= 64) warp_sort_tile<2>(keys, n, out, ids);
    else if (n <= 128) warp_sort_tile<4>(keys, n, out, ids);
    else if (n <= 256) warp_sort_tile<8>(keys, n, out, ids);
    else if (n <= 512) warp_sort_tile<16>(keys, n, out, ids);
    else warp_sort_tile<32>(keys, n, out, ids);
}

// One 32-thread CTA (one warp) per tile: the list length is uniform over the CTA, so the compiler knows every shuffle
// of the register network is convergent (inside a per-warp branch of a larger CTA each one is wrapped in
// WARPSYNC / ENDCOLLECTIVE).  Lists too long for one warp were queued by the tile scan as 1024-key chunks; after its
// own tile every CTA takes its share of that queue (grid-stride from blockIdx: uniform control flow again) and sorts
// the chunks in place, as keys, for the merge kernel.
__global__ void __launch_bounds__(32, GSB_TS_CTAS) tile_sort_kernel(const uint2* __restrict__ ranges, unsigned long long* __restrict__ pairs,
                                                           uint32_t* __restrict__ point_list, uint32_t* __restrict__ tile_count,
                                                           uint32_t* __restrict__ tile_max, uint32_t tiles,
                                                           const uint2* __restrict__ chunk_list, const GsbCtrl* ctrl,
                                                           const uint32_t* __restrict__ tile_order) {
    if (blockIdx.x < tiles) {
        const uint32_t tile = gsb_lpt_tile(ctrl, tile_order, tiles, blockIdx.x);     // longest lists first
        // housekeeping for the kernels around this one: the count / cursor of this tile starts the next view pass at
        // zero, and the render forward takes the maximum of n_contrib over the tile with atomicMax
        if (threadIdx.x == 0) { tile_count[tile] = 0u; tile_max[tile] = 0u; }
        const uint2 range = ranges[tile];
        const int n = (int)(range.y - range.x);
        if (n > 0 && n <= TS_WARP_MAX) warp_sort_dispatch(pairs + range.x, n, point_list + range.x, true);
    }
    const uint32_t count = ctrl->num_chunks;
    for (uint32_t q = blockIdx.x; q < count; q += gridDim.x) {
        const uint2 e = chunk_list[q];
        const uint2 range = ranges[e.x];
        const uint32_t off = e.y * TS_WARP_MAX;
        const int m = (int)min((uint32_t)TS_WARP_MAX, range.y - range.x - off);
        unsigned long long* keys = pairs + range.x + off;
        warp_sort_tile<32>(keys, m, keys, false);     // (one class only: all chunks but a list's last are full, and code size matters here)
    }
}

// Merge rounds of the queued long lists, one list per CTA at a time.  Round w merges runs [2pw, 2pw + w) and
// [2pw + w, 2pw + 2w) of `src` into `dst`; every thread finds the merge-path split of its diagonal by binary search
// (keys are distinct), stages the at most MG_ITEMS keys it can need from each run in shared memory with independent
// loads, and merges MG_ITEMS outputs from there.  The last round stores the splat ids into the point list.
__global__ void __launch_bounds__(MG_THREADS) tile_merge_kernel(const uint2* __restrict__ ranges, unsigned long long* __restrict__ pairs,
                                                               unsigned long long* __restrict__ pairs2, uint32_t* __restrict__ point_list,
                                                               const uint32_t* __restrict__ big_list, const GsbCtrl* ctrl) {
    __shared__ unsigned long long s_a[MG_ITEMS][MG_THREADS], s_b[MG_ITEMS][MG_THREADS];
    const int tid = threadIdx.x;
    const uint32_t count = ctrl->num_big;
    for (uint32_t q = blockIdx.x; q < count; q += gridDim.x) {
        const uint2 rg = ranges[big_list[q]];
        const int n = (int)(rg.y - rg.x);
        const unsigned long long* src = pairs + rg.x;
        unsigned long long* dst = pairs2 + rg.x;
        uint32_t* ids = point_list + rg.x;
        for (int w = TS_WARP_MAX; w < n; w <<= 1) {
            const bool last = 2 * w >= n;
            for (int o0 = tid * MG_ITEMS; o0 < n; o0 += MG_THREADS * MG_ITEMS) {
                const int pbase = (o0 / (2 * w)) * (2 * w);          // 2w is a multiple of MG_ITEMS: no straddling
                const unsigned long long* A = src + pbase;
                const int la = min(w, n - pbase), lb = min(w, n - pbase - la);
                const unsigned long long* B = A + la;
                const int d = o0 - pbase;
                int lo = max(0, d - lb), hi = min(d, la);
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (A[mid] < B[d - 1 - mid]) lo = mid + 1; else hi = mid;
                }
                int i = lo, j = d - lo;
                const int cnt = min(MG_ITEMS, la + lb - d);
#pragma unroll
                for (int k = 0; k < MG_ITEMS; k++) {
                    s_a[k][tid] = (i + k < la) ? A[i + k] : ~0ull;
                    s_b[k][tid] = (j + k < lb) ? B[j + k] : ~0ull;
                }
                int ia = 0, ib = 0;
                unsigned long long a = s_a[0][tid], b = s_b[0][tid];
#pragma unroll
                for (int k = 0; k < MG_ITEMS; k++) {
                    const bool take_a = a < b;
                    const unsigned long long v = take_a ? a : b;
                    if (take_a) { ia++; a = ia < MG_ITEMS ? s_a[ia][tid] : ~0ull; }
                    else { ib++; b = ib < MG_ITEMS ? s_b[ib][tid] : ~0ull; }
                    if (k < cnt) {
                        if (last) ids[o0 + k] = (uint32_t)v;
                        else dst[o0 + k] = v;
                    }
                }
            }
            __syncthreads();              // the round's stores are visible to the whole CTA; shared staging is per thread
            const unsigned long long* t = src; src = dst; dst = const_cast<unsigned long long*>(t);
        }
    }
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
                                                                a.num_rendered_host, a.big_list, a.chunk_list, a.tile_order);
}

void launch_queue_long_lists(const BinningArgs& a, cudaStream_t s) {
    const uint32_t tiles = (uint32_t)(a.grid_x * a.grid_y);
    queue_long_lists_kernel<<<gsb_div_up(tiles, 256), 256, 0, s>>>(a.ranges, tiles, a.big_list, a.chunk_list, a.ctrl);
}

void launch_emit(const BinningArgs& a, cudaStream_t s) {
    const uint32_t P = (uint32_t)a.P;
    if (P == 0 || a.r_cap == 0) return;
    emit_kernel<<<gsb_div_up(P, 256), 256, 0, s>>>(a.tiles_touched, a.radii, a.depths, a.recs, P, a.grid_x, a.grid_y, a.r_cap,
                                                   a.tile_count, a.pairs);
}

void launch_tile_sort(const BinningArgs& a, cudaStream_t s) {
    const uint32_t tiles = (uint32_t)(a.grid_x * a.grid_y);
    tile_sort_kernel<<<max(tiles, 148u * (uint32_t)GSB_TS_CTAS), 32, 0, s>>>(a.ranges, a.pairs, a.point_list, a.tile_count, a.tile_max, tiles, a.chunk_list, a.ctrl, a.tile_order);
    // the merge kernel runs over the device-side queue of long lists with a grid-stride loop (nearly empty at C4)
    tile_merge_kernel<<<min(tiles, 148u * 4u), MG_THREADS, 0, s>>>(a.ranges, a.pairs, a.pairs2, a.point_list, a.big_list, a.ctrl);
}

void launch_sorted_keys(const uint2* ranges, const uint32_t* point_list, const float* depths, uint32_t tiles, uint64_t* out,
                        cudaStream_t s) {
    sorted_keys_kernel<<<tiles, 256, 0, s>>>(ranges, point_list, depths, out);
}
