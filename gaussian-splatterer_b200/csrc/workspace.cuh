// Private layout of the three rasterizer workspaces (geometry / binning / image).
// Counterpart of GeometryState / BinningState / ImageState and obtain()/required()
// (reference cuda_rasterizer/rasterizer_impl.h:21-27,30-73, rasterizer_impl.cu:155-194): the caller
// allocates opaque blobs of the size this library asks for, the library carves 128-byte aligned
// sub-arrays out of them, and backward re-derives the same pointers from the same blob.
#pragma once
#include "common.cuh"

struct Carver {
    char* p;
    explicit Carver(char* base) : p(base) {}
    template <typename T>
    T* take(size_t count) {
        const uintptr_t a = (reinterpret_cast<uintptr_t>(p) + 127) & ~(uintptr_t)127;
        T* r = reinterpret_cast<T*>(a);
        p = reinterpret_cast<char*>(r + count);
        return r;
    }
};

// Per-splat state of one view.
struct GeomWs {
    GsbCtrl* ctrl;
    int* radii;
    float* depths;
    uint32_t* tiles_touched;
    SplatRec* recs;
    float* cov3D;
    uint8_t* clamped;
    GradRec* grads;

    char* carve(char* base, size_t P) {
        Carver c(base);
        ctrl = c.take<GsbCtrl>(1);
        radii = c.take<int>(P);
        depths = c.take<float>(P);
        tiles_touched = c.take<uint32_t>(P);
        recs = c.take<SplatRec>(P);
        cov3D = c.take<float>(6 * P);
        clamped = c.take<uint8_t>(P);
        grads = c.take<GradRec>(P);
        return c.p;
    }
    static size_t required(size_t P) {
        GeomWs w;
        return (size_t)(w.carve(nullptr, P) - (char*)nullptr) + 128;
    }
};

// Per-duplicate state of one view.
struct BinWs {
    unsigned long long* pairs;      // [R] depth_bits << 32 | splat, bucketed by tile
    uint32_t* point_list;           // [R] sorted splat ids
    unsigned long long* pairs2;     // [R] ping-pong partner of `pairs` in the merge rounds of lists longer than 1024
    uint2* chunk_list;              // [R / 512 + 2] queue of (tile, chunk): sum of ceil(n / 1024) over lists with n > 1024

    char* carve(char* base, size_t R) {
        Carver c(base);
        pairs = c.take<unsigned long long>(R);
        point_list = c.take<uint32_t>(R);
        pairs2 = c.take<unsigned long long>(R);
        chunk_list = c.take<uint2>(R / 512 + 2);
        return c.p;
    }
    static size_t required(size_t R) {
        BinWs w;
        return (size_t)(w.carve(nullptr, R) - (char*)nullptr) + 128;
    }
};

// Per-pixel / per-tile state of one view.
struct ImageWs {
    float* final_T;
    uint32_t* n_contrib;
    uint2* ranges;
    uint32_t* tile_max;
    uint32_t* tile_count;           // per-tile duplicate count / emit cursor; zero between view passes
    uint32_t* big_list;             // queue of long-list tiles (tile_sort_kernel -> tile_sort_big_kernel)
    uint32_t* tile_order;           // [GSB_LPT_BUCKETS][tiles] tiles by list-length class (heavy-first order, written by the tile scan)
    unsigned long long* lookback;   // chained-scan status words of tile_scan_kernel (2048 tiles per word)

    static size_t lookback_words(size_t tiles) { return (tiles + 2047) / 2048 + 1; }
    // bytes from `tile_max` to the end of the carved arrays: what must be zero before the first pass
    size_t zero_bytes(size_t tiles) const {
        return (size_t)(reinterpret_cast<const char*>(lookback + lookback_words(tiles)) - reinterpret_cast<const char*>(tile_max));
    }

    char* carve(char* base, size_t N, size_t tiles) {
        Carver c(base);
        final_T = c.take<float>(N);
        n_contrib = c.take<uint32_t>(N);
        ranges = c.take<uint2>(tiles);
        tile_max = c.take<uint32_t>(tiles);
        tile_count = c.take<uint32_t>(tiles);
        big_list = c.take<uint32_t>(tiles);
        lookback = c.take<unsigned long long>(lookback_words(tiles));
        tile_order = c.take<uint32_t>((size_t)GSB_LPT_BUCKETS * tiles);
        return c.p;
    }
    static size_t required(size_t N, size_t tiles) {
        ImageWs w;
        return (size_t)(w.carve(nullptr, N, tiles) - (char*)nullptr) + 128;
    }
};

int gsb_tile_bits(uint32_t n);
