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
    uint32_t* pkeys[2];
    uint32_t* pvals[2];
    uint32_t* offsets;
    GradRec* grads;
    unsigned long long* lookback;

    // scan tiles are 2048 items, sort tiles >= 1024: the sort bounds the tile count
    static size_t lookback_tiles(size_t P) { return (P + 1023) / 1024 + 1; }

    char* carve(char* base, size_t P) {
        Carver c(base);
        ctrl = c.take<GsbCtrl>(1);
        radii = c.take<int>(P);
        depths = c.take<float>(P);
        tiles_touched = c.take<uint32_t>(P);
        recs = c.take<SplatRec>(P);
        cov3D = c.take<float>(6 * P);
        clamped = c.take<uint8_t>(P);
        pkeys[0] = c.take<uint32_t>(P);
        pkeys[1] = c.take<uint32_t>(P);
        pvals[0] = c.take<uint32_t>(P);
        pvals[1] = c.take<uint32_t>(P);
        offsets = c.take<uint32_t>(P);
        grads = c.take<GradRec>(P);
        lookback = c.take<unsigned long long>(lookback_tiles(P) * 256);
        return c.p;
    }
    static size_t required(size_t P) {
        GeomWs w;
        return (size_t)(w.carve(nullptr, P) - (char*)nullptr) + 128;
    }
};

// Per-duplicate state of one view.
struct BinWs {
    uint32_t* rkeys[2];
    uint32_t* rvals[2];
    unsigned long long* lookback;

    static size_t lookback_tiles(size_t R) { return (R + 1023) / 1024 + 1; }

    char* carve(char* base, size_t R) {
        Carver c(base);
        rkeys[0] = c.take<uint32_t>(R);
        rkeys[1] = c.take<uint32_t>(R);
        rvals[0] = c.take<uint32_t>(R);
        rvals[1] = c.take<uint32_t>(R);
        lookback = c.take<unsigned long long>(lookback_tiles(R) * 256);
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

    char* carve(char* base, size_t N, size_t tiles) {
        Carver c(base);
        final_T = c.take<float>(N);
        n_contrib = c.take<uint32_t>(N);
        ranges = c.take<uint2>(tiles);
        tile_max = c.take<uint32_t>(tiles);
        return c.p;
    }
    static size_t required(size_t N, size_t tiles) {
        ImageWs w;
        return (size_t)(w.carve(nullptr, N, tiles) - (char*)nullptr) + 128;
    }
};

int gsb_tile_bits(uint32_t n);
