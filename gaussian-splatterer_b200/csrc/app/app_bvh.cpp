// Host-side BVH build for the truth ray caster (stands in for owlGroupBuildAccel, src/rtx/RtxHost.cpp:96-99).
// Binned surface-area heuristic, 16 bins, leaves of <= 4 triangles, children stored adjacently.
#include "rtx.cuh"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <stdexcept>

namespace gsbrtx {

namespace {

struct Box {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void grow(const float* p) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void grow(const Box& b) { grow(b.lo); grow(b.hi); }
    float area() const {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx < 0 ? 0.0f : 2.0f * (dx * dy + dy * dz + dz * dx);
    }
};

struct Item { Box box; float centroid[3]; int prim; };

struct Builder {
    std::vector<Item> items;
    std::vector<Node>& nodes;
    float pad;
    Builder(std::vector<Node>& n) : nodes(n), pad(0) {}

    void set_node_box(int idx, const Box& b) {
        for (int k = 0; k < 3; k++) { nodes[(size_t)idx].bmin[k] = b.lo[k] - pad; nodes[(size_t)idx].bmax[k] = b.hi[k] + pad; }
    }

    void build(int idx, int begin, int end) {
        Box box, cbox;
        for (int i = begin; i < end; i++) { box.grow(items[(size_t)i].box); cbox.grow(items[(size_t)i].centroid); }
        set_node_box(idx, box);
        const int n = end - begin;
        auto leaf = [&] { nodes[(size_t)idx].first = begin; nodes[(size_t)idx].count = n; };
        if (n <= 4) { leaf(); return; }
        constexpr int BINS = 16;
        int best_axis = -1, best_split = -1;
        float best_cost = FLT_MAX;
        for (int axis = 0; axis < 3; axis++) {
            const float lo = cbox.lo[axis], ext = cbox.hi[axis] - lo;
            if (!(ext > 0.0f)) continue;
            Box bins[BINS];
            int counts[BINS] = {0};
            for (int i = begin; i < end; i++) {
                int b = (int)(BINS * ((items[(size_t)i].centroid[axis] - lo) / ext));
                b = std::min(std::max(b, 0), BINS - 1);
                bins[b].grow(items[(size_t)i].box);
                counts[b]++;
            }
            float right_area[BINS];
            int right_count[BINS];
            Box acc;
            int cnt = 0;
            for (int b = BINS - 1; b > 0; b--) { acc.grow(bins[b]); cnt += counts[b]; right_area[b] = cnt ? acc.area() : 0.0f; right_count[b] = cnt; }
            Box lacc;
            int lcnt = 0;
            for (int b = 0; b < BINS - 1; b++) {
                if (counts[b]) lacc.grow(bins[b]);
                lcnt += counts[b];
                if (lcnt == 0 || right_count[b + 1] == 0) continue;
                const float cost = lacc.area() * lcnt + right_area[b + 1] * right_count[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = b; }
            }
        }
        int mid;
        if (best_axis < 0) {
            if (n <= 16) { leaf(); return; }              // coincident centroids: give up splitting small sets
            mid = begin + n / 2;
        } else {
            const float lo = cbox.lo[best_axis], ext = cbox.hi[best_axis] - lo;
            auto it = std::partition(items.begin() + begin, items.begin() + end, [&](const Item& t) {
                int b = (int)(BINS * ((t.centroid[best_axis] - lo) / ext));
                b = std::min(std::max(b, 0), BINS - 1);
                return b <= best_split;
            });
            mid = (int)(it - items.begin());
            if (mid == begin || mid == end) mid = begin + n / 2;
        }
        const int left = (int)nodes.size();
        nodes.push_back(Node{});
        nodes.push_back(Node{});
        nodes[(size_t)idx].first = left;
        nodes[(size_t)idx].count = 0;
        build(left, begin, mid);
        build(left + 1, mid, end);
    }
};

}  // namespace

void build_bvh(const std::vector<float>& vertices, const std::vector<int>& triangles, std::vector<Node>& nodes, std::vector<Tri>& tris) {
    nodes.clear();
    tris.clear();
    const size_t nt = triangles.size() / 3, nv = vertices.size() / 3;
    Builder b(nodes);
    b.items.reserve(nt);
    Box scene;
    for (size_t t = 0; t < nt; t++) {
        Item it;
        it.prim = (int)t;
        for (int k = 0; k < 3; k++) {
            const int vi = triangles[t * 3 + k];
            if (vi < 0 || (size_t)vi >= nv) throw std::runtime_error("triangle references a vertex that does not exist");
            it.box.grow(&vertices[(size_t)vi * 3]);
        }
        bool finite = true;
        for (int k = 0; k < 3; k++) {
            it.centroid[k] = 0.5f * (it.box.lo[k] + it.box.hi[k]);
            finite = finite && std::isfinite(it.centroid[k]);
        }
        if (!finite) throw std::runtime_error("mesh has a non-finite vertex");
        scene.grow(it.box);
        b.items.push_back(it);
    }
    nodes.push_back(Node{});
    if (nt == 0) {
        for (int k = 0; k < 3; k++) { nodes[0].bmin[k] = 0.0f; nodes[0].bmax[k] = 0.0f; }
        nodes[0].first = 0; nodes[0].count = 0;
        return;
    }
    float ext = 0.0f;
    for (int k = 0; k < 3; k++) ext = std::max(ext, std::max(std::fabs(scene.lo[k]), std::fabs(scene.hi[k])));
    b.pad = 4e-5f * ext + 1e-7f;
    b.build(0, 0, (int)nt);
    tris.resize(nt);
    for (size_t i = 0; i < nt; i++) {
        const int prim = b.items[i].prim;
        const float* v0 = &vertices[(size_t)triangles[(size_t)prim * 3] * 3];
        const float* v1 = &vertices[(size_t)triangles[(size_t)prim * 3 + 1] * 3];
        const float* v2 = &vertices[(size_t)triangles[(size_t)prim * 3 + 2] * 3];
        Tri& tr = tris[i];
        tr.prim = prim; tr.pad0 = tr.pad1 = 0.0f;
        for (int k = 0; k < 3; k++) { tr.v0[k] = v0[k]; tr.e1[k] = v1[k] - v0[k]; tr.e2[k] = v2[k] - v0[k]; }
    }
}

}  // namespace gsbrtx
