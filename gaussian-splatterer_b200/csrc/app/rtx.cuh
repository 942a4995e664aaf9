// Shared between the BVH builder (host, app_bvh.cpp) and the path-tracing kernel (rtx_trace.cu).
#pragma once

#include <cstdint>
#include <vector>

namespace gsbrtx {

// 32-byte BVH node, read as two 16-byte loads.  count > 0: leaf over tris[first .. first+count); count == 0: inner node
// whose children are nodes[first] and nodes[first + 1].
struct Node {
    float bmin[3];
    int first;
    float bmax[3];
    int count;
};

// 48-byte triangle record in leaf order: v0, e1 = v1 - v0, e2 = v2 - v0; `prim` is the index in the caller's triangle
// list (optixGetPrimitiveIndex), used for texture coordinates and for the deterministic tie-break between equal hits.
struct Tri {
    float v0[3];
    int prim;
    float e1[3];
    float pad0;
    float e2[3];
    float pad1;
};

// Binned-SAH build on the host (meshes are loaded once per project).  Boxes are padded by a few 1e-5 of the scene extent
// so that the slab test can never cull a hit the triangle test would accept.
void build_bvh(const std::vector<float>& vertices, const std::vector<int>& triangles, std::vector<Node>& nodes,
               std::vector<Tri>& tris);

}  // namespace gsbrtx
