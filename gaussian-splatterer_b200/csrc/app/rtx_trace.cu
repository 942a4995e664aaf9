// Truth ray caster: the reference's RtxDevice.cu programs (ray generation RD:53-99, closest hit RD:102-143, miss RD:146-157,
// tracePath RD:17-50) and RtxHost (RH) as ONE CUDA kernel over a BVH, without OptiX/OWL.  SURVEY.md 8 f1.
//
// One thread per pixel; each thread runs the reference's per-pixel program verbatim in structure: `samples` jittered primary
// rays, up to 50 path segments each, one LCG<4> stream per pixel shared by jitter, alpha test, scattering decision and the
// unit-sphere rejection loop, in the order the reference draws them.  Compiled with -fmad=false and IEEE division / square
// root, with every expression written in the reference's (owl vec3f) evaluation order, so the CPU restatement
// (oracle/oracle_rtx.cpp: brute-force intersection, no BVH) reproduces the frame bit for bit.
//
// Where the reference's behaviour is undefined or lives in OptiX hardware, the choice made here is stated once:
//   * LCG seed: init(pixel.x, pixel.y) (the reference never seeds PerRayData::random);
//   * vec3f(random(), random(), random()) is evaluated x, y, z;
//   * triangle test: Moeller-Trumbore, two-sided, barycentrics (u, v) weighting vertices 1 and 2 like
//     optixGetTriangleBarycentrics; of several hits at exactly the same t the lowest primitive index wins;
//   * tex2D point sampling with wrap addressing on normalised coordinates: texel = floor(frac(u) * size), value / 255.
#include <cuda_runtime.h>

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "app_common.h"
#include "rtx.cuh"

namespace {

using gsbrtx::Node;
using gsbrtx::Tri;

struct RtxParams {
    uint32_t* fb0;               // first frame (background bg0)
    uint32_t* fb1;               // second frame (background bg1) or null
    int W, H, samples;
    float bg0[3], bg1[3];
    const float4* nodes;
    const float4* tris;
    const float2* texcoords;     // 3 per primitive
    const uchar4* tex;
    int texW, texH;
    float cam[3];
    float M[16];                 // inverse(proj * view), column-major
    int n_orbs;
    const float* orbs;           // splat camera locations, 3 floats each
    unsigned long long* ray_count;
    int n_tris;
};

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
__device__ __forceinline__ V3 normalize(V3 v) { return v * (1.f / sqrtf(dot(v, v))); }      // owl vec.h:355, owl-common.h:262

// owl::common::LCG<4> (owl/common/math/random.h:46-84)
struct LCG {
    uint32_t state;
    __device__ __forceinline__ void init(uint32_t val0, uint32_t val1) {
        uint32_t v0 = val0, v1 = val1, s0 = 0;
        for (int n = 0; n < 4; n++) {
            s0 += 0x9e3779b9;
            v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4);
            v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761e);
        }
        state = v0;
    }
    __device__ __forceinline__ float operator()() {
        state = 1664525u * state + 1013904223u;
        return (state & 0x00FFFFFF) / (float)0x01000000;
    }
};

struct Hit { float t, u, v; int prim, slot; };

// Entry distance of the ray into a node's box clipped to (tmin, tbest]; a negative result means "missed".
// fminf / fmaxf drop the NaN of 0 * inf, and the far side is widened by 2 ulp so rounding never culls a real hit.
__device__ __forceinline__ float box_entry(const float4 lo, const float4 hi, V3 o, V3 inv, float tmin, float tbest) {
    const float tx0 = (lo.x - o.x) * inv.x, tx1 = (hi.x - o.x) * inv.x;
    const float ty0 = (lo.y - o.y) * inv.y, ty1 = (hi.y - o.y) * inv.y;
    const float tz0 = (lo.z - o.z) * inv.z, tz1 = (hi.z - o.z) * inv.z;
    const float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), tmin));
    const float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), tbest));
    return tn <= tf * 1.0000003f ? tn : -1.f;
}

// Closest intersection of origin + t * dir with tmin < t < tmax over the BVH (the role of optixTrace + the hardware
// triangle test).  Both children of an inner node are tested together; the nearer one is visited first.
__device__ __forceinline__ bool closest_hit(const RtxParams& p, V3 o, V3 d, float tmin, float tmax, Hit& hit) {
    hit.t = tmax; hit.prim = 0x7fffffff; hit.slot = 0; hit.u = hit.v = 0.f;
    bool found = false;
    const V3 inv = mk(1.f / d.x, 1.f / d.y, 1.f / d.z);
    int stack[48];
    float stack_t[48];
    int sp = 0;
    int node = 0;
    while (true) {
        const int first = __float_as_int(__ldg(&p.nodes[2 * node]).w), count = __float_as_int(__ldg(&p.nodes[2 * node + 1]).w);
        int next = -1;
        if (count == 0) {
            if (p.n_tris > 0) {
                const float e0 = box_entry(__ldg(&p.nodes[2 * first]), __ldg(&p.nodes[2 * first + 1]), o, inv, tmin, hit.t);
                const float e1 = box_entry(__ldg(&p.nodes[2 * first + 2]), __ldg(&p.nodes[2 * first + 3]), o, inv, tmin, hit.t);
                if (e0 >= 0.f && e1 >= 0.f) {
                    const bool swap = e1 < e0;
                    next = swap ? first + 1 : first;
                    if (sp < 48) { stack[sp] = swap ? first : first + 1; stack_t[sp] = swap ? e0 : e1; sp++; }
                } else if (e0 >= 0.f) next = first;
                else if (e1 >= 0.f) next = first + 1;
            }
        } else {
            for (int k = 0; k < count; k++) {
                const int slot = first + k;
                const float4 t0 = __ldg(&p.tris[3 * slot]), t1 = __ldg(&p.tris[3 * slot + 1]), t2 = __ldg(&p.tris[3 * slot + 2]);
                const V3 v0 = mk(t0.x, t0.y, t0.z), e1v = mk(t1.x, t1.y, t1.z), e2v = mk(t2.x, t2.y, t2.z);
                const int prim = __float_as_int(t0.w);
                const V3 pvec = cross(d, e2v);
                const float det = dot(e1v, pvec);
                if (det == 0.f) continue;
                const float idet = 1.f / det;
                const V3 tvec = o - v0;
                const float u = dot(tvec, pvec) * idet;
                if (!(u >= 0.f && u <= 1.f)) continue;
                const V3 qvec = cross(tvec, e1v);
                const float v = dot(d, qvec) * idet;
                if (!(v >= 0.f && u + v <= 1.f)) continue;
                const float t = dot(e2v, qvec) * idet;
                if (!(t > tmin && t < tmax)) continue;
                if (t < hit.t || (t == hit.t && prim < hit.prim)) { hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; hit.slot = slot; found = true; }
            }
        }
        while (next < 0) {
            if (sp == 0) return found;
            --sp;
            if (stack_t[sp] <= hit.t) next = stack[sp];      // boxes entered beyond the best hit cannot hold a closer (or equal) one
        }
        node = next;
    }
}

__device__ __forceinline__ uint32_t make_8bit(float f) { return (uint32_t)min(255, max(0, (int)(f * 256.f))); }      // owl_device.h:79-82
__device__ __forceinline__ uint32_t make_rgba(V3 c) { return make_8bit(c.x) | (make_8bit(c.y) << 8) | (make_8bit(c.z) << 16) | (0xffU << 24); }
__device__ __forceinline__ V3 clamp01(V3 c) { return mk(fmaxf(fminf(c.x, 1.0f), 0.0f), fmaxf(fminf(c.y, 1.0f), 0.0f), fmaxf(fminf(c.z, 1.0f), 0.0f)); }

__global__ void __launch_bounds__(128) rtx_trace_kernel(const RtxParams p) {
    const int px = blockIdx.x * 16 + (threadIdx.x & 15), py = blockIdx.y * 8 + (threadIdx.x >> 4);
    unsigned long long rays = 0;
    if (px < p.W && py < p.H) {
        LCG random;
        random.init((uint32_t)px, (uint32_t)py);
        V3 color0 = mk(0.f, 0.f, 0.f), color1 = mk(0.f, 0.f, 0.f);
        bool touches_orb = false;
        const V3 cam = mk(p.cam[0], p.cam[1], p.cam[2]);
        // The reference's two nested loops (samples, RD:60; path segments, RD:23) run here as ONE loop over path segments: a
        // lane that finishes a sample early starts its next one at once instead of idling until the longest path of its
        // warp ends.  Per pixel the sequence of operations and random draws is unchanged.
        V3 origin = cam, dir = cam, atten = mk(1.f, 1.f, 1.f);
        float tmin = 0.f, tmax = 1e30f;
        bool has_reflected = false;
        int s = 0, i = 0;
        bool fresh = true;
        while (true) {
            if (fresh) {
                if (s == p.samples) break;
                // ---- ray generation, RD:63-77
                const float r0 = random(), r1 = random();
                const float fx = ((float)px + r0) + 0.5f, fy = ((float)py + r1) + 0.5f;
                const float vx = (float)(((double)fx * 2.0 / (double)p.W) - 1.0);
                const float vy = (float)(((double)fy * 2.0 / (double)p.H) - 1.0);
                const float vz = 1.0f;
                const float* M = p.M;
                const float hx = vx * M[0] + vy * M[4] + vz * M[8] + M[12];
                const float hy = vx * M[1] + vy * M[5] + vz * M[9] + M[13];
                const float hz = vx * M[2] + vy * M[6] + vz * M[10] + M[14];
                const float hw = vx * M[3] + vy * M[7] + vz * M[11] + M[15];
                origin = cam;
                dir = normalize(mk(hx / hw, hy / hw, hz / hw) - cam);
                tmin = 0.f; tmax = 1e30f;                                // owl::Ray defaults, owl_device.h:147
                atten = mk(1.f, 1.f, 1.f);                               // tracePath, RD:17-21
                has_reflected = false;
                i = 0;
                fresh = false;
            }
            Hit h;
            const bool hit = closest_hit(p, origin, dir, tmin, tmax, h);
            rays++;
            V3 prd_color, hit_origin = mk(0.f, 0.f, 0.f), bounce = mk(0.f, 0.f, 0.f);
            bool reflected = false;
            if (hit) {                                                  // closest hit program, RD:102-143
                const float2 tc0 = __ldg(&p.texcoords[3 * h.prim]), tc1 = __ldg(&p.texcoords[3 * h.prim + 1]), tc2 = __ldg(&p.texcoords[3 * h.prim + 2]);
                const float4 q1 = __ldg(&p.tris[3 * h.slot + 1]), q2 = __ldg(&p.tris[3 * h.slot + 2]);
                const V3 normal = normalize(cross(mk(q1.x, q1.y, q1.z), mk(q2.x, q2.y, q2.z)));     // RD:112-113
                const float w0 = 1.0f - h.u - h.v;
                const float tcx = w0 * tc0.x + h.u * tc1.x + h.v * tc2.x;
                const float tcy = w0 * tc0.y + h.u * tc1.y + h.v * tc2.y;
                const float su = tcx, sv = 1.0f - tcy;
                const float fu = su - floorf(su), fv = sv - floorf(sv);
                const int ix = min((int)floorf(fu * (float)p.texW), p.texW - 1), iy = min((int)floorf(fv * (float)p.texH), p.texH - 1);
                const uchar4 tx = __ldg(&p.tex[iy * p.texW + ix]);
                const float cr = (float)tx.x / 255.f, cg = (float)tx.y / 255.f, cb = (float)tx.z / 255.f, ca = (float)tx.w / 255.f;
                hit_origin = origin + h.t * dir;
                if (ca > random()) {
                    V3 refl = dir - 2.0f * dot(dir, normal) * normal;
                    if (random() > 0.0f) {                              // MATERIAL_REFLECTIVITY 0, RD(h):14
                        V3 v;
                        do {
                            const float a = random(), b = random(), c = random();
                            v = 2.0f * mk(a, b, c) - mk(1.0f, 1.0f, 1.0f);
                        } while (dot(v, v) >= 1.0f);
                        refl = normal + v;
                    }
                    bounce = refl;
                    prd_color = mk(cr, cg, cb);
                    reflected = true;
                } else {
                    bounce = dir;
                    prd_color = mk(1.f, 1.f, 1.f);
                }
            } else {                                                    // miss program, RD:146-157
                const V3 n = normalize(dir);
                const float sky = fminf(1.0f, 1.0f + n.y);
                prd_color = mk(sky, sky, sky);
            }
            atten = atten * prd_color;
            if (reflected) has_reflected = true;
            if (i == 0) {                                               // camera indicator orbs, RD:33-47
                for (int c = 0; c < p.n_orbs; c++) {
                    const V3 orb = mk(p.orbs[3 * c], p.orbs[3 * c + 1], p.orbs[3 * c + 2]);
                    if (hit && dot(dir, orb - origin) > dot(dir, hit_origin - origin)) continue;
                    const V3 closest = origin + dir * dot(dir, orb - origin);
                    const V3 delta = orb - closest;
                    if (dot(delta, delta) < 0.025f * 0.025f) { touches_orb = true; break; }
                }
            }
            if (!hit) {
                const V3 out0 = has_reflected ? atten : mk(p.bg0[0], p.bg0[1], p.bg0[2]);
                const V3 out1 = has_reflected ? atten : mk(p.bg1[0], p.bg1[1], p.bg1[2]);
                color0 = color0 + clamp01(out0 * 1.0f);                 // PROGRAM_EXPOSURE_FACTOR 1
                color1 = color1 + clamp01(out1 * 1.0f);
                s++; fresh = true;
            } else if (++i == 50) {
                s++; fresh = true;                                      // 50 segments exceeded: the sample adds (0,0,0), RD:49
            } else {
                origin = hit_origin; dir = bounce; tmin = 1e-3f; tmax = 1e10f;
            }
        }
        const float ns = (float)p.samples;
        V3 c0 = mk(color0.x / ns, color0.y / ns, color0.z / ns), c1 = mk(color1.x / ns, color1.y / ns, color1.z / ns);
        if (touches_orb) { c0 = mk(1.f, 1.f, 1.f) - c0; c1 = mk(1.f, 1.f, 1.f) - c1; }
        p.fb0[px + p.W * py] = make_rgba(c0);
        if (p.fb1) p.fb1[px + p.W * py] = make_rgba(c1);
    }
    // one atomic per warp for the ray counter
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_down_sync(0xffffffffu, rays, off);
    if ((threadIdx.x & 31) == 0 && rays && p.ray_count) atomicAdd(p.ray_count, rays);
}

}  // namespace

// ================================================================================================== RtxHost (host side)
struct gsb_rtx {
    gsbapp::ObjMesh mesh;                         // bufferedVertices / bufferedTriangles / bufferedTextureCoords, RH(h):20-22
    std::vector<uint8_t> tex;                     // bufferedTextures[TEXTURE_DIFFUSE], RGBA8
    int texW = 0, texH = 0;
    bool model_present = false;
    // device copies
    float4* d_nodes = nullptr;
    float4* d_tris = nullptr;
    float2* d_texcoords = nullptr;
    uchar4* d_tex = nullptr;
    float* d_orbs = nullptr;
    int orbs_cap = 0;
    unsigned long long* d_rays = nullptr;
    int n_tris = 0;
    unsigned long long last_rays = 0;
};

namespace {

#define RTX_CUDA_OK(expr)                                                                       \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) return gsbapp::fail(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

void blank_texture(gsb_rtx* r) {                  // createBlankTexture, RH:23-36: 8x8 of (0x80, 0x80, 0x80, 0xFF)
    r->texW = r->texH = 8;
    r->tex.assign(8 * 8 * 4, 0x80);
    for (size_t i = 3; i < r->tex.size(); i += 4) r->tex[i] = 0xFF;
}

int upload_texture(gsb_rtx* r) {
    if (r->d_tex) { cudaFree(r->d_tex); r->d_tex = nullptr; }
    RTX_CUDA_OK(cudaMalloc(&r->d_tex, r->tex.size()));
    RTX_CUDA_OK(cudaMemcpy(r->d_tex, r->tex.data(), r->tex.size(), cudaMemcpyHostToDevice));
    return GSB_OK;
}

int buffer_geometry(gsb_rtx* r) {                 // RtxHost::bufferGeometry, RH:77-105
    std::vector<gsbrtx::Node> nodes;
    std::vector<gsbrtx::Tri> tris;
    try {
        gsbrtx::build_bvh(r->mesh.vertices, r->mesh.triangles, nodes, tris);
    } catch (const std::exception& e) {
        return gsbapp::fail(e.what());
    }
    cudaFree(r->d_nodes); cudaFree(r->d_tris); cudaFree(r->d_texcoords);
    r->d_nodes = nullptr; r->d_tris = nullptr; r->d_texcoords = nullptr;
    r->n_tris = (int)tris.size();
    static_assert(sizeof(gsbrtx::Node) == 32 && sizeof(gsbrtx::Tri) == 48, "record sizes are part of the kernel's load pattern");
    RTX_CUDA_OK(cudaMalloc(&r->d_nodes, nodes.size() * sizeof(gsbrtx::Node)));
    RTX_CUDA_OK(cudaMemcpy(r->d_nodes, nodes.data(), nodes.size() * sizeof(gsbrtx::Node), cudaMemcpyHostToDevice));
    if (!tris.empty()) {
        RTX_CUDA_OK(cudaMalloc(&r->d_tris, tris.size() * sizeof(gsbrtx::Tri)));
        RTX_CUDA_OK(cudaMemcpy(r->d_tris, tris.data(), tris.size() * sizeof(gsbrtx::Tri), cudaMemcpyHostToDevice));
        RTX_CUDA_OK(cudaMalloc(&r->d_texcoords, r->mesh.texcoords.size() * sizeof(float)));
        RTX_CUDA_OK(cudaMemcpy(r->d_texcoords, r->mesh.texcoords.data(), r->mesh.texcoords.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    if (!r->d_tex) { if (int e = upload_texture(r)) return e; }
    r->model_present = true;
    return GSB_OK;
}

int launch(gsb_rtx* r, uint32_t* fb0, uint32_t* fb1, int width, int height, const gsb_camera_desc* camera, const float bg0[3],
           const float bg1[3], int samples, const float* orbs, int n_orbs, cudaStream_t s) {
    if (!r || !fb0 || !camera || width <= 0 || height <= 0 || samples < 0 || (n_orbs > 0 && !orbs))
        return gsbapp::fail("gsb_rtx_render: invalid argument");
    const size_t bytes = (size_t)width * height * sizeof(uint32_t);
    if (!r->model_present) {                       // RH:220
        RTX_CUDA_OK(cudaMemsetAsync(fb0, 0, bytes, s));
        if (fb1) RTX_CUDA_OK(cudaMemsetAsync(fb1, 0, bytes, s));
        r->last_rays = 0;
        return GSB_OK;
    }
    if (!r->d_rays) RTX_CUDA_OK(cudaMalloc(&r->d_rays, sizeof(unsigned long long)));
    RTX_CUDA_OK(cudaMemsetAsync(r->d_rays, 0, sizeof(unsigned long long), s));
    RtxParams p;
    memset(&p, 0, sizeof(p));
    p.fb0 = fb0; p.fb1 = fb1; p.W = width; p.H = height; p.samples = samples;
    for (int k = 0; k < 3; k++) { p.bg0[k] = bg0[k]; p.bg1[k] = bg1 ? bg1[k] : bg0[k]; p.cam[k] = camera->location[k]; }
    p.nodes = r->d_nodes; p.tris = r->d_tris; p.texcoords = r->d_texcoords; p.tex = r->d_tex; p.texW = r->texW; p.texH = r->texH;
    p.n_tris = r->n_tris;
    if (int e = gsb_camera_inverse_projview(camera, (float)width / (float)height, p.M)) return e;      // RH:212
    if (n_orbs > 0) {
        if (n_orbs > r->orbs_cap) {
            cudaFree(r->d_orbs); r->d_orbs = nullptr;
            RTX_CUDA_OK(cudaMalloc(&r->d_orbs, (size_t)n_orbs * 3 * sizeof(float)));
            r->orbs_cap = n_orbs;
        }
        RTX_CUDA_OK(cudaMemcpyAsync(r->d_orbs, orbs, (size_t)n_orbs * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
        p.orbs = r->d_orbs; p.n_orbs = n_orbs;
    }
    p.ray_count = r->d_rays;
    const dim3 grid((unsigned)((width + 15) / 16), (unsigned)((height + 7) / 8));
    rtx_trace_kernel<<<grid, 128, 0, s>>>(p);
    RTX_CUDA_OK(cudaGetLastError());
    RTX_CUDA_OK(cudaMemcpyAsync(&r->last_rays, r->d_rays, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    RTX_CUDA_OK(cudaStreamSynchronize(s));         // owlRayGenLaunch2D is synchronous too
    return GSB_OK;
}

}  // namespace

extern "C" {

int gsb_rtx_create(gsb_rtx** out) {
    if (!out) return gsbapp::fail("gsb_rtx_create: invalid argument");
    gsb_rtx* r = new gsb_rtx();
    blank_texture(r);
    *out = r;
    return GSB_OK;
}

int gsb_rtx_destroy(gsb_rtx* r) {
    if (!r) return GSB_OK;
    cudaFree(r->d_nodes); cudaFree(r->d_tris); cudaFree(r->d_texcoords); cudaFree(r->d_tex); cudaFree(r->d_orbs); cudaFree(r->d_rays);
    delete r;
    return GSB_OK;
}

int gsb_rtx_reset(gsb_rtx* r) {
    if (!r) return gsbapp::fail("gsb_rtx_reset: invalid argument");
    r->model_present = false;
    blank_texture(r);
    if (r->d_tex) { cudaFree(r->d_tex); r->d_tex = nullptr; }
    return GSB_OK;
}

int gsb_rtx_load_model(gsb_rtx* r, const char* obj_path) {
    if (!r || !obj_path) return gsbapp::fail("gsb_rtx_load_model: invalid argument");
    if (!*obj_path) return GSB_OK;                 // RH:108: empty or missing path is silently ignored
    {
        FILE* f = fopen(obj_path, "rb");
        if (!f) return GSB_OK;
        fclose(f);
    }
    gsbapp::ObjMesh mesh;
    try {
        gsbapp::parse_obj(obj_path, mesh, true);
    } catch (const std::exception& e) {
        return gsbapp::fail(e.what());
    }
    r->mesh = std::move(mesh);
    return buffer_geometry(r);
}

int gsb_rtx_set_mesh(gsb_rtx* r, const float* vertices, int nv, const int* triangles, int nt, const float* texcoords) {
    if (!r || nv < 0 || nt < 0 || (nv > 0 && !vertices) || (nt > 0 && !triangles)) return gsbapp::fail("gsb_rtx_set_mesh: invalid argument");
    r->mesh.vertices.assign(vertices, vertices + (size_t)nv * 3);
    r->mesh.triangles.assign(triangles, triangles + (size_t)nt * 3);
    if (texcoords) r->mesh.texcoords.assign(texcoords, texcoords + (size_t)nt * 6);
    else r->mesh.texcoords.assign((size_t)nt * 6, 0.0f);
    return buffer_geometry(r);
}

int gsb_rtx_set_texture_diffuse(gsb_rtx* r, const uint8_t* rgba, int width, int height) {
    if (!r || !rgba || width <= 0 || height <= 0) return gsbapp::fail("gsb_rtx_set_texture_diffuse: invalid argument");
    r->tex.assign(rgba, rgba + (size_t)width * height * 4);
    r->texW = width; r->texH = height;
    return upload_texture(r);                      // RH:188: `if (modelPresent) bufferGeometry()` only re-binds the texture
}

int gsb_rtx_load_texture_diffuse(gsb_rtx* r, const char* png_path) {
    if (!r || !png_path) return gsbapp::fail("gsb_rtx_load_texture_diffuse: invalid argument");
    if (!*png_path) return GSB_OK;                 // RH:185
    {
        FILE* f = fopen(png_path, "rb");
        if (!f) return GSB_OK;
        fclose(f);
    }
    uint8_t* px = nullptr;
    int w = 0, h = 0;
    if (int e = gsb_png_read(png_path, &px, &w, &h)) return e;
    const int rc = gsb_rtx_set_texture_diffuse(r, px, w, h);
    gsb_free(px);
    return rc;
}

int gsb_rtx_model_present(gsb_rtx* r) { return r && r->model_present ? 1 : 0; }

int gsb_rtx_render(gsb_rtx* r, uint32_t* dev_frame_buffer, int width, int height, const gsb_camera_desc* camera,
                   const float background[3], int samples, const float* splat_cameras, int n_splat_cameras, void* stream) {
    if (!background) return gsbapp::fail("gsb_rtx_render: invalid argument");
    return launch(r, dev_frame_buffer, nullptr, width, height, camera, background, nullptr, samples, splat_cameras, n_splat_cameras,
                  (cudaStream_t)stream);
}

int gsb_rtx_render_pair(gsb_rtx* r, uint32_t* dev_frame_white, uint32_t* dev_frame_black, int width, int height,
                        const gsb_camera_desc* camera, int samples, void* stream) {
    if (!dev_frame_black) return gsbapp::fail("gsb_rtx_render_pair: invalid argument");
    const float white[3] = {1.0f, 1.0f, 1.0f}, black[3] = {0.0f, 0.0f, 0.0f};
    return launch(r, dev_frame_white, dev_frame_black, width, height, camera, white, black, samples, nullptr, 0, (cudaStream_t)stream);
}

int gsb_rtx_last_ray_count(gsb_rtx* r, unsigned long long* out) {
    if (!r || !out) return gsbapp::fail("gsb_rtx_last_ray_count: invalid argument");
    *out = r->last_rays;
    return GSB_OK;
}

}  // extern "C"
