// TEST INFRASTRUCTURE -- CPU restatement of the reference's truth ray tracer (src/rtx/RtxDevice.cu), used only by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker of gaussian-splatterer_b200/csrc/app/rtx_trace.cu.
// The product never links or loads this file.
//
// PARITY UNPINNED: the reference's own implementation runs inside OptiX (hardware triangle test, OWL launch) and cannot be
// built or run here (OptiX SDK absent), it ships no fixtures, and its random sequence is undefined (PerRayData::random is
// never seeded, RtxDevice.cu:57 / owl random.h:52-55).  What this file pins is the *algorithm*: the per-pixel program of
// RtxDevice.cu followed line by line, with an exhaustive (no BVH) Moeller-Trumbore intersection, so that the CUDA kernel's
// acceleration structure and traversal can be checked bit for bit.  Stated choices (same as the kernel's header): LCG
// seeded with (pixel.x, pixel.y); vec3f(random(), random(), random()) evaluated x, y, z; ties at equal t go to the lowest
// primitive index; tex2D = nearest texel of frac(u) * size, value / 255.  Compile with -ffp-contract=off.
#include <cmath>
#include <cstdint>
#include <cstring>

namespace {

struct V3 { float x, y, z; };
inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                       // owl vec functors: left to right
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline V3 normalize(V3 v) { return v * (1.f / std::sqrt(dot(v, v))); }                          // owl vec.h:355 + owl-common.h:262

struct LCG {                                                                                     // owl random.h:46-84, N = 4
    uint32_t state;
    void init(uint32_t val0, uint32_t val1) {
        uint32_t v0 = val0, v1 = val1, s0 = 0;
        for (unsigned n = 0; n < 4; n++) {
            s0 += 0x9e3779b9;
            v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4);
            v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761e);
        }
        state = v0;
    }
    float operator()() {
        state = 1664525u * state + 1013904223u;
        return (state & 0x00FFFFFF) / (float)0x01000000;
    }
};

struct Scene {
    const float* vertices; const int* triangles; int nt; const float* texcoords; const uint8_t* tex; int texW, texH;
};

struct Hit { float t, u, v; int prim; };

// optixTrace stand-in: every triangle is tested (this is the checker of the kernel's BVH)
bool closest_hit(const Scene& sc, V3 o, V3 d, float tmin, float tmax, Hit& hit) {
    hit.t = tmax; hit.prim = 0x7fffffff; hit.u = hit.v = 0.f;
    bool found = false;
    for (int prim = 0; prim < sc.nt; prim++) {
        const float* a = sc.vertices + 3 * sc.triangles[3 * prim];
        const float* b = sc.vertices + 3 * sc.triangles[3 * prim + 1];
        const float* c = sc.vertices + 3 * sc.triangles[3 * prim + 2];
        const V3 v0 = mk(a[0], a[1], a[2]), e1 = mk(b[0] - a[0], b[1] - a[1], b[2] - a[2]), e2 = mk(c[0] - a[0], c[1] - a[1], c[2] - a[2]);
        const V3 pvec = cross(d, e2);
        const float det = dot(e1, pvec);
        if (det == 0.f) continue;
        const float idet = 1.f / det;
        const V3 tvec = o - v0;
        const float u = dot(tvec, pvec) * idet;
        if (!(u >= 0.f && u <= 1.f)) continue;
        const V3 qvec = cross(tvec, e1);
        const float v = dot(d, qvec) * idet;
        if (!(v >= 0.f && u + v <= 1.f)) continue;
        const float t = dot(e2, qvec) * idet;
        if (!(t > tmin && t < tmax)) continue;
        if (t < hit.t || (t == hit.t && prim < hit.prim)) { hit.t = t; hit.u = u; hit.v = v; hit.prim = prim; found = true; }
    }
    return found;
}

inline uint32_t make_8bit(float f) { int v = (int)(f * 256.f); v = v < 0 ? 0 : v; v = v > 255 ? 255 : v; return (uint32_t)v; }   // owl_device.h:79-82
inline uint32_t make_rgba(V3 c) { return make_8bit(c.x) | (make_8bit(c.y) << 8) | (make_8bit(c.z) << 16) | (0xffU << 24); }
inline float clamp01(float x) { return std::fmax(std::fmin(x, 1.0f), 0.0f); }

}  // namespace

extern "C" {

// One frame of RtxHost::render (src/rtx/RtxHost.cpp:191-221) for the pixel rectangle [x0, x1) x [y0, y1) (so a bounded
// sample of a large frame can be timed / compared); pixels outside the rectangle are left untouched.  M = inverse(proj * view)
// column-major, cam = camera location.  Returns the number of ray segments traced.
unsigned long long oracle_rtx_render(uint32_t* frame, int W, int H, int x0, int y0, int x1, int y1, const float* cam3, const float* M,
                                     const float* bg, int samples, const float* vertices, const int* triangles, int nt,
                                     const float* texcoords, const uint8_t* tex, int texW, int texH, const float* orbs, int n_orbs) {
    const Scene sc{vertices, triangles, nt, texcoords, tex, texW, texH};
    const V3 cam = mk(cam3[0], cam3[1], cam3[2]);
    unsigned long long rays = 0;
    for (int py = y0; py < y1; py++)
        for (int px = x0; px < x1; px++) {
            LCG random;
            random.init((uint32_t)px, (uint32_t)py);
            V3 color = mk(0.f, 0.f, 0.f);
            bool touches = false;
            for (int s = 0; s < samples; s++) {
                // rayGenProgram, RtxDevice.cu:63-77
                const float r0 = random(), r1 = random();
                const float fx = ((float)px + r0) + 0.5f, fy = ((float)py + r1) + 0.5f;
                const float vx = (float)(((double)fx * 2.0 / (double)W) - 1.0);
                const float vy = (float)(((double)fy * 2.0 / (double)H) - 1.0);
                const float vz = 1.0f;
                const float hx = vx * M[0] + vy * M[4] + vz * M[8] + M[12];
                const float hy = vx * M[1] + vy * M[5] + vz * M[9] + M[13];
                const float hz = vx * M[2] + vy * M[6] + vz * M[10] + M[14];
                const float hw = vx * M[3] + vy * M[7] + vz * M[11] + M[15];
                V3 origin = cam;
                V3 dir = normalize(mk(hx / hw, hy / hw, hz / hw) - cam);
                float tmin = 0.f, tmax = 1e30f;
                // tracePath, RtxDevice.cu:17-50
                V3 atten = mk(1.f, 1.f, 1.f);
                bool has_reflected = false;
                V3 result = mk(0.f, 0.f, 0.f);                        // "Max bounces exceeded"
                for (int i = 0; i < 50; i++) {
                    Hit h;
                    const bool hit = closest_hit(sc, origin, dir, tmin, tmax, h);
                    rays++;
                    V3 prd_color, hit_origin = mk(0.f, 0.f, 0.f), bounce = mk(0.f, 0.f, 0.f);
                    bool reflected = false;
                    if (hit) {                                          // TriangleMesh closest hit, RtxDevice.cu:102-143
                        const float* a = vertices + 3 * triangles[3 * h.prim];
                        const float* b = vertices + 3 * triangles[3 * h.prim + 1];
                        const float* c = vertices + 3 * triangles[3 * h.prim + 2];
                        const V3 normal = normalize(cross(mk(b[0] - a[0], b[1] - a[1], b[2] - a[2]), mk(c[0] - a[0], c[1] - a[1], c[2] - a[2])));
                        const float* tc = texcoords + 6 * h.prim;
                        const float w0 = 1.0f - h.u - h.v;
                        const float tcx = w0 * tc[0] + h.u * tc[2] + h.v * tc[4];
                        const float tcy = w0 * tc[1] + h.u * tc[3] + h.v * tc[5];
                        const float su = tcx, sv = 1.0f - tcy;
                        const float fu = su - std::floor(su), fv = sv - std::floor(sv);
                        int ix = (int)std::floor(fu * (float)texW), iy = (int)std::floor(fv * (float)texH);
                        ix = ix < texW - 1 ? ix : texW - 1; iy = iy < texH - 1 ? iy : texH - 1;
                        const uint8_t* tx = tex + 4 * ((size_t)iy * texW + ix);
                        const float cr = (float)tx[0] / 255.f, cg = (float)tx[1] / 255.f, cb = (float)tx[2] / 255.f, ca = (float)tx[3] / 255.f;
                        hit_origin = origin + h.t * dir;
                        if (ca > random()) {
                            V3 refl = dir - 2.0f * dot(dir, normal) * normal;
                            if (random() > 0.0f) {
                                V3 v;
                                do {
                                    const float ra = random(), rb = random(), rc = random();
                                    v = 2.0f * mk(ra, rb, rc) - mk(1.0f, 1.0f, 1.0f);
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
                    } else {                                            // miss, RtxDevice.cu:146-157
                        const V3 n = normalize(dir);
                        const float sky = std::fmin(1.0f, 1.0f + n.y);
                        prd_color = mk(sky, sky, sky);
                    }
                    atten = atten * prd_color;
                    if (reflected) has_reflected = true;
                    if (i == 0) {
                        for (int c = 0; c < n_orbs; c++) {
                            const V3 orb = mk(orbs[3 * c], orbs[3 * c + 1], orbs[3 * c + 2]);
                            if (hit && dot(dir, orb - origin) > dot(dir, hit_origin - origin)) continue;
                            const V3 closest = origin + dir * dot(dir, orb - origin);
                            const V3 delta = orb - closest;
                            if (dot(delta, delta) < 0.025f * 0.025f) { touches = true; break; }
                        }
                    }
                    if (!hit) { result = has_reflected ? atten : mk(bg[0], bg[1], bg[2]); break; }
                    origin = hit_origin; dir = bounce; tmin = 1e-3f; tmax = 1e10f;
                }
                result = result * 1.0f;                                 // PROGRAM_EXPOSURE_FACTOR
                color = color + mk(clamp01(result.x), clamp01(result.y), clamp01(result.z));
            }
            const float ns = (float)samples;
            color = mk(color.x / ns, color.y / ns, color.z / ns);
            if (touches) color = mk(1.f, 1.f, 1.f) - color;
            frame[px + (size_t)W * py] = make_rgba(color);
        }
    return rays;
}

}  // extern "C"
