"""Deterministic synthetic scenes for the parity tests and bench.py (no meshes, no network).

Counter-based RNG (splitmix64 -> float), so a scene depends only on (seed, size): the same arrays
are fed to this library, to the compiled reference (oracle/_ref) and to the CPU oracle.
Distributions follow SURVEY.md section 8(d): positions uniform in the reference's own init box
[-4,4]^3 (src/ui/UiFrame.cpp:140-141), log-uniform scales between the cull threshold and the init
value, random unit quaternions stored (w,x,y,z), opacity U[0.05,1], SH degree 1 (4 coefficients).
"""
import math

import numpy as np

from .camera import Camera, CameraSphere, Project

SH_C0 = 0.28209479177387814


def _splitmix64(idx, seed):
    with np.errstate(over="ignore"):
        z = idx.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform(n, seed, stream):
    """n doubles in [0,1), stream selects an independent sequence."""
    idx = np.arange(n, dtype=np.uint64) + np.uint64(stream) * np.uint64(1 << 40)
    return (_splitmix64(idx, seed) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def normal(n, seed, stream):
    u1 = np.maximum(uniform(n, seed, stream), 1e-300)
    u2 = uniform(n, seed, stream + 1)
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * math.pi * u2)


def splat_cloud(count, seed, sh_coeffs=4, box=4.0, scale_lo=0.005, scale_hi=0.05):
    """Returns dict of flat float32 arrays: locations[3P], shs[3MP], scales[3P], opacities[P], rotations[4P]."""
    P = int(count)
    loc = (uniform(3 * P, seed, 0) * 2.0 - 1.0) * box
    sc = np.exp(math.log(scale_lo) + uniform(3 * P, seed, 2) * (math.log(scale_hi) - math.log(scale_lo)))
    q = normal(4 * P, seed, 4).reshape(P, 4)
    q = q / np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-12)
    op = 0.05 + 0.95 * uniform(P, seed, 6)
    sh = np.zeros((P, sh_coeffs, 3), dtype=np.float64)
    sh[:, 0, :] = (uniform(3 * P, seed, 8).reshape(P, 3) - 0.5) / SH_C0
    if sh_coeffs > 1:
        sh[:, 1:, :] = 0.1 * normal(3 * P * (sh_coeffs - 1), seed, 10).reshape(P, sh_coeffs - 1, 3)
    f = np.float32
    return {"locations": loc.astype(f), "shs": sh.reshape(-1).astype(f), "scales": sc.astype(f),
            "opacities": op.astype(f), "rotations": q.reshape(-1).astype(f)}


def project_for(config):
    """Project settings of the BASELINE.json configs (cameras, learning rates)."""
    p = Project()
    if config == "C1":
        p.sphere1 = CameraSphere(count=1, distance=10.0, fovDeg=60.0)
        p.sphere2 = CameraSphere(count=0, fovDeg=30.0)
    elif config == "C2":
        p.sphere1 = CameraSphere(count=8, distance=10.0, fovDeg=60.0)
        p.sphere2 = CameraSphere(count=0, fovDeg=30.0)
    elif config == "C3":
        p.sphere1 = CameraSphere(count=16, distance=10.0, fovDeg=60.0)
        p.sphere2 = CameraSphere(count=16, distance=10.0, fovDeg=20.0)
        p.lrSh = p.lrOpacity = 0.01
    elif config == "C4":
        p.sphere1 = CameraSphere(count=32, distance=10.0, fovDeg=60.0)
        p.sphere2 = CameraSphere(count=32, distance=10.0, fovDeg=20.0)
    elif config == "C5":
        p.sphere1 = CameraSphere(count=64, distance=10.0, fovDeg=60.0)
        p.sphere2 = CameraSphere(count=64, distance=10.0, fovDeg=20.0)
    else:
        raise ValueError(config)
    return p


CONFIGS = {   # name: (width, height, splats)
    "C1": (256, 256, 10000), "C2": (512, 512, 10000), "C3": (1024, 1024, 50000),
    "C4": (1920, 1080, 500000), "C5": (1920, 1080, 3000000),
}


def shard_cameras(cameras, n1, rank, world):
    """Camera partition of SURVEY 8(e): rank r owns cameras {i : i = r mod world} of each sphere, so
    every rank gets the same wide / close-up mix."""
    s1, s2 = cameras[:n1], cameras[n1:]
    return s1[rank::world] + s2[rank::world]


def textured_mesh(rings=24, segments=48, seed=7, radius=2.5, bumps=0.35, floor=True):
    """Procedural textured mesh for the truth ray tracer (north_star: "a procedural textured mesh"): a bumpy UV sphere
    (radius modulated by a few low-frequency lobes) over an optional floor quad.  Returns (vertices[nv,3] f32,
    triangles[nt,3] i32 zero-based, texcoords[nt,3,2] f32 per corner) -- the three arrays RtxHost buffers
    (src/rtx/RtxHost.h:20-22)."""
    rng_phase = uniform(6, seed, 40) * 2.0 * math.pi
    th = np.linspace(0.0, math.pi, rings + 1)
    ph = np.linspace(0.0, 2.0 * math.pi, segments + 1)
    T, Pp = np.meshgrid(th, ph, indexing="ij")
    r = radius * (1.0 + bumps * (0.5 * np.sin(3 * T + rng_phase[0]) * np.cos(2 * Pp + rng_phase[1]) +
                                 0.3 * np.sin(5 * Pp + rng_phase[2]) * np.sin(T) + 0.2 * np.cos(4 * T + rng_phase[3])))
    x, y, z = r * np.sin(T) * np.cos(Pp), r * np.cos(T), r * np.sin(T) * np.sin(Pp)
    verts = np.stack([x, y, z], -1).reshape(-1, 3)
    uv = np.stack([Pp / (2.0 * math.pi) * 3.0, 1.0 - T / math.pi], -1).reshape(-1, 2)      # u repeats 3x: exercises wrap
    tris, tcs = [], []
    idx = lambda i, j: i * (segments + 1) + j
    for i in range(rings):
        for j in range(segments):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
            for t in ((a, b, c), (a, c, d)):
                tris.append(t)
                tcs.append([uv[t[0]], uv[t[1]], uv[t[2]]])
    if floor:
        n0 = verts.shape[0]
        fy = -radius * (1.0 + bumps) - 0.25
        verts = np.concatenate([verts, np.array([[-8, fy, -8], [8, fy, -8], [8, fy, 8], [-8, fy, 8]], np.float64)])
        fuv = np.array([[0, 0], [4, 0], [4, 4], [0, 4]], np.float64)
        for t in ((0, 2, 1), (0, 3, 2)):
            tris.append((n0 + t[0], n0 + t[1], n0 + t[2]))
            tcs.append([fuv[t[0]], fuv[t[1]], fuv[t[2]]])
    return verts.astype(np.float32), np.asarray(tris, np.int32), np.asarray(tcs, np.float32)


def procedural_texture(size=64, seed=11):
    """RGBA8 [size, size, 4]: coloured checker with smooth gradients; every 4th cell is see-through (alpha 0) and a band is
    half transparent, so the alpha test of the closest-hit program (src/rtx/RtxDevice.cu:124) takes both branches."""
    yy, xx = np.mgrid[0:size, 0:size]
    cell = (xx // (size // 8) + yy // (size // 8))
    base = uniform(3 * 64, seed, 50).reshape(64, 3)
    rgb = base[(cell * 7 + (xx // (size // 8))) % 64] * 0.6 + 0.35
    rgb = rgb * (0.75 + 0.25 * np.sin(xx / size * 2 * math.pi)[..., None])
    alpha = np.full((size, size), 255.0)
    alpha[(cell % 4) == 3] = 0.0
    alpha[size // 2:size // 2 + size // 16] = 128.0
    out = np.concatenate([np.clip(rgb, 0, 1) * 255.0, alpha[..., None]], -1)
    return np.ascontiguousarray(out.astype(np.uint8))


def write_obj(path, vertices, triangles, texcoords):
    """The mesh as an OBJ file the reference's loader reads (v / vt / f a/ta b/tb c/tc, src/rtx/RtxHost.cpp:119-160)."""
    with open(path, "w") as f:
        for v in vertices:
            f.write("v %.9g %.9g %.9g\n" % tuple(v))
        flat = texcoords.reshape(-1, 2)
        for t in flat:
            f.write("vt %.9g %.9g\n" % tuple(t))
        for k, t in enumerate(triangles):
            f.write("f %d/%d %d/%d %d/%d\n" % (t[0] + 1, 3 * k + 1, t[1] + 1, 3 * k + 2, t[2] + 1, 3 * k + 3))
