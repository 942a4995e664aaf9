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
