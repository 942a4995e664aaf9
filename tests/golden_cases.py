"""Definitions of the golden cases shared by tools/make_golden.py (which runs the reference's CUDA build
on a B200 and writes tests/golden/*.npz) and by the tests that check the CPU oracle and the CUDA
product against those vectors.  Everything is derived from seeds, so the inputs are not stored."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
gsb = importlib.import_module("gaussian-splatterer_b200")

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# rasterizer-level cases: one view
RASTER_CASES = {
    #  name      W    H    P     seed  ncam cam fov   bg               scale_hi  sh_deg
    "tiny":   dict(W=64, H=48, P=300, seed=11, ncam=8, cam=0, fov=60.0, bg=(1.0, 1.0, 1.0), scale_hi=0.05),
    "odd":    dict(W=250, H=130, P=5000, seed=7, ncam=8, cam=0, fov=60.0, bg=(0.0, 0.0, 0.0), scale_hi=0.05),
    "fat":    dict(W=160, H=96, P=1500, seed=3, ncam=8, cam=2, fov=60.0, bg=(0.3, 0.6, 0.9), scale_hi=1.5),
    "close":  dict(W=128, H=128, P=4000, seed=5, ncam=8, cam=5, fov=20.0, bg=(1.0, 1.0, 1.0), scale_hi=0.05),
}

# trainer-level cases: a few cameras, one iteration + apply + densify
TRAINER_CASES = {
    "t1": dict(W=96, H=96, P=2000, seed=21, n1=3, n2=0, fov1=60.0, fov2=20.0),
    "t2": dict(W=112, H=80, P=3000, seed=22, n1=2, n2=2, fov1=60.0, fov2=20.0),
}


def raster_inputs(name):
    c = RASTER_CASES[name]
    cloud = gsb.synth.splat_cloud(c["P"], c["seed"], scale_hi=c["scale_hi"])
    project = gsb.Project()
    project.sphere1 = gsb.CameraSphere(count=c["ncam"], distance=10.0, fovDeg=c["fov"])
    project.sphere2 = gsb.CameraSphere(count=0)
    cam = gsb.Camera.getCameras(project)[c["cam"]]
    consts = cam.view_constants(c["W"], c["H"], square_fov=(c["W"] == c["H"]))
    n = 3 * c["W"] * c["H"]
    dL = (gsb.synth.uniform(n, c["seed"], 77) - 0.5).astype(np.float32)
    return c, cloud, consts, dL


def trainer_inputs(name):
    c = TRAINER_CASES[name]
    cloud = gsb.synth.splat_cloud(c["P"], c["seed"])
    teacher = gsb.synth.splat_cloud(c["P"] // 2, c["seed"] + 1)
    project = gsb.Project()
    project.sphere1 = gsb.CameraSphere(count=c["n1"], distance=10.0, fovDeg=c["fov1"])
    project.sphere2 = gsb.CameraSphere(count=c["n2"], distance=10.0, fovDeg=c["fov2"])
    # learning rates large enough for the update to be visible in float32
    project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.01, 0.01, 0.001, 0.01, 0.001
    cams = gsb.Camera.getCameras(project)
    return c, cloud, teacher, project, cams


def densify_thresholds(project, cloud, var, avg_loc):
    """Thresholds that make cull / split / clone all occur, placed in the widest gap near the target
    quantile so that no splat sits within rounding distance of a decision boundary."""
    def gap_threshold(values, q):
        v = np.sort(np.asarray(values, dtype=np.float64))
        k = int(q * (len(v) - 1))
        lo, hi = max(1, k - 20), min(len(v) - 1, k + 20)
        j = lo + int(np.argmax(v[lo + 1:hi + 1] - v[lo:hi]))
        return float(0.5 * (v[j] + v[j + 1]))
    score = np.asarray(var, np.float64) - np.linalg.norm(np.asarray(avg_loc, np.float64).reshape(-1, 3), axis=1)
    project.paramDensifyVariance = gap_threshold(score, 0.85)
    project.paramCullOpacity = gap_threshold(cloud["opacities"], 0.08)
    project.paramSplitSize = gap_threshold(np.linalg.norm(cloud["scales"].reshape(-1, 3), axis=1), 0.5)
    return project


def load(name):
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    if not os.path.exists(path):
        return None
    return dict(np.load(path))
