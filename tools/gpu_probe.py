"""Developer probe (GPU box): run this library and the compiled reference on the same scene and print a
field-by-field diff with diagnostics.  Not part of the test-suite; tests/ holds the asserting version.

usage: python tools/gpu_probe.py [config ...]   e.g. C1 C2 odd
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers as H  # noqa: E402
from oracle import refshim  # noqa: E402

gsb = H.gsb


def probe(name, W, Hh, P, seed, cam_idx=0, fov=60.0, ncam=8, bg=(1.0, 1.0, 1.0), backward=True, scale_hi=0.05):
    dev = "cuda"
    cloud = gsb.synth.splat_cloud(P, seed, scale_hi=scale_hi)
    sc = H.scene_tensors(cloud, dev)
    proj = gsb.Project()
    proj.sphere1 = gsb.CameraSphere(count=ncam, distance=10.0, fovDeg=fov)
    proj.sphere2 = gsb.CameraSphere(count=0)
    cam = gsb.Camera.getCameras(proj)[cam_idx]
    consts = cam.view_constants(W, Hh, square_fov=(W == Hh))
    vw = H.view_tensors(consts, bg, dev)
    D, M = 1, 4
    res = {"name": name, "W": W, "H": Hh, "P": P}

    out_ref, R_ref = H.run_forward_ref(sc, vw, P, D, M, W, Hh)
    st_ref = refshim.state(P, W, Hh, R_ref, dev)
    torch.cuda.synchronize()
    ws, out, R = H.run_forward_gsb(sc, vw, P, D, M, W, Hh)
    torch.cuda.synchronize()
    st = gsb.state(ws, P, W, Hh, R)
    res["R"] = [R, R_ref]
    vis = st_ref["radii"] > 0
    res["visible"] = [int((st["radii"] > 0).sum()), int(vis.sum())]

    def mism(a, b, mask=None):
        d = (a != b)
        if mask is not None:
            d = d & (mask if d.dim() == 1 else mask[:, None])
        return int(d.sum())

    res["radii_mismatch"] = mism(st["radii"], st_ref["radii"])
    res["tiles_touched_mismatch"] = mism(st["tiles_touched"], st_ref["tiles_touched"])
    res["depth_bits_mismatch"] = mism(st["depths"].view(torch.int32), st_ref["depths"].view(torch.int32), vis)
    res["means2D_bits_mismatch"] = mism(st["means2D"].view(torch.int32), st_ref["means2D"].view(torch.int32), vis)
    res["cov3D_mismatch"] = mism(st["cov3D"], st_ref["cov3D"], vis)
    res["conic_opacity_mismatch"] = mism(st["conic_opacity"], st_ref["conic_opacity"], vis)
    res["rgb_maxabs"] = float((st["rgb"] - st_ref["rgb"])[vis].abs().max()) if vis.any() else 0.0
    res["rgb_bits_mismatch"] = mism(st["rgb"].view(torch.int32), st_ref["rgb"].view(torch.int32), vis)
    cl = torch.stack([(st["clamped"] >> k) & 1 for k in range(3)], dim=1)
    res["clamped_mismatch"] = mism(cl, st_ref["clamped"], vis)
    if R == R_ref and R > 0:
        res["point_list_mismatch"] = mism(st["point_list"], st_ref["point_list"])
        res["keys_mismatch"] = mism(st["point_list_keys"], st_ref["point_list_keys"])
        res["ranges_mismatch"] = mism(st["ranges"], st_ref["ranges"])
    res["n_contrib_mismatch"] = mism(st["n_contrib"], st_ref["n_contrib"])
    res["final_T_maxabs"] = float((st["final_T"] - st_ref["final_T"]).abs().max())
    res["image_maxabs"] = float((out - out_ref).abs().max())
    res["image_bits_mismatch"] = mism(out.view(torch.int32), out_ref.view(torch.int32))
    if res["radii_mismatch"]:
        bad = torch.nonzero(st["radii"] != st_ref["radii"]).flatten()[:5]
        res["radii_bad"] = [(int(i), int(st["radii"][i]), int(st_ref["radii"][i])) for i in bad]
    if res.get("cov3D_mismatch"):
        bad = torch.nonzero(((st["cov3D"] != st_ref["cov3D"]) & vis[:, None]).any(dim=1)).flatten()[:3]
        res["cov3D_bad"] = [(int(i), st["cov3D"][i].tolist(), st_ref["cov3D"][i].tolist()) for i in bad]
    if res.get("conic_opacity_mismatch"):
        bad = torch.nonzero(((st["conic_opacity"] != st_ref["conic_opacity"]) & vis[:, None]).any(dim=1)).flatten()[:3]
        res["conic_bad"] = [(int(i), st["conic_opacity"][i].tolist(), st_ref["conic_opacity"][i].tolist()) for i in bad]

    if backward:
        torch.manual_seed(seed)
        dL = (torch.rand(3 * W * Hh, device=dev) - 0.5)
        g_ref = H.run_backward_ref(sc, vw, P, D, M, R_ref, W, Hh, dL)
        g_ref2 = H.run_backward_ref(sc, vw, P, D, M, R_ref, W, Hh, dL)
        g = H.run_backward_gsb(ws, sc, vw, P, D, M, R, W, Hh, dL)
        torch.cuda.synchronize()
        res["grad_rel_l2"] = {n: H.rel_l2(g[n], g_ref[n]) for n in H.GRAD_NAMES}
        res["ref_self_rel_l2"] = {n: H.rel_l2(g_ref2[n], g_ref[n]) for n in H.GRAD_NAMES}

    # timing (device events, warm)
    def timeit(fn, n=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    res["ms_fwd_ref_pooled"] = timeit(lambda: H.run_forward_ref(sc, vw, P, D, M, W, Hh))
    res["ms_fwd_gsb_compat"] = timeit(lambda: H.run_forward_gsb(sc, vw, P, D, M, W, Hh))
    if backward:
        res["ms_bwd_ref"] = timeit(lambda: H.run_backward_ref(sc, vw, P, D, M, R_ref, W, Hh, dL))
        res["ms_bwd_gsb"] = timeit(lambda: H.run_backward_gsb(ws, sc, vw, P, D, M, R, W, Hh, dL))
    return res


CASES = {
    "tiny": dict(W=64, Hh=48, P=300, seed=11),
    "C1": dict(W=256, Hh=256, P=10000, seed=0x5EED0001, ncam=1),
    "odd": dict(W=250, Hh=130, P=5000, seed=7),
    "C2": dict(W=512, Hh=512, P=10000, seed=0x5EED0002, cam_idx=3),
    "C3": dict(W=1024, Hh=1024, P=50000, seed=0x5EED0003, cam_idx=5, ncam=16),
    "C4": dict(W=1920, Hh=1080, P=500000, seed=0x5EED0004, cam_idx=5, ncam=32),
    "C4close": dict(W=1920, Hh=1080, P=500000, seed=0x5EED0004, cam_idx=5, ncam=32, fov=20.0),
    "big": dict(W=640, Hh=360, P=20000, seed=3, scale_hi=1.5),
}

if __name__ == "__main__":
    names = sys.argv[1:] or ["tiny", "C1", "odd"]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    allres = []
    for n in names:
        t0 = time.time()
        r = probe(n, **CASES[n])
        r["wall_s"] = time.time() - t0
        print(json.dumps(r), flush=True)
        allres.append(r)
    with open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w") as f:
        json.dump(allres, f, indent=1)
