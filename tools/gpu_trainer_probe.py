"""Developer probe (GPU box): Trainer-level parity against the reference replay (oracle/_ref):
running means after one iteration, parameters after apply, densify result as a multiset.

usage: python tools/gpu_trainer_probe.py [config] [cams_per_sphere]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import helpers as H  # noqa: E402
from gpu_stage_probe import make_truths  # noqa: E402
from oracle import refshim  # noqa: E402

gsb = H.gsb
NAMES = ("var", "loc", "sh", "scale", "opacity", "rot")


def split_avg(avg, P, M):
    sizes = [P, 3 * P, 3 * M * P, 3 * P, P, 4 * P]
    out, o = [], 0
    for s in sizes:
        out.append(avg[o:o + s])
        o += s
    return out


def sort_rows(model, M):
    """Canonical row order for multiset comparison: lexicographic on (location, scale, opacity)."""
    c = model["opacities"].shape[0]
    rows = np.concatenate([model["locations"].reshape(c, 3), model["scales"].reshape(c, 3), model["opacities"].reshape(c, 1),
                           model["rotations"].reshape(c, 4), model["shs"].reshape(c, 3 * M)], axis=1)
    order = np.lexsort(rows[:, ::-1].T)
    return rows[order]


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
    nper = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    W, Hh, P = gsb.synth.CONFIGS[cfg]
    dev = "cuda"
    project = gsb.synth.project_for(cfg)
    project.sphere1.count = nper
    project.sphere2.count = nper if project.sphere2.count > 0 else 0
    cams = gsb.Camera.getCameras(project)
    seed = 0x5EED0000 + int(cfg[1])
    cloud = gsb.synth.splat_cloud(P, seed)
    teacher = gsb.synth.splat_cloud(P // 2, seed + 1)
    tw, tb, consts = make_truths(teacher, cams, W, Hh, dev)
    host = gsb.ModelSplatsHost.from_arrays(cloud["locations"], cloud["shs"], cloud["scales"], cloud["opacities"],
                                           cloud["rotations"])
    M = 4
    p13 = project.params13()

    rt = refshim.RefTrainer(W, Hh, dev)
    rt.set_model(cloud, host.capacity, 1, M)
    rt.set_truths(tw, tb, consts)
    rt.train(p13, as_shipped=False, apply=False)
    ref_avg = rt.buffers()
    rt.train(p13, as_shipped=False, apply=False)
    ref_avg2 = rt.buffers()
    res = {"cfg": cfg, "cams": len(cams), "ref_self": {n: H.rel_l2(a, b) for n, a, b in zip(NAMES, ref_avg2, ref_avg)}}

    for fuse in (False, True):
        tr = gsb.Trainer(W, Hh, dev)
        tr.set_fuse_backgrounds(fuse)
        tr.model = gsb.ModelSplatsDevice(host, dev)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
        g = tr.grad_tensor()
        tr.accumulate()
        torch.cuda.synchronize()
        mine = split_avg(g.clone(), P, M)
        res["fused" if fuse else "unfused"] = {n: H.rel_l2(a, b) for n, a, b in zip(NAMES, mine, ref_avg)}
        res["stats_" + ("fused" if fuse else "unfused")] = tr.stats()
        tr.close()
    print(json.dumps(res), flush=True)

    # apply + densify parity: thresholds calibrated so that all three classes occur
    var, gl = ref_avg[0].cpu().numpy(), ref_avg[1].cpu().numpy().reshape(-1, 3)
    score = var - np.linalg.norm(gl, axis=1)
    project.paramDensifyVariance = float(np.quantile(score, 0.90))
    project.paramCullOpacity = 0.10
    project.paramSplitSize = float(np.median(np.linalg.norm(cloud["scales"].reshape(-1, 3), axis=1)))
    project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.01, 0.01, 0.001, 0.01, 0.001
    p13 = project.params13()
    rt.train(p13, as_shipped=False, apply=True)
    ref_after_apply = rt.get_model()
    rt.densify(p13)
    ref_model = rt.get_model()

    tr = gsb.Trainer(W, Hh, dev)
    tr.set_fuse_backgrounds(False)
    tr.model = gsb.ModelSplatsDevice(host, dev)
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    tr.train(project, False)
    mine_after_apply = gsb.ModelSplatsHost.from_device(tr.model)
    d = {}
    for k in ("locations", "shs", "scales", "opacities", "rotations"):
        a, b = getattr(mine_after_apply, k), ref_after_apply[k]
        d[k] = float(np.abs(a - b).max())
    res2 = {"apply_maxabs": d}
    project.iterations = 0
    tr.densify(project)
    mh = gsb.ModelSplatsHost.from_device(tr.model)
    res2["count"] = [mh.count, int(ref_model["opacities"].shape[0]), P]
    if mh.count == ref_model["opacities"].shape[0]:
        mine_rows = sort_rows({k: getattr(mh, k) for k in ("locations", "shs", "scales", "opacities", "rotations")}, M)
        ref_rows = sort_rows(ref_model, M)
        diff = np.abs(mine_rows - ref_rows)
        finite = np.isfinite(mine_rows) & np.isfinite(ref_rows)
        res2["densify_maxabs"] = float(diff[finite].max())
        res2["nonfinite"] = [int((~np.isfinite(mine_rows)).sum()), int((~np.isfinite(ref_rows)).sum())]
    print(json.dumps(res2), flush=True)
    rt.close()


if __name__ == "__main__":
    main()
