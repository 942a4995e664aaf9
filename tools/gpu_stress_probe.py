"""Developer probe (GPU box): BASELINE config C5 shape on one GPU -- 3 M splats, 1920x1080, densify / split / clone / cull
enabled on every iteration, a reduced camera set.  Prints count, time and memory per iteration as JSON lines.

usage: python tools/gpu_stress_probe.py [cams_per_sphere] [iterations]"""
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

gsb = H.gsb


def main():
    nper = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    W, Hh, P = gsb.synth.CONFIGS["C5"]
    dev = "cuda"
    project = gsb.synth.project_for("C5")
    project.sphere1.count = project.sphere2.count = nper
    cams = gsb.Camera.getCameras(project)
    cloud = gsb.synth.splat_cloud(P, 0x5EED0005)
    tw, tb, _ = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0006), cams, W, Hh, dev)
    host = gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")])
    tr = gsb.Trainer(W, Hh, dev, square_fov=False)
    tr.model = gsb.ModelSplatsDevice(host, dev)
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    # thresholds that make every class fire on this cloud: a few per cent split / cloned / culled per densify
    project.paramDensifyVariance = 0.0002
    project.paramCullOpacity = 0.06
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    for it in range(iters):
        before = tr.model.count
        densify = it >= 1
        e0.record()
        tr.train(project, densify)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        ok = bool(torch.isfinite(tr.model.devLocations[: 3 * tr.model.count]).all())
        print(json.dumps({"iteration": it + 1, "densify": densify, "count_before": before, "count_after": tr.model.count,
                          "ms": round(ms, 2), "views_per_s": round(2 * len(cams) / ms * 1000, 1), "finite": ok,
                          "retries": tr.stats()["retries"], "rendered_total": tr.stats()["rendered_total"],
                          "mem_GB": round(torch.cuda.memory_allocated() / 2**30, 2),
                          "mem_total_GB": round((torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 2**30, 2)}),
              flush=True)
    tr.close()


if __name__ == "__main__":
    main()
