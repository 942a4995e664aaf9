"""A/B on the GPU box: the shipped fused render backward against the skeleton of the "lanes = list entries" mapping
(csrc/measure.cu) on the state of a few C4 cameras.  Prints one JSON line (kept under profiles/)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402

import helpers as H  # noqa: E402
from gpu_stage_probe import make_truths  # noqa: E402

gsb = H.gsb
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
nper = int(sys.argv[2]) if len(sys.argv) > 2 else 3
W, Hh, P = gsb.synth.CONFIGS[cfg]
project = gsb.synth.project_for(cfg)
project.sphere1.count = nper
project.sphere2.count = nper if project.sphere2.count > 0 else 0
cams = gsb.Camera.getCameras(project)
seed = 0x5EED0000 + int(cfg[1])
cloud = gsb.synth.splat_cloud(P, seed)
tw, tb, _ = make_truths(gsb.synth.splat_cloud(P // 2, seed + 1), cams, W, Hh, "cuda")
tr = gsb.Trainer(W, Hh, "cuda")
tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")]), "cuda")
tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
tr.accumulate()
torch.cuda.synchronize()
rows = []
for i in range(len(cams)):
    ms = (C.c_float * 2)()
    gsb._lib.check(gsb.lib().gsb_trainer_measure_bwd_skeleton(tr._h, i, 5, ms))
    pc = tr.count_pairs(i)
    rows.append({"camera": i, "fov": cams[i].fovDegY, "render_bwd_kernel<2>_ms": round(ms[0], 4), "lanes_are_splats_skeleton_ms": round(ms[1], 4),
                 "R": pc["R"], "eval_b": pc["eval_b"], "blend_b": pc["blend_b"]})
mean = lambda k: sum(r[k] for r in rows) / len(rows)
print(json.dumps({"config": cfg, "cameras": len(cams), "mean_render_bwd_ms": round(mean("render_bwd_kernel<2>_ms"), 4),
                  "mean_skeleton_ms": round(mean("lanes_are_splats_skeleton_ms"), 4), "rows": rows,
                  "note": "skeleton = lower bound of a per-splat backward (no forward checkpoints, no exact blend decisions, no loss); "
                          "both kernels time one fused white+black launch on identical state, best of 5, CUDA events"}))
tr.close()
