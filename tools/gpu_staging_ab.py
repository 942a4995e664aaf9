"""A/B on the GPU box: staging of the render kernels' records with 16-byte cp.async (what ships)
against the sm_100 TMA row gather (cp.async.bulk.tensor.2d tile::gather4), csrc/measure.cu staging_ab_kernel, over the tile
lists of a few C4 cameras.  Prints one JSON line (kept under profiles/)."""
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
nper = int(sys.argv[2]) if len(sys.argv) > 2 else 2
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
    ms = (C.c_float * 3)()
    gsb._lib.check(gsb.lib().gsb_trainer_measure_staging(tr._h, i, 5, ms))
    pc = tr.count_pairs(i)
    rows.append({"camera": i, "fov": cams[i].fovDegY, "R": pc["R"], "cp_async_ms": round(ms[0], 4),
                 "tma_gather4_8_lanes_ms": round(ms[1], 4), "tma_gather4_1_lane_ms": round(ms[2], 4)})
hist = [0] * 40
for i in range(len(cams)):
    o = (C.c_ulonglong * 40)()
    gsb._lib.check(gsb.lib().gsb_trainer_reduction_stats(tr._h, i, o))
    for q in range(40):
        hist[q] += int(o[q])
reductions = sum(hist[1:33])
stats = {"cameras": len(cams), "reductions_per_camera": reductions / len(cams), "lanes_histogram_share": [round(h / max(reductions, 1), 4) for h in hist[1:33]],
         "direct_share(<=4 lanes)": round(sum(hist[1:5]) / max(reductions, 1), 4), "panel_per_camera": hist[33] / len(cams),
         "panel_confined_rows_half": round(hist[34] / max(hist[33], 1), 4), "panel_confined_cols_half": round(hist[35] / max(hist[33], 1), 4),
         "panel_confined_either": round(hist[36] / max(hist[33], 1), 4), "blended_pairs_per_camera": hist[37] / len(cams),
         "box_survivors_per_camera": hist[38] / len(cams), "mean_lanes_per_reduction": round(sum(q * hist[q] for q in range(1, 33)) / max(reductions, 1), 2)}
mean = lambda k: round(sum(r[k] for r in rows) / len(rows), 4)
print(json.dumps({"config": cfg, "cameras": len(cams), "mean_cp_async_ms": mean("cp_async_ms"),
                  "mean_tma_gather4_8_lanes_ms": mean("tma_gather4_8_lanes_ms"), "mean_tma_gather4_1_lane_ms": mean("tma_gather4_1_lane_ms"),
                  "rows": rows, "reduction_stats": stats,
                  "note": "staging only: every warp of a tile walks the tile's whole list in 32-entry chunks, double-buffered, and reads one "
                          "word of each record; best of 5 launches, CUDA events; -1 = the copies of that mode did not complete"}))
tr.close()
