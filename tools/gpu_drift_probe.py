"""Developer probe (GPU box): drift of the trainer parameters against the reference replay, per iteration.
usage: GSB_LIB=... python tools/gpu_drift_probe.py [iters]"""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import helpers as H
from oracle import refshim
from gpu_stage_probe import make_truths
gsb = H.gsb
DEV = "cuda"
KEYS = ("locations", "shs", "scales", "opacities", "rotations")
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
W, Hh, P = 512, 512, 10000
project = gsb.synth.project_for("C2")
project.sphere1.count = 4
project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.002, 0.01, 0.0005, 0.01, 0.0005
cams = gsb.Camera.getCameras(project)
cloud = gsb.synth.splat_cloud(P, 0x5EED0002)
tw, tb, consts = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0003), cams, W, Hh, DEV)
rt = refshim.RefTrainer(W, Hh, DEV); rt.set_model(cloud, 1000000, 1, 4); rt.set_truths(tw, tb, consts)
rt2 = refshim.RefTrainer(W, Hh, DEV); rt2.set_model(cloud, 1000000, 1, 4); rt2.set_truths(tw, tb, consts)
tr = gsb.Trainer(W, Hh, DEV)
tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
for it in range(iters):
    rt.train(project.params13(), as_shipped=True); rt2.train(project.params13(), as_shipped=True)
    tr.train(project, False)
    ref, ref2 = rt.get_model(), rt2.get_model()
    mine = gsb.ModelSplatsHost.from_device(tr.model)
    row = {}
    for k in KEYS:
        move = ref[k].astype(np.float64) - cloud[k]
        d = getattr(mine, k).astype(np.float64) - ref[k]
        d2 = ref2[k].astype(np.float64) - ref[k]
        row[k] = (round(np.linalg.norm(d) / np.linalg.norm(move), 6), round(np.linalg.norm(d2) / np.linalg.norm(move), 6))
    print(it + 1, json.dumps(row), flush=True)
d = (mine.locations.astype(np.float64) - ref["locations"]).reshape(P, 3)
n = np.linalg.norm(d, axis=1); top = np.argsort(-n)[:8]
print("top location offenders", [(int(i), float(n[i])) for i in top], "total", float(np.linalg.norm(n)))
for i in top[:4]:
    print(i, "scale", cloud["scales"].reshape(P, 3)[i], "op", cloud["opacities"][i], "moved", np.linalg.norm((ref["locations"].reshape(P, 3)[i] - cloud["locations"].reshape(P, 3)[i])))
