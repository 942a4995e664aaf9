import sys, os, json, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/tools")
import helpers as H
from gpu_stage_probe import make_truths
gsb = H.gsb
W, Hh, P = gsb.synth.CONFIGS["C4"]
project = gsb.synth.project_for("C4"); project.sphere1.count = project.sphere2.count = 8
cams = gsb.Camera.getCameras(project)
cloud = gsb.synth.splat_cloud(P, 0x5EED0004)
tw, tb, _ = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0005), cams, W, Hh, "cuda")
for loss in (False, True, False, True):
    tr = gsb.Trainer(W, Hh, "cuda")
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")]), "cuda")
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    tr.enable_loss(loss)
    for _ in range(3): tr.train(project, False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(6): tr.train(project, False)
    e1.record(); torch.cuda.synchronize()
    print("loss", loss, "ms/iter", e0.elapsed_time(e1) / 6, flush=True)
    tr.close()
