"""Job run under compute-sanitizer (tools/gpu_round2.sh): __graft_entry__.smoke() (Rasterizer::forward / backward and a
Trainer iteration with densify at 250x130) followed by one fused and one unfused 2-camera trainer step at config C1's
shape (256x256, 10 k splats) and a Trainer::render -- every kernel of the hot path runs at least once."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402

import __graft_entry__ as G  # noqa: E402

G.smoke()
import helpers as H  # noqa: E402
from gpu_stage_probe import make_truths  # noqa: E402

gsb = H.gsb
W, Hh, P = gsb.synth.CONFIGS["C1"]
project = gsb.synth.project_for("C1")
project.sphere1.count = 2
cams = gsb.Camera.getCameras(project)
cloud = gsb.synth.splat_cloud(P, 0x5EED0001)
tw, tb, _ = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0002), cams, W, Hh, "cuda")
KEYS = ("locations", "shs", "scales", "opacities", "rotations")
for fuse in (True, False):
    tr = gsb.Trainer(W, Hh, "cuda")
    tr.set_fuse_backgrounds(fuse)
    tr.enable_loss(True)
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), "cuda")
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    tr.train(project, False)
    project.paramDensifyVariance = -1.0
    tr.train(project, True)
    fb = torch.zeros(W * Hh, dtype=torch.int32, device="cuda")
    tr.render(fb, W, Hh, 1.0, cams[0])
    torch.cuda.synchronize()
    print("sanitize_job:", "fused" if fuse else "unfused", "count", tr.model.count, "loss", tr.stats()["sum_sq_residual"])
    tr.close()
print("sanitize_job done")
