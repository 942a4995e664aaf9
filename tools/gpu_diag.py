"""Developer probe: noise floor of the reference vs this library on the cases the tolerance tests use."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import helpers as H
from oracle import refshim, pyoracle
from gpu_stage_probe import make_truths
gsb = H.gsb
DEV = "cuda"
KEYS = ("locations", "shs", "scales", "opacities", "rotations")

def degenerate():
    W, Hh, P = 192, 128, 1200
    cloud = gsb.synth.splat_cloud(P, 99)
    sc, rot, op, loc = (cloud[k].reshape(P, -1) for k in ("scales", "rotations", "opacities", "locations"))
    sc[0:50, 0] = 0.0; sc[50:80] = 0.0; rot[80:120] = 0.0; rot[120:160] *= 3.0; op[160:200] = 1.0
    op[200:240] = 1.0 / 255.0; op[240:260] = 0.0; sc[260:262] = 40.0; loc[262:300, 2] = 9.85
    loc[300:320] = (0.0, 0.0, 11.0); sc[320:360] = (0.3, 0.004, 0.004)
    for cam, bg in ((gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0), (1.0, 1.0, 1.0)), (gsb.Camera((0.0, 0.0, 4.5), (0, 0, 0), 25.0), (0.0, 0.0, 0.0))):
        consts = cam.view_constants(W, Hh)
        s = H.scene_tensors(cloud, DEV); vw = H.view_tensors(consts, bg, DEV)
        dL = torch.from_numpy((gsb.synth.uniform(3 * W * Hh, 9, 3) - 0.5).astype(np.float32)).to(DEV)
        ref_out, R_ref = H.run_forward_ref(s, vw, P, 1, 4, W, Hh)
        g_ref = H.run_backward_ref(s, vw, P, 1, 4, R_ref, W, Hh, dL)
        g_ref2 = H.run_backward_ref(s, vw, P, 1, 4, R_ref, W, Hh, dL)
        ws, out, R = H.run_forward_gsb(s, vw, P, 1, 4, W, Hh)
        g = H.run_backward_gsb(ws, s, vw, P, 1, 4, R, W, Hh, dL)
        o_out, v = pyoracle.forward(P, 1, 4, bg, W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"], cloud["scales"], 1.0,
                                    cloud["rotations"], None, consts["view"], consts["projview"], consts["location"], consts["tan_fovx"], consts["tan_fovy"])
        og = v.backward(dL.cpu().numpy(), 4)
        res = {}
        for k in H.GRAD_NAMES:
            o = torch.from_numpy(og[k]).to(DEV)
            res[k] = {"mine_vs_ref": H.rel_l2(g[k], g_ref[k]), "ref_vs_ref": H.rel_l2(g_ref2[k], g_ref[k]),
                      "mine_vs_oracle": H.rel_l2(g[k], o), "ref_vs_oracle": H.rel_l2(g_ref[k], o)}
        print("degenerate", json.dumps(res), flush=True)
        # which splats carry the dL_drot difference
        d = (g["dL_drot"] - g_ref["dL_drot"]).view(P, 4).double().norm(dim=1)
        top = torch.topk(d, 5)
        print(" top drot diff idx", top.indices.tolist(), [float(x) for x in top.values], "ref norms", [float(g_ref["dL_drot"].view(P, 4)[i].double().norm()) for i in top.indices],
              "total", float(g_ref["dL_drot"].double().norm()), flush=True)

def iterations():
    W, Hh, P = 512, 512, 10000
    project = gsb.synth.project_for("C2"); project.sphere1.count = 4
    project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.002, 0.01, 0.0005, 0.01, 0.0005
    cams = gsb.Camera.getCameras(project)
    cloud = gsb.synth.splat_cloud(P, 0x5EED0002)
    tw, tb, consts = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0003), cams, W, Hh, DEV)
    models = []
    for rep in range(2):
        rt = refshim.RefTrainer(W, Hh, DEV); rt.set_model(cloud, 1000000, 1, 4); rt.set_truths(tw, tb, consts)
        for _ in range(5): rt.train(project.params13(), as_shipped=True)
        models.append(rt.get_model()); rt.close()
    for fuse in (True, False):
        tr = gsb.Trainer(W, Hh, DEV); tr.set_fuse_backgrounds(fuse)
        tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
        for _ in range(5): tr.train(project, False)
        mine = gsb.ModelSplatsHost.from_device(tr.model); tr.close()
        res = {}
        for k in KEYS:
            ref, ref2, m = models[0][k].astype(np.float64), models[1][k].astype(np.float64), getattr(mine, k).astype(np.float64)
            mv = ref - cloud[k]
            res[k] = {"moved_max": float(np.abs(mv).max()), "mine_max": float(np.abs(m - ref).max()), "refref_max": float(np.abs(ref2 - ref).max()),
                      "mine_relL2_of_move": float(np.linalg.norm(m - ref) / np.linalg.norm(mv)), "refref_relL2_of_move": float(np.linalg.norm(ref2 - ref) / np.linalg.norm(mv))}
        print("iterations fuse=%s" % fuse, json.dumps(res), flush=True)

degenerate()
iterations()
