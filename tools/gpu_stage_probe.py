"""Developer probe (GPU box): per-stage times of the trainer path next to the reference replay.

usage: python tools/gpu_stage_probe.py [config] [cams_per_sphere] [fwd_ppt] [bwd_ppt] [--noref]
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

gsb = H.gsb


def make_truths(cloud_t, cams, W, Hh, dev, use_ref=False):
    """Teacher cloud rendered on white and black, quantised like imageFloatToInt."""
    from oracle import refshim
    sc = H.scene_tensors(cloud_t, dev)
    P = cloud_t["opacities"].shape[0]
    tw, tb, consts = [], [], []
    for cam in cams:
        c = cam.view_constants(W, Hh, square_fov=(W == Hh))
        consts.append(c)
        for bg, dst in (((1.0, 1.0, 1.0), tw), ((0.0, 0.0, 0.0), tb)):
            vw = H.view_tensors(c, bg, dev)
            if use_ref:
                out, _ = H.run_forward_ref(sc, vw, P, 1, 4, W, Hh)
            else:
                _, out, _ = H.run_forward_gsb(sc, vw, P, 1, 4, W, Hh)
            fb = torch.empty(W * Hh, dtype=torch.int32, device=dev)
            gsb.lib().gsb_image_float_to_int(out.data_ptr(), fb.data_ptr(), W, Hh, None)
            dst.append(fb)
    torch.cuda.synchronize()
    return tw, tb, consts


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    noref = "--noref" in sys.argv
    cfg = args[0] if len(args) > 0 else "C4"
    nper = int(args[1]) if len(args) > 1 else 4
    fwd_ppt = int(args[2]) if len(args) > 2 else 2
    bwd_ppt = int(args[3]) if len(args) > 3 else 2
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
    tr = gsb.Trainer(W, Hh, dev)
    tr.model = gsb.ModelSplatsDevice(host, dev)
    fuse = "--nofuse" not in sys.argv
    tr.set_fuse_backgrounds(fuse)
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    for _ in range(3):
        tr.train(project, False)
    torch.cuda.synchronize()
    K = 5
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(K):
        tr.train(project, False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    nv = 2 * len(cams)
    print(json.dumps({"impl": "gsb", "cfg": cfg, "cams": len(cams), "ms_per_iter": ms, "ms_per_view": ms / nv,
                      "views_per_s": nv / ms * 1000, "stats": tr.stats(), "ppt": [fwd_ppt, bwd_ppt]}), flush=True)
    tot = {}
    for i in range(nv // 2 if fuse else nv):
        st, cnt = tr.profile_pass(i)
        print(f"pass {i:2d} R={cnt['R']:8d} vis={cnt['visible']:7d} " + " ".join(f"{k}={v * 1000:6.0f}us" for k, v in st.items()),
              flush=True)
        for k, v in st.items():
            tot[k] = tot.get(k, 0.0) + v
    s = sum(tot.values())
    print(("mean/view (fused pairs): " if fuse else "mean/view: ") +
          " ".join(f"{k}={v / nv * 1000:.0f}us({100 * v / s:.0f}%)" for k, v in tot.items()) +
          f" total={s / nv * 1000:.0f}us", flush=True)

    if not noref:
        from oracle import refshim
        rt = refshim.RefTrainer(W, Hh, dev)
        rt.set_model(cloud, host.capacity, 1, 4)
        rt.set_truths(tw, tb, consts)
        p13 = project.params13()
        for mode, name in ((True, "ref_as_shipped"), (False, "ref_best_case")):
            for _ in range(2):
                rt.train(p13, as_shipped=mode)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record()
            for _ in range(K):
                rt.train(p13, as_shipped=mode)
            e1.record()
            torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) / K * 1000
            msr = e0.elapsed_time(e1) / K
            print(json.dumps({"impl": name, "ms_per_iter": msr, "wall_ms_per_iter": wall, "ms_per_view": msr / nv,
                              "views_per_s": nv / msr * 1000, "speedup_gsb": msr / ms}), flush=True)
        rt.close()


if __name__ == "__main__":
    main()
