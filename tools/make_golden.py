"""Generate tests/golden/*.npz by running the UNMODIFIED reference CUDA build (oracle/_ref/libgsref.so,
compiled from /root/reference by oracle/Makefile) on a B200.

    gpurun -- python tools/make_golden.py          # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/       # commit them

Inputs are not stored: tests/golden_cases.py rebuilds them from seeds.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_cases as GC  # noqa: E402
import helpers as H  # noqa: E402
from oracle import refshim  # noqa: E402

gsb = GC.gsb
OUT = os.path.join(ROOT, "gpurun_out", "golden")


def raster_case(name):
    c, cloud, consts, dL = GC.raster_inputs(name)
    dev = "cuda"
    sc = H.scene_tensors(cloud, dev)
    vw = H.view_tensors(consts, c["bg"], dev)
    P, W, Hh = c["P"], c["W"], c["H"]
    out, R = H.run_forward_ref(sc, vw, P, 1, 4, W, Hh)
    st = refshim.state(P, W, Hh, R, dev)
    dLt = torch.from_numpy(dL).to(dev)
    # the reference's atomics are order-nondeterministic: store the mean of several runs
    acc = None
    runs = 8
    for _ in range(runs):
        g = H.run_backward_ref(sc, vw, P, 1, 4, R, W, Hh, dLt)
        acc = {k: v.double() for k, v in g.items()} if acc is None else {k: acc[k] + g[k].double() for k in g}
    grads = {k: (v / runs).float().cpu().numpy() for k, v in acc.items()}
    vis = (st["radii"] > 0).cpu().numpy()
    data = {"R": np.int64(R), "out_color": out.cpu().numpy(), "visible": vis}
    for k in ("radii", "tiles_touched", "point_offsets", "point_list_keys", "point_list", "ranges", "n_contrib", "final_T",
              "depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped"):
        a = st[k].cpu().numpy()
        if k in ("depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped"):
            a = np.where(vis.reshape((-1,) + (1,) * (a.ndim - 1)), a, 0)   # culled slots are uninitialised memory
        data[k] = a
    present = torch.zeros(P, dtype=torch.bool, device=dev)
    refshim.mark_visible(P, sc["locations"], vw["view"], vw["projview"], present)
    data["present"] = present.cpu().numpy()
    for k, v in grads.items():
        data[k] = v
    np.savez_compressed(os.path.join(OUT, f"raster_{name}.npz"), **data)
    print(name, "R", R, "visible", int(vis.sum()))


def trainer_case(name):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from gpu_stage_probe import make_truths
    c, cloud, teacher, project, cams = GC.trainer_inputs(name)
    dev = "cuda"
    W, Hh, P = c["W"], c["H"], c["P"]
    tw, tb, consts = make_truths(teacher, cams, W, Hh, dev, use_ref=True)
    rt = refshim.RefTrainer(W, Hh, dev)
    rt.set_model(cloud, 1000000, 1, 4)
    rt.set_truths(tw, tb, consts)
    acc = None
    runs = 8
    for _ in range(runs):
        rt.train(project.params13(), as_shipped=False, apply=False)
        b = [x.double() for x in rt.buffers()]
        acc = b if acc is None else [a + x for a, x in zip(acc, b)]
    avg = [(a / runs).float().cpu().numpy() for a in acc]
    data = {f"avg_{n}": a for n, a in zip(("var", "loc", "sh", "scale", "opacity", "rot"), avg)}
    data["truthW"] = torch.stack(tw).cpu().numpy()
    data["truthB"] = torch.stack(tb).cpu().numpy()
    rt.train(project.params13(), as_shipped=True, apply=True)
    after = rt.get_model()
    for k, v in after.items():
        data["applied_" + k] = v
    var_last, loc_last = [x.cpu().numpy() for x in rt.buffers()[:2]]
    GC.densify_thresholds(project, after, var_last, loc_last)
    data["densify_params13"] = project.params13()
    data["densify_var"] = var_last
    data["densify_avg_loc"] = loc_last
    rt.densify(project.params13())
    dens = rt.get_model()
    for k, v in dens.items():
        data["densified_" + k] = v
    data["densified_count"] = np.int64(rt.count())
    np.savez_compressed(os.path.join(OUT, f"trainer_{name}.npz"), **data)
    print(name, "count after densify", rt.count(), "of", P)
    rt.close()


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for n in GC.RASTER_CASES:
        raster_case(n)
    for n in GC.TRAINER_CASES:
        trainer_case(n)
