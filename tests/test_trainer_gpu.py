"""GPU: the Trainer path (gsb_trainer_* through the Python mirror of Trainer / ModelSplatsDevice) against
golden vectors of the reference's Trainer::train replay (tests/golden/trainer_*.npz), the CPU oracle, and
oracle/_ref where present.  Bars: running means within 1e-3 relative L2 per tensor, parameters after
applyGradients within 2e-5 absolute, densify result identical as a multiset of splats."""
import numpy as np
import pytest
import torch

import golden_cases as GC
import helpers as H
from oracle import pyoracle, refshim

gsb = GC.gsb
pytestmark = pytest.mark.gpu
DEV = "cuda"
NAMES = ("var", "loc", "sh", "scale", "opacity", "rot")
KEYS = ("locations", "shs", "scales", "opacities", "rotations")


def _split(avg, P, M=4):
    sizes, out, o = [P, 3 * P, 3 * M * P, 3 * P, P, 4 * P], [], 0
    for s in sizes:
        out.append(avg[o:o + s]); o += s
    return out


def _trainer(c, cloud, cams, g, fuse=True):
    host = gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS])
    tr = gsb.Trainer(c["W"], c["H"], DEV)
    tr.set_fuse_backgrounds(fuse)
    tr.model = gsb.ModelSplatsDevice(host, DEV)
    tr.truthFrameBuffersW = [torch.from_numpy(np.ascontiguousarray(x)).to(DEV) for x in g["truthW"]]
    tr.truthFrameBuffersB = [torch.from_numpy(np.ascontiguousarray(x)).to(DEV) for x in g["truthB"]]
    tr.truthCameras = list(cams)
    return tr


def _rows(m, M=4):
    n = m["opacities"].shape[0]
    r = np.concatenate([m["locations"].reshape(n, 3), m["scales"].reshape(n, 3), m["opacities"].reshape(n, 1),
                        m["rotations"].reshape(n, 4), m["shs"].reshape(n, 3 * M)], axis=1)
    return r[np.lexsort(np.round(r[:, ::-1], 4).T)]


@pytest.mark.parametrize("name", list(GC.TRAINER_CASES))
@pytest.mark.parametrize("fuse", [False, True])
def test_running_means_match_reference_golden(name, fuse):
    g = GC.load("trainer_" + name)
    c, cloud, _t, project, cams = GC.trainer_inputs(name)
    tr = _trainer(c, cloud, cams, g, fuse)
    avg = tr.grad_tensor()
    tr.accumulate()
    torch.cuda.synchronize()
    for k, mine in zip(NAMES, _split(avg.cpu(), c["P"])):
        rel = H.rel_l2(mine, torch.from_numpy(g["avg_" + k]))
        assert rel <= 1e-3, (k, rel)
    s = tr.stats()
    assert s["views"] == 2 * len(cams) and s["retries"] == 0 and s["kernel_launches"] > 0
    tr.close()


@pytest.mark.parametrize("name", list(GC.TRAINER_CASES))
def test_train_apply_and_densify_match_reference_golden(name):
    g = GC.load("trainer_" + name)
    c, cloud, _t, project, cams = GC.trainer_inputs(name)
    tr = _trainer(c, cloud, cams, g)
    it0 = project.iterations
    tr.train(project, False)
    assert project.iterations == it0 + 1                      # src/Trainer.cu:255
    after = gsb.ModelSplatsHost.from_device(tr.model)
    for k in KEYS:
        np.testing.assert_allclose(getattr(after, k), g["applied_" + k], atol=2e-5)
    # densify with the golden thresholds (they sit in gaps of the decision variables)
    for n, v in zip(gsb.Project.PARAM_ORDER, g["densify_params13"]):
        setattr(project, n, float(v))
    tr.densify(project)
    dens = gsb.ModelSplatsHost.from_device(tr.model)
    assert dens.count == int(g["densified_count"]) != c["P"]
    assert tr.model.devLocations.numel() == 3 * dens.count     # arrays sized to count (src/ModelSplatsDevice.cpp:11-15)
    mine = _rows({k: getattr(dens, k) for k in KEYS})
    ref = _rows({k: g["densified_" + k] for k in KEYS})
    np.testing.assert_allclose(mine, ref, atol=5e-5)
    # the trainer keeps working on the replaced model (count changed -> buffers re-sized, T:259-301)
    tr.train(project, False)
    assert gsb.ModelSplatsHost.from_device(tr.model).count == dens.count
    tr.close()


def test_densify_matches_oracle_canonical_order_and_capacity():
    g = GC.load("trainer_t1")
    c, cloud, _t, project, cams = GC.trainer_inputs("t1")
    tr = _trainer(c, cloud, cams, g)
    avg = tr.grad_tensor()
    tr.train(project, False)
    after = gsb.ModelSplatsHost.from_device(tr.model)
    a = _split(avg.cpu().numpy(), c["P"])
    applied = {k: getattr(after, k).copy() for k in KEYS}
    GC.densify_thresholds(project, applied, a[0], a[1])
    for cap in (1000000, c["P"] + 37):                    # second: capacity binds (src/Trainer.cu:460,500)
        tr2 = _trainer(c, applied, cams, g)
        tr2.model.capacity = cap
        tr2.grad_tensor().copy_(avg)
        tr2.densify(project)
        mine = gsb.ModelSplatsHost.from_device(tr2.model)
        ref = pyoracle.densify(applied, c["P"], cap, 4, a[0], a[1], project.params13())
        assert mine.count == ref["opacities"].shape[0]
        for k in KEYS:                                     # same canonical order, not just the same multiset
            np.testing.assert_allclose(getattr(mine, k), ref[k], atol=1e-5, err_msg=k)
        tr2.close()
    tr.close()


def test_train_without_truths_raises_like_reference():
    tr = gsb.Trainer(64, 64, DEV)
    cloud = gsb.synth.splat_cloud(10, 1)
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
    with pytest.raises(RuntimeError, match="no truth data"):      # src/Trainer.cu:253
        tr.train(gsb.Project(), False)
    with pytest.raises(gsb.GsbError, match="no truth data"):
        tr.accumulate()
    tr.close()


def test_render_preview_matches_oracle_and_reference_quantisation():
    W, Hh, P = 150, 90, 2000
    cloud = gsb.synth.splat_cloud(P, 31)
    cam = gsb.Camera((4.0, 3.0, 8.0), (0, 0, 0), 55.0)
    tr = gsb.Trainer(96, 96, DEV)                     # preview size differs from the training size (T:159-167)
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
    fb = torch.zeros(W * Hh, dtype=torch.int32, device=DEV)
    tr.render(fb, W, Hh, 1.7, cam)
    cs = cam.view_constants(W, Hh, square_fov=False)
    out, _ = pyoracle.forward(P, 1, 4, (0, 0, 0), W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"], cloud["scales"],
                              1.7, cloud["rotations"], None, cs["view"], cs["projview"], cs["location"], cs["tan_fovx"], cs["tan_fovy"])
    ref = pyoracle.image_float_to_int(out, W, Hh)
    mine = fb.cpu().numpy().view(np.uint32)
    assert np.all(mine >> 24 == 0xFF)
    diff = np.abs(((mine[:, None] >> np.array([0, 8, 16])) & 0xFF).astype(int) - ((ref[:, None] >> np.array([0, 8, 16])) & 0xFF).astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() <= 1e-3     # int(c * 256) may flip on a 1-ulp colour difference
    tr.close()


def test_binning_arena_overflow_retries_and_stays_correct():
    # 1500 splats that each cover most of a 640x640 image: R is several million, far above the initial arena
    W, Hh, P = 640, 640, 1500
    cloud = gsb.synth.splat_cloud(P, 8, scale_hi=3.0, scale_lo=1.0)
    cloud["opacities"][:] = 0.02
    project = gsb.Project()
    project.sphere1 = gsb.CameraSphere(count=2)
    project.sphere2 = gsb.CameraSphere(count=0)
    cams = gsb.Camera.getCameras(project)
    truth = [torch.full((W * Hh,), 0x7F7F7F7F, dtype=torch.int32, device=DEV) for _ in cams]
    runs = []
    for _ in range(2):
        tr = gsb.Trainer(W, Hh, DEV)
        tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = truth, truth, cams
        avg = tr.grad_tensor()
        tr.accumulate()
        first = tr.stats()
        a1 = avg.clone()
        tr.accumulate()                                  # arena already large enough now
        assert tr.stats()["retries"] == 0
        assert H.rel_l2(avg, a1) <= 1e-4
        runs.append((first, a1))
        tr.close()
    assert runs[0][0]["retries"] >= 1 and runs[0][0]["rendered_total"] > 2 * (1 << 20)
    assert H.rel_l2(runs[0][1], runs[1][1]) <= 1e-4


@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref/libgsref.so not present")
def test_iterations_track_reference_replay_over_several_steps():
    """Five full iterations (forward + loss + backward + accumulate + apply) of both implementations from
    the same start: parameters stay within float noise of each other (config C2 shape, fewer cameras)."""
    import sys, os
    sys.path.insert(0, os.path.join(GC.ROOT, "tools"))
    from gpu_stage_probe import make_truths
    W, Hh, P = 512, 512, 10000
    project = gsb.synth.project_for("C2")
    project.sphere1.count = 4
    project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.002, 0.01, 0.0005, 0.01, 0.0005
    cams = gsb.Camera.getCameras(project)
    cloud = gsb.synth.splat_cloud(P, 0x5EED0002)
    tw, tb, consts = make_truths(gsb.synth.splat_cloud(P // 2, 0x5EED0003), cams, W, Hh, DEV)
    rt = refshim.RefTrainer(W, Hh, DEV)
    rt.set_model(cloud, 1000000, 1, 4)
    rt.set_truths(tw, tb, consts)
    # a second replay of the reference from the same start measures its own run-to-run spread: its atomics are
    # unordered, and once in a while a rounding difference flips a blend / clamp decision of one splat, after which
    # two runs of the REFERENCE are further apart (1.4e-3 observed at step 5) than the budget below
    rt2 = refshim.RefTrainer(W, Hh, DEV)
    rt2.set_model(cloud, 1000000, 1, 4)
    rt2.set_truths(tw, tb, consts)
    tr = gsb.Trainer(W, Hh, DEV)
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
    tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
    for _ in range(5):
        rt.train(project.params13(), as_shipped=True)
        rt2.train(project.params13(), as_shipped=True)
        tr.train(project, False)
    ref, ref2 = rt.get_model(), rt2.get_model()
    mine = gsb.ModelSplatsHost.from_device(tr.model)
    for k in KEYS:
        # measured against how far the parameters moved: 1e-3 relative L2 (the gradient budget of north_star;
        # two runs of the reference itself usually differ by ~2e-5 here), no single parameter off by more than 1 %;
        # where the reference's own two runs are further apart than that, twice their spread
        move = ref[k].astype(np.float64) - cloud[k]
        diff = getattr(mine, k).astype(np.float64) - ref[k]
        spread = ref2[k].astype(np.float64) - ref[k]
        assert np.abs(move).max() > 0
        assert np.linalg.norm(diff) <= max(1e-3 * np.linalg.norm(move), 2 * np.linalg.norm(spread)), k
        assert np.abs(diff).max() <= max(1e-2 * np.abs(move).max(), 5e-6, 2 * np.abs(spread).max()), k
    rt.close(); rt2.close(); tr.close()


def test_overlapped_pipeline_equals_serial_execution():
    """The two-stream pipeline (forward chain of pass i+1 under the backward chain of pass i, alternating
    workspace slots) must give the running means of the serial order; also exercised with an odd pass count."""
    import sys, os
    sys.path.insert(0, os.path.join(GC.ROOT, "tools"))
    from gpu_stage_probe import make_truths
    W, Hh, P = 320, 200, 6000
    project = gsb.synth.project_for("C2")
    project.sphere1.count = 5
    cams = gsb.Camera.getCameras(project)
    cloud = gsb.synth.splat_cloud(P, 31, scale_hi=0.12)
    tw, tb, _ = make_truths(gsb.synth.splat_cloud(P // 2, 32, scale_hi=0.12), cams, W, Hh, DEV)
    res = {}
    for fuse in (True, False):
        for overlap in (False, True):
            tr = gsb.Trainer(W, Hh, DEV, square_fov=False)
            tr.set_fuse_backgrounds(fuse)
            tr.set_overlap(overlap)
            tr.enable_loss(True)
            tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS]), DEV)
            tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, cams
            avg = tr.grad_tensor()
            for _ in range(3):                      # repeated: slots and events are reused across iterations
                tr.accumulate()
            res[(fuse, overlap)] = (avg.clone(), tr.stats())
            tr.close()
    for fuse in (True, False):
        a, sa = res[(fuse, False)]
        b, sb = res[(fuse, True)]
        assert sa["rendered_total"] == sb["rendered_total"]
        assert abs(sa["sum_sq_residual"] - sb["sum_sq_residual"]) <= 1e-9 * sa["sum_sq_residual"]
        assert H.rel_l2(b, a) <= 1e-5
    # the fused loss (one double atomic per warp, spread over GSB_LOSS_SLOTS addresses and added up by the host) against an
    # independent value: imageIntToLoss's residuals (src/Trainer.cu:38-42) of every (camera, background) pass, squared
    # and summed in double by torch from Rasterizer::forward's image
    sc = H.scene_tensors(cloud, DEV)
    want = 0.0
    for cam, fw, fb in zip(cams, tw, tb):
        c = cam.view_constants(W, Hh, square_fov=False)
        for bg, truth in (((1.0, 1.0, 1.0), fw), ((0.0, 0.0, 0.0), fb)):
            _, out, _ = H.run_forward_gsb(sc, H.view_tensors(c, bg, DEV), P, 1, 4, W, Hh)
            t = truth.view(torch.int32)
            chan = torch.stack([(t >> s) & 0xFF for s in (0, 8, 16)]).to(torch.float32) / 255.0
            want += float(((chan.reshape(3, -1) - out.reshape(3, -1)).double() ** 2).sum())
    for key, (_, st) in res.items():
        assert abs(st["sum_sq_residual"] - want) <= 1e-5 * want, (key, st["sum_sq_residual"], want)
