"""GPU: the headless UiFrame (gsb_app_*, SURVEY.md 8 f2 + f4): scheduler semantics of UiFrame::update
(src/ui/UiFrame.cpp:266-298), Trainer::captureTruths through the CUDA ray caster (src/Trainer.cu:218-250), training through
the same C ABI the compat Trainer uses, .gobj / settings.json round trips from device state, and the screenshot paths
(src/ui/tools/UiPanelToolsView.cpp:227-259)."""
import ctypes
import importlib

import numpy as np
import pytest
import torch

import helpers as H
from test_app_host import _png_decode

gsb = H.gsb
app = importlib.import_module("gaussian-splatterer_b200.app")
pytestmark = pytest.mark.gpu
DEV = "cuda"
libc = ctypes.CDLL("libc.so.6")


def _mesh_app(w=128, h=128, cameras=4, samples=3):
    a = app.App(w, h)
    v, t, tc = gsb.synth.textured_mesh(rings=12, segments=24, seed=9)
    a.rtx.set_mesh(v, t, tc)
    a.rtx.set_texture(gsb.synth.procedural_texture(32, 2))
    a.project.d.sphere1.count = cameras
    a.project.rtSamples = samples
    return a


def test_new_app_is_a_new_project_with_the_grid_field():
    a = app.App(64, 64)
    assert a.project.dumps() == app.Project(new=True).dumps()
    assert a.model_count == 17 ** 3 and a.truth_count == 0 and not a.rtx.modelPresent
    md = a.model_device()
    assert (md["capacity"], md["shDegree"], md["shCoeffs"]) == (1000000, 1, 4)
    with pytest.raises(gsb.GsbError, match="Can't run training iteration, no truth data available!"):      # src/Trainer.cu:253
        a.train()
    assert a.project.iterations == 0
    a.initFieldMono()
    assert a.model_count == 1
    a.close()


def test_capture_truths_renders_both_frames_of_every_camera():
    a = _mesh_app()
    a.captureTruths()
    assert a.truth_count == 4
    cams = a.project.cameras()
    for i in (0, 3):
        w, b, cam = a.truth_frames(i)
        assert list(cam.location) == list(cams[i].location) and cam.fovDegY == cams[i].fovDegY
        w2, b2 = torch.zeros_like(w), torch.zeros_like(b)
        a.rtx.render(w2, (128, 128), cams[i], (1.0, 1.0, 1.0), 3)                                   # src/Trainer.cu:230-236
        a.rtx.render(b2, (128, 128), cams[i], (0.0, 0.0, 0.0), 3)
        assert torch.equal(w, w2) and torch.equal(b, b2)
    a.project.d.sphere1.count = 6                   # camera count changed: frames are re-allocated (src/Trainer.cu:219-227)
    a.captureTruths()
    assert a.truth_count == 6
    a.close()


def test_update_scheduler_follows_uiframe_update():
    a = _mesh_app(96, 96, cameras=3, samples=2)
    p = a.project
    p.intervalCapture, p.intervalDensify = 3, 4
    assert a.update(1.0) == 0 and p.previewTimer == 1.0 and p.iterations == 0         # not auto-training: only the timer moves
    a.set_auto_training(True)
    assert a.update(0.004) == 0                     # budget 0.4 < 1 (AUTO_TRAIN_BUDGET 100 / s, src/Config.h:10)
    assert a.update(0.004) == 0                     # 0.8
    libc.srand(12345)
    want_rot = [np.float32(np.float32(libc.rand()) / np.float32(2147483647)) * np.float32(360.0) for _ in range(4)]
    libc.srand(12345)
    flags = [a.update(0.004)]                       # 1.2 -> capped at 1 -> iteration 0: capture + densify
    assert flags[0] == a.RAN_TRAINED | a.RAN_CAPTURED | a.RAN_DENSIFIED and p.iterations == 1
    got_rot = [p.d.sphere1.rotX, p.d.sphere1.rotY, p.d.sphere2.rotX, p.d.sphere2.rotY]
    assert got_rot == [float(x) for x in want_rot]                                   # onButtonRandomRotate: four rand() in this order
    assert a.update(0.009) == 0                     # the budget restarted from 0: 0.9
    for _ in range(8):
        flags.append(a.update(0.5))                 # a big delta buys ONE iteration, never more (min(1, ...), UF:279)
    want = []
    for it in range(9):                             # the tests use the pre-increment iteration counter (UF:283-284)
        want.append(1 | (2 if it % 3 == 0 else 0) | (4 if it % 4 == 0 else 0))
    assert flags == want and p.iterations == 9
    p.intervalCapture = p.intervalDensify = 0       # "never"
    assert a.update(1.0) == a.RAN_TRAINED
    a.set_auto_training(False)
    assert a.update(1.0) == 0 and p.iterations == 10
    a.close()


def test_training_through_the_app_equals_the_trainer_api():
    W = Hh = 128
    a = _mesh_app(W, Hh, cameras=3, samples=4)
    cloud = gsb.synth.splat_cloud(3000, 0xA44, scale_lo=0.03, scale_hi=0.25)
    host = app.ModelSplatsHostC.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")])
    a.set_model(host)
    a.captureTruths()
    frames = [a.truth_frames(i) for i in range(3)]
    # the same three iterations through the Trainer mirror (what the compat Trainer class calls)
    proj = gsb.synth.project_for("C2")
    proj.sphere1.count = 3
    tr = gsb.Trainer(W, Hh, DEV, square_fov=True)
    tr.model = gsb.ModelSplatsDevice(gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")]), DEV)
    tr.truthFrameBuffersW = [f[0] for f in frames]
    tr.truthFrameBuffersB = [f[1] for f in frames]
    tr.truthCameras = gsb.Camera.getCameras(proj)
    for c, f in zip(tr.truthCameras, frames):
        assert np.array_equal(c.location, np.array(f[2].location[:], np.float32)) and c.fovDegY == f[2].fovDegY
    for _ in range(3):
        a.train(False)
        tr.train(proj, False)
    got = a.get_model().arrays()
    for k, t in (("locations", tr.model.devLocations), ("shs", tr.model.devShs), ("scales", tr.model.devScales),
                 ("opacities", tr.model.devOpacities), ("rotations", tr.model.devRotations)):
        ref = t.cpu().numpy().reshape(-1)
        moved = np.linalg.norm(ref - np.asarray(cloud[k], np.float32).reshape(-1))
        assert moved > 0 and np.linalg.norm(got[k] - ref) <= 1e-3 * moved, k      # summation order only
    assert a.project.iterations == 3
    a.project.paramDensifyVariance = -1.0           # force split / clone
    a.train(True)
    assert a.model_count > 3000
    tr.close()
    a.close()


def test_splats_and_settings_round_trip_through_files(tmp_path):
    a = _mesh_app(64, 64)
    cloud = gsb.synth.splat_cloud(500, 0xF11E)
    a.set_model(app.ModelSplatsHostC.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")]))
    a.project.iterations = 77
    a.project.lrSh = 0.01
    a.saveSplats(tmp_path / "m.gobj")
    a.saveSettings(tmp_path / "settings.json")
    b = app.App(64, 64)
    b.loadSplats(tmp_path / "m.gobj")
    b.loadSettings(tmp_path / "settings.json")
    assert b.model_count == 500 and b.project.dumps() == a.project.dumps() and b.project.iterations == 77
    want = app.ModelSplatsHostC.load_gobj(tmp_path / "m.gobj").arrays()
    got = b.get_model().arrays()
    assert all(np.array_equal(got[k], want[k]) for k in want)
    with pytest.raises(gsb.GsbError, match="Failed to load splats file"):
        b.loadSplats(tmp_path / "none.gobj")
    assert b.model_count == 500                     # a failed load leaves the model in place
    a.close(); b.close()


def test_screenshots_are_flipped_pngs_of_the_preview_camera(tmp_path):
    a = _mesh_app(64, 64)
    cloud = gsb.synth.splat_cloud(4000, 0x5C4EE, scale_lo=0.03, scale_hi=0.2)
    a.set_model(app.ModelSplatsHostC.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")]))
    p = a.project
    p.renderResX, p.renderResY, p.previewSplatScale, p.previewTimer, p.rtSamples = 200, 120, 0.8, 3.5, 4
    a.render_splats_png(tmp_path / "splats.png")
    cam = p.preview_camera()
    fb = torch.zeros(200 * 120, dtype=torch.int32, device=DEV)
    a.render_splats(fb, 200, 120, 0.8, cam)
    torch.cuda.synchronize()
    want = fb.cpu().numpy().view(np.uint8).reshape(120, 200, 4)
    dec, _ = _png_decode(tmp_path / "splats.png")
    assert np.array_equal(dec, want[::-1]) and want[..., :3].any()
    # the same frame through the rasterizer-level API with Trainer::render's camera convention (src/Trainer.cu:148-216)
    cc = app.camera_constants(cam, 200, 120, square_fov=False)
    consts = {"view": np.array(cc.view[:], np.float32), "projview": np.array(cc.projview[:], np.float32),
              "location": np.array(cc.location[:], np.float32), "tan_fovx": cc.tan_fovx, "tan_fovy": cc.tan_fovy}
    sc = H.scene_tensors(cloud, DEV)
    vw = H.view_tensors(consts, (0.0, 0.0, 0.0), DEV)
    ws, out, R = H.run_forward_gsb(sc, vw, 4000, 1, 4, 200, 120, scale_modifier=0.8)
    q = (out.reshape(3, 120, 200).clamp(0, 1) * 256.0).to(torch.int32).clamp(0, 255).permute(1, 2, 0).cpu().numpy()
    assert np.array_equal(want[..., :3], q.astype(np.uint8)) and (want[..., 3] == 255).all()
    a.render_rtx_png(tmp_path / "rtx.png")
    a.rtx.render(fb, (200, 120), cam, (0.0, 0.0, 0.0), 4)
    dec, _ = _png_decode(tmp_path / "rtx.png")
    assert np.array_equal(dec, fb.cpu().numpy().view(np.uint8).reshape(120, 200, 4)[::-1])
    a.close()
