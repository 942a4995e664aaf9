"""GPU: the CUDA truth ray caster (gaussian-splatterer_b200/csrc/app/rtx_trace.cu through gsb_rtx_*)
against its CPU restatement (oracle/oracle_rtx.cpp, exhaustive intersection, no BVH).

The bar is bit-exact RGBA8 frames and equal ray-segment counts: all arithmetic is IEEE float in a fixed order on both
sides, so the BVH, its traversal order and the tie-break are what these tests check.  Parity against the reference's own
OptiX programs is unpinned (see oracle/oracle_rtx.cpp's header); its observable contract -- alpha 0xff, background where
nothing reflected, sky shading, int(c * 256) packing, cleared frame without a model -- is checked as properties."""
import importlib

import numpy as np
import pytest
import torch

import helpers as H
from oracle import pyoracle

gsb = H.gsb
app = importlib.import_module("gaussian-splatterer_b200.app")
pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def scene():
    v, t, tc = gsb.synth.textured_mesh(rings=20, segments=40, seed=3)
    tex = gsb.synth.procedural_texture(64, 5)
    rtx = app.RtxHost()
    rtx.set_mesh(v, t, tc)
    rtx.set_texture(tex)
    yield rtx, v, t, tc, tex
    rtx.close()


def _oracle(scn, cam, W, Hh, bg, samples, rect=None, orbs=None):
    _, v, t, tc, tex = scn
    inv = app.camera_inverse_projview(cam, W / Hh)
    return pyoracle.rtx_render(W, Hh, np.array(cam.location[:], np.float32), inv, np.asarray(bg, np.float32), samples, v, t, tc, tex,
                               rect=rect, orbs=orbs)


@pytest.mark.parametrize("cam_args,size", [(((6.0, 3.0, 7.0), (0, 0, 0), 45.0), (96, 64)),
                                           (((0.5, 9.0, 0.5), (0, 0, 0), 60.0), (64, 64)),       # looking straight down
                                           (((-2.0, -1.0, 3.2), (0.3, 0, 0), 75.0), (80, 48))])  # close, grazing the floor
def test_frame_is_bit_exact_with_cpu_restatement(scene, cam_args, size):
    rtx = scene[0]
    W, Hh = size
    cam = app.camera_desc(*cam_args)
    fb = torch.zeros(W * Hh, dtype=torch.int32, device=DEV)
    for bg in ((1.0, 1.0, 1.0), (0.0, 0.0, 0.0), (0.2, 0.5, 0.9)):
        rtx.render(fb, (W, Hh), cam, bg, 6)
        want, rays = _oracle(scene, cam, W, Hh, bg, 6)
        got = fb.cpu().numpy().view(np.uint32).reshape(Hh, W)
        assert np.array_equal(got, want), (bg, int((got != want).sum()))
        assert rtx.last_ray_count() == rays
    assert rays > 6 * W * Hh                                      # bounces happened


def test_pair_render_equals_two_single_renders(scene):
    rtx = scene[0]
    W, Hh = 128, 72
    cam = app.camera_desc((5.0, 2.0, -6.0), (0, 0, 0), 50.0)
    w1, b1, w2, b2 = [torch.zeros(W * Hh, dtype=torch.int32, device=DEV) for _ in range(4)]
    rtx.render(w1, (W, Hh), cam, (1.0, 1.0, 1.0), 8)
    rtx.render(b1, (W, Hh), cam, (0.0, 0.0, 0.0), 8)
    rtx.render_pair(w2, b2, (W, Hh), cam, 8)
    assert torch.equal(w1, w2) and torch.equal(b1, b2)
    a = w1.cpu().numpy().view(np.uint8).reshape(-1, 4).astype(int)
    b = b1.cpu().numpy().view(np.uint8).reshape(-1, 4).astype(int)
    assert np.all(a[:, 3] == 255) and np.all(b[:, 3] == 255)      # make_rgba alpha, owl_device.h:89
    assert np.all(a[:, :3] >= b[:, :3])                           # backgrounds only ever add light
    assert (a[:, :3] != b[:, :3]).any() and (a[:, :3] == b[:, :3]).all(axis=1).any()     # both kinds of pixel exist


def test_camera_orbs_invert_pixels_like_reference(scene):
    rtx = scene[0]
    W, Hh = 96, 96
    cam = app.camera_desc((0.0, 1.0, 9.0), (0, 0, 0), 60.0)
    orbs = [app.camera_desc((0.0, 3.5, 0.0)), app.camera_desc((1.5, 0.0, 5.0)), app.camera_desc((0.0, 0.0, -6.0))]   # last one is hidden
    fb, plain = [torch.zeros(W * Hh, dtype=torch.int32, device=DEV) for _ in range(2)]
    rtx.render(fb, (W, Hh), cam, (0.0, 0.0, 0.0), 4, cameras=orbs)
    rtx.render(plain, (W, Hh), cam, (0.0, 0.0, 0.0), 4)
    want, _ = _oracle(scene, cam, W, Hh, (0, 0, 0), 4, orbs=np.array([list(o.location) for o in orbs], np.float32))
    got = fb.cpu().numpy().view(np.uint32).reshape(Hh, W)
    assert np.array_equal(got, want)
    changed = int((fb != plain).sum())
    assert 0 < changed < W * Hh // 20                             # a few small discs, nothing else


def test_no_model_clears_frame_and_reset(scene):
    r = app.RtxHost()
    assert not r.modelPresent
    fb = torch.full((64 * 64,), 7, dtype=torch.int32, device=DEV)
    r.render(fb, (64, 64), app.camera_desc((0, 0, 5)), (1, 1, 1), 2)
    assert int(fb.abs().sum()) == 0                               # cudaMemset(frameBuffer, 0, ...), RtxHost.cpp:220
    v, t, tc = gsb.synth.textured_mesh(rings=6, segments=8, floor=False)
    r.set_mesh(v, t, None)                                        # blank 8x8 grey texture, alpha 255 (RtxHost.cpp:23-36)
    r.render(fb, (64, 64), app.camera_desc((0, 0, 8)), (0, 0, 0), 3)
    px = fb.cpu().numpy().view(np.uint8).reshape(64, 64, 4)
    centre = px[28:36, 28:36, :3].astype(int)
    assert centre.max() <= 129 and centre.min() >= 0 and (px[..., 3] == 255).all()      # grey 0x80/255 attenuates every path
    assert px[0, 0, :3].tolist() == [0, 0, 0]                     # corner: nothing hit, black background
    r.reset()
    assert not r.modelPresent
    r.close()


def test_obj_and_png_loading_path_equals_in_memory_path(scene, tmp_path):
    rtx, v, t, tc, tex = scene
    gsb.synth.write_obj(tmp_path / "mesh.obj", v, t, tc)
    app.png_write(tmp_path / "tex.png", tex.reshape(-1).view(np.uint32), tex.shape[1], tex.shape[0], flip_vertically=False)
    r = app.RtxHost()
    r.loadModel(tmp_path / "missing.obj")                         # silently ignored, RtxHost.cpp:108
    assert not r.modelPresent
    r.loadModel(tmp_path / "mesh.obj")
    r.loadTextureDiffuse(tmp_path / "tex.png")
    assert r.modelPresent
    W, Hh = 72, 56
    cam = app.camera_desc((4.0, 4.0, 6.0), (0, 0, 0), 55.0)
    a, b = [torch.zeros(W * Hh, dtype=torch.int32, device=DEV) for _ in range(2)]
    r.render(a, (W, Hh), cam, (1, 1, 1), 5)
    rtx.render(b, (W, Hh), cam, (1, 1, 1), 5)
    assert torch.equal(a, b)
    r.close()


def test_truth_resolution_frame_properties_and_sampled_parity(scene):
    """The reference's truth resolution (1024 x 1024, src/Config.h:13-14) with rtSamples = 100 is too much for the CPU
    restatement; a pixel rectangle of it is compared exactly, the whole frame through properties."""
    rtx = scene[0]
    W = Hh = 1024
    cam = app.Project(new=True).cameras()[3]
    w, b = [torch.zeros(W * Hh, dtype=torch.int32, device=DEV) for _ in range(2)]
    rtx.render_pair(w, b, (W, Hh), cam, 100)
    rays = rtx.last_ray_count()
    assert rays >= 100 * W * Hh
    wn = w.cpu().numpy().view(np.uint32).reshape(Hh, W)
    rect = (500, 500, 524, 512)
    want, _ = _oracle(scene, cam, W, Hh, (1, 1, 1), 100, rect=rect)
    assert np.array_equal(wn[500:512, 500:524], want[500:512, 500:524])
    px = wn.view(np.uint8).reshape(Hh, W, 4)
    assert (px[..., 3] == 255).all()
    assert (px[0, 0, :3] == 255).all()                            # frame corner sees only background on the white frame
    bn = b.cpu().numpy().view(np.uint8).reshape(Hh, W, 4)
    assert (bn[0, 0, :3] == 0).all() and (px[..., :3].astype(int) >= bn[..., :3].astype(int)).all()
