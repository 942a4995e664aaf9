"""GPU: the ends of the size range the application allows -- preview resolutions from 16 to 8192
(reference src/ui/tools/UiPanelToolsView.cpp:120), a tile grid that needs a third pass of the tile sort, and the
3 M-splat cloud of BASELINE config C5 -- each against the compiled reference on identical device buffers."""
import numpy as np
import pytest
import torch

import golden_cases as GC
import helpers as H
from oracle import refshim

gsb = GC.gsb
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refshim.available(), reason="oracle/_ref/libgsref.so not present")]
DEV = "cuda"


def _compare_forward(cloud, P, W, Hh, cam, bg=(1.0, 1.0, 1.0)):
    consts = cam.view_constants(W, Hh, square_fov=False)
    sc = H.scene_tensors(cloud, DEV)
    vw = H.view_tensors(consts, bg, DEV)
    ws, out, R = H.run_forward_gsb(sc, vw, P, 1, 4, W, Hh)
    st = gsb.state(ws, P, W, Hh, R)
    ref_out, R_ref = H.run_forward_ref(sc, vw, P, 1, 4, W, Hh)
    ref = refshim.state(P, W, Hh, R_ref, DEV)
    assert R == R_ref
    assert torch.equal(st["radii"], ref["radii"])
    assert torch.equal(st["point_list_keys"], ref["point_list_keys"])
    assert torch.equal(st["point_list"], ref["point_list"])
    assert torch.equal(st["ranges"], ref["ranges"])
    assert torch.equal(st["n_contrib"], ref["n_contrib"])
    assert float((out - ref_out).abs().max()) <= 1e-4          # north_star: images within 1e-4 max-abs
    return sc, vw, ws, R, R_ref


@pytest.mark.parametrize("W,Hh", [(16, 16), (17, 33), (8192, 4608)])
def test_preview_resolution_limits(W, Hh):
    """16 x 16 is one tile; 17 x 33 has a one-pixel column and row of partial tiles; 8192 x 4608 has 147 456 tiles
    (18 tile-id bits: three 8-bit passes of the tile sort, 50 key bits in the reference's single sort)."""
    P = 40000
    cloud = gsb.synth.splat_cloud(P, 0x5EED0077, scale_lo=0.01, scale_hi=0.12)
    cam = gsb.Camera((0.0, 1.5, 9.0), (0, 0, 0), 55.0)
    _, _, _, R, _ = _compare_forward(cloud, P, W, Hh, cam)
    if W == 8192:
        assert R > 1000000


def test_three_million_splats_forward_and_backward():
    """BASELINE config C5's cloud (3 M splats, 1920 x 1080): keys / ranges / n_contrib bit for bit, image within 1e-4,
    every gradient tensor within 1e-3 relative L2 of the reference (its atomics are unordered)."""
    W, Hh, P = gsb.synth.CONFIGS["C5"]
    cloud = gsb.synth.splat_cloud(P, 0x5EED0005)
    cams = gsb.Camera.getCameras(gsb.synth.project_for("C5"))
    sc, vw, ws, R, R_ref = _compare_forward(cloud, P, W, Hh, cams[3])
    dL = torch.from_numpy((gsb.synth.uniform(3 * W * Hh, 11, 3) - 0.5).astype(np.float32)).to(DEV)
    g = H.run_backward_gsb(ws, sc, vw, P, 1, 4, R, W, Hh, dL)
    g_ref = H.run_backward_ref(sc, vw, P, 1, 4, R_ref, W, Hh, dL)
    for k in H.GRAD_NAMES:
        assert H.rel_l2(g[k], g_ref[k]) <= 1e-3, k
