"""GPU: the CUDA rasterizer (through the C ABI) against
  * the golden vectors of the reference's CUDA build (tests/golden/raster_*.npz),
  * the CPU oracle on fresh scenes,
  * the compiled reference itself (oracle/_ref/libgsref.so) where it travelled to the box,
plus the edge cases of SURVEY.md section 7.4.

Bars: sort keys, tile ranges, radii / tiles_touched / R bit-exact; images within
1e-4 max-abs (in practice they are bit-identical); every gradient tensor within 1e-3 relative L2."""
import ctypes as C

import numpy as np
import pytest
import torch

import golden_cases as GC
import helpers as H
from oracle import pyoracle, refshim

gsb = GC.gsb
pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(cloud, consts, bg, W, Hh, D=1, M=4, **kw):
    sc = H.scene_tensors(cloud, DEV)
    vw = H.view_tensors(consts, bg, DEV)
    P = cloud["opacities"].shape[0]
    ws, out, R = H.run_forward_gsb(sc, vw, P, D, M, W, Hh, **kw)
    torch.cuda.synchronize()
    return sc, vw, ws, out, R, gsb.state(ws, P, W, Hh, R)


@pytest.mark.parametrize("name", list(GC.RASTER_CASES))
def test_forward_matches_reference_golden(name):
    g = GC.load("raster_" + name)
    c, cloud, consts, _ = GC.raster_inputs(name)
    _, _, _, out, R, st = _run(cloud, consts, c["bg"], c["W"], c["H"])
    vis = torch.from_numpy(g["visible"]).to(DEV)
    eq = lambda a, b: bool(torch.equal(a.cpu(), torch.from_numpy(np.ascontiguousarray(b))))
    assert R == int(g["R"])
    assert eq(st["radii"], g["radii"])
    assert eq(st["tiles_touched"], g["tiles_touched"].astype(np.int32))
    assert eq(st["point_list_keys"], g["point_list_keys"].astype(np.int64))
    assert eq(st["point_list"], g["point_list"].astype(np.int32))
    assert eq(st["ranges"], g["ranges"].astype(np.int32))
    assert eq(st["n_contrib"], g["n_contrib"].astype(np.int32))
    assert eq(st["depths"][vis].view(torch.int32), g["depths"][g["visible"]].view(np.int32))
    assert eq(st["means2D"][vis].view(torch.int32), g["means2D"][g["visible"]].view(np.int32))
    assert eq(st["cov3D"][vis], g["cov3D"][g["visible"]])
    assert eq(st["conic_opacity"][vis], g["conic_opacity"][g["visible"]])
    cl = torch.stack([(st["clamped"] >> k) & 1 for k in range(3)], dim=1)
    assert eq(cl[vis], g["clamped"][g["visible"]].astype(np.uint8))
    assert (st["rgb"][vis].cpu() - torch.from_numpy(g["rgb"][g["visible"]])).abs().max() <= 1e-6
    assert (out.cpu() - torch.from_numpy(g["out_color"])).abs().max() <= 1e-4
    assert (st["final_T"].cpu() - torch.from_numpy(g["final_T"])).abs().max() <= 1e-4
    assert int(st["counters"][0]) == R and int(st["counters"][1]) == int(g["visible"].sum())


@pytest.mark.parametrize("name", list(GC.RASTER_CASES))
def test_tile_order_is_a_heavy_first_permutation(name):
    """The one-CTA-per-tile grids (tile sort, render forward / backward) take their tiles heaviest list first
    (csrc/common.cuh gsb_lpt_tile): every tile exactly once, list-length classes (two per octave from 32 entries up) never
    increasing along the order.  The reference launches its tile grid in raster order (forward.cu:261-280); the order
    changes no result, which the golden tests above check."""
    c, cloud, consts, _ = GC.raster_inputs(name)
    _, _, ws, _, R, st = _run(cloud, consts, c["bg"], c["W"], c["H"])
    order = gsb.tile_order(ws, cloud["locations"].shape[0], c["W"], c["H"]).astype(np.int64)
    tiles = ((c["W"] + 15) // 16) * ((c["H"] + 15) // 16)
    assert order.shape == (tiles,) and np.array_equal(np.sort(order), np.arange(tiles))
    ranges = st["ranges"].cpu().numpy().astype(np.int64)
    n = (ranges[:, 1] - ranges[:, 0])[order]

    def cls(v):
        if v < 32:
            return 0
        e = int(v).bit_length() - 1
        return min(15, 2 * (e - 5) + ((int(v) >> (e - 1)) & 1) + 1)
    k = np.array([cls(v) for v in n])
    assert np.all(np.diff(k) <= 0), (k[:20], n[:20])


@pytest.mark.parametrize("name", list(GC.RASTER_CASES))
def test_backward_matches_reference_golden(name):
    g = GC.load("raster_" + name)
    c, cloud, consts, dL = GC.raster_inputs(name)
    sc, vw, ws, _, R, _ = _run(cloud, consts, c["bg"], c["W"], c["H"])
    grads = H.run_backward_gsb(ws, sc, vw, c["P"], 1, 4, R, c["W"], c["H"], torch.from_numpy(dL).to(DEV))
    for k in H.GRAD_NAMES:
        rel = H.rel_l2(grads[k].cpu(), torch.from_numpy(g[k]))
        assert rel <= 1e-3, (k, rel)
    # unused slots stay untouched: mean2D.z, conic.z (backward.cu:545-552 never writes them)
    assert float(grads["dL_dmean2D"].view(-1, 3)[:, 2].abs().max()) == 0.0
    assert float(grads["dL_dconic"].view(-1, 4)[:, 2].abs().max()) == 0.0


def test_forward_backward_match_cpu_oracle_on_fresh_scene():
    W, Hh, P = 200, 120, 3000
    cloud = gsb.synth.splat_cloud(P, 12345, scale_hi=0.2)
    cam = gsb.Camera((3.0, 2.0, -8.0), (0, 0, 0), 50.0)
    consts = cam.view_constants(W, Hh)
    bg = (0.2, 0.4, 0.8)
    sc, vw, ws, out, R, st = _run(cloud, consts, bg, W, Hh)
    ref_out, v = pyoracle.forward(P, 1, 4, bg, W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"], cloud["scales"],
                                  1.0, cloud["rotations"], None, consts["view"], consts["projview"], consts["location"],
                                  consts["tan_fovx"], consts["tan_fovy"])
    assert R == v.R
    assert np.array_equal(st["radii"].cpu().numpy(), v.radii)
    assert np.array_equal(st["point_list_keys"].cpu().numpy().view(np.uint64), v.point_list_keys)
    assert np.array_equal(st["point_list"].cpu().numpy().astype(np.uint32), v.point_list)
    assert np.array_equal(st["ranges"].cpu().numpy().astype(np.uint32), v.ranges)
    npix = W * Hh
    assert (st["n_contrib"].cpu().numpy().astype(np.uint32) != v.n_contrib).sum() <= max(1, int(1e-4 * npix))
    d = np.abs(out.cpu().numpy() - ref_out)
    assert d.max() <= 5e-3 and (d > 1e-5).sum() <= max(1, int(1e-4 * d.size))
    dL = (gsb.synth.uniform(3 * npix, 5, 1) - 0.5).astype(np.float32)
    grads = H.run_backward_gsb(ws, sc, vw, P, 1, 4, R, W, Hh, torch.from_numpy(dL).to(DEV))
    og = v.backward(dL, 4)
    for k in H.GRAD_NAMES:
        assert H.rel_l2(grads[k].cpu(), torch.from_numpy(og[k])) <= 1e-3, k


needs_ref = pytest.mark.skipif(not refshim.available(), reason="oracle/_ref/libgsref.so not present")


def _compare_with_reference(cloud, consts, bg, W, Hh, D, M, backward=True, grad_tol={}, skip_grads=(), spread_aware=(), **kw):
    sc = H.scene_tensors(cloud, DEV)
    vw = H.view_tensors(consts, bg, DEV)
    P = cloud["opacities"].shape[0]
    for k in ("colors_precomp", "cov3D_precomp"):
        if kw.get(k) is not None:
            kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k], dtype=np.float32)).to(DEV)
    ref_out, R_ref = H.run_forward_ref(sc, vw, P, D, M, W, Hh, **kw)
    ref = refshim.state(P, W, Hh, R_ref, DEV)
    if kw.get("radii") is not None:
        # caller-supplied radii: the reference writes THAT array and leaves its internal one untouched
        # (rasterizer_impl.cu:233-236), so the caller's copy is the one to diff
        ref["radii"] = kw["radii"].clone()
        kw["radii"].zero_()
    ws, out, R = H.run_forward_gsb(sc, vw, P, D, M, W, Hh, **kw)
    st = gsb.state(ws, P, W, Hh, R)
    assert R == R_ref
    assert torch.equal(st["radii"], ref["radii"])
    if kw.get("radii") is not None:
        assert torch.equal(kw["radii"], ref["radii"])
    assert torch.equal(st["tiles_touched"], ref["tiles_touched"])
    assert torch.equal(st["point_list_keys"], ref["point_list_keys"])
    assert torch.equal(st["point_list"], ref["point_list"])
    assert torch.equal(st["ranges"], ref["ranges"])
    assert torch.equal(st["n_contrib"], ref["n_contrib"])
    assert float((out - ref_out).abs().max()) <= 1e-4
    if backward and kw.get("colors_precomp") is None and kw.get("cov3D_precomp") is None:
        dL = torch.from_numpy((gsb.synth.uniform(3 * W * Hh, 9, 3) - 0.5).astype(np.float32)).to(DEV)
        sm = kw.get("scale_modifier", 1.0)
        g_ref = H.run_backward_ref(sc, vw, P, D, M, R_ref, W, Hh, dL, scale_modifier=sm)
        g = H.run_backward_gsb(ws, sc, vw, P, D, M, R, W, Hh, dL, scale_modifier=sm)
        if spread_aware:
            # ill-conditioned tensors: the bar is what the REFERENCE itself achieves, measured here and asserted --
            # its run-to-run spread (unordered atomics, backward.cu:523-554; largest over three more runs) and its own
            # distance from the double-precision CPU oracle
            more = [H.run_backward_ref(sc, vw, P, D, M, R_ref, W, Hh, dL, scale_modifier=sm) for _ in range(3)]
            _, v = pyoracle.forward(P, D, M, bg, W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"], cloud["scales"],
                                    sm, cloud["rotations"], None, consts["view"], consts["projview"], consts["location"],
                                    consts["tan_fovx"], consts["tan_fovy"])
            exact = v.backward(dL.cpu().numpy(), M)
            v.close()
        for k in H.GRAD_NAMES:
            if k in skip_grads:
                continue
            tol = grad_tol.get(k, 1e-3)
            if k in spread_aware:
                spread = max(H.rel_l2(m[k], g_ref[k]) for m in more)
                ex = torch.from_numpy(exact[k].reshape(-1)).to(g[k].device)
                # the reference's distance from exact is itself a random variable (worst of its four runs here); one run
                # of either implementation lands within that plus the run-to-run spread
                ref_err = max(H.rel_l2(m[k], ex) for m in [g_ref] + more)
                my_err = H.rel_l2(g[k], ex)
                assert my_err <= max(1e-3, 1.5 * ref_err + spread), (k, my_err, ref_err, spread)
                tol = max(1e-3, 2.0 * spread, 2.0 * ref_err)
            assert H.rel_l2(g[k], g_ref[k]) <= tol, (k, H.rel_l2(g[k], g_ref[k]), tol)
    return st, out


@needs_ref
@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_degrees_against_compiled_reference(deg):
    M = (deg + 1) ** 2
    W, Hh, P = 256, 256, 10000          # BASELINE config C1 shape
    cloud = gsb.synth.splat_cloud(P, 0x5EED0001 + deg, sh_coeffs=M)
    cam = gsb.Camera.getCameras(gsb.synth.project_for("C1"))[0]
    _compare_with_reference(cloud, cam.view_constants(W, Hh, True), (1.0, 1.0, 1.0), W, Hh, deg, M)


@needs_ref
def test_api_surface_precomputed_inputs_scale_modifier_and_radii():
    W, Hh, P = 160, 112, 2500
    cloud = gsb.synth.splat_cloud(P, 77)
    cam = gsb.Camera((-5.0, 4.0, 7.0), (0.5, 0, 0), 40.0)
    consts = cam.view_constants(W, Hh)
    colors = gsb.synth.uniform(3 * P, 3, 0).astype(np.float32)
    radii = torch.zeros(P, dtype=torch.int32, device=DEV)
    st, _ = _compare_with_reference(cloud, consts, (0, 0, 0), W, Hh, 1, 4, colors_precomp=colors, radii=radii)
    assert torch.equal(radii, st["radii"])            # caller-supplied radii array is filled (rasterizer.h:54)
    _compare_with_reference(cloud, consts, (0, 0, 0), W, Hh, 1, 4, scale_modifier=2.5)
    _compare_with_reference(cloud, consts, (0, 0, 0), W, Hh, 1, 4, cov3D_precomp=st["cov3D"].cpu().numpy().reshape(-1))
    _compare_with_reference(cloud, consts, (0.5, 0.5, 0.5), W, Hh, 1, 4, debug=True)


@needs_ref
def test_degenerate_and_extreme_splats_against_compiled_reference():
    W, Hh, P = 192, 128, 1200
    cloud = gsb.synth.splat_cloud(P, 99)
    sc, rot, op, loc = (cloud[k].reshape(P, -1) for k in ("scales", "rotations", "opacities", "locations"))
    sc[0:50, 0] = 0.0                      # flat on one axis
    sc[50:80] = 0.0                        # zero covariance: only the 0.3 px low-pass remains
    rot[80:120] = 0.0                      # zero quaternion (never normalised, forward.cu:127)
    rot[120:160] *= 3.0                    # unnormalised quaternion
    op[160:200] = 1.0                      # alpha saturates at 0.99
    op[200:240] = 1.0 / 255.0              # at the skip threshold
    op[240:260] = 0.0
    sc[260:262] = 40.0                     # covers every tile
    loc[262:300, 2] = 9.85                 # just in front of the z = 0.2 cull plane (camera at z = 10 looks down -z)
    loc[300:320] = (0.0, 0.0, 11.0)        # behind the camera
    sc[320:360] = (0.3, 0.004, 0.004)      # needle: strongest anisotropy the trainer's clamps allow
    cam = gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0)
    # The needle and scale-40 splats make the covariance-path gradients ill-conditioned in float32: two runs
    # of the REFERENCE differ by ~3e-3 (dL_drot) / ~1e-3 (dL_dcov3D) here because its atomics sum in a different
    # order each time (backward.cu:523-554).  For those three tensors the bar is therefore measured inside the test
    # (spread_aware): no further from the double-precision oracle than 1.5 x the reference's worst run plus its spread, and within twice the
    # reference's own spread / oracle distance of the reference.  Every other tensor keeps 1e-3.
    loose = ("dL_drot", "dL_dcov3D", "dL_dscale")
    _compare_with_reference(cloud, cam.view_constants(W, Hh), (1.0, 1.0, 1.0), W, Hh, 1, 4, spread_aware=loose)
    # splats far outside the 1.3 tan(fov) clamp of computeCov2D (forward.cu:82-87)
    cam2 = gsb.Camera((0.0, 0.0, 4.5), (0, 0, 0), 25.0)
    _compare_with_reference(cloud, cam2.view_constants(W, Hh), (0.0, 0.0, 0.0), W, Hh, 1, 4, spread_aware=loose)


@needs_ref
def test_equal_depth_ties_keep_splat_index_order():
    # the reference's own initial field: a regular grid seen head-on (src/ui/UiFrame.cpp:137-160) -> many equal depths
    n = 17
    axis = np.linspace(-4, 4, n, dtype=np.float32)
    gx, gy, gz = np.meshgrid(axis, axis, axis, indexing="ij")
    P = n ** 3
    cloud = {"locations": np.stack([gx, gy, gz], -1).reshape(-1).astype(np.float32),
             "scales": np.full(3 * P, 0.05, np.float32), "opacities": np.ones(P, np.float32),
             "rotations": np.tile(np.array([1, 0, 0, 0], np.float32), P), "shs": np.zeros(12 * P, np.float32)}
    cam = gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0)
    # isotropic splats: the covariance does not depend on the rotation, dL/drot is pure rounding noise in both
    st, _ = _compare_with_reference(cloud, cam.view_constants(256, 256, True), (0, 0, 0), 256, 256, 1, 4,
                                    skip_grads=("dL_drot",))
    keys = st["point_list_keys"].cpu().numpy()
    ties = keys[1:] == keys[:-1]
    assert ties.any()
    pl = st["point_list"].cpu().numpy()
    assert np.all(pl[1:][ties] > pl[:-1][ties])        # stable sort: equal keys stay in splat order


def test_empty_and_culled_inputs():
    W, Hh = 100, 60
    cam = gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0)
    consts = cam.view_constants(W, Hh)
    bg = (0.25, 0.5, 0.75)
    expect = torch.tensor(bg).repeat_interleave(W * Hh)
    # P = 0 (the reference would read point_offsets[-1], rasterizer_impl.cu:281)
    empty = {k: np.zeros(0, np.float32) for k in ("locations", "shs", "scales", "opacities", "rotations")}
    sc = {k: torch.zeros(1, device=DEV) for k in empty}
    vw = H.view_tensors(consts, bg, DEV)
    ws, out, R = H.run_forward_gsb(sc, vw, 0, 1, 4, W, Hh)
    assert R == 0 and torch.equal(out.cpu(), expect)
    # everything behind the camera: R = 0, image = background, backward leaves gradients at zero
    P = 500
    cloud = gsb.synth.splat_cloud(P, 5)
    cloud["locations"].reshape(P, 3)[:, 2] += 20.0
    sc, vw, ws, out, R, st = _run(cloud, consts, bg, W, Hh)
    assert R == 0 and int(st["radii"].abs().sum()) == 0 and torch.equal(out.cpu(), expect)
    assert int(st["n_contrib"].sum()) == 0 and float((st["final_T"] - 1).abs().max()) == 0.0
    grads = H.run_backward_gsb(ws, sc, vw, P, 1, 4, R, W, Hh, torch.ones(3 * W * Hh, device=DEV))
    assert all(float(g.abs().sum()) == 0.0 for g in grads.values())


def test_single_splat_and_mark_visible():
    W, Hh = 64, 64
    cloud = {"locations": np.array([0, 0, 0], np.float32), "scales": np.array([0.4, 0.2, 0.1], np.float32),
             "opacities": np.array([0.9], np.float32), "rotations": np.array([0.9, 0.1, 0.3, 0.2], np.float32),
             "shs": np.array([1.0, 0.5, -0.2] + [0.0] * 9, np.float32)}
    cam = gsb.Camera((0.0, 0.0, 5.0), (0, 0, 0), 60.0)
    consts = cam.view_constants(W, Hh, True)
    _, vw, _, out, R, st = _run(cloud, consts, (0, 0, 0), W, Hh)
    ref_out, v = pyoracle.forward(1, 1, 4, (0, 0, 0), W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"],
                                  cloud["scales"], 1.0, cloud["rotations"], None, consts["view"], consts["projview"],
                                  consts["location"], consts["tan_fovx"], consts["tan_fovy"])
    assert R == v.R and int(st["radii"][0]) == int(v.radii[0]) > 0
    assert np.abs(out.cpu().numpy() - ref_out).max() <= 1e-5
    pts = torch.tensor([[0, 0, 0], [0, 0, 4.9], [0, 0, 6.0], [0, 0, 4.7]], dtype=torch.float32, device=DEV).reshape(-1)
    present = torch.zeros(4, dtype=torch.bool, device=DEV)
    gsb.Rasterizer.markVisible(4, pts, vw["view"], vw["projview"], present)
    assert present.cpu().tolist() == list(pyoracle.mark_visible(4, pts.cpu().numpy(), consts["view"]))
    assert present.cpu().tolist() == [True, False, False, True]       # z_view > 0.2 only (auxiliary.h:152-154)


def test_errors_are_reported_not_swallowed():
    L = gsb.lib()
    lib_mod = __import__("importlib").import_module("gaussian-splatterer_b200._lib")
    null_alloc = lib_mod.ALLOC_FN(lambda n, u: 0)
    n = C.c_int(0)
    one = torch.zeros(16, device=DEV)
    rc = L.gsb_rasterizer_forward(null_alloc, None, null_alloc, None, null_alloc, None, 1, 1, 4, one.data_ptr(), 32, 32,
                                  one.data_ptr(), one.data_ptr(), None, one.data_ptr(), one.data_ptr(), 1.0, one.data_ptr(), None,
                                  one.data_ptr(), one.data_ptr(), one.data_ptr(), 0.5, 0.5, 0, one.data_ptr(), None, 0, None, C.byref(n))
    assert rc != 0 and b"allocator" in L.gsb_last_error()
    ws = gsb.Workspaces(DEV)
    with pytest.raises(gsb.GsbError):      # neither SHs nor precomputed colours (rasterizer_impl.cu:242-245 throws too)
        gsb.Rasterizer.forward(ws, 1, 1, 4, one, 32, 32, one, None, None, one, one, 1.0, one, None, one, one, one, 0.5, 0.5, False, one)
