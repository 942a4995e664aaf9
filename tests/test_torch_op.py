"""The PyTorch operator boundary: gaussian-splatterer_b200/diff_gaussian_rasterization.py
against the reference's torch binding semantics (lib/diff-gaussian-rasterization/rasterize_points.cu:35-217,
diff_gaussian_rasterization/__init__.py:20-220).

CPU part: the module surface (names, NamedTuple fields, argument validation, empty-P behaviour).
GPU part: forward tuple / backward tuple / autograd against the compiled reference rasterizer called the way
rasterize_points.cu calls it (SH degrees 0-3, colors_precomp, cov3D_precomp), and against the CPU oracle.
Bars: num_rendered and radii bit-exact, colour <= 1e-4 max-abs, every gradient <= 1e-3 relative L2."""
import importlib
import inspect

import numpy as np
import pytest
import torch

import helpers as H
from oracle import pyoracle, refshim

gsb = H.gsb
dgr = importlib.import_module("gaussian-splatterer_b200.diff_gaussian_rasterization")
needs_ref = pytest.mark.skipif(not refshim.available(), reason="oracle/_ref/libgsref.so not built")
DEV = "cuda"


# ----------------------------------------------------------------------------------------------- CPU
def test_module_surface_matches_reference_binding():
    # __init__.py:154-166
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    # ext.cpp:15-18
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(dgr._C, name))
    # rasterize_points.cu:36-55 and :121-142: positional order of the C++ functions
    assert list(inspect.signature(dgr._C.rasterize_gaussians).parameters) == [
        "background", "means3D", "colors", "opacity", "scales", "rotations", "scale_modifier", "cov3D_precomp",
        "viewmatrix", "projmatrix", "tan_fovx", "tan_fovy", "image_height", "image_width", "sh", "degree", "campos",
        "prefiltered", "debug"]
    assert list(inspect.signature(dgr._C.rasterize_gaussians_backward).parameters) == [
        "background", "means3D", "radii", "colors", "scales", "rotations", "scale_modifier", "cov3D_precomp",
        "viewmatrix", "projmatrix", "tan_fovx", "tan_fovy", "dL_dout_color", "sh", "degree", "campos", "geomBuffer", "R",
        "binningBuffer", "imageBuffer", "debug"]
    # __init__.py:187: keyword names of GaussianRasterizer.forward
    assert list(inspect.signature(dgr.GaussianRasterizer.forward).parameters)[1:] == [
        "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]


def _settings(W, Hh, consts, bg, deg, device, debug=False):
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32), device=device)
    return dgr.GaussianRasterizationSettings(
        image_height=Hh, image_width=W, tanfovx=consts["tan_fovx"], tanfovy=consts["tan_fovy"], bg=t(bg),
        scale_modifier=1.0, viewmatrix=t(consts["view"]).reshape(4, 4), projmatrix=t(consts["projview"]).reshape(4, 4),
        sh_degree=deg, campos=t(consts["location"]), prefiltered=False, debug=debug)


def test_argument_validation_like_reference():
    cam = gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0)
    rs = _settings(64, 64, cam.view_constants(64, 64), (0, 0, 0), 1, "cpu")
    r = dgr.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):     # __init__.py:191-192
        r(m, m, torch.ones(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.ones(4, 1), shs=torch.zeros(4, 4, 3), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):             # __init__.py:194-195
        r(m, m, torch.ones(4, 1), colors_precomp=m)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, torch.ones(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):   # rasterize_points.cu:56-58
        dgr._C.rasterize_gaussians(rs.bg, torch.zeros(4, 2), m, m, m, m, 1.0, m, rs.viewmatrix, rs.projmatrix, 1.0, 1.0,
                                   64, 64, m, 1, rs.campos, False, False)
    # no CPU fallback: CPU tensors with P > 0 are refused, loudly
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.ones(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))


def test_empty_point_set_needs_no_device():
    """P == 0: the reference skips the rasterizer (rasterize_points.cu:83,156) and returns zero-filled tensors."""
    cam = gsb.Camera((0.0, 0.0, 10.0), (0, 0, 0), 60.0)
    rs = _settings(32, 16, cam.view_constants(32, 16), (0, 0, 0), 0, "cpu")
    e = torch.zeros(0, 3)
    n, color, radii, gb, bb, ib = dgr._C.rasterize_gaussians(rs.bg, e, e, torch.zeros(0, 1), e, torch.zeros(0, 4), 1.0,
                                                             torch.Tensor([]), rs.viewmatrix, rs.projmatrix, 1.0, 1.0,
                                                             16, 32, torch.Tensor([]), 0, rs.campos, False, False)
    assert n == 0 and color.shape == (3, 16, 32) and float(color.abs().sum()) == 0.0
    assert radii.shape == (0,) and radii.dtype == torch.int32 and gb.numel() == bb.numel() == ib.numel() == 0
    assert dgr._C.mark_visible(e, rs.viewmatrix, rs.projmatrix).shape == (0,)
    g = dgr._C.rasterize_gaussians_backward(rs.bg, e, radii, e, e, torch.zeros(0, 4), 1.0, torch.Tensor([]), rs.viewmatrix,
                                            rs.projmatrix, 1.0, 1.0, torch.zeros(3, 16, 32), torch.Tensor([]), 0, rs.campos,
                                            gb, 0, bb, ib, False)
    assert [tuple(x.shape) for x in g] == [(0, 3), (0, 3), (0, 1), (0, 3), (0, 6), (0, 0, 3), (0, 3), (0, 4)]


# ----------------------------------------------------------------------------------------------- GPU
def _scene(P, seed, deg, W, Hh, cam=None):
    M = (deg + 1) ** 2
    cloud = gsb.synth.splat_cloud(P, seed, sh_coeffs=M, scale_lo=0.02, scale_hi=0.15)
    cam = cam or gsb.Camera((3.0, 2.5, 8.0), (0.2, 0, 0), 50.0)
    consts = cam.view_constants(W, Hh, square_fov=False)
    t = lambda k, shape: torch.from_numpy(np.ascontiguousarray(cloud[k], dtype=np.float32)).to(DEV).reshape(shape)
    ten = {"means3D": t("locations", (P, 3)), "shs": t("shs", (P, M, 3)), "scales": t("scales", (P, 3)),
           "opacities": t("opacities", (P, 1)), "rotations": t("rotations", (P, 4))}
    return cloud, consts, ten, M


def _reference_op(rs, ten, M, dL, colors_precomp=None, cov3D_precomp=None):
    """What the reference's _C.rasterize_gaussians + _backward return, by calling the compiled reference rasterizer
    exactly as rasterize_points.cu does (zero-filled outputs, null for absent inputs)."""
    P = ten["means3D"].shape[0]
    W, Hh = rs.image_width, rs.image_height
    out = torch.zeros(3 * Hh * W, dtype=torch.float32, device=DEV)
    radii = torch.zeros(P, dtype=torch.int32, device=DEV)
    shs = None if colors_precomp is not None else ten["shs"]
    sca = None if cov3D_precomp is not None else ten["scales"]
    rot = None if cov3D_precomp is not None else ten["rotations"]
    Mr = 0 if shs is None else M
    R = refshim.forward(P, rs.sh_degree, Mr, rs.bg, W, Hh, ten["means3D"], shs, colors_precomp, ten["opacities"], sca,
                        rs.scale_modifier, rot, cov3D_precomp, rs.viewmatrix, rs.projmatrix, rs.campos, rs.tanfovx,
                        rs.tanfovy, False, out, radii, False)
    g = H.grad_buffers(P, max(Mr, 1), DEV)
    refshim.backward(P, rs.sh_degree, Mr, R, rs.bg, W, Hh, ten["means3D"], shs, colors_precomp, sca, rs.scale_modifier, rot,
                     cov3D_precomp, rs.viewmatrix, rs.projmatrix, rs.campos, rs.tanfovx, rs.tanfovy, radii, dL.reshape(-1),
                     *[g[n] for n in H.GRAD_NAMES])
    torch.cuda.synchronize()
    return R, out.reshape(3, Hh, W), radii, g


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_autograd_matches_compiled_reference_for_every_sh_degree(deg):
    W, Hh, P = 200, 136, 4000
    cloud, consts, ten, M = _scene(P, 0x7E57 + deg, deg, W, Hh)
    rs = _settings(W, Hh, consts, (0.3, 0.6, 0.1), deg, DEV)
    leaves = {k: v.clone().requires_grad_(True) for k, v in ten.items()}
    means2D = torch.zeros(P, 3, device=DEV, requires_grad=True)
    color, radii = dgr.GaussianRasterizer(rs)(leaves["means3D"], means2D, leaves["opacities"], shs=leaves["shs"],
                                               scales=leaves["scales"], rotations=leaves["rotations"])
    assert color.shape == (3, Hh, W) and radii.shape == (P,) and radii.dtype == torch.int32 and not radii.requires_grad
    dL = torch.from_numpy((gsb.synth.uniform(3 * W * Hh, 11, deg) - 0.5).astype(np.float32)).to(DEV).reshape(3, Hh, W)
    color.backward(dL)
    R_ref, c_ref, radii_ref, g = _reference_op(rs, ten, M, dL)
    assert torch.equal(radii, radii_ref)
    assert float((color.detach() - c_ref).abs().max()) <= 1e-4
    pairs = {"means3D": "dL_dmean3D", "shs": "dL_dsh", "scales": "dL_dscale", "opacities": "dL_dopacity",
             "rotations": "dL_drot"}
    for leaf, name in pairs.items():
        assert leaves[leaf].grad.shape == leaves[leaf].shape
        assert H.rel_l2(leaves[leaf].grad.reshape(-1), g[name]) <= 1e-3, (leaf, deg)
    assert H.rel_l2(means2D.grad.reshape(-1), g["dL_dmean2D"]) <= 1e-3        # screen-space gradient (densification stat)


@pytest.mark.gpu
@needs_ref
def test_precomputed_colors_and_covariances_forward_and_backward():
    W, Hh, P = 176, 144, 3000
    cloud, consts, ten, M = _scene(P, 0xC0105, 1, W, Hh)
    rs = _settings(W, Hh, consts, (0.0, 0.0, 0.0), 1, DEV)
    colors = torch.from_numpy(gsb.synth.uniform(3 * P, 5, 1).astype(np.float32)).to(DEV).reshape(P, 3)
    # cov3D of the same splats, taken from the reference's own geometry state
    R0, _, _, _ = _reference_op(rs, ten, M, torch.zeros(3, Hh, W, device=DEV))
    cov = refshim.state(P, W, Hh, R0, DEV)["cov3D"].contiguous()
    dL = torch.from_numpy((gsb.synth.uniform(3 * W * Hh, 12, 0) - 0.5).astype(np.float32)).to(DEV).reshape(3, Hh, W)

    # (a) colors_precomp + scale/rotation
    lv = {k: ten[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations")}
    col = colors.clone().requires_grad_(True)
    m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)
    color, radii = dgr.GaussianRasterizer(rs)(lv["means3D"], m2, lv["opacities"], colors_precomp=col, scales=lv["scales"],
                                               rotations=lv["rotations"])
    color.backward(dL)
    _, c_ref, radii_ref, g = _reference_op(rs, ten, M, dL, colors_precomp=colors)
    assert torch.equal(radii, radii_ref) and float((color.detach() - c_ref).abs().max()) <= 1e-4
    for leaf, name in (("means3D", "dL_dmean3D"), ("opacities", "dL_dopacity"), ("scales", "dL_dscale"),
                       ("rotations", "dL_drot")):
        assert H.rel_l2(lv[leaf].grad.reshape(-1), g[name]) <= 1e-3, leaf
    assert H.rel_l2(col.grad.reshape(-1), g["dL_dcolor"]) <= 1e-3

    # (b) SH colours + cov3D_precomp
    lv = {k: ten[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs")}
    cv = cov.clone().requires_grad_(True)
    m2 = torch.zeros(P, 3, device=DEV, requires_grad=True)
    color, radii = dgr.GaussianRasterizer(rs)(lv["means3D"], m2, lv["opacities"], shs=lv["shs"], cov3D_precomp=cv)
    color.backward(dL)
    _, c_ref, radii_ref, g = _reference_op(rs, ten, M, dL, cov3D_precomp=cov)
    assert torch.equal(radii, radii_ref) and float((color.detach() - c_ref).abs().max()) <= 1e-4
    assert H.rel_l2(cv.grad.reshape(-1), g["dL_dcov3D"]) <= 1e-3
    assert H.rel_l2(lv["means3D"].grad.reshape(-1), g["dL_dmean3D"]) <= 1e-3
    assert H.rel_l2(lv["shs"].grad.reshape(-1), g["dL_dsh"]) <= 1e-3


@pytest.mark.gpu
def test_raw_tuples_against_cpu_oracle_and_workspace_tensors():
    """_C.rasterize_gaussians / _backward tuple layout (rasterize_points.cu:117,193) checked against the CPU oracle;
    the three workspaces come back as uint8 CUDA tensors sized by the allocator callbacks."""
    W, Hh, P, deg = 120, 90, 1500, 1
    cloud, consts, ten, M = _scene(P, 0x0AC1E, deg, W, Hh)
    rs = _settings(W, Hh, consts, (1.0, 1.0, 1.0), deg, DEV, debug=True)
    E = torch.Tensor([])
    n, color, radii, gb, bb, ib = dgr._C.rasterize_gaussians(
        rs.bg, ten["means3D"], E, ten["opacities"], ten["scales"], ten["rotations"], 1.0, E, rs.viewmatrix, rs.projmatrix,
        rs.tanfovx, rs.tanfovy, Hh, W, ten["shs"], deg, rs.campos, False, True)
    for b in (gb, bb, ib):
        assert b.dtype == torch.uint8 and b.is_cuda and b.numel() > 0
    o_out, v = pyoracle.forward(P, deg, M, (1.0, 1.0, 1.0), W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"],
                                cloud["scales"], 1.0, cloud["rotations"], None, consts["view"], consts["projview"],
                                consts["location"], consts["tan_fovx"], consts["tan_fovy"])
    assert n == v.R
    assert np.array_equal(radii.cpu().numpy(), v.radii)
    assert float(np.abs(color.cpu().numpy().reshape(-1) - o_out.reshape(-1)).max()) <= 1e-4
    dL = (gsb.synth.uniform(3 * W * Hh, 13, 0) - 0.5).astype(np.float32)
    out = dgr._C.rasterize_gaussians_backward(
        rs.bg, ten["means3D"], radii, E, ten["scales"], ten["rotations"], 1.0, E, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
        rs.tanfovy, torch.from_numpy(dL).to(DEV).reshape(3, Hh, W), ten["shs"], deg, rs.campos, gb, n, bb, ib, True)
    assert [tuple(x.shape) for x in out] == [(P, 3), (P, 3), (P, 1), (P, 3), (P, 6), (P, M, 3), (P, 3), (P, 4)]
    og = v.backward(dL, M)
    names = ("dL_dmean2D", "dL_dcolor", "dL_dopacity", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot")
    for t, k in zip(out, names):
        a, b = t.cpu().numpy().astype(np.float64).reshape(-1), og[k].astype(np.float64).reshape(-1)
        assert np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30) <= 1e-3, k
    v.close()
    # markVisible through the module == radii-independent frustum test of the oracle
    vis = dgr.GaussianRasterizer(rs).markVisible(ten["means3D"])
    assert vis.dtype == torch.bool and np.array_equal(vis.cpu().numpy(), pyoracle.mark_visible(P, cloud["locations"], consts["view"]))


@pytest.mark.gpu
def test_optimizer_step_through_the_module_reduces_the_loss():
    """An INRIA-style loop: render, L2 loss against a target render, Adam on SH + opacity; the loss must go down and
    the call must honour the current (non-default) torch stream."""
    W, Hh, P, deg = 128, 128, 2000, 1
    _, consts, ten, M = _scene(P, 0x0071, deg, W, Hh)
    rs = _settings(W, Hh, consts, (0.0, 0.0, 0.0), deg, DEV)
    r = dgr.GaussianRasterizer(rs)
    with torch.no_grad():
        target, _ = r(ten["means3D"], torch.zeros(P, 3, device=DEV), ten["opacities"], shs=ten["shs"], scales=ten["scales"],
                      rotations=ten["rotations"])
    shs = (ten["shs"] * 0.5).clone().requires_grad_(True)
    opa = (ten["opacities"] * 0.8).clone().requires_grad_(True)
    opt = torch.optim.Adam([shs, opa], lr=0.02)
    losses = []
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(12):
            opt.zero_grad()
            img, _ = r(ten["means3D"], torch.zeros(P, 3, device=DEV), opa, shs=shs, scales=ten["scales"],
                       rotations=ten["rotations"])
            loss = ((img - target) ** 2).mean()
            loss.backward()
            opt.step()
            losses.append(float(loss.detach()))
    torch.cuda.synchronize()
    assert losses[-1] < 0.5 * losses[0], losses
