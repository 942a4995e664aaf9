"""CPU: the scalar oracle (oracle/oracle.cpp) against golden vectors produced by the reference's own CUDA
build on a B200 (tools/make_golden.py).  This is what pins the oracle; the GPU tests then use the
oracle (and the goldens, and oracle/_ref where present) to check the CUDA product.

Tolerances: everything the reference discretises must be bit-exact.  Floating-point images / T within
1e-5 except for isolated pixels (the oracle's exp2f is not the GPU's MUFU.EX2, so a borderline 1/255
or 1e-4 decision may flip): at most 1e-4 of the pixels may exceed it and none may exceed 5e-3.
Gradients within 1e-3 relative L2 (the north-star tolerance; the goldens are the mean of 8 reference
runs because its atomics are order-nondeterministic).
"""
import numpy as np
import pytest

import golden_cases as GC
from oracle import pyoracle
from oracle_train import oracle_running_means


def _forward(c, cloud, consts):
    return pyoracle.forward(c["P"], 1, 4, c["bg"], c["W"], c["H"], cloud["locations"], cloud["shs"], None, cloud["opacities"],
                            cloud["scales"], 1.0, cloud["rotations"], None, consts["view"], consts["projview"],
                            consts["location"], consts["tan_fovx"], consts["tan_fovy"])


@pytest.mark.parametrize("name", list(GC.RASTER_CASES))
def test_oracle_forward_matches_reference(name):
    g = GC.load("raster_" + name)
    assert g is not None, "golden vectors missing: run tools/make_golden.py on a GPU box"
    c, cloud, consts, _ = GC.raster_inputs(name)
    out, v = _forward(c, cloud, consts)
    vis = g["visible"]
    assert v.R == int(g["R"])
    assert np.array_equal(v.radii, g["radii"])
    assert np.array_equal(v.tiles_touched, g["tiles_touched"].astype(np.uint32))
    assert np.array_equal(v.point_offsets, g["point_offsets"].astype(np.uint32))
    assert np.array_equal(v.depths.view(np.uint32)[vis], g["depths"].view(np.uint32)[vis])
    assert np.array_equal(v.means2D.view(np.uint32)[vis], g["means2D"].view(np.uint32)[vis])
    assert np.array_equal(v.point_list_keys, g["point_list_keys"].view(np.uint64))
    assert np.array_equal(v.point_list, g["point_list"].astype(np.uint32))
    assert np.array_equal(v.ranges, g["ranges"].astype(np.uint32))
    assert np.array_equal(v.clamped[vis], g["clamped"][vis])
    np.testing.assert_allclose(v.cov3D[vis], g["cov3D"][vis], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(v.conic_opacity[vis], g["conic_opacity"][vis], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(v.rgb[vis], g["rgb"][vis], atol=1e-6)
    npix = c["W"] * c["H"]
    assert (v.n_contrib != g["n_contrib"].astype(np.uint32)).sum() <= max(1, int(1e-4 * npix))
    for mine, ref in ((out, g["out_color"]), (v.final_T, g["final_T"])):
        d = np.abs(mine - ref)
        assert d.max() <= 5e-3
        assert (d > 1e-5).sum() <= max(1, int(1e-4 * d.size))


@pytest.mark.parametrize("name", list(GC.RASTER_CASES))
def test_oracle_backward_matches_reference(name):
    g = GC.load("raster_" + name)
    c, cloud, consts, dL = GC.raster_inputs(name)
    _, v = _forward(c, cloud, consts)
    grads = v.backward(dL, 4)
    for k, mine in grads.items():
        ref = g[k].astype(np.float64)
        rel = np.linalg.norm(mine.astype(np.float64) - ref) / max(np.linalg.norm(ref), 1e-30)
        assert rel <= 1e-3, (k, rel)


def test_oracle_mark_visible():
    g = GC.load("raster_close")
    c, cloud, consts, _ = GC.raster_inputs("close")
    assert np.array_equal(pyoracle.mark_visible(c["P"], cloud["locations"], consts["view"]), g["present"])


@pytest.mark.parametrize("name", list(GC.TRAINER_CASES))
def test_oracle_trainer_iteration_apply_densify(name):
    g = GC.load("trainer_" + name)
    assert g is not None
    c, cloud, _teacher, project, cams = GC.trainer_inputs(name)
    P, M = c["P"], 4
    avg = oracle_running_means(c, cloud, cams, g["truthW"], g["truthB"])
    for k, mine in zip(("var", "loc", "sh", "scale", "opacity", "rot"), avg):
        ref = g["avg_" + k].astype(np.float64)
        rel = np.linalg.norm(mine.astype(np.float64) - ref) / max(np.linalg.norm(ref), 1e-30)
        assert rel <= 1e-3, (k, rel)
    # applyGradients (src/Trainer.cu:81-101)
    params = [cloud[k].copy() for k in ("locations", "shs", "scales", "opacities", "rotations")]
    pyoracle.apply(params, avg[1:], [project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation],
                   project.paramScaleMax, M, P)
    for k, mine in zip(("locations", "shs", "scales", "opacities", "rotations"), params):
        np.testing.assert_allclose(mine, g["applied_" + k], atol=2e-5)
    # densify (src/Trainer.cu:437-542) on the reference's own post-apply model and running means: exact decisions
    applied = {k: g["applied_" + k] for k in ("locations", "shs", "scales", "opacities", "rotations")}
    dens = pyoracle.densify(applied, P, 1000000, M, g["densify_var"], g["densify_avg_loc"], g["densify_params13"])
    assert dens["opacities"].shape[0] == int(g["densified_count"])
    assert int(g["densified_count"]) != P

    def rows(m):
        n = m["opacities"].shape[0]
        r = np.concatenate([m["locations"].reshape(n, 3), m["scales"].reshape(n, 3), m["opacities"].reshape(n, 1),
                            m["rotations"].reshape(n, 4), m["shs"].reshape(n, 3 * M)], axis=1)
        return r[np.lexsort(np.round(r[:, ::-1], 4).T)]
    ref_rows = rows({k: g["densified_" + k] for k in ("locations", "shs", "scales", "opacities", "rotations")})
    np.testing.assert_allclose(rows(dens), ref_rows, atol=1e-5)


def test_oracle_image_conversions():
    rng = np.random.default_rng(3)
    w, h = 37, 19
    img = rng.uniform(-0.2, 1.3, 3 * w * h).astype(np.float32)
    fb = pyoracle.image_float_to_int(img, w, h)
    r = np.clip((img[:w * h] * np.float32(256.0)).astype(np.int32), 0, 255)
    assert np.array_equal(fb & 0xFF, r.astype(np.uint32))
    assert np.all(fb >> 24 == 0xFF)
    loss = pyoracle.image_int_to_loss(fb, img, w, h)
    np.testing.assert_allclose(loss[:w * h], (fb & 0xFF).astype(np.float32) / np.float32(255.0) - img[:w * h], atol=1e-7)
