"""Shared helpers of the test-suite: scenes on the device, calls into the product (through the C ABI)
and into the checkers (oracle/)."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

gsb = importlib.import_module("gaussian-splatterer_b200")


def scene_tensors(cloud, device="cuda"):
    import torch
    return {k: torch.from_numpy(np.ascontiguousarray(v)).to(device) for k, v in cloud.items()}


def view_tensors(consts, bg, device="cuda"):
    import torch
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32), device=device)
    return {"view": t(consts["view"]), "projview": t(consts["projview"]), "location": t(consts["location"]),
            "bg": t(bg), "tan_fovx": consts["tan_fovx"], "tan_fovy": consts["tan_fovy"]}


def run_forward_gsb(sc, vw, P, D, M, W, H, scale_modifier=1.0, radii=None, debug=False, colors_precomp=None,
                    cov3D_precomp=None):
    import torch
    ws = gsb.Workspaces(sc["locations"].device)
    out = torch.zeros(3 * H * W, dtype=torch.float32, device=sc["locations"].device)
    R = gsb.Rasterizer.forward(ws, P, D, M, vw["bg"], W, H, sc["locations"], None if colors_precomp is not None else sc["shs"],
                               colors_precomp, sc["opacities"], None if cov3D_precomp is not None else sc["scales"],
                               scale_modifier, None if cov3D_precomp is not None else sc["rotations"], cov3D_precomp,
                               vw["view"], vw["projview"], vw["location"], vw["tan_fovx"], vw["tan_fovy"], False, out,
                               radii, debug)
    return ws, out, R


def run_forward_ref(sc, vw, P, D, M, W, H, scale_modifier=1.0, radii=None, debug=False, colors_precomp=None,
                    cov3D_precomp=None):
    import torch
    from oracle import refshim
    out = torch.zeros(3 * H * W, dtype=torch.float32, device=sc["locations"].device)
    R = refshim.forward(P, D, M, vw["bg"], W, H, sc["locations"], None if colors_precomp is not None else sc["shs"],
                        colors_precomp, sc["opacities"], None if cov3D_precomp is not None else sc["scales"],
                        scale_modifier, None if cov3D_precomp is not None else sc["rotations"], cov3D_precomp,
                        vw["view"], vw["projview"], vw["location"], vw["tan_fovx"], vw["tan_fovy"], False, out, radii,
                        debug)
    return out, R


GRAD_NAMES = ("dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale",
              "dL_drot")


def grad_buffers(P, M, device):
    import torch
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=device)
    return {"dL_dmean2D": z(3 * P), "dL_dconic": z(4 * P), "dL_dopacity": z(P), "dL_dcolor": z(3 * P),
            "dL_dmean3D": z(3 * P), "dL_dcov3D": z(6 * P), "dL_dsh": z(3 * M * P), "dL_dscale": z(3 * P),
            "dL_drot": z(4 * P)}


def run_backward_gsb(ws, sc, vw, P, D, M, R, W, H, dL_dpix, scale_modifier=1.0):
    g = grad_buffers(P, M, dL_dpix.device)
    gsb.Rasterizer.backward(ws, P, D, M, R, vw["bg"], W, H, sc["locations"], sc["shs"], None, sc["scales"], scale_modifier,
                            sc["rotations"], None, vw["view"], vw["projview"], vw["location"], vw["tan_fovx"],
                            vw["tan_fovy"], None, dL_dpix, *[g[n] for n in GRAD_NAMES])
    return g


def run_backward_ref(sc, vw, P, D, M, R, W, H, dL_dpix, scale_modifier=1.0):
    from oracle import refshim
    g = grad_buffers(P, M, dL_dpix.device)
    refshim.backward(P, D, M, R, vw["bg"], W, H, sc["locations"], sc["shs"], None, sc["scales"], scale_modifier,
                     sc["rotations"], None, vw["view"], vw["projview"], vw["location"], vw["tan_fovx"], vw["tan_fovy"],
                     None, dL_dpix, *[g[n] for n in GRAD_NAMES])
    return g


def rel_l2(a, b):
    import torch
    a = a.double(); b = b.double()
    den = torch.linalg.norm(b).item()
    num = torch.linalg.norm(a - b).item()
    return num / den if den > 0 else num
