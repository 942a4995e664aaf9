"""Python mirror of CudaRasterizer::Rasterizer (reference
lib/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:24-84) over the C ABI.

Arguments keep the reference's names, order and meaning; device arrays are torch CUDA tensors
(torch is only the device-memory plumbing here).  The three workspace allocators are Python
callables `size -> uint8 tensor`, exactly like the std::function<char*(size_t)> lambdas the
reference's callers pass (src/Trainer.cu:335-337, rasterize_points.cu:27-33).
"""
import ctypes as C

import torch

from ._lib import ALLOC_FN, StateView, check, lib


def _ptr(t):
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "device arrays must be contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


class Workspaces:
    """Holds the three opaque blobs of one forward call (geomBuffer / binningBuffer / imgBuffer)."""

    def __init__(self, device="cuda"):
        self.device = device
        self.geom = self.binning = self.image = None

    def _make(self, name):
        def alloc(nbytes, _user):
            buf = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=self.device)
            setattr(self, name, buf)
            return buf.data_ptr()
        return ALLOC_FN(alloc)


class Rasterizer:
    """Static methods named after the reference's."""

    @staticmethod
    def forward(ws, P, D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales,
                scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                prefiltered, out_color, radii=None, debug=False, stream=None):
        """Rasterizer::forward (rasterizer.h:33-56).  Returns num_rendered."""
        cbs = (ws._make("geom"), ws._make("binning"), ws._make("image"))
        n = C.c_int(0)
        check(lib().gsb_rasterizer_forward(
            cbs[0], None, cbs[1], None, cbs[2], None, P, D, M, _ptr(background), width, height, _ptr(means3D),
            _ptr(shs), _ptr(colors_precomp), _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations),
            _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos), tan_fovx, tan_fovy,
            int(bool(prefiltered)), _ptr(out_color), _ptr(radii), int(bool(debug)), stream, C.byref(n)))
        return n.value

    @staticmethod
    def backward(ws, P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales, scale_modifier,
                 rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, dL_dpix,
                 dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
                 debug=False, stream=None):
        """Rasterizer::backward (rasterizer.h:58-83) on the workspaces of the matching forward."""
        check(lib().gsb_rasterizer_backward(
            P, D, M, R, _ptr(background), width, height, _ptr(means3D), _ptr(shs), _ptr(colors_precomp), _ptr(scales),
            scale_modifier, _ptr(rotations), _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos),
            tan_fovx, tan_fovy, _ptr(radii), _ptr(ws.geom), _ptr(ws.binning), _ptr(ws.image), _ptr(dL_dpix),
            _ptr(dL_dmean2D), _ptr(dL_dconic), _ptr(dL_dopacity), _ptr(dL_dcolor), _ptr(dL_dmean3D), _ptr(dL_dcov3D),
            _ptr(dL_dsh), _ptr(dL_dscale), _ptr(dL_drot), int(bool(debug)), stream))

    @staticmethod
    def markVisible(P, means3D, viewmatrix, projmatrix, present, stream=None):
        """Rasterizer::markVisible (rasterizer.h:24-29); `present` is a bool tensor [P]."""
        check(lib().gsb_rasterizer_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix), _ptr(present), stream))


def _view(ptr, shape, dtype, device):
    """Wrap a raw device pointer of the private workspace as a tensor (parity tests only)."""
    n = 1
    for s in shape:
        n *= s
    if ptr is None or n == 0:
        return torch.empty(shape, dtype=dtype, device=device)
    esz = torch.empty((), dtype=dtype).element_size()

    class _Holder:
        __cuda_array_interface__ = {"shape": (n * esz,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    raw = torch.as_tensor(_Holder(), device=device)
    return raw.view(dtype).reshape(shape).clone()


def state(ws, P, width, height, R):
    """Copies of the private per-view state, named like the reference's Geometry/Binning/ImageState
    fields (rasterizer_impl.h:30-62) so parity tests can diff them one to one."""
    sv = StateView()
    check(lib().gsb_rasterizer_state(P, width, height, R, _ptr(ws.geom), _ptr(ws.binning) if R > 0 else None,
                                     _ptr(ws.image), C.byref(sv)))
    dev = ws.geom.device
    tiles = ((width + 15) // 16) * ((height + 15) // 16)
    recs = _view(sv.recs, (P, 12), torch.float32, dev)
    out = {
        "radii": _view(sv.radii, (P,), torch.int32, dev),
        "depths": _view(sv.depths, (P,), torch.float32, dev),
        "tiles_touched": _view(sv.tiles_touched, (P,), torch.int32, dev),
        "means2D": recs[:, 0:2].contiguous(),
        "conic_opacity": torch.cat([recs[:, 2:5], recs[:, 5:6]], dim=1).contiguous(),
        "rgb": recs[:, 6:9].contiguous(),
        "cov3D": _view(sv.cov3D, (P, 6), torch.float32, dev),
        "clamped": _view(sv.clamped, (P,), torch.uint8, dev),
        "point_list": _view(sv.point_list, (R,), torch.int32, dev),
        "ranges": _view(sv.ranges, (tiles, 2), torch.int32, dev),
        "n_contrib": _view(sv.n_contrib, (height * width,), torch.int32, dev),
        "final_T": _view(sv.final_T, (height * width,), torch.float32, dev),
        "tile_max": _view(sv.tile_max, (tiles,), torch.int32, dev),
        "counters": _view(sv.counters, (3,), torch.int32, dev),
    }
    keys = torch.empty(max(R, 1), dtype=torch.int64, device=dev)
    if R > 0:
        check(lib().gsb_rasterizer_sorted_keys(P, width, height, R, _ptr(ws.geom), _ptr(ws.binning), _ptr(ws.image), _ptr(keys), None))
    out["point_list_keys"] = keys[:R]
    return out


def tile_order(ws, P, width, height):
    """CTA number -> tile of the one-CTA-per-tile grids after a forward pass (heaviest list-length class first)."""
    import numpy as np
    tiles = ((width + 15) // 16) * ((height + 15) // 16)
    out = np.zeros(tiles, dtype=np.uint32)
    check(lib().gsb_rasterizer_tile_order(P, width, height, _ptr(ws.geom), _ptr(ws.image), out.ctypes.data_as(C.c_void_p), None))
    return out
