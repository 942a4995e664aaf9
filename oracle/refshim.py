"""TEST INFRASTRUCTURE: ctypes wrapper around oracle/_ref/libgsref.so, the UNMODIFIED reference
rasterizer compiled from /root/reference for sm_100 (see oracle/Makefile, oracle/ref_shim.cu).

Only tests/, __graft_entry__.smoke() and bench.py's reference arm may import this module; the
product (gaussian-splatterer_b200) never does.
"""
import ctypes as C
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libgsref.so")

_lib = None


def available():
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{REF_LIB} missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(REF_LIB)
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.gsref_last_error.restype = C.c_char_p
        L.gsref_forward.argtypes = [i, i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, i, vp, vp, i]
        L.gsref_forward.restype = i
        L.gsref_backward.argtypes = [i, i, i, i, vp, i, i, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp, vp,
                                     vp, vp, vp, vp, vp, vp, vp, vp, vp, i]
        L.gsref_backward.restype = i
        L.gsref_mark_visible.argtypes = [i, vp, vp, vp, vp]
        L.gsref_mark_visible.restype = None
        L.gsref_release.restype = None
        L.gsref_state.argtypes = [i, i, i, i, C.POINTER(vp)]
        L.gsref_state.restype = None
        L.gsref_trainer_create.argtypes = [i, i]
        L.gsref_trainer_create.restype = vp
        L.gsref_trainer_destroy.argtypes = [vp]
        L.gsref_trainer_destroy.restype = None
        L.gsref_trainer_set_model.argtypes = [vp, i, i, i, i, vp, vp, vp, vp, vp]
        L.gsref_trainer_set_model.restype = None
        L.gsref_trainer_count.argtypes = [vp]
        L.gsref_trainer_count.restype = i
        L.gsref_trainer_get_model.argtypes = [vp, vp, vp, vp, vp, vp]
        L.gsref_trainer_get_model.restype = None
        L.gsref_trainer_buffers.argtypes = [vp, C.POINTER(vp)]
        L.gsref_trainer_buffers.restype = None
        L.gsref_trainer_set_truths.argtypes = [vp, i, C.POINTER(vp), C.POINTER(vp), vp, vp, vp, vp, vp]
        L.gsref_trainer_set_truths.restype = None
        L.gsref_trainer_rendered_total.argtypes = [vp]
        L.gsref_trainer_rendered_total.restype = C.c_longlong
        L.gsref_trainer_train.argtypes = [vp, vp, i, i]
        L.gsref_trainer_train.restype = i
        L.gsref_trainer_densify.argtypes = [vp, vp]
        L.gsref_trainer_densify.restype = i
        L.gsref_image_float_to_int.argtypes = [vp, vp, i, i]
        L.gsref_image_float_to_int.restype = None
        L.gsref_image_int_to_loss.argtypes = [vp, vp, vp, i, i]
        L.gsref_image_int_to_loss.restype = None
        _lib = L
    return _lib


REF_CAM_LIB = os.path.join(_HERE, "_ref", "libgsrefcam.so")


def ref_cameras(project):
    """Camera::getCameras of the reference's own src/Camera.cpp (oracle/_ref/libgsrefcam.so) for a Project-like object
    with sphere1 / sphere2 (count, distance, fovDeg, rotX, rotY).  Returns [(location xyz float32[3], fovDegY)]."""
    L = C.CDLL(REF_CAM_LIB)
    L.gsrefcam_get_cameras.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    L.gsrefcam_get_cameras.restype = C.c_int
    sp = [np.array([s.count, s.distance, s.fovDeg, s.rotX, s.rotY], np.float32) for s in (project.sphere1, project.sphere2)]
    cap = int(project.sphere1.count + project.sphere2.count) + 1
    out = np.zeros(4 * cap, np.float32)
    n = L.gsrefcam_get_cameras(sp[0].ctypes.data_as(C.c_void_p), sp[1].ctypes.data_as(C.c_void_p),
                               out.ctypes.data_as(C.c_void_p), cap)
    return [(out[4 * i:4 * i + 3].copy(), float(out[4 * i + 3])) for i in range(n)]


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _wrap(ptr, shape, dtype, device):
    n = int(np.prod(shape)) if len(shape) else 1
    if not ptr or n == 0:
        return torch.empty(shape, dtype=dtype, device=device)
    esz = torch.empty((), dtype=dtype).element_size()

    class _H:
        __cuda_array_interface__ = {"shape": (n * esz,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(_H(), device=device).view(dtype).reshape(shape).clone()


def forward(P, D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
            rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color,
            radii=None, debug=False, pooled=True):
    """CudaRasterizer::Rasterizer::forward of the reference build. Returns num_rendered."""
    r = lib().gsref_forward(int(pooled), P, D, M, _p(background), width, height, _p(means3D), _p(shs), _p(colors_precomp),
                            _p(opacities), _p(scales), scale_modifier, _p(rotations), _p(cov3D_precomp), _p(viewmatrix),
                            _p(projmatrix), _p(cam_pos), tan_fovx, tan_fovy, int(prefiltered), _p(out_color), _p(radii),
                            int(debug))
    if r < 0:
        raise RuntimeError(lib().gsref_last_error().decode())
    return r


def backward(P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
             cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii, dL_dpix, dL_dmean2D, dL_dconic,
             dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug=False):
    r = lib().gsref_backward(P, D, M, R, _p(background), width, height, _p(means3D), _p(shs), _p(colors_precomp),
                             _p(scales), scale_modifier, _p(rotations), _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix),
                             _p(campos), tan_fovx, tan_fovy, _p(radii), _p(dL_dpix), _p(dL_dmean2D), _p(dL_dconic),
                             _p(dL_dopacity), _p(dL_dcolor), _p(dL_dmean3D), _p(dL_dcov3D), _p(dL_dsh), _p(dL_dscale),
                             _p(dL_drot), int(debug))
    if r != 0:
        raise RuntimeError(lib().gsref_last_error().decode())


def mark_visible(P, means3D, viewmatrix, projmatrix, present):
    lib().gsref_mark_visible(P, _p(means3D), _p(viewmatrix), _p(projmatrix), _p(present))


def state(P, width, height, R, device):
    """Copies of the reference's Geometry/Binning/ImageState arrays after forward()."""
    ptrs = (C.c_void_p * 16)()
    lib().gsref_state(P, width, height, R, ptrs)
    N = width * height
    tiles = ((width + 15) // 16) * ((height + 15) // 16)
    t = torch
    out = {
        "depths": _wrap(ptrs[0], (P,), t.float32, device),
        "clamped": _wrap(ptrs[1], (P, 3), t.uint8, device),
        "radii": _wrap(ptrs[2], (P,), t.int32, device),
        "means2D": _wrap(ptrs[3], (P, 2), t.float32, device),
        "cov3D": _wrap(ptrs[4], (P, 6), t.float32, device),
        "conic_opacity": _wrap(ptrs[5], (P, 4), t.float32, device),
        "rgb": _wrap(ptrs[6], (P, 3), t.float32, device),
        "point_offsets": _wrap(ptrs[7], (P,), t.int32, device),
        "tiles_touched": _wrap(ptrs[8], (P,), t.int32, device),
        "point_list_keys_unsorted": _wrap(ptrs[9], (R,), t.int64, device),
        "point_list_keys": _wrap(ptrs[10], (R,), t.int64, device),
        "point_list_unsorted": _wrap(ptrs[11], (R,), t.int32, device),
        "point_list": _wrap(ptrs[12], (R,), t.int32, device),
        "ranges": _wrap(ptrs[13], (N, 2), t.int32, device)[:tiles].contiguous(),
        "n_contrib": _wrap(ptrs[14], (N,), t.int32, device),
        "final_T": _wrap(ptrs[15], (N,), t.float32, device),
    }
    return out


def release():
    lib().gsref_release()


class RefTrainer:
    """Replay of Trainer::train (src/Trainer.cu:252-435) around the reference rasterizer."""

    def __init__(self, width, height, device="cuda"):
        self.W, self.H = width, height
        self.device = torch.device(device)
        self._h = lib().gsref_trainer_create(width, height)
        self._keep = None

    def close(self):
        if self._h:
            lib().gsref_trainer_destroy(self._h)
            self._h = None

    def set_model(self, cloud, capacity, D, M):
        a = [np.ascontiguousarray(cloud[k], dtype=np.float32) for k in ("locations", "shs", "scales", "opacities", "rotations")]
        count = a[3].shape[0]
        lib().gsref_trainer_set_model(self._h, count, capacity, D, M, *[x.ctypes.data_as(C.c_void_p) for x in a])
        self.M = M

    def count(self):
        return lib().gsref_trainer_count(self._h)

    def get_model(self):
        c, M = self.count(), self.M
        out = {"locations": np.empty(3 * c, np.float32), "shs": np.empty(3 * M * c, np.float32),
               "scales": np.empty(3 * c, np.float32), "opacities": np.empty(c, np.float32),
               "rotations": np.empty(4 * c, np.float32)}
        lib().gsref_trainer_get_model(self._h, *[out[k].ctypes.data_as(C.c_void_p) for k in
                                                 ("locations", "shs", "scales", "opacities", "rotations")])
        return out

    def set_truths(self, truthW, truthB, consts):
        n = len(truthW)
        aw = (C.c_void_p * n)(*[t.data_ptr() for t in truthW])
        ab = (C.c_void_p * n)(*[t.data_ptr() for t in truthB])
        view = np.concatenate([c["view"] for c in consts]).astype(np.float32)
        pv = np.concatenate([c["projview"] for c in consts]).astype(np.float32)
        loc = np.concatenate([c["location"] for c in consts]).astype(np.float32)
        tx = np.array([c["tan_fovx"] for c in consts], np.float32)
        ty = np.array([c["tan_fovy"] for c in consts], np.float32)
        self._keep = (truthW, truthB)
        lib().gsref_trainer_set_truths(self._h, n, aw, ab, *[x.ctypes.data_as(C.c_void_p) for x in (view, pv, loc, tx, ty)])

    def train(self, params13, as_shipped=True, apply=True):
        p = np.ascontiguousarray(params13, dtype=np.float32)
        if lib().gsref_trainer_train(self._h, p.ctypes.data_as(C.c_void_p), int(as_shipped), int(apply)) != 0:
            raise RuntimeError(lib().gsref_last_error().decode())

    def densify(self, params13):
        p = np.ascontiguousarray(params13, dtype=np.float32)
        if lib().gsref_trainer_densify(self._h, p.ctypes.data_as(C.c_void_p)) != 0:
            raise RuntimeError(lib().gsref_last_error().decode())

    def rendered_total(self):
        return int(lib().gsref_trainer_rendered_total(self._h))

    def buffers(self):
        """var, avgLoc, avgSh, avgScale, avgOpacity, avgRot as tensors (copies)."""
        ptrs = (C.c_void_p * 8)()
        lib().gsref_trainer_buffers(self._h, ptrs)
        c, M = self.count(), self.M
        shapes = [(c,), (3 * c,), (3 * M * c,), (3 * c,), (c,), (4 * c,)]
        return [_wrap(ptrs[k], shapes[k], torch.float32, self.device) for k in range(6)]


def image_float_to_int(src, fb, w, h):
    lib().gsref_image_float_to_int(_p(src), _p(fb), w, h)


def image_int_to_loss(truth, rasterized, loss, w, h):
    lib().gsref_image_int_to_loss(_p(truth), _p(rasterized), _p(loss), w, h)
