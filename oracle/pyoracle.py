"""TEST INFRASTRUCTURE: ctypes wrapper of oracle/liboracle.so (oracle.cpp, the scalar CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "liboracle.so")
_lib = None

FIELDS = {  # name: (field id, dtype, width per element)
    "radii": (0, np.int32, 1), "depths": (1, np.float32, 1), "tiles_touched": (2, np.uint32, 1),
    "means2D": (3, np.float32, 2), "cov3D": (4, np.float32, 6), "conic_opacity": (5, np.float32, 4),
    "rgb": (6, np.float32, 3), "clamped": (7, np.uint8, 3), "point_offsets": (8, np.uint32, 1),
}


def build():
    subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.oracle_forward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp]
        L.oracle_forward.restype = vp
        L.oracle_view_free.argtypes = [vp]
        L.oracle_view_num_rendered.argtypes = [vp]
        L.oracle_view_num_rendered.restype = i
        L.oracle_view_field.argtypes = [vp, i]
        L.oracle_view_field.restype = vp
        L.oracle_view_pair_counts.argtypes = [vp, vp]
        L.oracle_backward.argtypes = [vp] * 11
        L.oracle_mark_visible.argtypes = [i, vp, vp, vp]
        L.oracle_image_float_to_int.argtypes = [vp, vp, i, i]
        L.oracle_image_int_to_loss.argtypes = [vp, vp, vp, i, i]
        L.oracle_accumulate.argtypes = [vp] * 11 + [f, i, i]
        L.oracle_apply.argtypes = [vp] * 10 + [f, f, f, f, f, f, i, i]
        L.oracle_densify.argtypes = [i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        L.oracle_densify.restype = i
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class View:
    """Result of a forward: image + the reference's Geometry / Binning / Image state, as numpy arrays."""

    def __init__(self, handle, P, W, H, keep):
        self._h, self.P, self.W, self.H, self._keep = handle, P, W, H, keep
        L = lib()
        self.R = L.oracle_view_num_rendered(handle)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)

        def grab(fid, dtype, n):
            ptr = L.oracle_view_field(handle, fid)
            if n == 0 or not ptr:
                return np.zeros(0, dtype)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), shape=(n,)).copy()
        for name, (fid, dt, w) in FIELDS.items():
            arr = grab(fid, dt, P * w)
            setattr(self, name, arr.reshape(P, w) if w > 1 else arr)
        self.point_list_keys = grab(9, np.uint64, self.R)
        self.point_list = grab(10, np.uint32, self.R)
        self.ranges = grab(11, np.uint32, 2 * tiles).reshape(tiles, 2)
        self.n_contrib = grab(12, np.uint32, W * H)
        self.final_T = grab(13, np.float32, W * H)
        pc = np.zeros(3, np.int64)
        L.oracle_view_pair_counts(handle, _p(pc))
        self.pair_counts = {"evaluated": int(pc[0]), "exp": int(pc[1]), "blended": int(pc[2])}

    def backward(self, dL_dpix, M):
        P = self.P
        g = {"dL_dmean2D": np.zeros(3 * P, np.float32), "dL_dconic": np.zeros(4 * P, np.float32),
             "dL_dopacity": np.zeros(P, np.float32), "dL_dcolor": np.zeros(3 * P, np.float32),
             "dL_dmean3D": np.zeros(3 * P, np.float32), "dL_dcov3D": np.zeros(6 * P, np.float32),
             "dL_dsh": np.zeros(3 * M * P, np.float32), "dL_dscale": np.zeros(3 * P, np.float32),
             "dL_drot": np.zeros(4 * P, np.float32)}
        d = _f32(dL_dpix)
        lib().oracle_backward(self._h, _p(d), *[_p(g[k]) for k in ("dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolor",
                                                                  "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot")])
        return g

    def close(self):
        if self._h:
            lib().oracle_view_free(self._h)
            self._h = None

    def __del__(self):
        self.close()


def forward(P, D, M, background, width, height, means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations,
            cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy):
    """Returns (out_color[3*H*W], View)."""
    a = [_f32(x) for x in (background, means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp, viewmatrix,
                           projmatrix, cam_pos)]
    bg, means, shs_, cp, op, sc, rot, c3, vm, pm, cam = a
    out = np.zeros(3 * width * height, np.float32)
    h = lib().oracle_forward(P, D, M, _p(bg), width, height, _p(means), _p(shs_), _p(cp), _p(op), _p(sc), float(scale_modifier),
                             _p(rot), _p(c3), _p(vm), _p(pm), _p(cam), float(tan_fovx), float(tan_fovy), _p(out))
    return out, View(h, P, width, height, a)


def mark_visible(P, means3D, viewmatrix):
    out = np.zeros(P, np.uint8)
    m, v = _f32(means3D), _f32(viewmatrix)
    lib().oracle_mark_visible(P, _p(m), _p(v), _p(out))
    return out.astype(bool)


def image_float_to_int(src, w, h):
    fb = np.zeros(w * h, np.uint32)
    s = _f32(src)
    lib().oracle_image_float_to_int(_p(s), _p(fb), w, h)
    return fb


def image_int_to_loss(truth, rasterized, w, h):
    loss = np.zeros(3 * w * h, np.float32)
    t, r = np.ascontiguousarray(truth, np.uint32), _f32(rasterized)
    lib().oracle_image_int_to_loss(_p(t), _p(r), _p(loss), w, h)
    return loss


def accumulate(avg, grads, samples, M, n):
    """avg = [var, loc, sh, scale, op, rot] (updated in place), grads = [loc, sh, scale, op, rot]."""
    lib().oracle_accumulate(*[_p(x) for x in avg], *[_p(x) for x in grads], float(samples), M, n)


def apply(params, avg, lrs, max_scale, M, count):
    """params = [loc, sh, scale, op, rot] in place; avg = matching mean gradients; lrs = 5 floats."""
    lib().oracle_apply(*[_p(x) for x in params], *[_p(x) for x in avg], *[float(x) for x in lrs], float(max_scale), M, count)


def densify(cloud, count, capacity, M, var, avg_loc, params13):
    """Returns a new cloud dict (canonical order) after the reference's densify rules."""
    bufs = {}
    for k, w in (("locations", 3), ("shs", 3 * M), ("scales", 3), ("opacities", 1), ("rotations", 4)):
        b = np.zeros(capacity * w, np.float32)
        b[:count * w] = cloud[k][:count * w]
        bufs[k] = b
    v, g, p = _f32(var), _f32(avg_loc), _f32(params13)
    n = lib().oracle_densify(count, capacity, M, _p(bufs["locations"]), _p(bufs["shs"]), _p(bufs["scales"]), _p(bufs["opacities"]),
                             _p(bufs["rotations"]), _p(v), _p(g), _p(p))
    return {k: bufs[k][:n * w].copy() for k, w in (("locations", 3), ("shs", 3 * M), ("scales", 3), ("opacities", 1), ("rotations", 4))}


def rtx_render(W, H, cam_location, inv_projview, bg, samples, vertices, triangles, texcoords, tex_rgba, rect=None,
               orbs=None, frame=None):
    """oracle_rtx.cpp: one frame of the truth ray tracer (RtxHost::render) on the CPU, brute-force intersection.
    rect = (x0, y0, x1, y1) limits the work to a pixel rectangle.  Returns (frame uint32[H, W], ray segments)."""
    L = lib()
    L.oracle_rtx_render.restype = C.c_ulonglong
    L.oracle_rtx_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_int, C.c_void_p, C.c_int]
    v = _f32(np.asarray(vertices).reshape(-1))
    t = np.ascontiguousarray(np.asarray(triangles).reshape(-1), dtype=np.int32)
    nt = t.size // 3
    tc = _f32(np.zeros(6 * nt) if texcoords is None else np.asarray(texcoords).reshape(-1))
    tex = np.ascontiguousarray(tex_rgba, dtype=np.uint8)
    th, tw = tex.shape[:2]
    if frame is None:
        frame = np.zeros((H, W), np.uint32)
    x0, y0, x1, y1 = rect or (0, 0, W, H)
    o = None if orbs is None else _f32(np.asarray(orbs).reshape(-1))
    rays = L.oracle_rtx_render(_p(frame), W, H, x0, y0, x1, y1, _p(_f32(cam_location)), _p(_f32(inv_projview)), _p(_f32(bg)),
                               samples, _p(v), _p(t), nt, _p(tc), _p(tex), tw, th, _p(o), 0 if o is None else o.size // 3)
    return frame, int(rays)
