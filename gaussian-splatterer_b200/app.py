"""ctypes mirror of include/gsb_app.h: the headless application layer.

Python here is only the test / bench harness over the C ABI; the logic (settings.json, cameras, .gobj, field
initialisers, PNG, the truth ray caster, the UiFrame::update scheduler) lives in C++ / CUDA inside libgsb200.so.
Names follow the reference: Project (src/Project.h), Camera (src/Camera.h), ModelSplatsHost (src/ModelSplatsHost.h),
UiFrame::saveSplats / loadSplats / saveSettings / loadSettings / initField* (src/ui/UiFrame.cpp).
"""
import ctypes as C

import numpy as np

from ._lib import CameraC, ProjectParams, check, lib


class CameraSphereC(C.Structure):
    _fields_ = [("count", C.c_int), ("distance", C.c_float), ("fovDeg", C.c_float), ("rotX", C.c_float), ("rotY", C.c_float)]


class ProjectData(C.Structure):
    """gsb_project_data: every non-string member of Project in declaration order."""
    _fields_ = [("sphere1", CameraSphereC), ("sphere2", CameraSphereC), ("rtSamples", C.c_int)] + \
        [(n, C.c_float) for n in ("lrLocation", "lrSh", "lrScale", "lrOpacity", "lrRotation", "paramScaleMax",
                                  "paramCullOpacity", "paramCullSize", "paramDensifyVariance", "paramSplitSize",
                                  "paramSplitDistance", "paramSplitScale", "paramCloneDistance")] + \
        [("iterations", C.c_int), ("intervalCapture", C.c_int), ("intervalDensify", C.c_int), ("previewTimer", C.c_float),
         ("previewRtSamples", C.c_int), ("previewSplatScale", C.c_float), ("previewTruth", C.c_int),
         ("previewTruthIndex", C.c_int), ("previewFreeOrbit", C.c_int), ("previewFreeOrbitSpeed", C.c_float),
         ("previewFreeDistance", C.c_float), ("previewFreeFovDeg", C.c_float), ("previewFreeRotX", C.c_float),
         ("previewFreeRotY", C.c_float), ("renderResX", C.c_int), ("renderResY", C.c_int)]


class CameraDesc(C.Structure):
    """gsb_camera_desc = Camera{location, target, fovDegY}."""
    _fields_ = [("location", C.c_float * 3), ("target", C.c_float * 3), ("fovDegY", C.c_float)]


class ModelHostView(C.Structure):
    _fields_ = [("capacity", C.c_int), ("shDegree", C.c_int), ("shCoeffs", C.c_int), ("count", C.c_int)] + \
        [(n, C.POINTER(C.c_float)) for n in ("locations", "shs", "scales", "opacities", "rotations")]


_bound = False


def L():
    """libgsb200.so with the gsb_app.h prototypes attached."""
    global _bound
    lb = lib()
    if _bound:
        return lb
    vp, i, f, sz, cp = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_char_p
    PD, CD = C.POINTER(ProjectData), C.POINTER(CameraDesc)
    FP = C.POINTER(C.c_float)
    sigs = {
        "gsb_project_create": ([C.POINTER(vp)], i), "gsb_project_init_new": ([vp], i), "gsb_project_destroy": ([vp], i),
        "gsb_project_fields": ([vp], PD), "gsb_project_get_string": ([vp, cp], cp), "gsb_project_set_string": ([vp, cp, cp], i),
        "gsb_project_to_json": ([vp, vp, sz, C.POINTER(sz)], i), "gsb_project_from_json": ([vp, cp], i),
        "gsb_project_save": ([vp, cp], i), "gsb_project_load": ([vp, cp], i),
        "gsb_project_train_params": ([PD, C.POINTER(ProjectParams)], i),
        "gsb_cameras_count": ([PD], i), "gsb_cameras_get": ([PD, CD, i], i), "gsb_camera_preview": ([PD, CD], i),
        "gsb_camera_constants": ([CD, i, i, i, C.POINTER(CameraC)], i), "gsb_camera_inverse_projview": ([CD, f, FP], i),
        "gsb_model_host_create": ([C.POINTER(vp), i, i, i], i),
        "gsb_model_host_from_arrays": ([C.POINTER(vp), FP, sz, FP, sz, FP, sz, FP, sz, FP, sz], i),
        "gsb_model_host_destroy": ([vp], i), "gsb_model_host_get": ([vp, C.POINTER(ModelHostView)], i),
        "gsb_model_host_push_back": ([vp, FP, FP, i, FP, f, FP], i), "gsb_model_host_copy": ([vp, i, i], i),
        "gsb_gobj_save": ([vp, cp], i), "gsb_gobj_load": ([C.POINTER(vp), cp], i),
        "gsb_field_grid": ([C.POINTER(vp)], i), "gsb_field_mono": ([C.POINTER(vp)], i), "gsb_field_model": ([C.POINTER(vp), cp], i),
        "gsb_png_write": ([cp, vp, i, i, i], i), "gsb_png_read": ([cp, C.POINTER(vp), C.POINTER(i), C.POINTER(i)], i),
        "gsb_free": ([vp], None),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(lb, name)
        fn.argtypes, fn.restype = args, res
    _bound = True
    return lb


APP_SYMBOLS_HOST = [
    "gsb_project_create", "gsb_project_init_new", "gsb_project_destroy", "gsb_project_fields", "gsb_project_get_string",
    "gsb_project_set_string", "gsb_project_to_json", "gsb_project_from_json", "gsb_project_save", "gsb_project_load",
    "gsb_project_train_params", "gsb_cameras_count", "gsb_cameras_get", "gsb_camera_preview", "gsb_camera_constants",
    "gsb_camera_inverse_projview", "gsb_model_host_create", "gsb_model_host_from_arrays", "gsb_model_host_destroy",
    "gsb_model_host_get", "gsb_model_host_push_back", "gsb_model_host_copy", "gsb_gobj_save", "gsb_gobj_load",
    "gsb_field_grid", "gsb_field_mono", "gsb_field_model", "gsb_png_write", "gsb_png_read", "gsb_free"]


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class Project:
    """Project (src/Project.h): numeric members through `.d` (live struct), the three strings as properties."""

    def __init__(self, new=False, _borrow=None):
        object.__setattr__(self, "_owned", _borrow is None)
        if _borrow is not None:                                 # a handle owned by an App
            object.__setattr__(self, "_h", C.c_void_p(_borrow))
            return
        self._h = C.c_void_p()
        check(L().gsb_project_create(C.byref(self._h)))
        if new:
            check(L().gsb_project_init_new(self._h))        # UiFrame::initProject

    def __del__(self):
        if getattr(self, "_h", None) and getattr(self, "_owned", False):
            L().gsb_project_destroy(self._h)
        object.__setattr__(self, "_h", None)

    @property
    def d(self):
        return L().gsb_project_fields(self._h).contents

    def __getattr__(self, name):
        if name in ("perspective", "pathModel", "pathTextureDiffuse"):
            return L().gsb_project_get_string(self._h, name.encode()).decode("utf-8")
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.d, name)

    def __setattr__(self, name, value):
        if name in ("_h", "_owned"):
            object.__setattr__(self, name, value)
        elif name in ("perspective", "pathModel", "pathTextureDiffuse"):
            check(L().gsb_project_set_string(self._h, name.encode(), value.encode("utf-8")))
        else:
            setattr(self.d, name, value)

    def dumps(self):
        need = C.c_size_t(0)
        check(L().gsb_project_to_json(self._h, None, 0, C.byref(need)))
        buf = C.create_string_buffer(need.value)
        check(L().gsb_project_to_json(self._h, buf, need.value, None))
        return buf.value.decode("utf-8")

    def loads(self, text):
        check(L().gsb_project_from_json(self._h, text.encode("utf-8")))

    def save(self, path):
        check(L().gsb_project_save(self._h, str(path).encode()))

    def load(self, path):
        check(L().gsb_project_load(self._h, str(path).encode()))

    def train_params(self):
        q = ProjectParams()
        check(L().gsb_project_train_params(C.byref(self.d), C.byref(q)))
        return q

    # ---- Camera statics (src/Camera.h:11-13)
    def cameras(self):
        n = L().gsb_cameras_count(C.byref(self.d))
        arr = (CameraDesc * max(n, 1))()
        check(L().gsb_cameras_get(C.byref(self.d), arr, n))
        return [arr[k] for k in range(n)]

    def preview_camera(self):
        c = CameraDesc()
        check(L().gsb_camera_preview(C.byref(self.d), C.byref(c)))
        return c


def camera_desc(location, target=(0.0, 0.0, 0.0), fov=60.0):
    c = CameraDesc()
    c.location[:] = [float(x) for x in location]
    c.target[:] = [float(x) for x in target]
    c.fovDegY = float(fov)
    return c


def camera_constants(desc, width, height, square_fov=False):
    out = CameraC()
    check(L().gsb_camera_constants(C.byref(desc), width, height, int(square_fov), C.byref(out)))
    return out


def camera_inverse_projview(desc, aspect):
    out = np.zeros(16, np.float32)
    check(L().gsb_camera_inverse_projview(C.byref(desc), float(aspect), _fp(out)))
    return out


class ModelSplatsHostC:
    """ModelSplatsHost held by the library (handle); arrays come back as numpy copies."""

    def __init__(self, handle):
        self._h = handle

    def __del__(self):
        if getattr(self, "_h", None):
            L().gsb_model_host_destroy(self._h)
            self._h = None

    @classmethod
    def create(cls, capacity, shDegree, shCoeffs):
        h = C.c_void_p()
        check(L().gsb_model_host_create(C.byref(h), capacity, shDegree, shCoeffs))
        return cls(h)

    @classmethod
    def from_arrays(cls, locations, shs, scales, opacities, rotations):
        a = [np.ascontiguousarray(x, dtype=np.float32).reshape(-1) for x in (locations, shs, scales, opacities, rotations)]
        h = C.c_void_p()
        args = []
        for x in a:
            args += [_fp(x), x.size]
        check(L().gsb_model_host_from_arrays(C.byref(h), *args))
        return cls(h)

    @classmethod
    def _make(cls, fn, *args):
        h = C.c_void_p()
        check(fn(C.byref(h), *args))
        return cls(h)

    @classmethod
    def load_gobj(cls, path):
        return cls._make(L().gsb_gobj_load, str(path).encode())

    @classmethod
    def field_grid(cls):
        return cls._make(L().gsb_field_grid)

    @classmethod
    def field_mono(cls):
        return cls._make(L().gsb_field_mono)

    @classmethod
    def field_model(cls, obj_path):
        return cls._make(L().gsb_field_model, str(obj_path).encode())

    def view(self):
        v = ModelHostView()
        check(L().gsb_model_host_get(self._h, C.byref(v)))
        return v

    def arrays(self):
        v = self.view()
        n, M = v.count, v.shCoeffs
        grab = lambda p, k: np.ctypeslib.as_array(p, shape=(k,)).copy() if k else np.zeros(0, np.float32)
        return {"locations": grab(v.locations, 3 * n), "shs": grab(v.shs, 3 * M * n), "scales": grab(v.scales, 3 * n),
                "opacities": grab(v.opacities, n), "rotations": grab(v.rotations, 4 * n)}

    def push_back(self, location, sh, scale, opacity, rotation):
        loc, s, sc, r = [np.ascontiguousarray(x, dtype=np.float32) for x in (location, sh, scale, rotation)]
        check(L().gsb_model_host_push_back(self._h, _fp(loc), _fp(s), s.size, _fp(sc), float(opacity), _fp(r)))

    def copy(self, to, frm):
        check(L().gsb_model_host_copy(self._h, to, frm))

    def save_gobj(self, path):
        check(L().gsb_gobj_save(self._h, str(path).encode()))


def png_write(path, rgba_u32, width, height, flip_vertically=True):
    a = np.ascontiguousarray(rgba_u32, dtype=np.uint32)
    assert a.size == width * height
    check(L().gsb_png_write(str(path).encode(), a.ctypes.data_as(C.c_void_p), width, height, int(flip_vertically)))


def png_read(path):
    p, w, h = C.c_void_p(), C.c_int(), C.c_int()
    check(L().gsb_png_read(str(path).encode(), C.byref(p), C.byref(w), C.byref(h)))
    try:
        buf = (C.c_uint8 * (w.value * h.value * 4)).from_address(p.value)
        return np.frombuffer(buf, dtype=np.uint8).reshape(h.value, w.value, 4).copy()
    finally:
        L().gsb_free(p)


# ------------------------------------------------------------------------------------------------ truth ray caster (f1)
APP_SYMBOLS_RTX = ["gsb_rtx_create", "gsb_rtx_destroy", "gsb_rtx_reset", "gsb_rtx_load_model", "gsb_rtx_set_mesh",
                   "gsb_rtx_load_texture_diffuse", "gsb_rtx_set_texture_diffuse", "gsb_rtx_model_present", "gsb_rtx_render",
                   "gsb_rtx_render_pair", "gsb_rtx_last_ray_count"]
_rtx_bound = False


def _LR():
    global _rtx_bound
    lb = L()
    if not _rtx_bound:
        vp, i, cp = C.c_void_p, C.c_int, C.c_char_p
        CD, FP = C.POINTER(CameraDesc), C.POINTER(C.c_float)
        sigs = {
            "gsb_rtx_create": ([C.POINTER(vp)], i), "gsb_rtx_destroy": ([vp], i), "gsb_rtx_reset": ([vp], i),
            "gsb_rtx_load_model": ([vp, cp], i), "gsb_rtx_set_mesh": ([vp, FP, i, C.POINTER(i), i, FP], i),
            "gsb_rtx_load_texture_diffuse": ([vp, cp], i), "gsb_rtx_set_texture_diffuse": ([vp, vp, i, i], i),
            "gsb_rtx_model_present": ([vp], i), "gsb_rtx_render": ([vp, vp, i, i, CD, FP, i, FP, i, vp], i),
            "gsb_rtx_render_pair": ([vp, vp, vp, i, i, CD, i, vp], i),
            "gsb_rtx_last_ray_count": ([vp, C.POINTER(C.c_ulonglong)], i),
        }
        for name, (args, res) in sigs.items():
            fn = getattr(lb, name)
            fn.argtypes, fn.restype = args, res
        _rtx_bound = True
    return lb


class RtxHost:
    """RtxHost (src/rtx/RtxHost.h:16-41): reset / loadModel / loadTextureDiffuse / render, plus in-memory setters."""

    def __init__(self, _borrow=None):
        self._owned = _borrow is None
        if _borrow is not None:                                 # a handle owned by an App
            self._h = C.c_void_p(_borrow)
            return
        self._h = C.c_void_p()
        check(_LR().gsb_rtx_create(C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._owned:
            _LR().gsb_rtx_destroy(self._h)
        self._h = None

    __del__ = close

    def reset(self):
        check(_LR().gsb_rtx_reset(self._h))

    def loadModel(self, path):
        check(_LR().gsb_rtx_load_model(self._h, str(path).encode()))

    def loadTextureDiffuse(self, path):
        check(_LR().gsb_rtx_load_texture_diffuse(self._h, str(path).encode()))

    def set_mesh(self, vertices, triangles, texcoords=None):
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1)
        t = np.ascontiguousarray(triangles, dtype=np.int32).reshape(-1)
        tc = None if texcoords is None else np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1)
        check(_LR().gsb_rtx_set_mesh(self._h, _fp(v), v.size // 3, t.ctypes.data_as(C.POINTER(C.c_int)), t.size // 3,
                                     None if tc is None else _fp(tc)))

    def set_texture(self, rgba):
        a = np.ascontiguousarray(rgba, dtype=np.uint8)
        check(_LR().gsb_rtx_set_texture_diffuse(self._h, a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0]))

    @property
    def modelPresent(self):
        return bool(_LR().gsb_rtx_model_present(self._h))

    def render(self, frameBuffer, size, camera, background, samples, cameras=(), stream=None):
        """RtxHost::render(frameBuffer, size, camera, background, samples, cameras): frameBuffer is an int32/uint32 CUDA
        tensor of size.x * size.y elements; `cameras` are CameraDesc whose locations become the indicator orbs."""
        bg = np.ascontiguousarray(background, dtype=np.float32)
        orbs = np.array([list(c.location) for c in cameras], np.float32).reshape(-1) if len(cameras) else None
        check(_LR().gsb_rtx_render(self._h, C.c_void_p(frameBuffer.data_ptr()), size[0], size[1], C.byref(camera), _fp(bg),
                                   int(samples), None if orbs is None else _fp(orbs), len(cameras), stream))

    def render_pair(self, fbW, fbB, size, camera, samples, stream=None):
        check(_LR().gsb_rtx_render_pair(self._h, C.c_void_p(fbW.data_ptr()), C.c_void_p(fbB.data_ptr()), size[0], size[1],
                                        C.byref(camera), int(samples), stream))

    def last_ray_count(self):
        n = C.c_ulonglong(0)
        check(_LR().gsb_rtx_last_ray_count(self._h, C.byref(n)))
        return n.value


# --------------------------------------------------------------------------------------- headless UiFrame (f2, f4)
APP_SYMBOLS_DRIVER = [
    "gsb_app_create", "gsb_app_destroy", "gsb_app_project", "gsb_app_rtx", "gsb_app_trainer", "gsb_app_new_project",
    "gsb_app_init_field", "gsb_app_set_model", "gsb_app_get_model", "gsb_app_model_count", "gsb_app_model_device",
    "gsb_app_load_input", "gsb_app_random_rotate", "gsb_app_capture_truths", "gsb_app_truth_count", "gsb_app_truth_frames",
    "gsb_app_set_truths", "gsb_app_train", "gsb_app_set_auto_training", "gsb_app_update", "gsb_app_save_settings",
    "gsb_app_load_settings", "gsb_app_save_splats", "gsb_app_load_splats", "gsb_app_render_splats",
    "gsb_app_render_splats_png", "gsb_app_render_rtx_png"]
_app_bound = False


def _LA():
    global _app_bound
    lb = _LR()
    if not _app_bound:
        vp, i, f, cp = C.c_void_p, C.c_int, C.c_float, C.c_char_p
        CD, IP = C.POINTER(CameraDesc), C.POINTER(C.c_int)
        sigs = {
            "gsb_app_create": ([C.POINTER(vp), i, i], i), "gsb_app_destroy": ([vp], i), "gsb_app_project": ([vp], vp),
            "gsb_app_rtx": ([vp], vp), "gsb_app_trainer": ([vp], vp), "gsb_app_new_project": ([vp], i),
            "gsb_app_init_field": ([vp, i], i), "gsb_app_set_model": ([vp, vp], i), "gsb_app_get_model": ([vp, C.POINTER(vp)], i),
            "gsb_app_model_count": ([vp], i), "gsb_app_model_device": ([vp, IP, IP, IP, IP, C.POINTER(vp)], i),
            "gsb_app_load_input": ([vp], i), "gsb_app_random_rotate": ([vp], i), "gsb_app_capture_truths": ([vp], i),
            "gsb_app_truth_count": ([vp], i), "gsb_app_truth_frames": ([vp, i, C.POINTER(vp), C.POINTER(vp), CD], i),
            "gsb_app_set_truths": ([vp, i, C.POINTER(vp), C.POINTER(vp), CD], i), "gsb_app_train": ([vp, i], i),
            "gsb_app_set_auto_training": ([vp, i], i), "gsb_app_update": ([vp, f, IP], i),
            "gsb_app_save_settings": ([vp, cp], i), "gsb_app_load_settings": ([vp, cp], i), "gsb_app_save_splats": ([vp, cp], i),
            "gsb_app_load_splats": ([vp, cp], i), "gsb_app_render_splats": ([vp, vp, i, i, f, CD], i),
            "gsb_app_render_splats_png": ([vp, cp], i), "gsb_app_render_rtx_png": ([vp, cp], i),
        }
        for name, (args, res) in sigs.items():
            fn = getattr(lb, name)
            fn.argtypes, fn.restype = args, res
        _app_bound = True
    return lb


class App:
    """The headless UiFrame (src/ui/UiFrame.{h,cpp}): project, trainer + model, rtx, update()."""
    RAN_TRAINED, RAN_CAPTURED, RAN_DENSIFIED = 1, 2, 4

    def __init__(self, train_width=1024, train_height=1024):
        self._h = C.c_void_p()
        check(_LA().gsb_app_create(C.byref(self._h), train_width, train_height))
        self.size = (train_width, train_height)
        self.project = Project(_borrow=_LA().gsb_app_project(self._h))
        self.rtx = RtxHost(_borrow=_LA().gsb_app_rtx(self._h))

    def close(self):
        if getattr(self, "_h", None):
            _LA().gsb_app_destroy(self._h)
            self._h = None

    __del__ = close

    def _call(self, name, *args):
        check(getattr(_LA(), name)(self._h, *args))

    def new_project(self): self._call("gsb_app_new_project")
    def initFieldGrid(self): self._call("gsb_app_init_field", 0)
    def initFieldMono(self): self._call("gsb_app_init_field", 1)
    def initFieldModel(self): self._call("gsb_app_init_field", 2)
    def load_input(self): self._call("gsb_app_load_input")
    def random_rotate(self): self._call("gsb_app_random_rotate")
    def captureTruths(self): self._call("gsb_app_capture_truths")
    def train(self, densify=False): self._call("gsb_app_train", int(bool(densify)))
    def set_auto_training(self, on): self._call("gsb_app_set_auto_training", int(bool(on)))
    def saveSettings(self, path): self._call("gsb_app_save_settings", str(path).encode())
    def loadSettings(self, path): self._call("gsb_app_load_settings", str(path).encode())
    def saveSplats(self, path): self._call("gsb_app_save_splats", str(path).encode())
    def loadSplats(self, path): self._call("gsb_app_load_splats", str(path).encode())
    def render_splats_png(self, path): self._call("gsb_app_render_splats_png", str(path).encode())
    def render_rtx_png(self, path): self._call("gsb_app_render_rtx_png", str(path).encode())

    def update(self, delta_seconds):
        ran = C.c_int(0)
        self._call("gsb_app_update", float(delta_seconds), C.byref(ran))
        return ran.value

    @property
    def model_count(self):
        return _LA().gsb_app_model_count(self._h)

    def set_model(self, host):
        self._call("gsb_app_set_model", host._h)

    def get_model(self):
        h = C.c_void_p()
        self._call("gsb_app_get_model", C.byref(h))
        return ModelSplatsHostC(h)

    def model_device(self):
        cnt, cap, deg, M = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        ptrs = (C.c_void_p * 5)()
        self._call("gsb_app_model_device", C.byref(cnt), C.byref(cap), C.byref(deg), C.byref(M), ptrs)
        return {"count": cnt.value, "capacity": cap.value, "shDegree": deg.value, "shCoeffs": M.value, "ptrs": list(ptrs)}

    @property
    def truth_count(self):
        return _LA().gsb_app_truth_count(self._h)

    def truth_frames(self, index):
        """(white, black) as torch int32 CUDA tensors (copies) and the camera of truth frame `index`."""
        from .rasterizer import _view
        import torch
        w, b, cam = C.c_void_p(), C.c_void_p(), CameraDesc()
        self._call("gsb_app_truth_frames", index, C.byref(w), C.byref(b), C.byref(cam))
        n = self.size[0] * self.size[1]
        return _view(w.value, (n,), torch.int32, "cuda"), _view(b.value, (n,), torch.int32, "cuda"), cam

    def set_truths(self, whites, blacks, cameras):
        n = len(whites)
        aw = (C.c_void_p * n)(*[t.data_ptr() for t in whites])
        ab = (C.c_void_p * n)(*[t.data_ptr() for t in blacks])
        cams = (CameraDesc * n)(*cameras)
        self._call("gsb_app_set_truths", n, aw, ab, cams)

    def render_splats(self, frameBuffer, width, height, splat_scale, camera):
        self._call("gsb_app_render_splats", C.c_void_p(frameBuffer.data_ptr()), width, height, float(splat_scale), C.byref(camera))
