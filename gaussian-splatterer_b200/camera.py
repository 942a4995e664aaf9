"""Camera maths and the Project parameter block, mirroring reference src/Camera.{h,cpp} and
src/Project.h so both implementations can be fed identical matrices.

All arithmetic is float32 in the operation order of glm 0.9.9.9 (lookAtRH, perspectiveRH_NO,
mat4*mat4, angleAxis -> mat4), matrices are column-major flat arrays of 16 floats like
glm::value_ptr.
"""
import math
from dataclasses import dataclass, field

import numpy as np

f32 = np.float32


@dataclass
class CameraSphere:
    """Project::CameraSphere (src/Project.h:14-22)."""
    count: int = 16
    distance: float = 10.0
    fovDeg: float = 60.0
    rotX: float = 0.0
    rotY: float = 0.0


@dataclass
class Project:
    """The fields of src/Project.h:6-62 the hot path reads (GUI / preview fields omitted)."""
    sphere1: CameraSphere = field(default_factory=CameraSphere)
    sphere2: CameraSphere = field(default_factory=lambda: CameraSphere(count=0, fovDeg=30.0))
    lrLocation: float = 0.00005
    lrSh: float = 0.0001
    lrScale: float = 0.00002
    lrOpacity: float = 0.0001
    lrRotation: float = 0.000025
    paramScaleMax: float = 0.3
    paramCullOpacity: float = 0.005
    paramCullSize: float = 0.004
    paramDensifyVariance: float = 2.0
    paramSplitSize: float = 0.04
    paramSplitDistance: float = 1.5
    paramSplitScale: float = 0.8
    paramCloneDistance: float = 1.6
    iterations: int = 0
    intervalCapture: int = 50
    intervalDensify: int = 200

    PARAM_ORDER = ("lrLocation", "lrSh", "lrScale", "lrOpacity", "lrRotation", "paramScaleMax", "paramCullOpacity",
                   "paramCullSize", "paramDensifyVariance", "paramSplitSize", "paramSplitDistance", "paramSplitScale",
                   "paramCloneDistance")

    def params13(self):
        return np.array([getattr(self, n) for n in self.PARAM_ORDER], dtype=f32)


def _normalize(v):
    # glm::normalize = v * inversesqrt(dot(v, v))
    return (v * (f32(1.0) / np.sqrt(np.dot(v, v), dtype=f32))).astype(f32)


def _cross(a, b):
    return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], dtype=f32)


def look_at_rh(eye, center, up):
    """glm::lookAtRH (ext/matrix_transform.inl). Returns 4x4 with m[col][row]."""
    f = _normalize((center - eye).astype(f32))
    s = _normalize(_cross(f, up))
    u = _cross(s, f)
    m = np.zeros((4, 4), dtype=f32)
    m[0][0], m[1][0], m[2][0] = s
    m[0][1], m[1][1], m[2][1] = u
    m[0][2], m[1][2], m[2][2] = -f
    m[3][0] = -np.dot(s, eye)
    m[3][1] = -np.dot(u, eye)
    m[3][2] = np.dot(f, eye)
    m[3][3] = 1.0
    return m


def perspective_rh_no(fovy, aspect, z_near, z_far):
    """glm::perspectiveRH_NO (ext/matrix_clip_space.inl)."""
    t = f32(math.tan(f32(fovy) / f32(2.0)))
    m = np.zeros((4, 4), dtype=f32)
    m[0][0] = f32(1.0) / (f32(aspect) * t)
    m[1][1] = f32(1.0) / t
    m[2][2] = -(f32(z_far) + f32(z_near)) / (f32(z_far) - f32(z_near))
    m[2][3] = -1.0
    m[3][2] = -(f32(2.0) * f32(z_far) * f32(z_near)) / (f32(z_far) - f32(z_near))
    return m


def mat_mul(a, b):
    """glm mat4 * mat4 with m[col][row]: result[c] = sum_k a[k] * b[c][k]."""
    r = np.zeros((4, 4), dtype=f32)
    for c in range(4):
        r[c] = a[0] * b[c][0] + a[1] * b[c][1] + a[2] * b[c][2] + a[3] * b[c][3]
    return r.astype(f32)


def angle_axis_mat(angle_rad, axis):
    """(glm::mat4) glm::angleAxis(angle, axis): quaternion -> rotation matrix, m[col][row]."""
    a = f32(angle_rad)
    s = f32(math.sin(a * f32(0.5)))
    w = f32(math.cos(a * f32(0.5)))
    x, y, z = (np.asarray(axis, dtype=f32) * s).astype(f32)
    m = np.zeros((4, 4), dtype=f32)
    m[0][0] = 1 - 2 * (y * y + z * z); m[0][1] = 2 * (x * y + w * z); m[0][2] = 2 * (x * z - w * y)
    m[1][0] = 2 * (x * y - w * z); m[1][1] = 1 - 2 * (x * x + z * z); m[1][2] = 2 * (y * z + w * x)
    m[2][0] = 2 * (x * z + w * y); m[2][1] = 2 * (y * z - w * x); m[2][2] = 1 - 2 * (x * x + y * y)
    m[3][3] = 1
    return m


def fibonacci_sphere(count, distance):
    """getFibonacciSphere (src/Camera.cpp:9-27)."""
    golden = (f32(1.0) + f32(math.sqrt(5.0))) / f32(2.0)
    step = f32(2.0) * f32(math.pi) * golden
    out = []
    for i in range(count):
        t = f32(i) / f32(count)
        a1 = f32(math.acos(f32(1.0) - f32(2.0) * t))
        a2 = f32(step * f32(i))
        # sinf / cosf results and every product rounded to float32, like the reference's float expressions
        s1, c1, s2, c2 = f32(math.sin(a1)), f32(math.cos(a1)), f32(math.sin(a2)), f32(math.cos(a2))
        d = f32(distance)
        out.append(np.array([f32(s1 * c2) * d, f32(s1 * s2) * d, c1 * d], dtype=f32))
    return out


class Camera:
    """Camera (src/Camera.h:8-26): location, target, fovDegY; getView() = -lookAt, getProjection(aspect)."""

    def __init__(self, location, target, fovDegY):
        self.location = np.asarray(location, dtype=f32)
        self.target = np.asarray(target, dtype=f32)
        self.fovDegY = float(fovDegY)

    def getView(self):
        return (-look_at_rh(self.location, self.target, np.array([0, 1, 0], dtype=f32))).astype(f32)

    def getProjection(self, aspect):
        return perspective_rh_no(math.radians(self.fovDegY), aspect, 0.1, 100.0)

    @staticmethod
    def getCamerasCount(project):
        return project.sphere1.count + project.sphere2.count

    @staticmethod
    def getCameras(project):
        """Camera::getCameras (src/Camera.cpp:33-58): two Fibonacci spheres with a rotation offset.

        Delegates to the C++ implementation behind gsb_cameras_get (csrc/app/app_camera.cpp), which runs the reference's
        float expressions with the same libm calls (acosf / sinf / cosf) and is pinned by tests/golden/cameras.json; a
        double-precision restatement in Python differs from it by an ulp for some camera counts."""
        import ctypes as C
        from . import app
        d = app.ProjectData()
        for name in ("sphere1", "sphere2"):
            src, dst = getattr(project, name), getattr(d, name)
            dst.count, dst.distance, dst.fovDeg, dst.rotX, dst.rotY = src.count, src.distance, src.fovDeg, src.rotX, src.rotY
        n = app.L().gsb_cameras_count(C.byref(d))
        arr = (app.CameraDesc * max(n, 1))()
        app.check(app.L().gsb_cameras_get(C.byref(d), arr, n))
        return [Camera(np.array(arr[k].location[:], dtype=f32), (0.0, 0.0, 0.0), arr[k].fovDegY) for k in range(n)]

    def view_constants(self, width, height, square_fov=False):
        """What Trainer hands to the rasterizer for this camera at width x height.

        square_fov=True is Trainer::train's convention (tan_fovx = tan_fovy, src/Trainer.cu:355-356);
        otherwise Trainer::render's aspect-aware one (src/Trainer.cu:155,196-197).
        """
        view = self.getView()
        projview = mat_mul(self.getProjection(f32(width) / f32(height)), view)
        tany = math.tan(math.radians(self.fovDegY) * 0.5)
        tanx = tany if square_fov else math.tan(math.radians(width * self.fovDegY / height) * 0.5)
        return {"view": view.reshape(16).copy(), "projview": projview.reshape(16).copy(),
                "location": self.location.copy(), "tan_fovx": float(f32(tanx)), "tan_fovy": float(f32(tany))}
