"""Python mirror of the reference's Trainer / ModelSplatsDevice / ModelSplatsHost interface
(src/Trainer.cuh:49-73, src/ModelSplatsDevice.h:5-30, src/ModelSplatsHost.h:8-39) over the C ABI.

The C++ mirror of the same classes lives in include/compat/; this one exists so tests and
bench.py read like the reference's call sites.  torch tensors are the device memory.
"""
import ctypes as C

import numpy as np
import torch

from ._lib import ALLREDUCE_FN, CameraC, ProjectParams, TrainerStats, check, lib
from .camera import Camera, Project


class ModelSplatsHost:
    """Host splat container: arrays sized to `capacity`, `count` used (src/ModelSplatsHost.h:11-37)."""

    def __init__(self, capacity, shDegree, shCoeffs):
        self.capacity, self.shDegree, self.shCoeffs = int(capacity), int(shDegree), int(shCoeffs)
        self.count = 0
        self.locations = np.zeros(self.capacity * 3, np.float32)
        self.shs = np.zeros(self.capacity * 3 * self.shCoeffs, np.float32)
        self.scales = np.zeros(self.capacity * 3, np.float32)
        self.opacities = np.zeros(self.capacity, np.float32)
        self.rotations = np.zeros(self.capacity * 4, np.float32)

    @classmethod
    def from_arrays(cls, locations, shs, scales, opacities, rotations):
        """ModelSplatsHost(vector...) (src/ModelSplatsHost.cpp:28-57): capacity 1e6 * 10^k."""
        count = len(locations) // 3
        capacity = 1000000
        while capacity < count:
            capacity *= 10
        sh_coeffs = len(shs) // (3 * count)
        if (len(locations) != count * 3 or len(shs) != count * 3 * sh_coeffs or len(scales) != count * 3
                or len(opacities) != count or len(rotations) != count * 4):
            raise RuntimeError("Inconsistent feature dimensions supplied when creating a host model!")
        h = cls.__new__(cls)
        h.capacity, h.shDegree, h.shCoeffs, h.count = capacity, (sh_coeffs - 1) // 3, sh_coeffs, count
        # storage is only materialised up to count here (the reference news `capacity` floats)
        h.locations = np.asarray(locations, np.float32).copy()
        h.shs = np.asarray(shs, np.float32).copy()
        h.scales = np.asarray(scales, np.float32).copy()
        h.opacities = np.asarray(opacities, np.float32).copy()
        h.rotations = np.asarray(rotations, np.float32).copy()
        return h

    @classmethod
    def from_device(cls, device_model):
        h = cls.__new__(cls)
        h.capacity, h.shDegree, h.shCoeffs = device_model.capacity, device_model.shDegree, device_model.shCoeffs
        h.count = device_model.count
        h.locations = device_model.devLocations.cpu().numpy().reshape(-1)
        h.shs = device_model.devShs.cpu().numpy().reshape(-1)
        h.scales = device_model.devScales.cpu().numpy().reshape(-1)
        h.opacities = device_model.devOpacities.cpu().numpy().reshape(-1)
        h.rotations = device_model.devRotations.cpu().numpy().reshape(-1)
        return h

    def pushBack(self, location, sh, scale, opacity, rotation):
        if self.count >= self.capacity:
            raise RuntimeError("Model ran out of capacity!")
        c, m = self.count, self.shCoeffs
        self.locations[3 * c:3 * c + 3] = location
        self.shs[3 * m * c:3 * m * (c + 1)] = sh
        self.scales[3 * c:3 * c + 3] = scale
        self.opacities[c] = opacity
        self.rotations[4 * c:4 * c + 4] = rotation   # (w, x, y, z)
        self.count += 1

    def copy(self, indexTo, indexFrom):
        if indexTo < 0 or indexTo >= self.count or indexFrom < 0 or indexFrom >= self.count:
            raise RuntimeError("Can't copy splat in model, incorrect bounds and/or no capacity!")
        m = self.shCoeffs
        self.locations[3 * indexTo:3 * indexTo + 3] = self.locations[3 * indexFrom:3 * indexFrom + 3]
        self.shs[3 * m * indexTo:3 * m * (indexTo + 1)] = self.shs[3 * m * indexFrom:3 * m * (indexFrom + 1)]
        self.scales[3 * indexTo:3 * indexTo + 3] = self.scales[3 * indexFrom:3 * indexFrom + 3]
        self.opacities[indexTo] = self.opacities[indexFrom]
        self.rotations[4 * indexTo:4 * indexTo + 4] = self.rotations[4 * indexFrom:4 * indexFrom + 4]


class ModelSplatsDevice:
    """Five device arrays sized to `count` (src/ModelSplatsDevice.h:8-19)."""

    def __init__(self, host=None, device="cuda"):
        if host is None:
            return
        self.capacity, self.shDegree, self.shCoeffs, self.count = host.capacity, host.shDegree, host.shCoeffs, host.count
        c, m = host.count, host.shCoeffs
        dev = torch.device(device)
        self.devLocations = torch.from_numpy(np.ascontiguousarray(host.locations[:3 * c])).to(dev)
        self.devShs = torch.from_numpy(np.ascontiguousarray(host.shs[:3 * m * c])).to(dev)
        self.devScales = torch.from_numpy(np.ascontiguousarray(host.scales[:3 * c])).to(dev)
        self.devOpacities = torch.from_numpy(np.ascontiguousarray(host.opacities[:c])).to(dev)
        self.devRotations = torch.from_numpy(np.ascontiguousarray(host.rotations[:4 * c])).to(dev)

    @classmethod
    def empty(cls, count, capacity, shDegree, shCoeffs, device):
        d = cls()
        d.capacity, d.shDegree, d.shCoeffs, d.count = capacity, shDegree, shCoeffs, count
        d.devLocations = torch.empty(count * 3, dtype=torch.float32, device=device)
        d.devShs = torch.empty(count * 3 * shCoeffs, dtype=torch.float32, device=device)
        d.devScales = torch.empty(count * 3, dtype=torch.float32, device=device)
        d.devOpacities = torch.empty(count, dtype=torch.float32, device=device)
        d.devRotations = torch.empty(count * 4, dtype=torch.float32, device=device)
        return d

    def clone(self):
        d = ModelSplatsDevice()
        d.capacity, d.shDegree, d.shCoeffs, d.count = self.capacity, self.shDegree, self.shCoeffs, self.count
        for n in ("devLocations", "devShs", "devScales", "devOpacities", "devRotations"):
            setattr(d, n, getattr(self, n).clone())
        return d


def _params_struct(project):
    p = ProjectParams()
    for n in Project.PARAM_ORDER:
        setattr(p, n, float(getattr(project, n)))
    return p


def _camera_key(cam):
    return (cam.location.tobytes(), cam.target.tobytes(), float(cam.fovDegY))


def camera_struct(consts):
    c = CameraC()
    c.view[:] = [float(x) for x in consts["view"]]
    c.projview[:] = [float(x) for x in consts["projview"]]
    c.location[:] = [float(x) for x in consts["location"]]
    c.tan_fovx, c.tan_fovy = float(consts["tan_fovx"]), float(consts["tan_fovy"])
    return c


class Comm:
    """gsb_comm: the library's own NCCL communicator (include/gsb.h).  One process per GPU: rank 0 creates the NCCL
    unique id, `exchange(bytes) -> bytes` ships it to the other ranks (any out-of-band channel)."""

    def __init__(self, handle, rank, world):
        self._h, self.rank, self.world = handle, rank, world

    @classmethod
    def init(cls, rank, world, exchange):
        uid = (C.c_ubyte * 128)()
        if rank == 0:
            check(lib().gsb_comm_get_unique_id(uid))
        raw = exchange(bytes(uid))
        buf = (C.c_ubyte * 128).from_buffer_copy(raw)
        h = C.c_void_p()
        check(lib().gsb_comm_init(C.byref(h), buf, rank, world))
        return cls(h, rank, world)

    @classmethod
    def from_torch_distributed(cls, dist, device):
        """Bootstrap over an initialised torch.distributed group (any backend): broadcast of the 128-byte id."""
        rank, world = dist.get_rank(), dist.get_world_size()

        def exchange(raw):
            on = device if dist.get_backend() == "nccl" else "cpu"
            t = torch.tensor(list(raw), dtype=torch.uint8, device=on)
            dist.broadcast(t, src=0)
            return bytes(t.cpu().tolist())
        return cls.init(rank, world, exchange)

    def allreduce_sum(self, tensor):
        check(lib().gsb_comm_allreduce_sum(self._h, tensor.data_ptr(), tensor.numel(), None))

    def window(self, floats):
        """COLLECTIVE: a peer window of `floats` floats (gsb_comm_window_create)."""
        return Window(self, floats)

    def close(self):
        if self._h:
            lib().gsb_comm_destroy(self._h)
            self._h = None


class Window:
    """gsb_window: `floats` floats on this rank that every other rank of the node maps; allreduce_sum() is the library's
    two-shot kernel over NVLink peer memory (ncclAllReduce inside when the peers could not be mapped)."""

    def __init__(self, comm, floats):
        h = C.c_void_p()
        check(lib().gsb_comm_window_create(comm._h, int(floats), C.byref(h)))
        self._h, self.comm = h, comm
        self.device = torch.device("cuda", int(lib().gsb_comm_device(comm._h)))
        self.tensor = _wrap_device_floats(lib().gsb_window_data(h), int(floats), self.device)

    @property
    def peer_mapped(self):
        return bool(lib().gsb_window_peer_mapped(self._h))

    def allreduce_sum(self, count=None):
        n = self.tensor.numel() if count is None else int(count)
        check(lib().gsb_window_allreduce_sum(self._h, n, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))

    def close(self):
        """COLLECTIVE; before Comm.close()."""
        if self._h:
            self.tensor = None
            check(lib().gsb_comm_window_destroy(self._h))
            self._h = None


class _DeviceFloats:
    """__cuda_array_interface__ holder so torch can view memory the library owns (no copy, no ownership)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def _wrap_device_floats(ptr, n, device):
    with torch.cuda.device(device):
        return torch.as_tensor(_DeviceFloats(ptr, n), device=device)


class Trainer:
    """Trainer (src/Trainer.cuh:9-75): public `model`, `truthFrameBuffersW/B`, `truthCameras`,
    `render`, `train`.  Resolution is a constructor argument (the reference hard-codes 1024x1024)."""

    def __init__(self, width=1024, height=1024, device="cuda", square_fov=None):
        self.width, self.height = int(width), int(height)
        self.device = torch.device(device)
        # Trainer::train assumes a square image (tan_fovx = tan_fovy); non-square configs use render()'s rule
        self.square_fov = (width == height) if square_fov is None else square_fov
        h = C.c_void_p()
        check(lib().gsb_trainer_create(C.byref(h), self.width, self.height, None))
        self._h = h
        self.model = None
        self.truthFrameBuffersW, self.truthFrameBuffersB, self.truthCameras = [], [], []
        self._bound = None
        self._truth_sig = None
        self._total_cameras = None
        self._allreduce_cb = None
        self._avg = None
        self._comm = None
        self._render_cams = {}

    def close(self):
        if getattr(self, "_h", None):
            lib().gsb_trainer_destroy(self._h)
            self._h = None

    def __del__(self):
        # With a communicator the destroy is collective (the running means live in a peer window every rank maps):
        # never from the garbage collector -- close() explicitly, on every rank, before the Comm.
        try:
            if getattr(self, "_comm", None) is None:
                self.close()
        except Exception:
            pass

    # -- plumbing ------------------------------------------------------------------------------
    def _bind(self):
        m = self.model
        sig = (m.count, m.devLocations.data_ptr(), m.devShs.data_ptr(), m.devScales.data_ptr(),
               m.devOpacities.data_ptr(), m.devRotations.data_ptr())
        if sig != self._bound:   # `model` is a public field callers replace (src/ui/UiFrame.cpp:157-158)
            check(lib().gsb_trainer_bind_model(self._h, m.count, m.capacity, m.shDegree, m.shCoeffs,
                                               m.devLocations.data_ptr(), m.devShs.data_ptr(), m.devScales.data_ptr(),
                                               m.devOpacities.data_ptr(), m.devRotations.data_ptr()))
            self._bound = sig
            if self._avg is not None:   # caller-visible gradient buffer must follow the splat count
                if self._comm is not None:
                    self._avg = None        # library-owned window: wrapped again on the next grad_tensor()
                else:
                    self._alloc_avg()

    def set_total_cameras(self, n):
        """Camera count over all ranks (S = 2n in accumulateGradients, src/Trainer.cu:419)."""
        self._total_cameras = int(n)
        self._truth_sig = None

    def _sync_truths(self):
        n = len(self.truthFrameBuffersW)
        # keyed on the camera VALUES (a Camera mutated in place, or a rebuilt list that re-uses CPython ids, must be seen)
        sig = (tuple(t.data_ptr() for t in self.truthFrameBuffersW), tuple(t.data_ptr() for t in self.truthFrameBuffersB),
               tuple(_camera_key(c) for c in self.truthCameras), self._total_cameras)
        if sig == self._truth_sig:
            return
        arr_w = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in self.truthFrameBuffersW])
        arr_b = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in self.truthFrameBuffersB])
        # the camera constants depend on the camera objects only: when a caller just swaps frame buffers (double-buffered
        # uploads) they are not recomputed
        cam_sig = (sig[2], self.width, self.height, self.square_fov)
        if getattr(self, "_cam_sig", None) != cam_sig:
            cams = (CameraC * max(n, 1))()
            for k, cam in enumerate(self.truthCameras):
                cams[k] = camera_struct(cam.view_constants(self.width, self.height, self.square_fov))
            self._cam_sig, self._cam_structs = cam_sig, cams
        cams = self._cam_structs
        check(lib().gsb_trainer_set_truths(self._h, n, arr_w, arr_b, cams, self._total_cameras or n))
        self._truth_sig = sig

    def grad_tensor(self):
        """The running-mean buffer [var | loc | sh | scale | opacity | rot] as a torch tensor the
        caller may all-reduce (torch.distributed) between accumulate() and apply()."""
        self._bind()
        if self._comm is not None:
            # with a communicator the buffer is the library's peer window (gsb_comm_window_create, allocated collectively
            # by the first accumulate after the count changed): wrap it, never replace it
            n = int(lib().gsb_trainer_grad_floats(self._h))
            ptr = lib().gsb_trainer_grad_buffer(self._h)
            if not ptr:
                raise RuntimeError("grad_tensor(): with a communicator the running means exist after the first train()/accumulate()")
            if self._avg is None or self._avg.data_ptr() != ptr or self._avg.numel() != n:
                self._avg = _wrap_device_floats(ptr, n, self.device)
            return self._avg
        if self._avg is None:
            self._alloc_avg()
        return self._avg

    def _alloc_avg(self):
        n = int(lib().gsb_trainer_grad_floats(self._h))
        self._avg = torch.zeros(n, dtype=torch.float32, device=self.device)
        check(lib().gsb_trainer_set_grad_buffer(self._h, self._avg.data_ptr()))

    def set_allreduce(self, fn):
        """fn(tensor) sums the gradient buffer across ranks in place (e.g. dist.all_reduce)."""
        if fn is None:
            self._allreduce_cb = None
            check(lib().gsb_trainer_set_allreduce(self._h, ALLREDUCE_FN(0), None))
            return
        self.grad_tensor()

        def hook(_dev, _count, _stream, _user):
            try:
                fn(self._avg)
                return 0
            except Exception:   # never unwind through C
                return 1
        self._allreduce_cb = ALLREDUCE_FN(hook)
        check(lib().gsb_trainer_set_allreduce(self._h, self._allreduce_cb, None))

    def set_comm(self, comm):
        """Native exchange: train() all-reduces the running means over `comm` (a Comm) inside libgsb200.so."""
        if self._avg is not None:
            torch.cuda.synchronize(self.device)
            check(lib().gsb_trainer_set_grad_buffer(self._h, None))
            self._avg = None
        self._comm = comm
        check(lib().gsb_trainer_set_comm(self._h, comm._h if comm is not None else None))

    def exchange_path(self):
        """"none" (one rank), "nccl" (ncclAllReduce) or "peer" (the library's two-shot kernel over NVLink peer memory)."""
        return ("none", "nccl", "peer")[int(lib().gsb_trainer_exchange_path(self._h))]

    def set_fuse_backgrounds(self, on=True):
        """White + black pass of a camera in one forward / one backward walk (default on)."""
        check(lib().gsb_trainer_set_fuse_backgrounds(self._h, int(on)))

    def enable_loss(self, on=True):
        check(lib().gsb_trainer_enable_loss(self._h, int(on)))

    def loss(self, iterations_back=0):
        """Sum of squared residuals of the latest iteration (0) or the one before (1); waits for that value only."""
        v = C.c_double(0.0)
        check(lib().gsb_trainer_get_loss(self._h, int(iterations_back), C.byref(v)))
        return v.value

    def stats(self):
        s = TrainerStats()
        check(lib().gsb_trainer_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in TrainerStats._fields_}

    STAGES = ("clear", "preprocess", "tile_scan", "emit", "tile_sort", "render_fwd", "render_bwd", "preprocess_bwd")

    def profile_pass(self, pass_index):
        """Per-stage milliseconds (CUDA events) and counters of one training view pass."""
        ms = (C.c_float * 8)()
        cnt = (C.c_uint * 2)()
        check(lib().gsb_trainer_profile_pass(self._h, pass_index, ms, cnt))
        return dict(zip(self.STAGES, [float(x) for x in ms])), {"R": int(cnt[0]), "visible": int(cnt[1])}

    def set_overlap(self, on=True):
        """Forward chain of the next view pass underneath the backward chain of the current one (default on)."""
        check(lib().gsb_trainer_set_overlap(self._h, int(on)))

    def set_stage_timing(self, on=True):
        """Bracket every stage of every view pass with CUDA events (resets the totals)."""
        check(lib().gsb_trainer_set_stage_timing(self._h, int(on)))

    def stage_times(self):
        """({stage: total ms}, view-pass launches) since set_stage_timing(True)."""
        ms = (C.c_double * 8)()
        n = C.c_longlong(0)
        check(lib().gsb_trainer_get_stage_times(self._h, ms, C.byref(n)))
        return dict(zip(self.STAGES, [float(x) for x in ms])), int(n.value)

    def count_pairs(self, camera_index):
        """Pair counts of one local camera under the reference's control flow (roofline FLOPs)."""
        out = (C.c_ulonglong * 8)()
        self._bind()
        self._sync_truths()
        check(lib().gsb_trainer_count_pairs(self._h, camera_index, out))
        return dict(zip(("eval_f", "exp_f", "blend_f", "eval_b", "exp_b", "blend_b", "R", "visible"), [int(x) for x in out]))

    # -- reference interface -------------------------------------------------------------------
    def render(self, frameBuffer, sizeX, sizeY, splatScale, camera):
        """Trainer::render (src/Trainer.cu:148-216): RGBA8 preview into `frameBuffer` (int32/uint32 tensor)."""
        self._bind()
        key = (_camera_key(camera), int(sizeX), int(sizeY))
        cs = self._render_cams.get(key)
        if cs is None:                      # the float32 glm-order matrix maths runs in Python: once per camera and size
            if len(self._render_cams) > 4096:
                self._render_cams.clear()
            cs = self._render_cams[key] = camera_struct(camera.view_constants(sizeX, sizeY, square_fov=False))
        check(lib().gsb_trainer_render(self._h, frameBuffer.data_ptr(), sizeX, sizeY, float(splatScale), C.byref(cs)))

    def accumulate(self):
        self._bind()
        self._sync_truths()
        check(lib().gsb_trainer_accumulate(self._h))

    def apply(self, project):
        p = _params_struct(project)
        check(lib().gsb_trainer_apply(self._h, C.byref(p)))

    def train(self, project, densify):
        """Trainer::train (src/Trainer.cu:252-543)."""
        if not self.truthFrameBuffersW:
            raise RuntimeError("Can't run training iteration, no truth data available!")
        project.iterations += 1
        self._bind()
        self._sync_truths()
        p = _params_struct(project)
        new_count = C.c_int(0)
        check(lib().gsb_trainer_train(self._h, C.byref(p), int(bool(densify)), C.byref(new_count)))
        if densify:
            self._commit_densify(new_count.value)

    def densify(self, project):
        """The densify block alone (src/Trainer.cu:437-542) on the current running means."""
        self._bind()
        p = _params_struct(project)
        new_count = C.c_int(0)
        check(lib().gsb_trainer_densify_plan(self._h, C.byref(p), C.byref(new_count)))
        self._commit_densify(new_count.value)

    def _commit_densify(self, new_count):
        old = self.model
        new = ModelSplatsDevice.empty(new_count, old.capacity, old.shDegree, old.shCoeffs, self.device)
        check(lib().gsb_trainer_densify_commit(self._h, new.devLocations.data_ptr(), new.devShs.data_ptr(),
                                               new.devScales.data_ptr(), new.devOpacities.data_ptr(),
                                               new.devRotations.data_ptr()))
        self.model = new          # `delete model; model = new ModelSplatsDevice(...)` (src/Trainer.cu:540-541)
        self._bound = None
        self._bind()
