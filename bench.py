#!/usr/bin/env python
"""bench.py -- views/s (forward + loss + backward + accumulate, apply once per iteration) of the
training hot path at BASELINE.json's config C4: 32+32 cameras, 1920x1080, 500k splats.

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N ...            # the reference's own CUDA build (oracle/_ref)

One step = one Trainer::train iteration over all 64 cameras x 2 backgrounds = 128 view passes
(reference src/Trainer.cu:252-435).  With N ranks the cameras are partitioned (rank r takes cameras
r, r+N, ... of each sphere), splats are replicated, and the 24 running-mean floats per splat are
summed with one NCCL all-reduce per iteration before applyGradients: total work is fixed, so
`scaling` is "strong".

value   views/s with every input resident in HBM, CUDA events around K steps, max over ranks.
e2e     the same through Trainer.train with the truth frames in pinned HOST memory: every step
        uploads that step's RGBA8 truth frames (double-buffered on a copy stream) and reads the
        loss scalar back.
roofline  dominant kernel = render backward: achieved FP32 FLOP/s from the reference's per-pair
        FLOP counts and the kernel's mean launch time measured live with CUDA events
        around every stage of every view pass in a second loop over the same K steps.  No stage of this path is a dense contraction or HBM-bound at the
        top, so the bound is the FP32 SIMT issue rate; the whole-step HBM figure is reported beside it.
cpu_baseline  the scalar C++ oracle (oracle/liboracle.so, one host thread) on ONE view pass of the
        same workload -- context only; the graded baseline is `--impl reference`: the reference has
        no CPU path, its own code for this path is CUDA (north_star).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "views/s fwd+bwd+update @500k splats 1080p"
UNIT = "views/s"


_JSON_OUT = None


def emit(obj):
    f = _JSON_OUT or sys.stdout
    f.write(json.dumps(obj) + "\n")
    f.flush()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("gsb", "reference"), default="gsb")
    ap.add_argument("--config", default="C4", help="BASELINE config (C1..C5); the metric is quoted on C4")
    ap.add_argument("--cameras", type=int, default=0, help="override cameras per sphere (debug; invalidates the metric)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fuse", action="store_true", help="one full pass per (camera, background) like the reference")
    ap.add_argument("--torch-allreduce", action="store_true", help="A/B: torch.distributed.all_reduce through the hook instead of gsb_comm_*")
    ap.add_argument("--nccl-exchange", action="store_true", help="A/B: ncclAllReduce instead of the library's peer-window kernel (GSB_PEER_EXCHANGE=0)")
    ap.add_argument("--no-mgpu-check", action="store_true", help="skip the N > 1 correctness job (tests/mgpu_check.py)")
    ap.add_argument("--densify-every", type=int, default=0, help="densify every k-th step inside the timed loop (C5 as written)")
    args = ap.parse_args()
    if args.nccl_exchange:
        os.environ["GSB_PEER_EXCHANGE"] = "0"
    return args


# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, gpu_id):
        self.gpu_id, self.proc, self.path = gpu_id, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_id), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=fd, stderr=subprocess.DEVNULL)
            os.close(fd)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        with open(self.path) as f:
            for line in f:
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
                except ValueError:
                    continue
                for name, v in zip(self.NAMES, c[5:9]):
                    if v == "Active":
                        reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


def gpu_identifier(torch, dev_index):
    try:
        u = str(torch.cuda.get_device_properties(dev_index).uuid)
        return u if u.startswith("GPU-") else "GPU-" + u
    except Exception:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            return vis.split(",")[dev_index]
        return dev_index


def init_dist(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: this path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist = dist_
        if not (args.impl == "reference"):
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    return torch, dist, rank, world, local


# ---------------------------------------------------------------------------------------------
def build_scene(gsb, args, rank, world, reference_cameras=False):
    """Student cloud, teacher cloud and the camera set of the config; cameras sharded by rank.
    reference_cameras: the camera set comes from the reference's own Camera.cpp (oracle/_ref/libgsrefcam.so) instead of the
    product's host code, so the reference arm never loads libgsb200.so (the two agree bit for bit: tests/test_host_logic.py)."""
    cfg = args.config
    W, H, P = gsb.synth.CONFIGS[cfg]
    project = gsb.synth.project_for(cfg)
    if args.cameras > 0:
        project.sphere1.count = args.cameras
        project.sphere2.count = args.cameras if project.sphere2.count > 0 else 0
    if reference_cameras:
        from oracle import refshim
        cams = [gsb.Camera(loc, (0.0, 0.0, 0.0), fov) for loc, fov in refshim.ref_cameras(project)]
    else:
        cams = gsb.Camera.getCameras(project)
    seed = 0x5EED0000 + int(cfg[1])
    cloud = gsb.synth.splat_cloud(P, seed)
    teacher = gsb.synth.splat_cloud(P // 2, seed + 1)
    mine = gsb.synth.shard_cameras(cams, project.sphere1.count, rank, world)
    return {"W": W, "H": H, "P": P, "project": project, "cams": cams, "mine": mine, "cloud": cloud, "teacher": teacher}


def render_truths(torch, H, gsb, scene, dev, forward_fn, quantise_fn):
    """Teacher cloud on white and black, quantised like imageFloatToInt (src/Trainer.cu:19-29): the
    stand-in for captureTruths (the OptiX truth tracer is unavailable on this box)."""
    W, Hh = scene["W"], scene["H"]
    sc = H.scene_tensors(scene["teacher"], dev)
    Pt = scene["teacher"]["opacities"].shape[0]
    tw, tb, consts = [], [], []
    for cam in scene["mine"]:
        c = cam.view_constants(W, Hh, square_fov=(W == Hh))
        consts.append(c)
        for bg, dst in (((1.0, 1.0, 1.0), tw), ((0.0, 0.0, 0.0), tb)):
            vw = H.view_tensors(c, bg, dev)
            out = forward_fn(sc, vw, Pt, 1, 4, W, Hh)
            fb = torch.empty(W * Hh, dtype=torch.int32, device=dev)
            quantise_fn(out, fb, W, Hh)
            dst.append(fb)
    torch.cuda.synchronize()
    return tw, tb, consts


def timed_steps(torch, dist, step_fn, steps):
    """K steps between barrier + synchronize, CUDA events on the launching (legacy default) stream;
    returns the max over ranks of the elapsed milliseconds."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist is not None and dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for k in range(steps):
        step_fn(k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None and dist.is_initialized():
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    return ms


STREAM_BYTES = {  # algorithmic bytes per view pass and stage group, SURVEY.md 8(d)
    "preprocess": lambda c: 100 * c["P"] + 67 * c["Pv"],
    "binning": lambda c: 8 * c["P"] + 20 * c["P"] + 12 * c["R"] + 152 * c["R"] + 8 * c["R"] + 16 * c["tiles"],
    "preprocess_bwd": lambda c: (92 + 243) * c["Pv"] + 284 * c["P"],
}


def workload_text(cfg, n_total, views_per_step, W, Hh, P):
    """config.workload, identical in both arms (the driver compares the strings)."""
    return (f"{cfg}: {n_total} cameras x 2 backgrounds = {views_per_step} view passes per step, {W}x{Hh}, {P} splats, "
            f"SH degree 1")


def fused_step_bytes(c):
    """B_fused of SURVEY.md 8(d): external inputs/outputs of one view pass of a fused design."""
    return 128 * c["P"] + 526 * c["Pv"] + 252 * c["R"] + 44 * c["N_pix"]


def fwd_flops(pc):
    return 11 * pc["eval_f"] + 3 * pc["exp_f"] + 11 * pc["blend_f"]


def bwd_flops(pc):
    return 11 * pc["eval_b"] + 3 * pc["exp_b"] + 64 * pc["blend_b"]


# ---------------------------------------------------------------------------------------------
def measure_render_mpix(torch, render_one, cams, W, Hh, reps=3):
    """BASELINE.json's second metric: Mpix/s of the Trainer::render path (forward + float -> RGBA8, src/Trainer.cu:148-216)
    over the local cameras.  render_one(cam) must leave the frame in device memory; CUDA events around reps x cameras."""
    for cam in cams[:2]:
        render_one(cam)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for cam in cams:
            render_one(cam)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    frames = reps * len(cams)
    return {"mpix_per_s": W * Hh * frames / (ms * 1e-3) / 1e6, "ms_per_frame": ms / frames, "frames": frames,
            "resolution": [W, Hh]}


def measure_truth_tracer(torch, gsb, dev, cams, size=1024, samples=100):
    """captureTruths through the CUDA ray caster: one white+black frame pair per camera at the
    reference's truth resolution and rtSamples default (src/Config.h:13-14, src/Project.h:24), procedural textured mesh."""
    import importlib
    app = importlib.import_module("gaussian-splatterer_b200.app")
    rtx = app.RtxHost()
    v, t, tc = gsb.synth.textured_mesh(rings=96, segments=192, seed=3)
    rtx.set_mesh(v, t, tc)
    rtx.set_texture(gsb.synth.procedural_texture(256, 5))
    w = torch.zeros(size * size, dtype=torch.int32, device=dev)
    b = torch.zeros(size * size, dtype=torch.int32, device=dev)
    descs = [app.camera_desc(c.location, (0.0, 0.0, 0.0), c.fovDegY) for c in cams]
    rtx.render_pair(w, b, (size, size), descs[0], 4)                      # warm-up
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = 0
    e0.record()
    for d in descs:
        rtx.render_pair(w, b, (size, size), d, samples)
        rays += rtx.last_ray_count()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    rtx.close()
    return {"mray_per_s": rays / (ms * 1e-3) / 1e6, "ms_per_camera_pair": ms / len(descs), "cameras": len(descs),
            "triangles": int(t.shape[0]), "resolution": [size, size], "samples": samples,
            "ray_segments_per_pixel_sample": rays / (len(descs) * size * size * samples),
            "note": "white and black truth frame of a camera come from one trace (gsb_rtx_render_pair); the reference's "
                    "OptiX/OWL tracer is unavailable on this box, so there is no reference number beside this one"}


# ---------------------------------------------------------------------------------------------
def cpu_baseline_port(gsb, scene, truth_white0):
    """One view pass (forward + loss + backward + accumulate) of the same workload on the scalar C++
    oracle, one host thread.  Context only."""
    import numpy as np
    from oracle import pyoracle
    W, Hh, P, cloud = scene["W"], scene["H"], scene["P"], scene["cloud"]
    cs = scene["mine"][0].view_constants(W, Hh, square_fov=(W == Hh))
    M = 4
    avg = [np.zeros(k, np.float32) for k in (P, 3 * P, 3 * M * P, 3 * P, P, 4 * P)]
    t0 = time.perf_counter()
    out, v = pyoracle.forward(P, 1, M, (1.0, 1.0, 1.0), W, Hh, cloud["locations"], cloud["shs"], None, cloud["opacities"],
                              cloud["scales"], 1.0, cloud["rotations"], None, cs["view"], cs["projview"], cs["location"],
                              cs["tan_fovx"], cs["tan_fovy"])
    loss = pyoracle.image_int_to_loss(truth_white0, out, W, Hh)
    g = v.backward(loss, M)
    pyoracle.accumulate(avg, [g["dL_dmean3D"], g["dL_dsh"], g["dL_dscale"], g["dL_dopacity"], g["dL_drot"]],
                        2.0 * len(scene["cams"]), M, P)
    dt = time.perf_counter() - t0
    v.close()
    return {"value": 1.0 / dt, "unit": UNIT, "cores": 1, "kind": "port", "host_cores_total": os.cpu_count(),
            "sample": f"1 view pass (camera 0, white background; forward+loss+backward+accumulate) of {P} splats at "
                      f"{W}x{Hh} on oracle/liboracle.so, {dt:.1f} s; the reference has no CPU path"}


def run_gsb(args):
    torch, dist, rank, world, local = init_dist(args)
    import numpy as np
    import helpers as H
    gsb = H.gsb
    dev = f"cuda:{local}"
    scene = build_scene(gsb, args, rank, world)
    W, Hh, P = scene["W"], scene["H"], scene["P"]
    project, mine = scene["project"], scene["mine"]
    n_local, n_total = len(mine), len(scene["cams"])
    views_per_step = 2 * n_total

    L = gsb.lib()
    tw, tb, _ = render_truths(torch, H, gsb, scene, dev,
                              lambda sc, vw, *a: H.run_forward_gsb(sc, vw, *a)[1],
                              lambda out, fb, w, h: L.gsb_image_float_to_int(out.data_ptr(), fb.data_ptr(), w, h, None))
    cloud = scene["cloud"]
    host = gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")])

    comm = gsb.Comm.from_torch_distributed(dist, dev) if (world > 1 and not args.torch_allreduce) else None

    def new_trainer():
        tr = gsb.Trainer(W, Hh, dev)
        tr.set_fuse_backgrounds(not args.no_fuse)
        tr.set_total_cameras(n_total)
        tr.model = gsb.ModelSplatsDevice(host, dev)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = list(tw), list(tb), mine
        if world > 1:
            if args.torch_allreduce:
                tr.set_allreduce(lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM))
            else:
                tr.set_comm(comm)                  # the library's own NCCL communicator
        return tr

    # ---- value: inputs resident in HBM ------------------------------------------------------
    tr = new_trainer()
    if args.densify_every > 0 and comm is None:
        tr.grad_tensor()     # caller-owned buffer must exist BEFORE the iterations whose means the thresholds are taken from
    for _ in range(args.warmup):
        tr.train(project, False)
    densify_info = None
    if args.densify_every > 0:
        # "densify/split/cull compaction enabled" (BASELINE config 5): thresholds at quantiles of the current decision
        # variables so that every densify splits / clones ~2 % and culls ~2 % of the splats.  Every rank holds the same
        # all-reduced means and the same parameters, so every rank derives the same thresholds.
        torch.cuda.synchronize()
        a = tr.grad_tensor().cpu().numpy()
        Pn = tr.model.count
        score = a[:Pn].astype(np.float64) - np.linalg.norm(a[Pn:4 * Pn].astype(np.float64).reshape(-1, 3), axis=1)
        project.paramDensifyVariance = float(np.quantile(score, 0.98))
        project.paramCullOpacity = float(np.quantile(tr.model.devOpacities.cpu().numpy(), 0.02))
        project.paramSplitSize = float(np.median(np.linalg.norm(tr.model.devScales.cpu().numpy().reshape(-1, 3), axis=1)))
        densify_info = {"every": args.densify_every, "count_before": int(Pn),
                        "thresholds": [project.paramDensifyVariance, project.paramCullOpacity, project.paramSplitSize]}
    do_densify = (lambda k: (k + 1) % args.densify_every == 0) if args.densify_every > 0 else (lambda k: False)
    launches0 = tr.stats()["kernel_launches"]
    sampler = ClockSampler(gpu_identifier(torch, local))
    if rank == 0:
        sampler.start()
    ms = timed_steps(torch, dist, lambda k: tr.train(project, do_densify(k)), args.steps)
    clocks = sampler.stop() if rank == 0 else None
    stats = tr.stats()
    launches = stats["kernel_launches"] - launches0
    value = views_per_step * args.steps / (ms / 1000.0)
    if densify_info is not None:
        densify_info["count_after"] = int(tr.model.count)
        if world > 1:       # replicas must still be bit-identical after densifying inside the timed loop
            import mgpu_check
            cs = torch.tensor(mgpu_check._checksum(torch, tr.model), dtype=torch.int64, device=dev)
            gathered = [torch.zeros_like(cs) for _ in range(world)]
            dist.all_gather(gathered, cs)
            densify_info["replicas_identical_after_run"] = all(bool(torch.equal(g, gathered[0])) for g in gathered)
        P = int(tr.model.count)

    # ---- per-kernel launch times, live: the same K steps again with every stage of every view pass
    # bracketed by CUDA events on the launching stream.  Stage timing serialises the kernels (in the
    # loop above the next pass's forward chain overlaps the current backward), so a stage's time is
    # its own and `ms_serial` is the un-overlapped step.
    tr.set_stage_timing(True)
    ms_serial = timed_steps(torch, dist, lambda k: tr.train(project, False), args.steps)
    stage_ms, passes = tr.stage_times()
    st_x = tr.stats()
    exchange_path = "torch" if (world > 1 and args.torch_allreduce) else tr.exchange_path()
    iteration_tail = {"exchange_path": exchange_path, "exchange_ms": st_x["exchange_ms"] / max(st_x["timed_iterations"], 1),
                      "apply_ms": st_x["apply_ms"] / max(st_x["timed_iterations"], 1),
                      "what": "per iteration: all-reduce of the running means (on the trainer's stream) and applyGradients; "
                              "with the first forward chain (nothing to overlap with) they are the serial section of a step"}
    tr.set_stage_timing(False)

    # ---- counters for the roofline (after the timed region) ----------------------------------
    pcs = [tr.count_pairs(i) for i in range(n_local)]
    mean = lambda k: float(sum(p[k] for p in pcs)) / max(len(pcs), 1)
    tiles = ((W + 15) // 16) * ((Hh + 15) // 16)
    cnt = {"P": P, "Pv": mean("visible"), "R": mean("R"), "N_pix": W * Hh, "tiles": tiles}
    pair = {k: mean(k) for k in ("eval_f", "exp_f", "blend_f", "eval_b", "exp_b", "blend_b")}
    fused = not args.no_fuse
    units = 2 if fused else 1                     # view passes one launch of a stage serves
    per_launch = {k: v / max(passes, 1) for k, v in stage_ms.items()}
    fp32_meas = (C.c_double * 2)()
    if L.gsb_measure_fp32_peak(fp32_meas, None) != 0:
        fp32_meas[0] = fp32_meas[1] = 0.0
    tr.close()

    # ---- the same steps with one full pass per (camera, background), the reference's order ----
    unfused = None
    if fused and not args.no_e2e and args.densify_every == 0:
        tr = new_trainer()
        tr.set_fuse_backgrounds(False)
        for _ in range(args.warmup):
            tr.train(project, False)
        k_u = max(2, args.steps // 2)
        ms_u = timed_steps(torch, dist, lambda k: tr.train(project, False), k_u)
        unfused = {"value": views_per_step * k_u / (ms_u / 1000.0), "unit": UNIT, "ms_per_step": ms_u / k_u, "steps": k_u,
                   "what": "one forward + one backward walk per (camera, background) in the reference's order "
                           "(gsb_trainer_set_fuse_backgrounds(0)); `value` above shares them between a camera's two passes"}
        tr.close()

    # ---- N > 1: the correctness job (bit-identical replicas, equal to N = 1, with a densify) ----------------
    mgpu = None
    if world > 1 and not args.no_mgpu_check and not args.torch_allreduce:
        import mgpu_check
        mgpu = mgpu_check.run_check(torch, dist, rank, world, local, config="C4", cameras=8)

    # ---- e2e: truth frames in pinned host memory, uploaded every step, loss read back --------
    e2e = None
    if not args.no_e2e:
        tr = new_trainer()
        tr.enable_loss(True)
        # one pinned host block and one device block per buffer set: the step's 2n frames travel as ONE cudaMemcpyAsync
        # (128 separate copies cost ~2 ms of host enqueue time per step, during which the compute stream sat idle)
        n_fr = len(tw) + len(tb)
        numa = importlib.import_module("gaussian-splatterer_b200.numa")
        props = torch.cuda.get_device_properties(local)
        try:
            bus = "%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        except Exception:
            bus = None
        with numa.local_to_gpu(bus) as placement:         # the pinned block lives on the GPU's own NUMA node
            host_all = torch.stack([t.cpu() for t in list(tw) + list(tb)]).pin_memory()
        dev_all = [torch.stack(list(tw) + list(tb)), torch.empty((n_fr,) + tuple(tw[0].shape), dtype=tw[0].dtype, device=tw[0].device)]
        dset = [([dev_all[w][i] for i in range(len(tw))], [dev_all[w][len(tw) + i] for i in range(len(tb))]) for w in (0, 1)]
        copy_stream = torch.cuda.Stream()
        up_ev = [torch.cuda.Event(), torch.cuda.Event()]
        done_ev = [torch.cuda.Event(), torch.cuda.Event()]
        losses = []

        def upload(which, after=None):
            # `after`: the end of the step that read the set about to be overwritten.  Without it the copy starts while
            # that step's last backward pass still reads its truth frames (a write-after-read race, harmless here only
            # because every step uploads the same frames) and runs underneath that step's exchange, which it slows down:
            # tools/e2e_bisect.py on 2 B200, 8 cameras per rank: 8.38 ms HBM-resident, 8.67 end to end unordered, 8.50 ordered
            if after is not None:
                copy_stream.wait_event(after)
            with torch.cuda.stream(copy_stream):
                dev_all[which].copy_(host_all, non_blocking=True)
                up_ev[which].record(copy_stream)

        def step(k, last=False):
            cur = k & 1
            torch.cuda.current_stream().wait_event(up_ev[cur])    # this step's inputs have landed
            upload(cur ^ 1, done_ev[cur ^ 1])                     # next step's inputs, overlapped with this step
            tr.truthFrameBuffersW, tr.truthFrameBuffersB = dset[cur]
            tr.train(project, False)
            done_ev[cur].record()     # legacy stream: behind everything the step enqueued (joined backward chains, exchange, apply)
            # device -> host read of a step's result, every step: the scalar of step k-1 is read while step k runs (the
            # library copies it to pinned memory at the end of each iteration), the last step's own before the clock stops
            losses.append(tr.loss(1 if not last else 0) if (len(losses) or last) else tr.loss(0))

        upload(0)
        for k in range(2 * ((args.warmup + 1) // 2)):             # even count: the timed loop starts on set 0
            step(k)
        ms_e = timed_steps(torch, dist, lambda k: step(k, last=(k == args.steps - 1)), args.steps)
        h2d = n_local * 2 * W * Hh * 4 + n_local * 148
        e2e = {"value": views_per_step * args.steps / (ms_e / 1000.0), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 8 * 128 + 4 * 2 * n_local,   # GSB_LOSS_SLOTS partial sums + R of every view pass
               "ms_per_step": ms_e / args.steps,
               "result": "sum of squared residuals", "last_loss": losses[-1],
               "how": "Trainer.train with RGBA8 truth frames in pinned host memory (allocated on the GPU's NUMA node); one "
                      "cudaMemcpyAsync of the next step's frames on a copy stream overlaps the current step (it starts when the step that "
                      "last read that device set has finished); the loss scalar of "
                      "step k-1 is read from pinned memory while step k runs, the last step's before the clock stops; per rank",
               "host_placement": placement.info}
        tr.close()

    render_stats = truth_stats = None
    if rank == 0 and world == 1 and not args.no_e2e:
        tr = new_trainer()
        fb = torch.empty(W * Hh, dtype=torch.int32, device=dev)
        render_stats = measure_render_mpix(torch, lambda cam: tr.render(fb, W, Hh, 1.0, cam), mine[:16], W, Hh)
        tr.close()
        truth_stats = measure_truth_tracer(torch, gsb, dev, mine[:4])

    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    sm_max = float((clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz", 1965.0))
    fp32_nominal = 2 * 128 * 148 * sm_max * 1e6 / 1e12
    fp32_peak = float(fp32_meas[0]) if fp32_meas[0] > 0 else fp32_nominal    # measured FFMA issue peak (csrc/measure.cu)
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        traffic = {}

    bwd_ms = per_launch["render_bwd"]
    bwd_fl = units * bwd_flops(pair)
    fwd_ms = per_launch["render_fwd"]
    stages = {}
    total_stage_ms = sum(per_launch.values())
    for k, v in per_launch.items():
        stages[k] = {"ms_per_launch": round(v, 4), "share": round(v / total_stage_ms, 4) if total_stage_ms else None}
    stages["render_fwd"].update({"tflops": fwd_flops(pair) / (fwd_ms * 1e-3) / 1e12 if fwd_ms else None,
                                 "frac_fp32": fwd_flops(pair) / (fwd_ms * 1e-3) / 1e12 / fp32_peak if fwd_ms else None})
    bin_ms = sum(per_launch[k] for k in ("tile_scan", "emit", "tile_sort"))
    for name, ms_k in (("preprocess", per_launch["preprocess"]), ("binning", bin_ms),
                       ("preprocess_bwd", per_launch["preprocess_bwd"])):
        by = STREAM_BYTES[name](cnt) * (units if name == "preprocess_bwd" else 1)
        gbs = by / (ms_k * 1e-3) / 1e9 if ms_k else None
        stages.setdefault(name, {}).update({"algorithmic_bytes": by, "gb_s": gbs, "frac_hbm": gbs / hbm_peak if gbs else None,
                                            "ms_per_launch": round(ms_k, 4)})
    step_bytes = views_per_step / world * fused_step_bytes(cnt)        # per rank and step
    step_gbs = step_bytes / (ms / args.steps * 1e-3) / 1e9

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(args.config, n_total, views_per_step, W, Hh, P),
                   "views_per_step": views_per_step, "cameras_per_rank": n_local,
                   "fused_white_black": fused,
                   "parallelism": f"cameras partitioned over {world} GPU(s), splats replicated; one all-reduce of 24 floats/splat "
                                  f"per iteration inside libgsb200.so: " + {"peer": "its own two-shot kernel over NVLink peer "
                                  "windows (gsb_window_allreduce_sum)", "nccl": "ncclAllReduce (gsb_comm_allreduce_sum)",
                                  "none": "none (one rank)", "torch": "torch.distributed hook"}[exchange_path],
                   "exchange_path": exchange_path,
                   "l2": "inputs larger than L2: every step streams > 1 GB of truth frames, parameters and per-view state "
                         "through the 126 MB L2; no explicit flush"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "counters_per_view": {**{k: round(v, 1) for k, v in cnt.items()}, **{k: round(v, 1) for k, v in pair.items()}},
        "roofline": {"bound": "fp32", "kernel": "render_bwd_kernel<%d>" % (2 if fused else 1),
                     "achieved": bwd_fl / (bwd_ms * 1e-3) / 1e12 if bwd_ms else None, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": bwd_fl / (bwd_ms * 1e-3) / 1e12 / fp32_peak if bwd_ms else None,
                     "traffic": traffic.get("render_bwd"),
                     "flops_per_launch": bwd_fl, "ms_per_launch": bwd_ms, "share_of_step": stages["render_bwd"]["share"],
                     "view_passes_per_launch": units,
                     "peak_source": "measured on this GPU in this run: FFMA-only kernel, 8 independent chains per thread "
                                    "(gsb_measure_fp32_peak, csrc/measure.cu); MEASURED_PEAKS.json holds no FP32 SIMT figure and no "
                                    "stage is HBM- or tensor-bound at the top (SURVEY.md 8d)" if fp32_meas[0] > 0 else "nominal",
                     "fp32_peak_measured": float(fp32_meas[0]), "fp32x2_peak_measured": float(fp32_meas[1]),
                     "fp32_peak_nominal": fp32_nominal,
                     "fp32_peak_nominal_how": f"2 x 128 lanes x 148 SMs x clocks.max.sm ({sm_max:.0f} MHz)",
                     "frac_is": "effective: the numerator counts the REFERENCE algorithm's FLOPs (SURVEY 8d) for the pairs it "
                                "evaluates; this kernel culls most of them with bounding boxes and never executes them",
                     "step_hbm": {"bound": "hbm", "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s",
                                  "frac": step_gbs / hbm_peak, "bytes_per_step_per_rank": step_bytes, "peak_source": hbm_src}},
        "stages": stages,
        "iteration_tail": iteration_tail,
        "stats": {"rendered_per_step": stats["rendered_total"], "retries": stats["retries"],
                  "ms_per_step_serialised_with_stage_events": ms_serial / args.steps},
    }
    if unfused is not None:
        out["unfused"] = unfused
    if mgpu is not None:
        out["replicas_identical"] = bool(mgpu.get("replicas_identical")) and bool(mgpu.get("rendered_match"))
        out["vs_n1_rel"] = mgpu.get("vs_n1_rel")
        out["vs_n1_rel_first_iteration"] = mgpu.get("vs_n1_rel_first_iteration")
        out["n1_rerun_rel"] = mgpu.get("n1_rerun_rel")     # one GPU against itself on the same job: the yardstick for vs_n1_rel
        out["exchange_microbench"] = mgpu.get("exchange_microbench")
        out["mgpu_check"] = mgpu
    if densify_info is not None:
        out["densify"] = densify_info
    if e2e is not None:
        out["e2e"] = e2e
    if render_stats is not None:
        out["render"] = render_stats              # "render Mpix/s" of BASELINE.json's metric, Trainer::render path
        out["truth_tracer"] = truth_stats
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_port(gsb, scene, tw[0].cpu().numpy().view(np.uint32))
    emit(out)


# ---------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own implementation of the path: its UNMODIFIED CUDA rasterizer compiled for
    sm_100 (oracle/_ref/libgsref.so) driven in Trainer::train's exact call order.  The reference has
    no CPU path and no multi-GPU path: rank 0 runs the whole job on one GPU, other ranks exit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("the reference's implementation of this path is CUDA: a GPU is required")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    import helpers as H
    from oracle import refshim
    gsb = H.gsb                                    # pure-Python scene generation only: libgsb200.so is never loaded here
    if not refshim.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libgsref.so was not built (needs /root/reference)"})
        return
    dev = f"cuda:{local}"
    scene = build_scene(gsb, args, 0, 1, reference_cameras=os.path.exists(refshim.REF_CAM_LIB))
    W, Hh, P = scene["W"], scene["H"], scene["P"]
    project = scene["project"]
    n_total = len(scene["cams"])
    views_per_step = 2 * n_total
    tw, tb, consts = render_truths(torch, H, gsb, scene, dev,
                                   lambda sc, vw, *a: H.run_forward_ref(sc, vw, *a)[0],
                                   lambda out, fb, w, h: refshim.image_float_to_int(out, fb, w, h))
    cloud = scene["cloud"]
    rt = refshim.RefTrainer(W, Hh, dev)
    rt.set_model(cloud, 1000000 if P <= 1000000 else 10000000, 1, 4)
    rt.set_truths(tw, tb, consts)
    p13 = project.params13()
    res = {}
    clocks = None
    for name, as_shipped in (("as_shipped", True), ("best_case", False)):
        for _ in range(args.warmup):
            rt.train(p13, as_shipped=as_shipped)
        sampler = ClockSampler(gpu_identifier(torch, local))
        if not as_shipped:
            sampler.start()
        ms = timed_steps(torch, None, lambda k: rt.train(p13, as_shipped=as_shipped), args.steps)
        if not as_shipped:
            clocks = sampler.stop()
        res[name] = {"views_per_s": views_per_step * args.steps / (ms / 1000.0), "ms_per_step": ms / args.steps}
    rt.close()
    # render Mpix/s of the reference's Trainer::render (src/Trainer.cu:148-216): forward(debug=true) + imageFloatToInt + sync
    sc_r = H.scene_tensors(cloud, dev)
    fb_r = torch.empty(W * Hh, dtype=torch.int32, device=dev)

    def ref_render(cam):
        vw = H.view_tensors(cam.view_constants(W, Hh, square_fov=False), (0.0, 0.0, 0.0), dev)
        out_c, _ = H.run_forward_ref(sc_r, vw, P, 1, 4, W, Hh, debug=True)
        refshim.image_float_to_int(out_c, fb_r, W, Hh)
        torch.cuda.synchronize()
    render_stats = measure_render_mpix(torch, ref_render, scene["cams"][:16], W, Hh)
    render_stats["variant"] = "forward(debug=true) + imageFloatToInt + cudaDeviceSynchronize as in Trainer::render, but with pooled workspaces (no per-call cudaMalloc/cudaFree): favourable to the reference"
    best = res["best_case"]
    out = {
        "impl": "reference", "metric": METRIC, "value": best["views_per_s"], "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": best["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(args.config, n_total, views_per_step, W, Hh, P),
                   "views_per_step": views_per_step,
                   "parallelism": "single GPU (the reference has no multi-GPU path)",
                   "variant": "best case for the reference: debug=false, pooled workspaces, no per-view cudaMalloc/cudaFree; "
                              "the as-shipped call sequence (debug=true, per-view malloc/free) is in `as_shipped`",
                   "l2": "inputs larger than L2"},
        "as_shipped": res["as_shipped"],
        "render": render_stats,
        "clocks": clocks,
        "cpu_baseline": {"value": best["views_per_s"], "unit": UNIT, "cores": 1, "kind": "reference",
                         "sample": "the reference has no CPU implementation of this path; this is its own CUDA code "
                                   "(lib/diff-gaussian-rasterization + the Trainer.cu kernels) recompiled for sm_100 and run "
                                   "on the same B200, one host thread, full workload"},
        "e2e": {"value": best["views_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(out)


def main():
    args = parse_args()
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gsb(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
