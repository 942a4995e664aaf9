"""Developer probe: where the end-to-end step (bench.py `e2e`) loses its ~1 ms per step against the HBM-resident step
(`value`).  One GPU, config C4 with `--cameras` cameras per sphere (4 -> 8 cameras = the per-rank load of the 8-GPU job,
where a fixed millisecond is 10 % of the step).  The e2e loop of bench.py is rebuilt with every ingredient switchable:

    loss     the fused loss reduction is on (gsb_trainer_enable_loss)
    read     the loss scalar of step k-1 is read every step
    upload   the next step's truth frames go up on a copy stream (one cudaMemcpyAsync from pinned memory)
    wait     the compute stream waits for this step's upload
    swap     Trainer.truthFrameBuffersW/B alternate between the two device sets
    order    the copy stream waits for the end of the step that read the set it overwrites (no write-after-read race on the
             truth frames; the upload then starts when that step's exchange and apply are done instead of under them)

and every combination named on the command line is timed (CUDA events around K steps).  Prints one JSON line.

    python tools/e2e_bisect.py --cameras 4 --steps 40 base loss loss+read upload upload+wait upload+wait+swap all
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cameras", type=int, default=4, help="cameras per sphere over all ranks (0: the config's 32 + 32)")
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--config", default="C4")
    ap.add_argument("variants", nargs="*", default=["base", "loss", "loss+read", "upload", "upload+wait", "upload+wait+swap", "all", "base"])
    a = ap.parse_args()
    a.impl, a.torch_allreduce = "gsb", False
    torch, dist, rank, world, local = bench.init_dist(a)
    import helpers as H
    gsb = H.gsb
    dev = f"cuda:{local}"
    scene = bench.build_scene(gsb, a, rank, world)
    comm = gsb.Comm.from_torch_distributed(dist, dev) if world > 1 else None
    W, Hh = scene["W"], scene["H"]
    project, mine = scene["project"], scene["mine"]
    L = gsb.lib()
    tw, tb, _ = bench.render_truths(torch, H, gsb, scene, dev, lambda sc, vw, *x: H.run_forward_gsb(sc, vw, *x)[1],
                                    lambda out, fb, w, h: L.gsb_image_float_to_int(out.data_ptr(), fb.data_ptr(), w, h, None))
    cloud = scene["cloud"]
    host = gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")])
    n = len(tw)
    host_all = torch.stack([t.cpu() for t in list(tw) + list(tb)]).pin_memory()
    dev_all = [torch.stack(list(tw) + list(tb)), torch.empty((2 * n,) + tuple(tw[0].shape), dtype=tw[0].dtype, device=dev)]
    dev_all[1].copy_(dev_all[0])
    dset = [([dev_all[w][i] for i in range(n)], [dev_all[w][n + i] for i in range(n)]) for w in (0, 1)]
    copy_stream = torch.cuda.Stream()
    up_ev = [torch.cuda.Event(), torch.cuda.Event()]
    n_total = len(scene["cams"])
    out = {"world": world, "cameras_per_rank": n, "views_per_step": 2 * n_total, "steps": a.steps, "variants": []}

    for name in a.variants:
        # "nocomm" (N > 1 only): no exchange, the ranks run as independent replicas on their own cameras
        on = set(name.split("+"))
        if "all" in on:
            on |= {"loss", "read", "upload", "wait", "swap"}
        tr = gsb.Trainer(W, Hh, dev)
        tr.set_total_cameras(n_total)
        if comm is not None and "nocomm" not in on:
            tr.set_comm(comm)
        tr.model = gsb.ModelSplatsDevice(host, dev)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = dset[0][0], dset[0][1], mine
        tr.enable_loss("loss" in on or "read" in on)
        host_ms = []

        done_ev = [torch.cuda.Event(), torch.cuda.Event()]

        def upload(which, after=None):
            if after is not None:
                copy_stream.wait_event(after)
            with torch.cuda.stream(copy_stream):
                dev_all[which].copy_(host_all, non_blocking=True)
                up_ev[which].record(copy_stream)

        def step(k, last=False):
            t0 = time.perf_counter()
            cur = k & 1
            if "wait" in on:
                torch.cuda.current_stream().wait_event(up_ev[cur])
            if "upload" in on:
                upload(cur ^ 1, done_ev[cur ^ 1] if "order" in on else None)
            if "swap" in on:
                tr.truthFrameBuffersW, tr.truthFrameBuffersB = dset[cur]
            tr.train(project, False)
            if "order" in on:
                done_ev[cur].record()        # legacy stream: everything step k enqueued (joined backward chains, exchange, apply)
            if "read" in on and k > 0:
                tr.loss(1)
            host_ms.append((time.perf_counter() - t0) * 1e3)

        upload(0)
        upload(1)
        torch.cuda.synchronize()
        for k in range(2 * ((a.warmup + 1) // 2)):
            step(k)
        host_ms.clear()
        ms = bench.timed_steps(torch, dist, step, a.steps)
        host_ms.sort()
        out["variants"].append({"name": name, "ms_per_step": ms / a.steps, "views_per_s": 2 * n_total * a.steps / (ms / 1e3),
                                "host_ms_median": host_ms[len(host_ms) // 2], "host_ms_max": host_ms[-1]})
        torch.cuda.synchronize()
        tr.close()
        if rank == 0:
            print(name, round(ms / a.steps, 4), "ms/step; host median", round(host_ms[len(host_ms) // 2], 3), file=sys.stderr)
    if rank == 0:
        print(json.dumps(out))
    if comm is not None:
        comm.close()
    if dist is not None and dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
