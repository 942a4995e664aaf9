"""Multi-GPU correctness of the CUDA path (SURVEY.md 7.4 last row; reference linearity src/Trainer.cu:47-77 and the
densify block :437-542).  Run under torchrun, one rank per GPU; also called by bench.py for every N > 1.

The job: k = 3 iterations of a camera-sharded trainer with the library's own NCCL exchange (gsb_comm_*), the second one
densifying (cull + split + clone all occur).  Checked:
  (i)   a 64-bit checksum of count + all five parameter arrays is bit-identical on every rank  -> replicas_identical
  (ii)  after the first iteration (sums of per-camera gradients + one linear update) every parameter array is within
        1e-5 relative L2 of the N = 1 run of the same job                                      -> vs_n1_rel_first_iteration
  (iii) after all iterations (same canonical densify order, equal counts) the distance to the N = 1 run -> vs_n1_rel,
        next to the distance between two N = 1 runs of the same job                            -> n1_rerun_rel
        (clones are displaced along normalize(mean location gradient), src/Trainer.cu:507-515, which turns the
        summation-order noise of float atomics into O(1e-4..1e-3) relative differences for splats whose gradients
        nearly cancel -- one GPU against itself shows the same, and that is the yardstick)
  (iv)  R summed over the ranks' view passes of the first iteration equals the N = 1 value    -> rendered_match

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/mgpu_check.py [--config C3] [--cameras 4]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

KEYS = ("locations", "shs", "scales", "opacities", "rotations")
DEVKEYS = ("devLocations", "devShs", "devScales", "devOpacities", "devRotations")


def _checksum(torch, model):
    """Position-weighted 64-bit sum of the raw float bits of every array, plus the count."""
    out = [int(model.count)]
    for k in DEVKEYS:
        v = getattr(model, k).view(torch.int32).to(torch.int64)
        w = (torch.arange(v.numel(), device=v.device, dtype=torch.int64) % 65521) + 1
        out.append(int((v * w).sum().item()))
    return out


def _gap_threshold(values, q):
    v = np.sort(np.asarray(values, dtype=np.float64))
    k = int(q * (len(v) - 1))
    lo, hi = max(1, k - 200), min(len(v) - 1, k + 200)
    j = lo + int(np.argmax(v[lo + 1:hi + 1] - v[lo:hi]))
    return float(0.5 * (v[j] + v[j + 1]))


def window_selftest(torch, dist, gsb, comm, dev, rank, world, floats=1000003):
    """The peer-window all-reduce on its own: random floats per rank, (a) bit-identical result on every rank,
    (b) equal to the float64 sum of the ranks' inputs rounded once, within the rounding of G-1 float32 additions,
    (c) a second exchange on the same window (flag sequence numbers) works too."""
    win = comm.window(floats)
    out = {"peer_mapped": win.peer_mapped, "floats": floats}
    worst, same = 0.0, True
    for rep in range(2):
        g = torch.Generator(device="cpu").manual_seed(1234 + 97 * rank + rep)
        x = (torch.rand(floats, generator=g) - 0.5).to(dev)
        exact = x.double().clone()
        dist.all_reduce(exact)
        win.tensor.copy_(x)
        torch.cuda.synchronize()
        win.allreduce_sum()
        torch.cuda.synchronize()
        y = win.tensor.clone()
        worst = max(worst, float((y.double() - exact).abs().max()))
        bits = y.view(torch.int32).to(torch.int64)
        cs = torch.stack([bits.sum(), (bits * (torch.arange(floats, device=dev) % 8191 + 1)).sum()])
        gathered = [torch.zeros_like(cs) for _ in range(world)]
        dist.all_gather(gathered, cs)
        same = same and all(bool(torch.equal(t, gathered[0])) for t in gathered)
    out.update({"max_abs_err_vs_f64": worst, "bit_identical": same, "ok": bool(same and worst <= world * 2.0 ** -24 * world)})
    win.close()
    return out


def exchange_microbench(torch, dist, gsb, comm, dev, world, floats, reps=20):
    """Back-to-back all-reduces of `floats` floats (rank skew amortised: each exchange re-synchronises the ranks):
    the library's peer-window kernel against ncclAllReduce on the same communicator.  ms per exchange, max over ranks."""
    win = comm.window(floats)
    other = torch.zeros(floats, dtype=torch.float32, device=dev)
    out = {"floats": int(floats), "mbytes": floats * 4 / 1e6, "reps": reps, "peer_mapped": win.peer_mapped}
    for name, fn in (("peer_ms", win.allreduce_sum), ("nccl_ms", lambda: comm.allreduce_sum(other))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name] = float(t.item())
    out["peer_algbw_gb_s"] = out["mbytes"] / out["peer_ms"]
    out["nccl_algbw_gb_s"] = out["mbytes"] / out["nccl_ms"]
    win.close()
    return out


def run_check(torch, dist, rank, world, local, config="C4", cameras=8, iterations=3, densify_at=1):
    import helpers as H
    from gpu_stage_probe import make_truths
    gsb = H.gsb
    dev = f"cuda:{local}"
    W, Hh, P = gsb.synth.CONFIGS[config]
    project = gsb.synth.project_for(config)
    project.sphere1.count = cameras
    project.sphere2.count = cameras if project.sphere2.count > 0 else 0
    # learning rates large enough for three iterations to move the parameters visibly in float32
    project.lrLocation, project.lrSh, project.lrScale, project.lrOpacity, project.lrRotation = 0.002, 0.01, 0.0005, 0.01, 0.0005
    cams = gsb.Camera.getCameras(project)
    n1 = project.sphere1.count
    seed = 0x5EED0000 + int(config[1])
    cloud = gsb.synth.splat_cloud(P, seed)
    teacher = gsb.synth.splat_cloud(P // 2, seed + 1)
    mine = gsb.synth.shard_cameras(cams, n1, rank, world)
    sq = (W == Hh)

    def trainer_for(camset, tw, tb, comm):
        tr = gsb.Trainer(W, Hh, dev, square_fov=sq)
        tr.set_total_cameras(len(cams))
        host = gsb.ModelSplatsHost.from_arrays(*[cloud[k] for k in KEYS])
        tr.model = gsb.ModelSplatsDevice(host, dev)
        tr.truthFrameBuffersW, tr.truthFrameBuffersB, tr.truthCameras = tw, tb, camset
        if comm is not None:
            tr.set_comm(comm)
        return tr

    import copy
    thr = torch.zeros(3, dtype=torch.float64, device=dev)
    all_tw = all_tb = None
    if rank == 0:
        # thresholds in gaps of the decision variables, taken from an N = 1 pre-run up to the densifying iteration
        all_tw, all_tb, _ = make_truths(teacher, cams, W, Hh, dev)
        pre = trainer_for(cams, all_tw, all_tb, None)
        avg = pre.grad_tensor()
        pj = copy.deepcopy(project)
        for _ in range(densify_at + 1):
            pre.train(pj, False)
        torch.cuda.synchronize()
        a = avg.cpu().numpy()
        m = gsb.ModelSplatsHost.from_device(pre.model)
        score = a[:P].astype(np.float64) - np.linalg.norm(a[P:4 * P].astype(np.float64).reshape(-1, 3), axis=1)
        thr[0] = _gap_threshold(score, 0.85)
        thr[1] = _gap_threshold(m.opacities, 0.08)
        thr[2] = _gap_threshold(np.linalg.norm(m.scales.reshape(-1, 3), axis=1), 0.5)
        pre.close()
    dist.broadcast(thr, src=0)
    project.paramDensifyVariance, project.paramCullOpacity, project.paramSplitSize = [float(x) for x in thr.cpu()]

    # ---- the N-rank job: camera shards, native NCCL exchange -------------------------------------------------
    comm = gsb.Comm.from_torch_distributed(dist, dev)
    selftest = window_selftest(torch, dist, gsb, comm, dev, rank, world)
    micro = exchange_microbench(torch, dist, gsb, comm, dev, world, 24 * P)
    if rank == 0:
        idx = [cams.index(c) for c in mine]
        tw, tb = [all_tw[i] for i in idx], [all_tb[i] for i in idx]
    else:
        tw, tb, _ = make_truths(teacher, mine, W, Hh, dev)
    tr = trainer_for(mine, tw, tb, comm)
    pj = copy.deepcopy(project)
    counts = []

    def run_job(t, p, counts_out):
        first, rendered_first = None, 0
        for it in range(iterations):
            t.train(p, it == densify_at)
            counts_out.append(int(t.model.count))
            if it == 0:
                torch.cuda.synchronize()
                first = {dk: getattr(t.model, dk).double().clone() for dk in DEVKEYS}
                rendered_first = int(t.stats()["rendered_total"])
        torch.cuda.synchronize()
        return first, rendered_first

    def rel_to(a, b):
        return float(torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double()))

    first, r_first = run_job(tr, pj, counts)
    path = tr.exchange_path()
    rendered = torch.tensor([r_first], dtype=torch.int64, device=dev)
    dist.all_reduce(rendered)
    cs = torch.tensor(_checksum(torch, tr.model), dtype=torch.int64, device=dev)
    gathered = [torch.zeros_like(cs) for _ in range(world)]
    dist.all_gather(gathered, cs)
    identical = all(bool(torch.equal(g, gathered[0])) for g in gathered)

    res = {"world": world, "config": config, "cameras_total": len(cams), "iterations": iterations,
           "densify_iteration": densify_at, "counts": counts, "replicas_identical": identical,
           "exchange_path": path, "window_selftest": selftest, "exchange_microbench": micro,
           "checksums": [[int(x) for x in g.cpu().tolist()] for g in gathered]}
    if rank == 0:
        # ---- the same job on ONE GPU --------------------------------------------------------------------------
        one = trainer_for(cams, all_tw, all_tb, None)
        counts1 = []
        first1, r_first1 = run_job(one, copy.deepcopy(project), counts1)
        again = trainer_for(cams, all_tw, all_tb, None)          # the yardstick: the same N = 1 job once more
        counts2 = []
        run_job(again, copy.deepcopy(project), counts2)
        rel, rerun = {}, {}
        rel_first = {k: rel_to(first[dk], first1[dk]) for k, dk in zip(KEYS, DEVKEYS)}
        if counts1 == counts:
            rel = {k: rel_to(getattr(tr.model, dk), getattr(one.model, dk)) for k, dk in zip(KEYS, DEVKEYS)}
        if counts1 == counts2:
            rerun = {k: rel_to(getattr(again.model, dk), getattr(one.model, dk)) for k, dk in zip(KEYS, DEVKEYS)}
        res.update({"counts_n1": counts1, "counts_n1_rerun": counts2,
                    "vs_n1_rel_first_iteration": max(rel_first.values()),
                    "vs_n1_rel": max(rel.values()) if rel else None, "vs_n1_rel_per_array": rel,
                    "n1_rerun_rel": max(rerun.values()) if rerun else None, "n1_rerun_rel_per_array": rerun,
                    "rendered_first_iteration": int(rendered.item()), "rendered_first_iteration_n1": r_first1,
                    "rendered_match": int(rendered.item()) == r_first1,
                    "densify_changed_count": counts[densify_at] != P,
                    "thresholds": [float(x) for x in thr.cpu()]})
        one.close()
        again.close()
    tr.close()
    comm.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C3")
    ap.add_argument("--cameras", type=int, default=4)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    res = run_check(torch, dist, rank, world, local, args.config, args.cameras)
    if rank == 0:
        print("MGPU_CHECK " + json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
