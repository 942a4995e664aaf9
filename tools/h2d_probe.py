"""Measures the host -> device ceiling the end-to-end bench runs against: one pinned block of
`--mb` MB per GPU copied to the device with one cudaMemcpyAsync per repetition, every GPU alone and all GPUs at once,
with the block allocated (a) wherever the process happens to run and (b) on the GPU's own NUMA node
(gaussian-splatterer_b200/numa.py).  One process, one thread + stream per GPU.  Prints one JSON line.

    python tools/h2d_probe.py --mb 133 --reps 20
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

numa = importlib.import_module("gaussian-splatterer_b200.numa")


def bus_id(i):
    p = torch.cuda.get_device_properties(i)
    try:
        return "%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    except Exception:
        return None


def alloc(i, nbytes, local):
    if local:
        with numa.local_to_gpu(bus_id(i)) as ctx:
            t = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            t.fill_(1)
            return t, ctx.info
    t = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    t.fill_(1)
    return t, None


def run(gpus, hosts, devs, reps):
    """All listed GPUs copy concurrently; returns GB/s per GPU (CUDA events per stream)."""
    out = {}
    barrier = threading.Barrier(len(gpus))

    def work(i):
        torch.cuda.set_device(i)
        st = torch.cuda.Stream(device=i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            devs[i].copy_(hosts[i], non_blocking=True)
            st.synchronize()
            barrier.wait()
            e0.record(st)
            for _ in range(reps):
                devs[i].copy_(hosts[i], non_blocking=True)
            e1.record(st)
            st.synchronize()
        out[i] = hosts[i].numel() * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    th = [threading.Thread(target=work, args=(i,)) for i in gpus]
    [t.start() for t in th]
    [t.join() for t in th]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=133.0)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    n = torch.cuda.device_count()
    nbytes = int(a.mb * 1e6)
    res = {"gpus": n, "mb_per_copy": a.mb, "reps": a.reps, "host_cpus": os.cpu_count(),
           "affinity": sorted(os.sched_getaffinity(0))[:4] + ["..."] + sorted(os.sched_getaffinity(0))[-2:]}
    try:
        res["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], stdout=subprocess.PIPE, text=True, timeout=20).stdout.splitlines()[:n + 2]
    except Exception:
        res["topo"] = None
    for label, local in (("default_placement", False), ("numa_local", True)):
        hosts, devs, infos = {}, {}, {}
        for i in range(n):
            hosts[i], infos[i] = alloc(i, nbytes, local)
            devs[i] = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{i}")
        alone = {}
        for i in range(n):
            alone.update(run([i], hosts, devs, a.reps))
        together = run(list(range(n)), hosts, devs, a.reps)
        res[label] = {"alone_gb_s": [round(alone[i], 1) for i in range(n)], "concurrent_gb_s": [round(together[i], 1) for i in range(n)],
                      "concurrent_total_gb_s": round(sum(together.values()), 1), "placement": [infos[i] for i in range(n)] if local else None}
        del hosts, devs
    print(json.dumps(res))


if __name__ == "__main__":
    main()
