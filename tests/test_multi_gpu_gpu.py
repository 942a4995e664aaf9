"""GPU, world size 2 (skipped with fewer devices): the N > 1 CUDA path -- camera shards, the library's own NCCL
exchange (gsb_comm_* / gsb_window_*: two-shot kernel over NVLink peer windows, or ncclAllReduce), densify on the reduced means -- gives bit-identical replicas that equal the N = 1 run
(tests/mgpu_check.py has the protocol).  The C++ flavour of the same statement: compat_demo --devices 2 drives two
GPUs from one process through include/compat/Trainer.cuh with GSB_COMPAT_DEVICES."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu
needs2 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")


@needs2
@pytest.mark.parametrize("config,cameras,exchange", [("C3", 4, "peer"), ("C2", 4, "peer"), ("C2", 4, "nccl")])
def test_two_ranks_bit_identical_and_equal_to_one_rank(config, cameras, exchange):
    port = 29600 + (os.getpid() % 300)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_check.py"), "--config", config, "--cameras", str(cameras)]
    # the exchange is the library's own two-shot kernel over NVLink peer windows; GSB_PEER_EXCHANGE=0 sends the same
    # window through ncclAllReduce instead (the fallback every rank takes together when a peer cannot be mapped)
    env = dict(os.environ, GSB_PEER_EXCHANGE="0" if exchange == "nccl" else "1")
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, cwd=ROOT, env=env)
    lines = [l for l in r.stdout.splitlines() if l.startswith("MGPU_CHECK ")]
    assert r.returncode == 0 and lines, r.stdout[-4000:]
    res = json.loads(lines[-1][len("MGPU_CHECK "):])
    assert res["exchange_path"] == exchange, res
    assert res["window_selftest"]["ok"] and res["window_selftest"]["peer_mapped"] == (exchange == "peer"), res
    assert res["replicas_identical"], res
    assert res["counts"] == res["counts_n1"] and res["densify_changed_count"], res
    res.pop("checksums")
    assert res["rendered_match"], res
    assert res["vs_n1_rel_first_iteration"] <= 1e-5, res                 # linear sums + one update: summation order only
    # after densify + one more iteration the yardstick is one GPU against itself (see mgpu_check.py (iii))
    assert res["vs_n1_rel"] is not None and res["n1_rerun_rel"] is not None, res
    assert res["vs_n1_rel"] <= max(1e-5, 5.0 * res["n1_rerun_rel"]), res
    assert res["rendered_match"], res


@needs2
def test_compat_trainer_two_devices_in_one_process():
    demo = os.path.join(ROOT, "tests", "compat", "compat_demo")
    if not os.path.exists(demo):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "compat")], stdout=subprocess.DEVNULL)
    r = subprocess.run([demo, "--devices", "2"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["count_1"] == out["count_G"] and max(out[k] for k in out if k.startswith("rel_")) <= 1e-5, out
