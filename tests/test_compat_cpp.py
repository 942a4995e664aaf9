"""The C++ drop-in boundary: include/compat/{rasterizer.h,Trainer.cuh,Trainer.cpp} compile with a plain
host compiler against stand-ins of the application's own headers (tests/compat/app/, same public
interfaces as src/Camera.h, src/Project.h, src/ModelSplats*.h) and link against libgsb200.so.  On the GPU
the resulting program trains through the reference's `Trainer` interface only."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO_DIR = os.path.join(ROOT, "tests", "compat")
DEMO = os.path.join(DEMO_DIR, "compat_demo")


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "gaussian-splatterer_b200")], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", DEMO_DIR], stdout=subprocess.DEVNULL)


def test_compat_headers_compile_and_link():
    _build()
    assert os.path.exists(DEMO)
    syms = subprocess.run(["nm", "-D", "--undefined-only", DEMO], stdout=subprocess.PIPE, text=True).stdout
    # the reference-shaped classes resolve to the C ABI and nothing else of ours
    for s in ("gsb_trainer_create", "gsb_trainer_train", "gsb_trainer_render", "gsb_trainer_densify_commit",
              "gsb_rasterizer_forward", "gsb_last_error"):
        assert s in syms, s


@pytest.mark.gpu
def test_compat_demo_trains_through_the_reference_interface():
    if not os.path.exists(DEMO):
        _build()
    r = subprocess.run([DEMO], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["err1"] < 0.8 * out["err0"] and out["count1"] > out["count0"] and out["threw_without_truths"] and out["finite"]
