"""Truth ray caster timing probe: python tools/gpu_rtx_probe.py  (prints Mray/s at the reference's truth resolution)."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
gsb = importlib.import_module("gaussian-splatterer_b200")
cams = gsb.Camera.getCameras(gsb.synth.project_for("C3"))
print(json.dumps(bench.measure_truth_tracer(torch, gsb, "cuda", cams[:4])))
