"""Writes tests/golden/oracle_small.json: digest + sparse samples of the oracle's output on the
seeded small scene used by tests/test_oracle_kat.py.  Run after an INTENDED change of the oracle
or of the synthetic generator:   python tests/golden/make_oracle_golden.py"""
import hashlib
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
from helpers import SMALL_CAM, run_oracle  # noqa: E402
from oracle import pm_oracle as po  # noqa: E402

ge.build()
ge.load_package()
synth = importlib.import_module(ge.PKG_NAME + ".synth")
pipeline = importlib.import_module(ge.PKG_NAME + ".pipeline")
scene = synth.Scene(SMALL_CAM, seed=7)
plan = pipeline.SequencePlan(scene, 8, 6)
images = {k: scene.render(k) for k in plan.ids}
ora = run_oracle(po, plan, images)
h = hashlib.sha256()
for k in plan.ids:
    for w in (po.GRAD, po.THETA, po.DEPTH, po.SIGMA, po.DEPTH_CHECKED, po.POINTS):
        h.update(ora.plane(k, w).tobytes())
mid = plan.ids[len(plan.ids) // 2]
d = ora.plane(mid, po.DEPTH_CHECKED)
ys, xs = np.nonzero(d > 1e-6)
pick = np.linspace(0, len(ys) - 1, 64).astype(int)
out = {
    "images_sha256": hashlib.sha256(b"".join(images[k].tobytes() for k in plan.ids)).hexdigest(),
    "planes_sha256": h.hexdigest(),
    "mid_valid": int((d > 1e-6).sum()),
    "mid_samples": [[int(ys[i]), int(xs[i]), float(d[ys[i], xs[i]])] for i in pick],
}
with open(os.path.join(HERE, "oracle_small.json"), "w") as f:
    json.dump(out, f, indent=1)
print("valid pixels in mid key frame:", out["mid_valid"], "of", d.size)
