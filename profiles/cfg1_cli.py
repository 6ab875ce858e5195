"""cfg1 through the command line: sample.cfg as shipped (V=800000, nb=10, K=100, batch 50000, mse + weight files,
20 epochs) on a 50 000-line training set made of the Criteo sample excerpt repeated 25 times (the shipped fixture's
size: 5 files x 10 000 lines), then `logistic`; `run_tffm.py train` is run as a user would and its own
"Average speed" line is what is reported.  Beside it: one step of the restated reference CPU ops at the same shape.
Run on the GPU box:  python profiles/cfg1_cli.py > gpurun_out/cfg1_cli.json
"""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

lines = open(os.path.join(ROOT, "tests", "golden", "criteo_excerpt.txt")).read().splitlines()
weights = open(os.path.join(ROOT, "tests", "golden", "criteo_excerpt_weights.txt")).read().splitlines()
out = {"config": "sample.cfg as shipped on 5 files x 10 000 lines (Criteo excerpt x 25), 20 epochs = 20 steps of 50 000"}
with tempfile.TemporaryDirectory() as tmp:
    for f in range(5):
        with open(os.path.join(tmp, "train_%d" % f), "w") as d, open(os.path.join(tmp, "weight_%d" % f), "w") as w:
            for r in range(5):
                d.write("\n".join(lines) + "\n")
                w.write("\n".join(weights) + "\n")
    with open(os.path.join(tmp, "test_0"), "w") as d:
        d.write("\n".join(lines[:1000]) + "\n")
    for loss_type in ("mse", "logistic", "mse, 400 epochs"):
        cfg = open(os.path.join(ROOT, "sample.cfg")).read()
        cfg = cfg.replace("./data/", tmp + "/").replace("./log", os.path.join(tmp, "log_" + loss_type[:3]))
        cfg = cfg.replace("loss_type = mse", "loss_type = " + loss_type.split(",")[0])
        if "epochs" in loss_type:      # a run long enough for the start-up costs to amortise
            cfg = cfg.replace("epoch_num = 20", "epoch_num = 400")
        path = os.path.join(tmp, loss_type[:3] + str(len(loss_type)) + ".cfg")
        open(path, "w").write(cfg)
        t0 = time.time()
        run = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "train", path], capture_output=True,
                             text=True, cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT))
        wall = time.time() - t0
        assert run.returncode == 0, run.stdout[-2000:] + run.stderr[-2000:]
        losses = [float(x) for x in re.findall(r"Avg loss: ([0-9.]+);", run.stdout)]
        if "epochs" in loss_type:
            out[loss_type] = {"cli_average_speed_examples_per_s": float(re.search(r"Average speed:\s+([0-9.e+]+)", run.stdout).group(1)),
                              "steps": len(losses), "loss_first_last": [losses[0], losses[-1]],
                              "train_process_wall_s": round(wall, 2)}
            continue
        speed = float(re.search(r"Average speed:\s+([0-9.e+]+)", run.stdout).group(1))
        t0 = time.time()
        pr = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "predict", path], capture_output=True,
                            text=True, cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT))
        assert pr.returncode == 0, pr.stdout[-2000:] + pr.stderr[-2000:]
        out[loss_type] = {"cli_average_speed_examples_per_s": speed, "steps": len(losses), "loss_first_last":
                          [losses[0], losses[-1]], "train_process_wall_s": round(wall, 2),
                          "predict_process_wall_s": round(time.time() - t0, 2)}

# the restated reference CPU ops on one step of the same shape (50 000 lines, K=100), all host cores
from oracle import oracle as O  # noqa: E402
from fast_tffm_b200.fm_ops import pack_lines  # noqa: E402
batch = (lines * 25)[:50000]
p = pack_lines(batch)
vocab, k = 800000, 100
rng = np.random.default_rng(0)
blocks = (rng.random((1, vocab + 1, k + 1), dtype=np.float32) - 0.5) * 0.02
acc = np.full_like(blocks, np.float32(0.1))
threads = os.cpu_count() or 1
best = None
for _ in range(2):
    t0 = time.perf_counter()
    _, stage, _ = O.train_step_faithful(p.text.numpy(), p.offsets.numpy(), vocab, False, 1, k, blocks, acc,
                                        np.float32(([1.0, 2.0] * 25000)), "mse", 0.0, 0.0, 0.01, 50000, threads)
    dt = time.perf_counter() - t0
    best = dt if best is None or dt < best else best
out["cpu_port_one_step"] = {"seconds": round(best, 3), "examples_per_s": round(50000 / best, 1), "cores": threads,
                            "stage_s": [round(float(x), 4) for x in stage]}
print(json.dumps(out, indent=1))
