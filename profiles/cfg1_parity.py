"""cfg1 (sample.cfg shape: V=800000, nb=10, K=100, lr 0.01, acc0 0.1, lambda 0; mse with weight files as shipped,
and logistic) on the Criteo sample excerpt: loss per step, GPU path vs the CPU oracle model, same initial table, fixed
batch order.  Writes one JSON object (run on the GPU box: python profiles/cfg1_parity.py > gpurun_out/cfg1_parity.json)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fast_tffm_b200.fm_model import Model  # noqa: E402
from oracle import oracle as O  # noqa: E402

lines = [l.rstrip("\n") for l in open(os.path.join(ROOT, "tests", "golden", "criteo_excerpt.txt"))]
weights = np.float32([float(l) for l in open(os.path.join(ROOT, "tests", "golden", "criteo_excerpt_weights.txt"))])
vocab, nb, k, bs, epochs = 800000, 10, 100, 500, 5
init = np.random.default_rng(0).uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
out = {"config": "sample.cfg shape on tests/golden/criteo_excerpt.txt (2000 lines), batch 500, %d epochs" % epochs}
for loss_type, use_w in (("mse", True), ("logistic", False)):
    ref = O.RefModel(vocab, nb, k, init, batch_size=bs, lr=0.01, init_acc=0.1, loss_type=loss_type)
    rows = {}
    for path in Model.PATHS:
        m = Model(vocabulary_size=vocab, vocabulary_block_num=nb, factor_num=k, batch_size=bs, learning_rate=0.01,
                  adagrad_init_accumulator=0.1, loss_type=loss_type).build("cuda:0", init_table=init)
        rows[path] = (m, [])
    exp = []
    for e in range(epochs):
        for i in range(0, len(lines), bs):
            w = weights[i:i + bs] if use_w else None
            exp.append(ref.train_step(lines[i:i + bs], w)["loss"])
            for path, (m, got) in rows.items():
                got.append(m.train_step(lines[i:i + bs], w, path=path))
    res = {"oracle_loss_first_last": [exp[0], exp[-1]], "steps": len(exp)}
    for path, (m, got) in rows.items():
        rel = np.abs(np.float64(got) - np.float64(exp)) / np.abs(np.float64(exp))
        tab = m.table[:, :k + 1].cpu().numpy()
        ref_tab = ref.flat_table()
        touched = np.abs(ref_tab - init).max(1) > 0
        res[path] = {"loss_first_last": [got[0], got[-1]], "max_rel_diff_loss": float(rel.max()),
                     "max_abs_diff_table": float(np.abs(tab - ref_tab).max()),
                     "max_rel_diff_table_touched_rows": float((np.abs(tab - ref_tab)[touched] /
                                                               (np.abs(ref_tab[touched]) + 1e-6)).max())}
    out[loss_type + ("+weights" if use_w else "")] = res
print(json.dumps(out, indent=1))
