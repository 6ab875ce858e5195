"""The reference's `generate` task, without TensorFlow.

`run_tffm.py generate cfg --export_path DIR` (reference: run_tffm.py:93-118, :196-202) restores the newest
checkpoint and writes a TF SavedModel whose graph is serving_parser + serving_scorer (tffm/fm_model.py:194-252):
input `data_lines`, lines of "id:val id:val ..." -- NO label, ids are numbers and are never hashed, every token has
its value -- output `scores`, the FM score written out densely
    0.5 * sum_k (sum_j v_jk x_j)^2 - 0.5 * sum_jk (v_jk x_j)^2 + sum_j b_j x_j
which is fm_scorer's sparse formula (cc/fm_scorer_op.h:60-86) on the same rows.  A SavedModel needs the TF runtime;
what can be kept is the artefact's content and its contract: export_serving() writes the weight table alone
(`fm_table.f32`, [vocabulary_size, K+1] float32 row-major, the vocab_block relabelling undone) beside a small
`fm_serving.json`, and ServingScorer loads that directory onto one GPU and scores `data_lines` through the same
parser and scorer kernels as `predict` (no Adagrad state, a third of the training footprint).
"""
import json
import os
import re

import numpy as np
import torch

from .fm_model import Model
from .fm_ops import FmParseError

META, TABLE = "fm_serving.json", "fm_table.f32"
_PAIRS = re.compile(r"[^ :]+:[^ :]+(?: [^ :]+:[^ :]+)*")      # "id:val id:val ...": one colon in every token
_ROWS_PER_WRITE = 1 << 20


def export_serving(model, export_path):
    """model: a built (and restored) Model.  Writes export_path/{fm_serving.json, fm_table.f32}."""
    os.makedirs(export_path, exist_ok=True)
    k, v = model.factor_num, model.vocabulary_size
    with open(os.path.join(export_path, TABLE + ".tmp"), "wb") as f:
        for r0 in range(0, v, _ROWS_PER_WRITE):          # bounded host copies: the table may be tens of GB
            f.write(model.table[r0:r0 + _ROWS_PER_WRITE, :k + 1].contiguous().cpu().numpy().tobytes())
    os.replace(os.path.join(export_path, TABLE + ".tmp"), os.path.join(export_path, TABLE))
    meta = {"format": 1, "vocabulary_size": v, "factor_num": k, "global_step": int(model.global_step),
            "signature": {"inputs": {"data_lines": "lines of 'id:val id:val ...' (no label, numeric ids, not hashed)"},
                          "outputs": {"scores": "float32 per line"}}}
    with open(os.path.join(export_path, META), "w") as f:
        json.dump(meta, f, indent=1)
    return meta


class ServingScorer:
    """scores = ServingScorer(export_dir).score(data_lines): the exported model's `predict` signature."""

    def __init__(self, export_path, device=None):
        with open(os.path.join(export_path, META)) as f:
            self.meta = json.load(f)
        v, k = int(self.meta["vocabulary_size"]), int(self.meta["factor_num"])
        flat = np.memmap(os.path.join(export_path, TABLE), dtype=np.float32, mode="r", shape=(v, k + 1))
        self.model = Model(vocabulary_size=v, factor_num=k, batch_size=1, hash_feature_id=False)
        self.model.build(device, seed=0, inference=True)
        for r0 in range(0, v, _ROWS_PER_WRITE):          # bounded host copies, as in export_serving
            rows = torch.from_numpy(np.array(flat[r0:r0 + _ROWS_PER_WRITE]))      # a writable copy of the slice
            self.model.table[r0:r0 + rows.shape[0], :k + 1] = rows.to(self.model.device)

    def score(self, data_lines):
        """data_lines: iterable of str.  Like serving_parser (tffm/fm_model.py:194-220): every token is "id:val";
        an empty line scores 0 (the dense formula over nothing)."""
        lines = []
        for i, line in enumerate(data_lines):
            line = " ".join(t for t in line.split(" ") if t)     # tf.string_split(' ') skips empty tokens
            if line and _PAIRS.fullmatch(line) is None:          # tf.reshape(split(':').values, [-1, 2]) fails on these
                raise FmParseError(2, i, line)
            lines.append("0 " + line if line else "0")
        if not lines:
            return torch.empty(0, dtype=torch.float32, device=self.model.device)
        return self.model.predict(lines)
