"""Short driver for ncu: a few cfg2 train steps (device-resident batches), nothing else.
usage: python profiles/run_step.py [steps] [dist] [path] [factor_num] [vocab_log2] [batch]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_data  # noqa: E402
from bench import CFG  # noqa: E402
from fast_tffm_b200.fm_model import Model  # noqa: E402
from fast_tffm_b200.fm_ops import PackedLines  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dist = sys.argv[2] if len(sys.argv) > 2 else "uniform"
path = sys.argv[3] if len(sys.argv) > 3 else "sorted"
k = int(sys.argv[4]) if len(sys.argv) > 4 else CFG["factor_num"]
vocab = 1 << int(sys.argv[5]) if len(sys.argv) > 5 else CFG["vocab"]
batch = int(sys.argv[6]) if len(sys.argv) > 6 else CFG["batch"]
dev = torch.device("cuda", 0)
m = Model(vocabulary_size=vocab, factor_num=k, batch_size=batch, hash_feature_id=True,
          loss_type=CFG["loss_type"], learning_rate=CFG["learning_rate"], adagrad_init_accumulator=CFG["init_acc"],
          factor_lambda=CFG["factor_lambda"], bias_lambda=CFG["bias_lambda"]).build(
              dev, seed=0, max_nnz_per_line=39, layout="pair" if path == "sorted" else "interleaved", path=path)
batches = []
for i in range(2):
    text, off, _ = bench_data.criteo_batch(batch, 1 + i, dist)
    batches.append(PackedLines(torch.from_numpy(bench_data.pad16(text)).to(dev), torch.from_numpy(off).to(dev)))
for s in range(steps):
    b = m.train_step_async(batches[s % 2], path=path)
torch.cuda.synchronize()
st = b["status"].cpu()
print("ok", int(st[0]), int(st[1]), int(st[4]), float(st[6:8].view(torch.float32)[0]))
