"""world_size-2 run of the vocab-sharded path on CPU (gloo): the host logic -- owner routing,
all-to-all split sizes, the order of owner-side combine and Adagrad -- must reproduce the
single-process reference model on the same global batch."""
import numpy as np
import pytest
import torch.multiprocessing as mp

import shard_worker
from oracle import oracle as O


def _reference(lines, weights, cfg_kw, init, steps, world):
    ref = O.RefModel(cfg_kw["vocabulary_size"], 1, cfg_kw["factor_num"], init, batch_size=cfg_kw["batch_size"],
                     hash_feature_id=cfg_kw["hash_feature_id"], factor_lambda=cfg_kw["factor_lambda"],
                     bias_lambda=cfg_kw["bias_lambda"], lr=cfg_kw["learning_rate"],
                     init_acc=cfg_kw["adagrad_init_accumulator"], loss_type=cfg_kw["loss_type"])
    bs = cfg_kw["batch_size"]
    losses = []
    for s in range(steps):
        sl = slice(s * bs, (s + 1) * bs)
        losses.append(ref.train_step(lines[sl], None if s % 2 else np.float32(weights[sl]))["loss"])
    return ref, losses


@pytest.mark.parametrize("loss_type,hashed", [("logistic", False), ("mse", True)])
def test_two_ranks_equal_single_process(criteo_lines, tmp_path, loss_type, hashed):
    lines, weights = criteo_lines
    rng = np.random.default_rng(3)
    cfg_kw = dict(vocabulary_size=200003, factor_num=6, batch_size=200, hash_feature_id=hashed, factor_lambda=2e-3,
                  bias_lambda=1e-3, learning_rate=0.05, adagrad_init_accumulator=0.1, loss_type=loss_type)
    init = rng.uniform(-0.01, 0.01, (cfg_kw["vocabulary_size"], 7)).astype(np.float32)
    steps, world = 3, 2
    port = 29600 + (1 if hashed else 0)
    mp.spawn(shard_worker.run, args=(world, port, "gloo", lines, weights, cfg_kw, init, steps, str(tmp_path)),
             nprocs=world, join=True)
    ref, exp_losses = _reference(lines, weights, cfg_kw, init, steps, world)
    np.testing.assert_allclose(np.load(tmp_path / "losses.npy"), exp_losses, rtol=1e-5)
    np.testing.assert_allclose(np.load(tmp_path / "table.npy"), ref.flat_table(), rtol=1e-4, atol=1e-7)
    exp_pred = ref.predict(lines[:200])
    got = np.concatenate([np.load(tmp_path / ("pred%d.npy" % r)) for r in range(world)])
    np.testing.assert_allclose(got, exp_pred, rtol=1e-4, atol=1e-6)
