"""The vocab-sharded path on real GPUs (NCCL, one process per GPU) against the single-process
oracle model.  Needs at least 2 GPUs; skipped otherwise."""
import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import shard_worker
from test_sharded_gloo import _reference

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["exchange", "peer"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_equal_single_process(criteo_lines, tmp_path, world, mode):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    lines, weights = criteo_lines
    rng = np.random.default_rng(5)
    cfg_kw = dict(vocabulary_size=200003, factor_num=16, batch_size=400, hash_feature_id=False, factor_lambda=2e-3,
                  bias_lambda=1e-3, learning_rate=0.05, adagrad_init_accumulator=0.1, loss_type="logistic")
    init = rng.uniform(-0.01, 0.01, (cfg_kw["vocabulary_size"], 17)).astype(np.float32)
    steps = 3
    port = 29700 + world + (50 if mode == "peer" else 0)
    mp.spawn(shard_worker.run, args=(world, port, "nccl", lines, weights, cfg_kw, init, steps, str(tmp_path), mode),
             nprocs=world, join=True)
    ref, exp_losses = _reference(lines, weights, cfg_kw, init, steps, world)
    np.testing.assert_allclose(np.load(tmp_path / "losses.npy"), exp_losses, rtol=1e-5)
    np.testing.assert_allclose(np.load(tmp_path / "table.npy"), ref.flat_table(), rtol=1e-4, atol=1e-7)
    got = np.concatenate([np.load(tmp_path / ("pred%d.npy" % r)) for r in range(world)])
    np.testing.assert_allclose(got, ref.predict(lines[:400]), rtol=1e-4, atol=1e-6)
