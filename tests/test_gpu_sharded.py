"""The vocab-sharded path on real GPUs (NCCL, one process per GPU) against the single-process
oracle model.  Needs at least 2 GPUs; skipped otherwise."""
import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import shard_worker
from parity_util import RTOL, close, pred_scale
from test_sharded_gloo import _reference

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["exchange", "peer", "merge"])
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
    port = 29700 + world + {"exchange": 0, "peer": 50, "merge": 25}[mode]
    mp.spawn(shard_worker.run, args=(world, port, "nccl", lines, weights, cfg_kw, init, steps, str(tmp_path), mode),
             nprocs=world, join=True)
    ref, exp_losses = _reference(lines, weights, cfg_kw, init, steps, world)
    np.testing.assert_allclose(np.load(tmp_path / "losses.npy"), exp_losses, rtol=RTOL)
    close(np.load(tmp_path / "table.npy"), ref.flat_table(), 0.05 * steps)
    got = np.concatenate([np.load(tmp_path / ("pred%d.npy" % r)) for r in range(world)])
    f = ref.forward(lines[:400])
    close(got, f["pred"], pred_scale(f))


@pytest.mark.parametrize("mode,k", [("peer", 5), ("peer", 100), ("peer", 400), ("exchange", 100), ("exchange", 400),
                                    ("merge", 5), ("merge", 100)])
def test_sharded_steps_other_factor_nums(criteo_lines, tmp_path, mode, k):
    """row widths beyond the benchmark's: 2 and 26 chunks per row, and the wide-row kernels (K > 378)"""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines, weights = criteo_lines
    rng = np.random.default_rng(50 + k)
    cfg_kw = dict(vocabulary_size=170003, factor_num=k, batch_size=200, hash_feature_id=False, factor_lambda=2e-3,
                  bias_lambda=1e-3, learning_rate=0.05, adagrad_init_accumulator=0.1, loss_type="mse")
    init = rng.uniform(-0.01, 0.01, (cfg_kw["vocabulary_size"], k + 1)).astype(np.float32)
    port = 29820 + (k % 97) + {"exchange": 0, "peer": 100, "merge": 200}[mode]
    mp.spawn(shard_worker.run, args=(2, port, "nccl", lines, weights, cfg_kw, init, 2, str(tmp_path), mode),
             nprocs=2, join=True)
    ref, exp_losses = _reference(lines, weights, cfg_kw, init, 2, 2)
    np.testing.assert_allclose(np.load(tmp_path / "losses.npy"), exp_losses, rtol=RTOL)
    close(np.load(tmp_path / "table.npy"), ref.flat_table(), 0.05 * 2)
    got = np.concatenate([np.load(tmp_path / ("pred%d.npy" % r)) for r in range(2)])
    f = ref.forward(lines[:200])
    close(got, f["pred"], pred_scale(f))


@pytest.mark.parametrize("mode", ["exchange", "peer", "merge"])
def test_parse_error_on_one_rank_voids_the_step_everywhere(criteo_lines, tmp_path, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines, _ = criteo_lines
    rng = np.random.default_rng(6)
    cfg_kw = dict(vocabulary_size=200003, factor_num=8, batch_size=200, hash_feature_id=False, factor_lambda=0.0,
                  bias_lambda=0.0, learning_rate=0.05, adagrad_init_accumulator=0.1, loss_type="logistic")
    init = rng.uniform(-0.01, 0.01, (cfg_kw["vocabulary_size"], 9)).astype(np.float32)
    port = 29790 + {"exchange": 0, "peer": 1, "merge": 2}[mode]
    mp.spawn(shard_worker.run_bad_line, args=(2, port, lines, cfg_kw, init, str(tmp_path), mode), nprocs=2, join=True)
    for r in range(2):
        raised, loss = np.load(tmp_path / ("bad%d.npy" % r))
        assert raised == 1.0 and np.isfinite(loss)
    np.testing.assert_array_equal(np.load(tmp_path / "after_bad.npy"), init)        # nothing moved
    ref, exp = _reference(lines, [1.0] * len(lines), cfg_kw, init, 0, 2)
    e = ref.train_step(lines[:200])
    close(np.load(tmp_path / "after_good.npy"), ref.flat_table(), 0.05)


def test_sharded_checkpoint_set(criteo_lines, tmp_path):
    """every rank writes block r of a mod-G partition in the background; a sharded model and a single-GPU
    Model both restore the table as it was when save() was called"""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lines, _ = criteo_lines
    rng = np.random.default_rng(8)
    cfg_kw = dict(vocabulary_size=170003, factor_num=10, batch_size=200, hash_feature_id=False, factor_lambda=1e-3,
                  bias_lambda=1e-3, learning_rate=0.05, adagrad_init_accumulator=0.1, loss_type="logistic",
                  vocabulary_block_num=7)
    init = rng.uniform(-0.01, 0.01, (cfg_kw["vocabulary_size"], 11)).astype(np.float32)
    mp.spawn(shard_worker.run_checkpoint, args=(2, 29871, lines, cfg_kw, init, str(tmp_path)), nprocs=2, join=True)
    at_save = np.load(tmp_path / "at_save.npy")
    assert not np.array_equal(at_save, init)
    np.testing.assert_array_equal(np.load(tmp_path / "restored.npy"), at_save)
    assert int(np.load(tmp_path / "step.npy")[0]) == 2
    from fast_tffm_b200.fm_model import Model
    m = Model(**cfg_kw).build("cuda:0", seed=3)
    m.restore(str(tmp_path / "ckpt" / "model.ckpt.pt"))
    np.testing.assert_array_equal(m.table[:, :11].cpu().numpy(), at_save)
