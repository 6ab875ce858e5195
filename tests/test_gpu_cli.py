"""run_tffm.py train / predict driven by a sample.cfg-shaped config (configs[0] in small)."""
import os
import re
import shutil
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _small_files(tmp_path):
    """2 x 1000 training lines + weights and 500 test lines cut from the committed excerpt of the reference's
    data/train_0 (the full fixtures in data/ are the cfg1 test's; these runs only need the command line to work)"""
    golden = os.path.join(ROOT, "tests", "golden")
    lines = open(os.path.join(golden, "criteo_excerpt.txt")).read().splitlines()
    weights = open(os.path.join(golden, "criteo_excerpt_weights.txt")).read().splitlines()
    for i in range(2):
        (tmp_path / ("train_%d" % i)).write_text("\n".join(lines[1000 * i:1000 * (i + 1)]) + "\n")
        (tmp_path / ("weight_%d" % i)).write_text("\n".join(weights[1000 * i:1000 * (i + 1)]) + "\n")
    (tmp_path / "test_0").write_text("\n".join(lines[1500:2000]) + "\n")


@pytest.mark.parametrize("loss_type", ["mse", "logistic"])
def test_train_then_predict_cli(tmp_path, loss_type):
    _small_files(tmp_path)
    cfg = open(os.path.join(ROOT, "sample.cfg")).read()
    cfg = cfg.replace("./data/", str(tmp_path) + "/").replace("./log", str(tmp_path / "log"))
    cfg = cfg.replace("batch_size = 50000", "batch_size = 400").replace("epoch_num = 20", "epoch_num = 3")
    cfg = cfg.replace("factor_num = 100", "factor_num = 20").replace("loss_type = mse", "loss_type = " + loss_type)
    cfg = cfg.replace("learning_rate = 0.01", "learning_rate = 0.05")
    (tmp_path / "t.cfg").write_text(cfg)
    env = dict(os.environ, PYTHONPATH=ROOT)
    # logistic run: the files stream through the GPU in segments (--stream-mb) instead of sitting in HBM
    extra = ["--stream-mb", "1"] if loss_type == "logistic" else []
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "train", str(tmp_path / "t.cfg"), "-m",
                          "-t", str(tmp_path / "trace")] + extra, capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    losses = [float(x) for x in re.findall(r"-- Global Step: \d+; Avg loss: ([0-9.]+);", out.stdout)]
    assert len(losses) == 15                                   # 3 epochs x 2000 lines / 400
    assert np.mean(losses[-3:]) < np.mean(losses[:3])
    assert "Average speed:" in out.stdout and "Model saved to" in out.stdout and "speed:" in out.stdout
    assert (tmp_path / "log" / "model.ckpt.pt").exists() and (tmp_path / "trace.json").exists()
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "predict", str(tmp_path / "t.cfg")],
                         capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    scores = [float(x) for x in open(tmp_path / "test_0_score")]
    assert len(scores) == 500 and np.isfinite(scores).all()
    # the restored model scores the same lines identically through the Python API
    import torch
    from tffm.fm_model import Model
    m = Model(str(tmp_path / "t.cfg")).build(seed=0)
    m.restore(str(tmp_path / "log" / "model.ckpt.pt"))
    lines = open(tmp_path / "test_0").read().splitlines()
    np.testing.assert_allclose(m.predict(lines).cpu().numpy(), np.float32(scores), rtol=1e-6)
    # generate: like the reference it wants --export_path; the exported table scores unlabeled "id:val ..." lines
    # (serving_parser / serving_scorer, tffm/fm_model.py:194-252) like predict scores the labeled ones
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "generate", str(tmp_path / "t.cfg")],
                         capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 2 and "Export path is not specified" in out.stdout
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "generate", str(tmp_path / "t.cfg"),
                          "--export_path", str(tmp_path / "export")], capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0 and "Exporting trained model to" in out.stdout, out.stdout + out.stderr
    from fast_tffm_b200.fm_ops import FmParseError
    from fast_tffm_b200.serving import ServingScorer
    srv = ServingScorer(str(tmp_path / "export"))
    assert srv.model.accum is None
    unlabeled = [ln.split(" ", 1)[1] for ln in lines]
    np.testing.assert_allclose(srv.score(unlabeled).cpu().numpy(), np.float32(scores), rtol=1e-6)
    assert float(srv.score([""])[0]) == 0.0
    with pytest.raises(FmParseError):
        srv.score(["3:1 4"])                      # the serving parser needs a value on every token
    with pytest.raises(RuntimeError):
        srv.model.train_step(lines[:4])


def test_validation_and_tolerance_stop_early(tmp_path):
    """validation_files + tolerance: the loop prints the validation loss every save_steps and stops under it
    (run_tffm.py:69-77 of the reference)"""
    _small_files(tmp_path)
    cfg = open(os.path.join(ROOT, "sample.cfg")).read()
    cfg = cfg.replace("./data/", str(tmp_path) + "/").replace("./log", str(tmp_path / "log"))
    cfg = cfg.replace("batch_size = 50000", "batch_size = 400").replace("epoch_num = 20", "epoch_num = 50")
    cfg = cfg.replace("factor_num = 100", "factor_num = 8").replace("save_steps = 10", "save_steps = 2")
    cfg = cfg.replace("#tolerance = 0.2", "tolerance = 1000.0")
    cfg = cfg.replace("#validation_files = " + str(tmp_path) + "/test_*", "validation_files = " + str(tmp_path) + "/test_*")
    (tmp_path / "t.cfg").write_text(cfg)
    assert "\nvalidation_files" in cfg and "\ntolerance" in cfg
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "train", str(tmp_path / "t.cfg")],
                         capture_output=True, text=True, env=dict(os.environ, PYTHONPATH=ROOT), cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "Preloading validation data..." in out.stdout
    assert re.search(r"validation loss at step 2: [0-9.]+", out.stdout)
    assert "Loss on validation data set is below tolerance. Training completed." in out.stdout
    assert len(re.findall(r"-- Global Step:", out.stdout)) == 2      # stopped at the first check


def test_train_under_torchrun_then_predict_on_one_gpu(tmp_path):
    """the replacement of the reference's --dist mode: torchrun, table sharded over peer memory;
    the checkpoint it writes is read back by the single-GPU predict"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _small_files(tmp_path)
    cfg = open(os.path.join(ROOT, "sample.cfg")).read()
    cfg = cfg.replace("./data/", str(tmp_path) + "/").replace("./log", str(tmp_path / "log"))
    cfg = cfg.replace("batch_size = 50000", "batch_size = 400").replace("epoch_num = 20", "epoch_num = 3")
    cfg = cfg.replace("factor_num = 100", "factor_num = 16").replace("learning_rate = 0.01", "learning_rate = 0.05")
    cfg = cfg.replace("#validation_files = " + str(tmp_path) + "/test_*", "validation_files = " + str(tmp_path) + "/test_*")
    cfg = cfg.replace("save_steps = 10", "save_steps = 5").replace("#tolerance = 0.2", "tolerance = 0.000001")
    (tmp_path / "t.cfg").write_text(cfg)
    env = dict(os.environ, PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29655", os.path.join(ROOT, "run_tffm.py"),
                          "train", str(tmp_path / "t.cfg")], capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    losses = [float(x) for x in re.findall(r"-- Global Step: \d+; Avg loss: ([0-9.]+);", out.stdout)]
    assert len(losses) == 15 and np.mean(losses[-3:]) < np.mean(losses[:3])     # printed once (rank 0), and learning
    v = [float(x) for x in re.findall(r"validation loss at step \d+: ([0-9.]+)", out.stdout)]
    assert len(v) == 3 and v[-1] < v[0]
    # every rank wrote its shard (= block r of a mod-2 partition); one GPU reads the set back
    assert all((tmp_path / "log" / ("model.ckpt.pt.shard-%d-of-2" % r)).exists() for r in range(2))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "predict", str(tmp_path / "t.cfg")],
                         capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    scores = [float(x) for x in open(tmp_path / "test_0_score")]
    assert len(scores) == 500 and np.isfinite(scores).all() and np.std(scores) > 0
    # a single-GPU run continues from the set and supersedes it with its own (newer) file
    out = subprocess.run([sys.executable, os.path.join(ROOT, "run_tffm.py"), "train", str(tmp_path / "t.cfg")],
                         capture_output=True, text=True, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    steps = [int(x) for x in re.findall(r"-- Global Step: (\d+);", out.stdout)]
    assert steps == list(range(16, 31)) and (tmp_path / "log" / "model.ckpt.pt").exists()
    from fast_tffm_b200.fm_model import latest_checkpoint
    assert latest_checkpoint(str(tmp_path / "log" / "model.ckpt.pt"))[0] == 30
