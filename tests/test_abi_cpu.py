"""CPU-only checks of the drop-in boundary: the library builds, loads and exports
every symbol include/fm_b200.h declares; the host-side pieces behave; and the
product refuses to run without CUDA instead of falling back."""
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from fast_tffm_b200 import _lib, build, fm_ops
from fast_tffm_b200.fm_model import Model

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "fm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fm_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_for_sm100a_and_exports_the_header():
    lib_path = build.build()
    assert os.path.exists(lib_path)
    declared = _declared_functions()
    assert "fm_score" in declared and "fm_parse_fill" in declared
    out = subprocess.check_output(["nm", "-D", "--defined-only", lib_path], text=True)
    exported = set(re.findall(r" T (fm_[a-z0-9_]+)", out))
    assert set(declared) <= exported, sorted(set(declared) - exported)
    assert set(_lib.exported_symbols()) == set(declared)
    lib = _lib.load()
    assert lib.fm_abi_version() == 1
    # the cubin inside is sm_100a
    dump = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in dump


def test_workspace_queries_need_no_gpu():
    lib = _lib.load()
    assert lib.fm_score_workspace(65536) >= 65536 // 8 * 4
    assert lib.fm_unique_workspace(1000) > 1000 * 4
    assert lib.fm_parse_workspace(10) > 0 and lib.fm_loss_workspace(10) > 0 and lib.fm_split_lines_workspace(100) > 0


def test_bad_arguments_are_rejected_before_any_launch():
    lib = _lib.load()
    rc = lib.fm_score(None, None, 4, None, None, 8, 16, 0.0, 0.0, None, None, None, 0, None)
    assert rc == -1 and b"fm_score" in lib.fm_last_error()
    rc = lib.fm_unique(None, -1, None, None, None, None, None, 0, None)
    assert rc == -1


def test_pack_lines_layout():
    p = fm_ops.pack_lines(["1 2:3", b"0 4", ""])
    assert p.offsets.tolist() == [0, 6, 10, 11]
    assert p.text.numel() % 16 == 0 and p.text.numel() >= 11 + 16
    assert bytes(p.text[:11].numpy().tobytes()) == b"1 2:3\n0 4\n\n"
    assert p.line(0) == "1 2:3" and p.line(2) == ""


def test_no_cpu_fallback():
    with pytest.raises(_lib.FmB200Error, match="no CPU fallback"):
        fm_ops.fm_scorer([0], torch.zeros(1, 3), [1.0], [0, 1], 0.0, 0.0)
    with pytest.raises(_lib.FmB200Error):
        fm_ops.fm_unique(torch.zeros(3, dtype=torch.int64))
    if not torch.cuda.is_available():
        with pytest.raises(_lib.FmB200Error, match="no CPU fallback"):
            fm_ops.fm_parser(["1 2:3"], 10)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libfm_b200.so")
    with pytest.raises(_lib.FmB200Error, match="no CPU fallback"):
        _lib.load()


def test_config_reader_matches_reference_keys(tmp_path):
    (tmp_path / "a_train").write_text("1 1:1\n")
    (tmp_path / "a_weight").write_text("1\n")
    cfg = tmp_path / "x.cfg"
    cfg.write_text("""[General]
vocabulary_size = 800000
vocabulary_block_num = 10
factor_num = 100
model_file = ./model
hash_feature_id = False
log_dir = ./log
save_summaries_steps = 1
[Train]
batch_size = 50000
init_value_range = 0.01
factor_lambda = 0
bias_lambda = 0
epoch_num = 20
learning_rate = 0.01
adagrad.initial_accumulator = 0.1
save_steps = 10
loss_type = mse
train_files = %s/a_train*
weight_files = %s/a_weight*
[Predict]
predict_files = %s/a_train*
""" % (tmp_path, tmp_path, tmp_path))
    m = Model(str(cfg))
    assert (m.vocabulary_size, m.vocabulary_block_num, m.factor_num) == (800000, 10, 100)
    assert m.hash_feature_id is False and m.loss_type == "mse" and m.num_epochs == 20
    assert m.adagrad_init_accumulator == 0.1 and m.save_steps == 10 and m.batch_size == 50000
    assert len(m.train_files) == 1 and len(m.weight_files) == 1 and len(m.predict_files) == 1


def test_numparse_matches_glibc_on_the_host(tmp_path):
    """the number scanner the CUDA parser uses, compiled for the host and fuzzed against glibc"""
    exe = tmp_path / "numparse_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", str(exe),
                           os.path.join(ROOT, "tests", "native", "numparse_check.cpp")])
    out = subprocess.run([str(exe), "11", "300000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert " 0 mismatches" in out.stdout
