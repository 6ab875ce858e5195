"""Host-side logic that needs no GPU: which checkpoint is restored, how training files are cut into
segments, the shuffle-buffer batcher, NUMA binding fallback."""
import os

import numpy as np
import pytest
import torch

from fast_tffm_b200 import data as fm_data
from fast_tffm_b200.fm_model import latest_checkpoint, load_checkpoint, shard_sets


def _blocks(nb, rows, k, step, base):
    out = {"global_step": step}
    for i in range(nb):
        out["vocab_block_%d" % i] = torch.full((rows, k + 1), float(base + i))
        out["vocab_block_%d/Adagrad" % i] = torch.full((rows, k + 1), float(base + i) + 0.5)
    return out


def test_newest_checkpoint_wins_and_half_written_sets_are_ignored(tmp_path):
    path = str(tmp_path / "model.ckpt.pt")
    assert latest_checkpoint(path) is None
    with pytest.raises(IOError):
        load_checkpoint(path)
    torch.save(_blocks(3, 4, 2, 10, 100), path)
    assert latest_checkpoint(path) == (10, [path])
    # a complete 2-GPU set from a later step: file r holds block r of a mod-2 partition
    for r in range(2):
        part = {"global_step": 20, "vocab_block_%d" % r: torch.full((6, 3), float(r)),
                "vocab_block_%d/Adagrad" % r: torch.full((6, 3), r + 0.5)}
        torch.save(part, "%s.shard-%d-of-2" % (path, r))
    step, files = latest_checkpoint(path)
    assert step == 20 and [os.path.basename(f) for f in files] == ["model.ckpt.pt.shard-0-of-2", "model.ckpt.pt.shard-1-of-2"]
    state = load_checkpoint(path)
    assert state["global_step"] == 20 and sorted(k for k in state if k != "global_step") == [
        "vocab_block_0", "vocab_block_0/Adagrad", "vocab_block_1", "vocab_block_1/Adagrad"]
    # an incomplete 4-GPU set and a set whose files are from different steps do not count
    torch.save({"global_step": 99, "vocab_block_0": torch.zeros(1, 3)}, path + ".shard-0-of-4")
    for r, st in enumerate((50, 51, 50)):
        torch.save({"global_step": st, "vocab_block_%d" % r: torch.zeros(1, 3)}, "%s.shard-%d-of-3" % (path, r))
    assert sorted(shard_sets(path)) == [2, 3]
    assert latest_checkpoint(path)[0] == 20
    # the single file again, newer
    torch.save(_blocks(3, 4, 2, 30, 200), path)
    assert latest_checkpoint(path) == (30, [path])
    assert float(load_checkpoint(path)["vocab_block_1"][0, 0]) == 201.0


def _segments(files, cap):
    ds = fm_data.SegmentedDataset.__new__(fm_data.SegmentedDataset)
    ds.train_files, ds.weight_files, ds.segment_bytes = files, [], cap
    ds._pinned = torch.zeros(cap + 16, dtype=torch.uint8)       # pageable stand-in: only the cutting is tested
    out = []
    for n, parts in ds._segments(list(range(len(files)))):
        out.append((bytes(ds._pinned[:n].numpy().tobytes()), [tuple(p) for p in parts]))
    return out


def test_segments_are_cut_at_line_ends_and_cover_every_byte(tmp_path):
    rng = np.random.default_rng(0)
    texts = []
    for f in range(4):
        lines = ["%d %s" % (rng.integers(0, 2), " ".join("%d:1" % t for t in rng.integers(0, 1000, rng.integers(0, 30))))
                 for _ in range([50, 0, 3, 200][f])]
        body = "\n".join(lines).rstrip(" ")
        texts.append(body + ("\n" if lines and f != 2 else ""))       # file 2 lacks the final newline, file 1 is empty
        (tmp_path / ("f%d" % f)).write_text(texts[-1])
    files = [str(tmp_path / ("f%d" % f)) for f in range(4)]
    whole = "".join(t if (not t or t.endswith("\n")) else t + "\n" for t in texts).encode()
    for cap in (256, 1024, 4096, 1 << 20):
        segs = _segments(files, cap)
        assert b"".join(s for s, _ in segs) == whole
        for s, parts in segs:
            assert s.endswith(b"\n") and len(s) <= cap
            assert sum(nb for _, nb in parts) == len(s)
            assert [fi for fi, _ in parts] == sorted(fi for fi, _ in parts)
        # bytes attributed to a file add up to that file (plus the newline it was given)
        per_file = {}
        for _, parts in segs:
            for fi, nb in parts:
                per_file[fi] = per_file.get(fi, 0) + nb
        for f, t in enumerate(texts):
            assert per_file.get(f, 0) == len(t.encode()) + (1 if t and not t.endswith("\n") else 0)
    (tmp_path / "long").write_text("1 " + "7:1 " * 200 + "\n")
    with pytest.raises(ValueError, match="longer than the segment"):
        _segments([str(tmp_path / "long")], 256)


def test_shuffle_buffer_batcher_keeps_lines_and_weights_together(tmp_path):
    n = [130, 75]
    for f in range(2):
        (tmp_path / ("d%d" % f)).write_text("".join("%d 1:1\n" % (1000 * f + i) for i in range(n[f])))
        (tmp_path / ("w%d" % f)).write_text("".join("%d\n" % (1000 * f + i) for i in range(n[f])))
    files = [str(tmp_path / ("d%d" % f)) for f in range(2)]
    wfiles = [str(tmp_path / ("w%d" % f)) for f in range(2)]
    got = list(fm_data.batches(files, wfiles, 32, 2, shuffle=True, seed=3))
    sizes = [len(b) for b, _ in got]
    assert sizes[:-1] == [32] * (len(sizes) - 1) and sum(sizes) == 2 * sum(n)       # only the last batch is smaller
    pairs = [(int(l.split()[0]), w) for b, ws in got for l, w in zip(b, ws)]
    assert all(label == w for label, w in pairs)
    assert sorted(p[0] for p in pairs) == sorted(2 * [1000 * f + i for f in range(2) for i in range(n[f])])
    plain = list(fm_data.batches(files, None, 50, 1, shuffle=False))
    assert [int(l.split()[0]) for b, _ in plain for l in b] == [1000 * f + i for f in range(2) for i in range(n[f])]
    with pytest.raises(ValueError):
        list(fm_data.batches(files, wfiles[:1], 32, 1))


def test_numa_binding_reports_instead_of_failing():
    before = os.sched_getaffinity(0)
    msg = fm_data.bind_host_to_gpu(0)          # no GPU here: nothing is bound, the reason comes back
    assert isinstance(msg, str) and msg
    assert os.sched_getaffinity(0) == before or "bound to" in msg


def test_nnz_bound_comes_from_the_batch_not_from_the_buffer_behind_it():
    """a batch selected from a resident file shares the file's text: its buffers are sized by its own lines"""
    import torch
    from fast_tffm_b200 import fm_ops
    whole = fm_ops.pack_lines(["1 2:3 4:5"] * 1000 + ["0 " + " ".join("%d:1" % i for i in range(50))])
    assert whole.nnz_bound() == whole.text.numel() // 2 + 1
    batch = whole.select(torch.arange(10))
    assert whole.max_line_bytes == len("0 " + " ".join("%d:1" % i for i in range(50))) + 1
    assert batch.nnz_bound() == 10 * (whole.max_line_bytes // 2) + 1 < whole.nnz_bound()
    assert batch.to("cpu").max_line_bytes == whole.max_line_bytes


def test_optimizer_key_is_an_extension_with_a_safe_default(tmp_path):
    import os
    from fast_tffm_b200.fm_model import Model
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cfg = open(os.path.join(root, "sample.cfg")).read()
    (tmp_path / "a.cfg").write_text(cfg)
    assert Model(str(tmp_path / "a.cfg")).optimizer == "adagrad"          # the reference's only optimizer
    (tmp_path / "b.cfg").write_text(cfg.replace("loss_type = mse", "loss_type = mse\noptimizer = sgd"))
    assert Model(str(tmp_path / "b.cfg")).optimizer == "sgd"
    (tmp_path / "c.cfg").write_text(cfg.replace("loss_type = mse", "loss_type = mse\noptimizer = adam"))
    with pytest.raises(ValueError, match="optimizer"):
        Model(str(tmp_path / "c.cfg"))
    (tmp_path / "d.cfg").write_text(cfg.replace("adagrad.initial_accumulator = 0.1", "adagrad.initial_accumulator = 0"))
    with pytest.raises(ValueError, match="initial_accumulator_value must be positive"):
        Model(str(tmp_path / "d.cfg"))


def test_claim_form_follows_the_measured_share_of_distinct_rows():
    """Model._claim_form: the records (distinct rows U, occurrences N) a row-form step left behind are read once their
    event has passed; U / N picks the form; the occurrence form looks again every FORM_PROBE_EVERY steps"""
    from fast_tffm_b200.fm_model import Model

    class Ev:
        def __init__(self, done):
            self.done = done

        def query(self):
            return self.done
    m = Model(vocabulary_size=1000, factor_num=4, batch_size=8)
    assert m.claim_form == "auto" and m._claim_form() == "rows"          # the first step measures
    pending = Ev(False)
    m._form_q.append((pending, torch.tensor([950, 1000])))
    assert m._claim_form() == "rows"                                      # not there yet: nothing is read, nothing blocks
    pending.done = True
    assert m._claim_form() == "occurrences" and not m._form_q             # U / N = 0.95: the occurrence walk
    m._form_q.append((Ev(True), torch.tensor([0, 0])))                    # a void step says nothing
    assert m._claim_form() == "occurrences"
    m.FORM_PROBE_EVERY = 4
    m._form_age = 0
    forms = [m._claim_form() for _ in range(8)]
    assert forms == ["occurrences"] * 3 + ["rows"] + ["occurrences"] * 3 + ["rows"]
    m._form_q.append((Ev(True), torch.tensor([100, 1000])))               # ids repeat: back to the row form, and it stays
    assert [m._claim_form() for _ in range(3)] == ["rows"] * 3
    m.claim_form = "occurrences"                                          # pinned: no measuring
    assert m._claim_form() == "occurrences"
