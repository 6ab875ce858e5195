"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the
same inputs.  Integer / index outputs are bit-exact; float outputs of the parser
are bit-exact (they are glibc strtof's bits); scores, losses and gradients are
held to 1e-5 relative (north_star), with an absolute floor scaled to the size of
the terms being summed because the second-order term cancels.
"""
import os
import random

import numpy as np
import pytest
import torch

import bench_data
from fast_tffm_b200 import fm_ops
from fast_tffm_b200.fm_model import Model
from oracle import dense_ref, oracle as O
from parity_util import RTOL, assert_model_state as _assert_model_state, close as _close, pred_scale as _pred_scale

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PATHS = list(Model.PATHS)


def dev(a, dtype=None):
    return torch.as_tensor(np.asarray(a), dtype=dtype).to(DEV)


def bits(t):
    return t.detach().cpu().numpy().view(np.uint32)


def assert_parse_equal(lines, vocab, hashed):
    exp = O.parse_lines(lines, vocab, hashed)
    got = fm_ops.fm_parser(lines, vocab, hashed)
    np.testing.assert_array_equal(bits(got[0]), exp[0].view(np.uint32))
    np.testing.assert_array_equal(got[1].cpu().numpy(), exp[1])
    np.testing.assert_array_equal(got[2].cpu().numpy(), exp[2])
    np.testing.assert_array_equal(bits(got[3]), exp[3].view(np.uint32))
    return got


# ---------------------------------------------------------------- parser ----
def test_parser_reference_vectors(golden_parser):
    """port of test/fm_parser_op_test.py:17-46"""
    g = golden_parser
    labels, sizes, ids, vals = fm_ops.fm_parser(g["examples"], g["vocab_size"])
    assert labels.dtype == torch.float32 and sizes.dtype == torch.int32
    assert ids.dtype == torch.int64 and vals.dtype == torch.float32
    np.testing.assert_allclose(labels.cpu(), g["labels"])
    np.testing.assert_array_equal(ids.cpu(), g["feature_ids"])
    np.testing.assert_allclose(vals.cpu(), np.float32(g["feature_vals"]))
    np.testing.assert_array_equal(sizes.cpu(), g["sizes"])
    labels, sizes, ids, vals = fm_ops.fm_parser(g["examples"], g["vocab_size"], True)
    np.testing.assert_array_equal(ids.cpu(), g["hashed_ids_1_to_10_mod_10000_selfgenerated"])
    np.testing.assert_array_equal(sizes.cpu(), g["sizes"])
    np.testing.assert_allclose(vals.cpu(), np.float32(g["feature_vals"]))


def test_parser_reference_errors(golden_parser):
    """port of test/fm_parser_op_test.py:48-81: same message prefixes"""
    for line, prefix in golden_parser["errors"]:
        with pytest.raises(ValueError) as e:
            fm_ops.fm_parser([line], golden_parser["vocab_size"], False)
        assert str(e.value).startswith(prefix) and str(e.value).endswith(line)
    with pytest.raises(ValueError, match="Invalid feature id"):
        fm_ops.fm_parser(["1 10000:1"], 10000)
    with pytest.raises(ValueError, match="Invalid feature id"):
        fm_ops.fm_parser(["1 -1:1"], 10000)
    with pytest.raises(ValueError, match="Label could not be read"):
        fm_ops.fm_parser(["1 1:1", ""], 10000)


def test_parser_first_failing_line_wins():
    lines = ["1 1:1"] * 700 + ["1 1:x"] + ["1 2:2"] * 300 + ["y 1:1"] + ["1 1:1"] * 50
    with pytest.raises(fm_ops.FmParseError) as e:
        fm_ops.fm_parser(lines, 100)
    assert e.value.line_no == 700 and e.value.kind == 4


def test_parser_libc_quirks_match_oracle():
    cases = [
        [" 2  7: 3", "1"],                      # whitespace skipping by strtof / strtoll
        ["1 2:3 ", "1 2:3  ".rstrip() + " "],   # trailing single space
        ["\t1 \t2:\n3"],
        [b"1 2:3\x00 garbage"],                 # NUL ends the line
        ["1 +5:+.5e1 6:-0 7:1e-3 8:0x1p3 9:inf 10:nan 11:-nan(12)"],
        ["1e2 1:1.5E+1 2:00012.5000"],
        ["0x10 0:1"],
        ["1 5"],
        ["-0 99:1e-46 98:3.4028236e38 97:1.17549435e-38"],
    ]
    for lines in cases:
        assert_parse_equal(lines, 100, False)
    for lines in [["1  a", "2 a:1 :2 b", "3 a\tb:1"], ["1 ab cd:2.5 ef: 3"], ["0 :1"], ["7 x:y".replace("y", "1")]]:
        assert_parse_equal(lines, 1000, True)


def test_parser_error_kinds_match_oracle():
    bad = ["", " ", "x", "1 a", "1 2:", "1 2: ", "1 2:3x", "1 2:3 4:", "1 200", "1 -3", "1 2:3\t4", "1,2", "1 2 3:",
           "1 99999999999999999999", "1:2", "1 2:3:4", "1 2::3", "nan(", "1 0x10", "1\t2"]
    for line in bad:
        for hashed in (False, True):
            try:
                O.parse_lines([line], 100, hashed)
                exp = None
            except O.ParseError as e:
                exp = e.kind
            if exp is None:
                assert_parse_equal([line], 100, hashed)
            else:
                with pytest.raises(fm_ops.FmParseError) as e:
                    fm_ops.fm_parser([line], 100, hashed)
                assert e.value.kind == exp, (line, hashed)


def test_parser_number_vectors_bit_exact(golden_numbers):
    """every strtof golden string as a label and as a feature value, against glibc's bits"""
    ok_lines, exp_bits = [], []
    for s, b, n, *quirk in golden_numbers["strtof"]:
        if "\x00" in s or n != len(s) or n == 0 or s[0] in " \t\n\r\v\f":
            continue
        ok_lines.append("%s 3:%s" % (s, s))
        exp_bits.append(b)
    labels, sizes, ids, vals = fm_ops.fm_parser(ok_lines, 10)
    np.testing.assert_array_equal(bits(labels), np.uint32(exp_bits))
    np.testing.assert_array_equal(bits(vals), np.uint32(exp_bits))
    assert sizes.cpu().tolist() == [1] * len(ok_lines)


def _random_line(rnd, hashed, vocab):
    def num():
        r = rnd.random()
        if r < 0.3:
            return "%.4f" % (rnd.random() * 10)
        if r < 0.5:
            return "%.6g" % (rnd.gauss(0, 1) * 10 ** rnd.randint(-8, 8))
        if r < 0.6:
            return repr(rnd.random() * 10 ** rnd.randint(-40, 38))
        if r < 0.7:
            return str(rnd.randint(-5, 5))
        if r < 0.75:
            return "%.17g" % rnd.random()
        if r < 0.8:
            return rnd.choice(["1e-45", "3e-46", "1e39", "0.0", "-0", "+1", ".5", "5.", "1e+2"])
        return "%d.%d" % (rnd.randint(0, 99), rnd.randint(0, 999999))
    toks = []
    for _ in range(rnd.choice([0, 1, 2, 5, 39, 39, 39, 70])):
        fid = ("f%x" % rnd.getrandbits(rnd.choice([8, 32, 64, 100]))) if hashed else str(rnd.randrange(vocab))
        toks.append(fid if rnd.random() < 0.4 else fid + ":" + num())
    line = " ".join([num()] + toks)
    if rnd.random() < 0.2:
        line += " "
    return line


@pytest.mark.parametrize("hashed", [False, True])
def test_parser_fuzz_bit_exact(hashed):
    rnd = random.Random(1234 + hashed)
    lines = [_random_line(rnd, hashed, 1 << 20) for _ in range(3000)]
    # a few irregular (but valid) lines sprinkled in: they take the sequential path
    for i in range(0, 3000, 97):
        cand = lines[i].replace(" ", "  ", 1) if not hashed else lines[i].replace(":", ": ", 1)
        try:
            O.parse_lines([cand], 1 << 20, hashed)
            lines[i] = cand
        except O.ParseError:
            pass
    assert_parse_equal(lines, 1 << 20, hashed)


def test_parser_long_and_ragged_lines():
    lines, _ = bench_data.ragged_batch(64, 5, max_nnz=512)
    lines.append(b"1 " + b" ".join(b"%d:%d" % (i, i) for i in range(5000)))   # 5000 features, ~50 KB
    lines.append(b"0")
    assert_parse_equal(lines, 1 << 24, False)
    assert_parse_equal(lines, 1 << 24, True)


def test_parser_criteo_excerpt(criteo_lines):
    lines, _ = criteo_lines
    assert_parse_equal(lines, 800000, False)


def test_parser_shuffled_selection_of_resident_text(criteo_lines):
    """a batch = gathered (begin, end) indices over text that stays where it is"""
    lines, _ = criteo_lines
    packed = fm_ops.pack_lines(lines[:1000]).to(DEV, False)
    perm = torch.randperm(1000, generator=torch.Generator().manual_seed(3))[:317]
    got = fm_ops.fm_parser(packed.select(perm.to(DEV)), 800000)
    exp = O.parse_lines([lines[i] for i in perm.tolist()], 800000)
    np.testing.assert_array_equal(got[2].cpu().numpy(), exp[2])
    np.testing.assert_array_equal(got[1].cpu().numpy(), exp[1])
    np.testing.assert_array_equal(bits(got[3]), exp[3].view(np.uint32))
    np.testing.assert_array_equal(bits(got[0]), exp[0].view(np.uint32))


def test_device_dataset_epochs(tmp_path, criteo_lines):
    from fast_tffm_b200.data import DeviceDataset
    lines, weights = criteo_lines
    files, wfiles = [], []
    for i, (lo, hi) in enumerate([(0, 300), (300, 1000)]):
        f, w = tmp_path / ("train_%d" % i), tmp_path / ("weight_%d" % i)
        f.write_text("\n".join(lines[lo:hi]) + ("\n" if i == 0 else ""))     # second file: no final newline
        w.write_text("\n".join("%g" % x for x in weights[lo:hi]) + "\n")
        files.append(str(f))
        wfiles.append(str(w))
    ds = DeviceDataset(files, wfiles, DEV)
    assert len(ds) == 1000
    np.testing.assert_array_equal(ds.weights.cpu().numpy(), np.float32(weights[:1000]))
    seen = []
    for packed, w in ds.batches(256, num_epochs=1, shuffle=True, seed=5):
        labels, sizes, ids, vals = fm_ops.fm_parser(packed, 800000)
        assert w.numel() == packed.n_lines == labels.numel()
        # weight == label + 1 in the reference's sample files: the pairing survived the shuffle
        np.testing.assert_array_equal(w.cpu().numpy(), labels.cpu().numpy() + 1)
        seen.append(sizes.cpu().numpy())
    assert [len(x) for x in seen] == [256, 256, 256, 232]
    exp_sizes = O.parse_lines(lines[:1000], 800000)[1]
    assert sorted(np.concatenate(seen).tolist()) == sorted(exp_sizes.tolist())
    order = [p for p, _ in ds.batches(1000, shuffle=False)]
    np.testing.assert_array_equal(fm_ops.fm_parser(order[0], 800000)[2].cpu().numpy(),
                                  O.parse_lines(lines[:1000], 800000)[2])


def test_segmented_dataset_streams_files_through_the_arena(tmp_path, criteo_lines):
    """files larger than a segment: every line arrives exactly once per epoch with its own weight, batches
    have batch_size lines except the last of the run, lines cross segment and file boundaries"""
    from fast_tffm_b200.data import SegmentedDataset
    lines, weights = criteo_lines
    cuts = [0, 700, 741, 741, 2000]                       # four files, one of them empty
    files, wfiles = [], []
    for i in range(4):
        part = lines[cuts[i]:cuts[i + 1]]
        text = "\n".join(part) + ("\n" if part and i != 1 else "")     # file 1: no newline at the end
        (tmp_path / ("d%d" % i)).write_text(text)
        (tmp_path / ("w%d" % i)).write_text("".join("%g\n" % w for w in weights[cuts[i]:cuts[i + 1]]))
        files.append(str(tmp_path / ("d%d" % i)))
        wfiles.append(str(tmp_path / ("w%d" % i)))
    ds = SegmentedDataset(files, wfiles, DEV, segment_bytes=48 * 1024)      # ~220 lines per segment
    # in order
    got, got_w, sizes = [], [], []
    for packed, w in ds.batches(128, num_epochs=2, shuffle=False):
        sizes.append(packed.n_lines)
        got += [packed.line(i) for i in range(packed.n_lines)]
        got_w += w.cpu().tolist()
    assert sizes == [128] * (4000 // 128) + [4000 % 128]
    assert got == lines + lines
    np.testing.assert_array_equal(np.float32(got_w), np.float32(weights + weights))
    # shuffled: a permutation per epoch (within segments), weights travel with their lines
    seen = []
    for packed, w in ds.batches(128, num_epochs=1, shuffle=True, seed=5):
        wl = w.cpu().tolist()
        seen += [(packed.line(i), wl[i]) for i in range(packed.n_lines)]
    assert [l for l, _ in seen] != lines
    assert sorted(seen) == sorted(zip(lines, [float(np.float32(x)) for x in weights]))
    # the parser accepts what the dataset yields (begins / ends over the two-half arena)
    packed, w = next(iter(ds.batches(300, 1, shuffle=False)))
    e = O.parse_lines(lines[:300], 1 << 20, False)
    labels, szs, ids, vals = fm_ops.fm_parser(packed, 1 << 20)
    np.testing.assert_array_equal(ids.cpu().numpy(), e[2])
    np.testing.assert_array_equal(szs.cpu().numpy(), e[1])
    with pytest.raises(ValueError, match="raise segment_bytes"):
        list(SegmentedDataset(files, (), DEV, segment_bytes=8 * 1024).batches(1000, 1, shuffle=False))


def test_parser_empty_batch():
    labels, sizes, ids, vals = fm_ops.fm_parser([], 10)
    assert labels.numel() == 0 and sizes.numel() == 0 and ids.numel() == 0 and vals.numel() == 0


def test_split_lines_matches_host_offsets():
    text, off, _ = bench_data.criteo_batch(1000, 3)
    padded = torch.from_numpy(bench_data.pad16(text)).to(DEV)
    got = fm_ops.fm_split_lines(padded, text.size, 2000)
    np.testing.assert_array_equal(got.cpu().numpy(), off)
    with pytest.raises(ValueError, match="more than"):
        fm_ops.fm_split_lines(padded, text.size, 999)
    with pytest.raises(ValueError, match="newline"):
        fm_ops.fm_split_lines(padded, text.size - 1, 2000)


def test_parser_synthetic_cfg2_full_size():
    """BASELINE config 2 shape at full batch size, bit-exact against the oracle"""
    text, off, nnz = bench_data.criteo_batch(65536, 0)
    exp = O.parse(text, off, 1 << 24, True)
    packed = fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off))
    got = fm_ops.fm_parser(packed, 1 << 24, True)
    assert got[2].numel() == nnz
    np.testing.assert_array_equal(got[2].cpu().numpy(), exp[2])
    np.testing.assert_array_equal(bits(got[3]), exp[3].view(np.uint32))
    np.testing.assert_array_equal(bits(got[0]), exp[0].view(np.uint32))
    np.testing.assert_array_equal(got[1].cpu().numpy(), exp[1])


# ---------------------------------------------------------------- unique ----
@pytest.mark.parametrize("n,hi", [(0, 10), (1, 10), (1000, 50), (100000, 1 << 24), (300000, 1000), (2555904, 1 << 24)])
def test_unique_first_occurrence_order(n, hi):
    rng = np.random.default_rng(n + hi)
    ids = rng.integers(0, hi, n)
    ori, local = fm_ops.fm_unique(dev(ids, torch.int64))
    e_ori, e_local = O.unique(ids)
    np.testing.assert_array_equal(ori.cpu().numpy(), e_ori)
    np.testing.assert_array_equal(local.cpu().numpy(), e_local)
    assert local.dtype == torch.int32 and ori.dtype == torch.int64


def test_unique_all_equal_and_large_ids():
    ids = np.full(5000, (1 << 62) + 12345, np.int64)
    ori, local = fm_ops.fm_unique(dev(ids))
    assert ori.cpu().tolist() == [(1 << 62) + 12345] and not local.any()


# ------------------------------------------------------ sort by row ----
def _sort_by_row(ids, vals, poses, vocab, n_ranks=1, capacity=None):
    from fast_tffm_b200 import _lib
    lib = _lib.load()
    n, b = len(ids), len(poses) - 1
    cap = n if capacity is None else capacity
    d_ids = dev(np.concatenate([ids, np.zeros(cap - n, np.int64)]), torch.int64)
    d_vals = dev(np.concatenate([vals, np.zeros(cap - n, np.float32)]), torch.float32)
    d_poses, d_n = dev(poses, torch.int32), dev([n], torch.int64)
    keys = torch.empty(max(cap, 1), dtype=torch.int32, device=DEV)
    pairs = torch.empty(max(cap, 1), 2, dtype=torch.int32, device=DEV)
    counts = torch.full((2,), 7, dtype=torch.int64, device=DEV)
    ws = _lib.workspace(lib.fm_sort_workspace(cap), torch.device(DEV))
    _lib.check(lib.fm_sort_by_row(_lib.ptr(d_ids), _lib.ptr(d_vals), _lib.ptr(d_poses), b, cap, _lib.ptr(d_n), vocab,
                                  n_ranks, _lib.ptr(keys), _lib.ptr(pairs), _lib.ptr(counts), None,
                                  _lib.ptr(ws), ws.numel(), 0, _lib.stream()), "fm_sort_by_row")
    return keys[:n].cpu().numpy().view(np.uint32), pairs[:n].cpu().numpy().view(np.uint32), counts.cpu().numpy()


@pytest.mark.parametrize("n,vocab,n_ranks,slack", [(0, 10, 1, 5), (1, 10, 1, 0), (5000, 37, 1, 0), (40000, 1 << 16, 1, 999),
                                                  (300000, 1 << 24, 1, 0), (300000, (1 << 27) + 5, 1, 17),
                                                  (200000, 1 << 24, 8, 0), (70000, 1000003, 3, 0),
                                                  (2555904, 1 << 24, 1, 0)])
def test_sort_by_row_is_a_stable_sort(n, vocab, n_ranks, slack):
    """keys ascending, occurrences of one row in batch order (bit-exact index work); counts[0] = n"""
    rng = np.random.default_rng(n + n_ranks)
    ids = rng.integers(0, vocab, n)
    if n > 1000:
        ids[rng.integers(0, n, n // 3)] = ids[0]              # one very hot row: its run is cut into many chunks
    vals = rng.normal(size=n).astype(np.float32)
    b = max(1, n // 39)
    cuts = np.sort(rng.integers(0, n + 1, b - 1))
    poses = np.concatenate([[0], cuts, [n]]).astype(np.int32)  # ragged examples, some empty
    keys, pairs, counts = _sort_by_row(ids, vals, poses, vocab, n_ranks, n + slack)
    if n_ranks > 1:
        rows = -(-vocab // n_ranks)
        row_bits = max(1, int(rows - 1).bit_length())
        exp_key = ((ids % n_ranks) << row_bits) | (ids // n_ranks)
    else:
        exp_key = ids
    order = np.argsort(exp_key, kind="stable")
    example = np.repeat(np.arange(b), np.diff(poses))
    np.testing.assert_array_equal(keys, exp_key[order].astype(np.uint32))
    np.testing.assert_array_equal(pairs[:, 0], example[order].astype(np.uint32))
    np.testing.assert_array_equal(pairs[:, 1], vals[order].view(np.uint32))
    assert counts.tolist() == [n, 0]


def test_sgd_step_extension(criteo_lines):
    """optimizer = sgd (not in the reference, which only builds Adagrad): w -= lr * G with the reference's own
    gradient (the oracle's FmGrad), two steps, hot and cold rows"""
    lines, weights = criteo_lines
    rng = np.random.default_rng(77)
    vocab, k, bs, lr = 200003, 16, 300, 0.05
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    kw = dict(factor_lambda=2e-3, bias_lambda=1e-3, loss_type="logistic")
    ref = O.RefModel(vocab, 1, k, flat, batch_size=bs, lr=lr, **kw)
    m = Model(vocabulary_size=vocab, factor_num=k, batch_size=bs, learning_rate=lr, optimizer="sgd", **kw).build(
        DEV, init_table=flat)
    gmax = 0.0
    for step in range(2):
        sl = slice(step * bs, (step + 1) * bs)
        w = np.float32(weights[sl])
        f = ref.forward(lines[sl])
        lv, pg = O.loss(f["pred"], f["labels"], w, "logistic")
        g = O.grad(f["local"], f["lp"], f["vals"], f["poses"], ref.lf, ref.lb, pg, 1.0 / bs)
        ref.blocks[0, f["ori"]] -= np.float32(lr) * g
        gmax = max(gmax, float(np.abs(g).max()))
        got = m.train_step(lines[sl], w)
        np.testing.assert_allclose(got, lv, rtol=RTOL)
    got_t, exp_t = m.table[:, :k + 1].cpu().numpy(), ref.flat_table()
    _close(got_t, exp_t, lr * gmax * 2)      # a weight moved by at most lr * |g| per step
    # and the movement itself (tiny next to the weights): within 1e-3 of the largest move
    move, exp_move = got_t.astype(np.float64) - flat, exp_t.astype(np.float64) - flat
    assert np.abs(exp_move).max() > 0 and np.abs(move - exp_move).max() <= 1e-3 * np.abs(exp_move).max()
    assert bool((m.accum[:, :k + 1] == m.adagrad_init_accumulator).all())      # SGD leaves the accumulator alone
    with pytest.raises(ValueError, match="sorted path only"):
        m.train_step(lines[:bs], path="claim")


# ------------------------------------------------------ scorer and grad ----
def _random_case(rng, b, rows, k, ld_pad, max_nnz):
    sizes = rng.integers(0, max_nnz + 1, b).astype(np.int32)
    if b > 3:
        sizes[1] = 0
        sizes[2] = max_nnz
    poses = O.poses(sizes)
    n = int(poses[-1])
    fids = rng.integers(0, rows, n).astype(np.int32)
    vals = rng.normal(size=n).astype(np.float32)
    params = np.zeros((rows, k + 1 + ld_pad), np.float32)
    params[:, :k + 1] = rng.uniform(-0.3, 0.3, (rows, k + 1))
    return fids, params, vals, poses


def test_scorer_reference_case(golden_scorer):
    """port of test/fm_scorer_op_test.py: pred, reg and d cost / d params (autograd through fm_scorer)"""
    g = golden_scorer
    params = dev(np.float32(g["params"])).requires_grad_(True)
    pred, reg = fm_ops.fm_scorer(g["feature_ids"], params, g["feature_vals"], g["feature_poses"],
                                 g["factor_lambda"], g["bias_lambda"])
    assert pred.shape == (2,) and reg.shape == ()
    np.testing.assert_allclose(pred.detach().cpu(), g["pred"], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(reg.detach().cpu(), g["reg"], rtol=RTOL)
    cost = ((dev(g["labels"], torch.float32) - pred) ** 2).sum() + reg
    cost.backward()
    np.testing.assert_allclose(params.grad.cpu(), g["grad"], rtol=RTOL, atol=1e-6)


@pytest.mark.parametrize("k,ld_pad", [(0, 0), (1, 0), (3, 1), (4, 0), (8, 0), (16, 0), (16, 3), (17, 2), (32, 0),
                                       (33, 0), (64, 0), (100, 0), (100, 3), (128, 0), (150, 1), (256, 0)])
def test_scorer_and_grad_vs_oracle(k, ld_pad):
    rng = np.random.default_rng(k * 7 + ld_pad)
    b, rows = 257, 500
    fids, params, vals, poses = _random_case(rng, b, rows, k, ld_pad, 70)
    lf, lb = 0.013, 0.4
    e_pred, e_reg = O.score(fids, params, vals, poses, lf, lb, k=k)
    lib_params, d_fids, d_vals, d_poses = dev(params), dev(fids), dev(vals), dev(poses)
    # the op derives K from the width of params (cc/fm_scorer_op.cc:51): pass a K+1-wide view
    view = lib_params[:, :k + 1] if ld_pad else lib_params
    if ld_pad:
        # a padded table row stride goes through the C ABI's params_ld; exercise it via the model-level call
        from fast_tffm_b200 import _lib
        lib = _lib.load()
        pred = torch.empty(b, device=DEV)
        reg = torch.empty((), device=DEV)
        ws = _lib.workspace(lib.fm_score_workspace(b), lib_params.device)
        _lib.check(lib.fm_score(_lib.ptr(d_fids), _lib.ptr(lib_params), lib_params.stride(0), _lib.ptr(d_vals),
                                _lib.ptr(d_poses), b, k, lf, lb, _lib.ptr(pred), _lib.ptr(reg), _lib.ptr(ws),
                                ws.numel(), _lib.stream()), "fm_score")
    else:
        pred, reg = fm_ops.fm_scorer(d_fids, view, d_vals, d_poses, lf, lb)
    # scale of the summed terms per example: sum |x| |b| + 0.5 sum_k (sum |x v|)^2
    absx = np.abs(vals)
    scale = np.zeros(b)
    for i in range(b):
        sl = slice(poses[i], poses[i + 1])
        r = np.abs(params[fids[sl]])
        scale[i] = (absx[sl] * r[:, 0]).sum() + 0.5 * ((absx[sl, None] * r[:, 1:k + 1]).sum(0) ** 2).sum()
    _close(pred.cpu().numpy(), e_pred, scale)
    np.testing.assert_allclose(reg.item(), e_reg, rtol=RTOL)

    pg = rng.normal(size=b).astype(np.float32)
    d_pg = dev(pg)
    e_grad = O.grad(fids, params, vals, poses, lf, lb, pg, 0.37, k=k)
    if ld_pad:
        from fast_tffm_b200 import _lib
        lib = _lib.load()
        grad = torch.full_like(lib_params, 7.0)
        _lib.check(lib.fm_grad(_lib.ptr(d_fids), _lib.ptr(lib_params), lib_params.stride(0), _lib.ptr(d_vals),
                               _lib.ptr(d_poses), b, k, lf, lb, _lib.ptr(d_pg), 0.37, None, rows, None,
                               _lib.ptr(grad), _lib.stream()), "fm_grad")
    else:
        grad = fm_ops.fm_grad(d_fids, view, d_vals, d_poses, lf, lb, d_pg, dev(np.float32(0.37)))
    assert grad.shape == lib_params.shape if ld_pad else grad.shape == (rows, k + 1)
    gscale = np.abs(e_grad).max() if e_grad.size else 0.0
    _close(grad.cpu().numpy(), e_grad, gscale)


def test_scorer_ragged_rows_up_to_512():
    rng = np.random.default_rng(99)
    fids, params, vals, poses = _random_case(rng, 64, 4000, 32, 0, 512)
    params *= 0.1
    e_pred, e_reg = O.score(fids, params, vals, poses, 1e-5, 1e-5)
    pred, reg = fm_ops.fm_scorer(dev(fids), dev(params), dev(vals), dev(poses), 1e-5, 1e-5)
    _close(pred.cpu().numpy(), e_pred, np.abs(e_pred).max() + 1.0)
    np.testing.assert_allclose(reg.item(), e_reg, rtol=RTOL)


def test_scorer_matches_dense_closed_form():
    """independent second oracle (test/fm_scorer_op_test.py:49-63 in torch float64)"""
    rng = np.random.default_rng(5)
    fids, params, vals, poses = _random_case(rng, 40, 60, 16, 0, 12)
    labels = rng.normal(size=40)
    dp, dr, dg = dense_ref.dense_cost_and_grad(fids, torch.from_numpy(params), vals, poses, 0.3, 0.7, labels)
    p = dev(params).requires_grad_(True)
    pred, reg = fm_ops.fm_scorer(dev(fids), p, dev(vals), dev(poses), 0.3, 0.7)
    (((dev(labels, torch.float32) - pred) ** 2).sum() + reg).backward()
    np.testing.assert_allclose(pred.detach().cpu(), dp, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(reg.item(), float(dr), rtol=RTOL)
    np.testing.assert_allclose(p.grad.cpu(), dg, rtol=1e-4, atol=1e-5)


def test_scorer_gradient_only_flows_to_params():
    p = torch.rand(5, 4, device=DEV, requires_grad=True)
    vals = torch.rand(3, device=DEV, requires_grad=True)
    pred, reg = fm_ops.fm_scorer([0, 1, 2], p, vals, [0, 2, 3], 0.1, 0.1)
    (pred.sum() + reg).backward()
    assert p.grad is not None and vals.grad is None            # tffm/fm_ops.py:19-27


# ------------------------------------------------- loss, lookup, adagrad ----
@pytest.mark.parametrize("loss_type", ["logistic", "mse"])
def test_loss_and_pred_grad(loss_type):
    rng = np.random.default_rng(8)
    b = 10007
    pred = (rng.normal(size=b) * 4).astype(np.float32)
    y = rng.integers(0, 2, b).astype(np.float32)
    w = rng.uniform(0.5, 2, b).astype(np.float32)
    for weights in (w, None):
        e_loss, e_pg = O.loss(pred, y, weights, loss_type)
        loss, pg = fm_ops.fm_loss_grad(dev(pred), dev(y), None if weights is None else dev(weights), loss_type)
        np.testing.assert_allclose(loss.item(), e_loss, rtol=RTOL)
        np.testing.assert_allclose(pg.cpu().numpy(), e_pg, rtol=RTOL, atol=1e-9)


def test_lookup_and_adagrad_vs_oracle():
    rng = np.random.default_rng(11)
    vocab, nb, k = 20011, 10, 16
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    ref = O.RefModel(vocab, nb, k, flat, batch_size=8)
    m = Model(vocabulary_size=vocab, vocabulary_block_num=nb, factor_num=k).build(DEV, init_table=flat)
    ids = rng.permutation(vocab)[:5000]
    rows = fm_ops.fm_lookup(m.table, dev(ids), k)
    np.testing.assert_array_equal(rows.cpu().numpy(), O.lookup_mod(ref.blocks, vocab, nb, ids))
    g = rng.normal(size=(5000, k + 1)).astype(np.float32) * 0.1
    g[7] = 0.0   # a zero gradient row is still "touched": acc unchanged, w unchanged
    O.adagrad_mod(ref.blocks, ref.acc, vocab, nb, ids, g, 0.01)
    fm_ops.fm_sparse_adagrad(m.table, m.accum, dev(ids), dev(g), k, 0.01)
    np.testing.assert_allclose(m.table[:, :k + 1].cpu().numpy(), ref.flat_table(), rtol=RTOL, atol=1e-8)
    np.testing.assert_allclose(m.accum[:, :k + 1].cpu().numpy(), ref.flat_acc(), rtol=RTOL)
    # checkpoint relabelling: vocab_block_i[id // nb] for id % nb == i
    st = m.blocks_state()
    assert st["vocab_block_3"].shape == (vocab // nb + 1, k + 1)
    np.testing.assert_allclose(st["vocab_block_3"].numpy(), ref.blocks[3], rtol=RTOL, atol=1e-8)
    m2 = Model(vocabulary_size=vocab, vocabulary_block_num=nb, factor_num=k).build(DEV, seed=1)
    m2.load_blocks_state(st)
    assert torch.equal(m2.table[:, :k + 1], m.table[:, :k + 1]) and torch.equal(m2.accum[:, :k + 1], m.accum[:, :k + 1])


# ------------------------------------------------------------ whole step ----
@pytest.mark.parametrize("path", PATHS)
@pytest.mark.parametrize("loss_type,hashed", [("mse", False), ("logistic", False), ("logistic", True)])
def test_train_steps_match_reference_model(criteo_lines, loss_type, hashed, path):
    """cfg1 shape in small: sample.cfg's pipeline on the Criteo excerpt, explicit initial table,
    fixed batch order; loss per step and the trained table against the oracle model"""
    lines, weights = criteo_lines
    rng = np.random.default_rng(21)
    vocab, nb, k, bs = 200003, 10, 12, 500
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    kw = dict(factor_lambda=2e-3, bias_lambda=1e-3, loss_type=loss_type)
    ref = O.RefModel(vocab, nb, k, flat, batch_size=bs, hash_feature_id=hashed, lr=0.05, init_acc=0.1, **kw)
    m = Model(vocabulary_size=vocab, vocabulary_block_num=nb, factor_num=k, batch_size=bs, hash_feature_id=hashed,
              learning_rate=0.05, adagrad_init_accumulator=0.1, **kw).build(DEV, init_table=flat)
    for step in range(4):
        sl = slice(step * bs, (step + 1) * bs)
        w = np.float32(weights[sl]) if step % 2 == 0 else None
        e = ref.train_step(lines[sl], w)
        got = m.train_step(lines[sl], w, path=path)
        np.testing.assert_allclose(got, e["loss"], rtol=RTOL)
        _close(m._bufs["pred"][:bs].cpu().numpy(), e["pred"], _pred_scale(e))
        np.testing.assert_allclose(float(m._bufs["scalars"][2]), e["reg"], rtol=RTOL)
    _assert_model_state(m, ref, k, 0.05, 4)
    pred = m.predict(lines[1000:1300], path="composed" if path == "composed" else "claim")
    f = ref.forward(lines[1000:1300])
    _close(pred.cpu().numpy(), f["pred"], _pred_scale(f))
    if path == "claim":   # table-shaped accumulators are left all zero by the update
        assert not m.gtab.any()


@pytest.mark.parametrize("k", [0, 1, 2, 5, 13, 29, 30, 61, 100, 122, 123, 128, 250, 257, 378, 379, 500, 1030])
def test_train_steps_any_factor_num(k):
    """the reference takes any factor_num: lane-group shapes (K+3 = 1..32 chunks), the 2- and 3-chunk-per-lane
    shapes, the 256-column slabs of fm_scorer / fm_grad and the wide-row kernels (K > 378), all paths and both
    table layouts against the oracle model; ids repeat inside an example and across examples"""
    rng = np.random.default_rng(100 + k)
    vocab, bs = 997, 64
    lines = []
    for i in range(2 * bs):
        n = int(rng.integers(0, 12))
        ids = rng.integers(0, 40 if i % 3 else vocab, n)        # a hot set: many repeats
        toks = ["%d:%s" % (t, "%.3f" % rng.normal()) if rng.random() < 0.7 else "%d" % t for t in ids]
        lines.append(("%d " % rng.integers(0, 2) + " ".join(toks)).rstrip().encode())
    flat = rng.uniform(-0.05, 0.05, (vocab, k + 1)).astype(np.float32)
    kw = dict(factor_lambda=3e-3, bias_lambda=2e-3, loss_type="logistic")
    w = rng.uniform(0.5, 2.0, bs).astype(np.float32)
    combos = [("claim", "interleaved"), ("claim", "split"), ("fused", "split"), ("composed", "split"), ("claim", "pair"),
              ("claim/rows", "interleaved"), ("claim/rows", "split")]
    if (k + 4) // 4 <= 96:      # the sorted path's lane groups cover rows of up to 96 chunks
        combos += [("sorted", "pair"), ("sorted", "interleaved"), ("sorted", "split")]
    for path, layout in combos:
        ref = O.RefModel(vocab, 7, k, flat, batch_size=bs, lr=0.05, init_acc=0.1, **kw)
        m = Model(vocabulary_size=vocab, vocabulary_block_num=7, factor_num=k, batch_size=bs, learning_rate=0.05,
                  adagrad_init_accumulator=0.1, **kw).build(DEV, init_table=flat, layout=layout)
        path, _, form = path.partition("/")
        m.claim_form = form or "occurrences"
        for step in range(2):
            sl = slice(step * bs, (step + 1) * bs)
            e = ref.train_step(lines[sl], w if step else None)
            got = m.train_step(lines[sl], w if step else None, path=path)
            np.testing.assert_allclose(got, e["loss"], rtol=RTOL, err_msg="%s %s" % (path, layout))
            _close(m._bufs["pred"][:bs].cpu().numpy(), e["pred"], _pred_scale(e))
        _assert_model_state(m, ref, k, 0.05, 2)
        pred = m.predict(lines[:bs], path="composed" if path == "composed" else "claim")
        f = ref.forward(lines[:bs])
        _close(pred.cpu().numpy(), f["pred"], _pred_scale(f))


# BASELINE.json shapes cut down only in batch (and, where the host cannot hold the oracle's copy of the table, in
# vocabulary): hashed 2^24 x K=16 with 39 features per line (cfg2), K=64 (cfg3: lane groups of 17), mse + weight file
# with 1..512 features per line at K=32 (cfg4: the per-example loop runs up to 16 rounds of 32 rows)
_SHAPES = {
    "cfg2": dict(vocab=1 << 24, k=16, bs=4096, hashed=True, loss_type="logistic", lam=1e-5),
    "cfg3": dict(vocab=1 << 22, k=64, bs=2048, hashed=True, loss_type="logistic", lam=1e-5),
    "cfg4": dict(vocab=1 << 22, k=32, bs=192, hashed=False, loss_type="mse", lam=1e-4),
}


@pytest.mark.parametrize("path", [p for p in PATHS if p != "composed"] + ["claim/rows", "claim/occurrences"])
@pytest.mark.parametrize("shape", sorted(_SHAPES))
def test_whole_step_at_baseline_shapes(shape, path):
    path, _, form = path.partition("/")              # the claim path's two forms, pinned (default: it measures)
    c = _SHAPES[shape]
    vocab, k, bs = c["vocab"], c["k"], c["bs"]
    rng = np.random.default_rng(7)
    flat = (rng.random((vocab, k + 1), dtype=np.float32) - 0.5) * 0.02
    kw = dict(factor_lambda=c["lam"], bias_lambda=c["lam"], loss_type=c["loss_type"])
    ref = O.RefModel(vocab, 1, k, flat, batch_size=bs, hash_feature_id=c["hashed"], lr=0.01, init_acc=0.1, **kw)
    m = Model(vocabulary_size=vocab, vocabulary_block_num=1, factor_num=k, batch_size=bs, hash_feature_id=c["hashed"],
              learning_rate=0.01, adagrad_init_accumulator=0.1, **kw).build(DEV, init_table=flat)
    if form:
        m.claim_form = form
    touched = []
    steps = 2
    for step in range(steps):
        if shape == "cfg4":
            lines, w = bench_data.ragged_batch(bs, 30 + step, max_nnz=512, vocab=vocab)
        else:
            text, off, _ = bench_data.criteo_batch(bs, 30 + step)
            lines, w = [bytes(text[off[i]:off[i + 1] - 1]) for i in range(bs)], None
        e = ref.train_step(lines, w)
        got = m.train_step(lines, w, path=path)
        np.testing.assert_allclose(got, e["loss"], rtol=RTOL)
        _close(m._bufs["pred"][:bs].cpu().numpy(), e["pred"], _pred_scale(e))
        np.testing.assert_allclose(float(m._bufs["scalars"][2]), e["reg"], rtol=RTOL)
        touched.append(e["ori"])
    # the rows the steps touched, and a sample of the ones they must have left alone
    rows = np.unique(np.concatenate(touched + [rng.integers(0, vocab, 20000)]))
    idx = torch.from_numpy(rows).to(DEV)
    _close(m.table[idx, :k + 1].cpu().numpy(), ref.blocks[0, rows], 0.01 * steps)
    _close(m.accum[idx, :k + 1].cpu().numpy(), ref.acc[0, rows], 0.0)
    f = ref.forward(lines)
    _close(m.predict(lines).cpu().numpy(), f["pred"], _pred_scale(f))


def _cfg1_files():
    import glob
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lines, weights = [], []
    for f, wf in zip(sorted(glob.glob(os.path.join(root, "data", "train_*"))),
                     sorted(glob.glob(os.path.join(root, "data", "weight_*")))):
        with open(f, "rb") as fh:
            part = fh.read().split(b"\n")
        if part and part[-1] == b"":
            part.pop()
        lines += part
        with open(wf) as fh:
            weights += [float(x) for x in fh.read().split()]
    return os.path.join(root, "sample.cfg"), lines, np.float32(weights)


@pytest.mark.parametrize("loss_type,steps", [("mse", 20), ("logistic", 6)])
def test_cfg1_sample_cfg_on_the_reference_fixtures(loss_type, steps, capsys, monkeypatch):
    """BASELINE.json config 1 as written: sample.cfg unmodified (V = 800 000 in 10 blocks, K = 100, batch 50 000,
    Adagrad(0.01, 0.1), mse with the weight files, 20 epochs = 20 steps over data/train_0..4) and once with logistic
    loss, files in sorted order, no shuffle, explicit initial table: every step's loss against the oracle model
    (tffm/fm_model.py:311-348 restated), then the trained table"""
    cfg, lines, weights = _cfg1_files()
    monkeypatch.chdir(os.path.dirname(cfg))          # sample.cfg names its files relative to the repository root
    m = Model(cfg)
    capsys.readouterr()
    assert (m.vocabulary_size, m.vocabulary_block_num, m.factor_num, m.batch_size) == (800000, 10, 100, 50000)
    assert m.loss_type == "mse" and m.num_epochs == 20 and len(m.train_files) == 5 and len(lines) == 50000
    m.loss_type = loss_type
    use_w = loss_type == "mse"          # sample.cfg pairs the weight files with the mse objective
    rng = np.random.default_rng(1)
    k = m.factor_num
    flat = rng.uniform(-m.init_value_range, m.init_value_range, (m.vocabulary_size, k + 1)).astype(np.float32)
    ref = O.RefModel(m.vocabulary_size, m.vocabulary_block_num, k, flat, batch_size=m.batch_size,
                     factor_lambda=m.factor_lambda, bias_lambda=m.bias_lambda, lr=m.learning_rate,
                     init_acc=m.adagrad_init_accumulator, loss_type=loss_type)
    m.build(DEV, init_table=flat)
    packed = fm_ops.pack_lines(lines).to(DEV, False)
    w = weights if use_w else None
    w_dev = None if w is None else dev(w, torch.float32)
    for step in range(steps):
        e = ref.train_step(lines, w)
        got = m.train_step(packed, w_dev)
        np.testing.assert_allclose(got, e["loss"], rtol=RTOL, err_msg="step %d" % step)
    _close(m._bufs["pred"][:50000].cpu().numpy(), e["pred"], _pred_scale(e))
    _assert_model_state(m, ref, k, m.learning_rate, steps)


def test_parse_ahead_pipeline_matches_inline_steps(criteo_lines):
    """parse of batch t+1 on a side stream under batch t's kernels: same table as step-by-step"""
    lines, _ = criteo_lines
    rng = np.random.default_rng(33)
    vocab, k, bs = 200003, 16, 250
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    kw = dict(vocabulary_size=vocab, factor_num=k, batch_size=bs, learning_rate=0.05, loss_type="logistic",
              factor_lambda=1e-3, bias_lambda=1e-3)
    a = Model(**kw).build(DEV, init_table=flat, layout="interleaved")
    b = Model(**kw).build(DEV, init_table=flat)
    batches = [fm_ops.pack_lines(lines[i * bs:(i + 1) * bs]).to(DEV, False) for i in range(6)]
    ahead = a.parse_ahead(batches[0])
    losses = []
    for i in range(6):
        nxt = a.parse_ahead(batches[i + 1]) if i + 1 < 6 else None
        out = a.train_step_async(batches[i], parsed=ahead)
        losses.append(out["status_snapshot"])
        ahead = nxt
    exp = [b.train_step(batches[i]) for i in range(6)]
    got = [float(s.cpu()[6:8].view(torch.float32)[0]) for s in losses]
    np.testing.assert_allclose(got, exp, rtol=1e-6)
    np.testing.assert_allclose(a.table[:, :k + 1].cpu().numpy(), b.table[:, :k + 1].cpu().numpy(), rtol=1e-5, atol=1e-8)


def test_background_checkpoint_is_a_consistent_snapshot(tmp_path, criteo_lines):
    """save(wait=False): the file holds the table as it was when save() was called, in the reference's
    vocab_block_i layout, although training went on while it was being written"""
    lines, _ = criteo_lines
    vocab, nb, k = 170003, 7, 9
    kw = dict(vocabulary_size=vocab, vocabulary_block_num=nb, factor_num=k, batch_size=200, learning_rate=0.05)
    m = Model(**kw).build(DEV, seed=1)
    m.train_step(lines[:200])
    table, accum, blocks = m.table[:, :k + 1].clone(), m.accum[:, :k + 1].clone(), m.blocks_state()
    path = str(tmp_path / "ckpt" / "model.ckpt.pt")
    m.save(path, wait=False)
    for i in range(1, 6):
        m.train_step(lines[200 * i:200 * (i + 1)])
    m.wait_for_save()
    assert not torch.equal(m.table[:, :k + 1], table)
    st = torch.load(path)
    assert st["global_step"] == 1 and st["vocab_block_3"].shape == (vocab // nb + 1, k + 1)
    for key, val in blocks.items():
        if key != "global_step":
            assert torch.equal(st[key], val), key
    m2 = Model(**kw).build(DEV, seed=2)
    m2.restore(path)
    assert torch.equal(m2.table[:, :k + 1], table) and torch.equal(m2.accum[:, :k + 1], accum) and m2.global_step == 1
    m.save(path)                                             # synchronous form, same file name
    m2.restore(path)
    assert torch.equal(m2.table[:, :k + 1], m.table[:, :k + 1]) and m2.global_step == 6


def test_host_batch_pipeline_matches_step_by_step(criteo_lines):
    """copy | parse | train pipeline over host batches (pinned and not, growing sizes, weights on some, one
    batch that does not parse): same losses and the same table as train_step on each batch in turn"""
    from fast_tffm_b200.pipeline import HostBatchPipeline
    lines, weights = criteo_lines
    rng = np.random.default_rng(44)
    vocab, k = 200003, 16
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    kw = dict(vocabulary_size=vocab, factor_num=k, batch_size=300, learning_rate=0.05, loss_type="logistic",
              factor_lambda=1e-3, bias_lambda=1e-3)
    a = Model(**kw).build(DEV, init_table=flat)
    b = Model(**kw).build(DEV, init_table=flat)
    sizes = [100, 300, 250, 700, 300, 50, 1, 150, 149]
    cuts = np.concatenate([[0], np.cumsum(sizes)])
    batches = []
    for i, n in enumerate(sizes):
        ls = list(lines[cuts[i]:cuts[i + 1]])
        if i == 4:
            ls[17] = ls[17] + " 12:oops"
        w = torch.from_numpy(np.float32(weights[cuts[i]:cuts[i + 1]])) if i % 2 else None
        batches.append((fm_ops.pack_lines(ls, pin_memory=(i % 3 != 0)), w, ls))
    pipe = HostBatchPipeline(a)
    results = list(pipe.run((p, w) for p, w, _ in batches))
    assert [r.n_lines for r in results] == sizes
    for i, (p, w, ls) in enumerate(batches):
        if i == 4:
            with pytest.raises(fm_ops.FmParseError, match="Invalid feature value") as e:
                results[i].loss()
            assert "12:oops" in str(e.value)
            with pytest.raises(fm_ops.FmParseError):
                b.train_step(ls, w)
            continue
        np.testing.assert_allclose(results[i].loss(), b.train_step(ls, None if w is None else w.numpy()), rtol=1e-6)
    pipe.drain()
    np.testing.assert_allclose(a.table[:, :k + 1].cpu().numpy(), b.table[:, :k + 1].cpu().numpy(), rtol=1e-5, atol=1e-8)
    assert a.global_step == b.global_step == len(sizes) - 1      # the void step is not counted
    # a second run on the same pipeline reuses the staging slots
    more = list(pipe.run((p, w) for p, w, _ in batches[:3]))
    for i, r in enumerate(more):
        np.testing.assert_allclose(r.loss(), b.train_step(batches[i][2], None if batches[i][1] is None else batches[i][1].numpy()),
                                   rtol=1e-6)


def test_claim_path_picks_its_form_from_the_ids_it_sees():
    """claim_form = "auto": batches whose ids repeat a lot run the row-list form, batches of distinct ids the
    occurrence walk; the switch follows the ids with a few steps' delay and never changes the result"""
    rng = np.random.default_rng(77)
    vocab, k, bs = 50021, 8, 256

    def batch(hot):
        out = []
        for _ in range(bs):
            ids = rng.integers(0, 30 if hot else vocab, 20)
            out.append(("1 " + " ".join("%d:%.2f" % (t, rng.normal()) for t in ids)).encode())
        return out
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    ref = O.RefModel(vocab, 1, k, flat, batch_size=bs, lr=0.05, init_acc=0.1, loss_type="logistic")
    m = Model(vocabulary_size=vocab, factor_num=k, batch_size=bs, learning_rate=0.05, adagrad_init_accumulator=0.1,
              loss_type="logistic").build(DEV, init_table=flat)
    assert m.claim_form == "auto"
    forms = []
    for step in range(12):
        lines = batch(hot=step < 6)
        e = ref.train_step(lines, None)
        np.testing.assert_allclose(m.train_step(lines), e["loss"], rtol=RTOL)     # train_step synchronises
        forms.append(m._form)
    assert forms[2:6] == ["rows"] * 4 and forms[-3:] == ["occurrences"] * 3, forms
    _assert_model_state(m, ref, k, 0.05, 12)
    m.FORM_PROBE_EVERY = 2                       # a look at the ids every other step: back to the row list
    for step in range(6):
        lines = batch(hot=True)
        e = ref.train_step(lines, None)
        np.testing.assert_allclose(m.train_step(lines), e["loss"], rtol=RTOL)
    assert m._form == "rows"
    _assert_model_state(m, ref, k, 0.05, 18)


def test_predict_with_parse_ahead_equals_predict(criteo_lines):
    """a stream of predict batches with the next batch's parse running ahead: the same scores as one at a time"""
    lines, _ = criteo_lines
    m = Model(vocabulary_size=200003, factor_num=16, batch_size=256, hash_feature_id=True).build(DEV, seed=5, layout="split")
    batches = [fm_ops.pack_lines(list(lines[i * 300:(i + 1) * 300 - 7 * i])).to(DEV, False) for i in range(5)]
    want = [m.predict(p).cpu().numpy() for p in batches]
    got = []
    ahead = m.parse_ahead(batches[0], sort=False)
    for i, p in enumerate(batches):
        nxt = m.parse_ahead(batches[i + 1], sort=False) if i + 1 < len(batches) else None
        b = m.predict_async(p, parsed=ahead)
        got.append(b["pred"][:p.n_lines].clone())     # before the set is refilled (stream-ordered: no sync needed)
        ahead = nxt
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g.cpu().numpy(), w)


def test_failed_parse_leaves_the_table_untouched():
    m = Model(vocabulary_size=100, factor_num=4, batch_size=2).build(DEV, seed=0)
    before = m.table.clone()
    with pytest.raises(fm_ops.FmParseError, match="Invalid feature value"):
        m.train_step(["1 2:3", "0 4:x"])
    assert torch.equal(m.table, before) and m.global_step == 0
    m.train_step(["1 2:3", "0 4:1"])
    assert not torch.equal(m.table, before) and m.global_step == 1


def test_full_size_step_properties():
    """cfg2 at BASELINE size: size-independent properties of one step"""
    text, off, nnz = bench_data.criteo_batch(65536, 7)
    m = Model(vocabulary_size=1 << 24, factor_num=16, batch_size=65536, hash_feature_id=True, loss_type="logistic",
              factor_lambda=1e-5, bias_lambda=1e-5).build(DEV, seed=3, max_nnz_per_line=39)
    packed = fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off)).to(DEV, False)
    table0 = m.table.clone()
    b = m.train_step_async(packed, path="fused")
    st = b["status"].cpu()
    assert int(st[0]) == -1 and int(st[1]) == nnz
    u = int(st[4])
    ids, ori, local = b["ids"][:nnz], b["ori"][:u], b["local"][:nnz]
    assert torch.equal(ori[local.long()], ids)                        # tf.unique invariant
    assert u == torch.unique(ids).numel()
    poses = b["poses"][:65537]
    assert int(poses[0]) == 0 and int(poses[-1]) == nnz and bool((poses[1:] - poses[:-1] == 39).all())
    changed = (m.table != table0).any(1).nonzero().flatten()
    touched = torch.zeros(1 << 24, dtype=torch.bool, device=DEV)
    touched[ori] = True
    assert bool(touched[changed].all())                                # only the batch's rows were updated
    assert changed.numel() > 0.9 * u                                   # (rows whose x == 0 move by less than an ulp)
    loss = float(st[6:8].view(torch.float32)[0])
    assert 0.6 < loss < 0.8                                            # ~log 2 for a fresh model


def test_sorted_path_at_full_size_is_reproducible():
    """cfg2 at BASELINE size through the sorted path: the counts the sort and the update report, the same rows
    touched as the claim path, and -- what the stable sort and the atomic-free segment reduce are for -- the same
    bits in the table when the same two steps are run again from the same start (zipf ids: hot rows, cut runs)"""
    kw = dict(vocabulary_size=1 << 24, factor_num=16, batch_size=65536, hash_feature_id=True, loss_type="logistic",
              factor_lambda=1e-5, bias_lambda=1e-5)
    batches = []
    for seed, dist in ((7, "uniform"), (8, "zipf")):
        text, off, nnz = bench_data.criteo_batch(65536, seed, dist)
        batches.append((fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off)).to(DEV, False), nnz))
    tables = []
    for run in range(2):
        m = Model(**kw).build(DEV, seed=3, max_nnz_per_line=39, path="sorted")
        for packed, nnz in batches:
            b = m.train_step_async(packed)
            st = b["status"].cpu()
            assert int(st[0]) == -1 and int(st[1]) == nnz and int(st[3]) == nnz
            assert int(st[4]) == torch.unique(b["ids"][:nnz]).numel()          # tf.unique's U, counted by the update
        tables.append((m.table.clone(), m.accum.clone()))
        del m
    assert torch.equal(tables[0][0], tables[1][0]) and torch.equal(tables[0][1], tables[1][1])
    c = Model(**kw).build(DEV, seed=3, max_nnz_per_line=39, path="claim")
    for packed, _ in batches:
        c.train_step_async(packed)
    torch.testing.assert_close(tables[0][0][:, :17], c.table[:, :17], rtol=1e-5, atol=1e-7)
    assert torch.equal((tables[0][1][:, :17] != 0.1).any(1), (c.accum[:, :17] != 0.1).any(1))   # the same rows moved
