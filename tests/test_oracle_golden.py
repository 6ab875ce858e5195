"""Pins the CPU oracle against the reference's own test vectors (CPU only).

Ports of test/fm_parser_op_test.py and test/fm_scorer_op_test.py with the
oracle standing where the TF op stood, plus the independent dense closed form.
"""
import numpy as np
import pytest
import torch

from oracle import dense_ref, oracle as O


def test_parser_no_hash(golden_parser):
    g = golden_parser
    labels, sizes, ids, vals = O.parse_lines(g["examples"], g["vocab_size"])
    np.testing.assert_allclose(labels, g["labels"], rtol=1e-6, atol=1e-6)
    np.testing.assert_array_equal(ids, g["feature_ids"])
    np.testing.assert_allclose(vals, np.float32(g["feature_vals"]), rtol=1e-6, atol=1e-6)
    np.testing.assert_array_equal(sizes, g["sizes"])


def test_parser_with_hash(golden_parser):
    g = golden_parser
    labels, sizes, ids, vals = O.parse_lines(g["examples"], g["vocab_size"], True)
    expect = [O.hash64(str(x)) % g["vocab_size"] for x in g["feature_ids"]]
    np.testing.assert_array_equal(ids, expect)
    np.testing.assert_array_equal(ids, g["hashed_ids_1_to_10_mod_10000_selfgenerated"])
    np.testing.assert_array_equal(sizes, g["sizes"])
    np.testing.assert_allclose(vals, np.float32(g["feature_vals"]), rtol=1e-6)


def test_parser_errors(golden_parser):
    for line, prefix in golden_parser["errors"]:
        with pytest.raises(ValueError, match=prefix.replace(".", r"\.")):
            O.parse_lines([line], golden_parser["vocab_size"], False)
    with pytest.raises(ValueError, match="Invalid feature id"):
        O.parse_lines(["1 10000:1"], 10000)
    with pytest.raises(ValueError, match="Invalid feature id"):
        O.parse_lines(["1 -1:1"], 10000)


def test_first_failing_line_wins():
    with pytest.raises(O.ParseError) as e:
        O.parse_lines(["1 1:1", "1 1:x", "y 1:1"], 100)
    assert e.value.line_no == 1 and e.value.kind == 4


def test_parser_quirks_inherited_from_libc():
    # strtof / strtoll skip leading whitespace (fm_parser_op.cc:69,86,98)
    labels, sizes, ids, vals = O.parse_lines([" 2  7: 3", "1"], 100)
    assert labels.tolist() == [2.0, 1.0] and sizes.tolist() == [1, 0]
    assert ids.tolist() == [7] and vals.tolist() == [3.0]
    # hash mode: an empty token is hashed, not rejected (:82-84)
    _, sizes, ids, _ = O.parse_lines(["1  a"], 1000, True)
    assert sizes.tolist() == [2] and ids[0] == O.hash64("") % 1000
    # embedded NUL ends the line (std::string::c_str())
    labels, sizes, _, _ = O.parse_lines([b"1 2:3\x00 garbage"], 100)
    assert sizes.tolist() == [1]


def test_hash64_known_answers(golden_parser):
    g = golden_parser
    for hexdata, hexhash in g["hash64_kat"]:
        assert O.hash64(bytes.fromhex(hexdata), g["hash64_kat_seed"]) == int(hexhash, 16)
    for s, hexhash in g["hash64_long_selfgenerated"]:
        assert O.hash64(s) == int(hexhash, 16)


def test_number_vectors_match_glibc(golden_numbers):
    for s, bits, n, *quirk in golden_numbers["strtof"]:
        v, consumed = O.strtof(s)
        if quirk:  # glibc hex-subnormal rounding bug: the fixture pins the IEEE value instead
            bits = quirk[0]["glibc_2_39_bits"]
        assert (int(np.float32(v).view(np.uint32)), consumed) == (bits, n), s
    for s, val, n in golden_numbers["strtoll"]:
        assert O.strtoll(s) == (val, n), s


def test_criteo_excerpt_against_split_parser(criteo_lines):
    lines, _ = criteo_lines
    labels, sizes, ids, vals = O.parse_lines(lines, 800000)
    e_ids, e_vals, e_sizes, e_labels = [], [], [], []
    for l in lines:
        toks = l.split(" ")
        e_labels.append(float(toks[0]))
        feats = [t for t in toks[1:] if t]
        e_sizes.append(len(feats))
        for t in feats:
            a, b = t.split(":")
            e_ids.append(int(a))
            e_vals.append(float(b))
    np.testing.assert_array_equal(ids, e_ids)
    np.testing.assert_array_equal(sizes, e_sizes)
    np.testing.assert_array_equal(vals, np.float32(e_vals))
    np.testing.assert_array_equal(labels, np.float32(e_labels))


def test_unique_and_poses():
    rng = np.random.default_rng(0)
    ids = rng.integers(0, 50, 1000)
    ori, local = O.unique(ids)
    _, first = np.unique(ids, return_index=True)
    np.testing.assert_array_equal(ori, ids[np.sort(first)])
    np.testing.assert_array_equal(ori[local], ids)
    sizes = rng.integers(0, 9, 100).astype(np.int32)
    np.testing.assert_array_equal(O.poses(sizes), np.concatenate([[0], np.cumsum(sizes)]))
    ori, local = O.unique(np.zeros(0, np.int64))
    assert ori.size == 0 and local.size == 0


def test_scorer_and_grad_reference_case(golden_scorer):
    """port of test/fm_scorer_op_test.py:65-78 (whole table + raw ids, no unique step)"""
    g = golden_scorer
    params = np.float32(g["params"])
    pred, reg = O.score(g["feature_ids"], params, g["feature_vals"], g["feature_poses"],
                        g["factor_lambda"], g["bias_lambda"])
    np.testing.assert_allclose(pred, g["pred"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(reg, g["reg"], rtol=1e-6, atol=1e-6)
    pred_grad = -2.0 * (np.float32(g["labels"]) - np.float64(g["pred"]))
    grad = O.grad(g["feature_ids"], params, g["feature_vals"], g["feature_poses"], g["factor_lambda"],
                  g["bias_lambda"], pred_grad, 1.0)
    np.testing.assert_allclose(grad, g["grad"], rtol=1e-6, atol=1e-6)
    # the faithful (float, powf, threaded) flavour is the same function
    for nt in (1, 4):
        p2, r2 = O.score(g["feature_ids"], params, g["feature_vals"], g["feature_poses"],
                         g["factor_lambda"], g["bias_lambda"], faithful=True, nthreads=nt)
        np.testing.assert_allclose(p2, g["pred"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(r2, g["reg"], rtol=1e-5)
        g2 = O.grad(g["feature_ids"], params, g["feature_vals"], g["feature_poses"], g["factor_lambda"],
                    g["bias_lambda"], pred_grad, 1.0, faithful=True, nthreads=nt)
        np.testing.assert_allclose(g2, g["grad"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("k,ld_pad", [(1, 0), (4, 3), (16, 0), (33, 3)])
def test_oracle_vs_dense_random(k, ld_pad):
    rng = np.random.default_rng(k)
    rows, b = 40, 17
    sizes = rng.integers(0, 12, b).astype(np.int32)
    sizes[3] = 0
    poses = O.poses(sizes)
    n = int(poses[-1])
    fids = rng.integers(0, rows, n).astype(np.int32)
    vals = rng.normal(size=n).astype(np.float32)
    params = np.zeros((rows, k + 1 + ld_pad), np.float32)
    params[:, :k + 1] = rng.uniform(-0.5, 0.5, (rows, k + 1))
    labels = rng.normal(size=b)
    pred, reg = O.score(fids, params, vals, poses, 0.3, 0.7, k=k)
    dp, dr, dg = dense_ref.dense_cost_and_grad(fids, torch.from_numpy(params[:, :k + 1]), vals, poses,
                                               0.3, 0.7, labels)
    np.testing.assert_allclose(pred, dp.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(reg, float(dr), rtol=1e-6)
    pg = -2.0 * (labels - dp.numpy())
    grad = O.grad(fids, params, vals, poses, 0.3, 0.7, pg, 1.0, k=k)
    np.testing.assert_allclose(grad[:, :k + 1], dg.numpy(), rtol=1e-5, atol=1e-6)
    assert not grad[:, k + 1:].any()


@pytest.mark.parametrize("loss_type", ["logistic", "mse"])
def test_loss_matches_torch(loss_type):
    rng = np.random.default_rng(1)
    pred = torch.tensor(rng.normal(size=64) * 3, dtype=torch.float64, requires_grad=True)
    y = torch.tensor(rng.integers(0, 2, 64), dtype=torch.float64)
    w = torch.tensor(rng.uniform(0.5, 2, 64), dtype=torch.float64)
    if loss_type == "logistic":
        l = (w * torch.nn.functional.binary_cross_entropy_with_logits(pred, y, reduction="none")).mean()
    else:
        l = (w * (pred - y) ** 2).mean()
    l.backward()
    lv, pg = O.loss(pred.detach().numpy(), y.numpy(), w.numpy(), loss_type)
    np.testing.assert_allclose(lv, float(l.detach()), rtol=1e-6)
    np.testing.assert_allclose(pg, pred.grad.numpy(), rtol=1e-5, atol=1e-8)


def test_mod_partition_lookup_and_adagrad():
    rng = np.random.default_rng(2)
    vocab, nb, k = 103, 10, 5
    flat = rng.normal(size=(vocab, k + 1)).astype(np.float32)
    m = O.RefModel(vocab, nb, k, flat, batch_size=4)
    assert m.blocks.shape == (nb, vocab // nb + 1, k + 1)
    ids = np.array([0, 9, 10, 57, 102])
    np.testing.assert_array_equal(O.lookup_mod(m.blocks, vocab, nb, ids), flat[ids])
    g = rng.normal(size=(ids.size, k + 1)).astype(np.float32)
    O.adagrad_mod(m.blocks, m.acc, vocab, nb, ids, g, 0.01)
    acc = 0.1 + g.astype(np.float64) ** 2
    np.testing.assert_allclose(m.flat_acc()[ids], acc, rtol=1e-6)
    np.testing.assert_allclose(m.flat_table()[ids], flat[ids] - 0.01 * g / np.sqrt(acc), rtol=1e-5, atol=1e-7)
    untouched = np.setdiff1d(np.arange(vocab), ids)
    np.testing.assert_array_equal(m.flat_table()[untouched], flat[untouched])


def test_train_step_decreases_loss_on_criteo_excerpt(criteo_lines):
    lines, weights = criteo_lines
    rng = np.random.default_rng(3)
    vocab, k = 200000, 8
    m = O.RefModel(vocab, 10, k, rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32),
                   batch_size=500, lr=0.05, loss_type="logistic")
    first = m.train_step(lines[:500], weights[:500])["loss"]
    for _ in range(5):
        last = m.train_step(lines[:500], weights[:500])["loss"]
    assert last < first


def test_faithful_train_step_matches_det(criteo_lines):
    lines, weights = criteo_lines
    rng = np.random.default_rng(4)
    vocab, k, nb = 200000, 4, 10
    init = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    m = O.RefModel(vocab, nb, k, init, batch_size=300, lr=0.05, loss_type="mse", factor_lambda=1e-3,
                   bias_lambda=1e-3)
    m2 = O.RefModel(vocab, nb, k, init, batch_size=300)
    text, off = O.pack_lines(lines[:300])
    out = m.train_step(lines[:300], weights[:300])
    lv, t, pred = O.train_step_faithful(text, off, vocab, False, nb, k, m2.blocks, m2.acc,
                                        np.float32(weights[:300]), "mse", 1e-3, 1e-3, 0.05, 300, 4)
    np.testing.assert_allclose(lv, out["loss"], rtol=1e-5)
    np.testing.assert_allclose(pred, out["pred"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(m2.blocks, m.blocks, rtol=1e-4, atol=1e-6)
    assert (t >= 0).all()
