"""ctypes/numpy face of oracle/fm_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module (as the checker or the timed
CPU baseline).  The product package fast_tffm_b200 never does.

Every function names the reference code it restates; see fm_oracle.c for the
line-by-line citations.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libfm_oracle.so")

ERR_PREFIX = {
    1: "Label could not be read in example: ",       # fm_parser_op.cc:70
    2: "Invalid format in example: ",                # fm_parser_op.cc:77,88
    3: "Invalid feature id. Should be in range [0, vocabulary_size).",  # :90-92
    4: "Invalid feature value. ",                    # fm_parser_op.cc:100
    5: "oracle: output capacity too small",
}


def build(force=False):
    src = os.path.join(_HERE, "fm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libfm_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.fmo_hash64.restype = C.c_uint64
        _lib.fmo_hash64_seed.restype = C.c_uint64
        _lib.fmo_unique.restype = C.c_int64
        _lib.fmo_count_spaces.restype = C.c_int64
        _lib.fmo_loss.restype = C.c_double
        _lib.fmo_loss_f32.restype = C.c_float
        _lib.fmo_train_step_faithful.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class ParseError(ValueError):
    def __init__(self, kind, line_no, line):
        self.kind, self.line_no, self.line = kind, line_no, line
        super().__init__(ERR_PREFIX[kind] + line)


def pack_lines(lines):
    """list of str/bytes -> (u8 text, i64 offsets[B+1]); line i is
    text[off[i]:off[i+1]-1], each line followed by one terminator byte."""
    bl = [s.encode() if isinstance(s, str) else bytes(s) for s in lines]
    off = np.zeros(len(bl) + 1, np.int64)
    if bl:
        off[1:] = np.cumsum(np.fromiter((len(b) + 1 for b in bl), np.int64, len(bl)))
    text = np.frombuffer(b"".join(b + b"\n" for b in bl), np.uint8).copy() if bl else np.zeros(0, np.uint8)
    return text, off


def hash64(data, seed=None):
    b = data.encode() if isinstance(data, str) else bytes(data)
    buf = (C.c_uint8 * max(len(b), 1)).from_buffer_copy(b + (b"\0" if not b else b""))
    if seed is None:
        return lib().fmo_hash64(buf, C.c_uint64(len(b)))
    return lib().fmo_hash64_seed(buf, C.c_uint64(len(b)), C.c_uint64(seed))


def strtof(s):
    """glibc strtof -> (float32, consumed bytes)"""
    b = (s.encode() if isinstance(s, str) else bytes(s)).split(b"\0")[0] + b"\0"
    out = C.c_float()
    n = lib().fmo_strtof(b, C.byref(out))
    return np.float32(out.value), n


def strtoll(s):
    b = (s.encode() if isinstance(s, str) else bytes(s)).split(b"\0")[0] + b"\0"
    out = C.c_int64()
    n = lib().fmo_strtoll(b, C.byref(out))
    return out.value, n


def parse(text, off, vocab_size, hash_feature_id=False, raise_on_error=True):
    """FmParser (cc/fm_parser_op.cc:29-109) -> labels, sizes, ids(i64), vals"""
    text = np.ascontiguousarray(text, np.uint8)
    off = np.ascontiguousarray(off, np.int64)
    b = len(off) - 1
    cap = int(lib().fmo_count_spaces(_p(text), C.c_int64(text.size))) + 1
    labels = np.zeros(b, np.float32)
    sizes = np.zeros(b, np.int32)
    ids = np.zeros(cap, np.int64)
    vals = np.zeros(cap, np.float32)
    nnz, bad = C.c_int64(0), C.c_int64(-1)
    rc = lib().fmo_parse(_p(text), _p(off), C.c_int64(b), C.c_int64(vocab_size),
                         C.c_int(1 if hash_feature_id else 0), _p(labels), _p(sizes), _p(ids),
                         _p(vals), C.c_int64(cap), C.byref(nnz), C.byref(bad))
    if rc != 0:
        i = bad.value
        line = bytes(text[off[i]:off[i + 1] - 1]).split(b"\0")[0].decode("utf-8", "replace")
        if raise_on_error:
            raise ParseError(rc, i, line)
        return rc, i
    n = nnz.value
    return labels, sizes, ids[:n].copy(), vals[:n].copy()


def parse_lines(lines, vocab_size, hash_feature_id=False):
    return parse(*pack_lines(lines), vocab_size, hash_feature_id)


def unique(ids):
    """tf.unique (tffm/fm_model.py:72) -> ori_ids i64[U], local i32[N]"""
    ids = np.ascontiguousarray(ids, np.int64)
    ori = np.zeros(max(ids.size, 1), np.int64)
    local = np.zeros(ids.size, np.int32)
    u = lib().fmo_unique(_p(ids), C.c_int64(ids.size), _p(ori), _p(local))
    return ori[:u].copy(), local


def poses(sizes):
    """concat([0], cumsum(sizes)) (tffm/fm_model.py:73)"""
    sizes = np.ascontiguousarray(sizes, np.int32)
    out = np.zeros(sizes.size + 1, np.int32)
    lib().fmo_poses(_p(sizes), C.c_int64(sizes.size), _p(out))
    return out


def block_rows(vocab, nb):
    return vocab // nb + 1


def lookup_mod(blocks, vocab, nb, ori_ids):
    """tf.nn.embedding_lookup(vocab_blocks, ids), mod strategy (fm_model.py:269-291).
    blocks: f32[nb, vocab//nb+1, K+1]"""
    assert blocks.dtype == np.float32 and blocks.flags.c_contiguous
    k = blocks.shape[2] - 1
    ori_ids = np.ascontiguousarray(ori_ids, np.int64)
    out = np.zeros((ori_ids.size, k + 1), np.float32)
    lib().fmo_lookup_mod(_p(blocks), C.c_int64(vocab), C.c_int64(nb), C.c_int64(k), _p(ori_ids),
                         C.c_int64(ori_ids.size), _p(out))
    return out


def score(fids, params, vals, poses_, factor_lambda, bias_lambda, k=None, faithful=False, nthreads=1):
    """FmScorer (cc/fm_scorer_op.h) -> pred f32[B], reg f32"""
    fids = np.ascontiguousarray(fids, np.int32)
    params = np.ascontiguousarray(params, np.float32)
    vals = np.ascontiguousarray(vals, np.float32)
    poses_ = np.ascontiguousarray(poses_, np.int32)
    ld = params.shape[1]
    k = ld - 1 if k is None else k
    b = poses_.size - 1
    pred = np.zeros(b, np.float32)
    reg = C.c_float(0)
    if faithful:
        assert ld == k + 1
        lib().fmo_score_faithful(_p(fids), _p(params), _p(vals), _p(poses_), C.c_int64(b), C.c_int64(k),
                                 C.c_float(factor_lambda), C.c_float(bias_lambda), _p(pred),
                                 C.byref(reg), C.c_int(nthreads))
    else:
        lib().fmo_score_det(_p(fids), _p(params), C.c_int64(ld), _p(vals), _p(poses_), C.c_int64(b),
                            C.c_int64(k), C.c_float(factor_lambda), C.c_float(bias_lambda),
                            _p(pred), C.byref(reg))
    return pred, np.float32(reg.value)


def grad(fids, params, vals, poses_, factor_lambda, bias_lambda, pred_grad, reg_grad, k=None,
         faithful=False, nthreads=1):
    """FmGrad (cc/fm_grad_op.h) -> params_grad, same shape as params"""
    fids = np.ascontiguousarray(fids, np.int32)
    params = np.ascontiguousarray(params, np.float32)
    vals = np.ascontiguousarray(vals, np.float32)
    poses_ = np.ascontiguousarray(poses_, np.int32)
    pred_grad = np.ascontiguousarray(pred_grad, np.float32)
    ld = params.shape[1]
    k = ld - 1 if k is None else k
    b = poses_.size - 1
    out = np.zeros_like(params)
    if faithful:
        assert ld == k + 1
        lib().fmo_grad_faithful(_p(fids), _p(params), _p(vals), _p(poses_), C.c_int64(b), C.c_int64(k),
                                C.c_float(factor_lambda), C.c_float(bias_lambda), _p(pred_grad),
                                C.c_float(reg_grad), C.c_int64(params.shape[0]), _p(out),
                                C.c_int(nthreads))
    else:
        lib().fmo_grad_det(_p(fids), _p(params), C.c_int64(ld), _p(vals), _p(poses_), C.c_int64(b),
                           C.c_int64(k), C.c_float(factor_lambda), C.c_float(bias_lambda),
                           _p(pred_grad), C.c_float(reg_grad), C.c_int64(params.shape[0]), _p(out))
    return out


LOSS_KIND = {"logistic": 0, "mse": 1}


def loss(pred, labels, weights, loss_type):
    """weighted mean loss + d loss / d pred (tffm/fm_model.py:311-326)"""
    pred = np.ascontiguousarray(pred, np.float32)
    labels = np.ascontiguousarray(labels, np.float32)
    weights = None if weights is None else np.ascontiguousarray(weights, np.float32)
    pg = np.zeros_like(pred)
    val = lib().fmo_loss(_p(pred), _p(labels), _p(weights), C.c_int64(pred.size),
                         C.c_int(LOSS_KIND[loss_type]), _p(pg))
    return val, pg


def adagrad_mod(blocks, acc_blocks, vocab, nb, ori_ids, grad_rows, lr):
    """TF1 SparseApplyAdagrad on the mod-partitioned table, in place"""
    k = blocks.shape[2] - 1
    ori_ids = np.ascontiguousarray(ori_ids, np.int64)
    grad_rows = np.ascontiguousarray(grad_rows, np.float32)
    assert grad_rows.shape == (ori_ids.size, k + 1)
    lib().fmo_adagrad_mod(_p(blocks), _p(acc_blocks), C.c_int64(vocab), C.c_int64(nb), C.c_int64(k),
                          _p(ori_ids), C.c_int64(ori_ids.size), _p(grad_rows), C.c_float(lr))


class RefModel:
    """The reference's local-mode train / predict step strung together the way
    tffm/fm_model.py does (:69-73 parse+unique+cumsum, :291-293 lookup+score,
    :311-348 loss, objective loss + reg / batch_size, Adagrad)."""

    def __init__(self, vocab, nb, k, init_table, *, batch_size, hash_feature_id=False,
                 factor_lambda=0.0, bias_lambda=0.0, lr=0.01, init_acc=0.1, loss_type="mse"):
        self.vocab, self.nb, self.k = vocab, nb, k
        rows = block_rows(vocab, nb)
        self.blocks = np.zeros((nb, rows, k + 1), np.float32)
        # init_table is a flat [vocab, K+1] table: row id lives in block id % nb, row id // nb
        ids = np.arange(vocab)
        self.blocks[ids % nb, ids // nb] = np.asarray(init_table, np.float32)[:, :k + 1]
        self.acc = np.full_like(self.blocks, np.float32(init_acc))
        self.batch_size, self.hash = batch_size, hash_feature_id
        self.lf, self.lb, self.lr, self.loss_type = factor_lambda, bias_lambda, lr, loss_type

    def flat_table(self):
        ids = np.arange(self.vocab)
        return self.blocks[ids % self.nb, ids // self.nb].copy()

    def flat_acc(self):
        ids = np.arange(self.vocab)
        return self.acc[ids % self.nb, ids // self.nb].copy()

    def forward(self, lines):
        labels, sizes, ids, vals = parse_lines(lines, self.vocab, self.hash)
        ori, local = unique(ids)
        ps = poses(sizes)
        lp = lookup_mod(self.blocks, self.vocab, self.nb, ori)
        pred, reg = score(local, lp, vals, ps, self.lf, self.lb)
        return dict(labels=labels, sizes=sizes, ids=ids, vals=vals, ori=ori, local=local, poses=ps,
                    lp=lp, pred=pred, reg=reg)

    def predict(self, lines):
        return self.forward(lines)["pred"]

    def train_step(self, lines, weights=None):
        f = self.forward(lines)
        lv, pg = loss(f["pred"], f["labels"], weights, self.loss_type)
        g = grad(f["local"], f["lp"], f["vals"], f["poses"], self.lf, self.lb, pg,
                 1.0 / self.batch_size)
        adagrad_mod(self.blocks, self.acc, self.vocab, self.nb, f["ori"], g, self.lr)
        f.update(loss=lv, pred_grad=pg, grad=g)
        return f


def train_step_faithful(text, off, vocab, hash_feature_id, nb, k, blocks, acc_blocks, weights,
                        loss_type, lam_f, lam_b, lr, cfg_batch, nthreads, do_update=True):
    """Timed CPU baseline: one reference step, float arithmetic, Eigen-style
    parallel-for over coefficients.  Returns (loss, per-stage seconds[7], pred)."""
    text = np.ascontiguousarray(text, np.uint8)
    off = np.ascontiguousarray(off, np.int64)
    b = off.size - 1
    cap = int(lib().fmo_count_spaces(_p(text), C.c_int64(text.size))) + 1
    t = np.zeros(7, np.float64)
    pred = np.zeros(b, np.float32)
    w = None if weights is None else np.ascontiguousarray(weights, np.float32)
    lv = lib().fmo_train_step_faithful(
        _p(text), _p(off), C.c_int64(b), C.c_int64(vocab), C.c_int(1 if hash_feature_id else 0),
        C.c_int64(nb), C.c_int64(k), _p(blocks), _p(acc_blocks), _p(w),
        C.c_int(LOSS_KIND[loss_type]), C.c_float(lam_f), C.c_float(lam_b), C.c_float(lr),
        C.c_int64(cfg_batch), C.c_int64(cap), C.c_int(nthreads), C.c_int(1 if do_update else 0),
        _p(t), _p(pred))
    return lv, t, pred
