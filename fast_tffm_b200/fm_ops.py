"""The reference's operator API (tffm/fm_ops.py) on B200.

Same three callables, argument order, dtypes and output tuples as the TF ops
``tf.load_op_library('lib/libfast_tffm.so')`` exposed:

    fm_parser(data_strings, vocab_size, hash_feature_id=False)
        -> labels f32[B], sizes i32[B], feature_ids i64[N], feature_vals f32[N]   (fm_ops.py:30-36)
    fm_scorer(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda)
        -> pred_score f32[B], reg_score f32[]                                    (fm_ops.py:39-56)
    fm_grad(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda,
            pred_grad, reg_grad) -> params_grad, shaped like feature_params       (fm_ops.py:22-25)

``fm_scorer`` is differentiable with respect to ``feature_params`` only and its
backward is ``fm_grad`` -- the registration ``@ops.RegisterGradient('FmScorer')``
made at fm_ops.py:11-27.  They are torch custom ops (``torch.ops.fast_tffm.*``)
over CUDA tensors, implemented by the sm_100a kernels behind the C ABI of
libfm_b200.so.  There is no CPU implementation: CPU tensors raise.

Plus the stock-TF ops the model runs between them (tffm/fm_model.py:72-73,
:291, :311-348): fm_unique, fm_lookup, fm_loss_grad, fm_sparse_adagrad.
"""
import ctypes as C
from typing import Tuple

import numpy as np
import torch

from . import _lib
from ._lib import check, load, ptr, require_cuda, stream, workspace

Tensor = torch.Tensor


class FmParseError(ValueError):
    """tf.errors.InvalidArgumentError stand-in; str() starts with the reference's message."""

    def __init__(self, kind, line_no, line):
        self.kind, self.line_no, self.line = kind, line_no, line
        super().__init__(_lib.PARSE_ERROR_PREFIX[kind] + line)


# --------------------------------------------------------------------------
# text packing (host side of the string tensor the TF op received)
# --------------------------------------------------------------------------
class PackedLines:
    """A batch of lines over one byte buffer (include/fm_b200.h, "Text layout").

    line i = text[begins[i] : ends[i] - 1].  Built from `offsets` (lines back to back, one
    terminator byte after every line: begins = offsets[:-1], ends = offsets[1:]) or from explicit
    begins / ends index tensors -- e.g. a shuffled selection of the lines of a file, see select().
    """

    def __init__(self, text: Tensor, offsets: Tensor = None, begins: Tensor = None, ends: Tensor = None,
                 max_line_bytes: int = None):
        self.text, self.offsets = text, offsets
        if offsets is not None:
            begins, ends = offsets[:-1], offsets[1:]
        self.begins, self.ends = begins, ends
        self.text_bytes = text.numel() // 16 * 16     # readable extent handed to the kernels
        # upper bound on ends[i] - begins[i] when known: a selection over a resident file shares the file's whole
        # text, so per-batch buffers are sized from the lines, not from the backing buffer (see nnz_bound)
        self.max_line_bytes = max_line_bytes

    def nnz_bound(self):
        """upper bound on the number of features in this batch: every feature costs at least two bytes (" x")"""
        bound = int(self.text.numel() // 2 + 1)
        if self.max_line_bytes is not None:
            bound = min(bound, self.n_lines * (int(self.max_line_bytes) // 2) + 1)
        return bound

    @property
    def n_lines(self):
        return self.begins.numel()

    def line(self, i):
        b0, e0 = int(self.begins[i]), int(self.ends[i]) - 1
        raw = bytes(self.text[b0:e0].cpu().numpy().tobytes())
        return raw.split(b"\0")[0].decode("utf-8", "replace")

    def select(self, index: Tensor):
        """the lines `index` (any order, e.g. a slice of a random permutation); the text is not copied"""
        if self.max_line_bytes is None and self.n_lines:     # once per parent (one device reduction + read-back)
            self.max_line_bytes = int((self.ends - self.begins).max())
        return PackedLines(self.text, begins=self.begins[index].contiguous(), ends=self.ends[index].contiguous(),
                           max_line_bytes=self.max_line_bytes)

    def to(self, device, non_blocking=True):
        text = self.text.to(device, non_blocking=non_blocking)
        if self.offsets is not None:
            return PackedLines(text, self.offsets.to(device, non_blocking=non_blocking),
                               max_line_bytes=self.max_line_bytes)
        return PackedLines(text, begins=self.begins.to(device, non_blocking=non_blocking),
                           ends=self.ends.to(device, non_blocking=non_blocking), max_line_bytes=self.max_line_bytes)


def pack_lines(lines, pin_memory=False) -> PackedLines:
    """list of str / bytes -> host PackedLines (text padded to a multiple of 16 bytes)"""
    bl = [s.encode() if isinstance(s, str) else bytes(s) for s in lines]
    off = np.zeros(len(bl) + 1, np.int64)
    if bl:
        np.cumsum(np.fromiter((len(b) + 1 for b in bl), np.int64, len(bl)), out=off[1:])
    nbytes = int(off[-1])
    buf = np.zeros((nbytes + 15) // 16 * 16 + 16, np.uint8)
    if bl:
        buf[:nbytes] = np.frombuffer(b"\n".join(bl) + b"\n", np.uint8)
    text, offsets = torch.from_numpy(buf), torch.from_numpy(off)
    if pin_memory:
        text, offsets = text.pin_memory(), offsets.pin_memory()
    return PackedLines(text, offsets)


def pack_text(data: bytes, pin_memory=False):
    """raw '\\n'-separated file bytes -> padded host u8 tensor for fm_split_lines
    (a final line without '\\n' gets one)"""
    if data and not data.endswith(b"\n"):
        data = data + b"\n"
    buf = np.zeros((len(data) + 15) // 16 * 16 + 16, np.uint8)
    buf[:len(data)] = np.frombuffer(data, np.uint8)
    t = torch.from_numpy(buf)
    return (t.pin_memory() if pin_memory else t), len(data)


def _status_tensor(device):
    return torch.zeros(4, dtype=torch.int64, device=device)


def _raise_on_parse_error(status_host, packed: PackedLines):
    first = int(status_host[0]) & 0xFFFFFFFFFFFFFFFF
    if first != _lib.NO_ERROR:
        kind, line_no = first & 7, first >> 3
        line = packed.line(line_no) if line_no < packed.n_lines else ""
        raise FmParseError(kind, line_no, line)


def fm_split_lines(text: Tensor, text_bytes: int, max_lines: int) -> Tensor:
    """device newline scan: padded u8 text -> line offsets i64[n_lines+1]"""
    require_cuda(text)
    lib = load()
    offsets = torch.empty(max_lines + 1, dtype=torch.int64, device=text.device)
    status = _status_tensor(text.device)
    ws_bytes = lib.fm_split_lines_workspace(text_bytes)
    ws = workspace(ws_bytes, text.device)
    check(lib.fm_split_lines(ptr(text), text_bytes, ptr(offsets), max_lines, ptr(status), ptr(ws), ws.numel(),
                             stream()), "fm_split_lines")
    st = status.cpu()
    first = int(st[0]) & 0xFFFFFFFFFFFFFFFF
    if first != _lib.NO_ERROR:
        if first & 7 == 5:
            raise ValueError("fm_split_lines: more than %d lines in the buffer" % max_lines)
        raise ValueError("fm_split_lines: the buffer does not end with a newline")
    return offsets[:int(st[2]) + 1]


# --------------------------------------------------------------------------
# custom ops
# --------------------------------------------------------------------------
@torch.library.custom_op("fast_tffm::fm_parser_packed", mutates_args=(), device_types="cuda")
def _fm_parser_packed(text: Tensor, begins: Tensor, ends: Tensor, vocab_size: int, hash_feature_id: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    lib = load()
    dev = text.device
    n_lines = begins.numel()
    text_bytes = text.numel() // 16 * 16
    labels = torch.empty(n_lines, dtype=torch.float32, device=dev)
    sizes = torch.empty(n_lines, dtype=torch.int32, device=dev)
    poses = torch.empty(n_lines + 1, dtype=torch.int32, device=dev)
    status = _status_tensor(dev)
    ws = workspace(lib.fm_parse_workspace(n_lines), dev)
    check(lib.fm_parse_count(ptr(text), text_bytes, ptr(begins), ptr(ends), -1, n_lines, vocab_size,
                             int(hash_feature_id), ptr(sizes), ptr(poses), ptr(status), ptr(ws), ws.numel(), stream()),
          "fm_parse_count")
    nnz = int(status[1].item())  # output shape is data dependent: one sync, as in any such op
    ids = torch.empty(nnz, dtype=torch.int64, device=dev)
    vals = torch.empty(nnz, dtype=torch.float32, device=dev)
    check(lib.fm_parse_fill(ptr(text), text_bytes, ptr(begins), ptr(ends), -1, n_lines, vocab_size,
                            int(hash_feature_id), ptr(poses), ptr(labels), ptr(ids), ptr(vals), nnz, ptr(status), ptr(ws),
                            ws.numel(), 0, stream()), "fm_parse_fill")
    return labels, sizes, ids, vals, status


def fm_parser(data_strings, vocab_size, hash_feature_id=False, device=None):
    """FmParser (cc/fm_parser_op.cc:8-15).  data_strings: a list of str/bytes (the TF
    string tensor), or a PackedLines already on the host or device."""
    if isinstance(data_strings, PackedLines):
        packed = data_strings
    else:
        if isinstance(data_strings, (str, bytes)):
            data_strings = [data_strings]
        packed = pack_lines(list(data_strings))
    if not packed.text.is_cuda:
        if not torch.cuda.is_available():
            raise _lib.FmB200Error("fm_parser needs a CUDA device; there is no CPU fallback")
        packed = packed.to(device or torch.device("cuda", torch.cuda.current_device()), non_blocking=False)
    labels, sizes, ids, vals, status = torch.ops.fast_tffm.fm_parser_packed(
        packed.text, packed.begins, packed.ends, int(vocab_size), bool(hash_feature_id))
    _raise_on_parse_error(status.cpu(), packed)
    return labels, sizes, ids, vals


def _as_float(x):
    return float(x.item()) if isinstance(x, Tensor) else float(x)


@torch.library.custom_op("fast_tffm::fm_scorer", mutates_args=(), device_types="cuda")
def _fm_scorer(feature_ids: Tensor, feature_params: Tensor, feature_vals: Tensor, feature_poses: Tensor,
               factor_lambda: float, bias_lambda: float) -> Tuple[Tensor, Tensor]:
    lib = load()
    params = feature_params.contiguous()
    batch = feature_poses.numel() - 1
    k = params.shape[1] - 1                                        # cc/fm_scorer_op.cc:50-51
    pred = torch.empty(batch, dtype=torch.float32, device=params.device)
    reg = torch.empty((), dtype=torch.float32, device=params.device)
    ws = workspace(lib.fm_score_workspace(batch), params.device)
    check(lib.fm_score(ptr(feature_ids), ptr(params), params.stride(0), ptr(feature_vals), ptr(feature_poses),
                       batch, k, factor_lambda, bias_lambda, ptr(pred), ptr(reg), ptr(ws), ws.numel(), stream()),
          "fm_score")
    return pred, reg


@_fm_scorer.register_fake
def _(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda):
    return (feature_params.new_empty(feature_poses.numel() - 1), feature_params.new_empty(()))


@torch.library.custom_op("fast_tffm::fm_grad", mutates_args=(), device_types="cuda")
def _fm_grad(feature_ids: Tensor, feature_params: Tensor, feature_vals: Tensor, feature_poses: Tensor,
             factor_lambda: float, bias_lambda: float, pred_grad: Tensor, reg_grad: Tensor) -> Tensor:
    lib = load()
    params = feature_params.contiguous()
    batch = feature_poses.numel() - 1
    k = params.shape[1] - 1
    out = torch.empty_like(params)                                 # cc/fm_grad_op.cc:67
    check(lib.fm_grad(ptr(feature_ids), ptr(params), params.stride(0), ptr(feature_vals), ptr(feature_poses),
                      batch, k, factor_lambda, bias_lambda, ptr(pred_grad), 0.0, ptr(reg_grad), params.shape[0],
                      None, ptr(out), stream()), "fm_grad")
    return out


@_fm_grad.register_fake
def _(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda, pred_grad, reg_grad):
    return torch.empty_like(feature_params)


def _scorer_setup(ctx, inputs, output):
    ids, params, vals, poses, lf, lb = inputs
    ctx.save_for_backward(ids, params, vals, poses)
    ctx.lambdas = (lf, lb)


def _scorer_backward(ctx, pred_grad, reg_grad):
    # tffm/fm_ops.py:11-27: only feature_params gets a gradient
    ids, params, vals, poses = ctx.saved_tensors
    lf, lb = ctx.lambdas
    if pred_grad is None:
        pred_grad = params.new_zeros(poses.numel() - 1)
    if reg_grad is None:
        reg_grad = params.new_zeros(())
    g = torch.ops.fast_tffm.fm_grad(ids, params, vals, poses, lf, lb, pred_grad.contiguous().float(),
                                    reg_grad.contiguous().float())
    return None, g, None, None, None, None


_fm_scorer.register_autograd(_scorer_backward, setup_context=_scorer_setup)


def _coerce(t, dtype, device):
    if not isinstance(t, Tensor):
        t = torch.as_tensor(t, dtype=dtype, device=device)
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def fm_scorer(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda):
    """FmScorer (cc/fm_scorer_op.cc:9-17), with the dtype coercion of tffm/fm_ops.py:39-56."""
    if not isinstance(feature_params, Tensor):
        feature_params = torch.as_tensor(feature_params, dtype=torch.float32)
    require_cuda(feature_params)
    dev = feature_params.device
    return torch.ops.fast_tffm.fm_scorer(
        _coerce(feature_ids, torch.int32, dev), feature_params.float(), _coerce(feature_vals, torch.float32, dev),
        _coerce(feature_poses, torch.int32, dev), _as_float(factor_lambda), _as_float(bias_lambda))


def fm_grad(feature_ids, feature_params, feature_vals, feature_poses, factor_lambda, bias_lambda, pred_grad,
            reg_grad):
    """FmGrad (cc/fm_grad_op.cc:15-24)"""
    require_cuda(feature_params)
    dev = feature_params.device
    return torch.ops.fast_tffm.fm_grad(
        _coerce(feature_ids, torch.int32, dev), feature_params.float(), _coerce(feature_vals, torch.float32, dev),
        _coerce(feature_poses, torch.int32, dev), _as_float(factor_lambda), _as_float(bias_lambda),
        _coerce(pred_grad, torch.float32, dev), _coerce(reg_grad, torch.float32, dev))


# --------------------------------------------------------------------------
# the stock-TF ops between the custom ops
# --------------------------------------------------------------------------
def fm_unique(ids: Tensor):
    """tf.unique (tffm/fm_model.py:72) -> ori_ids i64[U] (first-occurrence order), local i32[N]"""
    require_cuda(ids)
    lib = load()
    ids = ids.contiguous()
    n = ids.numel()
    ori = torch.empty(max(n, 1), dtype=torch.int64, device=ids.device)
    local = torch.empty(n, dtype=torch.int32, device=ids.device)
    n_unique = torch.zeros(1, dtype=torch.int64, device=ids.device)
    ws = workspace(lib.fm_unique_workspace(n), ids.device)
    check(lib.fm_unique(ptr(ids), n, None, ptr(ori), ptr(local), ptr(n_unique), ptr(ws), ws.numel(), stream()),
          "fm_unique")
    return ori[:int(n_unique.item())], local


def fm_poses(sizes: Tensor):
    """feature_poses = concat([0], cumsum(sizes)) (tffm/fm_model.py:73).  The parser already
    produces it on the device; this is for callers that only hold `sizes`."""
    z = torch.zeros(1, dtype=torch.int32, device=sizes.device)
    return torch.cat([z, torch.cumsum(sizes, 0, dtype=torch.int32)])


def fm_lookup(table: Tensor, ori_ids: Tensor, factor_num: int):
    """tf.nn.embedding_lookup (tffm/fm_model.py:291) -> f32[U, K+1]"""
    require_cuda(table, ori_ids)
    lib = load()
    u = ori_ids.numel()
    out = torch.empty(u, factor_num + 1, dtype=torch.float32, device=table.device)
    check(lib.fm_lookup(ptr(table), table.stride(0), ptr(ori_ids), u, None, factor_num, ptr(out), out.stride(0),
                        None, stream()), "fm_lookup")
    return out


def fm_loss_grad(pred: Tensor, labels: Tensor, weights, loss_type: str):
    """weighted mean loss and d loss / d pred (tffm/fm_model.py:311-326)"""
    require_cuda(pred, labels, weights)
    lib = load()
    b = pred.numel()
    loss = torch.empty((), dtype=torch.float32, device=pred.device)
    pg = torch.empty_like(pred)
    ws = workspace(lib.fm_loss_workspace(b), pred.device)
    check(lib.fm_loss_grad(ptr(pred), ptr(labels), ptr(weights), b, _lib.LOSS_KIND[loss_type], ptr(loss), ptr(pg),
                           ptr(ws), ws.numel(), stream()), "fm_loss_grad")
    return loss, pg


def fm_sparse_adagrad(table: Tensor, accum: Tensor, ori_ids: Tensor, grad: Tensor, factor_num: int, lr: float):
    """TF1 SparseApplyAdagrad, in place on table and accum (tffm/fm_model.py:341-348)"""
    require_cuda(table, accum, ori_ids, grad)
    lib = load()
    assert table.stride(0) == accum.stride(0)
    check(lib.fm_sparse_adagrad(ptr(table), ptr(accum), table.stride(0), ptr(ori_ids), ori_ids.numel(), None,
                                factor_num, ptr(grad), grad.stride(0), lr, None, stream()), "fm_sparse_adagrad")
