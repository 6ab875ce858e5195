"""Vocabulary table sharded across the GPUs of one box; one process per GPU.

Replaces the reference's parameter-server sharding (tffm/fm_model.py:269-291:
`vocabulary_block_num` variables placed round-robin on `ps` tasks by
`replica_device_setter`, run_tffm.py:181-206; rows fetched and sparse gradients pushed over
gRPC, asynchronously).  Here the split is explicit and synchronous:

    rank  id % G  owns table row  id // G        (TF's "mod" partition with nb = G)
    rank r parses examples [r*B/G, (r+1)*B/G) of the global batch

and one step is
    parse + unique (local)                     -> distinct ids, grouped by owner (fm_shard_bucket)
    all_to_all counts, all_to_all ids          -> every owner learns which rows are wanted
    owners gather rows (fm_lookup)             -> all_to_all rows back
    fm_fused_score_scatter on the received rows (loss over the GLOBAL batch size)
    all_to_all accumulator rows to the owners  -> owners add them into the accumulator row
                                                  that sits beside each table row
                                                  (fm_shard_combine), then
                                                  fm_fused_adagrad_claim walks the received id
                                                  list: the first entry of an id applies the
                                                  update, its repeats from other ranks leave
    all_reduce(loss)

so a G-GPU step equals the single-GPU step on the same global batch (sums are combined
before Adagrad squares them).  The exchange is torch.distributed over NCCL/NVLink; the
kernels on either side are ours (libfm_b200.so).  `kernels=` exists so that the host-side
orchestration can be exercised on CPU tensors with gloo in tests; the default is the CUDA
library and there is no CPU implementation in this package.
"""
import ctypes

import torch
import torch.distributed as dist

from . import _lib
from ._lib import check, load, ptr, stream
from .fm_model import Model, _ld_for, _stride_for
from .fm_ops import FmParseError, PackedLines, pack_lines


class CudaShardKernels:
    """the C ABI of libfm_b200.so behind the small set of operations ShardedModel needs"""

    def __init__(self):
        self.lib = load()

    # -- local side ------------------------------------------------------------------
    def parse_unique(self, packed, vocab, hashed):
        lib, dev, n_lines = self.lib, packed.text.device, packed.n_lines
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        labels = torch.empty(n_lines, dtype=f32, device=dev)
        sizes = torch.empty(n_lines, dtype=i32, device=dev)
        poses = torch.empty(n_lines + 1, dtype=i32, device=dev)
        status = torch.zeros(8, dtype=i64, device=dev)
        ws = _lib.workspace(lib.fm_parse_workspace(n_lines), dev)
        check(lib.fm_parse_count(ptr(packed.text), ptr(packed.offsets), n_lines, vocab, int(hashed), ptr(sizes),
                                 ptr(poses), ptr(status), ptr(ws), ws.numel(), stream()), "fm_parse_count")
        nnz = int(status[1].item())
        ids = torch.empty(nnz, dtype=i64, device=dev)
        vals = torch.empty(nnz, dtype=f32, device=dev)
        check(lib.fm_parse_fill(ptr(packed.text), ptr(packed.offsets), n_lines, vocab, int(hashed), ptr(poses),
                                ptr(labels), ptr(ids), ptr(vals), nnz, ptr(status), ptr(ws), ws.numel(), stream()),
              "fm_parse_fill")
        ori, local = self.unique(ids)
        first = int(status[0].item()) & 0xFFFFFFFFFFFFFFFF
        if first != _lib.NO_ERROR:
            raise FmParseError(first & 7, first >> 3, packed.line(first >> 3))
        return dict(labels=labels, poses=poses, ids=ids, vals=vals, ori=ori, local=local)

    def unique(self, ids):
        lib, n = self.lib, ids.numel()
        ori = torch.empty(max(n, 1), dtype=torch.int64, device=ids.device)
        local = torch.empty(n, dtype=torch.int32, device=ids.device)
        n_unique = torch.zeros(1, dtype=torch.int64, device=ids.device)
        ws = _lib.workspace(lib.fm_unique_workspace(n), ids.device)
        check(lib.fm_unique(ptr(ids), n, None, ptr(ori), ptr(local), ptr(n_unique), ptr(ws), ws.numel(), stream()),
              "fm_unique")
        return ori[:int(n_unique.item())], local

    def bucket(self, ori, n_ranks):
        lib, n = self.lib, ori.numel()
        send_ids = torch.empty(n, dtype=torch.int64, device=ori.device)
        pos = torch.empty(n, dtype=torch.int32, device=ori.device)
        counts = torch.zeros(n_ranks, dtype=torch.int64, device=ori.device)
        ws = _lib.workspace(lib.fm_shard_bucket_workspace(n, n_ranks), ori.device)
        check(lib.fm_shard_bucket(ptr(ori), n, n_ranks, ptr(send_ids), ptr(pos), ptr(counts), ptr(ws), ws.numel(),
                                  stream()), "fm_shard_bucket")
        return send_ids, pos, counts

    def remap(self, local, pos):
        n = local.numel()
        slot64 = torch.empty(n, dtype=torch.int64, device=local.device)
        slot32 = torch.empty(n, dtype=torch.int32, device=local.device)
        check(self.lib.fm_shard_remap(ptr(local), n, None, ptr(pos), ptr(slot64), ptr(slot32), stream()),
              "fm_shard_remap")
        return slot64, slot32

    def score_scatter(self, slot64, slot32, vals, poses, labels, weights, k, rows, lf, lb, loss_type, global_batch,
                      gacc):
        lib, b, dev = self.lib, labels.numel(), labels.device
        pred = torch.empty(b, dtype=torch.float32, device=dev)
        scal = torch.zeros(2, dtype=torch.float32, device=dev)
        ws = _lib.workspace(lib.fm_fused_workspace(b), dev)
        check(lib.fm_fused_score_scatter(
            ptr(slot64), ptr(slot32), ptr(vals), ptr(poses), ptr(labels), ptr(weights), b, k, ptr(rows),
            rows.stride(0), lf, lb, _lib.LOSS_KIND[loss_type], global_batch, ptr(pred), ptr(scal),
            ctypes.c_void_p(scal.data_ptr() + 4), ptr(gacc), gacc.stride(0), None, ptr(ws), ws.numel(), stream()),
            "fm_fused_score_scatter")
        return pred, scal

    def score(self, slot32, rows, vals, poses, k):
        lib, b, dev = self.lib, poses.numel() - 1, rows.device
        pred = torch.empty(b, dtype=torch.float32, device=dev)
        reg = torch.empty((), dtype=torch.float32, device=dev)
        ws = _lib.workspace(lib.fm_score_workspace(b), dev)
        check(lib.fm_score(ptr(slot32), ptr(rows), rows.stride(0), ptr(vals), ptr(poses), b, k, 0.0, 0.0, ptr(pred),
                           ptr(reg), ptr(ws), ws.numel(), stream()), "fm_score")
        return pred

    # -- owner side ------------------------------------------------------------------
    def local_rows(self, ids, n_ranks):
        rows = torch.empty_like(ids)
        check(self.lib.fm_shard_local_rows(ptr(ids), ids.numel(), n_ranks, ptr(rows), stream()), "fm_shard_local_rows")
        return rows

    def lookup(self, table, rows, k):
        """rows as they travel: K+1 floats rounded up to whole 16-byte chunks"""
        out = torch.empty(rows.numel(), 4 * ((k + 4) // 4), dtype=torch.float32, device=table.device)
        check(self.lib.fm_lookup(ptr(table), table.stride(0), ptr(rows), rows.numel(), None, k, ptr(out),
                                 out.stride(0), None, stream()), "fm_lookup")
        return out

    def combine(self, src, index, dst):
        check(self.lib.fm_shard_combine(ptr(src), src.stride(0), ptr(index), src.shape[0], src.shape[1], ptr(dst),
                                        dst.stride(0), stream()), "fm_shard_combine")

    def adagrad_claim(self, table, accum, rows, k, gtab, lr, lf, lb, reg_grad):
        check(self.lib.fm_fused_adagrad_claim(ptr(table), ptr(accum), table.stride(0), ptr(rows), rows.numel(), None,
                                              k, ptr(gtab), gtab.stride(0), lr, lf, lb, reg_grad, None, stream()),
              "fm_fused_adagrad_claim")

    def gacc_ld(self, k):
        return int(self.lib.fm_fused_gacc_ld(k))


def _exchange(send, send_counts, recv_counts, group):
    """all_to_all of row blocks: send[sum(send_counts), ...] -> recv[sum(recv_counts), ...]"""
    out = send.new_empty((sum(recv_counts),) + tuple(send.shape[1:]))
    dist.all_to_all_single(out, send.contiguous(), output_split_sizes=list(recv_counts),
                           input_split_sizes=list(send_counts), group=group)
    return out


class ShardedModel:
    """Model (same config object) with the table row-partitioned over the ranks of `group`."""

    def __init__(self, model: Model, group=None, kernels=None, device=None):
        self.cfg = model
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.kernels = kernels if kernels is not None else CudaShardKernels()
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.global_step = 0

    def rows_per_rank(self):
        return self.cfg.vocabulary_size // self.world + 1

    def build(self, init_table=None, seed=None):
        """shard[r][row] = full_table[row * G + r].  init_table: the FULL [V, K+1] table (parity runs)."""
        k, g, r = self.cfg.factor_num, self.world, self.rank
        self.ld = _ld_for(k)
        rows = self.rows_per_rank()
        self.table = torch.zeros(rows, self.ld, dtype=torch.float32, device=self.device)
        if init_table is not None:
            part = torch.as_tensor(init_table, dtype=torch.float32)[r::g, :k + 1]
            self.table[:part.shape[0], :k + 1] = part.to(self.device)
        else:
            gen = torch.Generator(device=self.device)
            gen.manual_seed((seed or 0) * 1000 + r)
            self.table[:, :k + 1].uniform_(-self.cfg.init_value_range, self.cfg.init_value_range, generator=gen)
        self.accum = torch.full_like(self.table, self.cfg.adagrad_init_accumulator)
        # gradient accumulator rows, one beside each table row of this shard (all zero between steps)
        self.gtab = torch.zeros(rows, _stride_for(self.kernels.gacc_ld(k)), dtype=torch.float32, device=self.device)
        return self

    def gather_full_table(self):
        """[V, K+1] on every rank (tests / checkpoints)"""
        k, g = self.cfg.factor_num, self.world
        parts = [torch.empty_like(self.table) for _ in range(g)]
        dist.all_gather(parts, self.table, group=self.group)
        full = torch.zeros(self.rows_per_rank() * g, k + 1, dtype=torch.float32, device=self.device)
        for r in range(g):
            full[r::g] = parts[r][:, :k + 1]
        return full[:self.cfg.vocabulary_size]

    # -- one step ----------------------------------------------------------------------
    def _route(self, lines, mark=lambda name: None):
        """parse + unique locally, exchange ids, fetch rows.  Returns the step's working set."""
        kn, g, k = self.kernels, self.world, self.cfg.factor_num
        packed = lines if isinstance(lines, PackedLines) else pack_lines(list(lines))
        if packed.text.device != self.device:
            packed = packed.to(self.device, non_blocking=False)
        p = kn.parse_unique(packed, self.cfg.vocabulary_size, self.cfg.hash_feature_id)
        mark("parse_unique")
        send_ids, pos, counts = kn.bucket(p["ori"], g)
        recv_counts_t = torch.empty_like(counts)
        dist.all_to_all_single(recv_counts_t, counts, group=self.group)
        send_counts, recv_counts = counts.tolist(), recv_counts_t.tolist()
        mark("bucket_counts")
        want_ids = _exchange(send_ids, send_counts, recv_counts, self.group)      # ids other ranks want from me
        mark("xchg_ids")
        own_rows = kn.local_rows(want_ids, g)                                     # my rows the others asked for
        rows_out = kn.lookup(self.table, own_rows, k)
        mark("lookup")
        rows_in = _exchange(rows_out, recv_counts, send_counts, self.group)       # my distinct ids' rows, bucket order
        mark("xchg_rows")
        slot64, slot32 = kn.remap(p["local"], pos)
        p.update(send_counts=send_counts, recv_counts=recv_counts, own_rows=own_rows, rows_in=rows_in,
                 slot64=slot64, slot32=slot32, n_lines=packed.n_lines)
        return p

    def train_step(self, lines, weights=None, global_batch=None, mark=lambda name: None):
        """lines / weights: this rank's slice of the global batch.  Returns the GLOBAL mean loss."""
        kn, g, k, cfg = self.kernels, self.world, self.cfg.factor_num, self.cfg
        p = self._route(lines, mark)
        if global_batch is None:
            nb = torch.tensor([p["n_lines"]], dtype=torch.int64, device=self.device)
            dist.all_reduce(nb, group=self.group)
            global_batch = int(nb.item())
        if weights is not None and not isinstance(weights, torch.Tensor):
            weights = torch.as_tensor(weights, dtype=torch.float32)
        if weights is not None:
            weights = weights.to(self.device)
        gld = kn.gacc_ld(k)
        gacc = torch.zeros(p["ori"].numel(), gld, dtype=torch.float32, device=self.device)
        pred, scal = kn.score_scatter(p["slot64"], p["slot32"], p["vals"], p["poses"], p["labels"], weights, k,
                                      p["rows_in"], cfg.factor_lambda, cfg.bias_lambda, cfg.loss_type, global_batch,
                                      gacc)
        mark("score_scatter")
        gacc_in = _exchange(gacc, p["send_counts"], p["recv_counts"], self.group)  # rows for ids I own
        mark("xchg_gacc")
        kn.combine(gacc_in, p["own_rows"], self.gtab)            # sums over ranks land beside the table rows
        kn.adagrad_claim(self.table, self.accum, p["own_rows"], k, self.gtab, cfg.learning_rate,
                         cfg.factor_lambda, cfg.bias_lambda, 1.0 / cfg.batch_size)
        mark("update")
        dist.all_reduce(scal, group=self.group)
        self.global_step += 1
        self.last = dict(pred=pred, reg=scal[1], n_unique=p["ori"].numel(), nnz=p["ids"].numel(),
                         sent_rows=sum(p["send_counts"]) - p["send_counts"][self.rank])
        return float(scal[0].item())

    def predict(self, lines):
        kn, k = self.kernels, self.cfg.factor_num
        p = self._route(lines)
        return kn.score(p["slot32"], p["rows_in"], p["vals"], p["poses"], k)
