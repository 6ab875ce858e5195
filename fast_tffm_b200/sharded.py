"""Vocabulary table sharded across the GPUs of one box; one process per GPU.

Replaces the reference's parameter-server sharding (tffm/fm_model.py:269-291:
`vocabulary_block_num` variables placed round-robin on `ps` tasks by
`replica_device_setter`, run_tffm.py:181-206; rows fetched and sparse gradients pushed over
gRPC, asynchronously).  Here the split is explicit and synchronous:

    rank  id % G  owns table row  id // G        (TF's "mod" partition with nb = G)
    rank r parses examples [r*B/G, (r+1)*B/G) of the global batch

and one step is
    parse + unique (local) + group the distinct ids by owner (fm_shard_bucket)
    all_gather of a small info vector (sizes, per-owner counts, parse status): the step's
                                               ONE host synchronisation
    all_to_all ids                             -> every owner learns which rows are wanted
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
import os

import torch
import torch.distributed as dist

from . import _lib
from ._lib import check, load, ptr, stream
from .fm_model import CheckpointWriter, Model, latest_checkpoint, load_checkpoint, _ld_for, _stride_for
from .fm_ops import FmParseError, PackedLines, pack_lines

INFO_NNZ, INFO_UNIQUE, INFO_ERROR, INFO_COUNTS = 0, 1, 2, 3   # layout of the per-rank info vector


def _at(t, index):
    return ctypes.c_void_p(t.data_ptr() + index * t.element_size())


class CudaShardKernels:
    """the C ABI of libfm_b200.so behind the small set of operations ShardedModel needs.
    Buffers are sized by capacity and reused; data-dependent sizes stay on the device."""

    def __init__(self, max_nnz_per_line=None):
        self.lib = load()
        self.max_nnz_per_line = max_nnz_per_line
        self._b = None

    def _buffers(self, n_lines, nnz_bound, n_ranks, dev):
        cap = int(nnz_bound)
        if self.max_nnz_per_line:
            cap = min(cap, int(n_lines * self.max_nnz_per_line))
        b = self._b
        if b is not None and b["lines"] >= n_lines and b["cap"] >= cap and b["ranks"] == n_ranks:
            return b
        lib = self.lib
        lines, cap = max(n_lines, b["lines"] if b else 0), max(cap, b["cap"] if b else 0)
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        e = lambda n, dt: torch.empty(n, dtype=dt, device=dev)
        self._b = b = dict(
            lines=lines, cap=cap, ranks=n_ranks, labels=e(lines, f32), sizes=e(lines, i32), poses=e(lines + 1, i32),
            ids=e(cap, i64), vals=e(cap, f32), ori=e(cap, i64), local=e(cap, i32), send_ids=e(cap, i64),
            pos=e(cap, i32), slot64=e(cap, i64), slot32=e(cap, i32),
            status=torch.zeros(8, dtype=i64, device=dev),                 # fm_parse_status + n_unique at [4]
            info=torch.zeros(INFO_COUNTS + n_ranks, dtype=i64, device=dev),
            ws_parse=_lib.workspace(lib.fm_parse_workspace(lines), dev),
            ws_unique=_lib.workspace(lib.fm_unique_workspace(cap), dev),
            ws_bucket=_lib.workspace(lib.fm_shard_bucket_workspace(cap, n_ranks), dev),
            ws_fused=_lib.workspace(lib.fm_fused_workspace(lines), dev),
            pred=e(lines, f32), scal=torch.zeros(2, dtype=f32, device=dev))
        return b

    # -- local side ------------------------------------------------------------------
    def parse_unique_bucket(self, packed, vocab, hashed, n_ranks):
        """parse -> unique -> group distinct ids by owner -> per-occurrence slots; enqueues only.
        Returns capacity-sized tensors and `info` = [nnz, n_unique, parse status, counts[G]] (device)."""
        lib, dev, n_lines, st = self.lib, packed.text.device, packed.n_lines, stream()
        b = self._buffers(n_lines, packed.nnz_bound(), n_ranks, dev)
        status, nnz_dev, nuniq_dev = ptr(b["status"]), _at(b["status"], 1), _at(b["status"], 4)
        check(lib.fm_parse_count(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                 vocab, int(hashed), ptr(b["sizes"]), ptr(b["poses"]), status, ptr(b["ws_parse"]),
                                 b["ws_parse"].numel(), st), "fm_parse_count")
        check(lib.fm_parse_fill(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                vocab, int(hashed), ptr(b["poses"]),
                                ptr(b["labels"]), ptr(b["ids"]), ptr(b["vals"]), b["cap"], status, ptr(b["ws_parse"]),
                                b["ws_parse"].numel(), 0, st), "fm_parse_fill")
        check(lib.fm_unique(ptr(b["ids"]), b["cap"], nnz_dev, ptr(b["ori"]), ptr(b["local"]), nuniq_dev,
                            ptr(b["ws_unique"]), b["ws_unique"].numel(), st), "fm_unique")
        check(lib.fm_shard_bucket(ptr(b["ori"]), b["cap"], nuniq_dev, n_ranks, ptr(b["send_ids"]), ptr(b["pos"]),
                                  _at(b["info"], INFO_COUNTS), ptr(b["ws_bucket"]), b["ws_bucket"].numel(), st),
              "fm_shard_bucket")
        check(lib.fm_shard_remap(ptr(b["local"]), b["cap"], nnz_dev, ptr(b["pos"]), ptr(b["slot64"]), ptr(b["slot32"]),
                                 st), "fm_shard_remap")
        b["info"][INFO_NNZ:INFO_COUNTS] = b["status"][[1, 4, 0]]      # nnz, n_unique, first_error
        return dict(labels=b["labels"][:n_lines], poses=b["poses"][:n_lines + 1], vals=b["vals"],
                    send_ids=b["send_ids"], slot64=b["slot64"], slot32=b["slot32"], info=b["info"])

    def score_scatter(self, slot64, slot32, vals, poses, labels, weights, k, rows, lf, lb, loss_type, global_batch,
                      gacc):
        lib, b, bufs = self.lib, labels.numel(), self._b
        pred, scal = bufs["pred"][:b], bufs["scal"]
        check(lib.fm_fused_score_scatter(
            ptr(slot64), ptr(slot32), ptr(vals), ptr(poses), ptr(labels), ptr(weights), b, k, ptr(rows),
            rows.stride(0), lf, lb, _lib.LOSS_KIND[loss_type], global_batch, ptr(pred), ptr(scal), _at(scal, 1),
            ptr(gacc), gacc.stride(0), None, ptr(bufs["ws_fused"]), bufs["ws_fused"].numel(), stream()),
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
                                              k, ptr(gtab), gtab.stride(0), lr, lf, lb, reg_grad, None, 0, stream()),
              "fm_fused_adagrad_claim")

    def gacc_ld(self, k):
        return int(self.lib.fm_fused_gacc_ld(k))


def _exchange(send, send_counts, recv_counts, group):
    """all_to_all of row blocks: send[sum(send_counts), ...] -> recv[sum(recv_counts), ...]"""
    out = send.new_empty((sum(recv_counts),) + tuple(send.shape[1:]))
    dist.all_to_all_single(out, send, output_split_sizes=list(recv_counts), input_split_sizes=list(send_counts),
                           group=group)
    return out


def check_sync_word(f):
    """fm_peer_sync gives up on a rank that does not arrive within tens of seconds (word >= 2^62): that is a dead or
    wedged process, not a parse error -- say so"""
    if f >= (1 << 62):
        raise RuntimeError("sharded step: rank %d never reached the barrier (fm_peer_sync timed out); the step was void"
                           % ((f >> 48) & 0xFF))


class ShardedModel:
    """Model (same config object) with the table row-partitioned over the ranks of `group`."""

    def __init__(self, model: Model, group=None, kernels=None, device=None, max_nnz_per_line=None):
        self.cfg = model
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.kernels = kernels if kernels is not None else CudaShardKernels(max_nnz_per_line)
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
            packed = packed.to(self.device, non_blocking=True)
        p = kn.parse_unique_bucket(packed, self.cfg.vocabulary_size, self.cfg.hash_feature_id, g)
        mark("parse_unique_bucket")
        # every rank learns every rank's sizes and parse status in one small collective + one D2H copy
        info_all = torch.empty(g * p["info"].numel(), dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(info_all, p["info"], group=self.group)
        host = info_all.cpu().view(g, -1)
        mark("sizes")
        for r in range(g):   # a parse error anywhere aborts the step on every rank, before any update
            first = int(host[r, INFO_ERROR]) & 0xFFFFFFFFFFFFFFFF
            if first != _lib.NO_ERROR:
                line = packed.line(first >> 3) if r == self.rank else "<rank %d, line %d>" % (r, first >> 3)
                raise FmParseError(first & 7, first >> 3, line)
        nnz, n_unique = int(host[self.rank, INFO_NNZ]), int(host[self.rank, INFO_UNIQUE])
        send_counts = host[self.rank, INFO_COUNTS:].tolist()
        recv_counts = host[:, INFO_COUNTS + self.rank].tolist()
        want_ids = _exchange(p["send_ids"][:n_unique], send_counts, recv_counts, self.group)  # ids asked of me
        mark("xchg_ids")
        own_rows = kn.local_rows(want_ids, g)
        rows_out = kn.lookup(self.table, own_rows, k)
        mark("lookup")
        rows_in = _exchange(rows_out, recv_counts, send_counts, self.group)   # my distinct ids' rows, owner order
        mark("xchg_rows")
        p.update(send_counts=send_counts, recv_counts=recv_counts, own_rows=own_rows, rows_in=rows_in, nnz=nnz,
                 n_unique=n_unique, n_lines=packed.n_lines, vals=p["vals"][:nnz], slot64=p["slot64"][:nnz],
                 slot32=p["slot32"][:nnz])
        return p

    def train_step_async(self, lines, weights=None, global_batch=None, mark=lambda name: None):
        """lines / weights: this rank's slice of the global batch.  Returns a device tensor
        [global mean loss, global reg_score]; nothing after the size exchange synchronises."""
        kn, k, cfg = self.kernels, self.cfg.factor_num, self.cfg
        p = self._route(lines, mark)
        if global_batch is None:
            nb = torch.tensor([p["n_lines"]], dtype=torch.int64, device=self.device)
            dist.all_reduce(nb, group=self.group)
            global_batch = int(nb.item())
        if weights is not None and not isinstance(weights, torch.Tensor):
            weights = torch.as_tensor(weights, dtype=torch.float32)
        if weights is not None:
            weights = weights.to(self.device)
        gacc = torch.zeros(p["n_unique"], kn.gacc_ld(k), dtype=torch.float32, device=self.device)
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
        out = scal.clone()
        dist.all_reduce(out, group=self.group)
        self.global_step += 1
        self.last = dict(pred=pred, n_unique=p["n_unique"], nnz=p["nnz"],
                         sent_rows=sum(p["send_counts"]) - p["send_counts"][self.rank])
        return out

    def train_step(self, lines, weights=None, global_batch=None, mark=lambda name: None):
        """Returns the GLOBAL mean loss as a python float."""
        return float(self.train_step_async(lines, weights, global_batch, mark)[0].item())

    def predict(self, lines):
        kn, k = self.kernels, self.cfg.factor_num
        p = self._route(lines)
        return kn.score(p["slot32"], p["rows_in"], p["vals"], p["poses"], k)


class PeerShardedModel:
    """The sharded table over NVLink peer memory: no id, row or gradient exchange.

    Every rank's table shard, gradient-accumulator shard and touched-row bitmap live in one
    symmetric allocation that is mapped into every process of the box (torch symmetric
    memory = CUDA VMM handles exchanged at start-up; plumbing).  A step is

        parse (local)
        fm_peer_score_scatter    rows read from whichever GPU owns them (LDG.128 over NVLink),
                                 gradient sums red.add-ed into the owner's accumulator rows,
                                 the rows' bits set in the owner's bitmap
        all_reduce (1 word)      barrier + "did any rank fail to parse"
        fm_peer_touched_rows     owner: bitmap -> list of touched rows
        fm_fused_adagrad_claim   owner: Adagrad on those rows (or, if any rank failed, the
                                 step is void everywhere: the sums are thrown away)
        all_reduce (loss, reg)   also fences the update from the next step's reads

    with no host synchronisation: train_step_async returns device tensors.  Results equal the
    single-GPU step on the same global batch (tests/test_gpu_sharded.py).  There is no CPU form
    of this path; the NCCL all-to-all form (ShardedModel) is the one that runs under gloo.
    """

    def __init__(self, model: Model, group=None, device=None, max_nnz_per_line=None, exchange="atomics"):
        """exchange: how the gradient sums reach the rows' owners.  "atomics": system-scope vector atomics from inside
        the scorer (fm_peer_score_scatter).  "records": the occurrences are sorted by (owner, row) ahead of the step,
        one record per run is written into the owner's inbox and the owners merge their G sorted inboxes
        (fm_peer_sorted_score / fm_peer_push_records / fm_peer_merge_update): 236 instead of 328 bytes on NVLink per
        remote occurrence."""
        if exchange not in ("atomics", "records"):
            raise ValueError("exchange must be atomics or records")
        self.exchange = exchange
        self._inbox = None
        self.cfg = model
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.max_nnz_per_line = max_nnz_per_line
        if getattr(model, "optimizer", "adagrad") != "adagrad":
            raise ValueError("the sharded step is Adagrad (as the reference); optimizer = %r runs on one GPU, sorted path"
                             % model.optimizer)
        self.lib = load()
        self.global_step = 0
        self._slots = {}
        self._ahead_slot = 2     # parse_ahead alternates between buffer sets 1 and 2; set 0 serves inline steps / predict
        self._parse_stream = None

    def rows_per_rank(self):
        return self.cfg.vocabulary_size // self.world + 1

    def build(self, init_table=None, seed=None):
        import torch.distributed._symmetric_memory as symm_mem
        k, g, r, dev = self.cfg.factor_num, self.world, self.rank, self.device
        rows = self.rows_per_rank()
        # per id one block [w | h | g] of whole 128-byte lines (K=16: 3 x 20 floats in 256 B), as in
        # Model(layout="interleaved"); peers read w (line 0) and add into g, h is only the owner's
        part = max((k + 1 + 3) // 4 * 4, int(self.lib.fm_fused_gacc_ld(k)))
        block = (3 * part + 31) // 32 * 32
        self.ld = self.gld = block
        n_store = rows * block
        n_bits = ((rows + 31) // 32 + 7) // 8 * 8           # keeps what follows 32-byte aligned
        n_sync = int(self.lib.fm_peer_sync_bytes()) // 4
        self._symm = symm_mem.empty(n_store + n_bits + n_sync, dtype=torch.float32, device=dev)
        self._hdl = symm_mem.rendezvous(self._symm, self.group)
        self._symm.zero_()
        store = self._symm[:n_store].view(rows, block)
        self.table, self.accum, self.gtab = store[:, :part], store[:, part:2 * part], store[:, 2 * part:3 * part]
        self.touched = self._symm[n_store:n_store + n_bits].view(torch.int32)
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        arr = ctypes.c_void_p * g
        self._table_bases = arr(*ptrs)
        self._gtab_bases = arr(*[p + 4 * 2 * part for p in ptrs])
        self._touched_bases = arr(*[p + 4 * n_store for p in ptrs])
        self._sync_bases = arr(*[p + 4 * (n_store + n_bits) for p in ptrs])
        self._epoch = 0
        self._sync_word = torch.zeros(1, dtype=torch.int64, device=dev)
        if init_table is not None:
            part = torch.as_tensor(init_table, dtype=torch.float32)[r::g, :k + 1]
            self.table[:part.shape[0], :k + 1] = part.to(dev)
        else:
            gen = torch.Generator(device=dev)
            gen.manual_seed((seed or 0) * 1000 + r)
            self.table[:, :k + 1].uniform_(-self.cfg.init_value_range, self.cfg.init_value_range, generator=gen)
        self.accum.fill_(self.cfg.adagrad_init_accumulator)
        self.row_list = torch.empty(rows, dtype=torch.int64, device=dev)
        self.n_rows = torch.zeros(1, dtype=torch.int64, device=dev)
        self.ws_touched = _lib.workspace(self.lib.fm_peer_touched_workspace(rows), dev)
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.group)          # every shard is initialised before anyone reads it
        return self

    def _gather_full(self, shard):
        k, g = self.cfg.factor_num, self.world
        mine = shard.contiguous()
        parts = [torch.empty_like(mine) for _ in range(g)]
        dist.all_gather(parts, mine, group=self.group)
        full = torch.zeros(self.rows_per_rank() * g, k + 1, dtype=torch.float32, device=self.device)
        for r in range(g):
            full[r::g] = parts[r][:, :k + 1]
        return full[:self.cfg.vocabulary_size]

    def gather_full_table(self):
        """[V, K+1] on every rank (tests / checkpoints)"""
        return self._gather_full(self.table)

    # Checkpoints: rank r's shard IS block r of the reference's variable layout with vocabulary_block_num = G
    # (vocab_block_r[id // G] for id % G == r), so every rank writes its own file `path.shard-r-of-G` --
    # no gather, written in the background like Model.save -- and any reader (a single-GPU Model, a run on a
    # different number of GPUs) merges the set with the same mod arithmetic (fm_model.load_checkpoint).
    def _writer(self):
        if getattr(self, "_ckpt_writer", None) is None:
            self._ckpt_writer = CheckpointWriter(self.device)
        return self._ckpt_writer

    def prepare_save(self):
        self._writer().prepare()

    def wait_for_save(self):
        self._writer().wait()
        dist.barrier(group=self.group)

    def save(self, path, wait=True, skip_if_busy=False):
        """collective (every rank at the same step).  skip_if_busy: dropped on all ranks if any rank's writer
        is still busy with the previous set (returns False)."""
        w, k, r, g = self._writer(), self.cfg.factor_num, self.rank, self.world
        if skip_if_busy:
            busy = torch.tensor([1 if w.busy() else 0], dtype=torch.int32, device=self.device)
            dist.all_reduce(busy, op=dist.ReduceOp.MAX, group=self.group)
            if int(busy.item()):
                return False
        w.wait()
        snap = torch.stack([self.table[:, :k + 1], self.accum[:, :k + 1]]).contiguous()
        step = self.global_step

        def state(host):
            return {"global_step": step, "vocab_block_%d" % r: host[0], "vocab_block_%d/Adagrad" % r: host[1]}

        w.write("%s.shard-%d-of-%d" % (path, r, g), snap, state, wait=wait)
        if wait:
            dist.barrier(group=self.group)
        return True

    def restore(self, path):
        """the newest checkpoint at `path`: this rank's own file if the set was written by as many GPUs,
        else (single-GPU file, other GPU count) the merged table, of which the rank keeps its rows"""
        v, k, g, r = self.cfg.vocabulary_size, self.cfg.factor_num, self.world, self.rank
        best = latest_checkpoint(path)
        if best is None:
            raise IOError("no checkpoint at %s" % path)
        if len(best[1]) == g and best[1][0] != path:      # a set written by as many GPUs
            state = torch.load(best[1][r], map_location="cpu")
            for name, shard in (("", self.table), ("/Adagrad", self.accum)):
                blk = state["vocab_block_%d%s" % (r, name)]
                shard[:blk.shape[0], :k + 1] = blk.to(self.device)
        else:
            state = load_checkpoint(path)
            nb = sum(1 for key in state if key.startswith("vocab_block_") and not key.endswith("/Adagrad"))
            for name, shard in (("", self.table), ("/Adagrad", self.accum)):
                full = torch.zeros(v, k + 1, dtype=torch.float32)
                for i in range(nb):
                    n = full[i::nb].shape[0]
                    full[i::nb] = state["vocab_block_%d%s" % (i, name)][:n]
                mine = full[r::g]
                shard[:mine.shape[0], :k + 1] = mine.to(self.device)
        self.global_step = int(best[0])
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def _buffers(self, n_lines, nnz_bound, slot=0):
        cap = int(nnz_bound)
        if self.max_nnz_per_line:
            cap = min(cap, int(n_lines * self.max_nnz_per_line))
        b = self._slots.get(slot)
        if b is not None and b["lines"] >= n_lines and b["cap"] >= cap:
            return b
        lib, dev = self.lib, self.device
        if b is not None:   # growing: the side stream may still be using the old set (see Model._buffers_one)
            torch.cuda.synchronize(dev)
        lines, cap = max(n_lines, b["lines"] if b else 0), max(cap, b["cap"] if b else 0)
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        e = lambda n, dt: torch.empty(n, dtype=dt, device=dev)
        self._slots[slot] = b = dict(
            lines=lines, cap=cap, labels=e(lines, f32), sizes=e(lines, i32), poses=e(lines + 1, i32),
            ids=e(cap, i64), vals=e(cap, f32), status=torch.zeros(8, dtype=i64, device=dev),
            pred=e(lines, f32), scal=torch.zeros(2, dtype=f32, device=dev),
            ws_parse=_lib.workspace(lib.fm_parse_workspace(lines), dev),
            ws_fused=_lib.workspace(lib.fm_fused_workspace(lines), dev))
        return b

    def parse_ahead(self, lines, after=None):
        """parse of the next batch on a side stream (see Model.parse_ahead)"""
        if self._parse_stream is None:
            self._parse_stream = torch.cuda.Stream(device=self.device, priority=int(os.environ.get("FM_AHEAD_PRIORITY", "0")))
        self._ahead_slot = 3 - self._ahead_slot
        main, side = torch.cuda.current_stream(self.device), self._parse_stream
        b = self._slots.get(self._ahead_slot)
        side.wait_stream(main) if b is None or b.get("free") is None else side.wait_event(b["free"])
        if after is not None:
            side.wait_event(after)
        with torch.cuda.stream(side):
            bg = int(os.environ.get("FM_AHEAD_BLOCKS", Model.BACKGROUND_PARSE_BLOCKS))
            packed, b = self._parse(lines, slot=self._ahead_slot, max_blocks=bg)
            if self.exchange == "records":
                self._sort(b, packed.n_lines, max_blocks=bg)
            done = torch.cuda.Event()
            done.record(side)
        return dict(b=b, packed=packed, done=done, sorted=self.exchange == "records")

    def _parse(self, lines, slot=0, max_blocks=0):
        lib, st, cfg = self.lib, stream(), self.cfg
        packed = lines if isinstance(lines, PackedLines) else pack_lines(list(lines))
        if packed.text.device != self.device:
            packed = packed.to(self.device, non_blocking=True)
        b = self._buffers(packed.n_lines, packed.nnz_bound(), slot)
        n_lines, hashed = packed.n_lines, int(cfg.hash_feature_id)
        check(lib.fm_parse_count(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                 cfg.vocabulary_size, hashed, ptr(b["sizes"]), ptr(b["poses"]), ptr(b["status"]),
                                 ptr(b["ws_parse"]), b["ws_parse"].numel(), st), "fm_parse_count")
        check(lib.fm_parse_fill(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                cfg.vocabulary_size, hashed, ptr(b["poses"]), ptr(b["labels"]), ptr(b["ids"]), ptr(b["vals"]), b["cap"],
                                ptr(b["status"]), ptr(b["ws_parse"]), b["ws_parse"].numel(), max_blocks, st),
              "fm_parse_fill")
        return packed, b

    def _sort(self, b, n_lines, max_blocks=0):
        """this rank's occurrences sorted by (owner, row): fm_sort_by_row with n_ranks = world; counts in status[3:5]"""
        lib, dev, k = self.lib, self.device, self.cfg.factor_num
        if b.get("skeys") is None:
            cap = b["cap"]
            b["skeys"] = torch.empty(cap, dtype=torch.int32, device=dev)
            b["spairs"] = torch.empty(cap, 2, dtype=torch.int32, device=dev)
            b["ws_sort"] = _lib.workspace(lib.fm_sort_workspace(cap), dev)
            b["rec_ld"] = int(lib.fm_sorted_record_ld(k))
            b["rec"] = torch.empty(b["lines"], b["rec_ld"], dtype=torch.float32, device=dev)
            b["ws_push"] = _lib.workspace(lib.fm_merge_push_workspace(self.world), dev)
        check(lib.fm_sort_by_row(ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), n_lines, b["cap"], _at(b["status"], 1),
                                 self.cfg.vocabulary_size, self.world, ptr(b["skeys"]), ptr(b["spairs"]),
                                 _at(b["status"], 3), ptr(b["status"]), ptr(b["ws_sort"]), b["ws_sort"].numel(),
                                 max_blocks, stream()), "fm_sort_by_row")

    def _ensure_inbox(self, cap):
        """the peer-mapped inboxes of the record exchange, allocated by all ranks together at the first training step
        (capacity: the largest per-rank id buffer + 25 %); a later batch that does not fit voids its step"""
        if self._inbox is not None:
            return
        import torch.distributed._symmetric_memory as symm_mem
        need = torch.tensor([int(cap)], dtype=torch.int64, device=self.device)
        dist.all_reduce(need, op=dist.ReduceOp.MAX, group=self.group)
        self._inbox_cap = int(need.item()) * 5 // 4 + 64
        nbytes = int(self.lib.fm_merge_inbox_bytes(self._inbox_cap, self.table.shape[0], self.cfg.factor_num, self.world))
        self._inbox = symm_mem.empty(nbytes // 4, dtype=torch.float32, device=self.device)
        hdl = symm_mem.rendezvous(self._inbox, self.group)
        self._inbox.zero_()
        self._inbox_hdl = hdl
        self._inbox_bases = (ctypes.c_void_p * self.world)(*[int(p) for p in hdl.buffer_ptrs])
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def _train_records(self, packed, b, weights, global_batch, mark, presorted):
        """the step with the record exchange (see __init__)"""
        lib, st, cfg, k, g = self.lib, stream(), self.cfg, self.cfg.factor_num, self.world
        if not presorted:
            self._sort(b, packed.n_lines)
        self._ensure_inbox(b["cap"])
        check(lib.fm_peer_sorted_score(
            ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights), packed.n_lines, k, g,
            self._table_bases, self.ld, cfg.factor_lambda, cfg.bias_lambda, _lib.LOSS_KIND[cfg.loss_type], global_batch,
            ptr(b["pred"]), ptr(b["scal"]), _at(b["scal"], 1), ptr(b["rec"]), b["rec_ld"], ptr(b["status"]),
            ptr(b["ws_fused"]), b["ws_fused"].numel(), st), "fm_peer_sorted_score")
        mark("score")
        check(lib.fm_peer_push_records(
            ptr(b["skeys"]), ptr(b["spairs"]), _at(b["status"], 3), b["cap"], ptr(b["rec"]), b["rec_ld"], k, self.rank, g,
            cfg.vocabulary_size, self.table.shape[0], self._inbox_bases, self._inbox_cap, ptr(b["status"]),
            ptr(b["ws_push"]), b["ws_push"].numel(), st), "fm_peer_push_records")
        mark("push")
        flag = torch.empty(1, dtype=torch.int64, device=self.device)
        self._sync(ptr(b["status"]), None, None, ptr(flag), None)      # every inbox is complete; did any rank fail?
        mark("barrier")
        check(lib.fm_peer_merge_update(ptr(self.table), ptr(self.accum), self.ld, self.table.shape[0], ptr(self._inbox),
                                       self._inbox_cap, k, g, cfg.learning_rate, cfg.factor_lambda, cfg.bias_lambda,
                                       1.0 / cfg.batch_size, ptr(flag), st), "fm_peer_merge_update")
        mark("update")
        out = torch.empty(2, dtype=torch.float32, device=self.device)
        self._sync(None, ptr(b["scal"]), _at(b["scal"], 1), None, ptr(out))
        mark("allreduce_loss")
        return out, flag

    def train_step_async(self, lines, weights=None, global_batch=None, mark=lambda name: None, parsed=None):
        """Returns (scalars, flag): scalars = device f32 [global mean loss, global reg_score];
        flag = device i64, -1 when every rank parsed its slice, else rank << 48 | line << 3 | kind
        of a failing rank (the step was then void on every rank).  Nothing synchronises.
        parsed: handle from parse_ahead(lines) (then `lines` is not parsed again)."""
        lib, st, cfg, k, g = self.lib, stream(), self.cfg, self.cfg.factor_num, self.world
        if parsed is not None:
            torch.cuda.current_stream(self.device).wait_event(parsed["done"])
            packed, b = parsed["packed"], parsed["b"]
        else:
            packed, b = self._parse(lines)
        mark("parse")
        if global_batch is None:
            global_batch = packed.n_lines * g
        if weights is not None and not isinstance(weights, torch.Tensor):
            weights = torch.as_tensor(weights, dtype=torch.float32)
        if weights is not None:
            weights = weights.to(self.device)
        if self.exchange == "records":
            out, flag = self._train_records(packed, b, weights, global_batch, mark,
                                            parsed is not None and parsed.get("sorted", False))
            if parsed is not None:
                b["free"] = torch.cuda.Event()
                b["free"].record(torch.cuda.current_stream(self.device))
            self.global_step += 1
            self.last = dict(pred=b["pred"][:packed.n_lines], packed=packed, status=b["status"])
            return out, flag
        check(lib.fm_peer_score_scatter(
            ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights), packed.n_lines, k, g,
            self._table_bases, self.ld, self._gtab_bases, self.gld, self._touched_bases, cfg.factor_lambda,
            cfg.bias_lambda, _lib.LOSS_KIND[cfg.loss_type], global_batch, ptr(b["pred"]), ptr(b["scal"]),
            _at(b["scal"], 1), ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st),
            "fm_peer_score_scatter")
        mark("score_scatter")
        # barrier that also tells every rank whether any rank's parse failed (flags in peer-mapped memory)
        flag = torch.empty(1, dtype=torch.int64, device=self.device)
        self._sync(ptr(b["status"]), None, None, ptr(flag), None)
        mark("barrier")
        check(lib.fm_peer_touched_rows(ptr(self.touched), self.table.shape[0], ptr(self.row_list),
                                       self.row_list.numel(), ptr(self.n_rows), ptr(self.ws_touched),
                                       self.ws_touched.numel(), st), "fm_peer_touched_rows")
        check(lib.fm_fused_adagrad_claim(ptr(self.table), ptr(self.accum), self.ld, ptr(self.row_list),
                                         self.row_list.numel(), ptr(self.n_rows), k, ptr(self.gtab), self.gld,
                                         cfg.learning_rate, cfg.factor_lambda, cfg.bias_lambda, 1.0 / cfg.batch_size,
                                         ptr(flag), 1, st), "fm_fused_adagrad_claim")
        mark("update")
        out = torch.empty(2, dtype=torch.float32, device=self.device)
        # global loss and reg; also: nobody reads rows of the next step before all updates ran
        self._sync(None, ptr(b["scal"]), _at(b["scal"], 1), None, ptr(out))
        mark("allreduce_loss")
        if parsed is not None:
            b["free"] = torch.cuda.Event()
            b["free"].record(torch.cuda.current_stream(self.device))
        self.global_step += 1
        self.last = dict(pred=b["pred"][:packed.n_lines], packed=packed, status=b["status"])
        return out, flag

    def _sync(self, word_in, a_in, b_in, word_out, sums_out):
        """fm_peer_sync on the current stream: barrier over the ranks + max / sums of three device values"""
        self._epoch += 1
        check(self.lib.fm_peer_sync(self._sync_bases, self.rank, self.world, self._epoch, word_in, a_in, b_in,
                                    word_out, sums_out, stream()), "fm_peer_sync")

    def train_step(self, lines, weights=None, global_batch=None):
        out, flag = self.train_step_async(lines, weights, global_batch)
        f = int(flag.item())
        check_sync_word(f)
        if f >= 0:
            self.global_step -= 1        # void on every rank: not a step
            rank, first = f >> 48, f & ((1 << 48) - 1)
            line = self.last["packed"].line(first >> 3) if rank == self.rank else "<rank %d, line %d>" % (rank, first >> 3)
            raise FmParseError(first & 7, first >> 3, line)
        return float(out[0].item())

    def predict_async(self, lines, parsed=None):
        """scores of this rank's lines in b["pred"], rows read from whichever GPU owns them; no host sync.
        parsed: handle from parse_ahead(lines) for THIS batch (its parse ran under the previous batch's scorer)"""
        lib, st, cfg, k, g = self.lib, stream(), self.cfg, self.cfg.factor_num, self.world
        if parsed is not None:
            torch.cuda.current_stream(self.device).wait_event(parsed["done"])
            packed, b = parsed["packed"], parsed["b"]
        else:
            packed, b = self._parse(lines)
        check(lib.fm_peer_score_scatter(
            ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), None, None, packed.n_lines, k, g, self._table_bases,
            self.ld, None, self.gld, None, cfg.factor_lambda, cfg.bias_lambda, 0, 0, ptr(b["pred"]), ptr(b["scal"]),
            _at(b["scal"], 1), ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st),
            "fm_peer_score_scatter")
        if parsed is not None:
            b["free"] = torch.cuda.Event()
            b["free"].record(torch.cuda.current_stream(self.device))
        return packed, b

    def predict(self, lines):
        packed, b = self.predict_async(lines)
        host = b["status"].cpu()
        first = int(host[0]) & 0xFFFFFFFFFFFFFFFF
        if first != _lib.NO_ERROR:
            raise FmParseError(first & 7, first >> 3, packed.line(first >> 3))
        return b["pred"][:packed.n_lines].clone()
