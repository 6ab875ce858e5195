"""Host mirror of the reference's model layer (tffm/fm_model.py) for the hot path.

``Model`` reads the same INI file (sample.cfg; keys of fm_model.py:367-475) and
strings the ops together the way the reference graph does:

    parse -> unique + cumsum -> embedding_lookup -> fm_scorer            (:69-73, :291-293)
    loss (+ reg / batch_size) -> fm_grad -> sparse Adagrad               (:311-348)

but as one sync-free chain of C-ABI calls on the current CUDA stream: every
size that depends on the data (nnz, number of unique ids) stays on the device,
buffers are allocated once for the configured batch, and the host reads one
small status block per step.

What is NOT here (TF runtime machinery, out of scope per SURVEY.md section 8):
queue runners, the gRPC parameter server, SavedModel export, summaries.
"""
import collections
import configparser
import ctypes
import glob
import os

import torch

from . import _lib
from ._lib import check, load, ptr, stream
from .fm_ops import FmParseError, PackedLines, pack_lines


def _no_mark(name):
    pass


def _stride_for(width):
    """row stride (floats) for rows of `width` floats: a multiple of 4 (16-byte aligned rows,
    128-bit loads); rows of up to 128 bytes get a power-of-two stride so that no row straddles
    a 128-byte line (ncu: an 80-byte row at an 80-byte stride costs ~190 B of DRAM per read,
    at a 128-byte stride 128 B)."""
    ld = (width + 3) // 4 * 4
    if ld <= 32:
        p = 4
        while p < ld:
            p *= 2
        ld = p
    return ld


def _ld_for(factor_num):
    return _stride_for(factor_num + 1)


class CheckpointWriter:
    """Writes checkpoint files off the training thread: a device-side snapshot is drained through two pinned
    32 MiB bounce buffers on its own stream (asynchronous DMA -- a pageable cudaMemcpy or a table-sized
    cudaHostAlloc from another thread holds the driver's lock and stalls every launch of the training
    thread), turned into a state dict and written with torch.save + an atomic rename.  One file in flight."""

    def __init__(self, device):
        self.device = device
        self._stream = self._bounce = self._thread = self._error = None

    def prepare(self):
        """page-locks the bounce buffers (do it before the training loop: it stalls other CUDA calls)"""
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=self.device)
            self._bounce = torch.empty(2, 8 << 20, dtype=torch.float32).pin_memory()

    def busy(self):
        return self._thread is not None and self._thread.is_alive()

    def wait(self):
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        err, self._error = self._error, None
        if err is not None:
            raise RuntimeError("writing the checkpoint failed: %r" % (err,))

    def write(self, path, snap, make_state, wait=True, after=None):
        """snap: contiguous f32 device tensor holding the snapshot; make_state(host copy of snap) -> the dict
        to save; after(): called once the file is in place (e.g. removes files it supersedes)"""
        import threading
        self.wait()
        self.prepare()
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(self.device))
        snap.record_stream(self._stream)
        stream, bounce = self._stream, self._bounce

        def work():
            try:
                host = torch.empty(snap.shape, dtype=torch.float32)
                src, dst, chunk = snap.view(-1), host.view(-1), bounce.shape[1]
                prev = None
                with torch.cuda.stream(stream):
                    stream.wait_event(ready)
                    for i, lo in enumerate(range(0, src.numel(), chunk)):
                        hi = min(src.numel(), lo + chunk)
                        bounce[i % 2, :hi - lo].copy_(src[lo:hi], non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(stream)
                        if prev is not None:
                            prev[0].synchronize()
                            dst[prev[1]:prev[2]].copy_(bounce[prev[3], :prev[2] - prev[1]])
                        prev = (ev, lo, hi, i % 2)
                    if prev is not None:
                        prev[0].synchronize()
                        dst[prev[1]:prev[2]].copy_(bounce[prev[3], :prev[2] - prev[1]])
                os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
                torch.save(make_state(host), path + ".tmp")
                os.replace(path + ".tmp", path)          # a reader never sees a half-written file
                if after is not None:
                    after()
            except BaseException as e:                   # surfaces in wait()
                self._error = e
                if wait:
                    raise

        if wait:
            work()
            self._error = None
        else:
            self._thread = threading.Thread(target=work, name="fm-checkpoint")
            self._thread.start()


def shard_sets(path):
    """complete sharded checkpoint sets next to `path`: {g: [file of rank 0, ..., file of rank g-1]}"""
    sets = {}
    for f in glob.glob(glob.escape(path) + ".shard-*-of-*"):
        try:
            r, g = f[len(path) + len(".shard-"):].split("-of-")
            sets.setdefault(int(g), {})[int(r)] = f
        except ValueError:
            continue
    return {g: [files[r] for r in range(g)] for g, files in sets.items()
            if len(files) == g and all(r in files for r in range(g))}


def _step_of(path):
    return int(torch.load(path, map_location="cpu", mmap=True).get("global_step", 0))


def latest_checkpoint(path):
    """(global_step, [files]) of the newest checkpoint at `path`: the single file Model.save writes, or a
    sharded set of a multi-GPU run (`path.shard-r-of-g`, file r = block r of a mod-g partition); None if none"""
    found = [(_step_of(path), [path])] if os.path.exists(path) else []
    for g, files in shard_sets(path).items():
        steps = {_step_of(f) for f in files}
        if len(steps) == 1:                              # a set caught half-way through a save is not a checkpoint
            found.append((steps.pop(), files))
    return max(found, key=lambda c: c[0]) if found else None


def load_checkpoint(path):
    """merged state dict of latest_checkpoint(path)"""
    best = latest_checkpoint(path)
    if best is None:
        raise IOError("no checkpoint at %s" % path)
    state = {}
    for f in best[1]:
        state.update(torch.load(f, map_location="cpu"))
    return state


class Model(object):
    # defaults of the reference class attributes (tffm/fm_model.py:10-32)
    vocabulary_size = 8000000
    vocabulary_block_num = 100
    factor_num = 10
    hash_feature_id = False
    log_dir = None
    model_file = None
    batch_size = 50000
    init_value_range = 0.01
    factor_lambda = 0.0
    bias_lambda = 0.0
    learning_rate = 0.01
    adagrad_init_accumulator = 0.1
    num_epochs = 10
    loss_type = 'mse'
    save_steps = 100
    save_summaries_steps = 100
    queue_size = 10000
    shuffle_threads = 1
    tolerance = None
    optimizer = 'adagrad'     # extension: 'sgd' (sorted path only); the reference always builds AdagradOptimizer

    def __init__(self, config_file=None, **overrides):
        self.train_files, self.weight_files, self.predict_files = [], [], []
        self.validation_data_files, self.validation_weight_files = [], []
        if config_file is not None:
            self._get_config(config_file)
        for k, v in overrides.items():
            if not hasattr(self, k):
                raise AttributeError("unknown model option %r" % k)
            setattr(self, k, v)
        self.table = self.accum = None
        self.global_step = 0
        self._bufs = None
        self._slots = {}
        self._ahead_slot = 2     # parse_ahead alternates between buffer sets 1 and 2; set 0 serves inline steps / predict
        self._parse_stream = None
        self._bg_blocks, self._late, self._late_q = None, 0.0, collections.deque()
        # claim path: which of its two forms runs (see _claim_form); "auto" measures, "rows" / "occurrences" pin it
        self.claim_form = "auto"
        self._form, self._form_q, self._form_age, self._rows_state = "rows", collections.deque(), 0, None

    # ---- INI config, same sections / keys as tffm/fm_model.py:367-475 -----------------
    def _get_config(self, config_file, verbose=True):
        cfg = configparser.ConfigParser(interpolation=None)
        if not cfg.read(config_file):
            raise IOError("cannot read config file %s" % config_file)
        G, T, P = 'General', 'Train', 'Predict'

        def read(section, option, not_null=True):
            if not cfg.has_option(section, option):
                if not_null:
                    raise ValueError("%s is undefined." % option)
                return None
            value = cfg.get(section, option)
            if verbose:
                print('  {0} = {1}'.format(option, value))
            return value

        def read_strs(section, option, not_null=True):
            val = read(section, option, not_null)
            return None if val is None else [s.strip() for s in val.split(',')]

        def expand(patterns):
            return sorted(sum((glob.glob(f) for f in patterns), []))

        if verbose:
            print('Config: ')
        self.vocabulary_size = int(read(G, 'vocabulary_size'))
        self.vocabulary_block_num = int(read(G, 'vocabulary_block_num'))
        self.factor_num = int(read(G, 'factor_num'))
        self.hash_feature_id = read(G, 'hash_feature_id').strip().lower() == 'true'
        self.log_dir = read(G, 'log_dir', False)
        self.model_file = read(G, 'model_file', False)
        if cfg.has_option(G, 'save_summaries_steps'):
            self.save_summaries_steps = int(read(G, 'save_summaries_steps'))
        self.batch_size = int(read(T, 'batch_size'))
        self.init_value_range = float(read(T, 'init_value_range'))
        self.factor_lambda = float(read(T, 'factor_lambda'))
        self.bias_lambda = float(read(T, 'bias_lambda'))
        self.num_epochs = int(read(T, 'epoch_num'))
        self.learning_rate = float(read(T, 'learning_rate'))
        self.adagrad_init_accumulator = float(read(T, 'adagrad.initial_accumulator'))
        if not self.adagrad_init_accumulator > 0:
            raise ValueError("initial_accumulator_value must be positive: %s" % self.adagrad_init_accumulator)
        self.loss_type = read(T, 'loss_type').strip().lower()
        if self.loss_type not in ('logistic', 'mse'):
            raise ValueError("loss_type must be logistic or mse")
        if cfg.has_option(T, 'optimizer'):      # not a key of the reference (tffm/fm_model.py:341-344 is Adagrad only)
            self.optimizer = read(T, 'optimizer').strip().lower()
            if self.optimizer not in ('adagrad', 'sgd'):
                raise ValueError("optimizer must be adagrad or sgd")
        v = read(T, 'save_steps', False)
        if v is not None:
            self.save_steps = int(v)
        if cfg.has_option(T, 'queue_size'):
            self.queue_size = int(read(T, 'queue_size'))
        if cfg.has_option(T, 'shuffle_threads'):
            self.shuffle_threads = int(read(T, 'shuffle_threads'))
        train_files = read_strs(T, 'train_files')
        self.train_files = expand(train_files)
        weight_files = read_strs(T, 'weight_files', False)
        if weight_files is not None:
            self.weight_files = expand(weight_files)
            if len(train_files) != len(weight_files):
                raise ValueError('The numbers of train files and weight files do not match.')
        vfiles = read_strs(T, 'validation_files', False)
        if vfiles is not None:
            self.validation_data_files = expand(vfiles)
            self.tolerance = float(read(T, 'tolerance'))
        vwfiles = read_strs(T, 'validation_weight_files', False)
        if vwfiles is not None:
            self.validation_weight_files = expand(vwfiles)
            if vfiles is None or len(vfiles) != len(vwfiles):
                raise ValueError('The numbers of validation data files and validation weight files do not match.')
        pfiles = read_strs(P, 'predict_files', False)
        self.predict_files = expand(pfiles) if pfiles is not None else []

    # ---- the vocabulary table ------------------------------------------------------
    def build(self, device=None, init_table=None, seed=None, max_nnz_per_line=None, ld=None,
              layout=None, path=None, inference=False):
        """Allocates table + Adagrad accumulator (+ table-shaped gradient accumulator) in HBM.
        The table is [vocabulary_size, ld] (row = feature id, see _stride_for); the reference's
        vocabulary_block_num blocks (fm_model.py:269-284) are a relabelling applied when
        checkpoints are written (blocks_state / load_blocks_state).
        layout "interleaved": one array whose row holds [w | h | g], the weight, Adagrad and the claim path's
        gradient-accumulator rows of one id side by side in whole 128-byte lines; "pair": [w | h] only -- what
        the sorted path touches per id, two thirds of the memory; "split": separate arrays.  Every path runs on
        every layout (a path that needs an accumulator table the layout lacks allocates it on first use).
        Default: the layout of the default path.
        path: the default of train_step / parse_ahead.  "claim" unless the optimizer is sgd: measured on one
        B200 at cfg2 it is the faster of the two (uniform ids 0.56 vs 0.63 ms per step, zipf 0.34 vs 0.57;
        profiles/README.md r2); "sorted" is the reproducible one and the one that needs no third table.
        init: uniform(-init_value_range, init_value_range) (:278-281) or `init_table`
        ([vocabulary_size, K+1], for parity runs).
        inference: the weight table alone (layout "split", no Adagrad accumulator): predict only."""
        lib = load()
        if not self.adagrad_init_accumulator > 0:      # tf.train.AdagradOptimizer raises the same way
            raise ValueError("initial_accumulator_value must be positive: %s" % self.adagrad_init_accumulator)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        k, v = self.factor_num, self.vocabulary_size
        self.ld = _ld_for(k) if ld is None else int(ld)
        assert self.ld >= k + 1 and self.ld % 4 == 0
        self.gld = _stride_for(int(lib.fm_fused_gacc_ld(k)))
        if self.optimizer not in ("adagrad", "sgd"):
            raise ValueError("optimizer must be adagrad or sgd")
        sortable = (k + 4) // 4 <= 96 and v <= 1 << 32
        if path is None:
            path = "sorted" if self.optimizer == "sgd" else "claim"
        if inference:
            layout = "split"
        if layout is None:
            layout = "pair" if path == "sorted" else "interleaved"
        if layout not in ("pair", "interleaved", "split"):
            raise ValueError("layout must be pair, interleaved or split")
        if path not in self.PATHS or (path == "sorted" and not sortable) or (self.optimizer == "sgd" and path != "sorted"):
            raise ValueError("path %r not available for factor_num %d / optimizer %s" % (path, k, self.optimizer))
        self.layout = layout
        self.default_path = path
        if layout == "pair":
            # [w | h] of one id packed at 16-byte granularity in whole 128-byte lines (K=16: 2 x 20 floats in
            # 256 B; the weight row, which the scorer reads alone, never straddles a line)
            part = (k + 1 + 3) // 4 * 4 if ld is None else int(ld)
            block = (2 * part + 31) // 32 * 32 if ld is None else 2 * part
            self._store = torch.zeros(v, block, dtype=torch.float32, device=self.device)
            self.table = self._store[:, :part]
            self.gtab = None
        elif layout == "interleaved":
            # [w | h | g] of one id packed at 16-byte granularity in a block that is a whole number of
            # 128-byte lines (K=16: 3 x 20 floats in 256 B = 2 lines; the weight row never straddles)
            part = max((k + 1 + 3) // 4 * 4, int(lib.fm_fused_gacc_ld(k)))
            block = (3 * part + 31) // 32 * 32
            if ld is not None:
                part, block = int(ld), 3 * int(ld)
            self._store = torch.zeros(v, block, dtype=torch.float32, device=self.device)
            self.table = self._store[:, :part]
            self.gtab = self._store[:, 2 * part:3 * part]
        else:
            self.table = torch.zeros(v, self.ld, dtype=torch.float32, device=self.device)
            self.gtab = None  # allocated on the first step that needs it
        if init_table is not None:
            self.table[:, :k + 1] = torch.as_tensor(init_table, dtype=torch.float32)[:, :k + 1].to(self.device)
        else:
            gen = torch.Generator(device=self.device)
            if seed is not None:
                gen.manual_seed(seed)
            self.table[:, :k + 1].uniform_(-self.init_value_range, self.init_value_range, generator=gen)
        if inference:
            self.accum = None
        elif layout in ("pair", "interleaved"):
            self.accum = self._store[:, part:2 * part]
            self.accum.fill_(self.adagrad_init_accumulator)
        else:
            self.accum = torch.full_like(self.table, self.adagrad_init_accumulator)
        self._bufs = None
        self._slots = {}
        self._max_nnz_per_line = max_nnz_per_line
        return self

    def parse_ahead(self, packed: PackedLines, after=None, sort=None):
        """Enqueues the parse of a device-resident batch on a side stream, into the buffer set the
        current step is not using, and returns a handle for train_step_async(parsed=...).  The
        parser is instruction-bound while the table kernels are memory-bound, so the two overlap."""
        if self._parse_stream is None:
            self._parse_stream = torch.cuda.Stream(device=self.device)
        self._ahead_slot = 3 - self._ahead_slot
        slot = self._ahead_slot
        b = self._buffers(packed.n_lines, packed.nnz_bound(), slot=slot)
        main = torch.cuda.current_stream(self.device)
        side = self._parse_stream
        side.wait_stream(main) if b.get("free") is None else side.wait_event(b["free"])
        if after is not None:      # e.g. the event of the H2D copy that brings `packed` in
            side.wait_event(after)
        if self._bg_blocks is None:
            self._bg_blocks = self.BACKGROUND_PARSE_BLOCKS
        sort = self.default_path == "sorted" if sort is None else sort
        with torch.cuda.stream(side):
            self._parse_unique(packed, b, unique=False, max_blocks=self._bg_blocks)
            if sort:        # the sort needs only the ids: it runs ahead with the parse
                self._sort(b, packed.n_lines, max_blocks=self._bg_blocks)
            done = torch.cuda.Event(enable_timing=True)
            done.record(side)
        return dict(b=b, done=done, packed=packed, sorted=sort)

    def _adapt_parse_grid(self, need, done):
        """The run-ahead parse is a thin background kernel (BACKGROUND_PARSE_BLOCKS blocks) so that it takes
        little from the memory-bound table kernels.  When those are short (few distinct ids per batch) the
        thin parse finishes after the step needs it: `need` (main stream, just before it waits) then precedes
        `done` (side stream, end of the parse).  Read a few steps later, without blocking; if the parse keeps
        being late its grid is doubled, up to the full grid."""
        q = self._late_q
        q.append((need, done))
        while q and q[0][0].query() and q[0][1].query():
            n, d = q.popleft()
            self._late = 0.8 * self._late + 0.2 * n.elapsed_time(d)      # ms, > 0: the step waited
        if self._late > 0.02 and self._bg_blocks:
            self._bg_blocks = self._bg_blocks * 2 if self._bg_blocks < 1024 else 0    # 0 = no cap
            self._late = 0.0

    # ---- the two forms of the claim path ------------------------------------------------------------------
    # "occurrences": fm_fused_score_scatter + fm_fused_adagrad_claim over the N occurrences (each tries to claim its
    #                row; a row that occurs m times costs m - 1 lost claims of three row loads each);
    # "rows":        the scorer also sets one bit per touched row (fm_peer_score_scatter with one rank), the bitmap
    #                becomes an ascending row list (fm_peer_touched_rows) and the update walks the U distinct rows.
    # One B200, cfg2 (profiles/r2_claim_forms_probe.txt): uniform ids (U = 0.93 N) 0.489 against 0.499 ms for the
    # two table kernels, skewed ids (U = 0.10 N) 0.239 against 0.199 -- the bitmap scan costs what it costs, the lost
    # claims scale with N - U.  So the model measures: a "rows" step leaves U and N behind, read a few steps later
    # without blocking; U / N above ROWS_FORM_BELOW switches to "occurrences", and every FORM_PROBE_EVERY steps one
    # "rows" step looks again.
    ROWS_FORM_BELOW = 0.7
    FORM_PROBE_EVERY = 512

    def _claim_form(self):
        if self.claim_form != "auto":
            return self.claim_form
        q = self._form_q
        while q and q[0][0].query():
            _, rec = q.popleft()
            n_rows, nnz = int(rec[0]), int(rec[1])
            if nnz > 0:
                self._form = "rows" if n_rows < self.ROWS_FORM_BELOW * nnz else "occurrences"
        if self._form == "occurrences":
            self._form_age += 1
            if self._form_age >= self.FORM_PROBE_EVERY:
                self._form_age = 0
                return "rows"
        return self._form

    def _rows_buffers(self, cap):
        """bitmap, row list and scan scratch of the "rows" form (one set per model; grown with the batch capacity)"""
        r = self._rows_state
        if r is not None and r["cap"] >= cap:
            return r
        if r is not None:
            torch.cuda.synchronize(self.device)          # enqueued kernels may still use the old set
        lib, dev, v = load(), self.device, self.vocabulary_size
        r = dict(cap=cap, touched=torch.zeros((v + 31) // 32, dtype=torch.int32, device=dev),
                 rows=torch.empty(min(cap, v), dtype=torch.int64, device=dev),
                 n_rows=torch.zeros(1, dtype=torch.int64, device=dev),
                 ws=_lib.workspace(lib.fm_peer_touched_workspace(v), dev),
                 rec=[torch.zeros(2, dtype=torch.int64).pin_memory() for _ in range(4)], at=0)
        self._rows_state = r
        return r

    def _gtab(self):
        if self.gtab is None:
            self.gtab = torch.zeros(self.vocabulary_size, self.gld, dtype=torch.float32, device=self.device)
        return self.gtab

    def _buffers(self, n_lines, nnz_bound, slot=0):
        """per-batch device buffers, allocated once and grown on demand (two sets: the second one
        lets parse_ahead() fill the next batch while the current one trains).
        nnz_bound: PackedLines.nnz_bound() of the batch -- sized from the batch's own lines, not from the
        buffer they live in (a batch selected from a resident file shares the file's whole text)"""
        b = self._buffers_one(self._slots.get(slot), n_lines, nnz_bound)
        self._slots[slot] = b
        self._bufs = b          # the set used last (tests and callers read results from it)
        return b

    def _buffers_one(self, b, n_lines, nnz_bound):
        cap = int(nnz_bound)
        if self._max_nnz_per_line:
            cap = min(cap, int(n_lines * self._max_nnz_per_line))
        if b is not None and b["lines"] >= n_lines and b["cap"] >= cap:
            return b
        lib, dev = load(), self.device
        if b is not None:
            # growing: work enqueued on either stream may still use the old set, whose memory goes back to
            # the allocator below (it only tracks the stream that allocated it)
            torch.cuda.synchronize(dev)
        lines = max(n_lines, b["lines"] if b else 0)
        cap = max(cap, b["cap"] if b else 0)
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        b = dict(
            lines=lines, cap=cap,
            labels=torch.empty(lines, dtype=f32, device=dev), sizes=torch.empty(lines, dtype=i32, device=dev),
            poses=torch.empty(lines + 1, dtype=i32, device=dev), ids=torch.empty(cap, dtype=i64, device=dev),
            vals=torch.empty(cap, dtype=f32, device=dev), ori=None, local=None, ws_unique=None, rows=None,
            grad=None, pred=torch.empty(lines, dtype=f32, device=dev),
            gacc_ld=int(lib.fm_fused_gacc_ld(self.factor_num)),
            gacc=None,
            ws_fused=_lib.workspace(lib.fm_fused_workspace(lines), dev),
            pred_grad=torch.empty(lines, dtype=f32, device=dev),
            # [first_error, nnz, n_lines, reserved] (fm_parse_status), n_unique, then loss and reg as f32 bits
            status=torch.zeros(8, dtype=i64, device=dev),
            ws_parse=_lib.workspace(lib.fm_parse_workspace(lines), dev),
            ws_score=_lib.workspace(lib.fm_score_workspace(lines), dev),
            ws_loss=_lib.workspace(lib.fm_loss_workspace(lines), dev),
            skeys=None,      # the sorted path's buffers, allocated on first use (_sort)
        )
        b["scalars"] = b["status"][6:8].view(torch.float32)  # [loss, -, reg, -]
        return b

    BACKGROUND_PARSE_BLOCKS = 2 * 148   # grid cap of the fill pass when it runs ahead on the side stream

    def _parse_unique(self, packed, b, mark=_no_mark, unique=True, max_blocks=0):
        """parse (-> unique) (+ cumsum), all sizes device-resident.
        mark(name) is called after each stage is enqueued (bench.py records CUDA events there)."""
        lib, st = load(), stream()
        n_lines = packed.n_lines
        hashed = int(self.hash_feature_id)
        status, nnz_dev, nuniq_dev = ptr(b["status"]), _off(b["status"], 1), _off(b["status"], 4)
        check(lib.fm_parse_count(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                 self.vocabulary_size, hashed, ptr(b["sizes"]), ptr(b["poses"]), status,
                                 ptr(b["ws_parse"]), b["ws_parse"].numel(), st), "fm_parse_count")
        mark("parse_count")
        check(lib.fm_parse_fill(ptr(packed.text), packed.text_bytes, ptr(packed.begins), ptr(packed.ends), -1, n_lines,
                                self.vocabulary_size, hashed, ptr(b["poses"]), ptr(b["labels"]), ptr(b["ids"]), ptr(b["vals"]), b["cap"], status,
                                ptr(b["ws_parse"]), b["ws_parse"].numel(), max_blocks, st), "fm_parse_fill")
        mark("parse_fill")
        if not unique:
            return
        if b["ws_unique"] is None:   # only the fused / composed paths de-duplicate
            dev = self.device
            b["ori"] = torch.empty(b["cap"], dtype=torch.int64, device=dev)
            b["local"] = torch.empty(b["cap"], dtype=torch.int32, device=dev)
            b["ws_unique"] = _lib.workspace(lib.fm_unique_workspace(b["cap"]), dev)
        check(lib.fm_unique(ptr(b["ids"]), b["cap"], nnz_dev, ptr(b["ori"]), ptr(b["local"]), nuniq_dev,
                            ptr(b["ws_unique"]), b["ws_unique"].numel(), st), "fm_unique")
        mark("unique")

    def _forward(self, packed, b, mark=_no_mark):
        """parse -> unique -> lookup -> score: the reference's op chain, one C-ABI call per op"""
        lib, st = load(), stream()
        n_lines, k = packed.n_lines, self.factor_num
        status, nuniq_dev = ptr(b["status"]), _off(b["status"], 4)
        if b["rows"] is None:
            b["rows"] = torch.empty(b["cap"], self.ld, dtype=torch.float32, device=self.device)
        self._parse_unique(packed, b, mark)
        check(lib.fm_lookup(ptr(self.table), self.table.stride(0), ptr(b["ori"]), b["cap"], nuniq_dev, k,
                            ptr(b["rows"]), self.ld, status, st), "fm_lookup")
        mark("lookup")
        check(lib.fm_score(ptr(b["local"]), ptr(b["rows"]), self.ld, ptr(b["vals"]), ptr(b["poses"]), n_lines, k,
                           self.factor_lambda, self.bias_lambda, ptr(b["pred"]), _off(b["status"], 7),
                           ptr(b["ws_score"]), b["ws_score"].numel(), st), "fm_score")
        mark("score")

    def _sort(self, b, n_lines, mark=_no_mark, max_blocks=0):
        """occurrences of the parsed batch in `b` sorted by table row (fm_sort_by_row); counts land in
        status[3] = nnz sorted, status[4] = n_unique (once the update has run)"""
        lib, st, dev = load(), stream(), self.device
        if b.get("skeys") is None:
            cap = b["cap"]
            b["skeys"] = torch.empty(cap, dtype=torch.int32, device=dev)
            b["spairs"] = torch.empty(cap, 2, dtype=torch.int32, device=dev)
            b["ws_sort"] = _lib.workspace(lib.fm_sort_workspace(cap), dev)
            b["rec_ld"] = int(lib.fm_sorted_record_ld(self.factor_num))
            b["rec"] = torch.empty(b["lines"], b["rec_ld"], dtype=torch.float32, device=dev)
            b["ws_sorted"] = _lib.workspace(lib.fm_sorted_workspace(cap, self.factor_num), dev)
        check(lib.fm_sort_by_row(ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), n_lines, b["cap"], _off(b["status"], 1),
                                 self.vocabulary_size, 1, ptr(b["skeys"]), ptr(b["spairs"]),
                                 _off(b["status"], 3), ptr(b["status"]), ptr(b["ws_sort"]), b["ws_sort"].numel(),
                                 max_blocks, st), "fm_sort_by_row")
        mark("sort")

    PATHS = ("sorted", "claim", "fused", "composed")

    def train_step_async(self, packed: PackedLines, weights=None, mark=_no_mark, path=None, parsed=None,
                         prefetch=None):
        """One training step on a device-resident batch; enqueues only.  Returns the
        buffers dict: status[0] parse status word, status[1] nnz, status[4] n_unique,
        scalars[0] loss, scalars[2] reg.
        path "sorted"   (default) parse -> fm_sort_by_row -> fm_sorted_score -> fm_sorted_update: FmGrad's
                        scatter-add and the optimizer step as one segment reduce over the occurrences
                        sorted by row; no atomics, no accumulator rows, reproducible sums;
             "claim"    parse -> fm_fused_score_scatter into table-shaped accumulator
                        rows -> fm_fused_adagrad_claim: no tf.unique, each table row read once;
             "fused"    parse -> unique -> fm_fused_score_scatter (compact accumulators) ->
                        fm_fused_adagrad;
             "composed" the reference's op-by-op chain (unique, lookup, scorer, loss, grad, Adagrad).
        parsed:   handle from parse_ahead() for THIS batch (claim path);
        prefetch: the NEXT batch (device-resident): its parse is started on the side stream right
                  after this batch's scorer kernel, i.e. under the DRAM-bound update kernel, and the
                  handle comes back as result["ahead"]."""
        path = self.default_path if path is None else path
        if path not in self.PATHS:
            raise ValueError("path must be one of %s" % (self.PATHS,))
        if self.accum is None:
            raise RuntimeError("this model was built with inference=True: it has no optimizer state to train with")
        if parsed is not None and path not in ("claim", "sorted"):
            raise ValueError("parse_ahead() feeds the sorted and claim paths only")
        if self.optimizer != "adagrad" and path != "sorted":
            raise ValueError("optimizer %r runs on the sorted path only" % self.optimizer)
        b = parsed["b"] if parsed is not None else self._buffers(packed.n_lines, packed.nnz_bound())
        lib, st = load(), stream()
        n_lines, k = packed.n_lines, self.factor_num
        tstride = self.table.stride(0)
        if path == "sorted":
            if parsed is not None:
                main = torch.cuda.current_stream(self.device)
                need = torch.cuda.Event(enable_timing=True)
                need.record(main)
                main.wait_event(parsed["done"])
                self._adapt_parse_grid(need, parsed["done"])
                mark("parse_count")
                mark("parse_fill")
                if parsed.get("sorted"):
                    mark("sort")
                else:
                    self._sort(b, n_lines, mark)
            else:
                self._parse_unique(packed, b, mark, unique=False)
                self._sort(b, n_lines, mark)
            check(lib.fm_sorted_score(
                ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights), n_lines, k,
                ptr(self.table), tstride, self.factor_lambda, self.bias_lambda, _lib.LOSS_KIND[self.loss_type], 0,
                ptr(b["pred"]), _off(b["status"], 6), _off(b["status"], 7), ptr(b["rec"]), b["rec_ld"],
                ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st), "fm_sorted_score")
            mark("score")
            check(lib.fm_sorted_update(
                ptr(self.table), ptr(self.accum), tstride, ptr(b["skeys"]), ptr(b["spairs"]),
                _off(b["status"], 3), b["cap"], ptr(b["rec"]), b["rec_ld"], k, _lib.OPTIMIZER[self.optimizer],
                self.learning_rate, self.factor_lambda, self.bias_lambda, 1.0 / self.batch_size, 0,
                ptr(b["status"]), ptr(b["ws_sorted"]), b["ws_sorted"].numel(), st), "fm_sorted_update")
            mark("update")
            if parsed is not None:   # the buffer set may be refilled once this step is through with it
                b["status_snapshot"] = b["status"].clone()
                b["free"] = torch.cuda.Event()
                b["free"].record(torch.cuda.current_stream(self.device))
            self.global_step += 1
            return b
        if path == "claim":
            gtab = self._gtab()
            if parsed is not None:
                main = torch.cuda.current_stream(self.device)
                need = torch.cuda.Event(enable_timing=True)
                need.record(main)
                main.wait_event(parsed["done"])
                self._adapt_parse_grid(need, parsed["done"])
                mark("parse_count")
                mark("parse_fill")
            else:
                self._parse_unique(packed, b, mark, unique=False)
            form = self._claim_form()
            rs = self._rows_buffers(b["cap"]) if form == "rows" else None
            if rs is not None:
                one = ctypes.c_void_p * 1                # the sharded scorer with a single rank: this GPU's arrays
                check(lib.fm_peer_score_scatter(
                    ptr(b["ids"]), ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights), n_lines, k, 1,
                    one(self.table.data_ptr()), tstride, one(gtab.data_ptr()), gtab.stride(0),
                    one(rs["touched"].data_ptr()), self.factor_lambda,
                    self.bias_lambda, _lib.LOSS_KIND[self.loss_type], 0, ptr(b["pred"]), _off(b["status"], 6),
                    _off(b["status"], 7), ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st),
                    "fm_peer_score_scatter")
            else:
                check(lib.fm_fused_score_scatter(
                    ptr(b["ids"]), None, ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights),
                    n_lines, k, ptr(self.table), tstride, self.factor_lambda, self.bias_lambda,
                    _lib.LOSS_KIND[self.loss_type], 0, ptr(b["pred"]), _off(b["status"], 6), _off(b["status"], 7),
                    ptr(gtab), gtab.stride(0), ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st),
                    "fm_fused_score_scatter")
            mark("score_scatter")
            ahead = None
            if prefetch is not None:
                scored = torch.cuda.Event()
                scored.record(torch.cuda.current_stream(self.device))
                ahead = self.parse_ahead(prefetch, after=scored)
            if rs is not None and (k + 6) // 4 <= 96:
                # the update walks the bitmap itself (skewed ids: 0.031 ms against 0.015 + 0.027 for scan + list walk);
                # a batch that did not parse scattered nothing: the bitmap is empty, the update a no-op
                auto = self.claim_form == "auto"
                if auto:
                    rs["n_rows"].zero_()
                check(lib.fm_fused_adagrad_bitmap(
                    ptr(self.table), ptr(self.accum), tstride, ptr(rs["touched"]), self.vocabulary_size, k,
                    ptr(gtab), gtab.stride(0), self.learning_rate, self.factor_lambda, self.bias_lambda,
                    1.0 / self.batch_size, ptr(b["status"]), 0, ptr(rs["n_rows"]) if auto else None, st),
                    "fm_fused_adagrad_bitmap")
            elif rs is not None:                         # rows too wide for the bitmap kernel: bitmap -> list -> update
                check(lib.fm_peer_touched_rows(ptr(rs["touched"]), self.vocabulary_size, ptr(rs["rows"]),
                                               rs["rows"].numel(), ptr(rs["n_rows"]), ptr(rs["ws"]), rs["ws"].numel(), st),
                      "fm_peer_touched_rows")
                check(lib.fm_fused_adagrad_claim(
                    ptr(self.table), ptr(self.accum), tstride, ptr(rs["rows"]), rs["rows"].numel(), ptr(rs["n_rows"]), k,
                    ptr(gtab), gtab.stride(0), self.learning_rate, self.factor_lambda, self.bias_lambda,
                    1.0 / self.batch_size, ptr(b["status"]), 0, st), "fm_fused_adagrad_claim")
            else:
                check(lib.fm_fused_adagrad_claim(
                    ptr(self.table), ptr(self.accum), tstride, ptr(b["ids"]), b["cap"], _off(b["status"], 1), k,
                    ptr(gtab), gtab.stride(0), self.learning_rate, self.factor_lambda, self.bias_lambda,
                    1.0 / self.batch_size, ptr(b["status"]), 0, st), "fm_fused_adagrad_claim")
            if rs is not None and self.claim_form == "auto":     # U and N of this step, for _claim_form a few steps on
                rec = rs["rec"][rs["at"] % len(rs["rec"])]
                rs["at"] += 1
                rec[0:1].copy_(rs["n_rows"], non_blocking=True)
                rec[1:2].copy_(b["status"][1:2], non_blocking=True)
                seen = torch.cuda.Event()
                seen.record(torch.cuda.current_stream(self.device))
                self._form_q.append((seen, rec))
            mark("adagrad")
            b["ahead"] = ahead
            if parsed is not None:   # the buffer set may be refilled once this step is through with it
                b["status_snapshot"] = b["status"].clone()
                b["free"] = torch.cuda.Event()
                b["free"].record(torch.cuda.current_stream(self.device))
            self.global_step += 1
            return b
        if path == "fused":
            if b["gacc"] is None:
                b["gacc"] = torch.zeros(b["cap"], b["gacc_ld"], dtype=torch.float32, device=self.device)
            self._parse_unique(packed, b, mark)
            check(lib.fm_fused_score_scatter(
                ptr(b["ids"]), ptr(b["local"]), ptr(b["vals"]), ptr(b["poses"]), ptr(b["labels"]), ptr(weights),
                n_lines, k, ptr(self.table), tstride, self.factor_lambda, self.bias_lambda,
                _lib.LOSS_KIND[self.loss_type], 0, ptr(b["pred"]), _off(b["status"], 6), _off(b["status"], 7),
                ptr(b["gacc"]), b["gacc_ld"], ptr(b["status"]), ptr(b["ws_fused"]), b["ws_fused"].numel(), st),
                "fm_fused_score_scatter")
            mark("score_scatter")
            check(lib.fm_fused_adagrad(
                ptr(self.table), ptr(self.accum), tstride, ptr(b["ori"]), b["cap"], _off(b["status"], 4), k,
                ptr(b["gacc"]), b["gacc_ld"], self.learning_rate, self.factor_lambda, self.bias_lambda,
                1.0 / self.batch_size, ptr(b["status"]), st), "fm_fused_adagrad")
            mark("adagrad")
            self.global_step += 1
            return b
        if b["grad"] is None:
            b["grad"] = torch.empty(b["cap"], self.ld, dtype=torch.float32, device=self.device)
        self._forward(packed, b, mark)
        nuniq_dev = _off(b["status"], 4)
        check(lib.fm_loss_grad(ptr(b["pred"]), ptr(b["labels"]), ptr(weights), n_lines,
                               _lib.LOSS_KIND[self.loss_type], _off(b["status"], 6), ptr(b["pred_grad"]),
                               ptr(b["ws_loss"]), b["ws_loss"].numel(), st), "fm_loss_grad")
        mark("loss")
        # objective = loss + reg_score / batch_size with the CONFIGURED batch size (fm_model.py:345-347)
        check(lib.fm_grad(ptr(b["local"]), ptr(b["rows"]), self.ld, ptr(b["vals"]), ptr(b["poses"]), n_lines, k,
                          self.factor_lambda, self.bias_lambda, ptr(b["pred_grad"]), 1.0 / self.batch_size, None,
                          b["cap"], nuniq_dev, ptr(b["grad"]), st), "fm_grad")
        mark("grad")
        check(lib.fm_sparse_adagrad(ptr(self.table), ptr(self.accum), tstride, ptr(b["ori"]), b["cap"], nuniq_dev, k,
                                    ptr(b["grad"]), self.ld, self.learning_rate, ptr(b["status"]), st),
              "fm_sparse_adagrad")
        mark("adagrad")
        self.global_step += 1
        return b

    def train_step(self, lines, weights=None, path=None):
        """lines: list of str / PackedLines (host or device); weights: per-example floats or None.
        Returns the batch loss (python float); raises FmParseError like the op would."""
        packed, weights = self._to_device(lines, weights)
        b = self.train_step_async(packed, weights, path=path)
        host = b["status"].cpu()
        if int(host[0]) & 0xFFFFFFFFFFFFFFFF != _lib.NO_ERROR:
            self.global_step -= 1        # the guard voided the step: the reference's global_step does not advance either
        _raise(host, packed)
        return float(host[6:8].view(torch.float32)[0])

    def predict_async(self, packed: PackedLines, path="claim", parsed=None):
        """scores of a device-resident batch in b["pred"].  Default: parse, then the scorer reads
        the rows straight from the table by feature id (no unique, no gathered copy);
        path="composed" runs the reference's chain parse -> unique -> lookup -> fm_scorer.
        parsed: handle from parse_ahead(packed, sort=False) for THIS batch: its parse ran on the side stream, under
        the previous batch's scorer (the parser is instruction-bound, the scorer memory-bound); read the scores
        before the buffer set comes round again (two sets: before the parse_ahead after the next one)."""
        if parsed is not None:
            if path == "composed":
                raise ValueError("parse_ahead() feeds the default predict path only")
            b = parsed["b"]
            main = torch.cuda.current_stream(self.device)
            need = torch.cuda.Event(enable_timing=True)
            need.record(main)
            main.wait_event(parsed["done"])
            self._adapt_parse_grid(need, parsed["done"])
            self._bufs = b
        else:
            b = self._buffers(packed.n_lines, packed.nnz_bound())
        if path == "composed":
            self._forward(packed, b)
            return b
        lib, st = load(), stream()
        if parsed is None:
            self._parse_unique(packed, b, unique=False)
        check(lib.fm_fused_score_scatter(
            ptr(b["ids"]), None, ptr(b["vals"]), ptr(b["poses"]), None, None, packed.n_lines, self.factor_num,
            ptr(self.table), self.table.stride(0), self.factor_lambda, self.bias_lambda, 0, 0, ptr(b["pred"]),
            _off(b["status"], 6), _off(b["status"], 7), None, 4 * ((self.factor_num + 6) // 4), ptr(b["status"]),
            ptr(b["ws_fused"]), b["ws_fused"].numel(), st), "fm_fused_score_scatter")
        if parsed is not None:       # the buffer set may be refilled once the scorer is through with it
            b["free"] = torch.cuda.Event()
            b["free"].record(torch.cuda.current_stream(self.device))
        return b

    def predict(self, lines, path="claim"):
        """raw pred_score per line (no sigmoid), as tffm/fm_model.py:128-142"""
        packed, _ = self._to_device(lines, None)
        b = self.predict_async(packed, path)
        host = b["status"].cpu()
        _raise(host, packed)
        return b["pred"][:packed.n_lines].clone()

    def _to_device(self, lines, weights):
        packed = lines if isinstance(lines, PackedLines) else pack_lines(list(lines))
        if not packed.text.is_cuda:
            packed = packed.to(self.device, non_blocking=False)
        if weights is not None and not (isinstance(weights, torch.Tensor) and weights.is_cuda):
            weights = torch.as_tensor(weights, dtype=torch.float32).to(self.device)
        return packed, weights

    # ---- checkpoint layout of the reference: vocab_block_i[id // nb] for id % nb == i ----
    def blocks_state(self):
        """{'vocab_block_%d': f32[V//nb+1, K+1], 'vocab_block_%d/Adagrad': ...} on the host"""
        nb, v, k = self.vocabulary_block_num, self.vocabulary_size, self.factor_num
        rows = v // nb + 1
        out = {"global_step": self.global_step}
        for name, src in (("", self.table), ("/Adagrad", self.accum)):
            for i in range(nb):
                blk = torch.zeros(rows, k + 1, dtype=torch.float32)
                part = src[i::nb, :k + 1].cpu()
                blk[:part.shape[0]] = part
                out["vocab_block_%d%s" % (i, name)] = blk
        return out

    def load_blocks_state(self, state):
        """any mod partition: the number of blocks is the file's, not the config's (a sharded run of G GPUs
        writes G blocks)"""
        k = self.factor_num
        nb = sum(1 for key in state if key.startswith("vocab_block_") and not key.endswith("/Adagrad"))
        for name, dst in (("", self.table), ("/Adagrad", self.accum)):
            for i in range(nb):
                n = dst[i::nb].shape[0]
                dst[i::nb, :k + 1] = state["vocab_block_%d%s" % (i, name)][:n].to(self.device)
        self.global_step = int(state.get("global_step", 0))

    def _blocks_snapshot(self):
        """device copy of the table and the accumulator, already in block order: f32[2, nb, V//nb+1, K+1];
        None when HBM has no room for it"""
        nb, v, k = self.vocabulary_block_num, self.vocabulary_size, self.factor_num
        rows = v // nb + 1
        need = 2 * 2 * nb * rows * (k + 1) * 4          # the copy + the transposed copy
        if torch.cuda.mem_get_info(self.device)[0] < need + (1 << 30):
            return None
        dense = torch.zeros(2, rows * nb, k + 1, dtype=torch.float32, device=self.device)
        dense[0, :v] = self.table[:, :k + 1]
        dense[1, :v] = self.accum[:, :k + 1]
        return dense.view(2, rows, nb, k + 1).transpose(1, 2).contiguous()   # [which, id % nb, id // nb, :]

    @staticmethod
    def _blocks_dict(host, global_step):
        out = {"global_step": global_step}
        for w, name in enumerate(("", "/Adagrad")):
            for i in range(host.shape[1]):
                out["vocab_block_%d%s" % (i, name)] = host[w, i]       # views of one storage: written once
        return out

    def _writer(self):
        if getattr(self, "_ckpt_writer", None) is None:
            self._ckpt_writer = CheckpointWriter(self.device)
        return self._ckpt_writer

    def save(self, path, wait=True, skip_if_busy=False):
        """Writes the checkpoint (reference variable layout, see blocks_state).  wait=False: the table is
        snapshotted on the device now (one strided device copy) and the D2H copy + file write run on a
        background thread while training goes on (CheckpointWriter).  One save is in flight at a time: a new
        one waits for it, or -- skip_if_busy, for periodic saves that must never hold training up -- is
        dropped (returns False)."""
        w = self._writer()
        if skip_if_busy and w.busy():
            return False
        w.wait()
        snap = self._blocks_snapshot()
        if snap is None:                                 # no room for a snapshot: block by block, synchronously
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            torch.save(self.blocks_state(), path + ".tmp")
            os.replace(path + ".tmp", path)
            return True
        step = self.global_step
        w.write(path, snap, lambda host: self._blocks_dict(host, step), wait=wait)
        return True

    def prepare_save(self):
        self._writer().prepare()

    def wait_for_save(self):
        self._writer().wait()

    def restore(self, path):
        """a file written by Model.save, or the sharded set of a multi-GPU run at the same path"""
        self.load_blocks_state(load_checkpoint(path))


def _off(t, index):
    """device pointer to element `index` of a tensor"""
    import ctypes
    return ctypes.c_void_p(t.data_ptr() + index * t.element_size())


def _raise(status_host, packed):
    first = int(status_host[0]) & 0xFFFFFFFFFFFFFFFF
    if first != _lib.NO_ERROR:
        kind, line_no = first & 7, first >> 3
        raise FmParseError(kind, line_no, packed.line(line_no) if line_no < packed.n_lines else "")
