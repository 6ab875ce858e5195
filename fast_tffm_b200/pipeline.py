"""Host text batches -> training steps, three stages deep.

The reference feeds its parser from host memory (TF queues, tffm/fm_model.py:34-126).  Here a
batch that starts in host memory goes through three overlapped stages, each on its own stream:

    copy stream   H2D copy of batch s+2 into one of three device staging slots
    side stream   parse of batch s+1 (Model.parse_ahead: the parser is instruction-bound)
    main stream   table kernels of batch s (memory-bound), then a small D2H record of its outcome

Nothing blocks the host: a step's outcome is read later through the StepResult the generator
yields (loss() waits for that step's record only).  Works for the single-GPU Model and for
PeerShardedModel (every rank runs its own pipeline over its slice of the global batches; the
collectives inside the step keep the ranks in lock step).
"""
import collections

import torch

from . import _lib
from .fm_ops import FmParseError, PackedLines


def check_sync_word(f):
    from .sharded import check_sync_word as impl      # sharded imports torch.distributed: only when it is used
    impl(f)


class StepResult:
    """Outcome of one pipelined step.  loss() blocks until that step has finished on the device and
    raises FmParseError if its batch did not parse (the step was then void, as in train_step: the table is
    untouched and the model's global_step does not count it).  The steps enqueued behind it have run by then --
    the reference stops at the failing step; a pipelined caller learns of it up to RING steps later."""

    def __init__(self, index, n_lines, done, decode, on_void=None):
        self.index, self.n_lines = index, n_lines
        self._done, self._decode, self._value, self._on_void = done, decode, None, on_void

    def resolve(self):
        if self._decode is not None:
            self._done.synchronize()
            decode, self._decode = self._decode, None
            try:
                self._value = decode()
            except FmParseError as e:
                self._value = e
                if self._on_void is not None:
                    self._on_void()
        return self

    def loss(self):
        self.resolve()
        if isinstance(self._value, Exception):
            raise self._value
        return self._value


class HostBatchPipeline:
    """pipe = HostBatchPipeline(model); for res in pipe.run(batches): ...

    batches yields (PackedLines, weights or None[, global_batch]).  Host batches are copied in (pinned
    host tensors make the copies asynchronous); device-resident batches (DeviceDataset, SegmentedDataset)
    skip the copy stage and still get the parse of batch s+1 under the table kernels of batch s.
    `global_batch` (PeerShardedModel only) is the number of lines of the whole step over all ranks;
    default: this rank's lines x world size.
    """

    SLOTS = 3
    RING = 128

    def __init__(self, model):
        self.model = model
        self.device = model.device
        self.peer = hasattr(model, "world")                 # PeerShardedModel: (scalars, flag) instead of a buffer set
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._slots = [dict(text=None, offsets=None, weights=None, ready=None, free=None) for _ in range(self.SLOTS)]
        self._rec_status = torch.zeros(self.RING, 8, dtype=torch.int64).pin_memory()
        self._rec_scal = torch.zeros(self.RING, 2, dtype=torch.float32).pin_memory()
        self._ring = [None] * self.RING
        self._count = 0

    # ---- stage 1: host -> device staging slot (copy stream) ------------------------
    def _stage(self, index, host, weights):
        if host.text.is_cuda:       # already resident (DeviceDataset / SegmentedDataset batches): nothing to copy
            if weights is not None and not (isinstance(weights, torch.Tensor) and weights.is_cuda):
                weights = torch.as_tensor(weights, dtype=torch.float32).to(self.device)
            return dict(index=index, host=host, packed=host, weights=weights, slot=dict(ready=None, free=None))
        slot = self._slots[index % self.SLOTS]
        n_text, n_off = host.text.numel(), host.offsets.numel()
        grow = slot["text"] is None or slot["text"].numel() < n_text or slot["offsets"].numel() < n_off
        if grow:
            if slot["free"] is not None:
                slot["free"].synchronize()          # the step that used the old buffers is through
            cap_t = max(n_text, 0 if slot["text"] is None else slot["text"].numel())
            cap_o = max(n_off, 0 if slot["offsets"] is None else slot["offsets"].numel())
            slot["text"] = torch.zeros((cap_t + cap_t // 8 + 15) // 16 * 16 + 16, dtype=torch.uint8, device=self.device)
            slot["offsets"] = torch.zeros(cap_o + cap_o // 8 + 1, dtype=torch.int64, device=self.device)
            slot["weights"] = torch.zeros(slot["offsets"].numel(), dtype=torch.float32, device=self.device)
            self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))   # the fills above
        w_dev = None
        with torch.cuda.stream(self.copy_stream):
            if slot["free"] is not None:
                self.copy_stream.wait_event(slot["free"])
            slot["text"][:n_text].copy_(host.text, non_blocking=True)
            slot["offsets"][:n_off].copy_(host.offsets, non_blocking=True)
            if weights is not None:
                w_host = weights if isinstance(weights, torch.Tensor) else torch.as_tensor(weights, dtype=torch.float32)
                w_dev = slot["weights"][:w_host.numel()]
                w_dev.copy_(w_host, non_blocking=True)
            slot["ready"] = torch.cuda.Event()
            slot["ready"].record(self.copy_stream)
        packed = PackedLines(slot["text"][:n_text], slot["offsets"][:n_off])
        return dict(index=index, host=host, packed=packed, weights=w_dev, slot=slot)

    # ---- stage 3: the step itself + its record (main stream) -----------------------
    def _train(self, item, handle, global_batch):
        m, main = self.model, torch.cuda.current_stream(self.device)
        r = self._count % self.RING
        if self._ring[r] is not None:
            self._ring[r].resolve()                 # about to reuse its pinned record
        host, n_lines = item["host"], item["packed"].n_lines
        if self.peer:
            out, flag = m.train_step_async(item["packed"], item["weights"], global_batch=global_batch, parsed=handle)
            self._rec_status[r, 0:1].copy_(flag, non_blocking=True)
            self._rec_scal[r].copy_(out, non_blocking=True)
            rank = m.rank

            def decode(rec=self._rec_status[r], scal=self._rec_scal[r]):
                f = int(rec[0])
                check_sync_word(f)
                if f >= 0:
                    bad_rank, first = f >> 48, f & ((1 << 48) - 1)
                    line = host.line(first >> 3) if bad_rank == rank else "<rank %d, line %d>" % (bad_rank, first >> 3)
                    raise FmParseError(first & 7, first >> 3, line)
                return float(scal[0])
        else:
            b = m.train_step_async(item["packed"], item["weights"], parsed=handle)
            self._rec_status[r].copy_(b["status_snapshot"], non_blocking=True)

            def decode(rec=self._rec_status[r]):
                first = int(rec[0]) & 0xFFFFFFFFFFFFFFFF
                if first != _lib.NO_ERROR:
                    raise FmParseError(first & 7, first >> 3, host.line(first >> 3))
                return float(rec[6:8].view(torch.float32)[0])
        done = torch.cuda.Event()
        done.record(main)
        item["slot"]["free"] = done                 # staging slot (text, weights) reusable after this step
        res = StepResult(self._count, n_lines, done, decode, on_void=self._uncount_step)
        self._ring[r] = res
        self._count += 1
        return res

    def _uncount_step(self):
        self.model.global_step -= 1      # a void step is not a step (the save_steps cadence follows the reference's)

    def run(self, batches):
        """generator of StepResult, one per batch, yielded as soon as the step is enqueued"""
        it = iter(batches)
        staged, parsed = collections.deque(), collections.deque()
        state = dict(index=0, more=True)

        def pull():
            if not state["more"]:
                return
            try:
                entry = next(it)
            except StopIteration:
                state["more"] = False
                return
            host, weights = entry[0], entry[1]
            item = self._stage(state["index"], host, weights)
            item["global_batch"] = entry[2] if len(entry) > 2 else None
            state["index"] += 1
            staged.append(item)

        def parse_next():
            if staged:
                item = staged.popleft()
                handle = self.model.parse_ahead(item["packed"], after=item["slot"]["ready"])
                parsed.append((item, handle))

        pull()
        parse_next()
        pull()
        while parsed:
            item, handle = parsed.popleft()
            parse_next()        # batch s+1: its copy was issued one iteration ago
            pull()              # batch s+2
            yield self._train(item, handle, item["global_batch"])

    def drain(self):
        """resolve every outstanding result (e.g. before reading the table)"""
        for res in self._ring:
            if res is not None:
                res.resolve()
