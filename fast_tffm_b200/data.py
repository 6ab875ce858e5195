"""Line reader + shuffling batcher: what the reference's input pipeline does between the
files and the parser (tffm/fm_model.py:34-126), without the TF queue machinery.

Semantics kept: data and weight files are iterated in the same (per-epoch shuffled) order,
`epoch_num` passes, lines pass through a shuffle buffer that holds at least 3 batches before it
starts emitting and at most 4.5 (min_after_dequeue / capacity, :44-45), batches have
`batch_size` lines and the last one may be smaller (allow_smaller_final_batch, :57).
`shuffle=False` gives files in sorted order and lines in file order (parity runs).
"""
import random


def _read_lines(path):
    with open(path, "rb") as f:
        data = f.read()
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


def batches(train_files, weight_files, batch_size, num_epochs, shuffle=True, seed=None):
    """yields (lines: list[bytes], weights: list[float] | None)"""
    rnd = random.Random(seed)
    have_w = bool(weight_files)
    if have_w and len(weight_files) != len(train_files):
        raise ValueError('The numbers of train files and weight files do not match.')
    buf_lines, buf_w = [], []
    low, high = 3 * batch_size, int(3 * batch_size + 1.5 * batch_size)

    def emit():
        n = min(batch_size, len(buf_lines))
        if shuffle:  # draw n random elements (RandomShuffleQueue dequeue_many)
            for i in range(n):
                j = rnd.randrange(i, len(buf_lines))
                buf_lines[i], buf_lines[j] = buf_lines[j], buf_lines[i]
                if have_w:
                    buf_w[i], buf_w[j] = buf_w[j], buf_w[i]
        out = (buf_lines[:n], buf_w[:n] if have_w else None)
        del buf_lines[:n]
        if have_w:
            del buf_w[:n]
        return out

    for _ in range(num_epochs):
        order = list(range(len(train_files)))
        if shuffle:
            rnd.shuffle(order)
        for fi in order:
            lines = _read_lines(train_files[fi])
            ws = [float(x) for x in _read_lines(weight_files[fi])] if have_w else None  # tf.string_to_number, :60
            if have_w and len(ws) != len(lines):
                raise ValueError("%s and %s differ in line count" % (train_files[fi], weight_files[fi]))
            pos = 0
            while pos < len(lines):
                take = min(len(lines) - pos, max(high - len(buf_lines), 0)) if shuffle else len(lines) - pos
                if take > 0:
                    buf_lines.extend(lines[pos:pos + take])
                    if have_w:
                        buf_w.extend(ws[pos:pos + take])
                    pos += take
                threshold = low + batch_size if shuffle else batch_size
                while len(buf_lines) >= threshold:
                    yield emit()
    while buf_lines:
        yield emit()


class DeviceDataset:
    """Training files resident in HBM; batches are index selections, the text is never moved.

    The B200 form of the reference's reader + shuffle queue (tffm/fm_model.py:34-126): every
    file is copied to the device once (pinned host buffer -> H2D), `fm_split_lines` finds the
    line starts on the device, and a batch is `PackedLines.select(indices)` -- two gathered
    index arrays over the resident text -- fed straight to the parser.  Shuffling is a device
    `randperm` over all lines of the epoch (the reference shuffles within a 3-4.5 batch window;
    a full permutation is the limit of that).  Weight files go through the same parser as
    label-only lines, so they are read with strtof exactly like `tf.string_to_number` would
    (tffm/fm_model.py:60).  Smaller final batch allowed (:57).
    """

    def __init__(self, train_files, weight_files=(), device=None):
        import numpy as np
        import torch
        from . import fm_ops
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if weight_files and len(weight_files) != len(train_files):
            raise ValueError('The numbers of train files and weight files do not match.')
        blobs, counts = [], []
        for path in train_files:
            with open(path, "rb") as f:
                data = f.read()
            if data and not data.endswith(b"\n"):
                data += b"\n"
            blobs.append(data)
            counts.append(data.count(b"\n"))
        total = sum((len(b) + 15) // 16 * 16 for b in blobs)
        host = torch.zeros(total + 16, dtype=torch.uint8).pin_memory()
        view = host.numpy()
        bases, pos = [], 0
        for b in blobs:
            view[pos:pos + len(b)] = np.frombuffer(b, np.uint8)
            bases.append(pos)
            pos += (len(b) + 15) // 16 * 16
        self.text = host.to(self.device, non_blocking=True)
        begins, ends = [], []
        for base, b, n in zip(bases, blobs, counts):
            off = fm_ops.fm_split_lines(self.text[base:], len(b), n) + base
            begins.append(off[:-1])
            ends.append(off[1:])
        self.begins = torch.cat(begins) if begins else torch.zeros(0, dtype=torch.int64, device=self.device)
        self.ends = torch.cat(ends) if ends else torch.zeros(0, dtype=torch.int64, device=self.device)
        longest = int((self.ends - self.begins).max()) if self.begins.numel() else 0
        self.lines = fm_ops.PackedLines(self.text, begins=self.begins, ends=self.ends, max_line_bytes=longest)
        self.weights = None
        if weight_files:
            parts = []
            for path, n in zip(weight_files, counts):
                with open(path, "rb") as f:
                    wlines = f.read().split(b"\n")
                if wlines and wlines[-1] == b"":
                    wlines.pop()
                if len(wlines) != n:        # as the host batcher: a misaligned weight file is an error, not a truncation
                    raise ValueError("%s has %d lines, its data file %d" % (path, len(wlines), n))
                wl = fm_ops.pack_lines(wlines)
                labels, sizes, _, _ = fm_ops.fm_parser(wl.to(self.device, non_blocking=False), 1)
                if int(sizes.sum()) != 0 or labels.numel() != n:
                    raise ValueError("%s: expected %d lines holding one number each" % (path, n))
                parts.append(labels)
            self.weights = torch.cat(parts)

    def __len__(self):
        return self.begins.numel()

    def batches(self, batch_size, num_epochs=1, shuffle=True, seed=None):
        """yields (PackedLines on the device, weights tensor or None)"""
        import torch
        gen = torch.Generator(device=self.device)
        if seed is not None:
            gen.manual_seed(int(seed))
        n = len(self)
        for _ in range(num_epochs):
            order = torch.randperm(n, device=self.device, generator=gen) if shuffle else \
                torch.arange(n, device=self.device)
            for i in range(0, n, batch_size):
                idx = order[i:i + batch_size]
                yield self.lines.select(idx), (None if self.weights is None else self.weights[idx])


class SegmentedDataset:
    """DeviceDataset for training files that do not fit in HBM (or should not sit there): the files are
    streamed through the device in segments of about `segment_bytes` of text.

    A segment is read straight into pinned host memory (cut at its last newline; the cut-off tail opens
    the next segment; files follow each other inside a segment), copied into one half of a device arena,
    split into lines there (`fm_split_lines`) and shuffled by a device `randperm`; batches are index
    selections over the arena, exactly `batch_size` lines each -- the lines a segment leaves over open the
    next segment's first batch, which is why the arena has two halves -- and only the very last batch of
    the run may be smaller (tffm/fm_model.py:57).  The shuffle window is a segment (>= the reference's
    3 - 4.5 batches whenever a segment holds that many lines).  Weight files are read per file through
    the parser, as in DeviceDataset.
    """

    def __init__(self, train_files, weight_files=(), device=None, segment_bytes=256 << 20):
        import torch
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if weight_files and len(weight_files) != len(train_files):
            raise ValueError('The numbers of train files and weight files do not match.')
        self.train_files, self.weight_files = list(train_files), list(weight_files)
        self.segment_bytes = (int(segment_bytes) + 15) // 16 * 16
        self._arena = self._pinned = None

    def _file_weights(self, path):
        import torch
        from . import fm_ops
        with open(path, "rb") as f:
            lines = f.read().split(b"\n")
        if lines and lines[-1] == b"":
            lines.pop()
        labels, sizes, _, _ = fm_ops.fm_parser(fm_ops.pack_lines(lines).to(self.device, non_blocking=False), 1)
        if int(sizes.sum()) != 0:
            raise ValueError("%s: expected one number per line" % path)
        return labels

    def _segments(self, order):
        """yields (n_bytes, [(file index, complete lines of that file in this segment are counted later)]):
        fills self._pinned with whole lines of the files in `order`, one segment at a time"""
        import numpy as np
        cap = self.segment_bytes
        view = self._pinned.numpy()
        fill, parts = 0, []          # parts: [file index, bytes of that file in this segment]
        for fi in order:
            with open(self.train_files[fi], "rb") as f:
                last = 10
                while True:
                    if fill == cap:   # full: cut at the last newline, the tail opens the next segment
                        tail = view[max(0, cap - (1 << 20)):cap].tobytes()
                        pos = tail.rfind(b"\n")
                        if pos < 0:
                            tail = view[:cap].tobytes()
                            pos = tail.rfind(b"\n")
                            if pos < 0:
                                raise ValueError("a line of %s is longer than the segment (%d bytes)"
                                                 % (self.train_files[fi], cap))
                        keep = cap - len(tail) + pos + 1
                        carry = view[keep:cap].copy()
                        parts[-1][1] -= cap - keep
                        yield keep, parts
                        view[:carry.size] = carry
                        fill, parts = carry.size, [[fi, carry.size]]
                    got = f.readinto(memoryview(view[fill:cap]))
                    if not got:
                        break
                    last = int(view[fill + got - 1])
                    if parts and parts[-1][0] == fi:
                        parts[-1][1] += got
                    else:
                        parts.append([fi, got])
                    fill += got
                if last != 10:        # a final line without newline gets one (room: the arena half is cap + 16)
                    if fill == cap:
                        raise ValueError("%s: the last line does not end with a newline at a segment boundary; "
                                         "end the file with a newline" % self.train_files[fi])
                    view[fill] = 10
                    parts[-1][1] += 1
                    fill += 1
        if fill:
            yield fill, parts

    def batches(self, batch_size, num_epochs=1, shuffle=True, seed=None):
        """yields (PackedLines on the device, weights tensor or None)"""
        import torch
        from . import fm_ops
        dev, cap = self.device, self.segment_bytes
        if self._arena is None:
            self._arena = torch.zeros(2 * (cap + 16), dtype=torch.uint8, device=dev)
            self._pinned = torch.zeros(cap + 16, dtype=torch.uint8).pin_memory()
        arena = self._arena
        gen = torch.Generator(device=dev)
        if seed is not None:
            gen.manual_seed(int(seed))
        rnd = random.Random(seed)
        i64 = dict(dtype=torch.int64, device=dev)
        pend_b, pend_e = torch.zeros(0, **i64), torch.zeros(0, **i64)
        pend_w = torch.zeros(0, dtype=torch.float32, device=dev) if self.weight_files else None
        wcache, wpos = {}, {}
        seg_no = 0
        longest = 0       # longest line seen so far (lines carried over from the previous segment included)
        for _ in range(num_epochs):
            order = list(range(len(self.train_files)))
            if shuffle:
                rnd.shuffle(order)
            wpos = {fi: 0 for fi in order}
            for n_bytes, parts in self._segments(order):
                base = (seg_no % 2) * (cap + 16)
                seg_no += 1
                half = arena[base:base + cap + 16]
                padded = (n_bytes + 15) // 16 * 16
                # whatever still reads this half (a consumer may parse on a side stream) is through first
                torch.cuda.synchronize(dev)
                half[:n_bytes].copy_(self._pinned[:n_bytes], non_blocking=True)
                half[n_bytes:padded + 16].zero_()
                n_lines = int((half[:n_bytes] == 10).sum())      # also: the copy is done, the pinned buffer is free
                off = fm_ops.fm_split_lines(half, n_bytes, n_lines)
                begins, ends = off[:-1] + base, off[1:] + base
                longest = max(longest, int((off[1:] - off[:-1]).max())) if n_lines else longest
                w = None
                if self.weight_files:
                    # lines of each file in this segment: newlines inside that file's byte range
                    ws, at = [], 0
                    for fi, nb in parts:
                        cnt = int(torch.searchsorted(off, at + nb, right=True) - torch.searchsorted(off, at, right=True))
                        if fi not in wcache:
                            wcache.clear()          # one file's weights at a time
                            wcache[fi] = self._file_weights(self.weight_files[fi])
                        if wpos[fi] + cnt > wcache[fi].numel():
                            raise ValueError("%s has fewer lines than %s" % (self.weight_files[fi], self.train_files[fi]))
                        ws.append(wcache[fi][wpos[fi]:wpos[fi] + cnt])
                        wpos[fi] += cnt
                        at += nb
                    w = torch.cat(ws)
                if shuffle:
                    perm = torch.randperm(n_lines, device=dev, generator=gen)
                    begins, ends = begins[perm], ends[perm]
                    w = None if w is None else w[perm]
                if pend_b.numel() and pend_b.numel() + n_lines < batch_size:
                    raise ValueError("a segment of %d bytes holds fewer than batch_size = %d lines: raise segment_bytes"
                                     % (cap, batch_size))
                all_b, all_e = torch.cat([pend_b, begins]), torch.cat([pend_e, ends])
                all_w = None if w is None else torch.cat([pend_w, w])
                k = 0
                while all_b.numel() - k >= batch_size:
                    sl = slice(k, k + batch_size)
                    yield (fm_ops.PackedLines(arena, begins=all_b[sl].contiguous(), ends=all_e[sl].contiguous(),
                                              max_line_bytes=longest),
                           None if all_w is None else all_w[sl].contiguous())
                    k += batch_size
                pend_b, pend_e = all_b[k:], all_e[k:]
                pend_w = None if all_w is None else all_w[k:]
        if pend_b.numel():
            yield (fm_ops.PackedLines(arena, begins=pend_b.contiguous(), ends=pend_e.contiguous(),
                                      max_line_bytes=longest),
                   None if pend_w is None else pend_w.contiguous())


def _bind_by_nvml(bus):
    """NVML's CPU affinity of the GPU with PCI address `bus` (nvidia-smi topo's "CPU Affinity" column)"""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + bit for w, m in enumerate(mask) for bit in range(64) if (int(m) >> bit) & 1}
    except Exception as e:      # no NVML, or it does not know either
        return "gpu %s: numa node not reported (sysfs -1; NVML: %s)" % (bus, e)
    allowed = os.sched_getaffinity(0) & cpus
    if not allowed or allowed == os.sched_getaffinity(0):
        return "gpu %s: NVML affinity covers %d cpus, nothing to narrow" % (bus, len(cpus))
    os.sched_setaffinity(0, allowed)
    return "gpu %s: bound to %d cpus (NVML cpu affinity)" % (bus, len(allowed))


def bind_host_to_gpu(index=None):
    """Pin the calling thread (and the threads it starts later) to the CPUs of the NUMA node the GPU hangs
    off, so that pinned staging buffers allocated afterwards are local to the GPU's PCIe root and the
    H2D copies of one process per GPU do not all cross the socket interconnect.  Call it before the
    first pinned allocation.  Returns a one-line description of what was done (nothing is done, and
    the reason returned, when the topology is not visible -- e.g. inside some containers)."""
    import os
    import torch
    index = torch.cuda.current_device() if index is None else index
    try:
        p = torch.cuda.get_device_properties(index)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as f:
            node = int(f.read())
        if node < 0:      # sysfs does not say (common in containers): ask the driver for the GPU's ideal CPUs
            return _bind_by_nvml(bus)
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if not allowed:
            return "gpu %s: none of numa node %d's cpus are available to this process" % (bus, node)
        os.sched_setaffinity(0, allowed)
        return "gpu %s: bound to %d cpus of numa node %d" % (bus, len(allowed), node)
    except (OSError, ValueError, AttributeError, RuntimeError, AssertionError) as e:
        return "not bound (%s)" % e
