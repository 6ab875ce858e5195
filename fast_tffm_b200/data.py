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
        self.lines = fm_ops.PackedLines(self.text, begins=self.begins, ends=self.ends)
        self.weights = None
        if weight_files:
            parts = []
            for path, n in zip(weight_files, counts):
                with open(path, "rb") as f:
                    wl = fm_ops.pack_lines(f.read().split(b"\n")[:n])
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
        if node < 0:
            return "gpu %s: numa node not reported" % bus
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
    except (OSError, ValueError, AttributeError) as e:
        return "not bound (%s)" % e
