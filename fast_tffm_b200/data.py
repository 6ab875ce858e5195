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
