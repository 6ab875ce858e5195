"""Synthetic Criteo-shaped text batches.

Bench / test infrastructure, not part of the product package.  Lines are built
with vectorised numpy so that a 65 536-line batch takes well under a second.

cfg2 line:  <label> then 39 tokens "%02d_%08x" (field, value); fields 0-12 carry an
explicit ":%.4f" value (log1p of one of 64 buckets), fields 13-38 are categorical
with the value omitted (implicit 1.0, cc/fm_parser_op.cc:102-104).  Label "0"/"1"
Bernoulli(0.25).  Every line has the same width, so a batch is a [B, W] byte matrix.
"""
import numpy as np

N_FIELDS, N_NUMERIC = 39, 13
_HEX = np.frombuffer(b"0123456789abcdef", np.uint8)


def _field_cardinalities():
    # categorical field cardinalities log-spaced 10 .. 1e7 (zipf mode)
    return np.round(np.logspace(1, 7, N_FIELDS - N_NUMERIC)).astype(np.int64)


_ZIPF_CDF = {}


def _zipf_bounded(rng, n, card, alpha=1.05):
    """Zipf(alpha) restricted to ranks 1..card (inverse CDF on the finite support) -> values 0..card-1"""
    cdf = _ZIPF_CDF.get(card)
    if cdf is None:
        w = np.arange(1, card + 1, dtype=np.float64) ** -alpha
        cdf = np.cumsum(w)
        cdf /= cdf[-1]
        _ZIPF_CDF[card] = cdf
    return np.minimum(np.searchsorted(cdf, rng.random(n)), card - 1).astype(np.uint64)


def criteo_batch(batch, seed, dist="uniform"):
    """-> (text u8[B*W] incl. one '\\n' per line, offsets i64[B+1], nnz)"""
    rng = np.random.Generator(np.random.PCG64(seed))
    b = batch
    tok_num = 1 + 11 + 1 + 6   # " ff_vvvvvvvv:d.dddd"
    tok_cat = 1 + 11           # " ff_vvvvvvvv"
    width = 1 + N_NUMERIC * tok_num + (N_FIELDS - N_NUMERIC) * tok_cat + 1
    m = np.empty((b, width), np.uint8)
    m[:, 0] = ord("0") + (rng.random(b) < 0.25)
    m[:, -1] = ord("\n")
    buckets = np.arange(64)
    num_strs = np.array([("%.4f" % np.log1p(x)).encode() for x in buckets])
    num_tab = np.frombuffer(b"".join(num_strs), np.uint8).reshape(64, 6)
    card = _field_cardinalities()
    col = 1
    for f in range(N_FIELDS):
        if f < N_NUMERIC:
            bucket = rng.integers(0, 64, b)
        if dist == "uniform":
            val = rng.integers(0, 1 << 32, b, dtype=np.uint64)
        elif dist == "zipf":
            c = 64 if f < N_NUMERIC else int(card[f - N_NUMERIC])
            val = _zipf_bounded(rng, b, c)
        else:
            raise ValueError(dist)
        if f < N_NUMERIC and dist == "uniform":
            pass  # value field independent of the bucket: ids stay uniform
        m[:, col] = ord(" ")
        m[:, col + 1] = ord("0") + f // 10
        m[:, col + 2] = ord("0") + f % 10
        m[:, col + 3] = ord("_")
        for d in range(8):
            m[:, col + 4 + d] = _HEX[((val >> np.uint64(4 * (7 - d))) & np.uint64(15)).astype(np.int64)]
        col += 12
        if f < N_NUMERIC:
            m[:, col] = ord(":")
            m[:, col + 1:col + 7] = num_tab[bucket]
            col += 7
    assert col == width - 1
    off = np.arange(b + 1, dtype=np.int64) * width
    return m.reshape(-1), off, b * N_FIELDS


def ragged_batch(batch, seed, max_nnz=512, vocab=1 << 24):
    """cfg4: 1..max_nnz integer-id features per line, every token "id:%.6g" ~ N(0,1),
    label "%.6g" ~ N(0,1); plus per-line weights "%.3f" ~ U(0.5, 2).
    Returns (lines: list[bytes], weights: float32[B])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    lines = []
    for _ in range(batch):
        n = int(rng.integers(1, max_nnz + 1))
        ids = rng.integers(0, vocab, n)
        vals = rng.normal(size=n)
        toks = " ".join("%d:%.6g" % (i, v) for i, v in zip(ids, vals))
        lines.append(("%.6g %s" % (rng.normal(), toks)).encode())
    weights = np.array([float("%.3f" % w) for w in rng.uniform(0.5, 2.0, batch)], np.float32)
    return lines, weights


def pad16(text):
    """pad a u8 array to the 16-byte granularity + slack the parser's loads need"""
    out = np.zeros((text.size + 15) // 16 * 16 + 16, np.uint8)
    out[:text.size] = text
    return out
