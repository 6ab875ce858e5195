"""CPU stand-in for CudaShardKernels, for tests only: numpy + the oracle on CPU tensors.

Lets the host-side orchestration of fast_tffm_b200.sharded.ShardedModel (routing,
all-to-all split sizes, owner-side combine order) run under gloo with world_size 2 on a box
without GPUs.  It is also an independent (numpy, float64) statement of the fused path's
factorised gradient sums.  Never imported by the product package.
"""
import numpy as np
import torch

from oracle import oracle as O


def _t(a, dtype=None):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype)


class OracleShardKernels:
    def parse_unique_bucket(self, packed, vocab, hashed, n_ranks):
        assert packed.offsets is not None
        text, off = packed.text.numpy(), packed.offsets.numpy()
        rc = O.parse(text, off, vocab, hashed, raise_on_error=False)
        if len(rc) == 2:   # (kind, line): report it the way the device status word does
            kind, line = rc
            info = np.zeros(3 + n_ranks, np.int64)
            info[2] = (line << 3) | kind
            z = lambda dt: torch.zeros(0, dtype=dt)
            return dict(labels=z(torch.float32), poses=z(torch.int32), vals=z(torch.float32), send_ids=z(torch.int64),
                        slot64=z(torch.int64), slot32=z(torch.int32), info=_t(info))
        labels, sizes, ids, vals = rc
        ori, local = O.unique(ids)
        owner = (ori % n_ranks).astype(np.int64)
        perm = np.argsort(owner, kind="stable")                 # stable counting sort by owner
        pos = np.empty(ori.size, np.int32)
        pos[perm] = np.arange(ori.size, dtype=np.int32)
        slot = pos[local]
        info = np.concatenate([[ids.size, ori.size, -1], np.bincount(owner, minlength=n_ranks)]).astype(np.int64)
        return dict(labels=_t(labels), poses=_t(O.poses(sizes)), vals=_t(vals), send_ids=_t(ori[perm]),
                    slot64=_t(slot.astype(np.int64)), slot32=_t(slot.astype(np.int32)), info=_t(info))

    def score_scatter(self, slot64, slot32, vals, poses, labels, weights, k, rows, lf, lb, loss_type, global_batch,
                      gacc):
        fids, x, ps = slot32.numpy(), vals.numpy().astype(np.float64), poses.numpy()
        params = rows.numpy()
        b = labels.numel()
        pred, reg = O.score(fids, params, vals.numpy(), ps, lf, lb, k=k)
        w = None if weights is None else weights.numpy()
        loss, pg = O.loss(pred, labels.numpy(), w, loss_type)
        scale = b / float(global_batch)
        loss, pg = loss * scale, pg.astype(np.float64) * scale
        ex = np.repeat(np.arange(b), np.diff(ps))
        v = params[fids, 1:k + 1].astype(np.float64)
        s = np.zeros((b, k))
        np.add.at(s, ex, x[:, None] * v)
        pgx = pg[ex] * x
        acc = np.zeros((gacc.shape[0], gacc.shape[1]))
        np.add.at(acc[:, 0], fids, pgx)
        np.add.at(acc[:, 1:k + 1], fids, pgx[:, None] * s[ex])
        np.add.at(acc[:, k + 1], fids, pgx * x)
        np.add.at(acc[:, k + 2], fids, 1.0)
        gacc += _t(acc, torch.float32)
        return _t(pred), _t(np.float32([loss, reg]))

    def score(self, slot32, rows, vals, poses, k):
        pred, _ = O.score(slot32.numpy(), rows.numpy(), vals.numpy(), poses.numpy(), 0.0, 0.0, k=k)
        return _t(pred)

    def local_rows(self, ids, n_ranks):
        return ids // n_ranks

    def lookup(self, table, rows, k):
        out = torch.zeros(rows.numel(), 4 * ((k + 4) // 4))
        out[:, :k + 1] = table[rows, :k + 1]
        return out

    def combine(self, src, index, dst):
        acc = dst.numpy().astype(np.float64)
        np.add.at(acc[:, :src.shape[1]], index.numpy(), src.numpy().astype(np.float64))
        dst.copy_(_t(acc, torch.float32))

    def adagrad_claim(self, table, accum, rows, k, gtab, lr, lf, lb, reg_grad):
        r = np.unique(rows.numpy())            # every listed row once, whoever lists it first
        a = gtab.numpy()[r].astype(np.float64)
        w = table.numpy()[r, :k + 1].astype(np.float64)
        cnt, d = a[:, k + 2], a[:, k + 1]
        g = np.empty_like(w)
        g[:, 0] = a[:, 0] + cnt * reg_grad * lb * w[:, 0]
        g[:, 1:] = w[:, 1:] * (cnt * reg_grad * lf - d)[:, None] + a[:, 1:k + 1]
        h = accum.numpy()[r, :k + 1].astype(np.float64) + g * g
        rows = _t(r)
        accum[rows, :k + 1] = _t(h, torch.float32)
        table[rows, :k + 1] = _t(w - lr * g / np.sqrt(h.astype(np.float32)), torch.float32)
        gtab[rows] = 0.0

    def gacc_ld(self, k):
        return (k + 3 + 3) // 4 * 4
