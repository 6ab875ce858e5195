"""Single GPU: the claim path's two kernels as Model runs them (the update walks the N occurrences and claims rows)
against the sharded step's form run with ONE rank (the scorer also sets a bit per touched row; the bitmap becomes a
row list; the update walks the U distinct rows) and against the update walking the bitmap itself.  Question: which form pays on one GPU -- for uniform ids
(U = 0.93 N) and for skewed ones (U << N)?  (r2_claim_forms_probe.txt also has a third form, tried and removed.)
    python profiles/claim_bitmap_probe.py [uniform|zipf] [reps]"""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_data  # noqa: E402
from bench import CFG  # noqa: E402
from fast_tffm_b200 import _lib, fm_ops  # noqa: E402
from fast_tffm_b200._lib import check, ptr, stream  # noqa: E402

dist = sys.argv[1] if len(sys.argv) > 1 else "uniform"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
lib, dev = _lib.load(), torch.device("cuda", 0)
k, b, vocab = CFG["factor_num"], CFG["batch"], CFG["vocab"]
rows = vocab + 1
part = max((k + 1 + 3) // 4 * 4, int(lib.fm_fused_gacc_ld(k)))
block = (3 * part + 31) // 32 * 32
n_store, n_bits = rows * block, ((rows + 31) // 32 + 7) // 8 * 8
buf = torch.zeros(n_store + n_bits, dtype=torch.float32, device=dev)
store = buf[:n_store].view(rows, block)
store[:, :k + 1].uniform_(-0.01, 0.01)
store[:, part:2 * part] = 0.1
table, accum, gtab = store[:, :part], store[:, part:2 * part], store[:, 2 * part:3 * part]
touched = buf[n_store:]
arr = ctypes.c_void_p * 1
at = lambda t, i: ctypes.c_void_p(t.data_ptr() + i * t.element_size())
table_bases, gtab_bases, touched_bases = arr(table.data_ptr()), arr(gtab.data_ptr()), arr(touched.data_ptr())
batches = []
for i in range(4):
    text, off, nnz = bench_data.criteo_batch(b, 1001 + i, dist)
    packed = fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off)).to(dev, False)
    labels, sizes, ids, vals = fm_ops.fm_parser(packed, vocab, True)
    poses = torch.zeros(b + 1, dtype=torch.int32, device=dev)
    poses[1:] = torch.cumsum(sizes, 0)
    batches.append(dict(labels=labels, ids=ids.contiguous(), vals=vals.contiguous(), poses=poses, nnz=nnz,
                        nnz_dev=torch.tensor([nnz], dtype=torch.int64, device=dev), uniq=int(torch.unique(ids).numel())))
pred, scal = torch.empty(b, device=dev), torch.zeros(4, device=dev)
ws = _lib.workspace(lib.fm_fused_workspace(b), dev)
row_list = torch.empty(rows, dtype=torch.int64, device=dev)
n_rows = torch.zeros(1, dtype=torch.int64, device=dev)
ws_touched = _lib.workspace(lib.fm_peer_touched_workspace(rows), dev)
ok = torch.tensor([-1], dtype=torch.int64, device=dev)      # a guard word that says "no error"


def occurrences(d):
    check(lib.fm_fused_score_scatter(ptr(d["ids"]), None, ptr(d["vals"]), ptr(d["poses"]), ptr(d["labels"]), None, b, k,
                                     ptr(table), block, 1e-5, 1e-5, 0, 0, ptr(pred), ptr(scal), at(scal, 2), ptr(gtab),
                                     block, None, ptr(ws), ws.numel(), stream()), "score_scatter")
    yield "score_scatter"
    check(lib.fm_fused_adagrad_claim(ptr(table), ptr(accum), block, ptr(d["ids"]), d["ids"].numel(), ptr(d["nnz_dev"]), k,
                                     ptr(gtab), block, 0.01, 1e-5, 1e-5, 1.0 / b, None, 0, stream()), "claim")
    yield "update"


def row_listed(d):
    check(lib.fm_peer_score_scatter(ptr(d["ids"]), ptr(d["vals"]), ptr(d["poses"]), ptr(d["labels"]), None, b, k, 1,
                                    table_bases, block, gtab_bases, block, touched_bases, 1e-5, 1e-5, 0, b, ptr(pred),
                                    ptr(scal), at(scal, 2), None, ptr(ws), ws.numel(), stream()), "peer_score_scatter")
    yield "score_scatter"
    check(lib.fm_peer_touched_rows(ptr(touched), rows, ptr(row_list), row_list.numel(), ptr(n_rows), ptr(ws_touched),
                                   ws_touched.numel(), stream()), "touched_rows")
    yield "touched_rows"
    check(lib.fm_fused_adagrad_claim(ptr(table), ptr(accum), block, ptr(row_list), row_list.numel(), ptr(n_rows), k,
                                     ptr(gtab), block, 0.01, 1e-5, 1e-5, 1.0 / b, None, 0, stream()), "claim")
    yield "update"


def bitmap_walk(d):
    check(lib.fm_peer_score_scatter(ptr(d["ids"]), ptr(d["vals"]), ptr(d["poses"]), ptr(d["labels"]), None, b, k, 1,
                                    table_bases, block, gtab_bases, block, touched_bases, 1e-5, 1e-5, 0, b, ptr(pred),
                                    ptr(scal), at(scal, 2), None, ptr(ws), ws.numel(), stream()), "peer_score_scatter")
    yield "score_scatter"
    check(lib.fm_fused_adagrad_bitmap(ptr(table), ptr(accum), block, ptr(touched), rows, k, ptr(gtab), block, 0.01, 1e-5,
                                      1e-5, 1.0 / b, None, 0, None, stream()), "adagrad_bitmap")
    yield "update"


def timed(form):
    for d in batches:                      # warm-up
        for _ in form(d):
            pass
    torch.cuda.synchronize()
    total, n = {}, 0
    for r in range(reps):
        d = batches[r % len(batches)]
        e = [torch.cuda.Event(enable_timing=True)]
        e[0].record()
        names = []
        for name in form(d):
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            e.append(ev)
            names.append(name)
        torch.cuda.synchronize()
        for i, name in enumerate(names):
            total[name] = total.get(name, 0.0) + e[i].elapsed_time(e[i + 1])
        n += 1
    return {k_: round(v / n, 4) for k_, v in total.items()}


out = {"ids": dist, "nnz": batches[0]["nnz"], "distinct_rows": batches[0]["uniq"],
       "walk_occurrences": timed(occurrences), "walk_row_list": timed(row_listed), "walk_bitmap": timed(bitmap_walk)}
for f in ("walk_occurrences", "walk_row_list", "walk_bitmap"):
    out[f]["sum"] = round(sum(out[f].values()), 4)
print(json.dumps(out))
