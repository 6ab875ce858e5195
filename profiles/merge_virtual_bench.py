"""The record exchange's sender and owner kernels timed on ONE GPU with G virtual ranks (all shards and inboxes in
local memory): what fm_peer_push_records and fm_peer_merge_update cost before NVLink enters the picture.
    python profiles/merge_virtual_bench.py [G] [reps]"""
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

g = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
lib, dev = _lib.load(), torch.device("cuda", 0)
k, b, vocab = CFG["factor_num"], CFG["batch"], CFG["vocab"]
rows = vocab // g + 1
part = (k + 1 + 3) // 4 * 4
block = (3 * part + 31) // 32 * 32
at = lambda t, i: ctypes.c_void_p(t.data_ptr() + i * t.element_size())
stores = [torch.zeros(rows, block, device=dev) for _ in range(g)]
for s in stores:
    s[:, :k + 1].uniform_(-0.01, 0.01)
    s[:, part:2 * part] = 0.1
cap = b * 39
icap = cap * 5 // 4 // g * 2 + 64          # a segment holds about cap / g positions
inboxes = [torch.zeros(lib.fm_merge_inbox_bytes(icap, rows, k, g), dtype=torch.uint8, device=dev) for _ in range(g)]
arr = ctypes.c_void_p * g
table_bases, inbox_bases = arr(*[s.data_ptr() for s in stores]), arr(*[x.data_ptr() for x in inboxes])
rec_ld = int(lib.fm_sorted_record_ld(k))
ranks = []
for r in range(g):
    text, off, nnz = bench_data.criteo_batch(b, 1000 + r)
    packed = fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off)).to(dev, False)
    labels, sizes, ids, vals = fm_ops.fm_parser(packed, vocab, True)
    poses = torch.zeros(b + 1, dtype=torch.int32, device=dev)
    poses[1:] = torch.cumsum(sizes, 0)
    status = torch.zeros(8, dtype=torch.int64, device=dev)
    status[0], status[1] = -1, nnz
    d = dict(labels=labels, ids=ids.contiguous(), vals=vals.contiguous(), poses=poses, status=status,
             skeys=torch.empty(cap, dtype=torch.int32, device=dev), spairs=torch.empty(cap, 2, dtype=torch.int32, device=dev),
             ws_sort=_lib.workspace(lib.fm_sort_workspace(cap), dev), rec=torch.empty(b, rec_ld, device=dev),
             pred=torch.empty(b, device=dev), ws=_lib.workspace(lib.fm_fused_workspace(b), dev),
             ws_push=_lib.workspace(lib.fm_merge_push_workspace(g), dev))
    ranks.append(d)


def timed(fn, n):
    fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / n


def sort(d):
    check(lib.fm_sort_by_row(ptr(d["ids"]), ptr(d["vals"]), ptr(d["poses"]), b, cap, at(d["status"], 1), vocab, g,
                             ptr(d["skeys"]), ptr(d["spairs"]), at(d["status"], 3), ptr(d["status"]), ptr(d["ws_sort"]),
                             d["ws_sort"].numel(), 0, stream()), "sort")


def score(d):
    check(lib.fm_peer_sorted_score(ptr(d["ids"]), ptr(d["vals"]), ptr(d["poses"]), ptr(d["labels"]), None, b, k, g,
                                   table_bases, block, 1e-5, 1e-5, 0, b * g, ptr(d["pred"]), at(d["status"], 6),
                                   at(d["status"], 7), ptr(d["rec"]), rec_ld, ptr(d["status"]), ptr(d["ws"]),
                                   d["ws"].numel(), stream()), "score")


def push(r):
    d = ranks[r]
    check(lib.fm_peer_push_records(ptr(d["skeys"]), ptr(d["spairs"]), at(d["status"], 3), cap, ptr(d["rec"]), rec_ld, k, r, g,
                                   vocab, rows, inbox_bases, icap, ptr(d["status"]), ptr(d["ws_push"]), d["ws_push"].numel(),
                                   stream()), "push")


flag = torch.tensor([-1], dtype=torch.int64, device=dev)


def merge(r):
    check(lib.fm_peer_merge_update(ptr(stores[r]), at(stores[r], part), block, rows, ptr(inboxes[r]), icap, k, g, 0.01, 1e-5,
                                   1e-5, 1.0 / (b * g), ptr(flag), stream()), "merge")


for r in range(g):
    sort(ranks[r]); score(ranks[r]); push(r)
torch.cuda.synchronize()
assert all(int(d["status"][0]) == -1 for d in ranks)
out = {"virtual_ranks": g, "sort_ms": timed(lambda: sort(ranks[0]), reps), "score_local_ms": timed(lambda: score(ranks[0]), reps),
       "push_local_ms": timed(lambda: push(0), reps), "merge_ms": timed(lambda: merge(0), reps),
       "tile_rows": int(lib.fm_merge_tile_rows(g)), "shard_rows": rows}
print(json.dumps(out))
