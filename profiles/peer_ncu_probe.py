"""One process, two GPUs: fm_peer_score_scatter of one cfg2 batch on GPU 0 over a table whose rows id % 2 == 1 live on
GPU 1, so that ncu (which must not wrap a multi-rank command) can count what crosses NVLink for the kernel as rank 0 of
a 2-GPU run sees it: nvltx / nvlrx bytes, user against protocol.

    python profiles/peer_ncu_probe.py [train|score]
    ncu --metrics nvltx__bytes.sum,nvlrx__bytes.sum,... -k regex:fused_score_scatter python profiles/peer_ncu_probe.py
"""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_data  # noqa: E402
from bench import CFG  # noqa: E402
from fast_tffm_b200 import _lib, fm_ops  # noqa: E402
from fast_tffm_b200._lib import check, ptr, stream  # noqa: E402

train = (sys.argv[1] if len(sys.argv) > 1 else "train") == "train"
assert torch.cuda.device_count() >= 2 and torch.cuda.can_device_access_peer(0, 1)
lib = _lib.load()
k, b, world = CFG["factor_num"], CFG["batch"], 2
rows = CFG["vocab"] // world + 1
part = max((k + 1 + 3) // 4 * 4, int(lib.fm_fused_gacc_ld(k)))
block = (3 * part + 31) // 32 * 32                       # [w | h | g] per row, as PeerShardedModel lays it out
n_store, n_bits = rows * block, ((rows + 31) // 32 + 7) // 8 * 8
shards = []
for d in range(world):
    buf = torch.zeros(n_store + n_bits, dtype=torch.float32, device="cuda:%d" % d)
    buf[:n_store].view(rows, block)[:, :k + 1].uniform_(-0.01, 0.01)
    shards.append(buf)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
rt = ctypes.CDLL("libcudart.so.12")                     # kernels of GPU 0 dereference GPU 1's memory: peer access, both ways
for me, peer in ((0, 1), (1, 0)):
    assert rt.cudaSetDevice(me) == 0
    rc = rt.cudaDeviceEnablePeerAccess(peer, 0)
    assert rc in (0, 704), rc                            # 704 = already enabled
    rt.cudaGetLastError()
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
arr = ctypes.c_void_p * world
bases = [s.data_ptr() for s in shards]
table_bases, gtab_bases = arr(*bases), arr(*[p + 4 * 2 * part for p in bases])
touched_bases = arr(*[p + 4 * n_store for p in bases])

text, off, nnz = bench_data.criteo_batch(b, 1001)
packed = fm_ops.PackedLines(torch.from_numpy(bench_data.pad16(text)), torch.from_numpy(off)).to(dev, False)
labels, sizes, ids, vals = fm_ops.fm_parser(packed, CFG["vocab"], True)
poses = torch.zeros(b + 1, dtype=torch.int32, device=dev)
poses[1:] = torch.cumsum(sizes, 0)
pred = torch.empty(b, device=dev)
scal = torch.zeros(2, device=dev)
ws = _lib.workspace(lib.fm_fused_workspace(b), dev)
remote = int((ids % world != 0).sum())
for _ in range(3):
    check(lib.fm_peer_score_scatter(ptr(ids), ptr(vals), ptr(poses), ptr(labels) if train else None, None, b, k, world,
                                    table_bases, block, gtab_bases if train else None, block,
                                    touched_bases if train else None, 1e-5, 1e-5, 0, b * world, ptr(pred), ptr(scal),
                                    ctypes.c_void_p(scal.data_ptr() + 4), None, ptr(ws), ws.numel(), stream()), "peer")
torch.cuda.synchronize(0)
print("ok", "train" if train else "score", "nnz", nnz, "remote occurrences", remote, "loss", float(scal[0]))
