"""The record exchange of the sharded step (fm_merge.cu) with G virtual ranks on ONE GPU: every rank's shard, inbox and
batch slice are separate buffers on the same device, the kernels are called through the C ABI in the order a real
run calls them (sort -> pull-score -> push -> [barrier] -> merge), and the result must equal the single-process
oracle on the whole batch.  The 2+ GPU runs of the same path are in tests/test_gpu_sharded.py (mode "merge")."""
import ctypes

import numpy as np
import pytest
import torch

from fast_tffm_b200 import _lib, fm_ops
from fast_tffm_b200._lib import check, ptr, stream
from oracle import oracle as O
from parity_util import RTOL, close, pred_scale

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _at(t, index):
    return ctypes.c_void_p(t.data_ptr() + index * t.element_size())


def _virtual_step(lines, weights, flat, g, k, vocab, hashed, loss_type, lr, lf, lb, cfg_batch, inbox_cap=None):
    lib, dev = _lib.load(), torch.device(DEV)
    rows = vocab // g + 1
    part = (k + 1 + 3) // 4 * 4
    stores = []
    for r in range(g):
        st = torch.zeros(rows, 2 * part, dtype=torch.float32, device=dev)
        mine = torch.as_tensor(flat[r::g, :k + 1])
        st[:mine.shape[0], :k + 1] = mine.to(dev)
        st[:, part:] = 0.1
        stores.append(st)
    per = len(lines) // g
    parsed, cap = [], 0
    for r in range(g):
        labels, sizes, ids, vals = fm_ops.fm_parser(lines[r * per:(r + 1) * per], vocab, hashed)
        poses = torch.zeros(per + 1, dtype=torch.int32, device=dev)
        poses[1:] = torch.cumsum(sizes, 0)
        parsed.append((labels, ids, vals, poses))
        cap = max(cap, ids.numel())
    cap = cap + 5
    icap = cap if inbox_cap is None else inbox_cap
    inbox_bytes = lib.fm_merge_inbox_bytes(icap, rows, k, g)
    inboxes = [torch.zeros(inbox_bytes, dtype=torch.uint8, device=dev) for _ in range(g)]
    arr = ctypes.c_void_p * g
    table_bases = arr(*[s.data_ptr() for s in stores])
    inbox_bases = arr(*[b.data_ptr() for b in inboxes])
    rec_ld = int(lib.fm_sorted_record_ld(k))
    preds, loss_sum, statuses = [], 0.0, []
    for r in range(g):                                   # every rank: sort, score by pulling rows, push records
        labels, ids, vals, poses = parsed[r]
        n = ids.numel()
        ids_c = torch.zeros(cap, dtype=torch.int64, device=dev); ids_c[:n] = ids
        vals_c = torch.zeros(cap, dtype=torch.float32, device=dev); vals_c[:n] = vals
        status = torch.zeros(8, dtype=torch.int64, device=dev); status[0] = -1; status[1] = n
        skeys = torch.empty(cap, dtype=torch.int32, device=dev)
        spairs = torch.empty(cap, 2, dtype=torch.int32, device=dev)
        ws_sort = _lib.workspace(lib.fm_sort_workspace(cap), dev)
        check(lib.fm_sort_by_row(ptr(ids_c), ptr(vals_c), ptr(poses), per, cap, _at(status, 1), vocab, g, ptr(skeys),
                                 ptr(spairs), _at(status, 3), ptr(status), ptr(ws_sort), ws_sort.numel(), 0, stream()),
              "fm_sort_by_row")
        rec = torch.empty(per, rec_ld, dtype=torch.float32, device=dev)
        pred = torch.empty(per, dtype=torch.float32, device=dev)
        ws = _lib.workspace(lib.fm_fused_workspace(per), dev)
        w = None if weights is None else torch.as_tensor(weights[r * per:(r + 1) * per], dtype=torch.float32).to(dev)
        check(lib.fm_peer_sorted_score(ptr(ids_c), ptr(vals_c), ptr(poses), ptr(labels), ptr(w), per, k, g, table_bases,
                                       2 * part, lf, lb, _lib.LOSS_KIND[loss_type], len(lines), ptr(pred),
                                       _at(status, 6), _at(status, 7), ptr(rec), rec_ld, ptr(status), ptr(ws), ws.numel(),
                                       stream()), "fm_peer_sorted_score")
        ws_push = _lib.workspace(lib.fm_merge_push_workspace(g), dev)
        check(lib.fm_peer_push_records(ptr(skeys), ptr(spairs), _at(status, 3), cap, ptr(rec), rec_ld, k, r, g, vocab, rows,
                                       inbox_bases, icap, ptr(status), ptr(ws_push), ws_push.numel(), stream()),
              "fm_peer_push_records")
        torch.cuda.synchronize()
        preds.append(pred)
        loss_sum += float(status[6:8].view(torch.float32)[0])
        statuses.append(int(status[0]))
    flag = torch.tensor([max(statuses)], dtype=torch.int64, device=dev)     # what fm_peer_sync would hand out
    for r in range(g):                                   # every owner: merge its inbox
        check(lib.fm_peer_merge_update(ptr(stores[r]), _at(stores[r], part), 2 * part, rows, ptr(inboxes[r]), icap, k, g,
                                       lr, lf, lb, 1.0 / cfg_batch, ptr(flag), stream()), "fm_peer_merge_update")
    torch.cuda.synchronize()
    full = torch.zeros(rows * g, k + 1, dtype=torch.float32, device=dev)
    for r in range(g):
        full[r::g] = stores[r][:, :k + 1]
    return torch.cat(preds).cpu().numpy(), loss_sum, full[:vocab].cpu().numpy(), statuses


@pytest.mark.parametrize("g,k,vocab,hashed,loss_type", [(1, 16, 200003, False, "logistic"), (2, 16, 200003, False, "mse"),
                                                        (3, 5, 200003, True, "logistic"), (4, 16, 37, True, "logistic"),
                                                        (8, 64, 5003, True, "logistic"), (3, 100, 200003, False, "mse")])
def test_virtual_ranks_equal_single_process(criteo_lines, g, k, vocab, hashed, loss_type):
    lines, weights = criteo_lines
    n = 120 * g
    lines, weights = lines[:n], np.float32(weights[:n])
    rng = np.random.default_rng(g * 100 + k)
    flat = rng.uniform(-0.01, 0.01, (vocab, k + 1)).astype(np.float32)
    ref = O.RefModel(vocab, 1, k, flat, batch_size=n, hash_feature_id=hashed, factor_lambda=2e-3, bias_lambda=1e-3,
                     lr=0.05, init_acc=0.1, loss_type=loss_type)
    e = ref.train_step(lines, weights)
    pred, loss, table, statuses = _virtual_step(lines, weights, flat, g, k, vocab, hashed, loss_type, 0.05, 2e-3, 1e-3, n)
    assert statuses == [-1] * g
    np.testing.assert_allclose(loss, e["loss"], rtol=RTOL)
    close(pred, e["pred"], pred_scale(e))
    close(table, ref.flat_table(), 0.05)


def test_inbox_too_small_voids_the_step(criteo_lines):
    lines, _ = criteo_lines
    flat = np.random.default_rng(1).uniform(-0.01, 0.01, (200003, 9)).astype(np.float32)
    pred, loss, table, statuses = _virtual_step(lines[:200], None, flat, 2, 8, 200003, False, "logistic", 0.05, 0.0, 0.0,
                                                200, inbox_cap=100)
    assert all(s & 7 == 5 for s in statuses if s != -1) and any(s != -1 for s in statuses)   # FM_PARSE_E_CAPACITY
    np.testing.assert_array_equal(table, flat)           # nobody merged: the flag said the step was void


@pytest.mark.parametrize("rows,density", [(1, 1.0), (31, 0.5), (1000, 0.0), (4096 * 32 + 5, 0.14), (1 << 24, 0.14),
                                           (3_000_001, 0.9), (70_000, 1.0)])
def test_touched_rows_lists_the_bitmap_in_ascending_order(rows, density):
    """fm_peer_touched_rows (owner side of the atomics exchange): bitmap -> ascending row list, bitmap cleared"""
    lib, dev = _lib.load(), torch.device(DEV)
    rng = np.random.default_rng(rows)
    want = np.flatnonzero(rng.random(rows) < density).astype(np.int64)
    words = (rows + 31) // 32
    host = np.zeros(words, dtype=np.uint32)
    np.bitwise_or.at(host, want >> 5, np.uint32(1) << (want & 31).astype(np.uint32))
    bits = torch.from_numpy(host.view(np.int32)).to(dev)
    ws = _lib.workspace(lib.fm_peer_touched_workspace(rows), dev)
    for cap in (rows, max(1, len(want) // 2)):
        b = bits.clone()
        out = torch.full((cap,), -7, dtype=torch.int64, device=dev)
        n = torch.zeros(1, dtype=torch.int64, device=dev)
        ws.fill_(0x55)                                   # whatever the last call left behind
        check(lib.fm_peer_touched_rows(ptr(b), rows, ptr(out), cap, ptr(n), ptr(ws), ws.numel(), stream()), "touched_rows")
        torch.cuda.synchronize()
        assert int(n) == len(want)
        kept = min(cap, len(want))
        np.testing.assert_array_equal(out[:kept].cpu().numpy(), want[:kept])
        assert (out[kept:] == -7).all() and int(b.abs().sum()) == 0


@pytest.mark.parametrize("k", [4, 16, 64])
def test_bitmap_update_equals_list_update_and_honours_the_guard(k):
    """fm_fused_adagrad_bitmap against fm_peer_touched_rows + fm_fused_adagrad_claim on the same accumulated rows;
    a tripped guard with discard_on_guard throws the sums away and leaves the table alone"""
    lib, dev = _lib.load(), torch.device(DEV)
    rows, part = 70_001, (k + 3 + 3) // 4 * 4
    rng = np.random.default_rng(k)
    touched_rows = np.unique(rng.integers(0, rows, 9000))
    words = (rows + 31) // 32
    host_bits = np.zeros(words, dtype=np.uint32)
    np.bitwise_or.at(host_bits, touched_rows >> 5, np.uint32(1) << (touched_rows & 31).astype(np.uint32))

    def fresh():
        g = torch.Generator(device=dev).manual_seed(5)
        store = torch.zeros(rows, 3 * part, device=dev)
        store[:, :k + 1].uniform_(-0.1, 0.1, generator=g)
        store[:, part:2 * part] = 0.1
        idx = torch.from_numpy(touched_rows).to(dev)
        sums = torch.zeros(len(touched_rows), part, device=dev)
        sums[:, :k + 3].uniform_(-1.0, 1.0, generator=g)
        sums[:, k + 2] = torch.randint(1, 5, (len(touched_rows),), generator=g, device=dev).float()      # cnt
        store[idx, 2 * part:] = sums
        return store, torch.from_numpy(host_bits.view(np.int32)).to(dev)
    args = (0.05, 2e-3, 1e-3, 1.0 / 256)
    a, bits_a = fresh()
    lst = torch.empty(rows, dtype=torch.int64, device=dev)
    n = torch.zeros(1, dtype=torch.int64, device=dev)
    ws = _lib.workspace(lib.fm_peer_touched_workspace(rows), dev)
    check(lib.fm_peer_touched_rows(ptr(bits_a), rows, ptr(lst), rows, ptr(n), ptr(ws), ws.numel(), stream()), "touched")
    check(lib.fm_fused_adagrad_claim(ptr(a), _at(a, part), 3 * part, ptr(lst), rows, ptr(n), k, _at(a, 2 * part), 3 * part,
                                     *args, None, 0, stream()), "claim")
    b, bits_b = fresh()
    seen = torch.zeros(1, dtype=torch.int64, device=dev)
    check(lib.fm_fused_adagrad_bitmap(ptr(b), _at(b, part), 3 * part, ptr(bits_b), rows, k, _at(b, 2 * part), 3 * part,
                                      *args, None, 0, ptr(seen), stream()), "bitmap")
    torch.cuda.synchronize()
    assert int(seen) == len(touched_rows) == int(n) and int(bits_b.abs().sum()) == 0
    assert torch.equal(a, b) and not b[:, 2 * part:].any()
    c, bits_c = fresh()
    before = c.clone()
    bad = torch.tensor([5], dtype=torch.int64, device=dev)          # a status word that says "a line failed"
    check(lib.fm_fused_adagrad_bitmap(ptr(c), _at(c, part), 3 * part, ptr(bits_c), rows, k, _at(c, 2 * part), 3 * part,
                                      *args, ptr(bad), 0, None, stream()), "bitmap, guard: no-op")
    torch.cuda.synchronize()
    assert torch.equal(c, before) and int(bits_c.abs().sum()) > 0
    check(lib.fm_fused_adagrad_bitmap(ptr(c), _at(c, part), 3 * part, ptr(bits_c), rows, k, _at(c, 2 * part), 3 * part,
                                      *args, ptr(bad), 1, None, stream()), "bitmap, guard: discard")
    torch.cuda.synchronize()
    assert torch.equal(c[:, :2 * part], before[:, :2 * part]) and not c[:, 2 * part:].any() and int(bits_c.abs().sum()) == 0
