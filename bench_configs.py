#!/usr/bin/env python
"""Measurements for the BASELINE.json configs that are not bench.py's headline line.

    python bench_configs.py cfg4            mse + per-example weights, ragged rows (1-512 nnz), K=32
    python bench_configs.py cfg5 [--quick]  predict-only sweep: batch 4K-1M x factor 8-128, 1 GPU
    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 bench_configs.py cfg5 [--quick]
                                            the same sweep on G GPUs, every rank scoring its own batch (weak
                                            scaling): table sharded over NVLink peer memory, then replicated

Prints one JSON line per measurement (CUDA events, device-resident inputs, >= 3 warm-up runs).
cfg3 (2^27 vocab, K=64, 8 GPUs) is `bench.py --gpus 8 --workload cfg3` under torchrun.
"""
import argparse
import json
import sys
import time

import numpy as np
import torch

import bench_data
from fast_tffm_b200.fm_model import Model
from fast_tffm_b200.fm_ops import PackedLines, pack_lines


def ragged_text(batch, seed, max_nnz=512, vocab=1 << 24):
    """cfg4 lines: label '%.6g', 1..max_nnz tokens 'id:%.6g' (values from a pool of N(0,1) draws)"""
    rng = np.random.default_rng(seed)
    n = rng.integers(1, max_nnz + 1, batch)
    total = int(n.sum())
    pool = np.array(["%.6g" % v for v in rng.normal(size=65536)])
    toks = np.char.add(np.char.add(np.char.mod("%d", rng.integers(0, vocab, total)), ":"),
                       pool[rng.integers(0, pool.size, total)])
    labels = pool[rng.integers(0, pool.size, batch)]
    off = np.concatenate([[0], np.cumsum(n)])
    lines = [labels[i] + " " + " ".join(toks[off[i]:off[i + 1]]) for i in range(batch)]
    weights = np.round(rng.uniform(0.5, 2.0, batch), 3).astype(np.float32)
    return lines, weights, total


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / steps


def cfg4(args):
    dev = torch.device("cuda", 0)
    b, k, vocab = 65536, 32, 1 << 24
    t = time.time()
    batches = []
    for i in range(1):   # one batch (16.8 M features, ~290 MB of text: far beyond L2), reused
        lines, weights, nnz = ragged_text(b, 10 + i)
        batches.append((pack_lines(lines).to(dev, False), torch.from_numpy(weights).to(dev), nnz))
    batches.append(batches[0])
    gen_s = time.time() - t
    m = Model(vocabulary_size=vocab, factor_num=k, batch_size=b, hash_feature_id=False, loss_type="mse",
              learning_rate=0.01, adagrad_init_accumulator=0.1, factor_lambda=1e-5, bias_lambda=1e-5).build(
        dev, seed=0, max_nnz_per_line=512, path=args.path)
    it = [0]

    def step():
        p, w, _ = batches[it[0] % 2]
        it[0] += 1
        return m.train_step_async(p, w)
    ms_inline = timed(step, args.steps)
    # per-stage times of the inline step (CUDA events between the stages)
    names, ev = [], []

    def mark(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        names.append(name)
        ev.append(e)
    e0 = torch.cuda.Event(enable_timing=True)
    e0.record()
    m.train_step_async(batches[0][0], batches[0][1], mark=mark)
    torch.cuda.synchronize()
    stages, prev = {}, e0
    for n, e in zip(names, ev):
        stages[n] = round(stages.get(n, 0.0) + prev.elapsed_time(e), 4)
        prev = e
    # the step as the pipeline runs it: the parse of the next batch on the side stream under this batch's table kernels
    state = {"ahead": m.parse_ahead(batches[0][0])}

    def step_ahead():
        p, w, _ = batches[it[0] % 2]
        it[0] += 1
        nxt = m.parse_ahead(batches[it[0] % 2][0])
        out = m.train_step_async(p, w, parsed=state["ahead"])
        state["ahead"] = nxt
        return out
    ms = timed(step_ahead, args.steps)
    torch.cuda.synchronize()
    bufs = step()
    st = bufs["status"].cpu()
    assert int(st[0]) == -1
    text_bytes = int(batches[0][0].offsets[-1])
    print(json.dumps({"config": "cfg4: mse, per-example weights, 1-512 nnz/row (mean 256.5), K=32, V=2^24, B=65536, 1xB200",
                      "metric": "fm_train_examples_per_sec", "value": b / (ms * 1e-3), "ms_per_step": ms,
                      "ms_per_step_parse_inline": ms_inline, "stage_ms_inline": stages,
                      "nnz_per_step": int(st[1]), "text_mb": round(text_bytes / 1e6, 1), "path": args.path,
                      "loss": float(st[6:8].view(torch.float32)[0]), "data_gen_s": round(gen_s, 1)}))


def cfg5_multi(args):
    """cfg5 at G > 1: (a) the table row-partitioned id % G, rows read over NVLink; (b) one full table per GPU"""
    import os
    import torch.distributed as dist
    from fast_tffm_b200.sharded import PeerShardedModel
    world, rank = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    vocab = 1 << 24
    sizes = [65536, 1048576] if args.quick else [4096, 16384, 65536, 262144, 1048576]
    ks = [16, 64] if args.quick else [8, 16, 32, 64, 128]
    data = {}
    for b in sizes:
        text, off, _ = bench_data.criteo_batch(b, 5 + 100 * rank)
        data[b] = PackedLines(torch.from_numpy(bench_data.pad16(text)).to(dev), torch.from_numpy(off).to(dev))
    for mode in ("sharded", "replicated"):
        for k in ks:
            cfg = Model(vocabulary_size=vocab, factor_num=k, batch_size=65536, hash_feature_id=True)
            if mode == "sharded":
                m = PeerShardedModel(cfg, device=dev, max_nnz_per_line=39).build(seed=0)
                run, status = (lambda b: m.predict_async(data[b])), (lambda: m._slots[0]["status"])
                if not args.no_parse_ahead:      # a stream of batches: the next one's parse under this one's scorer
                    os.environ.setdefault("FM_AHEAD_BLOCKS", "0")
                    state = {}

                    def run(b, m=m, state=state):
                        if state.get("b") != b:
                            state.update(b=b, ahead=m.parse_ahead(data[b]))
                        nxt = m.parse_ahead(data[b])
                        m.predict_async(data[b], parsed=state["ahead"])
                        state["ahead"] = nxt
                    status = lambda: m._slots[1]["status"]
            else:
                m = cfg.build(dev, seed=0, max_nnz_per_line=39, layout="split")
                run, status = (lambda b: m.predict_async(data[b])), (lambda: m._bufs["status"])
                if not args.no_parse_ahead:
                    m._bg_blocks = 0
                    state = {}

                    def run(b, m=m, state=state):
                        if state.get("b") != b:
                            state.update(b=b, ahead=m.parse_ahead(data[b], sort=False))
                        nxt = m.parse_ahead(data[b], sort=False)
                        m.predict_async(data[b], parsed=state["ahead"])
                        state["ahead"] = nxt
            for b in sizes:
                steps = max(3, min(args.steps, int(2e6 // b) + 3))
                for _ in range(3):
                    run(b)
                dist.barrier()
                torch.cuda.synchronize()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                for _ in range(steps):
                    run(b)
                t1.record()
                torch.cuda.synchronize()
                assert int(status().cpu()[0]) == -1
                ms = torch.tensor([t0.elapsed_time(t1) / steps], device=dev)
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(json.dumps({"config": "cfg5: predict-only, 39 nnz, V=2^24, %dxB200, table %s" % (world, mode),
                                      "metric": "fm_predict_examples_per_sec", "n_gpus": world, "table": mode,
                                      "batch_per_gpu": b, "factor_num": k, "value": b * world / (float(ms) * 1e-3),
                                      "ms_per_batch": float(ms)}))
                    sys.stdout.flush()
            del m, run, status
            torch.cuda.empty_cache()
            dist.barrier()
    dist.destroy_process_group()


def cfg5(args):
    import os
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return cfg5_multi(args)
    dev = torch.device("cuda", 0)
    vocab = 1 << 24
    sizes = [4096, 65536] if args.quick else [4096, 16384, 65536, 262144, 1048576]
    ks = [16, 64] if args.quick else [8, 16, 32, 64, 128]
    data = {}
    for b in sizes:
        text, off, nnz = bench_data.criteo_batch(b, 5)
        data[b] = PackedLines(torch.from_numpy(bench_data.pad16(text)).to(dev), torch.from_numpy(off).to(dev))
    for k in ks:
        m = Model(vocabulary_size=vocab, factor_num=k, batch_size=65536, hash_feature_id=True).build(
            dev, seed=0, max_nnz_per_line=39, layout="split")
        for b in sizes:
            steps = max(3, min(args.steps, int(2e6 // b) + 3))
            ms_inline = timed(lambda: m.predict_async(data[b]), steps)
            st = m._bufs["status"].cpu()
            assert int(st[0]) == -1
            # a stream of batches: the parse of the next one runs on the side stream under this one's scorer
            m._bg_blocks = int(os.environ.get("FM_PREDICT_AHEAD_BLOCKS", "0"))     # predict is parse-bound: no grid cap
            state = {"ahead": m.parse_ahead(data[b], sort=False)}

            def step_ahead():
                nxt = m.parse_ahead(data[b], sort=False)
                out = m.predict_async(data[b], parsed=state["ahead"])
                state["ahead"] = nxt
                return out
            ms = timed(step_ahead, steps)
            torch.cuda.synchronize()
            assert int(m._bufs["status"].cpu()[0]) == -1
            print(json.dumps({"config": "cfg5: predict-only (parse -> score from the table), 39 nnz, V=2^24, 1xB200",
                              "metric": "fm_predict_examples_per_sec", "batch": b, "factor_num": k,
                              "value": b / (min(ms, ms_inline) * 1e-3), "ms_per_batch": min(ms, ms_inline),
                              "ms_per_batch_parse_inline": ms_inline, "ms_per_batch_parse_ahead": ms}))
            sys.stdout.flush()
        del m
        torch.cuda.empty_cache()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["cfg4", "cfg5"])
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--no-parse-ahead", action="store_true", help="cfg5 at G > 1: parse each batch inline")
    ap.add_argument("--path", default="claim", choices=["claim", "sorted"], help="cfg4: training path")
    a = ap.parse_args()
    {"cfg4": cfg4, "cfg5": cfg5}[a.which](a)
