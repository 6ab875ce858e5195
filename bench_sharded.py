"""bench.py's multi-GPU leg: vocab-sharded FM training, one process per GPU (torchrun).

Weak scaling: every rank parses its own 65 536-line cfg2 batch per step (global batch =
65 536 x N); the 2^24-row table and its Adagrad accumulator are row-partitioned id % N.
Timing: CUDA events on each rank around the K timed steps, barrier + synchronize on both
sides, MAX over ranks.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

import bench_data


def parity_check(cls, cfg, world, rank, dev):
    """One small step through the sharded model on every rank and through the single-GPU Model on rank 0, same
    global batch, same explicit initial table: loss, scores and the trained table must agree to 1e-5 (relative, over
    the magnitude of the summed terms).  Raises if they do not; the returned dict goes into the JSON line, so the
    scaling runs carry their own evidence that G GPUs compute what one GPU computes."""
    from fast_tffm_b200.fm_model import Model
    vocab, k, per = 1 << 20, cfg["factor_num"], 512
    gen = torch.Generator().manual_seed(1234)
    init = (torch.rand(vocab, k + 1, generator=gen) - 0.5) * 0.02
    kw = dict(vocabulary_size=vocab, factor_num=k, batch_size=per * world, hash_feature_id=True,
              loss_type=cfg["loss_type"], learning_rate=0.05, adagrad_init_accumulator=cfg["init_acc"],
              factor_lambda=1e-3, bias_lambda=1e-3)
    text, off, _ = bench_data.criteo_batch(per * world, 777, "zipf")      # hot ids: rows shared between ranks
    lines = [bytes(text[off[i]:off[i + 1] - 1]) for i in range(per * world)]
    sm = cls(Model(**kw), device=dev).build(init_table=init)
    loss = sm.train_step(lines[rank * per:(rank + 1) * per])
    pred = sm.predict(lines[rank * per:(rank + 1) * per])
    preds = [torch.empty_like(pred) for _ in range(world)]
    dist.all_gather(preds, pred)
    full = sm.gather_full_table()
    out = None
    if rank == 0:
        one = Model(**kw).build(dev, init_table=init)
        loss1 = one.train_step(lines)
        pred1 = one.predict(lines)
        t1, tg = one.table[:, :k + 1].double(), full.double()
        moved = (t1 != init.to(dev).double()).any(1)
        # scale of a score: sum |x b| + 0.5 sum_k (sum |x v_k|)^2 per example (the second-order term cancels)
        bufs = one._bufs
        nnz = int(bufs["status"][1])
        ids, x = bufs["ids"][:nnz], bufs["vals"][:nnz].double().abs()
        poses = bufs["poses"][:per * world + 1].long()
        ex = torch.repeat_interleave(torch.arange(per * world, device=dev), poses[1:] - poses[:-1])
        rows = t1[ids].abs()
        lin = torch.zeros(per * world, dtype=torch.float64, device=dev).index_add_(0, ex, x * rows[:, 0])
        fac = torch.zeros(per * world, k, dtype=torch.float64, device=dev).index_add_(0, ex, x[:, None] * rows[:, 1:])
        pscale = lin + 0.5 * (fac ** 2).sum(1)
        perr = float(((torch.cat(preds).double() - pred1.double()).abs() / (pred1.double().abs() + pscale)).max())
        terr = float(((tg - t1).abs() / (t1.abs() + 0.05)).max())
        lerr = abs(loss - loss1) / abs(loss1)
        out = {"ok": bool(perr <= 1e-5 and terr <= 1e-5 and lerr <= 1e-5), "global_batch": per * world,
               "rel_err_loss": lerr, "max_err_over_scale_pred": perr, "max_err_over_scale_table": terr,
               "rows_moved": int(moved.sum()), "tolerance": 1e-5,
               "what": "sharded step + predict on %d ranks vs single-GPU Model on rank 0, same batch and table" % world}
        assert out["ok"], "multi-GPU parity check failed: %r" % (out,)
    del sm
    dist.barrier()
    return out


def run_sharded(args, cfg, workload_config):
    from bench import METRIC, UNIT, ClockSampler, algorithmic_bytes, measured_peak_gbs
    from fast_tffm_b200.fm_model import Model
    from fast_tffm_b200.fm_ops import PackedLines
    from fast_tffm_b200.sharded import PeerShardedModel, ShardedModel

    world, rank = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    from fast_tffm_b200.data import bind_host_to_gpu
    affinity = "off" if args.no_numa_bind else bind_host_to_gpu(local_rank)   # before the pinned staging buffers
    b, k = cfg["batch"], cfg["factor_num"]
    model = Model(vocabulary_size=cfg["vocab"], factor_num=k, batch_size=b * world, hash_feature_id=True,
                  loss_type=cfg["loss_type"], learning_rate=cfg["learning_rate"],
                  adagrad_init_accumulator=cfg["init_acc"], factor_lambda=cfg["factor_lambda"],
                  bias_lambda=cfg["bias_lambda"])
    peer = args.shard_mode in ("peer", "merge")
    if peer:
        def cls(mdl, device=None, max_nnz_per_line=None):
            return PeerShardedModel(mdl, device=device, max_nnz_per_line=max_nnz_per_line,
                                    exchange="records" if args.shard_mode == "merge" else "atomics")
    else:
        cls = ShardedModel
    parity = parity_check(cls, cfg, world, rank, dev)
    m = cls(model, device=dev, max_nnz_per_line=cfg["nnz_per_line"]).build(seed=0)
    host, devb = [], []
    for i in range(args.batches):
        text, off, _ = bench_data.criteo_batch(b, 1000 * (rank + 1) + i, args.dist)
        ht, ho = torch.from_numpy(bench_data.pad16(text)).pin_memory(), torch.from_numpy(off).pin_memory()
        host.append(PackedLines(ht, ho))
        devb.append(PackedLines(ht.to(dev), ho.to(dev)))

    stage_events = []
    ahead_on = peer and not args.no_parse_ahead

    def timed(batches, steps, base, stages=False):
        dist.barrier()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        loss = None
        ahead = m.parse_ahead(batches[base % len(batches)]) if ahead_on else None
        for s in range(steps):
            evs = []

            def mark(name, evs=evs):
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                evs.append((name, e))
            if stages:
                mark("start")
            kw = {}
            if ahead_on:   # parse of the next batch under this batch's NVLink-bound kernels
                nxt = m.parse_ahead(batches[(base + s + 1) % len(batches)])
                kw["parsed"] = ahead
            loss = m.train_step_async(batches[(base + s) % len(batches)], global_batch=b * world,
                                      mark=mark if stages else (lambda name: None), **kw)
            if ahead_on:
                ahead = nxt
            if stages:
                mark("allreduce_loss")
                stage_events.append(evs)
        t1.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if peer:
            assert int(loss[1].item()) == -1, "parse error in the benchmark data"
            loss = loss[0]
        return float(ms.item()), float(loss[0].item())

    timed(devb, args.warmup, 0)      # warm-up through the same pipelined loop (buffer sets, side stream, NCCL)
    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms, loss = timed(devb, args.steps, args.warmup, stages=True)
    clocks = sampler.stop()
    stage_ms = {}
    for evs in stage_events:
        for (_, e0), (name, e1) in zip(evs[:-1], evs[1:]):
            stage_ms[name] = stage_ms.get(name, 0.0) + e0.elapsed_time(e1) / len(stage_events)
    # ---- end to end: pinned host text -> H2D on a copy stream (double-buffered) -> step -> loss back ----
    from fast_tffm_b200.pipeline import HostBatchPipeline
    pipe = HostBatchPipeline(m) if ahead_on else None
    stage_in = [PackedLines(torch.empty_like(devb[0].text), torch.empty_like(devb[0].offsets)) for _ in range(2)]
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    pinned_loss = torch.zeros(args.steps + args.warmup, 2, dtype=torch.float32).pin_memory()
    last_result = [None]

    def e2e_loop(n_steps, base):
        if pipe is not None:   # the package's host-batch pipeline: copy s+2 | parse s+1 | table kernels s
            for res in pipe.run((host[(base + s) % len(host)], None, b * world) for s in range(n_steps)):
                last_result[0] = res
            return
        main = torch.cuda.current_stream()
        for s in range(n_steps):
            slot = s % 2
            with torch.cuda.stream(copy_stream):
                if s >= 2:
                    copy_stream.wait_event(consumed[slot])
                hb = host[(base + s) % len(host)]
                stage_in[slot].text.copy_(hb.text, non_blocking=True)
                stage_in[slot].offsets.copy_(hb.offsets, non_blocking=True)
                ready[slot].record(copy_stream)
            main.wait_event(ready[slot])
            out = m.train_step_async(stage_in[slot], global_batch=b * world)
            consumed[slot] = torch.cuda.Event()
            consumed[slot].record(main)
            pinned_loss[base + s].copy_(out[0] if peer else out, non_blocking=True)

    e2e_loop(args.warmup, 0)
    dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    e2e_loop(args.steps, args.warmup)
    t1.record()
    torch.cuda.synchronize()
    dist.barrier()
    e2e_t = torch.tensor([t0.elapsed_time(t1)], device=dev)
    dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_t.item())
    if pipe is not None:
        assert np.isfinite(last_result[0].loss())

    if peer:
        st = m.last["status"].cpu()
        nnz_local = int(st[1])
        remote = nnz_local * (world - 1) / world
        if args.shard_mode == "merge":      # no touched-row list in this mode: this rank's own distinct ids
            bufs = next(v for v in m._slots.values() if v["status"] is m.last["status"])
            n_rows = int(torch.unique(bufs["ids"][:nnz_local]).numel())
        else:
            n_rows = int(m.n_rows.item())
        stats = torch.tensor([nnz_local, n_rows, remote], dtype=torch.float64, device=dev)
    else:
        stats = torch.tensor([m.last["nnz"], m.last["n_unique"], m.last["sent_rows"]], dtype=torch.float64, device=dev)
    dist.all_reduce(stats)
    if rank == 0:
        nnz, nuniq, sent = (float(x) / world for x in stats.tolist())
        text_bytes = int(host[0].offsets[-1])
        alg, _ = algorithmic_bytes(nnz, nuniq, b, k, text_bytes, "fused")
        peak, peak_kind = measured_peak_gbs()
        step_s = total_ms / args.steps * 1e-3
        gld = (k + 3 + 3) // 4 * 4
        ld = (k + 1 + 3) // 4 * 4
        # per rank per step over NVLink: exchange mode moves id + row + accumulator row per remote distinct id
        # each way; peer mode reads a row and pushes an accumulator row per remote occurrence
        xbytes = sent * (8 + 4 * ld + 4 * gld)
        if args.shard_mode == "merge":      # a row pulled per remote occurrence; a key word + (at most) a record pushed
            xbytes = sent * (4 * ld + 4 + 4 * gld)
        line = {
            "metric": METRIC, "value": args.steps * b * world / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world, b),
            # the step is bound by what crosses NVLink: bytes per rank per direction over the whole step
            # time, against the peer-copy bandwidth B200_PROFILING.md states as measured on this pool (770 GB/s per
            # direction per GPU; 900 nominal).  The HBM view of the same step stays beside it.
            "roofline": {"bound": "nvlink", "kernel": "whole step (per rank, per direction)",
                         "achieved": round(xbytes / step_s / 1e9, 1), "peak": 770.0,
                         "peak_kind": "measured peer copy per direction (B200_PROFILING.md); nominal 900",
                         "unit": "GB/s", "frac": round(xbytes / step_s / 1e9 / 770.0, 4), "traffic": None,
                         "bytes_per_rank_per_direction": int(xbytes),
                         "hbm": {"achieved": round(sum(alg.values()) / step_s / 1e9, 1), "peak": peak,
                                 "peak_kind": peak_kind, "frac": round(sum(alg.values()) / step_s / 1e9 / peak, 4)}},
            "parity_check": parity,
            "cpu_baseline": None,
            "e2e": {"value": args.steps * b * world / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(host[0].text.numel() + host[0].offsets.numel() * 8) * world,
                    "d2h_bytes_per_step": 16 * world, "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": ({"peer": 9, "merge": 19, "exchange": 20}[args.shard_mode]) * args.steps, "shard_mode": args.shard_mode, "parse_ahead": ahead_on,
            "stage_ms_rank0": {k_: round(v, 4) for k_, v in stage_ms.items()},
            "host_affinity_rank0": affinity, "clocks": clocks, "final_loss": loss, "nnz_per_rank": nnz, "unique_per_rank": nuniq,
        }
        print(json.dumps(line))
    dist.barrier()
    dist.destroy_process_group()
