#!/usr/bin/env python
"""bench.py -- FM train examples/sec on synthetic Criteo-shaped data (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--dist uniform|zipf]

A step = one pass of the whole hot path (parse -> unique -> lookup -> score -> loss ->
grad -> sparse Adagrad) over one batch of synthetic text (cfg2: 39 fields, 2^24 hashed
vocab, factor 16, batch 65536 per GPU).  Prints ONE JSON line (see the task contract):
value = device-resident throughput, e2e = through the host-buffer API with the H2D copy
of each batch's text and the D2H read of its loss inside the timed region, roofline =
the dominant kernel against the measured HBM peak, cpu_baseline = the oracle's
faithful-threaded restatement of the reference CPU ops on this box's host cores.

--impl reference times only the CPU restatement (the TF-hosted original cannot run here).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import bench_data  # noqa: E402

CFG = dict(vocab=1 << 24, factor_num=16, batch=65536, nnz_per_line=39, hash_feature_id=True,
           loss_type="logistic", learning_rate=0.01, init_acc=0.1, factor_lambda=1e-5, bias_lambda=1e-5)
CFG3 = dict(CFG, vocab=1 << 27, factor_num=64, batch=32768)   # BASELINE config 3 (meant for 8 GPUs)
METRIC, UNIT = "fm_train_examples_per_sec", "examples/s"
STAGES = {  # stage marks per step, in order
    "sorted": ["parse_count", "parse_fill", "sort", "score", "update"],
    "claim": ["parse_count", "parse_fill", "score_scatter", "adagrad"],
    "fused": ["parse_count", "parse_fill", "unique", "score_scatter", "adagrad"],
    "composed": ["parse_count", "parse_fill", "unique", "lookup", "score", "loss", "grad", "adagrad"],
}
# my kernels per step.  parse: count, count-slow, poses scan, fill, fill-slow (5); unique: insert, rank
# scan, finish (3); claim: + score_scatter, reduce2, adagrad_claim; fused: + unique + score_scatter,
# reduce2, adagrad; composed: + unique + lookup | score, reduce | loss, reduce | zero rows, grad | adagrad
# sorted: + expand, histogram, one radix pass per 8 key bits | score, reduce2 | update, finish
KERNELS_PER_STEP = {"sorted": 19, "claim": 8, "fused": 11, "composed": 16}


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def algorithmic_bytes(n, u, b, k, text_bytes, fused):
    """Compulsory bytes per stage (fp32; ids i64 at the parser output) -> (strict, as_implemented).

    strict: SURVEY.md section 8(d) to the letter -- each tensor that crosses a stage boundary once per write and
    once per read, each DISTINCT row once per stage that needs it, and a fused stage DROPS the intermediates it
    eliminates (gathered rows, pred_grad, the dense gradient) without adding anything of its own.  This is the
    numerator of roofline.frac.
    as_implemented: strict plus the traffic the chosen algorithm adds by design (the K+3-wide accumulator rows of
    the claim / fused paths: one read-modify-write in the scatter, one read + one clearing write in the update;
    the per-occurrence id list the claim update walks; the sort's key/payload passes and the [B, K+1] example
    records of the sorted path).  Reported as frac_as_implemented."""
    rb = 4 * (k + 1)
    ab = 4 * (k + 3)
    parse = text_bytes + 12 * n + 8 * b                # count + fill together
    if fused == "claim":   # no unique; the update walks the N occurrences' ids
        strict = {"parse": parse,
                  "score_scatter": 12 * n + 4 * (b + 1) + 8 * b + 2 * u * rb,     # rows read, gradient rows written
                  "adagrad": 8 * u + 5 * u * rb}
        impl = {"parse": parse,
                "score_scatter": 12 * n + 4 * (b + 1) + 8 * b + u * rb + 2 * u * ab,
                "adagrad": 8 * n + 2 * u * ab + 4 * u * rb}
        return strict, impl
    if fused == "sorted":  # run-ahead sort by id; the gradient rows are never materialised
        strict = {"parse": parse,
                  "sort": 12 * n + 8 * u + 4 * (b + 1),                           # section 8(d)'s unique + cumsum
                  "score": 12 * n + 4 * (b + 1) + 8 * b + u * rb,
                  "update": 8 * u + 4 * u * rb}
        passes = 3 if k <= 16 else 4
        impl = {"parse": parse,
                "sort": 12 * n + 4 * n + passes * 2 * 12 * n + 12 * n,            # keys + payload per pass, heads
                "score": 12 * n + 4 * (b + 1) + 8 * b + u * rb + b * rb,          # + the [B, K+1] example records
                "update": 12 * n + 4 * u + n * rb + 4 * u * rb}                   # sorted pairs, record gathers (L2)
        return strict, impl
    if fused == "fused":
        strict = {"parse": parse,
                  "unique": 12 * n + 8 * u + 4 * (b + 1),
                  "score_scatter": 16 * n + 4 * (b + 1) + 8 * b + 2 * u * rb,
                  "adagrad": 8 * u + 5 * u * rb}
        impl = dict(strict, score_scatter=16 * n + 4 * (b + 1) + 8 * b + u * rb + 2 * u * ab,
                    adagrad=8 * u + 2 * u * ab + 4 * u * rb)
        return strict, impl
    strict = {
        "parse": parse,
        "unique": 12 * n + 8 * u + 4 * (b + 1),
        "lookup": 8 * u + 2 * u * rb,
        "score": 8 * n + 4 * (b + 1) + u * rb + 4 * b,
        "loss": 16 * b,
        "grad": 8 * n + 4 * (b + 1) + 4 * b + 2 * u * rb,
        "adagrad": 8 * u + 5 * u * rb,
    }
    return strict, dict(strict)


def kernel_traffic(workload, dist, path, layout):
    """DRAM bytes per launch per kernel from the committed `ncu --set full` capture of this very workload
    (profiles/traffic_r2.json, else round 1's); {} when there is none for this workload / path / layout"""
    for name in ("traffic_r2.json", "traffic_r1.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(tpath):
            continue
        doc = json.load(open(tpath))
        for entry in doc.get("captures", [doc]):
            if (entry.get("workload", "cfg2") == workload and entry.get("dist", "uniform") == dist
                    and entry.get("path", "claim") == path and entry.get("layout", "interleaved") == layout):
                return entry["dram_bytes_per_launch"]
    return {}


STAGE_KERNELS = {   # the kernels behind each stage mark (ncu names)
    "sorted": {"update": ["sorted_update_kernel", "sorted_finish_kernel"], "score": ["fused_score_scatter"],
               "parse": ["parse_count_kernel", "parse_fill_kernel"]},
    "claim": {"adagrad": ["fused_adagrad_claim"], "score_scatter": ["fused_score_scatter"],
              "parse": ["parse_count_kernel", "parse_fill_kernel"]},
}


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU while the timed region runs (NVML)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.05)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def cpu_baseline(steps, warmup, sample_lines, dist, threads):
    """the oracle's faithful-threaded restatement of the reference CPU ops, one step per batch"""
    from oracle import oracle as O
    k, vocab = CFG["factor_num"], CFG["vocab"]
    rng = np.random.default_rng(0)
    nb = 10                        # sample.cfg's vocabulary_block_num: Adagrad runs one task per block
    blocks = (rng.random((nb, vocab // nb + 1, k + 1), dtype=np.float32) - 0.5) * 0.02
    acc = np.full_like(blocks, np.float32(CFG["init_acc"]))
    batches = [bench_data.criteo_batch(sample_lines, 100 + i, dist)[:2] for i in range(min(2, steps + warmup))]
    stage = np.zeros(7)
    total = 0.0
    for i in range(warmup + steps):
        text, off = batches[i % len(batches)]
        t0 = time.perf_counter()
        _, ts, _ = O.train_step_faithful(text, off, vocab, True, nb, k, blocks, acc, None, CFG["loss_type"],
                                         CFG["factor_lambda"], CFG["bias_lambda"], CFG["learning_rate"],
                                         CFG["batch"], threads)
        dt = time.perf_counter() - t0
        if i >= warmup:
            total += dt
            stage += ts
    names = ["parse", "unique", "lookup", "score", "loss", "grad", "adagrad"]
    return {"value": steps * sample_lines / total, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d steps of %d-line cfg2 batches (%s ids), all stages, oracle faithful-threaded build: parser "
                      "and tf.unique serial as in the reference, scorer and grad generators on all threads, Adagrad one "
                      "task per vocabulary block (vocabulary_block_num = 10, sample.cfg's) as the TF runtime "
                      "schedules it, lookup a serial gather (faster here than per-block tasks)"
                      % (steps, sample_lines, dist),
            "ms_per_step": 1e3 * total / steps,
            "stage_ms": {n: round(1e3 * s / steps, 3) for n, s in zip(names, stage)}}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (restated; kind 'port')"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = CFG["batch"]        # the benchmark's own batch: same config as the GPU arm
    base = cpu_baseline(args.steps, max(args.warmup, 1), sample, args.dist, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1, sample),
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "stage_ms")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, n_gpus, batch):
    cfg = CFG3 if args.workload == "cfg3" else CFG
    return {"workload": "%s: synthetic Criteo shape, 39 fields, 2^%d hashed vocab, factor %d, "
                        "batch %d per GPU, logistic loss, Adagrad(0.01, 0.1), lambda 1e-5"
                        % (args.workload, cfg["vocab"].bit_length() - 1, cfg["factor_num"], batch),
            "ids": args.dist, "batch_per_gpu": batch, "global_batch": batch * n_gpus, "factor_num": cfg["factor_num"],
            "vocabulary_size": cfg["vocab"], "distinct_batches": args.batches,
            "l2_policy": "inputs larger than L2: %d distinct text batches (~37 MB each) + random rows of a "
                         "1.3 GB table / accumulator per step" % args.batches,
            "parallelism": "single GPU" if n_gpus == 1 else "vocab-sharded x%d (id %% G), %s" % (
                n_gpus, {"peer": "NVLink peer loads + atomics", "merge": "NVLink peer loads + sorted records merged by the "
                         "owners", "exchange": "NCCL all-to-all"}[args.shard_mode])}


def run_ours(args):
    import torch
    if args.layout is None:
        args.layout = "pair" if args.path == "sorted" else "interleaved"
    from fast_tffm_b200.fm_model import Model
    from fast_tffm_b200.fm_ops import PackedLines

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        from bench_sharded import run_sharded  # vocab-sharded path
        return run_sharded(args, CFG3 if args.workload == "cfg3" else CFG, workload_config)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if args.parse_blocks is not None:
        Model.BACKGROUND_PARSE_BLOCKS = args.parse_blocks
    # cfg3 on ONE GPU: the 2^27-row K=64 table, its Adagrad and gradient accumulators are 120 GB interleaved --
    # they fit the 180 GB of one B200, so the whole global batch of 262 144 can also run unsharded
    cfg = dict(CFG3, batch=8 * CFG3["batch"]) if args.workload == "cfg3" else CFG
    b, k = cfg["batch"], cfg["factor_num"]
    m = Model(vocabulary_size=cfg["vocab"], factor_num=k, batch_size=b, hash_feature_id=True,
              loss_type=cfg["loss_type"], learning_rate=cfg["learning_rate"],
              adagrad_init_accumulator=cfg["init_acc"], factor_lambda=cfg["factor_lambda"],
              bias_lambda=cfg["bias_lambda"]).build(dev, seed=0, max_nnz_per_line=cfg["nnz_per_line"], ld=args.ld,
                                                     layout=args.layout, path=args.path)
    if args.l2_fetch:
        # experiment knob: cudaLimitMaxL2FetchGranularity (0x05)
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(args.l2_fetch))
        got = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(got), 5)
        print("l2 fetch granularity ->", args.l2_fetch, "rc", rc, "now", got.value, file=sys.stderr)

    # pinned staging buffers on the GPU's own NUMA node (the CPU baseline later gets every core back)
    from fast_tffm_b200.data import bind_host_to_gpu
    all_cpus = os.sched_getaffinity(0)
    affinity = "off" if args.no_numa_bind else bind_host_to_gpu(local_rank)
    host, devb = [], []
    for i in range(args.batches):
        text, off, nnz = bench_data.criteo_batch(b, 1 + i, args.dist)
        ht = torch.from_numpy(bench_data.pad16(text)).pin_memory()
        ho = torch.from_numpy(off).pin_memory()
        host.append(PackedLines(ht, ho))
        devb.append(PackedLines(ht.to(dev), ho.to(dev)))
    text_bytes = int(host[0].offsets[-1])

    # ---- device-resident throughput (value) + per-stage CUDA events ----------------
    fused = args.path
    stages = STAGES[fused]
    ahead_on = fused in ("claim", "sorted") and not args.no_parse_ahead

    def run_steps(n_steps, base, ev=None):
        """the timed loop; the warm-up runs the very same code so that every buffer set, stream and
        event the pipeline needs exists before the clock starts"""
        bufs = None
        ahead = m.parse_ahead(devb[base % len(devb)]) if ahead_on else None
        for s in range(n_steps):
            cursor = [0]
            if ev is not None:
                ev[s][0].record()

            def mark(name, s=s, cursor=cursor):
                if ev is not None:
                    cursor[0] += 1
                    ev[s][cursor[0]].record()
            # the parse of batch s+1 runs on a side stream under the table kernels of batch s
            nxt = m.parse_ahead(devb[(base + s + 1) % len(devb)]) if ahead_on else None
            bufs = m.train_step_async(devb[(base + s) % len(devb)], mark=mark, path=fused, parsed=ahead)
            ahead = nxt
        return bufs

    run_steps(args.warmup, 0)
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(stages) + 1)] for _ in range(args.steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()
    torch.cuda.synchronize()
    bufs = run_steps(args.steps, args.warmup, ev)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    total_ms = ev[0][0].elapsed_time(ev[-1][-1])
    stage_ms = {n: sum(ev[s][i].elapsed_time(ev[s][i + 1]) for s in range(args.steps)) / args.steps
                for i, n in enumerate(stages)}
    # kernel durations for the roofline: a second, untimed pass of the same steps with the parse inline, so that
    # no other kernel runs beside the one between two events (in the pipelined loop above the parse of the next
    # batch shares the SMs with the table kernels and the per-stage times include that interference)
    stage_ms_pipelined = stage_ms
    if ahead_on:
        iso = min(args.steps, 20)
        for s in range(2):      # the inline path has its own buffer set: allocate it outside the events
            m.train_step_async(devb[s % len(devb)], path=fused)
        ev2 = [[torch.cuda.Event(enable_timing=True) for _ in range(len(stages) + 1)] for _ in range(iso)]
        for s in range(iso):
            cursor = [0]
            ev2[s][0].record()

            def mark2(name, s=s, cursor=cursor):
                cursor[0] += 1
                ev2[s][cursor[0]].record()
            bufs = m.train_step_async(devb[(args.warmup + args.steps + s) % len(devb)], mark=mark2, path=fused)
        torch.cuda.synchronize()
        stage_ms = {n: sum(ev2[s][i].elapsed_time(ev2[s][i + 1]) for s in range(iso)) / iso
                    for i, n in enumerate(stages)}
    st = bufs["status"].cpu()
    assert int(st[0]) == -1, "parse error in the benchmark data"
    nnz = int(st[1])
    n_unique = int(st[4]) if fused != "claim" else int(torch.unique(bufs["ids"][:nnz]).numel())
    if fused == "sorted":
        assert n_unique == int(torch.unique(bufs["ids"][:nnz]).numel()) and int(st[3]) == nnz, "sort counts are off"
    loss = float(st[6:8].view(torch.float32)[0])
    value = args.steps * b / (total_ms * 1e-3)

    # ---- roofline of the dominant stage -------------------------------------------
    alg, alg_impl = algorithmic_bytes(nnz, n_unique, b, k, text_bytes, fused)
    merged = dict(stage_ms)
    merged["parse"] = merged.pop("parse_count") + merged.pop("parse_fill")
    peak, peak_kind = measured_peak_gbs()
    layout_used = getattr(m, "layout", args.layout)
    dram = kernel_traffic(args.workload, args.dist, fused, layout_used) if args.ld is None else {}
    per_stage = {}
    for n_, ms in merged.items():
        kernels = STAGE_KERNELS.get(fused, {}).get(n_, [])
        traffic_n = sum(dram[kn] for kn in kernels) if kernels and all(kn in dram for kn in kernels) else None
        gbs = alg[n_] / (ms * 1e-3) / 1e9 if ms > 0 else None
        per_stage[n_] = {"ms": round(ms, 4), "alg_mb": round(alg[n_] / 1e6, 2),
                         "alg_mb_as_implemented": round(alg_impl[n_] / 1e6, 2),
                         "gbs": round(gbs, 1) if gbs else None, "frac": round(gbs / peak, 4) if gbs else None,
                         "frac_as_implemented": round(alg_impl[n_] / (ms * 1e-3) / 1e9 / peak, 4) if ms > 0 else None,
                         "dram_mb": None if traffic_n is None else round(traffic_n / 1e6, 1),
                         "traffic_ratio": None if traffic_n is None else round(traffic_n / alg[n_], 3)}
    # the dominant GATHER/SCATTER kernel: the longest of the table-touching stages
    dom = max((n_ for n_ in merged if n_ not in ("parse", "sort")), key=lambda n_: merged[n_])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": per_stage[dom]["gbs"], "peak": peak,
                "peak_kind": peak_kind, "unit": "GB/s", "frac": per_stage[dom]["frac"],
                "frac_as_implemented": per_stage[dom]["frac_as_implemented"],
                "traffic": None if per_stage[dom]["dram_mb"] is None else int(per_stage[dom]["dram_mb"] * 1e6),
                "traffic_ratio": per_stage[dom]["traffic_ratio"],
                "bytes_rule": "frac = SURVEY 8(d) bytes (fused stages drop intermediates, add nothing) / kernel ms / "
                              "peak; frac_as_implemented adds the traffic the algorithm itself introduces",
                "step_frac": round(sum(alg.values()) / (total_ms / args.steps * 1e-3) / 1e9 / peak, 4),
                "stages": per_stage,
                "stages_note": "per-kernel ms from an untimed pass with the parse inline (nothing runs beside the "
                               "kernel timed); stages_pipelined_ms = the same events inside the timed, pipelined loop",
                "stages_pipelined_ms": {n_: round(v_, 4) for n_, v_ in stage_ms_pipelined.items()}}

    # ---- end to end through the host-buffer API -----------------------------------
    h2d = int(host[0].text.numel() + host[0].offsets.numel() * 8)
    d2h = int(bufs["status"].numel() * 8)
    # the package's own host-batch pipeline (fast_tffm_b200/pipeline.py): H2D copy of batch s+2, parse of batch
    # s+1 and the table kernels of batch s overlap; each step's status + loss record comes back to pinned memory
    from fast_tffm_b200.pipeline import HostBatchPipeline
    pipe = HostBatchPipeline(m) if ahead_on else None
    stage_in = [PackedLines(torch.empty_like(devb[0].text), torch.empty_like(devb[0].offsets)) for _ in range(2)]
    copy_stream = torch.cuda.Stream()
    pinned_status = torch.empty(args.steps + args.warmup, 8, dtype=torch.int64).pin_memory()
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    last_result = [None]

    def e2e_loop(n_steps, base):
        if pipe is not None:
            for res in pipe.run((host[(base + s) % len(host)], None) for s in range(n_steps)):
                last_result[0] = res
            return
        # other paths (--path fused / composed, --no-parse-ahead): copy of batch s+1 under the step of batch s
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
            bb = m.train_step_async(stage_in[slot], path=fused)
            consumed[slot] = torch.cuda.Event()
            consumed[slot].record(main)
            pinned_status[base + s].copy_(bb["status"], non_blocking=True)   # loss + status back to the host

    e2e_loop(args.warmup, 0)
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    e2e_loop(args.steps, args.warmup)
    t1.record()
    torch.cuda.synchronize()
    e2e_ms = t0.elapsed_time(t1)
    if pipe is not None:
        assert np.isfinite(last_result[0].loss())       # raises if the batch had not parsed
    else:
        assert int(pinned_status[-1, 0]) == -1
    e2e_value = args.steps * b / (e2e_ms * 1e-3)

    os.sched_setaffinity(0, all_cpus)
    base = cpu_baseline(16, 1, 65536, args.dist, os.cpu_count() or 1) if not (args.no_cpu_baseline or args.workload != "cfg2") else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, 1, b),
        "roofline": roofline,
        "cpu_baseline": None if base is None else {k_: base[k_] for k_ in ("value", "unit", "cores", "kind", "sample", "stage_ms")},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": KERNELS_PER_STEP[fused] * args.steps,      # either form of the claim path: the same count
        "claim_form": m._form if fused == "claim" else None,
        "path": fused, "layout": args.layout, "unique_per_step": n_unique, "parse_ahead": ahead_on,
        "host_affinity": affinity, "clocks": clocks,
        "final_loss": loss, "nnz_per_step": nnz,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dist", default="uniform", choices=["uniform", "zipf"])
    ap.add_argument("--batches", type=int, default=8, help="distinct synthetic batches cycled through")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3"],
                    help="cfg2 (default): 2^24 vocab, K=16, 65536 examples per GPU; cfg3: 2^27 vocab, K=64, "
                         "32768 examples per GPU (BASELINE config 3, 8 GPUs)")
    ap.add_argument("--path", default="claim", choices=["sorted", "claim", "fused", "composed"],
                    help="claim: no unique, table-shaped accumulators (default: the fastest measured); sorted: radix "
                         "sort by row + segment-reduce update; fused: unique + compact accumulators; composed: the "
                         "reference's op-by-op chain")
    ap.add_argument("--layout", default=None, choices=["pair", "split", "interleaved"],
                    help="default: pair ([w | h] per id) for the sorted path, interleaved ([w | h | g]) for the others")
    ap.add_argument("--no-parse-ahead", action="store_true",
                    help="parse each batch inline instead of on a side stream under the previous batch's kernels")
    ap.add_argument("--shard-mode", default="peer", choices=["peer", "merge", "exchange"],
                    help="N > 1: peer = rows and gradients move by NVLink loads / atomics inside the kernels; "
                         "merge = rows by NVLink loads, gradients as sorted records merged by the owners; "
                         "exchange = NCCL all-to-all of ids, rows and accumulator rows")
    ap.add_argument("--parse-blocks", type=int, default=None,
                    help="experiment: grid cap of the run-ahead parse (Model.BACKGROUND_PARSE_BLOCKS; 0 = full grid)")
    ap.add_argument("--no-numa-bind", action="store_true",
                    help="do not bind the process to the CPUs of the GPU's NUMA node before allocating pinned buffers")
    ap.add_argument("--ld", type=int, default=None, help="experiment: table row stride in floats")
    ap.add_argument("--l2-fetch", type=int, default=0, help="experiment: cudaLimitMaxL2FetchGranularity")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
