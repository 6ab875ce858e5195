#!/usr/bin/env python
"""run_tffm.py -- the reference's command line (run_tffm.py:123-231) over the B200 hot path.

    python run_tffm.py train   sample.cfg [-m] [-t TRACE.json]
    python run_tffm.py predict sample.cfg

Same config file, same stdout lines ("-- Global Step: %d; Avg loss: %.5f;",
"Average speed: ...  ex/s", "Model saved to ..."), same <file>_score outputs.
    python run_tffm.py generate sample.cfg --export_path DIR
writes the serving artefact of the newest checkpoint: not a TF SavedModel (no TF runtime here) but the same content
and contract, the weight table + a scorer for unlabeled "id:val ..." lines (fast_tffm_b200/serving.py).  Not carried
over: --dist / --protocol / --wait-for-workers (gRPC parameter server).  The replacement for --dist is

    python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 run_tffm.py train sample.cfg

one process per GPU, the table sharded over NVLink peer memory: every rank takes
its slice of each global batch of `batch_size` lines, rank 0 prints and writes the checkpoint (in the
single-GPU layout, so `predict` runs on one GPU afterwards).
"""
from __future__ import print_function

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

from tffm.fm_model import Model
from fast_tffm_b200 import data as fm_data
from fast_tffm_b200 import fm_ops
from fast_tffm_b200.fm_model import latest_checkpoint
from fast_tffm_b200.pipeline import HostBatchPipeline


def _ckpt_path(model):
    return os.path.join(model.log_dir, "model.ckpt.pt")


def _dataset(files, weight_files, device, stream_mb):
    """files resident in HBM when they fit comfortably, else streamed through it in segments"""
    total = sum(os.path.getsize(f) for f in files)
    free, _ = torch.cuda.mem_get_info(device)
    if stream_mb is None and total < free // 2:
        return fm_data.DeviceDataset(files, weight_files, device)
    return fm_data.SegmentedDataset(files, weight_files, device, segment_bytes=(stream_mb if stream_mb else 256) << 20)


def _training_set(cfg, device, stream_mb):
    return _dataset(cfg.train_files, cfg.weight_files, device, stream_mb)


def predict(model, stream_mb=None):
    for data_file in model.predict_files:
        lines = _dataset([data_file], (), model.device, stream_mb)      # split into lines on the device
        with open(data_file + '_score', 'w') as f:
            for batch, _ in lines.batches(model.batch_size, 1, shuffle=False):
                scores = model.predict(batch).cpu().numpy()
                f.write(''.join(s + '\n' for s in scores.astype(str).tolist()))   # str(np.float32), as run_tffm.py:14
    print('Done. Scores saved to same directory as predict files')


def _validation_loss(model, vdata):
    total, n = 0.0, 0
    for lines, weights in vdata:
        pred = model.predict(lines)
        labels = model._bufs["labels"][:lines.n_lines].clone()
        loss, _ = fm_ops.fm_loss_grad(pred, labels, weights, model.loss_type)
        total += float(loss) * lines.n_lines
        n += lines.n_lines
    return total / max(n, 1)


def train_sharded(cfg, monitor, stream_mb=None):
    """`train` under torchrun: synchronous data-parallel steps on the peer-sharded table"""
    import torch.distributed as dist
    from fast_tffm_b200.sharded import PeerShardedModel
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    device = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=device)
    model = PeerShardedModel(cfg, device=device).build(seed=int(time.time()) // 60)
    ckpt = None if cfg.log_dir is None else _ckpt_path(cfg)
    if ckpt is not None and latest_checkpoint(ckpt) is not None:
        model.restore(ckpt)
    say = print if rank == 0 else (lambda *a, **k: None)

    def cut(lines, weights):
        n = lines.n_lines
        lo, hi = rank * n // world, (rank + 1) * n // world
        return lines.select(torch.arange(lo, hi, device=device)), None if weights is None else weights[lo:hi]

    def validation_loss(vdata):
        acc = torch.zeros(2, dtype=torch.float64, device=device)
        for lines, weights in vdata:
            mine, w = cut(lines, weights)
            if mine.n_lines:
                pred = model.predict(mine)
                loss, _ = fm_ops.fm_loss_grad(pred, model._slots[0]["labels"][:mine.n_lines].clone(), w, cfg.loss_type)
                acc += torch.stack([loss.double() * mine.n_lines, torch.tensor(float(mine.n_lines), device=device)])
        dist.all_reduce(acc)
        return float(acc[0] / acc[1].clamp(min=1))

    vdata = None
    if cfg.validation_data_files:
        say('Preloading validation data...')
        vset = fm_data.DeviceDataset(cfg.validation_data_files, cfg.validation_weight_files, device)
        vdata = list(vset.batches(cfg.batch_size, 1, shuffle=False))
    say("========", "train", "(%d GPUs)" % world, "========")
    # every rank holds the files and walks the same shuffled order; a global batch is cut into equal slices
    seed = torch.tensor([int(time.time())], dtype=torch.int64, device=device)
    dist.broadcast(seed, 0)
    dataset = _training_set(cfg, device, stream_mb)
    if ckpt is not None:
        model.prepare_save()
    st = time.time()
    start_step = step_num = saved_step = model.global_step

    def slices():
        for lines, weights in dataset.batches(cfg.batch_size, cfg.num_epochs, shuffle=True, seed=int(seed.item())):
            n = lines.n_lines
            if n < world:       # a tail shorter than the number of GPUs cannot be cut into non-empty slices
                continue
            global_lines.append(n)
            yield cut(lines, weights) + (n,)

    # as in train(): results are read one step late, except at save / validation steps
    pipe = HostBatchPipeline(model)
    global_lines, pending = [], []

    def report(res, cur):
        loss = res.loss()
        if monitor:
            say('speed:', global_lines[res.index] / (time.time() - cur))
        say('-- Global Step: %d; Avg loss: %.5f;' % (start_step + res.index + 1, loss))

    cur = time.time()
    for res in pipe.run(slices()):
        step_num = model.global_step
        for old in pending:
            report(*old)
        pending = [(res, cur)]
        cur = time.time()
        if step_num % cfg.save_steps == 0:
            report(*pending.pop())
            cur = time.time()
            if ckpt is not None and model.save(ckpt, wait=False, skip_if_busy=True):   # every rank its own shard file
                saved_step = step_num
            if vdata is not None:
                v_loss = validation_loss(vdata)          # the same number on every rank: all stop together
                say('validation loss at step %d: %.8f' % (step_num, v_loss))
                if v_loss < cfg.tolerance:
                    say('Loss on validation data set is below tolerance. Training completed.')
                    break
    for old in pending:
        report(*old)
    total = time.time() - st
    say('Average speed: ', (step_num - start_step) * cfg.batch_size / total, ' ex/s')
    if ckpt is not None:
        if saved_step != step_num:
            model.save(ckpt, wait=False)
        model.wait_for_save()
    say('Model saved to ', cfg.log_dir)
    dist.destroy_process_group()


def train(model, monitor, trace, stream_mb=None):
    vdata = None
    if model.validation_data_files:
        print('Preloading validation data...')
        vset = fm_data.DeviceDataset(model.validation_data_files, model.validation_weight_files, model.device)
        vdata = list(vset.batches(model.batch_size, 1, shuffle=False))
    events = []
    # the training files live in HBM; a batch is a shuffled index selection over the resident text
    dataset = _training_set(model, model.device, stream_mb)
    if model.log_dir is not None:
        model.prepare_save()
    st = time.time()
    start_step = model.global_step
    step_num = saved_step = start_step
    # steps are enqueued ahead of their results: the parse of batch s+1 runs under the table kernels of batch
    # s, and the line of step s is printed while step s+1 runs -- except at save / validation steps, which
    # need the table exactly as step s left it
    pipe = HostBatchPipeline(model)
    pending = []

    def report(res, cur):
        loss = res.loss()                       # raises FmParseError like the op would
        tend = time.time()
        if trace and len(events) < 1000:
            events.append({"name": "train_step", "ph": "X", "ts": cur * 1e6, "dur": (tend - cur) * 1e6,
                           "pid": 0, "tid": 0, "args": {"step": start_step + res.index + 1, "examples": res.n_lines}})
        if monitor:
            print('speed:', res.n_lines / (tend - cur))
        print('-- Global Step: %d; Avg loss: %.5f;' % (start_step + res.index + 1, loss))

    cur = time.time()
    for res in pipe.run(dataset.batches(model.batch_size, model.num_epochs, shuffle=True, seed=int(st))):
        step_num = model.global_step
        for old in pending:
            report(*old)
        pending = [(res, cur)]
        cur = time.time()
        if step_num % model.save_steps == 0:
            report(*pending.pop())
            cur = time.time()
            if model.log_dir is not None:
                # device snapshot now, file written in the background; dropped if the last one is still being written
                if model.save(_ckpt_path(model), wait=False, skip_if_busy=True):
                    saved_step = step_num
            if vdata is not None:
                v_loss = _validation_loss(model, vdata)
                print('validation loss at step %d: %.8f' % (step_num, v_loss))
                if v_loss < model.tolerance:
                    print('Loss on validation data set is below tolerance. Training completed.')
                    break
    for old in pending:
        report(*old)
    total = time.time() - st
    print('Average speed: ', (step_num - start_step) * model.batch_size / total, ' ex/s')
    if model.log_dir is not None:
        if saved_step != step_num:
            model.save(_ckpt_path(model), wait=False)
        model.wait_for_save()
    print('Model saved to ', model.log_dir)
    if trace:
        if not trace.endswith('.json'):
            trace += '.json'
        with open(trace, 'w') as f:
            json.dump({"traceEvents": events}, f)


def main():
    parser = argparse.ArgumentParser()
    parser.add_argument("task", choices=['train', 'predict', 'generate'])
    parser.add_argument("config_file", type=str)
    parser.add_argument("--dist", nargs=4, metavar=('JOB_NAME', 'TASK_INDEX', 'PS_HOSTS', 'WORKER_HOSTS'),
                        default=None, help="(TensorFlow parameter-server mode: not available in this build)")
    parser.add_argument("--protocol", default='grpc', help="(unused: no gRPC runtime here)")
    parser.add_argument("--wait-for-workers", type=int, help="(unused: no parameter-server mode)")
    parser.add_argument("-t", "--trace", metavar='OUTPUT_FILE_NAME',
                        help="Stores per-step timings as a chrome trace file")
    parser.add_argument("-m,", "--monitor", action="store_true", help="Prints execution speed to screen")
    parser.add_argument("--export_path", help="Specifies the location to which the model is to be exported.")
    parser.add_argument("--stream-mb", type=int, default=None, metavar='MB',
                        help="Stream the training files through the GPU in segments of MB megabytes instead of "
                             "keeping them resident (automatic when they exceed half of the free GPU memory)")
    args = parser.parse_args()

    if args.dist is not None:
        print("This build replaces the hot path only: --dist (TF parameter server) is not provided. Multi-GPU "
              "training: python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1 run_tffm.py "
              "train <cfg>.")
        sys.exit(2)
    if args.task == 'generate' and args.export_path is None:
        print("Export path is not specified. Use --export_path.")
        sys.exit(2)

    model = Model(args.config_file)
    if args.task in ('predict', 'generate') and model.log_dir is None:
        print("Missing log directory. Must include a checkpoint file.")
        sys.exit(1)
    if not torch.cuda.is_available():
        print("No CUDA device: the B200 build has no CPU fallback.")
        sys.exit(3)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        if args.task != 'train':
            print("predict and generate run on one GPU (the checkpoint of a sharded run is in the single-GPU layout).")
            sys.exit(2)
        train_sharded(model, args.monitor, args.stream_mb)
        return
    model.build(seed=int(time.time()))
    if model.log_dir is not None and latest_checkpoint(_ckpt_path(model)) is not None:
        model.restore(_ckpt_path(model))     # MonitoredTrainingSession(checkpoint_dir=log_dir) restores
    elif args.task in ('predict', 'generate'):
        print("No checkpoint in", model.log_dir)
        sys.exit(1)
    if args.task == 'generate':
        from fast_tffm_b200.serving import export_serving
        print('Exporting trained model to', args.export_path)
        export_serving(model, args.export_path)
        return
    print("========", args.task, "========")
    if args.task == 'train':
        train(model, args.monitor, args.trace, args.stream_mb)
    else:
        predict(model, args.stream_mb)


if __name__ == '__main__':
    main()
