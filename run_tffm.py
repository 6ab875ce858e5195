#!/usr/bin/env python
"""run_tffm.py -- the reference's command line (run_tffm.py:123-231) over the B200 hot path.

    python run_tffm.py train   sample.cfg [-m] [-t TRACE.json]
    python run_tffm.py predict sample.cfg

Same config file, same stdout lines ("-- Global Step: %d; Avg loss: %.5f;",
"Average speed: ...  ex/s", "Model saved to ..."), same <file>_score outputs.  Not carried over
(TensorFlow-runtime features, out of scope): `generate` / --export_path (SavedModel), --dist /
--protocol / --wait-for-workers (gRPC parameter server; multi-GPU runs go through torchrun, see
DESIGN.md section 6).
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


def _ckpt_path(model):
    return os.path.join(model.log_dir, "model.ckpt.pt")


def predict(model):
    for data_file in model.predict_files:
        lines = fm_data._read_lines(data_file)
        with open(data_file + '_score', 'w') as f:
            for i in range(0, len(lines), model.batch_size):
                for score in model.predict(lines[i:i + model.batch_size]).cpu().numpy():
                    f.write(str(score) + '\n')
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


def train(model, monitor, trace):
    vdata = None
    if model.validation_data_files:
        print('Preloading validation data...')
        vset = fm_data.DeviceDataset(model.validation_data_files, model.validation_weight_files, model.device)
        vdata = list(vset.batches(model.batch_size, 1, shuffle=False))
    events = []
    st = time.time()
    start_step = model.global_step
    step_num = start_step
    # the training files live in HBM; a batch is a shuffled index selection over the resident text
    dataset = fm_data.DeviceDataset(model.train_files, model.weight_files, model.device)
    for lines, weights in dataset.batches(model.batch_size, model.num_epochs, shuffle=True, seed=int(st)):
        cur = time.time()
        loss = model.train_step(lines, weights)
        step_num = model.global_step
        tend = time.time()
        if trace and len(events) < 1000:
            events.append({"name": "train_step", "ph": "X", "ts": cur * 1e6, "dur": (tend - cur) * 1e6,
                           "pid": 0, "tid": 0, "args": {"step": step_num, "examples": lines.n_lines}})
        if monitor:
            print('speed:', lines.n_lines / (tend - cur))
        print('-- Global Step: %d; Avg loss: %.5f;' % (step_num, loss))
        if model.log_dir is not None and step_num % model.save_steps == 0:
            model.save(_ckpt_path(model))
        if vdata is not None and step_num % model.save_steps == 0:
            v_loss = _validation_loss(model, vdata)
            print('validation loss at step %d: %.8f' % (step_num, v_loss))
            if v_loss < model.tolerance:
                print('Loss on validation data set is below tolerance. Training completed.')
                break
    total = time.time() - st
    print('Average speed: ', (step_num - start_step) * model.batch_size / total, ' ex/s')
    if model.log_dir is not None:
        model.save(_ckpt_path(model))
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
    parser.add_argument("--export_path", help="(SavedModel export: not available in this build)")
    args = parser.parse_args()

    if args.task == 'generate' or args.dist is not None:
        print("This build replaces the hot path only: 'generate' (TF SavedModel) and --dist (TF parameter "
              "server) are not provided. Multi-GPU training: torchrun bench.py / fast_tffm_b200.sharded.")
        sys.exit(2)

    model = Model(args.config_file)
    if args.task == 'predict' and model.log_dir is None:
        print("Missing log directory. Must include a checkpoint file.")
        sys.exit(1)
    if not torch.cuda.is_available():
        print("No CUDA device: the B200 build has no CPU fallback.")
        sys.exit(3)
    model.build(seed=int(time.time()))
    if model.log_dir is not None and os.path.exists(_ckpt_path(model)):
        model.restore(_ckpt_path(model))     # MonitoredTrainingSession(checkpoint_dir=log_dir) restores
    elif args.task == 'predict':
        print("No checkpoint in", model.log_dir)
        sys.exit(1)
    print("========", args.task, "========")
    if args.task == 'train':
        train(model, args.monitor, args.trace)
    else:
        predict(model)


if __name__ == '__main__':
    main()
