"""worker for the world_size-2 tests of the sharded path (spawned processes import this)"""
import os

import numpy as np
import torch
import torch.distributed as dist


def run(rank, world, port, backend, lines, weights, cfg_kw, init, steps, out_dir, mode="exchange"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    if backend == "nccl":
        torch.cuda.set_device(rank)
        device, kernels = torch.device("cuda", rank), None
    else:
        from shard_backend_oracle import OracleShardKernels
        device, kernels = torch.device("cpu"), OracleShardKernels()
    dist.init_process_group(backend, rank=rank, world_size=world)
    from fast_tffm_b200.fm_model import Model
    from fast_tffm_b200.sharded import PeerShardedModel, ShardedModel
    if mode in ("peer", "merge"):
        m = PeerShardedModel(Model(**cfg_kw), device=device,
                             exchange="records" if mode == "merge" else "atomics").build(init_table=init)
    else:
        m = ShardedModel(Model(**cfg_kw), kernels=kernels, device=device).build(init_table=init)
    bs = cfg_kw["batch_size"]
    per = bs // world
    losses = []
    for s in range(steps):
        lo = s * bs + rank * per
        w = None if weights is None or s % 2 else np.float32(weights[lo:lo + per])
        losses.append(m.train_step(lines[lo:lo + per], w))
    pred = m.predict(lines[rank * per:(rank + 1) * per]).cpu().numpy()
    if mode == "peer":      # the same scores when the batch's parse ran ahead on the side stream
        mine = lines[rank * per:(rank + 1) * per]
        _, b = m.predict_async(mine, parsed=m.parse_ahead(mine))
        assert np.array_equal(b["pred"][:per].cpu().numpy(), pred)
    full = m.gather_full_table().cpu().numpy()
    if rank == 0:
        np.save(os.path.join(out_dir, "table.npy"), full)
        np.save(os.path.join(out_dir, "losses.npy"), np.float64(losses))
    np.save(os.path.join(out_dir, "pred%d.npy" % rank), pred)
    dist.barrier()
    dist.destroy_process_group()


def run_bad_line(rank, world, port, lines, cfg_kw, init, out_dir, mode):
    """rank 1's slice holds a line the parser rejects: the step must be void on EVERY rank"""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    from fast_tffm_b200.fm_model import Model
    from fast_tffm_b200.fm_ops import FmParseError
    from fast_tffm_b200.sharded import PeerShardedModel, ShardedModel
    if mode == "exchange":
        m = ShardedModel(Model(**cfg_kw), device=device).build(init_table=init)
    else:
        m = PeerShardedModel(Model(**cfg_kw), device=device,
                             exchange="records" if mode == "merge" else "atomics").build(init_table=init)
    per = cfg_kw["batch_size"] // world
    mine = list(lines[rank * per:(rank + 1) * per])
    if rank == 1:
        mine[7] = "1 3:oops"
    raised = False
    try:
        m.train_step(mine)
    except FmParseError as e:
        raised = "Invalid feature value" in str(e)
    after_bad = m.gather_full_table().cpu().numpy()
    loss = m.train_step(lines[rank * per:(rank + 1) * per])      # and the next, clean step works
    after_good = m.gather_full_table().cpu().numpy()
    np.save(os.path.join(out_dir, "bad%d.npy" % rank), np.array([float(raised), loss]))
    if rank == 0:
        np.save(os.path.join(out_dir, "after_bad.npy"), after_bad)
        np.save(os.path.join(out_dir, "after_good.npy"), after_good)
    dist.barrier()
    dist.destroy_process_group()


def run_checkpoint(rank, world, port, lines, cfg_kw, init, out_dir):
    """save in the background after step 2, keep training, restore into a fresh sharded model"""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    from fast_tffm_b200.fm_model import Model
    from fast_tffm_b200.sharded import PeerShardedModel
    m = PeerShardedModel(Model(**cfg_kw), device=device).build(init_table=init)
    per = cfg_kw["batch_size"] // world
    path = os.path.join(out_dir, "ckpt", "model.ckpt.pt")
    for s in range(3):
        lo = s * cfg_kw["batch_size"] + rank * per
        m.train_step(lines[lo:lo + per])
        if s == 1:
            at_save = m.gather_full_table().cpu().numpy()
            assert m.save(path, wait=False)
            assert m.save(path, wait=False, skip_if_busy=True) in (True, False)     # never blocks, never corrupts
    m.wait_for_save()
    m2 = PeerShardedModel(Model(**cfg_kw), device=device).build(seed=9)
    m2.restore(path)
    restored = m2.gather_full_table().cpu().numpy()
    if rank == 0:
        np.save(os.path.join(out_dir, "at_save.npy"), at_save)
        np.save(os.path.join(out_dir, "restored.npy"), restored)
        np.save(os.path.join(out_dir, "step.npy"), np.int64([m2.global_step]))
    dist.barrier()
    dist.destroy_process_group()
