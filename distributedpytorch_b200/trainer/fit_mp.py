"""Pipeline-parallel training (`python3 train.py -t MP`).

Reference: `fit(UNet(pipe=True))` (train.py:41,62-64): one process, encoder+mid on cuda:0, decoder+head on cuda:1,
two micro-batches, checkpoint written as `singleGPU.pth`.  Here `fit_MP` spawns one worker process per stage (the
launch command stays `python3 train.py -t MP`), each worker builds its stage of the pipeline (parallel/pipe.py),
rank 0 merges the per-stage parameters into one state_dict with the standard keys and saves `checkpoints/MP.pth`
(the reference's MP run overwrites singleGPU.pth)."""
from __future__ import annotations

import logging
import os
import socket
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch.utils.data import DataLoader

from ..parallel import pipe as P
from ..utils import data as D
from ..utils.checkpoint import save_checkpoint
from ..utils.metrics import LossRecorder
from ..utils.sched import PlateauScheduler
from .fit import LOG_EVERY, _dataset, _size


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank: int, world: int, port: int, model, criterion, epochs, batch_size, learning_rate, val_percent, n_micro,
            schedule, use_cuda):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["RANK"], os.environ["WORLD_SIZE"], os.environ["LOCAL_RANK"] = str(rank), str(world), str(rank)
    from ..utils.logging_utils import setup_logging

    setup_logging("MP", rank=rank)  # spawned workers start with an unconfigured root logger
    if use_cuda:
        configure_pipeline_runtime()
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world)  # lazy init: per-pair P2P communicators
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        _run_stage(rank, world, model, criterion, epochs, batch_size, learning_rate, val_percent, n_micro, schedule, use_cuda)
        dist.barrier()  # success path only: a stage that raised must not park in a barrier its peers never reach
    finally:
        dist.destroy_process_group()


def configure_pipeline_runtime() -> None:
    """Call before the process group and the first engine exist.  The persistent tcgen05 kernels own an SM each with a
    static tile assignment, so an NCCL send/recv kernel that lands on an SM stalls one CTA and with it the whole launch:
    leave DPT_SM_RESERVE SMs (default 16) to NCCL and cap its point-to-point kernels at that many channels (CTAs).
    Measured on 2 B200, batch 16, 4 micro-batches, 1F1B (profiles/modes_2gpu_sweep_r2.jsonl): no reserve 2.3x slower;
    reserve 8 / 4 channels 732 images/s, 8 / 8 925, 16 / 8 914, 16 / 16 1043, 24 / 12 986."""
    reserve = int(os.environ.get("DPT_SM_RESERVE", "16"))
    # 16 / 32 gave 1111 (profiles/modes_4gpu_r2.jsonl): NCCL's CTAs are small enough to share the reserved SMs two by two
    os.environ.setdefault("NCCL_MAX_NCHANNELS", str(max(2, 2 * reserve)))
    os.environ.setdefault("NCCL_MAX_P2P_NCHANNELS", str(max(2, 2 * reserve)))
    from ..ops import native as N

    if reserve > 0:
        N.check(N.lib().dpt_set_sm_reserve(reserve))


def _run_stage(rank, world, model, criterion, epochs, batch_size, learning_rate, val_percent, n_micro, schedule, use_cuda):
    newsize = _size()
    W, H = int(newsize[0]), int(newsize[1])
    dataset = _dataset(newsize)
    train_set, val_set = D.split_dataset(dataset, val_percent, seed=0)
    stage_of = P.partition([c for _, c in model.stage_costs()], world)
    groups = P.make_pair_groups(stage_of, list(range(world)),
                                device=torch.device("cuda", rank) if use_cuda else torch.device("cpu"))
    lo = stage_of.index(rank)
    hi = len(stage_of) - stage_of[::-1].index(rank)
    sizes = P.micro_batch_sizes(batch_size, n_micro)
    whole = schedule == "gpipe"
    if use_cuda:
        from ..engine.stage_engine import NativeStage

        be = NativeStage(model, lo, hi, sizes, H, W, f"cuda:{rank}", lr=learning_rate, loss_scale=float(batch_size),
                         whole_batch=whole)
        if rank == 0:
            logging.info(f"Pipeline: native stages ({model.variant} variant), cut {stage_of}, {len(sizes)} micro-batches, {schedule}.")
    else:
        be = P.TorchStage(model, lo, hi, sizes, H, W, "cpu", criterion=criterion, lr=learning_rate,
                          loss_scale=float(batch_size), whole_batch=whole)
    stage = P.PipelineStage(be, rank, stage_of, len(sizes), schedule, groups)
    is_first, is_last = rank == 0, rank == world - 1
    gen = torch.Generator().manual_seed(0)  # every stage walks the loader in the same order
    train_loader = DataLoader(train_set, batch_size=batch_size, shuffle=True, drop_last=True, generator=gen, num_workers=0)
    val_loader = DataLoader(val_set, batch_size=batch_size, shuffle=False, drop_last=True, num_workers=0)
    rec = LossRecorder("MP")

    def set_lr(lr):
        if hasattr(be, "set_lr"):
            be.set_lr(lr)
        elif be.opt is not None:
            for g in be.opt.param_groups:
                g["lr"] = lr

    sched = PlateauScheduler(learning_rate, set_lr=set_lr, patience=2, verbose=is_last)

    def feed(batch):
        off = 0
        for mb, b in enumerate(sizes):
            img = batch["image"][off:off + b] if is_first else None
            msk = batch["mask"][off:off + b] if is_last else None
            if msk is not None and use_cuda:
                msk = msk.to(torch.uint8) if model.variant == "milesial" else msk.float()
            be.load_micro_batch(mb, img, msk)
            off += b

    step, window = 0, []
    for epoch in range(epochs):
        for batch in train_loader:
            feed(batch)
            loss = stage.train_step()
            step += 1
            if is_last:
                window.append(loss.detach().clone())  # the engines return a view of ONE device buffer: snapshot it
                if step % LOG_EVERY == 0:
                    vals = [float(v) for v in window]
                    window = []
                    for k, v in enumerate(vals):
                        rec.train_step(step - len(vals) + k + 1, v)
        vals = []
        for batch in val_loader:
            feed(batch)
            stage.forward_only()
            if is_last:
                vals.append(float(be.eval_loss() if hasattr(be, "eval_loss") else be.loss()))
        t = torch.tensor([float(np.mean(vals)) if vals else float("nan")], dtype=torch.float64,
                         device=f"cuda:{rank}" if use_cuda else "cpu")
        dist.broadcast(t, src=world - 1)
        val_loss = float(t.item())
        if is_last:
            rec.validation(step, val_loss)
        if not np.isnan(val_loss):
            sched.step(val_loss)
    # merge the per-stage parameters into one standard state_dict on rank 0
    owned = be.owned_parameter_names() if use_cuda else {n for n, _ in model.named_parameters() if be._owns(n)}
    piece = {n: p.detach().cpu() for n, p in model.named_parameters() if n in owned}
    if model.variant == "milesial":  # BatchNorm buffers of the owned blocks travel with their parameters
        prefixes = {n.rsplit(".", 1)[0] for n in owned}
        piece.update({n: b.detach().cpu() for n, b in model.named_buffers() if n.rsplit(".", 1)[0] in prefixes})
    pieces = [None] * world
    dist.all_gather_object(pieces, piece)
    if is_last:
        rec.save()
    if rank == 0:
        merged = {}
        for p in pieces:
            merged.update(p)
        sd = {k: merged.get(k, v.detach().cpu()) for k, v in model.state_dict().items()}
        save_checkpoint(sd, "MP")
        logging.info(f"Pipeline ({world} stages {stage_of}, {len(sizes)} micro-batches, {schedule}) finished.")


def fit_MP(model, criterion, epochs: int = 10, batch_size: int = 4, learning_rate: float = 1e-4, val_percent: float = 10.0,
           n_stages: Optional[int] = None, n_micro: int = 2, schedule: str = "gpipe"):
    """Spawn one worker per pipeline stage and train.  Defaults reproduce the reference: 2 stages, 2 micro-batches."""
    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    use_cuda = n_gpu >= 2
    if n_stages is None:
        n_stages = int(os.environ.get("DPT_PIPE_STAGES", "2"))
    if use_cuda and n_stages > n_gpu:
        raise RuntimeError(f"{n_stages} pipeline stages need {n_stages} GPUs, found {n_gpu}")
    n_micro = int(os.environ.get("DPT_PIPE_MICRO", str(n_micro)))
    if getattr(model, "variant", "ref") != "ref" and not use_cuda:
        raise NotImplementedError("the torch-op pipeline backend (CPU plumbing path) covers the reference UNet variant only")
    schedule = os.environ.get("DPT_PIPE_SCHEDULE", schedule)
    model.cpu()
    port = _free_port()
    mp.spawn(_worker, args=(n_stages, port, model, criterion, epochs, batch_size, learning_rate, val_percent, n_micro,
                            schedule, use_cuda), nprocs=n_stages, join=True)
