"""Training loops: `fit` (single GPU), `fit_DP`, `fit_DDP`, `fit_MP` - same signatures and side effects as the
reference's utils/train_utils.py (`fit` :22-92, `fit_DP` :95-167, `fit_DDP` :170-248; MP reuses `fit`, train.py:63):
dataset -> seeded split -> loaders -> Adam + ReduceLROnPlateau('min', patience 2) -> epochs of
[batch -> step] with the mean of the last 10 losses recorded every 10 steps -> evaluate + scheduler per epoch ->
checkpoint `./checkpoints/<mode>.pth` + `./loss/<mode>/{train,val}_loss.pkl`.

What differs, deliberately: the step runs on the native engines; the loss is fetched from the
device every `LOG_EVERY` steps instead of `.item()` each step; in DDP the sampler gets `set_epoch`, the validation
loss is broadcast and EVERY rank steps the plateau scheduler (the reference steps rank 0 only, so replicas diverge
after the first LR drop); DP actually runs (the reference never moves the model to cuda:0 and forgets
`unsqueeze(1)`); `val_percent` is a percentage in every signature; output directories are created.
"""
from __future__ import annotations

import logging
import os
import sys
import time
from typing import Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist
from torch.utils.data import DataLoader

from ..utils import data as D
from ..utils.checkpoint import load_checkpoint, load_resume_state, save_checkpoint
from ..utils.metrics import LossRecorder
from ..utils.sched import PlateauScheduler
from .evaluate import evaluate
from .steppers import MaskAdapter, make_stepper, native_available

LOG_EVERY = 10


def _size() -> Sequence[int]:
    """Image size (W, H): the reference hard-codes [960, 640] (utils/train_utils.py:26); DPT_IMAGE_SIZE=WxH overrides it
    (tests, small GPUs)."""
    env = os.environ.get("DPT_IMAGE_SIZE")
    if env:
        w, h = env.lower().split("x")
        return (int(w), int(h))
    return D.DEFAULT_SIZE


def _uint8_input(model, device) -> bool:
    """DPT_UINT8_INPUT=1: the loader hands over raw uint8 pixels and uint8 masks, the native engine normalises them on
    the device (a quarter of the host->device bytes and no float conversion on the host).  Reference-variant engine only."""
    return (os.environ.get("DPT_UINT8_INPUT", "0") == "1" and getattr(model, "variant", "") == "ref"
            and native_available(model, device))


def _dataset(newsize, raw_uint8: bool = False):
    n_syn = int(os.environ.get("DPT_SYNTHETIC", "0"))
    return D.build_dataset(D.DIR_IMG, D.DIR_MASK, newsize, synthetic_length=n_syn, raw_uint8=raw_uint8)


def _loaders(train_set, val_set, batch_size: int, num_workers: int, drop_last_train: bool, sampler=None):
    args = dict(batch_size=batch_size, num_workers=num_workers, pin_memory=torch.cuda.is_available())
    train_loader = DataLoader(train_set, shuffle=sampler is None, sampler=sampler, drop_last=drop_last_train, **args)
    # drop_last like the reference's validation loader, unless that would leave nothing to evaluate
    val_loader = DataLoader(val_set, shuffle=False, drop_last=len(val_set) >= batch_size, **args)
    return train_loader, val_loader


def _resume_state(save_name: Optional[str], model, resume: bool):
    """`--resume`: reload `checkpoints/<mode>.pth` into `model` (BEFORE the engine copies the parameters into its
    arena) and return the side file's optimizer / scheduler / epoch / step / RNG state, or None."""
    if not resume or save_name is None:
        return None
    state = load_resume_state(save_name)
    if state is None:
        logging.info(f"--resume: no checkpoints/{save_name}.resume.pth, starting from scratch")
        return None
    load_checkpoint(model, save_name)
    return state


def _progress(iterable, rank: int, desc: str):
    """tqdm bar with images/s like the reference's loops (utils/train_utils.py:57,72), on rank 0 when stderr is a
    terminal or DPT_PROGRESS=1."""
    want = os.environ.get("DPT_PROGRESS")
    if rank != 0 or want == "0" or (want is None and not sys.stderr.isatty()):
        return iterable, None
    try:
        from tqdm import tqdm
    except ImportError:
        return iterable, None
    bar = tqdm(iterable, desc=desc, unit="batch", leave=False)
    return bar, bar


def _train_epochs(stepper, train_loader, val_loader, epochs: int, mode: str, learning_rate: float, *, rank: int = 0,
                  world: int = 1, sampler=None, save_name: Optional[str] = None, module_prefix: bool = False,
                  model=None, evaluate_on_all_ranks: bool = False, batch_size: int = 1, resume_state=None):
    """The epoch/step loop shared by all modes.  Returns (train_rows, val_rows)."""
    rec = LossRecorder(mode)
    sched = PlateauScheduler(learning_rate, set_lr=stepper.set_lr, mode="min", patience=2, verbose=rank == 0)
    global_step, first_epoch = 0, 0
    if resume_state is not None:
        # every rank restores the same state: optimizer moments, scheduler (and with it the learning rate), position
        if resume_state.get("optimizer") is not None and hasattr(stepper, "load_optimizer_state"):
            stepper.load_optimizer_state(resume_state["optimizer"])
        sched.load_state_dict(resume_state["scheduler"])
        global_step, first_epoch = int(resume_state["global_step"]), int(resume_state["epoch"])
        rec.train_rows, rec.val_rows = list(resume_state.get("train_rows", [])), list(resume_state.get("val_rows", []))
        if resume_state.get("rng") is not None:
            torch.set_rng_state(resume_state["rng"])
        if rank == 0:
            logging.info(f"Resumed {mode} at epoch {first_epoch}, step {global_step}, lr {sched.lr:.3e}")
    fixed_b = getattr(stepper, "B", None)
    partial_ok = bool(getattr(stepper, "supports_partial_batch", False))
    window = []
    for epoch in range(first_epoch, epochs):
        if sampler is not None and hasattr(sampler, "set_epoch"):
            sampler.set_epoch(epoch)
        if model is not None:
            model.train()
        epoch_t0, epoch_images = time.time(), 0
        loader, bar = _progress(train_loader, rank, f"epoch {epoch + 1}/{epochs}")
        for batch in loader:
            images, masks = batch["image"], batch["mask"]
            if fixed_b is not None and images.shape[0] != fixed_b and not partial_ok:
                continue  # fixed-shape DP / pipeline engines: the loader drops the incomplete last batch instead
            if getattr(stepper, "input_uint8", False):
                masks = masks.to(torch.uint8)
            elif getattr(stepper, "wants_float_masks", True) and fixed_b is not None:
                masks = masks.float()
            loss = stepper.train_step(images, masks)
            epoch_images += int(images.shape[0])
            global_step += 1
            if bar is not None and global_step % LOG_EVERY == 0:
                bar.set_postfix_str(f"{epoch_images * world / max(time.time() - epoch_t0, 1e-9):.1f} img/s")
            # the native engines return a view of ONE device buffer that the next step overwrites: snapshot it (a device
            # copy, no host sync) so the window really holds LOG_EVERY different losses (utils/train_utils.py:75-79)
            window.append(loss.detach().clone())
            if global_step % LOG_EVERY == 0:
                vals = [float(l) for l in window]  # one host sync per LOG_EVERY steps
                window = []
                if rank == 0:
                    for k, v in enumerate(vals):
                        rec.train_step(global_step - len(vals) + k + 1, v)
                if mode == "DDP" and global_step % (2 * LOG_EVERY) == 0:
                    print(f"At step {global_step} in rank {rank}, the training loss is {vals[-1]}")
        if torch.cuda.is_available():
            torch.cuda.synchronize()  # once per epoch: the native engines never block the host inside the loop
        train_dt = max(time.time() - epoch_t0, 1e-9)
        # validation after each epoch (rank 0 in the reference; broadcast here so every rank schedules alike)
        val_loss = float("nan")
        if rank == 0 or evaluate_on_all_ranks:
            val_loss = evaluate(stepper, val_loader)
        if world > 1 and not evaluate_on_all_ranks:
            t = torch.tensor([val_loss], dtype=torch.float64,
                             device="cuda" if dist.get_backend() == "nccl" else "cpu")
            dist.broadcast(t, src=0)
            val_loss = float(t.item())
        if rank == 0:
            rec.validation(global_step, val_loss)
            # observability the reference lacks: throughput of the epoch's training part (host clock, this
            # rank's samples x world), the validation loss and the learning rate in force, appended to logs/<mode>.log
            logging.info(f"epoch {epoch + 1}/{epochs}: {epoch_images * world / train_dt:.1f} train images/s "
                         f"(loader included), val loss {val_loss:.5f}, lr {sched.lr:.3e}, step {global_step}")
        if not np.isnan(val_loss):
            sched.step(val_loss)
        # resume point after every epoch: weights in the reference's format plus `<mode>.resume.pth` with the optimizer
        # moments, scheduler, position, curves and RNG state (the reference only ever writes weights, once, at the end)
        if save_name is not None and model is not None and os.environ.get("DPT_SAVE_EVERY_EPOCH", "1") == "1":
            opt = stepper.optimizer_state() if hasattr(stepper, "optimizer_state") else None  # collective in DDP
            if rank == 0:
                save_checkpoint(model, save_name, module_prefix=module_prefix, resume_state={
                    "epoch": epoch + 1, "global_step": global_step, "scheduler": sched.state_dict(), "optimizer": opt,
                    "train_rows": rec.train_rows, "val_rows": rec.val_rows, "rng": torch.get_rng_state()})
    if rank == 0:
        if save_name is not None and model is not None:
            save_checkpoint(model, save_name, module_prefix=module_prefix)
        rec.save()
    return rec.train_rows, rec.val_rows


# --------------------------------------------------------------------------------------------------------
def fit(model, criterion, epochs: int = 10, batch_size: int = 4, learning_rate: float = 1e-4, val_percent: float = 10.0,
        device=None, save_name: str = "singleGPU", mode: str = "singleGPU", resume: bool = False):
    """Single-device training (reference: utils/train_utils.py:22-92)."""
    newsize = _size()
    device = torch.device(device if device is not None else ("cuda:0" if torch.cuda.is_available() else "cpu"))
    state = _resume_state(save_name, model, resume)
    u8 = _uint8_input(model, device)
    dataset = _dataset(newsize, raw_uint8=u8)
    train_set, val_set = D.split_dataset(dataset, val_percent, seed=0)
    stepper, native = make_stepper(model, criterion, batch_size, newsize, device, lr=learning_rate, uint8_input=u8)
    # every batch is trained on, the incomplete last one too (utils/train_utils.py:40): the native engines build a second
    # program for the remainder size on first use
    train_loader, val_loader = _loaders(train_set, val_set, batch_size, num_workers=1 if len(dataset) > 64 else 0,
                                        drop_last_train=native and not getattr(stepper, "supports_partial_batch", False))
    logging.info(f"Using {device} ({'native sm_100a engine' if native else 'torch-op path'}) for {mode} training.")
    return _train_epochs(stepper, train_loader, val_loader, epochs, mode, learning_rate, save_name=save_name, model=model,
                         batch_size=batch_size, resume_state=state)


def fit_DP(model, criterion, epochs: int = 10, batch_size: int = 4, learning_rate: float = 1e-4,
           val_percent: float = 10.0, device_ids: Optional[Sequence[int]] = None, resume: bool = False):
    """Single-process data parallel training over `device_ids` (reference: utils/train_utils.py:95-167)."""
    newsize = _size()
    state = _resume_state("DP", model, resume)
    dataset = _dataset(newsize)
    train_set, val_set = D.split_dataset(dataset, val_percent, seed=0)
    n_dev = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if device_ids is None:
        device_ids = list(range(n_dev)) if n_dev else [0, 1]
    if n_dev >= 1 and native_available(model, "cuda:0"):
        from ..parallel.dp import NativeDP

        opt = "adam" if getattr(model, "variant", "ref") == "ref" else "rmsprop"
        stepper = NativeDP(model, batch_size, int(newsize[1]), int(newsize[0]), list(device_ids), lr=learning_rate, optimizer=opt)
        native = True
    else:
        from ..parallel.dp import TorchDP

        dev = torch.device("cuda:0" if n_dev else "cpu")
        model.to(dev)
        crit = MaskAdapter(criterion, binary_prob=not getattr(model, "returns_logits", False))
        stepper = TorchDP(model, crit, len(device_ids), lr=learning_rate, loss_scale=batch_size, device=dev)
        native = False
    train_loader, val_loader = _loaders(train_set, val_set, batch_size, num_workers=0, drop_last_train=native)
    logging.info(f"DataParallel over devices {list(device_ids)} ({'native' if native else 'torch-op'} path, "
                 f"{getattr(model, 'variant', 'ref')} variant).")
    return _train_epochs(stepper, train_loader, val_loader, epochs, "DP", learning_rate, save_name="DP", module_prefix=True,
                         model=model, batch_size=batch_size, resume_state=state)


def fit_DDP(world_size: int, model, criterion, epochs: int = 10, batch_size: int = 4, learning_rate: float = 1e-4,
            val_percent: float = 10.0, resume: bool = False):
    """One process per GPU (reference: utils/train_utils.py:170-248).  `world_size` is kept for signature
    compatibility; the process group's size is what is actually used (the reference passes the GPU count)."""
    state = _resume_state("DDP", model, resume)  # every rank reads the same files
    rank = dist.get_rank()
    world = dist.get_world_size()
    local = int(os.environ.get("LOCAL_RANK", rank))
    newsize = _size()
    u8 = torch.cuda.is_available() and dist.get_backend() == "nccl" and _uint8_input(model, f"cuda:{local}")
    dataset = _dataset(newsize, raw_uint8=u8)
    train_set, val_set = D.split_dataset(dataset, val_percent, seed=0)  # identical split on all ranks
    sampler = D.ShardedSampler(len(train_set), rank, world, shuffle=True, seed=0)
    lr = learning_rate * world  # utils/train_utils.py:199
    use_cuda = torch.cuda.is_available() and dist.get_backend() == "nccl"
    if use_cuda and native_available(model, f"cuda:{local}"):
        from ..parallel.ddp import FusedDDP

        dt = "uint8" if u8 else "float32"
        stepper = FusedDDP(model, batch_size, int(newsize[1]), int(newsize[0]), f"cuda:{local}", lr=lr, input_dtype=dt,
                           target_dtype=dt)
        native = True
    else:
        from ..parallel.ddp import TorchDDP

        dev = torch.device(f"cuda:{local}" if use_cuda else "cpu")
        model.to(dev)
        crit = MaskAdapter(criterion, binary_prob=not getattr(model, "returns_logits", False))
        stepper = TorchDDP(model, crit, lr=lr, loss_scale=batch_size, device=dev)
        native = False
    train_loader, val_loader = _loaders(train_set, val_set, batch_size, num_workers=0, sampler=sampler,
                                        drop_last_train=not getattr(stepper, "supports_partial_batch", False))
    if rank == 0:
        logging.info(f"DistributedDataParallel, world {world} ({'native fused' if native else 'torch-op'} path).")
    return _train_epochs(stepper, train_loader, val_loader, epochs, "DDP", lr, rank=rank, world=world, sampler=sampler,
                         save_name="DDP", module_prefix=True, model=model, batch_size=batch_size, resume_state=state)
