"""Pipeline model parallelism (`-t MP`) for 2..N stages with cross-stage skip connections.

Reference: `UNet(pipe=True)` (model/unet_model.py:14-53) hard-wires encoder+mid on cuda:0 and decoder+head on
cuda:1, splits the batch in two and relies on asynchronous launches from one Python thread for overlap; five
tensors per micro-batch cross the boundary through `.to('cuda:1')`, and backward is whatever autograd does.

Here the nine blocks of the UNet (enc1..enc4, mid, dec1..dec4+head) are cut into contiguous stages by a
cost-based partitioner (the reference's 35 % / 65 % split is one possible cut), one process per stage.  Every
tensor that crosses a cut - the main activation AND the skip tensors - goes DIRECTLY from the stage that
produces it to the stage that consumes it (NVSwitch is uniform, no hop-by-hop relays) over NCCL point-to-point
(`isend` / `irecv`) on a per-pair communicator, so a late consumer never blocks traffic to an early one; NCCL
runs on its own streams and is chained to the compute stream with events, so transfers overlap compute.
Schedules: GPipe (all forwards, then all backwards; the loss is taken over the WHOLE batch exactly as the
reference does after concatenating its micro-batch outputs) and 1F1B (per-micro-batch loss statistics).

The stage work is done by a backend: `TorchStage` (torch ops + autograd; CPU/gloo plumbing and oracle) or
`NativeStage` (engine/stage_engine.py, the sm_100a kernels).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

NUM_BLOCKS = 9  # enc1..enc4, mid, dec1..dec4(+head)
BLOCK_NAMES = ["enc1", "enc2", "enc3", "enc4", "mid", "dec1", "dec2", "dec3", "dec4"]


@dataclass(frozen=True)
class Flow:
    name: str       # "main3" = output of block 3, "skip1" = skip tensor of encoder level 1
    producer: int   # block index
    consumer: int   # block index


def flows() -> List[Flow]:
    out = [Flow(f"main{k}", k, k + 1) for k in range(NUM_BLOCKS - 1)]
    out += [Flow(f"skip{l}", l, NUM_BLOCKS - 1 - l) for l in range(4)]
    return out


def partition(costs: Sequence[float], n_stages: int) -> List[int]:
    """Contiguous partition of `costs` into n_stages minimising the most expensive stage (DP over cut points).
    Returns the stage index of every block."""
    n = len(costs)
    if not 1 <= n_stages <= n:
        raise ValueError(f"cannot cut {n} blocks into {n_stages} stages")
    pre = [0.0]
    for c in costs:
        pre.append(pre[-1] + c)
    INF = float("inf")
    best = [[INF] * (n + 1) for _ in range(n_stages + 1)]
    cut = [[0] * (n + 1) for _ in range(n_stages + 1)]
    best[0][0] = 0.0
    for s in range(1, n_stages + 1):
        for j in range(s, n + 1):
            for i in range(s - 1, j):
                v = max(best[s - 1][i], pre[j] - pre[i])
                if v < best[s][j]:
                    best[s][j], cut[s][j] = v, i
    bounds, j = [n], n
    for s in range(n_stages, 0, -1):
        j = cut[s][j]
        bounds.append(j)
    bounds = bounds[::-1]
    stage_of = []
    for s in range(n_stages):
        stage_of += [s] * (bounds[s + 1] - bounds[s])
    return stage_of


def reference_partition() -> List[int]:
    """The reference's fixed 2-stage cut: encoder + mid | decoder + head (model/unet_model.py:15-20)."""
    return [0] * 5 + [1] * 4


def micro_batch_sizes(batch: int, n_micro: int) -> List[int]:
    if batch < n_micro:
        n_micro = batch  # the reference crashes for B=1 (x.split(0)); we just use fewer micro-batches
    base, rem = divmod(batch, n_micro)
    return [base + (1 if i < rem else 0) for i in range(n_micro)]


class StageBackend:
    """Interface a pipeline stage implements (one instance per process)."""

    blocks: Tuple[int, int]

    def input_buffer(self, flow: str, mb: int) -> torch.Tensor: ...          # where an incoming activation lands
    def output_tensor(self, flow: str, mb: int) -> torch.Tensor: ...         # produced activation to send
    def grad_in_buffer(self, flow: str, mb: int) -> torch.Tensor: ...        # incoming gradient of an output
    def grad_out_tensor(self, flow: str, mb: int) -> torch.Tensor: ...       # produced gradient of an input
    def load_micro_batch(self, mb: int, images, masks) -> None: ...
    def forward(self, mb: int) -> None: ...
    def finalize_loss(self, mbs: Sequence[int]) -> None: ...                 # last stage only
    def backward(self, mb: int) -> None: ...
    def begin_step(self) -> None: ...
    def optimizer_step(self) -> None: ...
    def loss(self) -> torch.Tensor: ...


class PipelineStage:
    """Schedules one stage's micro-batches and moves boundary tensors over per-pair process groups.

    With a block-granular backend (NativeStage) a micro-batch's forward is run block by block: every incoming tensor is
    awaited right before the first block that reads it (a skip tensor is not needed until its decoder block) and every
    outgoing tensor is sent as soon as the block that produces it has been launched (the 37 MB skip tensor of enc1 is on
    the wire while enc2..mid still compute), and the same for gradients in backward.  Receives are posted at the start
    of the micro-batch in exactly the order the peer sends, so NCCL's per-pair FIFO matching is never out of step.
    `timeline=True` records CUDA events around every compute segment; `timeline_summary()` turns them into busy / idle
    milliseconds per step (idle = pipeline bubble + exposed transfers)."""

    def __init__(self, backend: StageBackend, stage: int, stage_of: Sequence[int], n_micro: int,
                 schedule: str = "gpipe", pair_groups: Optional[Dict[Tuple[int, int], dist.ProcessGroup]] = None,
                 stage_ranks: Optional[Sequence[int]] = None, timeline: bool = False):
        if schedule not in ("gpipe", "1f1b"):
            raise ValueError("schedule must be 'gpipe' or '1f1b'")
        self.be = backend
        self.stage = stage
        self.stage_of = list(stage_of)
        self.n_stages = max(stage_of) + 1
        self.n_micro = n_micro
        self.schedule = schedule
        self.pair_groups = pair_groups or {}
        self.stage_ranks = list(stage_ranks) if stage_ranks is not None else list(range(self.n_stages))
        fl = flows()
        self.recv_flows = [f for f in fl if self.stage_of[f.consumer] == stage and self.stage_of[f.producer] != stage]
        self.send_flows = [f for f in fl if self.stage_of[f.producer] == stage and self.stage_of[f.consumer] != stage]
        # issue order = production order on the sending side: forward by producer block, backward by descending consumer
        self.recv_flows.sort(key=lambda f: f.producer)
        self.send_flows.sort(key=lambda f: f.producer)
        self.blocks = [b for b in range(NUM_BLOCKS) if self.stage_of[b] == stage]
        self.granular = bool(getattr(backend, "block_granular", False))
        self.is_first = stage == 0
        self.is_last = stage == self.n_stages - 1
        self._pending: List = []
        self.timeline = timeline
        self._events: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    # ---- transfers -------------------------------------------------------------------------------------
    # Activations (forward) and gradients (backward) travel on DIFFERENT communicators.  NCCL executes the
    # point-to-point ops of one communicator in issue order and a send of more than the FIFO size only completes
    # once the peer's matching receive runs; with one communicator per pair, 1F1B interleaves "send activation k+1"
    # before "receive gradient k" on one side and "send gradient k" before "receive activation k+1" on the other:
    # a rendezvous deadlock at real tensor sizes (small test tensors slip through NCCL's eager FIFO).
    def _group(self, a: int, b: int, direction: str):
        if direction == "bwd" and "bwd" in self.pair_groups:
            return self.pair_groups["bwd"]
        return self.pair_groups.get((min(a, b), max(a, b)))

    def _irecv(self, t: torch.Tensor, peer: int, direction: str):
        return dist.irecv(t, src=self.stage_ranks[peer], group=self._group(self.stage, peer, direction))

    def _isend(self, t: torch.Tensor, peer: int, direction: str) -> None:
        self._pending.append(dist.isend(t, dst=self.stage_ranks[peer], group=self._group(self.stage, peer, direction)))

    def _segment(self, name: str, fn) -> None:
        if not self.timeline:
            fn()
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        self._events.append((name, e0, e1))

    def _fwd(self, mb: int, evaluate: bool = False):
        be = self.be
        # post every receive of this micro-batch now (the NCCL streams run ahead of the compute stream) ...
        works = {f.name: self._irecv(be.input_buffer(f.name, mb), self.stage_of[f.producer], "fwd") for f in self.recv_flows}
        if not self.granular:
            for w in works.values():
                w.wait()  # the compute stream waits for the transfer (NCCL: stream-ordered, does not block the host)
            self._segment(f"fwd{mb}", lambda: (be.eval_forward(mb) if evaluate and hasattr(be, "eval_forward") else be.forward(mb)))
            for f in self.send_flows:
                self._isend(be.output_tensor(f.name, mb), self.stage_of[f.consumer], "fwd")
            return
        for b in self.blocks:
            for f in self.recv_flows:  # ... and make the compute stream wait for each one where it is first read
                if f.consumer == b:
                    works[f.name].wait()
            self._segment(f"fwd{mb}.b{b}", lambda: be.forward_block(mb, b, evaluate))
            for f in self.send_flows:
                if f.producer == b:
                    self._isend(be.output_tensor(f.name, mb), self.stage_of[f.consumer], "fwd")

    def _bwd(self, mb: int):
        be = self.be
        # gradients arrive in the order the consumer stages produce them: descending consumer block
        order = sorted(self.send_flows, key=lambda f: -f.consumer)
        works = {f.name: self._irecv(be.grad_in_buffer(f.name, mb), self.stage_of[f.consumer], "bwd") for f in order}
        if not self.granular:
            for w in works.values():
                w.wait()
            self._segment(f"bwd{mb}", lambda: be.backward(mb))
            for f in sorted(self.recv_flows, key=lambda f: -f.consumer):
                self._isend(be.grad_out_tensor(f.name, mb), self.stage_of[f.producer], "bwd")
            return
        for b in reversed(self.blocks):
            for f in order:
                if f.producer == b:
                    works[f.name].wait()
            self._segment(f"bwd{mb}.b{b}", lambda: be.backward_block(mb, b))
            for f in self.recv_flows:
                if f.consumer == b:
                    self._isend(be.grad_out_tensor(f.name, mb), self.stage_of[f.producer], "bwd")

    # ---- schedules -------------------------------------------------------------------------------------
    def train_step(self) -> Optional[torch.Tensor]:
        """One optimizer step over the micro-batches already loaded into the backend."""
        self.be.begin_step()
        M = self.n_micro
        if self.schedule == "gpipe":
            for mb in range(M):
                self._fwd(mb)
            if self.is_last:
                self.be.finalize_loss(list(range(M)))  # whole-batch loss statistics, like the reference
            for mb in reversed(range(M)):
                self._bwd(mb)
        else:  # 1F1B: warm-up forwards, steady one-forward-one-backward, cool-down backwards
            warm = min(M, self.n_stages - 1 - self.stage)
            f = b = 0
            for _ in range(warm):
                self._fwd(f)
                f += 1
            while f < M:
                self._fwd(f)
                if self.is_last:
                    self.be.finalize_loss([f])
                f += 1
                self._bwd(b)
                b += 1
            while b < M:
                self._bwd(b)
                b += 1
        for w in self._pending:
            w.wait()
        self._pending = []
        self._segment("update", self.be.optimizer_step)
        return self.be.loss() if self.is_last else None

    def forward_only(self, evaluate: bool = True) -> None:
        for mb in range(self.n_micro):
            self._fwd(mb, evaluate=evaluate)
        if self.is_last:
            self.be.finalize_loss(list(range(self.n_micro)))
        for w in self._pending:
            w.wait()
        self._pending = []

    def timeline_summary(self, steps: int) -> Dict[str, float]:
        """Per-step busy ms (sum of the compute segments) by phase; call after synchronising.  The caller compares it
        with the measured step time: step - busy = idle (pipeline bubble + transfers the schedule could not hide)."""
        out: Dict[str, float] = {"fwd_ms": 0.0, "bwd_ms": 0.0, "update_ms": 0.0}
        for name, e0, e1 in self._events:
            key = "fwd_ms" if name.startswith("fwd") else "bwd_ms" if name.startswith("bwd") else "update_ms"
            out[key] += e0.elapsed_time(e1)
        self._events = []
        for k in out:
            out[k] /= max(steps, 1)
        out["busy_ms"] = out["fwd_ms"] + out["bwd_ms"] + out["update_ms"]
        return out


def stage_pairs(stage_of: Sequence[int]) -> List[Tuple[int, int]]:
    """Sorted list of stage pairs that exchange at least one tensor."""
    return sorted({(min(stage_of[f.producer], stage_of[f.consumer]), max(stage_of[f.producer], stage_of[f.consumer]))
                   for f in flows() if stage_of[f.producer] != stage_of[f.consumer]})


def make_pair_groups(stage_of: Sequence[int], stage_ranks: Sequence[int], explicit: bool = False,
                     device: Optional[torch.device] = None) -> Dict[Tuple[int, int], dist.ProcessGroup]:
    """Give every communicating pair of stages its own communicator (own NCCL stream), so a late consumer never
    blocks traffic to an early one.

    Default (`explicit=False`): point-to-point ops go through the world group, where ProcessGroupNCCL lazily creates
    one dedicated 2-rank communicator per (low, high) rank pair.  Those are created here, up front, by a tiny
    exchange over every pair in one global order (deadlock-free, and it keeps communicator creation out of the
    first training step).  The process group must NOT be eagerly initialised (no `device_id=`): in eager mode
    unbatched P2P ops share the group's single communicator and stream and serialise behind each other.
    `explicit=True` creates a real process group per pair instead (collective call, all ranks)."""
    pairs = stage_pairs(stage_of)
    if explicit:
        return {p: dist.new_group([stage_ranks[p[0]], stage_ranks[p[1]]]) for p in pairs}
    me = dist.get_rank()
    dev = device if device is not None else torch.device("cpu")
    # a second world-sized group carries the backward traffic: its lazily created per-pair communicators (and their
    # streams) are distinct from the forward ones
    bwd = dist.new_group(list(stage_ranks))
    for grp in (None, bwd):
        for a, b in pairs:  # same order on every rank
            ra, rb = stage_ranks[a], stage_ranks[b]
            if me == ra:
                dist.send(torch.ones(1, device=dev), dst=rb, group=grp)
                dist.recv(torch.empty(1, device=dev), src=rb, group=grp)
            elif me == rb:
                dist.recv(torch.empty(1, device=dev), src=ra, group=grp)
                dist.send(torch.ones(1, device=dev), dst=ra, group=grp)
    return {"bwd": bwd}


# --------------------------------------------------------------------------------------------------------
# torch-op backend (CPU / gloo plumbing path, oracle for the native backend)
# --------------------------------------------------------------------------------------------------------
class TorchStage(StageBackend):
    """Runs blocks [lo, hi) of a reference-variant UNet with torch ops; autograd graphs are kept per micro-batch."""

    def __init__(self, model, lo: int, hi: int, mb_sizes: Sequence[int], height: int, width: int, device,
                 criterion=None, lr: float = 1e-4, weight_decay: float = 1e-8, loss_scale: float = 1.0,
                 whole_batch: bool = True):
        from ..models.unet import center_crop

        self._crop = center_crop
        self.model = model
        self.blocks = (lo, hi)
        self.device = torch.device(device)
        self.mb_sizes = list(mb_sizes)
        self.H, self.W = height, width
        self.criterion = criterion
        self.loss_scale = loss_scale
        self.whole_batch = whole_batch
        self.is_last = hi == NUM_BLOCKS
        self.params = [p for n, p in model.named_parameters() if self._owns(n)]
        for p in self.params:
            p.data = p.data.to(self.device)
        self.opt = torch.optim.Adam(self.params, lr=lr, weight_decay=weight_decay) if self.params else None
        self.inputs: Dict[Tuple[str, int], torch.Tensor] = {}
        self.outputs: Dict[Tuple[str, int], torch.Tensor] = {}
        self.grad_in: Dict[Tuple[str, int], torch.Tensor] = {}
        self.images: Dict[int, torch.Tensor] = {}
        self.masks: Dict[int, torch.Tensor] = {}
        self.probs: Dict[int, torch.Tensor] = {}
        self._loss = torch.zeros((), device=self.device)
        self._loss_terms: Dict[int, torch.Tensor] = {}
        widths = [32, 64, 128, 256]
        self._shape: Dict[str, Tuple[int, int, int]] = {}  # flow -> (C, H, W)
        for l in range(4):
            self._shape[f"main{l}"] = (widths[l], height >> (l + 1), width >> (l + 1))
            self._shape[f"skip{l}"] = (widths[l], height >> l, width >> l)
        self._shape["main4"] = (512, height >> 4, width >> 4)
        for i in range(1, 4):
            self._shape[f"main{4 + i}"] = (widths[4 - i], height >> (4 - i), width >> (4 - i))

    def _owns(self, name: str) -> bool:
        lo, hi = self.blocks
        for b in range(lo, hi):
            bn = BLOCK_NAMES[b]
            if bn.startswith("enc") and name.startswith(f"encoder.conv{bn[3]}."):
                return True
            if bn == "mid" and name.startswith("mid."):
                return True
            if bn.startswith("dec") and (name.startswith(f"decoder.conv{bn[3]}.") or name.startswith(f"decoder.deconv{bn[3]}.")):
                return True
            if bn == "dec4" and name.startswith("segmap."):
                return True
        return False

    def _buf(self, flow: str, mb: int) -> torch.Tensor:
        c, h, w = self._shape[flow]
        return torch.zeros(self.mb_sizes[mb], c, h, w, device=self.device)

    def input_buffer(self, flow, mb):
        t = self._buf(flow, mb).requires_grad_(True)
        self.inputs[(flow, mb)] = t
        return t.data

    def output_tensor(self, flow, mb):
        return self.outputs[(flow, mb)].detach().contiguous()

    def grad_in_buffer(self, flow, mb):
        t = self._buf(flow, mb)
        self.grad_in[(flow, mb)] = t
        return t

    def grad_out_tensor(self, flow, mb):
        g = self.inputs[(flow, mb)].grad
        return (g if g is not None else torch.zeros_like(self.inputs[(flow, mb)])).contiguous()

    def load_micro_batch(self, mb, images, masks):
        if images is not None:
            self.images[mb] = images.to(self.device).float()
        if masks is not None:
            self.masks[mb] = masks.to(self.device)

    def begin_step(self):
        if self.opt is not None:
            self.opt.zero_grad(set_to_none=True)
        self.inputs.clear()
        self.outputs.clear()
        self.grad_in.clear()
        self.probs.clear()
        self._loss_terms.clear()

    def forward(self, mb):
        lo, hi = self.blocks
        m = self.model
        x = self.images[mb] if lo == 0 else self.inputs[(f"main{lo - 1}", mb)]
        for b in range(lo, hi):
            if b < 4:
                conv = getattr(m.encoder, f"conv{b + 1}")(x)
                self.outputs[(f"skip{b}", mb)] = conv
                x = m.encoder.maxpool(conv)
            elif b == 4:
                x = m.mid(x)
            else:
                i = b - 4
                lvl = 4 - i
                skip = self.outputs.get((f"skip{lvl}", mb))
                if skip is None:
                    skip = self.inputs[(f"skip{lvl}", mb)]
                up = getattr(m.decoder, f"deconv{i}")(x)
                skip = self._crop(skip, up.shape[2], up.shape[3])
                x = getattr(m.decoder, f"conv{i}")(torch.cat((skip, up), dim=1))
            if b < NUM_BLOCKS - 1:
                self.outputs[(f"main{b}", mb)] = x
        if self.is_last:
            self.probs[mb] = m.sigmoid(m.segmap(x))

    def finalize_loss(self, mbs):
        """Loss over the given micro-batches (all of them for GPipe = the reference's whole-batch loss) and its
        gradient with respect to each micro-batch's probabilities, so backward can run one micro-batch at a time."""
        probs = [self.probs[mb] for mb in mbs]
        cat = torch.cat(probs, dim=0)
        masks = torch.cat([self.masks[mb] for mb in mbs], dim=0).float().reshape(cat.shape)
        loss = self.criterion(cat, masks)
        weight = sum(self.mb_sizes[mb] for mb in mbs) / float(sum(self.mb_sizes))
        dps = torch.autograd.grad(self.loss_scale * weight * loss, probs)
        for mb, dp in zip(mbs, dps):
            self._loss_terms[mb] = dp
        self._loss = loss.detach() * weight + (self._loss if (mbs[0] != 0) else 0)

    def backward(self, mb):
        tensors, grads = [], []
        if self.is_last:
            tensors.append(self.probs[mb])
            grads.append(self._loss_terms[mb])
        for (flow, m2), t in self.outputs.items():
            if m2 == mb and (flow, mb) in self.grad_in:
                tensors.append(t)
                grads.append(self.grad_in[(flow, mb)])
        if tensors:
            torch.autograd.backward(tensors, grads)

    def optimizer_step(self):
        if self.opt is not None:
            self.opt.step()

    def loss(self):
        return self._loss
