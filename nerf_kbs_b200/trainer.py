"""Training-step driver of the ray-batch hot path: what nerfstudio's ``Trainer.train_iteration`` does
around ``model.get_outputs`` for the reference's configs (ref:train.py:42-101 — 4096 rays per step,
Adam lr 1e-3 eps 1e-15 on the "proposal_networks" and "fields" groups, no scheduler), as one CUDA
graph per step:

    model(ray_bundle) -> losses -> backward -> [gradient exchange] -> Adam (which re-zeroes the gradients)

Parameters and gradients live in one flat arena (optim.ParamArena): the backward kernels accumulate parameter
gradients straight into it, so the step has no zero_grad pass, no fill launches and no AccumulateGrad copies.

The callbacks of ref:nerfacto.py:365-394 (weight-anneal before the step, ``step_cb`` after it) stay on
the host: the anneal value reaches the captured kernels through a device scalar, and the proposal
"updated" gating selects between two captured graphs (with / without proposal gradients).

Data parallelism: one process per GPU, every rank steps its own ray batch; the one exchange of
the path is the gradient mean before Adam.  Default on several GPUs: reduce-scatter + Adam + all-gather fused
in one kernel over NVLink peer (symmetric) memory (dp_optim.ShardedAdam -> nkbs_dp_adam_step), NCCL carrying
only barriers.  ``NKBS_DP_FUSED=0``: one NCCL sum-allreduce per parameter group over the arena's flat gradient
slice, then the local Adam with 1/world folded in (the checker route; gloo on the CPU for the host-logic tests)."""
from __future__ import annotations

import os

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from . import _lib
from .optim import ArenaAdam, ParamArena
from .rays import RayBundle

BIG_TENSOR = 1 << 18  # elements; gradients at least this large are all-reduced in place, the rest bucketed


# ----------------------------------------------------------------------------------------------
# gradient exchange (device-agnostic: exercised with gloo on CPU tensors by the tests)
# ----------------------------------------------------------------------------------------------
def allreduce_mean_(tensors: Sequence[torch.Tensor], group=None, world_size: Optional[int] = None) -> None:
    """In-place mean over ranks of every tensor in ``tensors`` (DDP gradient semantics)."""
    if not dist.is_available() or not dist.is_initialized():
        return
    ws = world_size if world_size is not None else dist.get_world_size(group)
    if ws == 1:
        return
    small = [t for t in tensors if t.numel() < BIG_TENSOR]
    big = [t for t in tensors if t.numel() >= BIG_TENSOR]
    inv = 1.0 / ws
    for t in big:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        t.mul_(inv)
    if small:
        flat = torch.cat([t.reshape(-1) for t in small])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.mul_(inv)
        off = 0
        for t in small:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t))
            off += n


def _detach_tree(x):
    """Same nesting with every tensor detached; ray-sample containers keep their bin arrays only."""
    if isinstance(x, torch.Tensor):
        return x.detach()
    if isinstance(x, dict):
        return {k: _detach_tree(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_detach_tree(v) for v in x)
    return x


def shard_rays(num_rays_total: int, rank: int, world_size: int) -> range:
    """Strong-scaling split used by the reference's datamanager: ``train_num_rays_per_batch // world_size``
    rays per rank (ref:dataparser/datamanager_z.py:225-231)."""
    per = num_rays_total // world_size
    return range(rank * per, (rank + 1) * per)


# ----------------------------------------------------------------------------------------------
class Trainer:
    """One training step of ``model`` (NerfactoModel / SemanticNerfWModel) per call."""

    def __init__(self, model, num_rays: int, lr: float = 1e-3, eps: float = 1e-15, betas=(0.9, 0.999),
                 use_graph: bool = True, process_group=None, graph_warmup: int = 3, seed: Optional[int] = None,
                 fused_loss: Optional[bool] = None, inject_jitter: bool = False, keep_grads: bool = False,
                 deterministic: Optional[bool] = None):
        """``inject_jitter``: the per-ray jitter of the three sampling levels is an input of the step
        (``train_step(..., jitter=[t0, t1, t2])``, each [R,1]) instead of being drawn on the device — parity tests feed
        the oracle the same draws.  ``keep_grads``: debugging / tests — the optimizer leaves the gradients in the arena
        (``p.grad`` readable after the step) and the step starts with a fill instead.  ``deterministic`` (default: env
        ``NKBS_DETERMINISTIC=1``): parameter gradients are accumulated in 64-bit fixed point (optim.ParamArena), so a
        step's result does not depend on the arrival order of warps / CTAs / streams — two runs from the same state
        and inputs give the same bits; costs one flush pass over the shadow per optimizer step."""
        self.model = model
        # weighted total straight out of the loss kernel (models.get_loss_total); NKBS_FUSED_LOSS=0 keeps the loss-dict route
        self.fused_loss = (os.environ.get("NKBS_FUSED_LOSS", "1") != "0") if fused_loss is None else bool(fused_loss)
        p0 = next(model.parameters())
        if p0.device.type != "cuda":
            raise RuntimeError("nkbs Trainer needs the model on a CUDA device (no CPU fallback)")
        self.device = p0.device
        self.R = int(num_rays)
        groups = model.get_param_groups()
        # one Adam per group, as nerfstudio's Optimizers does (same hyper-parameters in the reference)
        self.group = process_group
        self.world_size = dist.get_world_size(process_group) if (dist.is_available() and dist.is_initialized()) else 1
        # several GPUs: reduce-scatter + Adam + all-gather fused over NVLink peer memory (dp_optim.py);
        # NKBS_DP_FUSED=0 keeps the NCCL allreduce + per-rank Adam route
        self.sharded, self.opt = None, None
        self.deterministic = (os.environ.get("NKBS_DETERMINISTIC", "0") == "1") if deterministic is None else bool(deterministic)
        if self.world_size > 1 and os.environ.get("NKBS_DP_FUSED", "1") != "0":
            from .dp_optim import ShardedAdam
            self.arena = ShardedAdam.make_arena(groups, process_group, deterministic=self.deterministic)
            self.sharded = ShardedAdam(self.arena, lr=lr, betas=betas, eps=eps, process_group=process_group)
        else:
            self.arena = ParamArena(groups, deterministic=self.deterministic)
            self.opt = ArenaAdam(self.arena, lr=lr, betas=betas, eps=eps)
        self.keep_grads = bool(keep_grads)
        # Deferred optimizer step of the "fields" group: its Adam (a 0.6 GB stream over the main table, gradient and
        # moments) runs at the START of the next step on a side stream, beside the proposal levels' forward — which
        # reads none of those parameters — and is joined before the main field's forward.  Same updates in the same
        # order; only the schedule changes.  flush() applies a pending update (called for you before any forward of the
        # model outside train_step, by close() and by state readers).  NKBS_DEFER_ADAM=0 keeps Adam at the step's end.
        self.defer_adam = os.environ.get("NKBS_DEFER_ADAM", "1") != "0" and not self.keep_grads
        self._pending = False
        self._in_step = False
        self._join_adam = False
        self._adam_stream = torch.cuda.Stream(p0.device, priority=0)   # lowest priority: fills what the step leaves idle
        self._step_stream = torch.cuda.Stream(p0.device, priority=int(os.environ.get("NKBS_STEP_PRIO", "-1")))  # captured steps run at high priority
        self._hooks = [model.register_forward_pre_hook(self._pre_model_forward),
                       model.field.register_forward_pre_hook(self._pre_field_forward)]
        if hasattr(model, "register_state_dict_pre_hook"):  # a checkpoint must see the update that is still pending
            self._hooks.append(model.register_state_dict_pre_hook(lambda module, prefix, keep_vars: self.flush()))
        self._inv_world = torch.full((), 1.0 / self.world_size, device=p0.device) if self.world_size > 1 else None
        self.use_graph = use_graph
        self.graph_warmup = graph_warmup
        self.step = 0
        self._graphs: Dict = {}  # (updated, pending deferred Adam) -> captured step
        self._graph_loss: Dict[bool, Dict[str, torch.Tensor]] = {}
        self._graph_outputs: Dict[bool, Dict] = {}
        self.outputs: Dict = {}  # model outputs of the last step (graph replays: the captured static tensors)
        self._pool = None
        self._eager_steps = 0
        self._calls_per_graph: Dict[bool, int] = {}
        self.nkbs_calls = 0  # nkbs entry-point launches issued so far (graph replays count what they captured)
        self._h2d_done: Optional[torch.cuda.Event] = None
        dev = self.device
        C = getattr(model.config, "num_semantic_classes", 0)
        self.needs_labels = bool(getattr(model.config, "use_semantic", False)) or model.__class__.__name__ == "SemanticNerfWModel"
        # static device-side inputs of the step (CUDA-graph replay reads these addresses): views of ONE flat buffer, so a
        # batch that arrives packed in the same layout (``pack_batch``) is one copy — one H2D transfer per step from a
        # pinned host batch, one D2D copy from a device-resident one — instead of one per tensor
        spec = [("origins", (self.R, 3), torch.float32), ("directions", (self.R, 3), torch.float32),
                ("camera_indices", (self.R, 1), torch.int64), ("pixel_area", (self.R, 1), torch.float32),
                ("image", (self.R, 3), torch.float32)]
        if self.needs_labels:
            spec.append(("semantics_label", (self.R,), torch.int64))
        self._layout, off = {}, 0
        for k, shape, dt in spec:
            n = int(torch.tensor(shape).prod()) * torch.empty((), dtype=dt).element_size()
            self._layout[k] = (off, n, shape, dt)
            off += (n + 255) // 256 * 256
        self._flat = torch.zeros(off, dtype=torch.uint8, device=dev)
        self._flat_pin = torch.zeros(off, dtype=torch.uint8).pin_memory()
        self.buf = self._views(self._flat)
        self.pinned = self._views(self._flat_pin)
        n_levels = model.proposal_sampler.num_proposal_network_iterations + 1
        self.jitter = torch.zeros(n_levels, self.R, 1, device=dev) if inject_jitter else None
        from . import ops
        ops.loss_scratch(dev)  # created outside any capture
        self._unit = torch.ones((), device=dev)  # seed of the backward pass (backward() would fill one per step)
        # the collider's near / far planes are static inputs of the captured step as well: the graphs bake in their
        # addresses, so they are owned here and not by the collider's (evictable) cache
        col = getattr(model, "collider", None)
        self._planes = None
        if col is not None and hasattr(col, "near_plane"):
            self._planes = (torch.full((self.R, 1), float(col.near_plane), device=dev),
                            torch.full((self.R, 1), float(col.far_plane), device=dev))
        self.anneal_t = torch.ones((), device=dev)
        model.proposal_sampler.pdf_sampler.anneal_dev = self.anneal_t
        self._loss_slots = [torch.zeros((), pin_memory=True), torch.zeros((), pin_memory=True)]
        self._slot, self._loss_pending = 0, None
        self._copy_stream, self._flat_next, self._prefetched, self._prefetch_done, self._d2d_done = None, None, None, None, None
        self.loss_dict: Dict[str, torch.Tensor] = {}
        if seed is not None:
            torch.cuda.manual_seed(seed)

    # -- inputs ------------------------------------------------------------------------------------
    def _views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {k: flat[o:o + n].view(dt).view(shape) for k, (o, n, shape, dt) in self._layout.items()}

    def h2d_bytes(self) -> int:
        return sum(n for _, n, _, _ in self._layout.values())

    def _sources(self, ray_bundle: RayBundle, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        src = {"origins": ray_bundle.origins, "directions": ray_bundle.directions,
               "camera_indices": ray_bundle.camera_indices, "pixel_area": ray_bundle.pixel_area,
               "image": batch["image"]}
        if self.needs_labels:
            src["semantics_label"] = batch["semantics_label"]
        for k, t in src.items():
            if t is not None and tuple(t.shape) != tuple(self._layout[k][2]):
                raise ValueError(f"{k}: expected shape {tuple(self._layout[k][2])}, got {tuple(t.shape)}")
        return src

    def pack_batch(self, ray_bundle: RayBundle, batch: Dict[str, torch.Tensor], device=None) -> torch.Tensor:
        """One flat uint8 tensor holding the batch in the layout of the step's input buffer: pinned host memory
        (``device=None``: what a data loader hands to ``train_step_host``) or device memory.  ``train_step(packed=...)``
        loads it with a single copy."""
        flat = (torch.zeros(self._flat.numel(), dtype=torch.uint8).pin_memory() if device is None
                else torch.zeros(self._flat.numel(), dtype=torch.uint8, device=device))
        views = self._views(flat)
        for k, t in self._sources(ray_bundle, batch).items():
            if t is not None:
                views[k].copy_(t)
        return flat

    def prefetch_packed(self, flat_host: torch.Tensor) -> None:
        """Start the host -> device transfer of a LATER step's pinned packed batch now, on a copy stream, into a second
        device buffer: it overlaps whatever the step stream is running.  ``train_step(packed=<the same tensor>)`` then
        waits for the transfer and moves the batch into the step's input buffer with one device-side copy."""
        if flat_host.dtype != torch.uint8 or flat_host.numel() != self._flat.numel() or not flat_host.is_pinned():
            raise ValueError("prefetch_packed takes a pinned batch from Trainer.pack_batch")
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
            self._flat_next = torch.empty_like(self._flat)
        if self._d2d_done is not None:  # the staging buffer's previous content has been consumed
            self._copy_stream.wait_event(self._d2d_done)
        with torch.cuda.stream(self._copy_stream):
            self._flat_next.copy_(flat_host, non_blocking=True)
            self._prefetch_done = torch.cuda.Event()
            self._prefetch_done.record()
        self._prefetched = flat_host

    def load_packed(self, flat: torch.Tensor) -> None:
        if flat.dtype != torch.uint8 or flat.numel() != self._flat.numel():
            raise ValueError("not a batch packed by Trainer.pack_batch for this trainer")
        if flat is self._prefetched:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._prefetch_done)
            self._flat.copy_(self._flat_next, non_blocking=True)
            self._d2d_done = torch.cuda.Event()
            self._d2d_done.record()
            self._prefetched = None
            return
        if flat.device.type == "cpu" and not flat.is_pinned():
            self._wait_staging()
            self._flat_pin.copy_(flat)
            flat = self._flat_pin
            self._flat.copy_(flat, non_blocking=True)
            self._h2d_done = torch.cuda.Event()
            self._h2d_done.record()
            return
        self._flat.copy_(flat, non_blocking=True)

    def _wait_staging(self) -> None:
        if self._h2d_done is not None:  # the staging buffer may still feed the previous step's copy
            self._h2d_done.synchronize()
            self._h2d_done = None

    def load_batch(self, ray_bundle: RayBundle, batch: Dict[str, torch.Tensor]) -> None:
        """Copy one ray batch (host or device tensors) into the step's static input buffers.  Host tensors are
        gathered in the pinned staging buffer and cross in ONE transfer; device tensors are copied one by one."""
        src = self._sources(ray_bundle, batch)
        if all(t is None or t.device.type == "cpu" for t in src.values()):
            self._wait_staging()
            for k, t in src.items():
                if t is not None:
                    self.pinned[k].copy_(t)
            self._flat.copy_(self._flat_pin, non_blocking=True)
            self._h2d_done = torch.cuda.Event()
            self._h2d_done.record()
            return
        for k, t in src.items():
            if t is not None:
                self.buf[k].copy_(t, non_blocking=True)

    def _bundle(self) -> RayBundle:
        b = self.buf
        nears, fars = self._planes if self._planes is not None else (None, None)
        return RayBundle(origins=b["origins"], directions=b["directions"], pixel_area=b["pixel_area"],
                         camera_indices=b["camera_indices"], nears=nears, fars=fars, metadata={})

    # -- the step ----------------------------------------------------------------------------------
    def zero_grad(self) -> None:
        """Zero the arena's gradient buffer (the step itself never needs this: the optimizer leaves it zeroed)."""
        self.arena.G.zero_()
        if self.arena.F is not None:
            self.arena.F.zero_()

    def active_groups(self, updated: bool) -> List[str]:
        """Parameter groups that receive gradients in a step: the proposal networks only when the sampler's
        ``updated`` gating lets them train; torch.optim.Adam skips them otherwise (``grad is None``)."""
        return [k for k in self.arena.layout if updated or k != "proposal_networks"]

    def _pre_model_forward(self, module, args):
        if self._pending and not self._in_step:
            self.flush()  # somebody evaluates the model between steps: the parameters must be up to date

    def _pre_field_forward(self, module, args):
        if self._join_adam:  # the main field reads the parameters the side-stream Adam is writing
            torch.cuda.current_stream(self.device).wait_stream(self._adam_stream)
            self._join_adam = False

    def flush(self) -> None:
        """Apply the deferred optimizer step of the previous ``train_step`` now (no-op when nothing is pending)."""
        if self._pending:
            self._run_adam(self.deferred_groups())
            self._pending = False

    def deferred_groups(self) -> List[str]:
        return [k for k in self.arena.layout if k != "proposal_networks"]

    def optimizer_step(self, updated: bool = True) -> None:
        """Adam on every group that received gradients (eager callers; train_step schedules the groups itself)."""
        self._run_adam(self.active_groups(updated))

    def _run_adam(self, active: List[str]) -> None:
        if not active:
            return
        if self.sharded is not None:
            self.sharded.step(active)
            return
        if self.world_size > 1:
            self.arena.flush_fixed(active)  # deterministic mode: the allreduce reads G
            for name in active:
                dist.all_reduce(self.arena.group_slice(name, self.arena.G), op=dist.ReduceOp.SUM, group=self.group)
        self.opt.step(active, inv_scale=self._inv_world, zero_grads=not self.keep_grads)

    def _step_body(self, updated: bool, pending: bool = False) -> Dict[str, torch.Tensor]:
        if self.keep_grads:
            self.zero_grad()
        if pending:
            # the previous step's Adam of the deferred groups, beside the proposal levels of this step
            cur = torch.cuda.current_stream(self.device)
            self._adam_stream.wait_stream(cur)
            with torch.cuda.stream(self._adam_stream):
                self._run_adam(self.deferred_groups())
            self._join_adam = True
        jitter = list(self.jitter.unbind(0)) if self.jitter is not None else None
        self._in_step = True
        try:
            outputs = self.model(self._bundle(), jitter=jitter)
        finally:
            self._in_step = False
        if self._join_adam:  # a model that never called its field
            self._pre_field_forward(None, None)
        batch = {"image": self.buf["image"]}
        if self.needs_labels:
            batch["semantics_label"] = self.buf["semantics_label"]
        fused = self.model.get_loss_total(outputs, batch) if (self.fused_loss and hasattr(self.model, "get_loss_total")) else None
        if fused is not None:
            # the loss kernel weights its own gradients and the total is the root of the backward pass: nothing is
            # launched between the loss kernel and the first backward kernel except the unit seed
            loss, loss_dict = fused
        else:
            loss_dict = self.model.get_loss_dict(outputs, batch)
            loss = sum(loss_dict.values())
        torch.autograd.backward([loss], [self._unit])
        # detached views only: a reference to the live outputs would keep this step's autograd graph — and with it the
        # parameters' AccumulateGrad nodes, bound to THIS step's stream — alive into the next step's capture
        self.outputs = _detach_tree(outputs)
        del outputs
        if self.defer_adam:
            self._run_adam([k for k in self.active_groups(updated) if k not in self.deferred_groups()])
        else:
            self._run_adam(self.active_groups(updated))
        out = {k: v.detach() for k, v in loss_dict.items()}
        out["loss"] = loss.detach()
        return out

    def _capture(self, key) -> None:
        g = torch.cuda.CUDAGraph()
        if self._pool is None:
            self._pool = torch.cuda.graph_pool_handle()
        torch.cuda.synchronize(self.device)
        c0 = sum(_lib.call_counts.values())
        with torch.cuda.graph(g, pool=self._pool, stream=self._step_stream):
            out = self._step_body(*key)
        self._calls_per_graph[key] = sum(_lib.call_counts.values()) - c0
        self._graphs[key] = g
        self._graph_loss[key] = out
        self._graph_outputs[key] = self.outputs  # static tensors of the graph's pool: rewritten by every replay

    def close(self) -> None:
        """Drop the captured graphs and detach the parameters from the gradient arena.  Call before tearing the
        process group down: a CUDA graph that captured a collective keeps its communicator referenced."""
        self.flush()
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._graphs.clear()
        self._graph_loss.clear()
        self._graph_outputs.clear()
        self.outputs, self.loss_dict = {}, {}
        self.arena.release()

    def prepare(self) -> None:
        """Run the eager warm-up steps and capture every CUDA-graph variant up front (these are real
        training steps on whatever the static buffers hold), so that no later step pays for a capture."""
        if not self.use_graph:
            return
        sampler = self.model.proposal_sampler
        while self._eager_steps < self.graph_warmup:
            self.train_step()
        gated = self.model.__class__.__name__ != "SemanticNerfWModel"  # no step_cb there: always updated
        for key in ((True, False) if gated else (True,)):
            if (key, self._pending) not in self._graphs:
                sampler.force_updated = key
                try:
                    self.train_step()
                finally:
                    sampler.force_updated = None

    def train_step(self, ray_bundle: Optional[RayBundle] = None, batch: Optional[Dict[str, torch.Tensor]] = None,
                   jitter: Optional[Sequence[torch.Tensor]] = None, packed: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Run one step on the given batch (``ray_bundle`` + ``batch``, or ``packed``: a batch from ``pack_batch``), or
        on whatever the static buffers hold.  Returns the total loss as a device scalar; ``self.loss_dict`` holds the
        individual terms."""
        if packed is not None:
            self.load_packed(packed)
        elif ray_bundle is not None:
            self.load_batch(ray_bundle, batch)
        if jitter is not None:
            if self.jitter is None:
                raise ValueError("Trainer(inject_jitter=True) is needed to feed the jitter draws")
            for dst, src in zip(self.jitter, jitter):
                dst.copy_(src.reshape(dst.shape), non_blocking=True)
        m = self.model
        m.set_step(self.step)
        self.anneal_t.fill_(float(m.proposal_sampler._anneal))  # scalar travels as a kernel argument
        updated = bool(m.proposal_sampler.will_update())
        key = (updated, self._pending)
        if self.use_graph and self._eager_steps >= self.graph_warmup:
            if key not in self._graphs:
                self._capture(key)
            self._graphs[key].replay()
            self.nkbs_calls += self._calls_per_graph[key]
            out = self._graph_loss[key]
            self.outputs = self._graph_outputs[key]
            if updated:  # the eager path resets this inside the sampler; the replay has no host side
                m.proposal_sampler._steps_since_update = 0
        else:
            c0 = sum(_lib.call_counts.values())
            if self.use_graph:
                # warm-up iterations run on a side stream, as CUDA-graph capture requires
                s = torch.cuda.Stream(self.device)
                s.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(s):
                    out = self._step_body(*key)
                torch.cuda.current_stream(self.device).wait_stream(s)
            else:
                out = self._step_body(*key)
            self._eager_steps += 1
            self.nkbs_calls += sum(_lib.call_counts.values()) - c0
        self._pending = self.defer_adam
        m.after_step(self.step)
        self.step += 1
        self.loss_dict = out
        return out["loss"]

    def train_step_host(self, ray_bundle: Optional[RayBundle] = None, batch: Optional[Dict[str, torch.Tensor]] = None,
                        packed: Optional[torch.Tensor] = None, lag: int = 0,
                        prefetch: Optional[torch.Tensor] = None) -> float:
        """Host in, host out: H2D of the batch, the step, D2H of the loss (what a datamanager-driven loop sees per
        iteration).  ``packed``: a pinned batch from ``pack_batch`` (one transfer).

        ``lag=0``: returns this step's loss (the host waits for the step).  ``lag=1``: double-buffered read-back — the
        loss of every step is still copied to the host, but the call returns the loss of the PREVIOUS step (NaN on the
        first call) and waits only for that one, so the host enqueues step k+1 while step k runs and the GPU never idles
        on the round trip; ``drain_host()`` returns the last step's loss.  ``prefetch``: the NEXT step's pinned packed
        batch; its transfer is started as soon as this step is enqueued (``prefetch_packed``)."""
        loss = self.train_step(ray_bundle, batch, packed=packed)
        if prefetch is not None:
            self.prefetch_packed(prefetch)
        slot = self._loss_slots[self._slot]
        self._slot ^= 1
        slot.copy_(loss, non_blocking=True)
        if lag == 0:
            torch.cuda.current_stream(self.device).synchronize()
            self._loss_pending = None
            return float(slot)
        ev = torch.cuda.Event()
        ev.record()
        prev, self._loss_pending = self._loss_pending, (slot, ev)
        if prev is None:
            return float("nan")
        prev[1].synchronize()
        return float(prev[0])

    def drain_host(self) -> float:
        """Wait for the loss read-back still in flight (``train_step_host(lag=1)``) and return it."""
        prev, self._loss_pending = self._loss_pending, None
        if prev is None:
            return float("nan")
        prev[1].synchronize()
        return float(prev[0])
