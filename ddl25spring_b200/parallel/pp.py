"""Pipeline (model) parallelism: schedules as data + a stage executor, and the PP x DP mesh.

Covers lab/tutorial_1b/PP/1F1B:

* ``sequential``  – one batch in flight (intro_PP_1F1B.py:47-98);
* ``gpipe``       – all-forward-then-all-backward over M micro-batches, loss / M on the last stage
                    (intro_PP_1F1B_MB.py:48-141);
* ``one_f_one_b`` – a CORRECT 1F1B: warm-up forwards, steady 1F/1B alternation, cool-down backwards.
                    The reference's attempt (intro_PP_1F1B_MP.py:86-206) interleaves on stage 1 only and then
                    runs a second backward phase, so its ranks disagree on the message order and hang at
                    iteration 0 (lab/homework-1.ipynb cell 48);
* :class:`Mesh`   – ``pipeline_id = rank // S``, ``stage_id = rank % S`` (intro_PP_1F1B_MP.py:28-29) with the
                    per-stage DP groups created COLLECTIVELY (every rank calls every ``new_group``; the
                    reference does not).

A schedule is a list of :class:`Op` per stage; the executor just walks it.  Deadlock-freedom is a
property of the data: :func:`check_schedule` simulates the per-stage lists against each other and
is run in the unit tests for every (schedule, stages, micro-batches) combination, and every
distributed test runs under a timeout.

Transport: ``torch.distributed`` p2p on the tensors' own device — NCCL device-to-device on GPUs (no
CPU bounce as in the reference's ``.to("cpu")`` before every send), gloo on CPU.  Receives for the
next micro-batch are posted before computing the current one, so transfer overlaps compute.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Sequence

import torch
import torch.distributed as dist
from torch import nn

F, B = "F", "B"


@dataclass(frozen=True)
class Op:
    kind: str      # "F" or "B"
    mb: int        # micro-batch index


def gpipe(n_stages: int, stage: int, n_micro: int) -> list[Op]:
    return [Op(F, i) for i in range(n_micro)] + [Op(B, i) for i in range(n_micro)]


def sequential(n_stages: int, stage: int, n_micro: int = 1) -> list[Op]:
    return gpipe(n_stages, stage, 1)


def one_f_one_b(n_stages: int, stage: int, n_micro: int) -> list[Op]:
    warm = min(n_stages - 1 - stage, n_micro)
    ops = [Op(F, i) for i in range(warm)]
    f, b = warm, 0
    while f < n_micro:                       # steady state: one forward, one backward
        ops.append(Op(F, f)); f += 1
        ops.append(Op(B, b)); b += 1
    while b < n_micro:                       # cool-down
        ops.append(Op(B, b)); b += 1
    return ops


SCHEDULES: dict[str, Callable[[int, int, int], list[Op]]] = {"gpipe": gpipe, "1f1b": one_f_one_b, "sequential": sequential}


def check_schedule(name: str, n_stages: int, n_micro: int) -> int:
    """Simulate all stages with blocking-receive semantics; returns the number of time steps, raises
    ``RuntimeError`` on deadlock or on a stage consuming a message that is never produced."""
    scheds = [list(SCHEDULES[name](n_stages, s, n_micro)) for s in range(n_stages)]
    pc = [0] * n_stages
    done_f = [set() for _ in range(n_stages)]      # forwards finished per stage
    done_b = [set() for _ in range(n_stages)]
    steps = 0
    while any(pc[s] < len(scheds[s]) for s in range(n_stages)):
        progressed = False
        snapshot_f = [set(x) for x in done_f]
        snapshot_b = [set(x) for x in done_b]
        for s in range(n_stages):
            if pc[s] >= len(scheds[s]):
                continue
            op = scheds[s][pc[s]]
            if op.kind == F:
                ready = s == 0 or op.mb in snapshot_f[s - 1]
            else:
                ready = op.mb in snapshot_f[s] and (s == n_stages - 1 or op.mb in snapshot_b[s + 1])
            if ready:
                (done_f if op.kind == F else done_b)[s].add(op.mb)
                pc[s] += 1
                progressed = True
        steps += 1
        if not progressed:
            raise RuntimeError(f"schedule {name!r} deadlocks with {n_stages} stages / {n_micro} micro-batches at {pc}")
    for s in range(n_stages):
        if done_f[s] != set(range(n_micro)) or done_b[s] != set(range(n_micro)):
            raise RuntimeError(f"schedule {name!r}: stage {s} did not process every micro-batch")
    return steps


def p2p_program(name: str, n_stages: int, stage: int, n_micro: int) -> list[tuple]:
    """The communication / compute actions :meth:`PipelineStage.run` issues, in host order, as data:
    ``("recv", dir, mb)``, ``("compute", kind, mb)``, ``("send", dir, mb)`` with ``dir`` "fwd" (activations) or "bwd"
    (gradients).  Receives are pre-posted one op ahead exactly like the executor does."""
    ops = SCHEDULES[name](n_stages, stage, n_micro)
    first, last = stage == 0, stage == n_stages - 1
    prog, posted = [], set()

    def post(op: Op):
        if op.kind == F and not first and ("fwd", op.mb) not in posted:
            posted.add(("fwd", op.mb)); prog.append(("recv", "fwd", op.mb))
        if op.kind == B and not last and ("bwd", op.mb) not in posted:
            posted.add(("bwd", op.mb)); prog.append(("recv", "bwd", op.mb))

    if ops:
        post(ops[0])
    for i, op in enumerate(ops):
        if i + 1 < len(ops):
            post(ops[i + 1])
        prog.append(("compute", op.kind, op.mb))
        if op.kind == F and not last:
            prog.append(("send", "fwd", op.mb))
        if op.kind == B and not first:
            prog.append(("send", "bwd", op.mb))
    return prog


def check_stream_ordered_p2p(name: str, n_stages: int, n_micro: int, per_direction: bool = True) -> int:
    """Simulate the executor under NCCL semantics: the host never blocks, every p2p operation of a rank is queued in issue
    order on ONE stream per communicator, and a transfer happens only when the send and the matching receive are both at
    the head of their queues (rendezvous) and the send's payload has been computed.  ``per_direction=False`` puts
    activations and gradients on the same communicator - the configuration that hung on the GPUs; ``True`` is what
    :class:`Mesh` sets up.  Returns the number of simulation steps, raises ``RuntimeError`` on deadlock."""
    from collections import deque
    progs = [p2p_program(name, n_stages, s, n_micro) for s in range(n_stages)]
    comm = (lambda d: d) if per_direction else (lambda d: "all")
    queues = [dict() for _ in range(n_stages)]                 # stage -> communicator -> deque of (type, dir, mb)
    computes = [deque() for _ in range(n_stages)]
    for s, prog in enumerate(progs):
        for act in prog:
            if act[0] == "compute":
                computes[s].append(act)
            else:
                queues[s].setdefault(comm(act[1]), deque()).append(act)
    received = [set() for _ in range(n_stages)]                # (dir, mb) that arrived at stage s
    computed = [set() for _ in range(n_stages)]                # (kind, mb) finished at stage s
    steps = 0
    while any(computes[s] or any(q for q in queues[s].values()) for s in range(n_stages)):
        progressed = False
        for s in range(n_stages):                              # compute stream: in order, gated by its input message
            if computes[s]:
                _, kind, mb = computes[s][0]
                need = ("fwd", mb) if kind == F else ("bwd", mb)
                has_input = (kind == F and s == 0) or (kind == B and s == n_stages - 1) or need in received[s]
                if has_input and (kind == F or (F, mb) in computed[s]):
                    computed[s].add((kind, mb)); computes[s].popleft(); progressed = True
        for s in range(n_stages):                              # transfers: both heads must match
            for c, q in queues[s].items():
                if not q or q[0][0] != "send":
                    continue
                _, d, mb = q[0]
                if ((F if d == "fwd" else B), mb) not in computed[s]:
                    continue
                peer = s + 1 if d == "fwd" else s - 1
                pq = queues[peer].get(c)
                if pq and pq[0] == ("recv", d, mb):
                    q.popleft(); pq.popleft(); received[peer].add((d, mb)); progressed = True
        steps += 1
        if not progressed:
            heads = [{c: (q[0] if q else None) for c, q in queues[s].items()} for s in range(n_stages)]
            raise RuntimeError(f"stream-ordered p2p deadlock: {name!r}, {n_stages} stages, {n_micro} micro-batches, queue heads {heads}")
    return steps


# ------------------------------------------------------------------------------------------ mesh
class Mesh:
    """``n_pipelines`` x ``n_stages`` process mesh with one DP group per stage."""

    def __init__(self, n_stages: int, n_pipelines: int = 1, rank: int | None = None):
        self.n_stages, self.n_pipelines = n_stages, n_pipelines
        self.rank = dist.get_rank() if rank is None else rank
        self.pipeline_id, self.stage_id = divmod(self.rank, n_stages)
        self.dp_group = None
        if n_pipelines > 1:
            for s in range(n_stages):        # collective: EVERY rank creates EVERY group, in the same order
                ranks = [p * n_stages + s for p in range(n_pipelines)]
                grp = dist.new_group(ranks)
                if s == self.stage_id:
                    self.dp_group = grp
        # NCCL p2p operations between two ranks execute in issue order on one stream per communicator.  In 1F1B a stage
        # issues "send activation k+1" before "receive gradient k" while its successor issues "send gradient k" before
        # "receive activation k+1": on a single communicator the two sends wait for each other's receive forever.  Forward
        # traffic (activations) and backward traffic (gradients) therefore get a communicator each; inside one direction
        # both ends issue transfers in micro-batch order, so each communicator's streams always agree.  (gloo matches
        # messages by tag and needs none of this.)
        self.fwd_group = self.bwd_group = None
        if dist.is_initialized() and dist.get_backend() == "nccl" and n_stages > 1:
            world = dist.get_world_size()
            self.fwd_group = dist.new_group(list(range(world)))
            self.bwd_group = dist.new_group(list(range(world)))
            # NCCL creates the communicator of a p2p pair lazily, in a blocking rendezvous of the two ranks, on their FIRST
            # transfer.  Inside a schedule that first transfer can sit behind a dependency on the peer (a stage posts its first
            # gradient receive before it sends the activation its peer is waiting for), so every pair is connected here, in
            # stage order, before any schedule runs.
            dev = torch.device("cuda", torch.cuda.current_device())
            token = torch.zeros(1, device=dev)
            for s in range(n_stages - 1):
                lo, hi = self.rank_of(s), self.rank_of(s + 1)
                if self.rank == lo:
                    dist.send(token, dst=hi, group=self.fwd_group)
                    dist.recv(token, src=hi, group=self.bwd_group)
                elif self.rank == hi:
                    dist.recv(token, src=lo, group=self.fwd_group)
                    dist.send(token, dst=lo, group=self.bwd_group)
            torch.cuda.synchronize()

    def rank_of(self, stage: int) -> int:
        return self.pipeline_id * self.n_stages + stage

    @property
    def prev_rank(self):
        return None if self.stage_id == 0 else self.rank_of(self.stage_id - 1)

    @property
    def next_rank(self):
        return None if self.stage_id == self.n_stages - 1 else self.rank_of(self.stage_id + 1)


# ------------------------------------------------------------------------------------------ executor
class PipelineStage:
    """Runs one stage of a pipeline according to a schedule.

    ``module``: this stage's layers.  Stage 0 gets the micro-batch input from ``inputs``; the last
    stage computes ``loss_fn(output, target) / n_micro`` and starts the backward.  Activations /
    gradients of shape ``act_shape`` travel between neighbouring stages.
    """

    def __init__(self, module: nn.Module, mesh: Mesh, act_shape: Sequence[int], *, schedule: str = "1f1b",
                 n_micro: int = 1, loss_fn=None, act_dtype=torch.float32, device=None, flat_arena: bool = True,
                 grad_reducer=None, graph: bool = False):
        self.module, self.mesh, self.n_micro = module, mesh, n_micro
        # graph=True (CUDA only): a micro-batch's compute is two CUDA graphs per stage - "forward" (no autograd state kept)
        # and "forward + backward" (the activations are RECOMPUTED inside the backward graph), replayed per op with static
        # input / output buffers; the last stage needs only the second one.  The lab-sized stages are launch-bound (a few
        # hundred ~5 us kernels per op), so one cudaGraphLaunch per op is worth far more than the extra forward costs.
        # NCCL p2p stays outside the graphs.  Needs a deterministic forward without batch statistics (no Dropout, no BatchNorm).
        self.graph = bool(graph) and torch.cuda.is_available()
        self._g_fwd = self._g_bwd = None
        if self.graph:
            for mod in module.modules():
                if isinstance(mod, nn.modules.batchnorm._BatchNorm) or (isinstance(mod, nn.Dropout) and mod.p > 0):
                    raise ValueError("PipelineStage(graph=True) recomputes the forward: the stage must not contain BatchNorm / Dropout")
            assert flat_arena, "graph mode accumulates into the static flat gradient arena"
        # the stage's parameters and gradients live in one flat arena each, so the per-stage DP reduction of PP x DP is ONE
        # collective over one buffer (NCCL all-reduce, or `grad_reducer`: parallel.symm_impl.SymmetricGradReducer built on an
        # arena of the stage's DP group) instead of one call per parameter (intro_PP_1F1B_MP.py:232-235)
        from ..core.arena import FlatArena
        self.arena = FlatArena.from_module(module, with_grad=True) if flat_arena else None
        self.grad_reducer = grad_reducer
        if grad_reducer is not None and self.arena is not None and hasattr(grad_reducer, "grad"):
            self.arena.grad = grad_reducer.grad[:self.arena.size]
            self.arena.grad.zero_()
            self.arena.bind_grads()
        self.schedule = SCHEDULES[schedule](mesh.n_stages, mesh.stage_id, n_micro)
        self.loss_fn = loss_fn
        self.act_shape, self.act_dtype = tuple(act_shape), act_dtype
        self.device = torch.device(device) if device is not None else next(module.parameters()).device
        self.is_first, self.is_last = mesh.stage_id == 0, mesh.stage_id == mesh.n_stages - 1
        self.losses: list[float] = []

    def _buf(self):
        return torch.empty(self.act_shape, dtype=self.act_dtype, device=self.device)

    # ------------------------------------------------------------------ CUDA-graph mode
    def _build_graphs(self, x_example: torch.Tensor, tgt_example: torch.Tensor | None) -> None:
        from ..utils.graphs import GraphedStep
        mod = self.module
        self._rebind_grads()
        saved_grad = self.arena.grad.clone()

        def fwd(x):
            with torch.no_grad():
                return mod(x)

        def fwd_bwd(x, gout):
            xin = x.detach().requires_grad_(x.is_floating_point())
            mod(xin).backward(gout)
            return xin.grad if xin.grad is not None else gout.new_zeros(1)

        def fwd_loss_bwd(x, tgt):
            xin = x.detach().requires_grad_(x.is_floating_point())
            loss = self.loss_fn(mod(xin), tgt) / self.n_micro
            loss.backward()
            return (xin.grad if xin.grad is not None else loss.new_zeros(1)), loss.detach()

        if self.is_last:
            self._g_bwd = GraphedStep(fwd_loss_bwd, (x_example, tgt_example))
        else:
            self._g_fwd = GraphedStep(fwd, (x_example,))
            gout = torch.zeros(self.act_shape, dtype=self.act_dtype, device=self.device)
            self._g_bwd = GraphedStep(fwd_bwd, (x_example, gout))
        torch.cuda.synchronize()
        with torch.no_grad():
            self.arena.grad.copy_(saved_grad)        # warm-up / capture runs accumulated gradients of the example batch

    def _run_graphed(self, inputs, targets) -> list[float]:
        m = self.mesh
        saved_in: dict[int, torch.Tensor] = {}
        pending_send: list = []
        recv_f: dict[int, tuple] = {}
        recv_b: dict[int, tuple] = {}
        loss_vals: list[torch.Tensor] = []
        if self._g_bwd is None:
            # capture before any transfer of this iteration is posted (no communication is in flight during a capture)
            ex = inputs[0].to(self.device) if self.is_first else torch.zeros(self.act_shape, dtype=self.act_dtype, device=self.device)
            self._build_graphs(ex, targets[0].to(self.device) if self.is_last else None)

        def post_recv(op: Op):
            if op.kind == F and not self.is_first and op.mb not in recv_f:
                t = self._buf()
                recv_f[op.mb] = (t, dist.irecv(t, src=m.prev_rank, group=m.fwd_group))
            if op.kind == B and not self.is_last and op.mb not in recv_b:
                t = self._buf()
                recv_b[op.mb] = (t, dist.irecv(t, src=m.next_rank, group=m.bwd_group))

        ops = self.schedule
        if ops:
            post_recv(ops[0])
        for i, op in enumerate(ops):
            if i + 1 < len(ops):
                post_recv(ops[i + 1])
            if op.kind == F:
                if self.is_first:
                    x = inputs[op.mb].to(self.device)
                else:
                    t, req = recv_f.pop(op.mb)
                    req.wait()
                    x = t
                saved_in[op.mb] = x
                if not self.is_last:                   # the last stage does forward + loss + backward in its B op
                    payload = self._g_fwd(x).clone()
                    pending_send.append((dist.isend(payload, dst=m.next_rank, group=m.fwd_group), payload))
            else:
                x = saved_in.pop(op.mb)
                if self.is_last:
                    gin, loss = self._g_bwd(x, targets[op.mb].to(self.device))
                    loss_vals.append(loss.clone())
                else:
                    t, req = recv_b.pop(op.mb)
                    req.wait()
                    gin = self._g_bwd(x, t)
                if not self.is_first:
                    payload = gin.clone()
                    pending_send.append((dist.isend(payload, dst=m.prev_rank, group=m.bwd_group), payload))
        for r, _keepalive in pending_send:
            r.wait()
        self.losses = (torch.stack(loss_vals) * self.n_micro).tolist() if loss_vals else []
        return self.losses

    def run(self, inputs: Sequence[torch.Tensor] | None = None, targets: Sequence[torch.Tensor] | None = None) -> list[float]:
        """One training iteration (all micro-batches, forward and backward).  Gradients accumulate in
        ``module.parameters()``; the caller zeroes them and steps its optimizer."""
        if self.graph and self.device.type == "cuda":
            return self._run_graphed(inputs, targets)
        m = self.mesh
        saved_in: dict[int, torch.Tensor] = {}
        saved_out: dict[int, torch.Tensor] = {}
        pending_send: list = []
        recv_f: dict[int, tuple] = {}
        recv_b: dict[int, tuple] = {}
        self.losses = []
        loss_vals: list[torch.Tensor] = []        # kept on the device: ONE read-back per iteration, not one per micro-batch

        def post_recv(op: Op):
            if op.kind == F and not self.is_first and op.mb not in recv_f:
                t = self._buf()
                recv_f[op.mb] = (t, dist.irecv(t, src=m.prev_rank, group=m.fwd_group))
            if op.kind == B and not self.is_last and op.mb not in recv_b:
                t = self._buf()
                recv_b[op.mb] = (t, dist.irecv(t, src=m.next_rank, group=m.bwd_group))

        ops = self.schedule
        if ops:
            post_recv(ops[0])
        for i, op in enumerate(ops):
            if i + 1 < len(ops):
                post_recv(ops[i + 1])                 # prefetch the next message while this op computes
            if op.kind == F:
                if self.is_first:
                    x = inputs[op.mb].to(self.device)
                else:
                    t, req = recv_f.pop(op.mb)
                    req.wait()
                    x = t.requires_grad_()
                out = self.module(x)
                saved_in[op.mb] = x
                if self.is_last:
                    loss = self.loss_fn(out, targets[op.mb].to(self.device)) / self.n_micro
                    saved_out[op.mb] = loss
                    loss_vals.append(loss.detach())
                else:
                    saved_out[op.mb] = out
                    payload = out.detach().contiguous()
                    pending_send.append((dist.isend(payload, dst=m.next_rank, group=m.fwd_group), payload))
            else:
                out = saved_out.pop(op.mb)
                if self.is_last:
                    out.backward()
                else:
                    t, req = recv_b.pop(op.mb)
                    req.wait()
                    out.backward(t)
                x = saved_in.pop(op.mb)
                if not self.is_first:
                    payload = x.grad.contiguous()
                    pending_send.append((dist.isend(payload, dst=m.prev_rank, group=m.bwd_group), payload))
        for r, _keepalive in pending_send:
            r.wait()
        if loss_vals:
            self.losses = (torch.stack(loss_vals) * self.n_micro).tolist()
        return self.losses

    def zero_grad(self) -> None:
        if self.arena is not None:
            self.arena.grad.zero_()
            self._rebind_grads()
        else:
            for p in self.module.parameters():
                p.grad = None

    def _rebind_grads(self) -> None:
        """Make every ``p.grad`` the arena view again (an optimizer's ``zero_grad(set_to_none=True)`` drops the binding;
        gradients autograd then allocated elsewhere are copied in)."""
        a = self.arena
        params = dict(self.module.named_parameters())
        for sp in a.specs:
            if sp.is_buffer:
                continue
            p = params[sp.name]
            view = a.grad[sp.offset:sp.offset + sp.numel].view(sp.shape)
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad)
            p.grad = view

    def allreduce_grads(self) -> None:
        """Average gradients over the stage's DP group (PP x DP; intro_PP_1F1B_MP.py:232-235): one collective over the
        stage's flat gradient arena, in place."""
        if self.mesh.dp_group is None:
            return
        if self.arena is None:
            grads = [p.grad for p in self.module.parameters() if p.grad is not None]
            for g in grads:
                dist.all_reduce(g, group=self.mesh.dp_group)
                g.div_(self.mesh.n_pipelines)
            return
        self._rebind_grads()
        g = self.arena.grad
        if self.grad_reducer is not None:
            self.grad_reducer(g, 0, g.numel()).wait()        # fused peer-memory reduce over the DP group's arena
        else:
            dist.all_reduce(g, group=self.mesh.dp_group)
        g.div_(self.mesh.n_pipelines)
