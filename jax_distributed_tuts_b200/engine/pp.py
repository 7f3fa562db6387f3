"""Pipeline engine: GPipe over a ``pipe`` mesh axis with every stage on the fused tier.

The reference's ``pipeline_parallel.py`` is a stub (imports and three aliases, pipeline_parallel.py:1-39:
it meant to reuse ``Batch`` / ``TrainState`` / gradient accumulation / ``fold_rng_over_axis`` /
``sync_gradients``); BASELINE.json fixes the algorithm: layers split into stages, the batch into
microbatches, activations shifted stage to stage for M + S - 1 ticks, loss on the last stage,
backward in reverse, then gradient sync + optimizer per stage, with the point-to-point traffic on
NCCL, overlapped with stage compute. ``parallel/pp.py`` is that algorithm on the generic tier
(autograd, eager launches); this module is the product path:

* every stage is a :class:`GPT2Engine` restricted to its blocks (flat buffers, our sm_100a kernels,
  one set of saved activations per microbatch), the forward and the backward of each microbatch are
  captured once as CUDA graphs and replayed;
* activations and activation gradients travel as ``ncclSend`` / ``ncclRecv`` on a communication stream
  of their own. Each adjacent stage pair has its own communicator, so a stage's receive from below
  never queues behind its send upwards; all receives of a phase are posted before its first
  microbatch is computed (into per-microbatch buffers), sends are issued the moment a microbatch's
  graph has been queued and are only waited for when their buffer is about to be reused;
* the last stage runs the loss and the backward of the final norm + tied head right behind each
  microbatch's forward (the logits buffer is never kept M times), the rest of its backward in the
  backward phase;
* the tied ``wte`` lives on the first and on the last stage; they exchange its gradient once per
  step and apply the same update (the generic tier's convention);
* layers are dealt to stages by cost, not by count: the tied head costs as much as
  ``vocab / (12 d_model)`` blocks (5.5 for GPT-2 small), so the last stage gets fewer blocks;
* a ``data`` axis, if the mesh has one, averages each stage's gradients with the fused
  reduce + AdamW kernel over that axis' symmetric-memory group (``sync_gradients`` semantics,
  param_sharding.py:293-322).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import os

import torch
import torch.distributed as dist

from .. import rng as _rng
from ..mesh import Mesh
from ..models.gpt2 import padded_vocab
from ..ops import _lib
from ..ops import kernels as K
from ..state import Batch
from .gpt2 import GPT2Engine


def balanced_split(n_layer: int, n_stages: int, head_cost: float, embed_cost: float = 0.0) -> List[range]:
    """Contiguous layer ranges that minimise the most loaded stage, where the last stage also carries
    the head (``head_cost`` in units of one block) and the first the embeddings. The last stage may
    end up with no block at all."""
    assert n_layer >= n_stages - 1, f"{n_stages} stages need at least {n_stages - 1} blocks, the model has {n_layer}"
    best, best_cut = None, None      # best = (largest load, sum of squared loads): even stages among equal maxima

    def rec(start: int, stage: int, cuts: List[int], worst: float, sq: float):
        nonlocal best, best_cut
        if best is not None and worst > best[0]:
            return
        if stage == n_stages - 1:
            load = (n_layer - start) + head_cost + (embed_cost if stage == 0 else 0.0)
            cand = (max(worst, load), sq + load * load)
            if best is None or cand < best:
                best, best_cut = cand, cuts + [n_layer]
            return
        # every stage but the last keeps at least one block (only the last owns something else: the head)
        for n in range(1, n_layer - start + 1):
            load = n + (embed_cost if stage == 0 else 0.0)
            rec(start + n, stage + 1, cuts + [start + n], max(worst, load), sq + load * load)

    rec(0, 0, [], 0.0, 0.0)
    out, start = [], 0
    for c in best_cut:
        out.append(range(start, c))
        start = c
    return out


def stage_sequence(schedule: str, n_stages: int, stage: int, n_micro: int) -> List[Tuple[str, int]]:
    """Order of forwards ("F", k) and backwards ("B", k) on one stage. "gpipe": every forward, then the
    backwards from the last microbatch down. "1f1b": S-1-s warm-up forwards, then one forward and one
    backward in turn (backwards in arrival order 0, 1, ...), then the remaining backwards. Every stage
    sends its forwards up and its backwards down in exactly this order, which is also the order in
    which its neighbours post the matching receives (checked by simulation in tests/test_cpu_pipeline.py)."""
    if schedule == "gpipe":
        return [("F", k) for k in range(n_micro)] + [("B", k) for k in range(n_micro - 1, -1, -1)]
    warm = min(n_stages - 1 - stage, n_micro)
    seq = [("F", k) for k in range(warm)]
    for i in range(n_micro - warm):
        seq += [("F", warm + i), ("B", i)]
    return seq + [("B", k) for k in range(n_micro - warm, n_micro)]


class PipelineEngine:
    """``train_step(batch) -> (loss_sum, token_count)`` for one (data, pipe) mesh coordinate. The mesh must
    cover the whole job; the engine's kernels run on the stream that is current when it is constructed."""

    def __init__(self, config, mesh: Mesh, *, batch_per_replica: int, seq_len: Optional[int] = None,
                 num_microbatches: int = 4, lr: float = 3e-4, init_params: Optional[Dict] = None,
                 pipe_axis: str = "pipe", data_axis: Optional[str] = "data", use_graph: bool = True,
                 layer_split: Optional[Sequence[range]] = None, schedule: str = "1f1b"):
        assert torch.cuda.is_available(), "the pipeline engine needs CUDA devices; use parallel/pp.py on CPU"
        self.mesh, self.cfg = mesh, config
        self.pipe_axis = pipe_axis
        self.data_axis = data_axis if (data_axis in mesh.axis_names) else None
        self.S, self.s = mesh.axis_size(pipe_axis), mesh.axis_index(pipe_axis)
        self.D = mesh.axis_size(self.data_axis) if self.data_axis else 1
        self.first, self.last = self.s == 0, self.s == self.S - 1
        self.Mb = num_microbatches
        self.B, self.T = batch_per_replica, seq_len or config.n_ctx
        assert self.B % self.Mb == 0
        self.mbB = self.B // self.Mb
        self.use_graph = use_graph
        # "gpipe": all forwards, then all backwards in reverse order. "1f1b": the same microbatches and the
        # same gradients, but a stage starts the backward of microbatch k as soon as its gradient arrives
        # (one forward, one backward in turn after a warm-up of S-1-s forwards): the stages below the head
        # work on backwards while the head stage is still busy with later microbatches.
        assert schedule in ("gpipe", "1f1b")
        self.schedule = os.environ.get("B200DIST_PP_SCHEDULE", schedule)
        assert self.schedule in ("gpipe", "1f1b")
        schedule = self.schedule
        head_cost = padded_vocab(config.vocab_size) / (12.0 * config.d_model)
        self.split = list(layer_split) if layer_split is not None else balanced_split(config.n_layer, self.S, head_cost)
        assert len(self.split) == self.S
        self.stage = GPT2Engine(config, mesh, mode="dp", batch_per_rank=self.B, seq_len=self.T, lr=lr,
                                num_minibatches=self.Mb, init_params=init_params, layer_ids=self.split[self.s],
                                has_embed=self.first, has_head=self.last, slots=self.Mb, use_graph=False, overlap=False,
                                comm_axes=(self.data_axis,) if self.data_axis else ())
        # programmatic dependent launch inside each microbatch graph: a stage's kernels at microbatch size are
        # short (tens of microseconds), the same regime as the MLP step (FusedEngine.pdl)
        self.pdl = os.environ.get("B200DIST_PP_PDL", "1") == "1"
        # per-stage RNG stream (the fold_rng_over_axis the reference's stub imports, pipeline_parallel.py:37)
        self.stage.rng_base = _rng.rebase(_rng.fold_in(_rng.PRNGKey(0), self.s)).base.to(self.stage.device)
        self.device = self.stage.device
        self.compute = torch.cuda.current_stream()
        self.comm_stream = torch.cuda.Stream()       # sends are issued here, behind the compute that produced their data
        self.recv_stream = torch.cuda.Stream()       # receives are issued here, behind nothing but the previous step
        self._setup_groups()
        self._bwd_order = list(range(self.Mb - 1, -1, -1)) if schedule == "gpipe" else list(range(self.Mb))
        self._first_bwd = self._bwd_order[0]
        self.graphs: Dict[Tuple[str, int], torch.cuda.CUDAGraph] = {}
        self._calls = 0
        self._send_works: List = []
        self.loss_vec = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.launches_per_step = 0
        if self.S > 1 and (self.first or self.last):
            sp = self.stage.layout.leaves["wte"]
            self._wte_grad = self.stage.g32[sp.offset: sp.offset + sp.rows * sp.ld]
            self._wte_other = torch.empty_like(self._wte_grad)

    # ------------------------------------------------------------------ communicators -------------
    def _setup_groups(self) -> None:
        """Two 2-rank communicators (one per direction) per adjacent stage pair of every pipeline and one for
        the (first, last) pair; every rank creates all of them, in the same order. The metrics use the
        communicator of the whole job."""
        m = self.mesh
        ai_p = m.axis_names.index(self.pipe_axis)
        others = [i for i in range(len(m.axis_names)) if i != ai_p]
        self.prev_rank = self.next_rank = None
        self.up_prev = self.up_next = self.down_prev = self.down_next = self.pair_ends = None
        self.ends_peer = None
        self._conn_plan = []          # (group, sender, receiver) in creation order, identical on every rank
        import itertools
        for fixed in itertools.product(*(range(m.shape[i]) for i in others)):
            def rank_of(stage):
                coords = [0] * len(m.axis_names)
                for i, v in zip(others, fixed):
                    coords[i] = v
                coords[ai_p] = stage
                return m._rank_of(coords)
            mine = all(m.coords[i] == v for i, v in zip(others, fixed))
            for st in range(self.S - 1):
                a, b = rank_of(st), rank_of(st + 1)
                # one communicator per direction: activations go up on one, gradients come down on the
                # other, so neither ever queues behind the other
                up = dist.new_group([a, b]) if m.size > 1 else None
                down = dist.new_group([a, b]) if m.size > 1 else None
                if m.size > 1:
                    self._conn_plan += [(up, a, b), (down, b, a)]
                if mine and st == self.s:
                    self.up_next, self.down_next, self.next_rank = up, down, b
                if mine and st + 1 == self.s:
                    self.up_prev, self.down_prev, self.prev_rank = up, down, a
            if self.S > 2:
                a, b = rank_of(0), rank_of(self.S - 1)
                g = dist.new_group([a, b])
                self._conn_plan += [(g, a, b), (g, b, a)]
                if mine and self.first:
                    self.pair_ends, self.ends_peer = g, b
                if mine and self.last:
                    self.pair_ends, self.ends_peer = g, a
            elif self.S == 2:
                a, b = rank_of(0), rank_of(1)
                g = dist.new_group([a, b])
                self._conn_plan += [(g, a, b), (g, b, a)]
                if mine:
                    self.pair_ends, self.ends_peer = g, (b if self.first else a)
        self.all_group = dist.group.WORLD if m.size > 1 else None
        self._warm_connections()

    def _warm_connections(self) -> None:
        """NCCL sets up the connection of a (communicator, peer) pair inside the first send / receive on it, and
        that handshake blocks the host until the peer issues its matching call. The step posts receives long
        before the matching sends, on several communicators at once, which would turn those first handshakes
        into a cycle. So every connection is opened here, one communicator after the other in the same order
        on every rank (a total order: no cycle), with a one-element message in the direction it will carry."""
        if self.mesh.size == 1:
            return
        tiny = torch.zeros(1, device=self.stage.device)
        me = self.mesh.rank
        for (g, src, dst) in self._conn_plan:
            if me == src:
                dist.send(tiny, dst, group=g)
            elif me == dst:
                dist.recv(tiny, src, group=g)
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ per-microbatch bodies -----
    def _tok(self, k: int) -> torch.Tensor:
        return self.stage.tokens[k * self.mbB:(k + 1) * self.mbB]

    def _tgt(self, k: int) -> torch.Tensor:
        return self.stage.targets[k * self.mbB:(k + 1) * self.mbB]

    def _fwd_body(self, k: int) -> None:
        st = self.stage
        x = st._forward(self._tok(k) if self.first else None, slot=k)
        self._fwd_out[k] = x
        if self.last:      # loss + backward of the final norm and the tied head right away: the logits buffer is reused
            st._head_backward(self._tgt(k), st.slot_dx_top[k], accumulate=k > 0)

    def _bwd_body(self, k: int) -> None:
        st = self.stage
        acc = k != self._first_bwd                 # the first backward of the step overwrites, the others accumulate
        dx_in = st.slot_dx_top[k]                  # from the head (last stage) or received from the stage above
        if not self.last and st.L:
            # the column sums of the incoming gradient are the bias gradient of this stage's top projection
            K.colsum(dx_in, st.grad1d(f"h_{st.layer_ids[-1]}/mlp/c_proj/bias"), accumulate=True)
        dx = st._layers_backward(dx_in, accumulate=acc, last=False, slot=k,
                                 final_out=None if self.first else st.slot_dx_out[k])
        if self.first:
            st._embed_backward(self._tok(k), dx, accumulate=acc, zero_wte=not self.last)
        else:
            self._bwd_out[k] = dx

    def _end_body(self) -> None:
        st = self.stage
        next_rng = _rng.split(_rng.Key(st.rng_base, 0), 2)[0]
        st._fold_end(next_rng)
        st._sync_all_buckets()
        st._end_step(next_rng)

    def _run(self, name: str, k: int, body) -> None:
        """Eager on the first step (lazy initialisation inside CUDA / NCCL happens there), a graph replay
        afterwards."""
        g = self.graphs.get((name, k))
        if g is not None:
            g.replay()
            return
        before = _lib.launch_count()
        _lib.set_pdl(1 if self.pdl else 0)
        try:
            body()
        finally:
            _lib.set_pdl(-1)
        if self._calls == 0:
            self.launches_per_step += _lib.launch_count() - before

    def _capture_all(self) -> None:
        """Captures every piece of the step (begin, forward and backward of each microbatch, update) as
        its own CUDA graph. Capturing launches nothing, and no message is in flight here, so the ranks
        need not be in step with each other."""
        torch.cuda.synchronize()
        cap = torch.cuda.Stream()
        bodies = [("begin", 0, lambda: self.stage._begin_step(None, None))]
        bodies += [("fwd", k, (lambda k=k: self._fwd_body(k))) for k in range(self.Mb)]
        if self.stage.L or self.first:
            bodies += [("bwd", k, (lambda k=k: self._bwd_body(k))) for k in range(self.Mb)]
        bodies += [("end", 0, self._end_body)]
        for name, k, body in bodies:
            g = torch.cuda.CUDAGraph()
            _lib.set_pdl(1 if self.pdl else 0)
            try:
                with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                    body()
            finally:
                _lib.set_pdl(-1)
            self.graphs[(name, k)] = g
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ the step -------------------
    def load_batch(self, batch: Batch) -> None:
        """Host -> device copy of the replica's batch (every stage keeps tokens and targets: the first
        needs the tokens, the last the targets, and shapes are shared)."""
        self.stage.tokens.copy_(batch.inputs, non_blocking=True)
        self.stage.targets.copy_(batch.labels, non_blocking=True)

    def _isend(self, t: torch.Tensor, dst: int, group) -> None:
        """Queues a send of ``t`` behind everything already on the compute stream."""
        self.comm_stream.wait_stream(self.compute)
        with torch.cuda.stream(self.comm_stream):
            self._send_works.append(dist.isend(t, dst, group=group))

    def _irecv(self, buf: torch.Tensor, src: int, group):
        with torch.cuda.stream(self.recv_stream):
            return dist.irecv(buf, src, group=group)

    def _sequence(self) -> List[Tuple[str, int]]:
        return stage_sequence(self.schedule, self.S, self.s, self.Mb)

    def run_step(self) -> None:
        st = self.stage
        if self.use_graph and self._calls == 1 and not self.graphs:
            self._capture_all()
        self._fwd_out: Dict[int, torch.Tensor] = getattr(self, "_fwd_out", {})
        self._bwd_out: Dict[int, torch.Tensor] = getattr(self, "_bwd_out", {})
        # receive buffers were last read by the previous step's kernels, all queued on the compute stream by now
        self.recv_stream.wait_stream(self.compute)
        self._run("begin", 0, lambda: st._begin_step(None, None))
        has_bwd = bool(st.L or self.first)
        posted: Dict[Tuple[str, int], object] = {}

        def post(op):
            """The receive that feeds ``op``. A receive is posted only once everything this rank contributes
            to the message's production has been queued (the operation before ``op`` and its send): a
            waiting receive can then never sit in front of work it depends on, however the streams are
            mapped onto hardware queues."""
            kind, k = op
            if kind == "F" and not self.first:
                posted[op] = self._irecv(st.slot_x0[k], self.prev_rank, self.up_prev)
            elif kind == "B" and not self.last:
                posted[op] = self._irecv(st.slot_dx_top[k], self.next_rank, self.down_next)

        seq = self._sequence()
        post(seq[0])
        for n, op in enumerate(seq):
            kind, k = op
            if op in posted:
                posted[op].wait()                         # the compute stream waits; the host does not
            if kind == "F":
                self._run("fwd", k, lambda: self._fwd_body(k))
                if not self.last:
                    self._isend(self._fwd_out[k], self.next_rank, self.up_next)
            else:
                if has_bwd:
                    self._run("bwd", k, lambda: self._bwd_body(k))
                if not self.first:
                    self._isend(self._bwd_out.get(k, st.slot_dx_top[k]), self.prev_rank, self.down_prev)
            if n + 1 < len(seq):
                post(seq[n + 1])                          # host runs ahead of the GPU: this is long before op n finishes
        # ---- tied embedding: first and last stage exchange their wte gradients and both keep the sum ----
        if self.S > 1 and (self.first or self.last):
            self.comm_stream.wait_stream(self.compute)
            with torch.cuda.stream(self.comm_stream):
                ops = [dist.P2POp(dist.isend, self._wte_grad, self.ends_peer, group=self.pair_ends),
                       dist.P2POp(dist.irecv, self._wte_other, self.ends_peer, group=self.pair_ends)]
                works = dist.batch_isend_irecv(ops)
            for w in works:
                w.wait()
            self._wte_grad.add_(self._wte_other)
        # ---- per-stage gradient sync over the data axis + AdamW (one fused launch) ----
        self._run("end", 0, self._end_body)
        # every send of this step must have left its buffer before the next step overwrites it
        for w in self._send_works:
            w.wait()
        self._send_works.clear()
        self._calls += 1

    def train_step(self, batch: Batch) -> Tuple[torch.Tensor, int]:
        """H2D of the batch, one GPipe step, and the step's loss sum (summed over the pipe and data axes,
        so every rank returns the same value) with the number of tokens it covers."""
        self.load_batch(batch)
        self.run_step()
        self.loss_vec.copy_(self.stage.step_metrics[:2])
        if self.mesh.size > 1:
            dist.all_reduce(self.loss_vec, group=self.all_group)
        return self.loss_vec[0], self.B * self.T * self.D

    # ------------------------------------------------------------------ inspection -----------------
    def params_tree(self) -> Dict:
        """This stage's parameters (global layout of its leaves)."""
        return self.stage.params_tree()

    def state_dict(self) -> Dict[str, object]:
        """The whole model's state in the checkpoint layout (``checkpoint.py``): every stage contributes its
        blocks, the dicts are merged over the pipe axis (tied ``wte`` once), Adam moments in the global
        per-leaf layout. Collective over the mesh; every rank gets the same dict."""
        from .. import checkpoint as ckpt
        st = self.stage
        og = st.opt_state_global()
        sd = {"step": ckpt._to_numpy(st.step_count).reshape(()).astype("int32"), "params": ckpt._plain(st.params_tree()),
              "opt_state": {"0": {"count": ckpt._to_numpy(og["count"]).reshape(()), "mu": ckpt._plain(og["mu"]),
                                  "nu": ckpt._plain(og["nu"])}, "1": {}, "2": {}},
              "rng": ckpt._to_numpy(st.rng_base)}
        return ckpt.merge_over_pipe(sd, self.mesh, self.pipe_axis)

    def load_state_dict(self, loaded: Dict[str, object]) -> None:
        """Inverse of ``state_dict`` (also accepts a checkpoint written by any other strategy): each stage
        keeps the leaves it owns."""
        import numpy as np
        st = self.stage
        st.load_params(loaded["params"])
        adam = loaded["opt_state"]["0"]
        st.load_opt_state_global(adam["mu"], adam["nu"], int(np.asarray(adam["count"]).reshape(-1)[0]))

    def report(self) -> Dict[str, object]:
        """Layer split, ideal GPipe bubble, and how long this stage's kernels are busy per step (its
        graphs replayed back to back without any communication), max / min over the stages: step time
        minus busy time is what the stage spends waiting for its neighbours (bubble + exposed p2p)."""
        busy = torch.zeros(1, dtype=torch.float64, device=self.device)
        if self.graphs:
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for (_name, _k), g in self.graphs.items():
                g.replay()
            b.record()
            torch.cuda.synchronize()
            busy[0] = a.elapsed_time(b)
        hi, lo = busy.clone(), busy.clone()
        if self.mesh.size > 1:
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.all_group)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.all_group)
        return {"stages": self.S, "microbatches": self.Mb, "schedule": self.schedule, "layers_per_stage": [len(r) for r in self.split],
                "ideal_bubble_frac": (self.S - 1) / (self.Mb + self.S - 1),
                "stage_busy_ms_max": float(hi.item()), "stage_busy_ms_min": float(lo.item()),
                "note": "busy = this stage's graphs replayed without communication (the measurement perturbs the training "
                        "state: it runs after every reported number was taken); step - busy = bubble + exposed p2p"}
