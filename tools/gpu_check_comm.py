"""Cross-rank protocol of the fused collective kernels, checked on ONE GPU.

`b200_fused_update_emulated` / `b200_fused_collective_emulated` run W virtual ranks in a single
cooperative launch (blockIdx.y = rank, all "peer" pointers are local buffers): the device code,
the epoch flags, the split-phase publish counters and the metric mailboxes are the ones the real
multi-GPU launch uses; only the multicast (NVLS) instructions are out of reach. Reference: plain
PyTorch fp32 mean-reduce + AdamW (data_paral.py:210-217, param_sharding.py:134-138,353).

    python tools/gpu_check_comm.py            # all checks
    python tools/gpu_check_comm.py --dead-peer   # child process of the dead-peer check
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _ptr_array(ptrs):
    arr = (C.c_void_p * len(ptrs))(*ptrs)
    return arr, C.cast(arr, C.POINTER(C.c_void_p))


def _adamw_ref(p, m, v, g, t, lr, b1, b2, eps, wd):
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    mhat = m / (1 - b1 ** t)
    vhat = v / (1 - b2 ** t)
    p.sub_(lr * (mhat / (vhat.sqrt() + eps) + wd * p))


class _Ranks:
    """W virtual ranks' buffers for the fused update, DP (broadcast) or FSDP-like (local shards)."""

    def __init__(self, W: int, n: int, broadcast: bool, publish_fp32: bool, blocks: int, pieces: int = 1):
        import torch
        from jax_distributed_tuts_b200.ops import _lib
        self.W, self.n, self.broadcast, self.blocks = W, n, broadcast, blocks
        self.sl = n // W
        d = "cuda"
        g = torch.Generator(device="cpu").manual_seed(7)
        p0 = torch.randn(n, generator=g)
        self.grad = [torch.zeros(n, device=d) for _ in range(W)]
        self.param = [p0.clone().to(d) for _ in range(W)]
        self.bf16 = [p0.to(torch.bfloat16).to(d) for _ in range(W)]
        self.signal = [torch.zeros(_lib.SIGNAL_WORDS, dtype=torch.int32, device=d) for _ in range(W)]
        self.m = [torch.zeros(self.sl, device=d) for _ in range(W)]
        self.v = [torch.zeros(self.sl, device=d) for _ in range(W)]
        self.p_loc = [p0[r * self.sl:(r + 1) * self.sl].clone().to(d) for r in range(W)]
        self.b_loc = [torch.zeros(self.sl, dtype=torch.bfloat16, device=d) for _ in range(W)]
        self.step = [torch.zeros(1, dtype=torch.int32, device=d) for _ in range(W)]
        self.key = [torch.full((1,), 1234, dtype=torch.int64, device=d) for _ in range(W)]
        self.done = [torch.zeros(1, dtype=torch.int32, device=d) for _ in range(W)]
        self.step_metrics = [torch.zeros(8, device=d) for _ in range(W)]
        self.metrics_acc = [torch.zeros(8, device=d) for _ in range(W)]
        self.hp = _lib.AdamWHParams(1e-2, 0.9, 0.999, 1e-8, 1e-4, 1.0 / W)
        self.keep = []
        self.args = (_lib.FusedUpdateArgs * W)()
        self.scratch = torch.zeros(int(_lib.lib().b200_fused_update_emulated_scratch_bytes()), dtype=torch.uint8, device=d)
        segs = []
        for r in range(W):
            # the rank's slice as `pieces` segments (multiples of 64 elements); the last one takes the
            # ordered peer-load path like a replicated leaf would
            cuts = [0] + [min(self.sl, (self.sl // pieces // 64 + 1) * 64 * (i + 1)) for i in range(pieces - 1)] + [self.sl]
            raw = b""
            for i in range(pieces):
                lo, hi = cuts[i], cuts[i + 1]
                sg = _lib.Seg(src_off=r * self.sl + lo, src_pitch=hi - lo, dst_off=(r * self.sl + lo) if broadcast else 0,
                              loc_off=lo, rows=1, row_len=hi - lo,
                              flags=_lib.SEG_ORDERED if (pieces > 1 and i == pieces - 1) else 0)
                raw += bytes(sg)
            segs.append(torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(d))
        self.segs = segs
        for r in range(W):
            ctx = _lib.CommCtx()
            ctx.rank, ctx.world = r, W
            for name, bufs in (("peer_grad", self.grad), ("peer_param", self.param), ("peer_bf16", self.bf16),
                               ("peer_signal", self.signal)):
                arr, p = _ptr_array([b.data_ptr() for b in bufs])
                self.keep.append(arr)
                setattr(ctx, name, p)
            self.keep.append(ctx)
            end = _lib.StepEnd(self.key[r].data_ptr(), 99, self.step[r].data_ptr(), self.done[r].data_ptr())
            self.keep.append(end)
            u = self.args[r]
            u.ctx = C.pointer(ctx)
            u.segs_dev, u.nseg, u.broadcast = segs[r].data_ptr(), pieces, int(broadcast)
            u.publish_fp32, u.nmetrics, u.blocks = int(publish_fp32), 2, blocks
            u.p_local = None if broadcast else self.p_loc[r].data_ptr()
            u.bf16_local = None if broadcast else self.b_loc[r].data_ptr()
            u.m, u.v = self.m[r].data_ptr(), self.v[r].data_ptr()
            u.hp = C.pointer(self.hp)
            u.step_count = self.step[r].data_ptr()
            u.step_metrics = self.step_metrics[r].data_ptr()
            u.metrics_acc = self.metrics_acc[r].data_ptr()
            u.end = C.pointer(end)
        self.ref_p, self.ref_m, self.ref_v = p0.clone().double(), torch.zeros(n).double(), torch.zeros(n).double()
        self.ref_metrics = torch.zeros(2, dtype=torch.float64)

    def launch(self):
        import torch
        from jax_distributed_tuts_b200.ops import _lib
        _lib.check(_lib.lib().b200_fused_update_emulated(self.args, self.W, self.scratch.data_ptr(),
                                                          torch.cuda.current_stream().cuda_stream))

    def wait_published(self):
        import torch
        from jax_distributed_tuts_b200.ops import _lib
        for r in range(self.W):
            _lib.check(_lib.lib().b200_comm_wait_published(self.signal[r].data_ptr(), r, self.W,
                                                            torch.cuda.current_stream().cuda_stream))


def check_fused_update(W: int, n: int, broadcast: bool, publish_fp32: bool, blocks: int, steps: int = 4,
                       pieces: int = 1) -> bool:
    import torch
    from jax_distributed_tuts_b200.ops import _lib
    R = _Ranks(W, n, broadcast, publish_fp32, blocks, pieces)
    gen = torch.Generator().manual_seed(11)
    for t in range(1, steps + 1):
        gs = [torch.randn(n, generator=gen) for _ in range(W)]
        ms = [torch.rand(2, generator=gen) for _ in range(W)]
        for r in range(W):
            R.grad[r].copy_(gs[r])
            R.step_metrics[r][:2].copy_(ms[r])
        R.launch()
        R.wait_published()
        gmean = torch.stack(gs).double().sum(0) / W
        _adamw_ref(R.ref_p, R.ref_m, R.ref_v, gmean, t, 1e-2, 0.9, 0.999, 1e-8, 1e-4)
        R.ref_metrics += torch.stack(ms).double().sum(0)
    torch.cuda.synchronize()
    ok = True
    sl = R.sl
    ref32 = R.ref_p.float()
    tol = 2e-5
    for r in range(W):
        own = slice(r * sl, (r + 1) * sl)
        if broadcast:
            e_own = (R.param[r][own].cpu() - ref32[own]).abs().max().item()
            ok &= e_own < tol
            if publish_fp32:
                ok &= (R.param[r].cpu() - ref32).abs().max().item() < tol
                ok &= all(torch.equal(R.param[r], R.param[0]) for _ in (0,))
            got_b = R.bf16[r].float().cpu()
            ok &= torch.equal(R.bf16[r], R.bf16[0])                       # replicas bit-identical
            ok &= (got_b - ref32).abs().max().item() <= 1.0 / 128 * ref32.abs().max().item() + 1e-6
            ok &= torch.equal(R.bf16[r][own], R.param[r][own].to(torch.bfloat16))
        else:
            ok &= (R.p_loc[r].cpu() - ref32[own]).abs().max().item() < tol
            ok &= torch.equal(R.b_loc[r], R.p_loc[r].to(torch.bfloat16))
        ok &= (R.m[r].cpu().double() - R.ref_m[own]).abs().max().item() < 1e-5
        ok &= (R.v[r].cpu().double() - R.ref_v[own]).abs().max().item() < 1e-5
        ok &= int(R.step[r].item()) == steps and int(R.done[r].item()) == 0 and int(R.key[r].item()) != 1234
        ok &= (R.metrics_acc[r][:2].cpu().double() - R.ref_metrics).abs().max().item() < 1e-4
    pub = [int(R.signal[r][_lib.SIG_PUB + p].item()) for r in range(W) for p in range(W)]     # SIG_PUB block
    ok &= len(set(pub)) == 1 and pub[0] > 0
    print(f"  [{'ok' if ok else 'FAIL'}] fused update emulated: W={W} n={n} {'dp' if broadcast else 'fsdp'} "
          f"publish_fp32={int(publish_fp32)} blocks={blocks} segments={pieces} steps={steps} publishes/peer={pub[0]}", flush=True)
    return bool(ok)


def check_collectives(W: int, n: int, blocks: int) -> bool:
    import torch
    from jax_distributed_tuts_b200.ops import _lib
    d = "cuda"
    gen = torch.Generator().manual_seed(5)
    xs = [torch.randn(n, generator=gen) for _ in range(W)]
    signal = [torch.zeros(_lib.SIGNAL_WORDS, dtype=torch.int32, device=d) for _ in range(W)]
    scratch = torch.zeros(int(_lib.lib().b200_fused_update_emulated_scratch_bytes()), dtype=torch.uint8, device=d)
    total = torch.stack(xs).double().sum(0)
    ok = True
    for mode, name in ((0, "all-reduce(mean)"), (1, "reduce-scatter"), (2, "all-gather")):
        for rep in range(2):      # twice: the epoch counters carry over
            bufs = [x.clone().to(d) for x in xs]
            out_n = {0: 1, 1: n // W, 2: n * W}[mode]
            outs = [torch.zeros(out_n, device=d) for _ in range(W)]
            keep = []
            args = (_lib.CollectiveArgs * W)()
            for r in range(W):
                a = args[r]
                a.rank, a.world, a.mode, a.blocks = r, W, mode, blocks
                arr, p = _ptr_array([b.data_ptr() for b in bufs]); keep.append(arr); a.peer_buf = p
                arr, p = _ptr_array([s.data_ptr() for s in signal]); keep.append(arr); a.peer_signal = p
                a.mc_buf = None
                a.out = outs[r].data_ptr() if mode else None
                a.n = n if mode != 2 else n // 4        # mode 2 counts 16-byte units
                a.scale = 1.0 / W if mode == 0 else 1.0
            _lib.check(_lib.lib().b200_fused_collective_emulated(args, W, scratch.data_ptr(),
                                                                  torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
            good = True
            for r in range(W):
                if mode == 0:
                    good &= (bufs[r].cpu().double() - total / W).abs().max().item() < 1e-5
                    good &= torch.equal(bufs[r], bufs[0])
                elif mode == 1:
                    sl = n // W
                    good &= (outs[r].cpu().double() - total[r * sl:(r + 1) * sl]).abs().max().item() < 1e-5
                else:
                    good &= torch.equal(outs[r].cpu(), torch.cat(xs))
            print(f"  [{'ok' if good else 'FAIL'}] collective emulated: {name} W={W} n={n} blocks={blocks} call {rep + 1}", flush=True)
            ok &= good
    return bool(ok)


def check_dead_peer() -> bool:
    """A rank whose peer never shows up must fail its launch after the timeout, not hang."""
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--dead-peer"], capture_output=True, text=True, timeout=120)
    out = r.stdout + r.stderr
    ok = r.returncode != 0 and "SURVIVED" not in out      # (the device-side message is not always flushed after a trap)
    print(f"  [{'ok' if ok else 'FAIL'}] dead peer: launch aborted with rc={r.returncode} "
          f"({'message seen' if 'gave up waiting' in out else 'no message'})", flush=True)
    if not ok:
        print(out[-1500:])
    return ok


def _dead_peer_child() -> None:
    import torch
    from jax_distributed_tuts_b200.ops import _lib
    lib = _lib.lib()
    lib.b200_comm_set_timeout_ms(300)
    R = _Ranks(2, 2048, True, False, 4)
    # launch rank 0 alone through the real (non-emulated) entry point: rank 1 never arrives
    _lib.check(lib.b200_fused_update(C.byref(R.args[0]), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()       # raises: the kernel trapped
    print("SURVIVED")


def group_comm_emulated(dead_peer: bool = True) -> bool:
    from jax_distributed_tuts_b200.ops import _lib
    _lib.lib().b200_comm_set_timeout_ms(5000)
    ok = True
    try:
        for (W, n, bc, p32, blocks) in [(2, 2 * 4096, True, False, 8), (4, 4 * 4096 * 3, True, True, 16),
                                        (8, 8 * 64 * 5, True, False, 32), (8, 8 * 64 * 1000, True, False, 32),
                                        (4, 4 * 4096, False, False, 16), (8, 8 * 64 * 33, False, False, 8)]:
            ok &= check_fused_update(W, n, bc, p32, blocks)
        # several segments per rank: spread over block ranges (small) or one after the other (large)
        ok &= check_fused_update(4, 4 * 4096, False, False, 16, pieces=3)
        ok &= check_fused_update(8, 8 * 64 * 1000, True, False, 8, pieces=4)
        ok &= check_collectives(4, 4 * 4 * 1000, 8)
        ok &= check_collectives(8, 8 * 4 * 64, 32)
    finally:
        _lib.lib().b200_comm_set_timeout_ms(0)
    if dead_peer:
        ok &= check_dead_peer()
    return bool(ok)


if __name__ == "__main__":
    if "--dead-peer" in sys.argv:
        _dead_peer_child()
        sys.exit(0)
    good = group_comm_emulated(dead_peer="--no-dead-peer" not in sys.argv)
    print("COMM EMULATED CHECK", "PASS" if good else "FAIL")
    sys.exit(0 if good else 1)
