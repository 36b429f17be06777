"""Data-parallel plumbing of the decision path: one process per GPU, torch.distributed for the exchange.

Inference shards battles with no collective (battles are independent, mlp_model.py:12: kernel size 1).  The
update needs exactly one exchange: the sum of the flat fp32 gradient vectors.  Each rank
computes d/dparams of (sum of its shard's Huber terms / GLOBAL batch size), so the sum over ranks is the
gradient of the reference's mean loss (trainer.py:187); the [-1, 1] clamp (trainer.py:193-194) is applied to
the reduced gradient, never to the per-rank pieces (clamp does not commute with the sum).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of n items for `rank` of `world`; shards are disjoint, ordered and cover
    range(n); sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in 0..{world - 1}")
    return (n * rank) // world, (n * (rank + 1)) // world


def active_group():
    """torch.distributed if a process group with more than one rank is initialised, else None."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


def allreduce_sum_(grad: torch.Tensor, loss: Optional[torch.Tensor] = None, dist=None) -> None:
    """In-place sum over ranks of the flat gradient (and the loss share).  NCCL on GPUs (one allreduce on one
    contiguous buffer: 175 KB for the default network, latency-bound), gloo in the CPU tests."""
    dist = dist if dist is not None else active_group()
    if dist is None:
        return
    dist.all_reduce(grad, op=dist.ReduceOp.SUM)
    if loss is not None:
        dist.all_reduce(loss, op=dist.ReduceOp.SUM)


class PeerGradExchange:
    """Gradient sum + clamp + Adam in one kernel over NVLink peer memory (pk_peer_clamp_adam, pk_update.cuh).

    Every rank's flat gradient lives in a buffer all peers have mapped (torch symmetric memory does the one-time
    handle exchange: plumbing only); each step the ranks write their shard gradient into the buffer of the step's
    parity and launch ONE kernel that waits on per-rank flags, sums the W copies in rank order, clamps and applies
    Adam.  No NCCL call per step; the reduced gradient and the new weights are bit-identical on every rank.

        px = PeerGradExchange.create(n_params, device)      # None when unavailable (every rank agrees)
        grad, loss = px.next_slot()                          # views: write the shard gradient / loss share here
        px.reduce_clamp_adam(flat, exp_avg, exp_avg_sq, n_layers, H, C, step, lr, betas, eps, prepared)
        px.loss_status -> tensor [summed loss, timeout flag]
    """

    FLAG_FLOATS = 64  # 256 B of flag words in front of the two gradient sets

    def __init__(self, n_params: int, device: torch.device, group=None):
        import ctypes as C
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self._C = C
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > 8:
            raise RuntimeError("PeerGradExchange supports up to 8 ranks (one NVSwitch box)")
        self.n = int(n_params)
        self.npad = (self.n + 1 + 63) // 64 * 64
        total = self.FLAG_FLOATS + 2 * self.npad
        self.buf = symm.empty(total, dtype=torch.float32, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.hdl.barrier()  # every rank's flags are zero before anyone publishes
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        assert len(ptrs) == self.world and ptrs[self.rank] == self.buf.data_ptr()
        self._flag_ptrs = (C.c_void_p * self.world)(*ptrs)
        base = [p + 4 * self.FLAG_FLOATS for p in ptrs]
        self._grad_ptrs = [(C.c_void_p * self.world)(*[b + 4 * self.npad * par for b in base]) for par in (0, 1)]
        self.epoch = 0
        self.loss_status = torch.zeros(2, dtype=torch.float32, device=device)
        self.device = device

    @classmethod
    def create(cls, n_params: int, device, group=None) -> Optional["PeerGradExchange"]:
        """Collective: every rank calls it; returns an exchange on all ranks or None on all ranks."""
        import os
        import torch.distributed as dist
        device = torch.device(device)
        px, ok = None, 1
        if os.environ.get("PK_PEER_ALLREDUCE", "1") == "0" or device.type != "cuda" or dist.get_backend(group) != "nccl":
            ok = 0
        else:
            try:
                px = cls(n_params, device, group)
            except Exception as e:  # no P2P mapping on this machine, older torch, ...
                import logging
                logging.getLogger(__name__).warning("peer gradient exchange unavailable, using NCCL allreduce: %s", e)
                ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=device if device.type == "cuda" else "cpu")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return px if int(flag.item()) == 1 else None

    def next_slot(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """Advance to the next epoch; (gradient (n,), loss share (1,)) views of this rank's buffer for it."""
        self.epoch += 1
        o = self.FLAG_FLOATS + (self.epoch & 1) * self.npad
        return self.buf[o:o + self.n], self.buf[o + self.n:o + self.n + 1]

    def reduce_clamp_adam(self, params: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, n_layers: int,
                          n_channels: int, in_channels: int, step: int, lr: float, betas, eps: float,
                          prepared: Optional[torch.Tensor] = None, grad_sum: Optional[torch.Tensor] = None) -> None:
        from . import _native as N
        assert self.epoch >= 1, "next_slot() first"
        assert params.numel() == self.n and params.is_cuda and params.dtype == torch.float32
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(N.lib().pk_peer_clamp_adam(
            params.data_ptr(), self._grad_ptrs[self.epoch & 1], self._flag_ptrs, self.rank, self.world, self.epoch,
            exp_avg.data_ptr(), exp_avg_sq.data_ptr(), int(n_layers), int(n_channels), int(in_channels), float(lr),
            float(betas[0]), float(betas[1]), float(eps), int(step),
            prepared.data_ptr() if prepared is not None else None,
            grad_sum.data_ptr() if grad_sum is not None else None, self.loss_status.data_ptr(), stream))

    def allreduce(self) -> None:
        """Sum this epoch's buffer (gradient + loss share, as written through `next_slot()`) over the ranks, in place on
        every rank: the sliced reduce-scatter / all-gather of pk_peer_allreduce (wide network).  `loss_status[1]` is
        raised when a peer did not arrive in time; the exchange is unusable afterwards."""
        from . import _native as N
        assert self.epoch >= 1, "next_slot() first"
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(N.lib().pk_peer_allreduce(self._grad_ptrs[self.epoch & 1], self._flag_ptrs, self.rank, self.world,
                                          self.epoch, self.npad, self.loss_status.data_ptr(), stream))

    # ---- graph-replayable form (pk_update_ctl) ------------------------------------------------------------------
    def slot(self, parity: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """(gradient, loss share) views of this rank's buffer of the given epoch parity (fixed addresses: what a
        captured graph of the step writes into)."""
        o = self.FLAG_FLOATS + (parity & 1) * self.npad
        return self.buf[o:o + self.n], self.buf[o + self.n:o + self.n + 1]

    def reduce_clamp_adam_ctl(self, parity: int, params: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
                              n_layers: int, n_channels: int, in_channels: int, ctl: torch.Tensor,
                              prepared: Optional[torch.Tensor] = None, grad_sum: Optional[torch.Tensor] = None) -> None:
        """`reduce_clamp_adam` with epoch, step and the Adam constants read from the device control block: capture
        [ops.update_ctl_advance(ctl), ops.dqn_grad(..., out_grad / out_loss = slot(p)), this] once per parity p and
        replay the two graphs alternately, parity 1 first (the epoch counts from 1)."""
        from . import _native as N
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(N.lib().pk_peer_clamp_adam_ctl(
            params.data_ptr(), self._grad_ptrs[parity & 1], self._flag_ptrs, self.rank, self.world, ctl.data_ptr(),
            exp_avg.data_ptr(), exp_avg_sq.data_ptr(), int(n_layers), int(n_channels), int(in_channels),
            prepared.data_ptr() if prepared is not None else None,
            grad_sum.data_ptr() if grad_sum is not None else None, self.loss_status.data_ptr(), stream))
