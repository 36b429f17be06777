"""Train-mode BatchNorm for MLPModel(bn=True): host side of the pk_bn_* entry points (include/pokeai_b200.h).

The reference's Trainer never calls .eval() on its model / target_model (pokeai/ai/generic_move_model/trainer.py:41-43),
so with ``bn: true`` every forward of ``_update`` (trainer.py:163, 174-175, 181-182) uses the statistics of the batch
it is given and moves the running statistics, and ``loss.backward()`` (:192) differentiates through them
(mlp_model.py:12-15, 24-26).  torch holds the memory; all arithmetic is in csrc/pk_bn.cuh.  There is no CPU path.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _native as N


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def param_floats(n_layers: int, H: int, C: int) -> int:
    return int(N.lib().pk_bn_param_floats(n_layers, H, C))


def flat_parameters(model, device) -> torch.Tensor:
    """The parameters in nn.Module.parameters() order as one fp32 vector on `device` (a copy)."""
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).to(device, torch.float32).contiguous()
    assert flat.numel() == param_floats(model.n_layers, model.n_channels, model.in_channels)
    return flat


def stacked_running(model, device) -> torch.Tensor:
    """(n_layers, 2, H): running_mean, running_var of every BatchNorm1d (a copy)."""
    return torch.stack([torch.stack([b.running_mean.detach(), b.running_var.detach()]) for b in model.bn_layers]) \
        .to(device, torch.float32).contiguous()


def workspace(n_layers: int, H: int, B: int, A: int, device) -> torch.Tensor:
    n = int(N.lib().pk_bn_workspace_floats(n_layers, H, B, A))
    return torch.empty(n, dtype=torch.float32, device=device)


def forward_train(flat: torch.Tensor, running: Optional[torch.Tensor], n_layers: int, H: int, feat: torch.Tensor,
                  ws: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """feat (B, C, A) fp32 CUDA -> q (B, A) with batch statistics; `running` (n_layers, 2, H) is moved in place.
    Returns (q, ws); ws holds what pk_bn_backward needs."""
    if not (feat.is_cuda and feat.dtype == torch.float32 and feat.is_contiguous() and feat.dim() == 3):
        raise ValueError("features must be a contiguous fp32 CUDA tensor (B, C, A)")
    B, C, A = feat.shape
    dev = feat.device
    assert flat.is_cuda and flat.is_contiguous() and flat.numel() == param_floats(n_layers, H, C)
    if running is not None:
        assert running.is_cuda and running.is_contiguous() and running.shape == (n_layers, 2, H)
    need = int(N.lib().pk_bn_workspace_floats(n_layers, H, B, A))
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.float32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    N.check(N.lib().pk_bn_forward_train(flat.data_ptr(), running.data_ptr() if running is not None else None, n_layers,
                                        H, C, feat.data_ptr(), B, A, ws.data_ptr(), q.data_ptr(), _stream(dev)))
    return q, ws


def forward_train_packed(flat: torch.Tensor, running: Optional[torch.Tensor], n_layers: int, H: int, packed: torch.Tensor,
                         A: int, ws: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """packed (B, 144) u8 CUDA records -> (q (B, A), mask (B, A) int32, ws): the same forward with layer 0 gathered from
    the records (no dense feature block)."""
    from .ops import _check_packed, n_channels
    B = _check_packed(packed)
    dev, C = packed.device, n_channels()
    assert flat.is_cuda and flat.is_contiguous() and flat.numel() == param_floats(n_layers, H, C)
    if running is not None:
        assert running.is_cuda and running.is_contiguous() and running.shape == (n_layers, 2, H)
    need = int(N.lib().pk_bn_packed_workspace_floats(n_layers, H, C, B, A))
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.float32, device=dev)
    q = torch.empty((B, A), dtype=torch.float32, device=dev)
    mask = torch.empty((B, A), dtype=torch.int32, device=dev)
    N.check(N.lib().pk_bn_forward_train_packed(flat.data_ptr(), running.data_ptr() if running is not None else None,
                                               n_layers, H, C, packed.data_ptr(), B, A, ws.data_ptr(), q.data_ptr(),
                                               mask.data_ptr(), _stream(dev)))
    return q, mask, ws


def backward_packed(flat: torch.Tensor, n_layers: int, H: int, packed: torch.Tensor, A: int, ws: torch.Tensor,
                    dq: torch.Tensor, grad: torch.Tensor, scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """backward() for a forward that ran on packed records (forward_train_packed with the same `packed` and `ws`)."""
    from .ops import _check_packed, n_channels
    B = _check_packed(packed)
    dev, C = packed.device, n_channels()
    need = int(N.lib().pk_bn_packed_backward_scratch_floats(n_layers, H, C, B, A))
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, dtype=torch.float32, device=dev)
    assert grad.is_cuda and grad.is_contiguous() and grad.numel() == flat.numel() and dq.is_contiguous()
    N.check(N.lib().pk_bn_backward_packed(flat.data_ptr(), n_layers, H, C, packed.data_ptr(), B, A, ws.data_ptr(),
                                          dq.data_ptr(), scratch.data_ptr(), grad.data_ptr(), _stream(dev)))
    return scratch


def dqn_dq(q: torch.Tensor, action: torch.Tensor, reward: torch.Tensor, slot: torch.Tensor,
           q_next_online: Optional[torch.Tensor], q_next_target: torch.Tensor, mask_next: torch.Tensor, n_global: int,
           gamma: float, double_dqn: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> dq (n, A), loss (1,).  trainer.py:163-187."""
    n, A = q.shape
    dev = q.device
    dq = torch.empty((n, A), dtype=torch.float32, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    for t, dt in ((action, torch.int32), (reward, torch.float32), (slot, torch.int32), (mask_next, torch.int32)):
        assert t.is_cuda and t.dtype == dt and t.is_contiguous()
    N.check(N.lib().pk_bn_dqn_dq(q.data_ptr(), action.data_ptr(), reward.data_ptr(), slot.data_ptr(),
                                 q_next_online.data_ptr() if q_next_online is not None else None,
                                 q_next_target.data_ptr(), mask_next.data_ptr(), n, A, int(n_global), float(gamma),
                                 1 if double_dqn else 0, dq.data_ptr(), loss.data_ptr(), _stream(dev)))
    return dq, loss


def backward(flat: torch.Tensor, n_layers: int, H: int, feat: torch.Tensor, ws: torch.Tensor, dq: torch.Tensor,
             grad: torch.Tensor, scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The flat gradient (parameters() order, not clamped) of sum(q * dq) through the forward that filled `ws`."""
    B, C, A = feat.shape
    dev = feat.device
    need = int(N.lib().pk_bn_backward_scratch_floats(n_layers, H, B, A))
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, dtype=torch.float32, device=dev)
    assert grad.is_cuda and grad.is_contiguous() and grad.numel() == flat.numel() and dq.is_contiguous()
    N.check(N.lib().pk_bn_backward(flat.data_ptr(), n_layers, H, C, feat.data_ptr(), B, A, ws.data_ptr(), dq.data_ptr(),
                                   scratch.data_ptr(), grad.data_ptr(), _stream(dev)))
    return scratch


def clamp_adam_flat(params: torch.Tensor, grad: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, step: int,
                    lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8) -> None:
    for t in (params, grad, exp_avg, exp_avg_sq):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == params.numel()
    N.check(N.lib().pk_clamp_adam_flat(params.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
                                       params.numel(), float(lr), float(betas[0]), float(betas[1]), float(eps), int(step),
                                       _stream(params.device)))


def model_forward_train(model, x: torch.Tensor) -> torch.Tensor:
    """MLPModel(bn=True).forward in training mode on dense CUDA input: batch statistics, running statistics and
    num_batches_tracked move as nn.BatchNorm1d's do (mlp_model.py:24-26)."""
    dev = x.device
    flat = flat_parameters(model, dev)
    running = stacked_running(model, dev)
    q, _ = forward_train(flat, running, model.n_layers, model.n_channels, x.contiguous().float())
    with torch.no_grad():
        for l, b in enumerate(model.bn_layers):
            b.running_mean.copy_(running[l, 0])
            b.running_var.copy_(running[l, 1])
            b.num_batches_tracked += 1
    return q
