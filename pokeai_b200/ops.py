"""Device-level entry points: torch CUDA tensors in, torch CUDA tensors out, every byte of
work done by libpokeai_b200.so (pokeai_b200/csrc).  torch supplies memory and the stream.

These are thin: argument checking, output allocation, one C-ABI call.  The reference
functions each one replaces are cited in include/pokeai_b200.h.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch

from . import _native as N
from .packed import PK_OBS_BYTES, PK_OBS_DTYPE
from .tables import load_tables


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _dev(t: torch.Tensor) -> torch.device:
    if not t.is_cuda:
        raise ValueError("pokeai_b200.ops works on CUDA tensors only (there is no CPU path)")
    N.ensure_init(t.device.index if t.device.index is not None else torch.cuda.current_device())
    return t.device


def packed_to_device(packed: np.ndarray, device, pinned: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(B,) PK_OBS_DTYPE records -> uint8 CUDA tensor (B, 144)."""
    assert packed.dtype == PK_OBS_DTYPE
    host = torch.from_numpy(np.ascontiguousarray(packed).view(np.uint8).reshape(-1, PK_OBS_BYTES))
    return host.to(device, non_blocking=False)


def _check_packed(packed: torch.Tensor) -> int:
    if packed.dtype != torch.uint8 or packed.dim() != 2 or packed.shape[1] != PK_OBS_BYTES or not packed.is_contiguous():
        raise ValueError("packed observations must be a contiguous uint8 tensor of shape (B, 144)")
    return packed.shape[0]


def n_channels() -> int:
    return load_tables().n_channels


def encode(packed: torch.Tensor, A: int, out_feat: Optional[torch.Tensor] = None,
           out_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """packed (B,144) u8 -> feat (B, C, A) fp32, mask (B, A) int32.  feature_extractor.py:40-52."""
    dev = _dev(packed)
    B = _check_packed(packed)
    Cn = n_channels()
    feat = out_feat if out_feat is not None else torch.empty((B, Cn, A), dtype=torch.float32, device=dev)
    mask = out_mask if out_mask is not None else torch.empty((B, A), dtype=torch.int32, device=dev)
    assert feat.is_contiguous() and mask.is_contiguous() and feat.shape == (B, Cn, A) and mask.shape == (B, A)
    N.check(N.lib().pk_encode(packed.data_ptr(), B, A, feat.data_ptr(), mask.data_ptr(), _stream_ptr(dev)))
    return feat, mask


class SelectSpec:
    """How to pick the action: argmax (AgentVal), epsilon-greedy (AgentTrain) or softmax sampling."""

    def __init__(self, mode: int = N.PK_SELECT_ARGMAX, eps: float = 0.0, temperature: float = 1.0,
                 u_explore: Optional[torch.Tensor] = None, u_pick: Optional[torch.Tensor] = None,
                 probs: Optional[torch.Tensor] = None):
        self.mode, self.eps, self.temperature = mode, eps, temperature
        self.u_explore, self.u_pick, self.probs = u_explore, u_pick, probs

    def to_c(self, B: int) -> N.pk_select_params:
        for t in (self.u_explore, self.u_pick):
            if t is not None:
                assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() >= B
        p = N.pk_select_params()
        p.mode, p.eps, p.temperature = int(self.mode), float(self.eps), float(self.temperature)
        p.u_explore = self.u_explore.data_ptr() if self.u_explore is not None else None
        p.u_pick = self.u_pick.data_ptr() if self.u_pick is not None else None
        p.probs = self.probs.data_ptr() if self.probs is not None else None
        return p


def select(q: torch.Tensor, mask: torch.Tensor, spec: Optional[SelectSpec] = None,
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q (B, A) fp32, mask (B, A) int32 -> action (B,) int32.  agent.py:32-45."""
    dev = _dev(q)
    B, A = q.shape
    assert q.dtype == torch.float32 and mask.dtype == torch.int32 and q.is_contiguous() and mask.is_contiguous()
    assert mask.shape == (B, A)
    action = out if out is not None else torch.empty((B,), dtype=torch.int32, device=dev)
    sp = (spec or SelectSpec()).to_c(B)
    N.check(N.lib().pk_select(q.data_ptr(), mask.data_ptr(), B, A, C.byref(sp), action.data_ptr(), _stream_ptr(dev)))
    return action


def mlp_forward(prepared: torch.Tensor, n_layers: int, n_channels_: int, feat: torch.Tensor,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Dense features (B, C, A) -> q (B, A).  mlp_model.py:21-30 as called by agent.py:23-30."""
    dev = _dev(feat)
    assert feat.dtype == torch.float32 and feat.dim() == 3 and feat.is_contiguous()
    B, Cn, A = feat.shape
    q = out if out is not None else torch.empty((B, A), dtype=torch.float32, device=dev)
    N.check(N.lib().pk_mlp_forward(prepared.data_ptr(), n_layers, n_channels_, Cn, feat.data_ptr(), B, A, q.data_ptr(),
                                   _stream_ptr(dev)))
    return q


def decide(packed: torch.Tensor, A: int, prepared: torch.Tensor, n_layers: int, n_channels_: int,
           spec: Optional[SelectSpec] = None, materialise: bool = True, out_feat: Optional[torch.Tensor] = None,
           out_mask: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
           out_action: Optional[torch.Tensor] = None):
    """One fused launch: packed -> (feat | None, mask, q, action).
    AgentVal.act / AgentTrain.act for a whole batch (agent_val.py:8-11, agent_train.py:24-37)."""
    dev = _dev(packed)
    B = _check_packed(packed)
    Cn = n_channels()
    feat = None
    if materialise:
        feat = out_feat if out_feat is not None else torch.empty((B, Cn, A), dtype=torch.float32, device=dev)
    mask = out_mask if out_mask is not None else torch.empty((B, A), dtype=torch.int32, device=dev)
    q = out_q if out_q is not None else torch.empty((B, A), dtype=torch.float32, device=dev)
    action = out_action if out_action is not None else torch.empty((B,), dtype=torch.int32, device=dev)
    sp = (spec or SelectSpec()).to_c(B)
    N.check(N.lib().pk_decide(packed.data_ptr(), B, A, prepared.data_ptr(), n_layers, n_channels_,
                              feat.data_ptr() if feat is not None else None, mask.data_ptr(), q.data_ptr(),
                              action.data_ptr(), C.byref(sp), _stream_ptr(dev)))
    return feat, mask, q, action


def prepare_weights(state_dict: dict, device, out: Optional[torch.Tensor] = None):
    """state_dict (reference keys, mlp_model.py:6-19) -> (prepared fp32 CUDA tensor, n_layers, n_channels, flat params).
    The tensors may live anywhere; they are copied to `device` as one flat fp32 vector first."""
    device = torch.device(device)
    N.ensure_init(device.index if device.index is not None else torch.cuda.current_device())
    n_layers = 0
    while f"layers.{n_layers}.weight" in state_dict:
        n_layers += 1
    if n_layers == 0 or "output.weight" not in state_dict:
        raise ValueError("state_dict does not look like an MLPModel (layers.N.weight / output.weight)")
    if any(k.startswith("bn_layers") for k in state_dict):
        raise NotImplementedError("fold BatchNorm into the preceding conv before preparing (MLPModel.fold_bn)")
    H, Cin = state_dict["layers.0.weight"].shape[0], state_dict["layers.0.weight"].shape[1]
    keys = []
    for i in range(n_layers):
        keys += [f"layers.{i}.weight", f"layers.{i}.bias"]
    keys += ["output.weight", "output.bias"]
    flat = torch.cat([state_dict[k].detach().to(device=device, dtype=torch.float32).reshape(-1) for k in keys])
    assert flat.numel() == N.lib().pk_mlp_param_floats(n_layers, H, Cin)
    prepared = prepare_from_flat(flat, n_layers, H, Cin, out)
    return prepared, n_layers, H, flat


def prepare_from_flat(flat: torch.Tensor, n_layers: int, H: int, Cin: int, out: Optional[torch.Tensor] = None):
    dev = _dev(flat)
    assert flat.dtype == torch.float32 and flat.is_contiguous()
    n = N.lib().pk_mlp_prepared_floats(n_layers, H, Cin)
    prepared = out if out is not None else torch.empty((n,), dtype=torch.float32, device=dev)
    m = N.pk_mlp()
    m.n_layers, m.n_channels, m.in_channels = n_layers, H, Cin
    off = 0
    base = flat.data_ptr()
    for i in range(n_layers):
        cin = Cin if i == 0 else H
        m.w[i] = base + 4 * off
        off += H * cin
        m.b[i] = base + 4 * off
        off += H
    m.w_out = base + 4 * off
    off += H
    m.b_out = base + 4 * off
    N.check(N.lib().pk_mlp_prepare(C.byref(m), prepared.data_ptr(), _stream_ptr(dev)))
    return prepared


# ---- DQN update (trainer.py:152-198) ------------------------------------------------------------------------
def dqn_grad(prepared: torch.Tensor, prepared_target: torch.Tensor, n_layers: int, n_channels_: int, A: int,
             state: torch.Tensor, next_state: torch.Tensor, action: torch.Tensor, reward: torch.Tensor,
             terminal: torch.Tensor, gamma: float, double_dqn: bool = True, n_global: Optional[int] = None,
             scratch: Optional[torch.Tensor] = None, out_grad: Optional[torch.Tensor] = None,
             out_loss: Optional[torch.Tensor] = None):
    """Packed transitions (device) -> (flat pre-clamp gradient in state_dict order, loss (1,) tensor).
    state / next_state: (n, 144) uint8; action (n,) int32; reward (n,) fp32; terminal (n,) uint8."""
    dev = _dev(state)
    n = _check_packed(state)
    assert _check_packed(next_state) == n
    assert action.dtype == torch.int32 and reward.dtype == torch.float32 and terminal.dtype == torch.uint8
    assert action.numel() == n and reward.numel() == n and terminal.numel() == n
    Cn = n_channels()
    lib = N.lib()
    if scratch is None:
        scratch = torch.empty(lib.pk_dqn_scratch_floats(n_layers, n_channels_, Cn, n), dtype=torch.float32, device=dev)
    n_par = lib.pk_mlp_param_floats(n_layers, n_channels_, Cn)
    grad = out_grad if out_grad is not None else torch.empty(n_par, dtype=torch.float32, device=dev)
    loss = out_loss if out_loss is not None else torch.zeros(1, dtype=torch.float32, device=dev)
    assert grad.is_cuda and grad.dtype == torch.float32 and grad.is_contiguous() and grad.numel() == n_par
    assert loss.is_cuda and loss.dtype == torch.float32 and loss.numel() == 1
    b = N.pk_batch()
    b.n, b.n_global = n, (n if n_global is None else int(n_global))
    b.state, b.next_state = state.data_ptr(), next_state.data_ptr()
    b.action, b.reward, b.terminal = action.data_ptr(), reward.data_ptr(), terminal.data_ptr()
    N.check(lib.pk_dqn_grad(prepared.data_ptr(), prepared_target.data_ptr(), n_layers, n_channels_, A, C.byref(b),
                            float(gamma), 1 if double_dqn else 0, grad.data_ptr(), loss.data_ptr(), scratch.data_ptr(),
                            _stream_ptr(dev)))
    return grad, loss


def clamp_adam(params: torch.Tensor, grad: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
               n_layers: int, n_channels_: int, step: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
               prepared: Optional[torch.Tensor] = None) -> None:
    """In place: clamp(grad, -1, 1), Adam step on the flat parameter vector, refresh of `prepared`."""
    dev = _dev(params)
    for t in (params, grad, exp_avg, exp_avg_sq):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == params.numel()
    N.check(N.lib().pk_clamp_adam(params.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
                                  n_layers, n_channels_, n_channels(), float(lr), float(betas[0]), float(betas[1]),
                                  float(eps), int(step), prepared.data_ptr() if prepared is not None else None,
                                  _stream_ptr(dev)))


def make_update_ctl(device, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, step: int = 0, epoch: int = 0):
    """A device-resident pk_update_ctl (include/pokeai_b200.h): what lets a captured CUDA graph of the update step be
    replayed -- Adam's step count and the peer-exchange epoch advance on the device (`update_ctl_advance`)."""
    host = np.zeros(8, dtype=np.uint32)
    host[0], host[1] = epoch, np.int32(step).view(np.uint32)
    host[2:6] = np.array([lr, betas[0], betas[1], eps], dtype=np.float32).view(np.uint32)
    return torch.from_numpy(host.view(np.int32)).to(device)


def update_ctl_advance(ctl: torch.Tensor) -> None:
    N.check(N.lib().pk_update_ctl_advance(ctl.data_ptr(), _stream_ptr(_dev(ctl))))


def clamp_adam_ctl(params: torch.Tensor, grad: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
                   n_layers: int, n_channels_: int, ctl: torch.Tensor, prepared: Optional[torch.Tensor] = None) -> None:
    """`clamp_adam` with the step count and the Adam constants read from the device control block `ctl`."""
    dev = _dev(params)
    for t in (params, grad, exp_avg, exp_avg_sq):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == params.numel()
    N.check(N.lib().pk_clamp_adam_ctl(params.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
                                      n_layers, n_channels_, n_channels(), ctl.data_ptr(),
                                      prepared.data_ptr() if prepared is not None else None, _stream_ptr(dev)))


def unflatten(flat: torch.Tensor, n_layers: int, H: int, Cin: int) -> dict:
    """Flat parameter / gradient vector -> dict with the reference's state_dict keys and shapes."""
    out, off = {}, 0
    for i in range(n_layers):
        cin = Cin if i == 0 else H
        out[f"layers.{i}.weight"] = flat[off:off + H * cin].view(H, cin, 1)
        off += H * cin
        out[f"layers.{i}.bias"] = flat[off:off + H]
        off += H
    out["output.weight"] = flat[off:off + H].view(1, H, 1)
    off += H
    out["output.bias"] = flat[off:off + 1]
    return out
