"""Host side of the wide-network (tensor-core) path: the same MLPModel (mlp_model.py:5-30) at n_channels that
are a multiple of 256 (BASELINE.json configs[4]: 4 x 2048) runs as bf16 tcgen05 GEMMs with fused bias + ReLU
(pokeai_b200/csrc/pk_wide.cuh) instead of the fp32 small-net kernel.  No CPU path: everything raises without the
CUDA library and a device.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _native as N
from .ops import SelectSpec, _check_packed, _dev, _stream_ptr, n_channels, packed_to_device

WIDE_MULTIPLE = 256


def is_wide(n_channels_: int) -> bool:
    """True when the network takes the tensor-core path (the fp32 small-net kernel stops at 128 channels)."""
    return n_channels_ >= WIDE_MULTIPLE and n_channels_ % WIDE_MULTIPLE == 0


def gemm_bf16(a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
              out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """act(a @ w.T + bias): a (M, K) bf16, w (N, K) bf16 (a state_dict weight cast to bf16), fp32 accumulation."""
    dev = _dev(a)
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and a.is_contiguous() and w.is_contiguous()
    M, K = a.shape
    Nn, K2 = w.shape
    assert K == K2
    N.ensure_init(dev.index if dev.index is not None else torch.cuda.current_device())
    out = torch.empty((M, Nn), dtype=out_dtype, device=dev)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == Nn and bias.is_cuda
    ob = out.data_ptr() if out_dtype == torch.bfloat16 else None
    of = out.data_ptr() if out_dtype == torch.float32 else None
    assert ob is not None or of is not None, "out_dtype must be bfloat16 or float32"
    N.check(N.lib().pk_gemm_bf16(a.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None, ob, of,
                                 M, Nn, K, 1 if relu else 0, _stream_ptr(dev)))
    return out


def prepare_wide(state_dict: dict, device, out: Optional[torch.Tensor] = None) -> tuple:
    """state_dict (reference key names) -> (prepared uint8 buffer on device, n_layers, n_channels)."""
    dev = torch.device(device)
    N.ensure_init(dev.index if dev.index is not None else torch.cuda.current_device())
    nl = 0
    while f"layers.{nl}.weight" in state_dict:
        nl += 1
    w0 = state_dict["layers.0.weight"]
    H, Cin = int(w0.shape[0]), int(w0.shape[1])
    lib = N.lib()
    nbytes = lib.pk_wide_prepared_bytes(nl, H, Cin)
    if nbytes == 0:
        raise ValueError(f"network {nl} x {H} is not supported by the tensor-core path (n_channels % 256 != 0)")
    keep = []
    m = N.pk_mlp()
    m.n_layers, m.n_channels, m.in_channels = nl, H, Cin
    for i in range(nl):
        w = state_dict[f"layers.{i}.weight"].detach().to(dev, torch.float32).contiguous()
        b = state_dict[f"layers.{i}.bias"].detach().to(dev, torch.float32).contiguous()
        keep += [w, b]
        m.w[i], m.b[i] = w.data_ptr(), b.data_ptr()
    wo = state_dict["output.weight"].detach().to(dev, torch.float32).contiguous()
    bo = state_dict["output.bias"].detach().to(dev, torch.float32).contiguous()
    keep += [wo, bo]
    m.w_out, m.b_out = wo.data_ptr(), bo.data_ptr()
    prepared = out if out is not None else torch.empty(nbytes, dtype=torch.uint8, device=dev)
    assert prepared.numel() == nbytes and prepared.dtype == torch.uint8
    N.check(lib.pk_wide_prepare(C.byref(m), prepared.data_ptr(), _stream_ptr(dev)))
    torch.cuda.current_stream(dev).synchronize()  # the fp32 staging copies in `keep` may be freed now
    return prepared, nl, H


def workspace(n_layers: int, H: int, B: int, A: int, device) -> torch.Tensor:
    nbytes = N.lib().pk_wide_workspace_bytes(n_layers, H, n_channels(), B, A)
    if nbytes == 0:
        raise ValueError("unsupported wide-network shape")
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def wide_forward(packed: torch.Tensor, A: int, prepared: torch.Tensor, n_layers: int, H: int,
                 ws: Optional[torch.Tensor] = None, spec: Optional[SelectSpec] = None,
                 out_mask: Optional[torch.Tensor] = None, out_q: Optional[torch.Tensor] = None,
                 out_action: Optional[torch.Tensor] = None):
    """Packed records (device) -> (mask (B, A) int32, q (B, A) fp32, action (B,) int32)."""
    dev = _dev(packed)
    B = _check_packed(packed)
    if ws is None:
        ws = workspace(n_layers, H, B, A, dev)
    mask = out_mask if out_mask is not None else torch.empty((B, A), dtype=torch.int32, device=dev)
    q = out_q if out_q is not None else torch.empty((B, A), dtype=torch.float32, device=dev)
    act = out_action if out_action is not None else torch.empty((B,), dtype=torch.int32, device=dev)
    sp = spec.to_c(B) if spec is not None else None
    N.check(N.lib().pk_wide_forward(packed.data_ptr(), B, A, prepared.data_ptr(), n_layers, H, ws.data_ptr(),
                                    mask.data_ptr(), q.data_ptr(), act.data_ptr(),
                                    C.byref(sp) if sp is not None else None, _stream_ptr(dev)))
    return mask, q, act


class WideRunner:
    """Batched decisions for a wide MLPModel: host records -> pinned staging -> H2D -> pk_wide_forward -> D2H.
    The counterpart of the native session the small network uses (Agent.q_and_action_batch dispatches here)."""

    def __init__(self, model, A: int, device):
        self.device = torch.device(device)
        self.A = A
        self.model = model
        self._prepared = None
        self._key = None
        self._cap = 0

    def _ensure(self, n: int) -> None:
        m = self.model
        key = (getattr(m, "_generation", 0),) + tuple((p.data_ptr(), p._version) for p in m.parameters())
        if self._prepared is None or key != self._key:
            self._prepared = prepare_wide(m.folded_state_dict(), self.device)[0]
            self._key = key
        if n > self._cap:
            cap = max(n, 2 * self._cap)
            self._pin = torch.empty((cap, N.PK_OBS_BYTES), dtype=torch.uint8, pin_memory=True)
            self._dpk = torch.empty((cap, N.PK_OBS_BYTES), dtype=torch.uint8, device=self.device)
            self._ws = workspace(m.n_layers, m.n_channels, cap, self.A, self.device)
            self._mask = torch.empty((cap, self.A), dtype=torch.int32, device=self.device)
            self._q = torch.empty((cap, self.A), dtype=torch.float32, device=self.device)
            self._act = torch.empty((cap,), dtype=torch.int32, device=self.device)
            self._hq = torch.empty((cap, self.A), dtype=torch.float32, pin_memory=True)
            self._ha = torch.empty((cap,), dtype=torch.int32, pin_memory=True)
            self._cap = cap

    def decide(self, records: np.ndarray, spec: Optional[SelectSpec] = None):
        n = len(records)
        self._ensure(n)
        self._pin.numpy()[:n] = np.ascontiguousarray(records).view(np.uint8).reshape(n, N.PK_OBS_BYTES)
        self._dpk[:n].copy_(self._pin[:n], non_blocking=True)
        m = self.model
        wide_forward(self._dpk[:n], self.A, self._prepared, m.n_layers, m.n_channels, ws=self._ws, spec=spec,
                     out_mask=self._mask[:n], out_q=self._q[:n], out_action=self._act[:n])
        self._hq[:n].copy_(self._q[:n], non_blocking=True)
        self._ha[:n].copy_(self._act[:n], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._hq.numpy()[:n], self._ha.numpy()[:n]


# ---- forward + backward (Trainer._update's loss.backward() for the wide network) ---------------------------------
def gemm_bf16_tn(a_km: torch.Tensor, w_kn: torch.Tensor) -> torch.Tensor:
    """a_km.T @ w_kn in fp32: a_km (K, M) bf16, w_kn (K, N) bf16 -- the weight-gradient contraction dW = dZ^T H."""
    dev = _dev(a_km)
    assert a_km.dtype == torch.bfloat16 and w_kn.dtype == torch.bfloat16 and a_km.is_contiguous() and w_kn.is_contiguous()
    K, M = a_km.shape
    K2, Nn = w_kn.shape
    assert K == K2
    out = torch.empty((M, Nn), dtype=torch.float32, device=dev)
    N.check(N.lib().pk_gemm_bf16_tn(a_km.data_ptr(), w_kn.data_ptr(), out.data_ptr(), M, Nn, K, _stream_ptr(dev)))
    return out


def train_workspace(n_layers: int, H: int, B: int, A: int, device) -> torch.Tensor:
    nbytes = N.lib().pk_wide_train_workspace_bytes(n_layers, H, n_channels(), B, A)
    if nbytes == 0:
        raise ValueError("unsupported wide-network shape")
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def train_forward(packed: torch.Tensor, A: int, prepared: torch.Tensor, n_layers: int, H: int, ws: torch.Tensor,
                  out_q: Optional[torch.Tensor] = None, out_mask: Optional[torch.Tensor] = None,
                  col_select: Optional[torch.Tensor] = None):
    """Forward that keeps the activations in `ws` for train_backward.  Returns (mask (B, A), q).
    col_select None: q is (B, A), every column is a GEMM row.  col_select (B,) int32: q is (B,), only that column of
    each battle is evaluated (the gather of trainer.py:163)."""
    dev = _dev(packed)
    B = _check_packed(packed)
    mask = out_mask if out_mask is not None else torch.empty((B, A), dtype=torch.int32, device=dev)
    if col_select is not None:
        assert col_select.dtype == torch.int32 and col_select.numel() == B and col_select.is_cuda
    q = out_q if out_q is not None else torch.empty((B,) if col_select is not None else (B, A), dtype=torch.float32,
                                                    device=dev)
    N.check(N.lib().pk_wide_train_forward(packed.data_ptr(), B, A,
                                          col_select.data_ptr() if col_select is not None else None,
                                          prepared.data_ptr(), n_layers, H, ws.data_ptr(), mask.data_ptr(), q.data_ptr(),
                                          _stream_ptr(dev)))
    return mask, q


def train_backward(prepared: torch.Tensor, n_layers: int, H: int, B: int, A: int, ws: torch.Tensor, dq: torch.Tensor,
                   out_grad: Optional[torch.Tensor] = None, chosen_only: bool = False,
                   layer_events: Optional[list] = None) -> torch.Tensor:
    """dq = d loss / d q ((B, A), or (B,) after a col_select forward) -> flat fp32 gradient in state_dict order.
    layer_events: n_layers torch.cuda.Event objects (each recorded at least once before); event l is recorded when the
    gradient of layers >= l and of the head is complete (see `layer_grad_buckets`)."""
    dev = _dev(dq)
    assert dq.dtype == torch.float32 and dq.is_contiguous() and dq.numel() == (B if chosen_only else B * A)
    n = N.lib().pk_mlp_param_floats(n_layers, H, n_channels())
    grad = out_grad if out_grad is not None else torch.empty(n, dtype=torch.float32, device=dev)
    evp = None
    if layer_events is not None:
        assert len(layer_events) == n_layers
        evp = (C.c_void_p * n_layers)(*[int(e.cuda_event) for e in layer_events])
        assert all(evp[i] for i in range(n_layers)), "record every layer event once before the first use"
    N.check(N.lib().pk_wide_train_backward_ev(prepared.data_ptr(), n_layers, H, B, A, 1 if chosen_only else 0,
                                              ws.data_ptr(), dq.data_ptr(), grad.data_ptr(), evp, _stream_ptr(dev)))
    return grad


def layer_grad_buckets(n_layers: int, H: int) -> list:
    """[(lo, hi)] per layer: the contiguous slice of the flat gradient that is complete when layer event l fires and
    was not complete before (layer l's weight + bias; the last layer's bucket also holds the head)."""
    Cn = n_channels()
    out, off = [], 0
    for l in range(n_layers):
        size = H * (Cn if l == 0 else H) + H
        out.append([off, off + size])
        off += size
    out[-1][1] = off + H + 1
    return [tuple(b) for b in out]


def dqn_dq(q_sa, q_next_online, q_next_target, mask_next, reward, terminal, n_global: int, gamma: float,
           double_dqn: bool = True):
    """Double-DQN target + Huber loss + dq for the chosen columns (trainer.py:163-187).  Returns (dq (n,), loss (1,))."""
    dev = _dev(q_sa)
    n, A = q_next_target.shape
    dq = torch.empty(n, dtype=torch.float32, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    N.check(N.lib().pk_wide_dqn_dq(q_sa.data_ptr(), q_next_online.data_ptr() if q_next_online is not None else None,
                                   q_next_target.data_ptr(), mask_next.data_ptr(), reward.data_ptr(), terminal.data_ptr(),
                                   n, A, int(n_global), float(gamma), 1 if double_dqn else 0, dq.data_ptr(),
                                   loss.data_ptr(), _stream_ptr(dev)))
    return dq, loss
