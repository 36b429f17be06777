"""Host-buffer decision session: pinned host records in, Q values and actions out.

This is what ``Agent.act`` / ``Agent.act_batch`` use: the caller's observations are packed
straight into a pinned staging array, one C-ABI call (``pk_session_decide_host``) copies
them to the device, runs the fused decision kernel and copies Q / actions back.  The
device buffers and the stream belong to the native session; torch only provides the
pinned host memory.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch

from . import _native as N
from .packed import PK_OBS_BYTES, PK_OBS_DTYPE


class DecisionSession:
    def __init__(self, max_batch: int, A: int, device=0, materialise_features: bool = True):
        device = torch.device(device if not isinstance(device, int) else f"cuda:{device}")
        if device.type != "cuda":
            raise ValueError("DecisionSession needs a CUDA device (pokeai_b200 has no CPU path)")
        self.device = device
        self.max_batch, self.A = int(max_batch), int(A)
        self.materialise = bool(materialise_features)
        N.ensure_init(device.index if device.index is not None else torch.cuda.current_device())
        self._h = C.c_void_p()
        N.check(N.lib().pk_session_create(C.byref(self._h), self.max_batch, self.A, int(self.materialise)))
        # pinned staging (torch owns it; numpy views for the packer and the caller)
        self._pin_packed = torch.empty((self.max_batch, PK_OBS_BYTES), dtype=torch.uint8, pin_memory=True)
        self._pin_q = torch.empty((self.max_batch, self.A), dtype=torch.float32, pin_memory=True)
        self._pin_action = torch.empty((self.max_batch,), dtype=torch.int32, pin_memory=True)
        self._pin_u = torch.empty((2, self.max_batch), dtype=torch.float32, pin_memory=True)
        self.packed = self._pin_packed.numpy().reshape(-1).view(PK_OBS_DTYPE)  # (max_batch,) records
        self.q = self._pin_q.numpy()
        self.action = self._pin_action.numpy()
        self.u_explore = self._pin_u.numpy()[0]
        self.u_pick = self._pin_u.numpy()[1]
        self._weights_key = None

    def close(self) -> None:
        if self._h:
            N.lib().pk_session_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_state_dict(self, sd: dict) -> None:
        """Host (or device) tensors with the reference's state_dict keys -> session weights."""
        n_layers = 0
        while f"layers.{n_layers}.weight" in sd:
            n_layers += 1
        if n_layers == 0:
            raise ValueError("state_dict does not look like an MLPModel")
        keep = []
        m = N.pk_mlp()
        H, Cin = sd["layers.0.weight"].shape[0], sd["layers.0.weight"].shape[1]
        m.n_layers, m.n_channels, m.in_channels = n_layers, H, Cin

        def host(t):
            t = t.detach().to("cpu", torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        for i in range(n_layers):
            m.w[i] = host(sd[f"layers.{i}.weight"])
            m.b[i] = host(sd[f"layers.{i}.bias"])
        m.w_out = host(sd["output.weight"])
        m.b_out = host(sd["output.bias"])
        N.check(N.lib().pk_session_load_weights(self._h, C.byref(m)))
        self._prepared_ref = None  # the session reads its own copy now

    def use_prepared(self, prepared: torch.Tensor, n_layers: int, n_channels: int) -> None:
        """Point the session at caller-owned prepared weights on the device (no copy)."""
        assert prepared.is_cuda and prepared.dtype == torch.float32 and prepared.is_contiguous()
        self._prepared_ref = prepared  # keep alive
        torch.cuda.current_stream(prepared.device).synchronize()  # the session runs on its own stream
        self._weights_key = None
        N.check(N.lib().pk_session_use_prepared(self._h, prepared.data_ptr(), int(n_layers), int(n_channels)))

    def load_model(self, model) -> None:
        """Load (or refresh, if any parameter changed) the weights of an MLPModel."""
        key = tuple((p.data_ptr(), p._version) for p in model.parameters()) + \
            tuple((b.data_ptr(), b._version) for b in model.buffers())
        if key != self._weights_key:
            self.load_state_dict(model.folded_state_dict() if hasattr(model, "folded_state_dict")
                                 else model.state_dict())
            self._weights_key = key

    def decide(self, B: int, mode: int = N.PK_SELECT_ARGMAX, eps: float = 0.0,
               temperature: float = 1.0) -> Tuple[np.ndarray, np.ndarray]:
        """Run the first B staged records.  Returns views (q (B, A), action (B,)) of the pinned
        result buffers, valid until the next call."""
        if not 0 <= B <= self.max_batch:
            raise ValueError(f"batch {B} exceeds the session capacity {self.max_batch}")
        N.check(N.lib().pk_session_decide_host(
            self._h, self._pin_packed.data_ptr(), B, int(mode), float(eps), float(temperature),
            self._pin_u.data_ptr(), self._pin_u.data_ptr() + 4 * self.max_batch,
            self._pin_q.data_ptr(), self._pin_action.data_ptr()))
        return self.q[:B], self.action[:B]

    def encode(self, B: int) -> Tuple[np.ndarray, np.ndarray]:
        """Encode the first B staged records: (feat (B, C, A) fp32, mask (B, A) int32) host arrays."""
        if not self.materialise:
            raise RuntimeError("this session does not materialise features")
        if not 0 <= B <= self.max_batch:
            raise ValueError(f"batch {B} exceeds the session capacity {self.max_batch}")
        Cn = N.lib().pk_n_channels()
        feat = np.empty((B, Cn, self.A), dtype=np.float32)
        mask = np.empty((B, self.A), dtype=np.int32)
        N.check(N.lib().pk_session_encode_host(self._h, self._pin_packed.data_ptr(), B, feat.ctypes.data,
                                               mask.ctypes.data))
        return feat, mask

    def features_to_host(self, B: int) -> Tuple[np.ndarray, np.ndarray]:
        """Copy the materialised features / masks of the last call back (debug / transform())."""
        if not self.materialise:
            raise RuntimeError("this session does not materialise features")
        Cn = N.lib().pk_n_channels()
        feat = np.empty((B, Cn, self.A), dtype=np.float32)
        mask = np.empty((B, self.A), dtype=np.int32)
        N.check(N.lib().pk_session_fetch(self._h, B, feat.ctypes.data, mask.ctypes.data))
        return feat, mask
