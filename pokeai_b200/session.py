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
               temperature: float = 1.0, packed: Optional[np.ndarray] = None) -> Tuple[np.ndarray, np.ndarray]:
        """Run the first B staged records -- or, with `packed`, the first B records of that C-contiguous PK_OBS_DTYPE
        array, read in place (zero-copy over PCIe when it is pinned, see `pinned_records`; a driver-side copy when it is
        pageable).  Returns views (q (B, A), action (B,)) of the pinned result buffers, valid until the next call."""
        if not 0 <= B <= self.max_batch:
            raise ValueError(f"batch {B} exceeds the session capacity {self.max_batch}")
        src = self._pin_packed.data_ptr()
        if packed is not None:
            if packed.dtype != PK_OBS_DTYPE or not packed.flags.c_contiguous or len(packed) < B:
                raise ValueError("packed must be a C-contiguous PK_OBS_DTYPE array of at least B records")
            src = packed.ctypes.data
        N.check(N.lib().pk_session_decide_host(
            self._h, src, B, int(mode), float(eps), float(temperature),
            self._pin_u.data_ptr(), self._pin_u.data_ptr() + 4 * self.max_batch,
            self._pin_q.data_ptr(), self._pin_action.data_ptr()))
        return self.q[:B], self.action[:B]

    def submit(self, B: int, mode: int = N.PK_SELECT_ARGMAX, eps: float = 0.0, temperature: float = 1.0,
               packed: Optional[np.ndarray] = None) -> None:
        """Queue the first B staged records on the session's stream and return at once (pk_session_submit_host).
        The staging and result buffers must not be touched until `wait`.  `packed`: read the records from this
        caller-owned array instead of the staging buffer (no host copy when it is pinned, see `pinned_records`)."""
        if not 0 <= B <= self.max_batch:
            raise ValueError(f"batch {B} exceeds the session capacity {self.max_batch}")
        src = self._pin_packed.data_ptr()
        if packed is not None:
            if packed.dtype != PK_OBS_DTYPE or not packed.flags.c_contiguous or len(packed) < B:
                raise ValueError("packed must be a C-contiguous PK_OBS_DTYPE array of at least B records")
            src = packed.ctypes.data
            self._keep = packed
        N.check(N.lib().pk_session_submit_host(
            self._h, src, B, int(mode), float(eps), float(temperature),
            self._pin_u.data_ptr(), self._pin_u.data_ptr() + 4 * self.max_batch,
            self._pin_q.data_ptr(), self._pin_action.data_ptr()))
        self._pending = B

    def wait(self) -> Tuple[np.ndarray, np.ndarray]:
        """Block until the submitted batch is done; views (q (B, A), action (B,)) as `decide` returns them."""
        B = self.__dict__.pop("_pending", None)
        if B is None:
            raise RuntimeError("wait() without a submitted batch")
        N.check(N.lib().pk_session_wait(self._h))
        self._keep = None
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


def pinned_records(n: int) -> np.ndarray:
    """(n,) zeroed PK_OBS_DTYPE records in page-locked host memory: a driver packs its requests straight into such an
    array (`protocol.pack_requests`-style field writes, or a byte copy) and hands it to `DecisionPipeline.submit(...,
    pinned=True)`, so the records cross PCIe once, read by the kernel, with no staging copy."""
    t = torch.zeros((int(n), PK_OBS_BYTES), dtype=torch.uint8).pin_memory()
    return t.numpy().reshape(-1).view(PK_OBS_DTYPE)  # the view keeps the tensor alive


class DecisionPipeline:
    """`depth` decision sessions used in rotation, so that a lock-step driver over several groups of battles
     keeps `depth` batches in flight: while one batch runs on the GPU the host packs the next and reads
    the previous one's actions.  Greedy decisions (AgentVal semantics: q[mask == 0] = -inf, argmax).

        pipe = DecisionPipeline(agent, max_batch=4096)
        for actions in pipe.map(batches):        # batches: iterable of observation lists / packed record arrays
            ...                                   # actions: np.int32 (B,), in submission order

    or by hand: `submit(batch)` ... `result()` (first in, first out; at most `depth` outstanding)."""

    def __init__(self, agent, max_batch: int, depth: int = 2, materialise_features: bool = True):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.agent = agent
        fe = agent._feature_extractor
        self._fe = fe
        device = fe.session(1).device
        self._sessions = [DecisionSession(max_batch, fe.output_dim, device, materialise_features) for _ in range(depth)]
        self._head = 0      # next session to submit on
        self._inflight = 0  # submitted, not yet collected
        self._prep = None
        self._version = None

    def _point_weights(self) -> None:
        m = self.agent._model
        ps = self.__dict__.get("_plist")
        if ps is None or self._plist_model is not m:
            ps = self._plist = list(m.parameters()) + list(m.buffers())
            self._plist_model = m
        version = [getattr(m, "_generation", 0)] + [p._version for p in ps] + [p.data_ptr() for p in ps]
        if version == self._version:
            return
        for s in self._sessions:  # batches in flight still read the old weights: let them finish first
            N.check(N.lib().pk_session_wait(s._h))
        if hasattr(m, "prepared"):
            prep = m.prepared(self._sessions[0].device)
            for s in self._sessions:
                s.use_prepared(prep, m.n_layers, m.n_channels)
            self._prep = prep
        else:
            for s in self._sessions:
                s.load_model(m)
        self._version = version

    def submit(self, obs_list, pinned: bool = False) -> None:
        """`pinned=True`: `obs_list` is a record array from `pinned_records` that the caller leaves untouched until
        the matching `result()`; the kernel reads it in place and no staging copy is made."""
        if self._inflight == len(self._sessions):
            raise RuntimeError("pipeline full: collect a result() first")
        self._point_weights()
        s = self._sessions[self._head]
        n = len(obs_list)
        if n > s.max_batch:
            raise ValueError(f"batch {n} exceeds the pipeline capacity {s.max_batch}")
        if pinned:
            s.submit(n, packed=obs_list)
        else:
            self._fe.pack(obs_list, out=s.packed)
            s.submit(n)
        self._head = (self._head + 1) % len(self._sessions)
        self._inflight += 1

    def result(self, with_q: bool = False):
        """Actions (and Q, -inf NOT applied) of the oldest outstanding batch; copies, safe to keep."""
        if self._inflight == 0:
            raise RuntimeError("nothing submitted")
        s = self._sessions[(self._head - self._inflight) % len(self._sessions)]
        q, a = s.wait()
        self._inflight -= 1
        return (q.copy(), a.copy()) if with_q else a.copy()

    def map(self, batches, pinned: bool = False):
        for batch in batches:
            if self._inflight == len(self._sessions):
                yield self.result()
            self.submit(batch, pinned)
        while self._inflight:
            yield self.result()

    def close(self) -> None:
        for s in self._sessions:
            s.close()
