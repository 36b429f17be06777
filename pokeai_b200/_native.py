"""ctypes binding of libpokeai_b200.so (include/pokeai_b200.h).

There is no CPU fallback: if the shared library is missing, import of this module's
``lib()`` raises, and every product entry point fails loudly.  torch is used by callers
for device memory and streams only; nothing here takes a torch type.
"""
from __future__ import annotations

import ctypes as C
import re
from pathlib import Path
from typing import List, Optional

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "_lib" / "libpokeai_b200.so"
HEADER_PATH = PKG.parent / "include" / "pokeai_b200.h"

PK_MAX_ACTIONS = 9
PK_MAX_LAYERS = 8
PK_OBS_BYTES = 144
PK_SELECT_ARGMAX, PK_SELECT_EPSGREEDY, PK_SELECT_SOFTMAX = 0, 1, 2


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"pokeai_b200 native call failed (code {code}): {msg}")
        self.code = code


class pk_tables(C.Structure):
    _fields_ = [("n_species", C.c_int32), ("n_moves", C.c_int32), ("n_items", C.c_int32),
                ("species_type_mask", C.POINTER(C.c_uint32))]


class pk_mlp(C.Structure):
    _fields_ = [("n_layers", C.c_int32), ("n_channels", C.c_int32), ("in_channels", C.c_int32),
                ("reserved", C.c_int32),
                ("w", C.c_void_p * PK_MAX_LAYERS), ("b", C.c_void_p * PK_MAX_LAYERS),
                ("w_out", C.c_void_p), ("b_out", C.c_void_p)]


class pk_select_params(C.Structure):
    _fields_ = [("mode", C.c_int32), ("eps", C.c_float), ("temperature", C.c_float), ("reserved", C.c_int32),
                ("u_explore", C.c_void_p), ("u_pick", C.c_void_p), ("probs", C.c_void_p)]


class pk_batch(C.Structure):
    _fields_ = [("n", C.c_int32), ("n_global", C.c_int32), ("state", C.c_void_p), ("next_state", C.c_void_p),
                ("action", C.c_void_p), ("reward", C.c_void_p), ("terminal", C.c_void_p)]


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_SIGNATURES = {
    "pk_abi_version": (C.c_int, []),
    "pk_last_error": (C.c_char_p, []),
    "pk_init": (C.c_int, [C.POINTER(pk_tables), _i]),
    "pk_debug_set_trace": (C.c_int, [_vp]),
    "pk_set_overlap_launches": (C.c_int, [_i]),
    "pk_debug_set_pairing": (C.c_int, [_i]),
    "pk_debug_set_decide_tc": (C.c_int, [_i]),
    "pk_debug_set_dense_tc": (C.c_int, [_i]),
    "pk_debug_set_hidden_tc": (C.c_int, [_i]),
    "pk_n_channels": (C.c_int, []),
    "pk_sm_count": (C.c_int, []),
    "pk_mlp_prepared_floats": (_sz, [_i, _i, _i]),
    "pk_mlp_param_floats": (_sz, [_i, _i, _i]),
    "pk_mlp_prepare": (C.c_int, [C.POINTER(pk_mlp), _vp, _vp]),
    "pk_encode": (C.c_int, [_vp, _i, _i, _vp, _vp, _vp]),
    "pk_mlp_forward": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "pk_select": (C.c_int, [_vp, _vp, _i, _i, C.POINTER(pk_select_params), _vp, _vp]),
    "pk_decide": (C.c_int, [_vp, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, C.POINTER(pk_select_params), _vp]),
    "pk_gemm_bf16": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "pk_gemm_bf16_tn": (C.c_int, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "pk_wide_train_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "pk_wide_train_forward": (C.c_int, [_vp, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "pk_wide_train_backward": (C.c_int, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "pk_wide_train_backward_ev": (C.c_int, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "pk_wide_dqn_dq": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _vp, _vp, _vp]),
    "pk_wide_prepared_bytes": (_sz, [_i, _i, _i]),
    "pk_wide_prepare": (C.c_int, [C.POINTER(pk_mlp), _vp, _vp]),
    "pk_wide_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "pk_wide_forward": (C.c_int, [_vp, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, C.POINTER(pk_select_params), _vp]),
    "pk_dqn_scratch_floats": (_sz, [_i, _i, _i, _i]),
    "pk_dqn_grad": (C.c_int, [_vp, _vp, _i, _i, _i, C.POINTER(pk_batch), _f, _i, _vp, _vp, _vp, _vp]),
    "pk_dqn_grad_static": (C.c_int, [_vp, _vp, _i, _i, _i, C.POINTER(pk_batch), _f, _i, _vp, _vp, _vp]),
    "pk_clamp_adam": (C.c_int, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _f, _f, _i, _vp, _vp]),
    "pk_update_ctl_advance": (C.c_int, [_vp, _vp]),
    "pk_clamp_adam_ctl": (C.c_int, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "pk_peer_clamp_adam_ctl": (C.c_int, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "pk_peer_allreduce": (C.c_int, [_vp, _vp, _i, _i, C.c_uint32, C.c_size_t, _vp, _vp]),
    "pk_peer_clamp_adam": (C.c_int, [_vp, _vp, _vp, _i, _i, C.c_uint32, _vp, _vp, _i, _i, _i, _f, _f, _f, _f, _i, _vp, _vp,
                                     _vp, _vp]),
    "pk_bn_param_floats": (_sz, [_i, _i, _i]),
    "pk_bn_workspace_floats": (_sz, [_i, _i, _i, _i]),
    "pk_bn_backward_scratch_floats": (_sz, [_i, _i, _i, _i]),
    "pk_bn_forward_train": (C.c_int, [_vp, _vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "pk_bn_dqn_dq": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _vp, _vp, _vp]),
    "pk_bn_backward": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "pk_bn_packed_workspace_floats": (_sz, [_i, _i, _i, _i, _i]),
    "pk_bn_forward_train_packed": (C.c_int, [_vp, _vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "pk_bn_packed_backward_scratch_floats": (_sz, [_i, _i, _i, _i, _i]),
    "pk_bn_backward_packed": (C.c_int, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "pk_clamp_adam_flat": (C.c_int, [_vp, _vp, _vp, _vp, _sz, _f, _f, _f, _f, _i, _vp]),
    "pk_session_create": (C.c_int, [C.POINTER(_vp), _i, _i, _i]),
    "pk_session_destroy": (None, [_vp]),
    "pk_session_load_weights": (C.c_int, [_vp, C.POINTER(pk_mlp)]),
    "pk_session_use_prepared": (C.c_int, [_vp, _vp, _i, _i]),
    "pk_session_decide_host": (C.c_int, [_vp, _vp, _i, _i, _f, _f, _vp, _vp, _vp, _vp]),
    "pk_session_submit_host": (C.c_int, [_vp, _vp, _i, _i, _f, _f, _vp, _vp, _vp, _vp]),
    "pk_session_wait": (C.c_int, [_vp]),
    "pk_session_encode_host": (C.c_int, [_vp, _vp, _i, _vp, _vp]),
    "pk_session_fetch": (C.c_int, [_vp, _i, _vp, _vp]),
    "pk_session_features": (_vp, [_vp]),
    "pk_session_mask": (_vp, [_vp]),
}

_lib: Optional[C.CDLL] = None


def declared_symbols() -> List[str]:
    """Every function include/pokeai_b200.h declares (parsed from the header)."""
    text = HEADER_PATH.read_text()
    return sorted(set(re.findall(r"^PK_API [\w \*]+?\b(pk_\w+)\(", text, flags=re.M)))


def lib() -> C.CDLL:
    """The loaded library.  Raises if it has not been built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m pokeai_b200.build` "
                               "(pokeai_b200 has no CPU or PyTorch fallback)")
        l = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise NativeError(rc, lib().pk_last_error().decode("utf-8", "replace"))


_inited_device: Optional[int] = None


def ensure_init(device_index: int) -> None:
    """pk_init on `device_index` with the packaged tables (once per process)."""
    global _inited_device
    if _inited_device == device_index:
        return
    if _inited_device is not None:
        raise RuntimeError(f"pokeai_b200 is bound to cuda:{_inited_device}; one process drives one GPU")
    import numpy as np
    from .tables import load_tables
    t = load_tables()
    tm = np.ascontiguousarray(t.species_type_mask, dtype=np.uint32)
    st = pk_tables(t.n_species, t.n_moves, t.n_items, tm.ctypes.data_as(C.POINTER(C.c_uint32)))
    check(lib().pk_init(C.byref(st), int(device_index)))
    assert lib().pk_n_channels() == t.n_channels
    _inited_device = device_index
