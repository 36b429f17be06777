"""CPU: the C-ABI library builds for sm_100a, loads, and exports everything include/pokeai_b200.h declares.
No compute call is made here (there is no GPU and no CPU fallback)."""
import ctypes as C

import numpy as np
import pytest


def test_library_builds_loads_and_exports_header():
    from pokeai_b200 import build as b
    from pokeai_b200 import _native as N
    path = b.build()
    assert path.exists()
    lib = N.lib()
    declared = N.declared_symbols()
    assert len(declared) >= 20
    for s in declared:
        assert hasattr(lib, s), f"{s} declared in the header but not exported"
    assert set(N._SIGNATURES) == set(declared)
    assert lib.pk_abi_version() == 1


def test_every_entry_point_is_mapped_to_the_reference_in_integration_md():
    """INTEGRATION.md's entry point table names every symbol the header declares (beside what it replaces)"""
    from pathlib import Path
    from pokeai_b200 import _native as N
    text = (Path(__file__).resolve().parents[1] / "INTEGRATION.md").read_text()
    assert [s for s in N.declared_symbols() if s not in text] == []


def test_layout_sizes():
    from pokeai_b200 import _native as N
    lib = N.lib()
    # 43 777 parameters for the default network
    assert lib.pk_mlp_param_floats(2, 64, 617) == 43777
    assert lib.pk_mlp_param_floats(4, 2048, 617) == 13856769
    # + the hidden matrix once more as tcgen05 operand tiles (hi / lo, 64 x 64 each) when the padded width is 64, and
    # layer 0 as 20 k-block tiles of 128 rows x 32 floats (hi rows, lo rows)
    tc0 = 20 * 128 * 32
    assert lib.pk_mlp_prepared_floats(2, 64, 617) == (617 + 1) * 64 + 64 + 64 * 64 + 64 + 64 + 4 + 64 * 64 + 2 * 64 * 64 + tc0
    assert lib.pk_mlp_prepared_floats(3, 48, 617) == (618 * 64 + 64 + 2 * (64 * 64 + 64) + 64 + 4 + 2 * 64 * 64 + 4 * 64 * 64
                                                      + tc0)
    assert lib.pk_mlp_prepared_floats(3, 16, 617) == 618 * 32 + 32 + 2 * (32 * 32 + 32) + 32 + 4 + 2 * 32 * 32
    assert lib.pk_mlp_prepared_floats(0, 64, 617) == 0


def test_bn_layout_matches_the_module():
    """pk_bn_param_floats == the parameter count of MLPModel(bn=True) (conv layers without bias + BatchNorm affine +
    head), workspace sizes grow with the batch, and bad shapes give 0."""
    from pokeai_b200 import _native as N
    from pokeai_b200.mlp_model import MLPModel
    lib = N.lib()
    for nl, H in ((2, 64), (3, 24), (1, 100)):
        m = MLPModel((617, 9), 9, n_layers=nl, n_channels=H, bn=True)
        assert lib.pk_bn_param_floats(nl, H, 617) == sum(p.numel() for p in m.parameters())
        assert [k for k, _ in m.named_parameters()][:nl] == [f"layers.{i}.weight" for i in range(nl)]
    assert lib.pk_bn_param_floats(0, 64, 617) == 0 and lib.pk_bn_param_floats(9, 64, 617) == 0
    small, big = lib.pk_bn_workspace_floats(2, 64, 32, 9), lib.pk_bn_workspace_floats(2, 64, 64, 9)
    assert big - small == 2 * 2 * 32 * 9 * 64  # pre-norm and post-ReLU activations of every layer
    assert lib.pk_bn_backward_scratch_floats(2, 64, 64, 9) - lib.pk_bn_backward_scratch_floats(2, 64, 32, 9) == 2 * 32 * 9 * 64


def test_calls_fail_loudly_without_init_or_device():
    import torch
    from pokeai_b200 import _native as N
    lib = N.lib()
    if torch.cuda.is_available():
        pytest.skip("covers the no-GPU behaviour")
    rc = lib.pk_encode(None, 4, 9, None, None, None)
    assert rc == 3 and b"pk_init" in lib.pk_last_error()
    tm = np.zeros(251, dtype=np.uint32)
    t = N.pk_tables(251, 267, 47, tm.ctypes.data_as(C.POINTER(C.c_uint32)))
    rc = lib.pk_init(C.byref(t), 0)
    assert rc == 2  # CUDA error: no device; never a silent CPU path
    from pokeai_b200 import ops
    with pytest.raises(ValueError):
        ops.encode(torch.zeros((1, 144), dtype=torch.uint8), 9)


def test_record_layout_matches_header():
    """the numpy dtype of the host packer is byte-for-byte the C struct"""
    from pokeai_b200.packed import PK_ACT_DTYPE, PK_OBS_DTYPE
    off = {n: PK_OBS_DTYPE.fields[n][1] for n in PK_OBS_DTYPE.names}
    assert off == {"remaining": 0, "total": 2, "hp_cur": 4, "hp_max": 8, "status": 12, "boost": 14, "opp_species": 28,
                   "weather": 29, "n_act": 30, "side_cond": 31, "act": 32, "user_tag": 140}
    aoff = {n: PK_ACT_DTYPE.fields[n][1] for n in PK_ACT_DTYPE.names}
    assert aoff == {"flags": 0, "species": 1, "item": 2, "n_moves": 3, "move": 4}
    assert PK_OBS_DTYPE.itemsize == 144
