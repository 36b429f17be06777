"""pk_dqn_grad_static: the DQN gradient entry point that allocates nothing (its scratch is a file-scope device array
inside the library, include/pokeai_b200.h).  It has no counterpart in the reference.
Checked here through the C ABI against pk_dqn_grad (caller-owned scratch) on the same inputs, on
two calls in a row that share the static scratch.  Collected last on purpose
(file name).
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
G_RTOL, G_ATOL = 1e-4, 1e-7      # the gradient tolerance of tests/test_gpu_update.py


def _transitions(n, ps, tables, seed):
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    rng = np.random.default_rng(seed)
    dev = torch.device(DEV)
    rec_s = synth_packed(n, seed=seed, party_size=ps, kind="ragged", tables=tables)
    rec_n = synth_packed(n, seed=seed + 1, party_size=ps, kind="ragged", tables=tables)
    action = (rng.random(n) * rec_s["n_act"]).astype(np.int32)
    reward = rng.uniform(-1, 1, n).astype(np.float32)
    reward[rng.random(n) < 0.1] *= 8.0
    terminal = (rng.random(n) < 0.05).astype(np.uint8)
    return (ops.packed_to_device(rec_s, dev), ops.packed_to_device(rec_n, dev), torch.from_numpy(action).to(dev),
            torch.from_numpy(reward).to(dev), torch.from_numpy(terminal).to(dev))


def _batch(N, tensors, n=None):
    s, sn, a, r, t = tensors
    b = N.pk_batch()
    b.n = b.n_global = int(s.shape[0]) if n is None else n
    b.state, b.next_state = s.data_ptr(), sn.data_ptr()
    b.action, b.reward, b.terminal = a.data_ptr(), r.data_ptr(), t.data_ptr()
    return b


@pytest.mark.parametrize("nl,H,ps", [(2, 64, 6), (3, 48, 3)])
def test_static_scratch_gradient_matches_caller_scratch(tables, nl, H, ps):
    from oracle import ref_port
    from pokeai_b200 import _native as N, ops
    dev = torch.device(DEV)
    A = 4 + ps - 1
    Cn = tables.n_channels
    prepared = ops.prepare_weights(ref_port.init_mlp_params(Cn, nl, H, seed=31), dev)[0]     # also runs pk_init
    prepared_t = ops.prepare_weights(ref_port.init_mlp_params(Cn, nl, H, seed=32), dev)[0]
    lib = N.lib()
    n_par = lib.pk_mlp_param_floats(nl, H, Cn)
    stream = torch.cuda.current_stream(dev).cuda_stream
    # two calls in a row on different batches (the second smaller: it must not see the first one's partials)
    for n, seed, double in ((3000, 41, 1), (517, 43, 1), (517, 43, 0)):
        tensors = _transitions(n, ps, tables, seed)
        b = _batch(N, tensors)
        scratch = torch.empty(lib.pk_dqn_scratch_floats(nl, H, Cn, n), dtype=torch.float32, device=dev)
        g_ref = torch.empty(n_par, dtype=torch.float32, device=dev)
        l_ref = torch.zeros(1, dtype=torch.float32, device=dev)
        N.check(lib.pk_dqn_grad(prepared.data_ptr(), prepared_t.data_ptr(), nl, H, A, C.byref(b), 0.95, double,
                                g_ref.data_ptr(), l_ref.data_ptr(), scratch.data_ptr(), stream))
        # outputs exist before the call; the call itself must not allocate: torch's allocator statistics stay put
        g_new = torch.full((n_par,), float("nan"), dtype=torch.float32, device=dev)
        l_new = torch.zeros(1, dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
        free_before = torch.cuda.mem_get_info(dev)[0]
        N.check(lib.pk_dqn_grad_static(prepared.data_ptr(), prepared_t.data_ptr(), nl, H, A, C.byref(b), 0.95, double,
                                       g_new.data_ptr(), l_new.data_ptr(), stream))
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
        assert torch.cuda.mem_get_info(dev)[0] >= free_before - (1 << 20)   # nothing was cudaMalloc'ed by the call
        assert bool(torch.isfinite(g_new).all()) and float(g_new.abs().max()) > 0
        np.testing.assert_allclose(g_new.cpu().numpy(), g_ref.cpu().numpy(), rtol=G_RTOL, atol=G_ATOL,
                                   err_msg=f"n={n} double={double}")
        assert abs(float(l_new.item()) - float(l_ref.item())) < 1e-6
        if nl == 2:  # owned accumulators in shared memory (fixed summation order): in fact the same bits; the deeper
            assert torch.equal(g_new, g_ref) and torch.equal(l_new, l_ref)   # net falls back to atomics (order varies)


def test_static_scratch_rejects_oversized_requests(tables):
    from oracle import ref_port
    from pokeai_b200 import _native as N, ops
    dev = torch.device(DEV)
    nl, H, ps = 2, 64, 6
    Cn = tables.n_channels
    prepared = ops.prepare_weights(ref_port.init_mlp_params(Cn, nl, H, seed=31), dev)[0]
    lib = N.lib()
    tensors = _transitions(64, ps, tables, 5)
    n_big = 1 << 20
    assert lib.pk_dqn_scratch_floats(nl, H, Cn, n_big) > (8 << 20)
    b = _batch(N, tensors, n=n_big)      # rejected on its size before anything is read or launched
    g = torch.zeros(lib.pk_mlp_param_floats(nl, H, Cn), dtype=torch.float32, device=dev)
    rc = lib.pk_dqn_grad_static(prepared.data_ptr(), prepared.data_ptr(), nl, H, 9, C.byref(b), 0.95, 1, g.data_ptr(), None,
                                torch.cuda.current_stream(dev).cuda_stream)
    assert rc == 1   # PK_ERR_INVALID
    msg = lib.pk_last_error().decode()
    assert "static scratch" in msg and "pk_dqn_grad" in msg
    torch.cuda.synchronize()
    assert float(g.abs().max()) == 0.0
