"""GPU parity: the CUDA decision path (through the C ABI) against the oracle and the golden
vectors produced by the real reference.  Features / masks / actions are bit-exact; Q values are
fp32 with a different summation order than torch-CPU, tolerance 1e-5 abs + 1e-5 rel."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN, bits

pytestmark = pytest.mark.gpu

Q_RTOL, Q_ATOL = 1e-5, 1e-5


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _to_dev(rec, dev):
    from pokeai_b200 import ops
    return ops.packed_to_device(rec, dev)


def _sd(z, prefix):
    return {k[len(prefix):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix)}


# ---------------------------------------------------------------------------------------------
# encode
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["full6", "ragged6", "ragged3", "ragged1", "full4"])
def test_encode_matches_reference_golden(synth_golden, dev, name):
    from pokeai_b200 import ops
    rec, ps, feat, mask = synth_golden[name]
    A = 4 + ps - 1
    f, m = ops.encode(_to_dev(rec, dev), A)
    assert f.shape == feat.shape and m.dtype == torch.int32
    assert np.array_equal(bits(f.cpu().numpy()), bits(feat))
    assert np.array_equal(m.cpu().numpy(), mask)


def test_encode_real_battle_log(fixture_golden, tables, dev):
    from pokeai_b200 import ops
    from pokeai_b200.observation import RLPolicyObservation, battle_status_from_json, possible_actions_from_json
    from pokeai_b200.packed import pack_observations
    dec, feat, mask = fixture_golden
    obs = [RLPolicyObservation(battle_status_from_json(d["battle_status"]), {},
                               possible_actions_from_json(d["possible_actions"])) for d in dec]
    rec = pack_observations(obs, tables, max_actions=6)
    f, m = ops.encode(_to_dev(rec, dev), 6)
    assert np.array_equal(bits(f.cpu().numpy()), bits(feat)) and np.array_equal(m.cpu().numpy(), mask)


@pytest.mark.parametrize("B,ps,kind", [(1, 6, "full"), (3, 6, "ragged"), (147, 6, "ragged"), (149, 5, "ragged"),
                                       (1000, 6, "ragged"), (4096, 6, "full"), (777, 2, "ragged"), (5000, 3, "ragged")])
def test_encode_matches_oracle_sizes(tables, dev, B, ps, kind):
    from oracle import packed_oracle
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(B, seed=1000 + B, party_size=ps, kind=kind, tables=tables)
    A = 4 + ps - 1
    f, m = ops.encode(_to_dev(rec, dev), A)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    assert np.array_equal(bits(f.cpu().numpy()), bits(fo))
    assert np.array_equal(m.cpu().numpy(), mo)


def test_encode_repeated_moves_and_edge_records(tables, dev):
    """duplicate move ids set the element once; n_act = 0; all state one-hot codes; fainted 0/100."""
    from oracle import packed_oracle
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(64, seed=5, party_size=6, kind="ragged", tables=tables)
    rec["act"]["move"][:, :, 1] = rec["act"]["move"][:, :, 0]  # duplicates inside every action
    rec["n_act"][0] = 0
    rec["status"][1] = (7, 7)
    rec["hp_cur"][1] = (0, 0)
    rec["hp_max"][1] = (100, 100)
    rec["boost"][2] = -6
    rec["boost"][3] = 6
    f, m = ops.encode(_to_dev(rec, dev), 9)
    fo, mo = packed_oracle.encode(rec, 6, tables)
    assert np.array_equal(bits(f.cpu().numpy()), bits(fo)) and np.array_equal(m.cpu().numpy(), mo)


def test_encode_empty_batch_and_bad_arguments(dev):
    from pokeai_b200 import ops
    from pokeai_b200._native import NativeError
    empty = torch.empty((0, 144), dtype=torch.uint8, device=dev)
    f, m = ops.encode(empty, 9)
    assert f.shape == (0, 617, 9) and m.shape == (0, 9)
    with pytest.raises(NativeError):
        ops.encode(torch.zeros((4, 144), dtype=torch.uint8, device=dev), 10)  # A > 9
    with pytest.raises(ValueError):
        ops.encode(torch.zeros((4, 100), dtype=torch.uint8, device=dev), 9)


def test_encode_full_size_properties(tables, dev):
    """65536 states (BASELINE config 3): size-independent properties instead of a dense oracle."""
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    B = 65536
    rec = synth_packed(B, seed=99, party_size=6, kind="full", tables=tables)
    f, m = ops.encode(_to_dev(rec, dev), 9)
    assert int(m.sum()) == B * 9
    # every column of a "full" record: 4 move columns with 3 ones, 5 switch columns with 7 ones
    ones = (f[:, 50:, :] == 1.0).sum(dim=1)
    assert torch.equal(ones[:, :4], torch.full((B, 4), 3, device=dev, dtype=ones.dtype))
    assert torch.equal(ones[:, 4:], torch.full((B, 5), 7, device=dev, dtype=ones.dtype))
    assert int((f[:, 50:, :] != 0).sum()) == B * (4 * 3 + 5 * 7)
    # state rows are broadcast over the columns
    assert torch.equal(f[:, :50, 0:1].expand(-1, -1, 9), f[:, :50, :])
    # idempotence: a second launch into the same buffer gives the same bytes
    f2, _ = ops.encode(_to_dev(rec, dev), 9, out_feat=f.clone())
    assert torch.equal(f.view(torch.int32), f2.view(torch.int32))
    # sample 512 rows against the oracle
    from oracle import packed_oracle
    idx = np.random.default_rng(0).choice(B, 512, replace=False)
    fo, _ = packed_oracle.encode(rec[idx], 6, tables)
    assert np.array_equal(bits(f[torch.from_numpy(idx).to(dev)].cpu().numpy()), bits(fo))


def test_decide_full_size_properties(tables, dev):
    """65536 states through the fused kernel (BASELINE configs[2]): with and without materialised features the Q
    values are the same bits; the action is the exact masked argmax of the kernel's own Q; the dense-input kernel on
    the materialised features agrees; 1024 sampled battles agree with the oracle; a second launch is bit-identical"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    B, A, nl, H = 65536, 9, 2, 64
    rec = synth_packed(B, seed=123, party_size=6, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=5)
    prepared = ops.prepare_weights(sd, dev)[0]
    pk = _to_dev(rec, dev)
    f, m, q, a = ops.decide(pk, A, prepared, nl, H)
    _, m2, q2, a2 = ops.decide(pk, A, prepared, nl, H, materialise=False)
    assert torch.equal(m, m2) and torch.equal(a, a2)
    # the two kernels (one / two battles per warp) order their sums differently: close, not identical
    assert float((q - q2).abs().max()) < 2e-5
    f3, m3, q3, a3 = ops.decide(pk, A, prepared, nl, H)
    assert torch.equal(q, q3) and torch.equal(a, a3) and torch.equal(f.view(torch.int32), f3.view(torch.int32))
    n_act = torch.from_numpy(rec["n_act"].astype(np.int64)).to(dev)
    legal = torch.arange(A, device=dev)[None, :] < n_act[:, None]
    assert torch.equal(m.bool(), legal)
    qm = torch.where(legal, q, torch.full_like(q, float("-inf")))
    assert torch.equal(a.long(), qm.argmax(dim=1))
    qd = ops.mlp_forward(prepared, nl, H, f)
    assert float((qd - q).abs().max()) < 2e-5
    idx = np.random.default_rng(1).choice(B, 1024, replace=False)
    fo, mo = packed_oracle.encode(rec[idx], 6, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    ti = torch.from_numpy(idx).to(dev)
    assert np.array_equal(bits(f[ti].cpu().numpy()), bits(fo)) and np.array_equal(m[ti].cpu().numpy(), mo)
    np.testing.assert_allclose(q[ti].cpu().numpy(), qo, rtol=Q_RTOL, atol=Q_ATOL)


# ---------------------------------------------------------------------------------------------
# forward
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["default", "small"])
def test_forward_matches_reference_golden(synth_golden, dev, name):
    """weights + Q produced by the real MLPModel (tools/make_golden.py)"""
    from pokeai_b200 import ops
    z = np.load(GOLDEN / "mlp_forward.npz")
    rec, ps, feat, mask = synth_golden[str(z[f"{name}_src"])]
    n = int(z[f"{name}_n"])
    A = 4 + ps - 1
    sd = _sd(z, f"{name}_sd_")
    prepared, nl, H, _ = ops.prepare_weights(sd, dev)
    assert nl == int(z[f"{name}_n_layers"]) and H == int(z[f"{name}_n_channels"])
    # dense-input kernel
    q_dense = ops.mlp_forward(prepared, nl, H, torch.from_numpy(feat[:n]).to(dev))
    np.testing.assert_allclose(q_dense.cpu().numpy(), z[f"{name}_q"], rtol=Q_RTOL, atol=Q_ATOL)
    # fused kernel from packed records, with and without materialised features
    for mat in (True, False):
        f, m, q, a = ops.decide(_to_dev(rec[:n], dev), A, prepared, nl, H, materialise=mat)
        np.testing.assert_allclose(q.cpu().numpy(), z[f"{name}_q"], rtol=Q_RTOL, atol=Q_ATOL)
        assert np.array_equal(m.cpu().numpy(), mask[:n])
        if mat:
            assert np.array_equal(bits(f.cpu().numpy()), bits(feat[:n]))


@pytest.mark.parametrize("ps,nl,H,B,kind", [(6, 2, 64, 4096, "full"), (6, 2, 64, 1500, "ragged"), (3, 3, 64, 700, "ragged"),
                                            (1, 3, 16, 300, "ragged"), (6, 1, 32, 300, "ragged"), (6, 2, 128, 400, "ragged"),
                                            (4, 4, 48, 333, "ragged")])
def test_decide_matches_oracle(tables, dev, ps, nl, H, B, kind):
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=7 * B + H, party_size=ps, kind=kind, tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H + nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    ao = ref_port.masked_argmax(qo, mo)
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    f, m, q, a = ops.decide(_to_dev(rec, dev), A, prepared, nl, H)
    assert np.array_equal(bits(f.cpu().numpy()), bits(fo)) and np.array_equal(m.cpu().numpy(), mo)
    qg = q.cpu().numpy()
    np.testing.assert_allclose(qg, qo, rtol=Q_RTOL, atol=Q_ATOL)
    # the action is the exact masked argmax of the kernel's own Q, and agrees with the oracle's
    # wherever the oracle's top two legal values are further apart than the tolerance
    assert np.array_equal(a.cpu().numpy(), ref_port.masked_argmax(qg, mo))
    qm = np.where(mo == 1, qo, -np.inf)
    top2 = np.sort(qm, axis=1)[:, -2:]
    clear = (top2[:, 1] - top2[:, 0]) > 1e-4
    assert np.array_equal(a.cpu().numpy()[clear], ao[clear])
    # dense kernel on the materialised features
    q2 = ops.mlp_forward(prepared, nl, H, f)
    np.testing.assert_allclose(q2.cpu().numpy(), qo, rtol=Q_RTOL, atol=Q_ATOL)


@pytest.mark.parametrize("ps,nl,H,B", [(6, 2, 64, 4097), (6, 2, 64, 1), (3, 3, 24, 301), (4, 1, 64, 77), (6, 4, 64, 500)])
def test_paired_and_single_kernels_agree_with_oracle(tables, dev, ps, nl, H, B):
    """pk_decide has two kernels for <= 64 channels (one battle per warp / two per warp, pk_decide2.cuh); both must
    give bit-exact features and masks and Q within tolerance, with and without materialised features, odd B included"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import _native as N, ops
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=B + H, party_size=ps, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H * nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    try:
        N.check(N.lib().pk_debug_set_decide_tc(0))
        for mode in (0, 2):
            N.check(N.lib().pk_debug_set_pairing(mode))
            for mat in (True, False):
                f, m, q, a = ops.decide(_to_dev(rec, dev), A, prepared, nl, H, materialise=mat)
                qg = q.cpu().numpy()
                np.testing.assert_allclose(qg, qo, rtol=Q_RTOL, atol=Q_ATOL, err_msg=f"pairing={mode} mat={mat}")
                assert np.array_equal(m.cpu().numpy(), mo)
                assert np.array_equal(a.cpu().numpy(), ref_port.masked_argmax(qg, mo))
                if mat:
                    assert np.array_equal(bits(f.cpu().numpy()), bits(fo)), f"pairing={mode}"
    finally:
        N.check(N.lib().pk_debug_set_pairing(1))
        N.check(N.lib().pk_debug_set_decide_tc(1))


@pytest.mark.parametrize("ps,nl,H,B", [(6, 2, 64, 4096), (6, 2, 64, 4097), (6, 2, 64, 1), (6, 2, 64, 13), (6, 3, 64, 700),
                                       (3, 2, 48, 301), (1, 4, 64, 65), (4, 2, 33, 100), (2, 3, 64, 2500),
                                       (6, 2, 64, 30011)])
def test_tcgen05_hidden_layers_match_fp32_oracle(tables, dev, ps, nl, H, B):
    """pk_decide_tc_kernel: the hidden layers of the eight battles of a warp group as one tcgen05.mma.kind::tf32
    contraction (3-term hi/lo split, TMEM accumulator; PK_DECIDE_TC, off by default).  Same bar as the FP32-pipe kernels: features / masks bit-exact, Q within
    1e-5, the action the exact masked argmax of the kernel's own Q; and the two paths agree closely.  Batch sizes
    cover a partly filled last round, a single battle, and more rounds than two."""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import _native as N, ops
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=13 * B + H, party_size=ps, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H + 5 * nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    qs = {}
    try:
        for mode in (2, 0):
            N.check(N.lib().pk_debug_set_decide_tc(mode))
            for mat in (True, False):
                f, m, q, a = ops.decide(_to_dev(rec, dev), A, prepared, nl, H, materialise=mat)
                qg = q.cpu().numpy()
                np.testing.assert_allclose(qg, qo, rtol=Q_RTOL, atol=Q_ATOL, err_msg=f"decide_tc={mode} mat={mat}")
                assert np.array_equal(m.cpu().numpy(), mo)
                assert np.array_equal(a.cpu().numpy(), ref_port.masked_argmax(qg, mo))
                if mat:
                    assert np.array_equal(bits(f.cpu().numpy()), bits(fo)), f"decide_tc={mode}"
                qs[(mode, mat)] = qg
    finally:
        N.check(N.lib().pk_debug_set_decide_tc(1))
    assert np.abs(qs[(2, True)] - qs[(0, True)]).max() < 5e-6
    assert np.abs(qs[(2, True)] - qs[(2, False)]).max() < 5e-6
    if nl <= 3 and 32 < H <= 64:  # the tensor-core kernel really ran (deeper nets fall back: shared memory)
        assert np.array_equal(qs[(2, True)], qs[(2, False)])  # same kernel with and without the feature stores


@pytest.mark.parametrize("ps,nl,H,B", [(6, 2, 64, 4096), (6, 3, 64, 513), (3, 2, 48, 200), (1, 4, 64, 65), (6, 2, 33, 100)])
def test_tensor_path_hidden_layer_matches_fp32_oracle(tables, dev, ps, nl, H, B):
    """the 64-channel hidden layers run on mma.sync tf32 with a 3-term hi/lo split (pk_hidden_tc); Q must stay inside
    the same fp32 tolerance as the FFMA path, features / masks bit-exact, and the two paths must agree closely"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import _native as N, ops
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=11 * B + H, party_size=ps, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H + 3 * nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    qs = {}
    try:
        N.check(N.lib().pk_debug_set_pairing(0))
        N.check(N.lib().pk_debug_set_decide_tc(0))
        for on in (1, 0):
            N.check(N.lib().pk_debug_set_hidden_tc(on))
            for mat in (True, False):
                f, m, q, a = ops.decide(_to_dev(rec, dev), A, prepared, nl, H, materialise=mat)
                qg = q.cpu().numpy()
                np.testing.assert_allclose(qg, qo, rtol=Q_RTOL, atol=Q_ATOL, err_msg=f"hidden_tc={on} mat={mat}")
                assert np.array_equal(m.cpu().numpy(), mo)
                assert np.array_equal(a.cpu().numpy(), ref_port.masked_argmax(qg, mo))
                if mat:
                    assert np.array_equal(bits(f.cpu().numpy()), bits(fo))
                qs[(on, mat)] = qg
    finally:
        N.check(N.lib().pk_debug_set_hidden_tc(0))
        N.check(N.lib().pk_debug_set_pairing(1))
        N.check(N.lib().pk_debug_set_decide_tc(1))
    assert np.abs(qs[(1, True)] - qs[(0, True)]).max() < 5e-6


def test_dense_forward_on_arbitrary_dense_input(tables, dev):
    """_calc_q_vector_batch accepts any float tensor, not only one-hot features"""
    from oracle import ref_port
    from pokeai_b200 import ops
    g = torch.Generator().manual_seed(3)
    x = torch.randn((37, tables.n_channels, 9), generator=g) * (torch.rand((37, tables.n_channels, 9), generator=g) < 0.3)
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=9)
    with torch.no_grad():
        qo = ref_port.mlp_forward(sd, x).numpy()
    prepared, nl, H, _ = ops.prepare_weights(sd, dev)
    q = ops.mlp_forward(prepared, nl, H, x.to(dev).contiguous())
    np.testing.assert_allclose(q.cpu().numpy(), qo, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("A,nl,H,B", [(9, 2, 64, 4099), (9, 2, 64, 1), (9, 2, 64, 14), (9, 2, 64, 15), (9, 3, 64, 700),
                                      (7, 2, 48, 301), (9, 4, 64, 65), (1, 2, 64, 100), (3, 2, 64, 77), (4, 3, 33, 129),
                                      (6, 2, 64, 2368), (8, 2, 64, 500)])
def test_dense_forward_on_tensor_cores_matches_fp32_oracle(tables, dev, A, nl, H, B):
    """pk_dense_tc_kernel (pk_mlp_forward's default for 64-channel nets): TMA-streamed dense blocks, operands in
    tensor memory, every layer a tcgen05.mma.kind::tf32 contraction with a 3-term hi/lo split.  Same bar as the FP32-pipe
    kernel (<= 1e-5 against the torch fp32 oracle, mlp_model.py:21-30) on sparse AND fully dense random input (so the lo
    terms matter), on blocks whose address is 16-, 4- and 12-byte aligned and that end at the last byte of their view,
    and the two kernels agree closely.  Sizes cover one battle, exactly one tile (14 at A = 9), one over, several tiles
    per CTA."""
    from oracle import ref_port
    from pokeai_b200 import _native as N, ops
    C = tables.n_channels
    g = torch.Generator().manual_seed(1000 * A + B)
    sd = ref_port.init_mlp_params(C, nl, H, seed=H + 3 * nl + A)
    prepared, _, _, _ = ops.prepare_weights(sd, dev)
    sparse = torch.randn((B, C, A), generator=g) * (torch.rand((B, C, A), generator=g) < 0.05)
    onehot = (torch.rand((B, C, A), generator=g) < 0.03).float()
    onehot[:, :50, :] = torch.rand((B, 50, 1), generator=g)      # ratio-valued state rows, the same in every column
    dense = torch.rand((B, C, A), generator=g) - 0.5
    for name, x in (("sparse", sparse), ("onehot", onehot), ("dense", dense)):
        with torch.no_grad():
            qo = ref_port.mlp_forward(sd, x).numpy()
        tol = dict(rtol=Q_RTOL, atol=Q_ATOL * max(1.0, float(np.abs(qo).max())))
        flat = torch.empty(B * C * A + 3, dtype=torch.float32, device=dev)
        qs = {}
        try:
            for tc in (1, 0):
                N.check(N.lib().pk_debug_set_dense_tc(tc))
                for off in (0, 1, 3):   # element offsets: 16-, 4- and 12-byte aligned blocks
                    xv = flat[off:off + B * C * A].view(B, C, A)
                    xv.copy_(x)
                    qg = ops.mlp_forward(prepared, nl, H, xv).cpu().numpy()
                    np.testing.assert_allclose(qg, qo, err_msg=f"{name} tc={tc} off={off}", **tol)
                    qs[(tc, off)] = qg
        finally:
            N.check(N.lib().pk_debug_set_dense_tc(1))
        assert np.array_equal(qs[(1, 0)], qs[(1, 1)]) and np.array_equal(qs[(1, 0)], qs[(1, 3)]), name
        assert np.abs(qs[(1, 0)] - qs[(0, 0)]).max() <= 2 * tol["atol"], name


# ---------------------------------------------------------------------------------------------
# select
# ---------------------------------------------------------------------------------------------
def _rand_q_mask(B, A, seed):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((B, A)).astype(np.float32)
    mask = (rng.random((B, A)) < 0.6).astype(np.int32)
    mask[: B // 2, 0] |= (mask[: B // 2].sum(1) == 0)  # second half keeps some all-illegal rows
    q[3, :] = 0.5
    q[5, 2] = np.inf
    return q, mask, rng


@pytest.mark.parametrize("A", [4, 6, 9])
def test_select_modes_match_oracle(dev, A):
    from oracle import ref_port
    from pokeai_b200 import ops
    from pokeai_b200._native import PK_SELECT_EPSGREEDY, PK_SELECT_SOFTMAX
    B = 2048
    q, mask, rng = _rand_q_mask(B, A, A)
    qd, md = torch.from_numpy(q).to(dev), torch.from_numpy(mask).to(dev)
    a = ops.select(qd, md).cpu().numpy()
    assert np.array_equal(a, ref_port.masked_argmax(q, mask))
    u1, u2 = rng.random(B).astype(np.float32), rng.random(B).astype(np.float32)
    u1d, u2d = torch.from_numpy(u1).to(dev), torch.from_numpy(u2).to(dev)
    e = ops.select(qd, md, ops.SelectSpec(PK_SELECT_EPSGREEDY, eps=0.3, u_explore=u1d, u_pick=u2d)).cpu().numpy()
    assert np.array_equal(e, ref_port.epsilon_greedy(q, mask, 0.3, u1, u2))
    finite = np.isfinite(np.where(mask == 1, q, 0)).all(axis=1)
    for T in (1.0, 0.25):
        probs = torch.zeros((B, A), dtype=torch.float32, device=dev)
        s = ops.select(qd, md, ops.SelectSpec(PK_SELECT_SOFTMAX, temperature=T, u_pick=u2d, probs=probs)).cpu().numpy()
        so, po = ref_port.softmax_sample(q, mask, T, u2)
        np.testing.assert_allclose(probs.cpu().numpy()[finite], po[finite], rtol=2e-6, atol=1e-7)
        # the draw agrees except where u lands within rounding of a CDF step
        cdf = np.cumsum(po, axis=1)
        near = (np.abs(cdf - u2[:, None]) < 1e-5).any(axis=1)
        ok = finite & ~near
        assert np.array_equal(s[ok], so[ok])
        legal_pick = mask[np.arange(B), s] == 1
        assert legal_pick[mask.sum(1) > 0].all()


def test_fused_select_modes(tables, dev):
    """epsilon-greedy and softmax inside the fused kernel equal the standalone kernel on the same Q"""
    from oracle import ref_port
    from pokeai_b200 import ops
    from pokeai_b200._native import PK_SELECT_EPSGREEDY, PK_SELECT_SOFTMAX
    from pokeai_b200.packed import synth_packed
    B, ps, A = 900, 6, 9
    rec = synth_packed(B, seed=77, party_size=ps, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=1)
    prepared, nl, H, _ = ops.prepare_weights(sd, dev)
    rng = np.random.default_rng(1)
    u1 = torch.from_numpy(rng.random(B).astype(np.float32)).to(dev)
    u2 = torch.from_numpy(rng.random(B).astype(np.float32)).to(dev)
    pk = _to_dev(rec, dev)
    for spec in (ops.SelectSpec(PK_SELECT_EPSGREEDY, eps=0.5, u_explore=u1, u_pick=u2),
                 ops.SelectSpec(PK_SELECT_SOFTMAX, temperature=0.7, u_pick=u2)):
        _, m, q, a = ops.decide(pk, A, prepared, nl, H, spec=spec, materialise=False)
        a2 = ops.select(q, m, spec)
        assert torch.equal(a, a2)
        e = ref_port.epsilon_greedy(q.cpu().numpy(), m.cpu().numpy(), 0.5, u1.cpu().numpy(), u2.cpu().numpy())
        if spec.mode == PK_SELECT_EPSGREEDY:
            assert np.array_equal(a.cpu().numpy(), e)


def test_session_host_buffers(tables, dev):
    """the plain-C host-buffer session (what a non-torch binding would call)"""
    import ctypes as C
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import _native as N
    from pokeai_b200.packed import synth_packed
    N.ensure_init(0)
    B, A = 1000, 9
    rec = synth_packed(B, seed=31, party_size=6, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=4)
    sess = C.c_void_p()
    N.check(N.lib().pk_session_create(C.byref(sess), B, A, 1))
    try:
        m = N.pk_mlp()
        m.n_layers, m.n_channels, m.in_channels = 2, 64, tables.n_channels
        keep = []
        for i in range(2):
            w = np.ascontiguousarray(sd[f"layers.{i}.weight"].numpy()); b = np.ascontiguousarray(sd[f"layers.{i}.bias"].numpy())
            keep += [w, b]
            m.w[i], m.b[i] = w.ctypes.data, b.ctypes.data
        wo = np.ascontiguousarray(sd["output.weight"].numpy()); bo = np.ascontiguousarray(sd["output.bias"].numpy())
        m.w_out, m.b_out = wo.ctypes.data, bo.ctypes.data
        N.check(N.lib().pk_session_load_weights(sess, C.byref(m)))
        q = np.empty((B, A), dtype=np.float32)
        a = np.empty((B,), dtype=np.int32)
        buf = np.ascontiguousarray(rec)
        N.check(N.lib().pk_session_decide_host(sess, buf.ctypes.data, B, 0, 0.0, 1.0, None, None, q.ctypes.data, a.ctypes.data))
        fo, mo = packed_oracle.encode(rec, 6, tables)
        with torch.no_grad():
            qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
        np.testing.assert_allclose(q, qo, rtol=Q_RTOL, atol=Q_ATOL)
        assert np.array_equal(a, ref_port.masked_argmax(q, mo))
    finally:
        N.lib().pk_session_destroy(sess)
