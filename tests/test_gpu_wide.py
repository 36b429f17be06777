"""GPU parity tests of the wide-network (tcgen05 tensor-core) path, through the C ABI.

Oracle: the reference's MLPModel arithmetic is torch CPU fp32 (oracle/ref_port.mlp_forward).  The tensor-core
path rounds operands and activations to bf16 and accumulates in fp32, so two checks are made:
  * against a bf16-EMULATING fp32 oracle (same roundings, fp32 accumulation on the CPU): tight tolerance, catches
    any layout / descriptor / pipeline error;
  * against the plain fp32 oracle: the bf16 tolerance of SURVEY.md 8c (|dq| <= 2e-2), argmax agreement reported.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _bf(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


@pytest.mark.parametrize("M,N,K,relu,bias", [(128, 256, 64, False, False), (300, 256, 128, True, True),
                                             (1000, 512, 640, True, True), (129, 2048, 2048, False, True),
                                             (4096, 1024, 1024, True, False), (64, 256, 96, False, True),
                                             (300, 96, 200, True, True), (777, 640, 2048, False, False)])
def test_gemm_bf16_matches_fp32_matmul_of_bf16_operands(dev, M, N, K, relu, bias):
    from pokeai_b200 import wide
    g = torch.Generator().manual_seed(M + N + K)
    a = _bf(torch.randn((M, K), generator=g))
    w = _bf(torch.randn((N, K), generator=g) / np.sqrt(K))
    b = torch.randn((N,), generator=g) if bias else None
    ref = a @ w.t()
    if b is not None:
        ref = ref + b
    if relu:
        ref = ref.clamp_min(0)
    out32 = wide.gemm_bf16(a.to(dev, torch.bfloat16), w.to(dev, torch.bfloat16),
                           b.to(dev) if b is not None else None, relu=relu, out_dtype=torch.float32)
    np.testing.assert_allclose(out32.cpu().numpy(), ref.numpy(), rtol=2e-4, atol=2e-4)
    out16 = wide.gemm_bf16(a.to(dev, torch.bfloat16), w.to(dev, torch.bfloat16),
                           b.to(dev) if b is not None else None, relu=relu, out_dtype=torch.bfloat16)
    np.testing.assert_allclose(out16.float().cpu().numpy(), ref.numpy(), rtol=1e-2, atol=1e-2)


def test_gemm_refuses_unsupported_shapes(dev):
    from pokeai_b200 import wide
    a = torch.zeros((64, 100), dtype=torch.bfloat16, device=dev)
    w = torch.zeros((256, 100), dtype=torch.bfloat16, device=dev)
    with pytest.raises(RuntimeError):
        wide.gemm_bf16(a, w)  # K not a multiple of 8 (TMA row pitch)
    a = torch.zeros((64, 64), dtype=torch.bfloat16, device=dev)
    w = torch.zeros((100, 64), dtype=torch.bfloat16, device=dev)
    with pytest.raises(RuntimeError):
        wide.gemm_bf16(a, w)  # N not a multiple of 32


def _emulated_forward(sd, feat, nl):
    """fp32 CPU forward with the tensor-core path's roundings: bf16 inputs, weights and hidden activations,
    fp32 accumulation, fp32 biases and head."""
    B, C, A = feat.shape
    x = _bf(torch.from_numpy(feat).permute(0, 2, 1).reshape(B * A, C))
    for i in range(nl):
        w = _bf(sd[f"layers.{i}.weight"][:, :, 0])
        x = _bf((x @ w.t() + sd[f"layers.{i}.bias"]).clamp_min(0))
    q = x @ sd["output.weight"][0, :, 0] + sd["output.bias"][0]
    return q.reshape(B, A).numpy()


@pytest.mark.parametrize("ps,nl,H,B,kind", [(6, 2, 256, 300, "ragged"), (6, 4, 2048, 512, "full"), (3, 3, 512, 77, "ragged")])
def test_wide_forward_matches_oracle(tables, dev, ps, nl, H, B, kind):
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import ops, wide
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=B + H, party_size=ps, kind=kind, tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H + nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    with torch.no_grad():
        q32 = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
        qem = _emulated_forward(sd, fo, nl)
    prepared, nl2, H2 = wide.prepare_wide(sd, dev)
    assert (nl2, H2) == (nl, H)
    m, q, a = wide.wide_forward(ops.packed_to_device(rec, dev), A, prepared, nl, H)
    torch.cuda.synchronize()
    qg = q.cpu().numpy()
    assert np.array_equal(m.cpu().numpy(), mo)                      # masks bit-exact
    np.testing.assert_allclose(qg, qem, rtol=3e-3, atol=3e-3)       # same roundings: only summation order differs
    np.testing.assert_allclose(qg, q32, rtol=0, atol=2e-2)          # bf16 tolerance against the fp32 reference
    assert np.array_equal(a.cpu().numpy(), ref_port.masked_argmax(qg, mo))  # exact argmax of its own Q
    agree = float((a.cpu().numpy() == ref_port.masked_argmax(q32, mo)).mean())
    assert agree > 0.9, f"argmax agreement with the fp32 reference only {agree:.3f}"


def test_wide_runner_host_records(tables, dev):
    """the host-buffer API (records in, numpy out) used by Agent.act_batch for wide models"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import wide
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    torch.manual_seed(1)
    model = MLPModel((tables.n_channels, 9), 9, n_layers=2, n_channels=256).eval()
    rec = synth_packed(100, seed=5, party_size=6, kind="ragged", tables=tables)
    r = wide.WideRunner(model, 9, dev)
    q, a = r.decide(rec)
    fo, mo = packed_oracle.encode(rec, 6, tables)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        q32 = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    np.testing.assert_allclose(q, q32, rtol=0, atol=2e-2)
    assert np.array_equal(a, ref_port.masked_argmax(q, mo))


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 640, 1000), (2048, 2048, 4104), (136, 32, 72)])
def test_gemm_tn_matches_fp32_matmul_of_bf16_operands(dev, M, N, K):
    """the contraction-over-rows GEMM (both operands MN-major in shared memory) used for weight gradients"""
    from pokeai_b200 import wide
    g = torch.Generator().manual_seed(M + N + K)
    a = _bf(torch.randn((K, M), generator=g))
    w = _bf(torch.randn((K, N), generator=g) / np.sqrt(K))
    ref = a.t() @ w
    out = wide.gemm_bf16_tn(a.to(dev, torch.bfloat16), w.to(dev, torch.bfloat16))
    np.testing.assert_allclose(out.cpu().numpy(), ref.numpy(), rtol=3e-4, atol=3e-4)


def _emulated_backward(sd, feat, nl, dq):
    """Gradients with the tensor-core path's roundings (bf16 activations, weights and dZ; fp32 accumulation),
    by hand: the same chain rule as autograd through mlp_forward."""
    B, C, A = feat.shape
    hs = [_bf(torch.from_numpy(feat).permute(0, 2, 1).reshape(B * A, C))]
    for i in range(nl):
        w = _bf(sd[f"layers.{i}.weight"][:, :, 0])
        hs.append(_bf((hs[-1] @ w.t() + sd[f"layers.{i}.bias"]).clamp_min(0)))
    d = torch.from_numpy(dq).reshape(-1)
    wout = sd["output.weight"][0, :, 0]
    g = {"output.weight": (d[:, None] * hs[-1]).sum(0)[None, :, None], "output.bias": d.sum()[None]}
    dz = _bf(d[:, None] * wout[None, :] * (hs[-1] > 0))
    for i in range(nl - 1, -1, -1):
        g[f"layers.{i}.weight"] = (dz.t() @ hs[i])[:, :, None]
        g[f"layers.{i}.bias"] = dz.sum(0)
        if i > 0:
            dz = _bf((dz @ _bf(sd[f"layers.{i}.weight"][:, :, 0])) * (hs[i] > 0))
    return g


def _autocast_grads(sd, feat, dq, nl, H, dev):
    """what eager PyTorch gives on the same GPU with bf16 autocast (the reference's MLPModel under torch.autocast):
    the yardstick for how far a bf16 tensor-core backward may be from fp32 autograd"""
    x = torch.from_numpy(feat).to(dev)
    params = {k: v.clone().to(dev).requires_grad_(True) for k, v in sd.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        h = x
        for i in range(nl):
            h = torch.relu(torch.nn.functional.conv1d(h, params[f"layers.{i}.weight"], params[f"layers.{i}.bias"]))
        q = torch.nn.functional.conv1d(h, params["output.weight"], params["output.bias"]).view(x.shape[0], -1)
    (q.float() * torch.from_numpy(dq).to(dev)).sum().backward()
    return {k: v.grad.float().cpu() for k, v in params.items()}


@pytest.mark.parametrize("ps,nl,H,B", [(6, 2, 256, 64), (6, 4, 512, 200), (4, 3, 256, 8), (6, 4, 2048, 96)])
def test_wide_backward_matches_oracle(tables, dev, ps, nl, H, B):
    """flat gradient of sum(dq * q) against (a) the hand-derived bf16-emulating oracle (tight), (b) torch CPU fp32
    autograd through the reference forward, with the error of eager PyTorch under bf16 autocast on the same GPU as the
    yardstick (a bf16 tensor-core backward cannot be closer to fp32 than bf16 rounding allows; ours must not be
    materially further than torch's own).  The benchmark's shape, 4 x 2048, is one of the cases; per-tensor errors are
    written to gpurun_out/ when that directory exists (profiles/r02_wide_grad_errors.json)."""
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import ops, wide
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    rec = synth_packed(B, seed=3 * B + H, party_size=ps, kind="ragged", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=H + nl)
    fo, mo = packed_oracle.encode(rec, ps, tables)
    rng = np.random.default_rng(B)
    dq = (rng.standard_normal((B, A)) * (rng.random((B, A)) < 0.3)).astype(np.float32)  # sparse, like the chosen column
    # (b) fp32 autograd
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    (ref_port.mlp_forward(params, torch.from_numpy(fo)) * torch.from_numpy(dq)).sum().backward()
    # (a) emulated
    gem = _emulated_backward(sd, fo, nl, dq)
    prepared, _, _ = wide.prepare_wide(sd, dev)
    ws = wide.train_workspace(nl, H, B, A, dev)
    m, q = wide.train_forward(ops.packed_to_device(rec, dev), A, prepared, nl, H, ws)
    grad = wide.train_backward(prepared, nl, H, B, A, ws, torch.from_numpy(dq).to(dev))
    torch.cuda.synchronize()
    with torch.no_grad():
        q32 = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    np.testing.assert_allclose(q.cpu().numpy(), q32, rtol=0, atol=2e-2)
    flat = grad.cpu()
    off = 0
    for k, v in sd.items():  # state_dict order
        n = v.numel()
        gk = flat[off:off + n].reshape(v.shape)
        off += n
        scale = float(gem[k].abs().max()) + 1e-12
        err_em = float((gk - gem[k]).abs().max()) / scale
        # the emulation rounds the same values to bf16 but sums in a different order: an activation whose fp32 sum sits
        # on a rounding boundary flips one bf16 ulp (2^-8 relative), more often the longer the sums (K = 2048)
        if H <= 512:
            assert err_em < 5e-3, f"{k}: {err_em:.2e} of the tensor's max against the bf16-emulating oracle"
        else:  # 2048-long sums: single elements flip by an ulp of bf16 here and there; bound the tensor as a whole
            fro = float((gk - gem[k]).norm() / (gem[k].norm() + 1e-12))
            assert fro < 3e-2, f"{k}: Frobenius {fro:.2e} (max {err_em:.2e}) against the emulating oracle"
        ref = params[k].grad
        # bf16 tolerance against fp32 autograd: operands, activations and dZ are rounded to bf16 (2^-9 relative each) and
        # ReLU gates of near-zero pre-activations can flip, so the bound is on the tensor as a whole
        rel = float((gk - ref).norm() / (ref.norm() + 1e-12))
        cos = float((gk * ref).sum() / (gk.norm() * ref.norm() + 1e-20))
        print(f"{k}: emulated-oracle err {err_em:.2e}, fp32 rel {rel:.3f}, cos {cos:.5f}")
        assert rel < 0.15 and cos > 0.99, f"{k}: relative error {rel:.3f}, cosine {cos:.4f} against fp32 autograd"
    assert off == flat.numel()
    # (c) the same errors for eager PyTorch with bf16 autocast; ours must stay within 1.3x of it (+ 0.01)
    gac = _autocast_grads(sd, fo, dq, nl, H, dev)
    report, off = {}, 0
    for k, v in sd.items():
        n = v.numel()
        gk = flat[off:off + n].reshape(v.shape)
        off += n
        ref = params[k].grad
        rel = float((gk - ref).norm() / (ref.norm() + 1e-12))
        rel_ac = float((gac[k] - ref).norm() / (ref.norm() + 1e-12))
        cos = float((gk * ref).sum() / (gk.norm() * ref.norm() + 1e-20))
        report[k] = {"rel_frobenius_vs_fp32": rel, "cosine_vs_fp32": cos, "torch_autocast_bf16_rel_vs_fp32": rel_ac,
                     "max_err_vs_bf16_emulating_oracle_over_max": float((gk - gem[k]).abs().max()) /
                     (float(gem[k].abs().max()) + 1e-12)}
        assert rel <= 1.3 * rel_ac + 0.01, f"{k}: {rel:.4f} against fp32, torch bf16 autocast gives {rel_ac:.4f}"
    import json
    from pathlib import Path
    out = Path(__file__).resolve().parents[1] / "gpurun_out"
    if out.is_dir():
        f = out / "r02_wide_grad_errors.json"
        cur = json.loads(f.read_text()) if f.exists() else {}
        cur[f"{nl}x{H} A={A} B={B}"] = report
        f.write_text(json.dumps(cur, indent=1))
    # bit-reproducible: a second backward gives the same bits
    grad2 = wide.train_backward(prepared, nl, H, B, A, ws, torch.from_numpy(dq).to(dev))
    assert torch.equal(grad, grad2)


def test_agent_dispatches_wide_models(tables, dev):
    """AgentVal.act_batch with a wide MLPModel goes through the tensor-core path and returns the masked argmax"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    fe = FeatureExtractor(6, device=dev)
    torch.manual_seed(2)
    model = MLPModel(fe.input_shape, fe.output_dim, n_layers=2, n_channels=256).eval()
    rec = synth_packed(64, seed=8, party_size=6, kind="ragged", tables=tables)
    agent = AgentVal(model, fe)
    q, a = agent.q_and_action_batch(rec)
    fo, mo = packed_oracle.encode(rec, 6, tables)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        q32 = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
    np.testing.assert_allclose(q, q32, rtol=0, atol=2e-2)
    assert np.array_equal(agent.act_batch(rec), ref_port.masked_argmax(q, mo))


def test_wide_chosen_column_forward_backward(tables, dev):
    """col_select: one GEMM row per transition (the gather of trainer.py:163); q equals the chosen column of the full
    forward, the gradient equals the full backward with dq scattered to the chosen columns; odd row counts work"""
    from oracle import ref_port
    from pokeai_b200 import ops, wide
    from pokeai_b200.packed import synth_packed
    B, A, nl, H = 77, 9, 2, 256
    rec = synth_packed(B, seed=4, party_size=6, kind="full", tables=tables)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=1)
    prepared, _, _ = wide.prepare_wide(sd, dev)
    pk = ops.packed_to_device(rec, dev)
    rng = np.random.default_rng(0)
    col = torch.from_numpy(rng.integers(0, A, B).astype(np.int32)).to(dev)
    dqc = torch.from_numpy(rng.standard_normal(B).astype(np.float32)).to(dev)
    ws = wide.train_workspace(nl, H, B, A, dev)
    _, q_full = wide.train_forward(pk, A, prepared, nl, H, ws)
    dq_full = torch.zeros((B, A), dtype=torch.float32, device=dev)
    dq_full[torch.arange(B, device=dev), col.long()] = dqc
    g_full = wide.train_backward(prepared, nl, H, B, A, ws, dq_full).clone()
    _, q_sel = wide.train_forward(pk, A, prepared, nl, H, ws, col_select=col)
    g_sel = wide.train_backward(prepared, nl, H, B, A, ws, dqc, chosen_only=True)
    torch.cuda.synchronize()
    assert torch.equal(q_sel, q_full[torch.arange(B, device=dev), col.long()])
    scale = float(g_full.abs().max())
    assert float((g_sel - g_full).abs().max()) <= 2e-3 * scale  # same terms, different summation grouping


def test_wide_trainer_update_matches_oracle(tables, dev):
    """Trainer._update for a wide network (3 forwards + chosen-column backward + clamp + Adam on the tensor-core path)
    against the oracle port of the reference update (torch CPU fp32 autograd + Adam): loss within the bf16 tolerance,
    parameters after the step within lr-scaled tolerance, and the step is deterministic"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    n, ps, A, nl, H = 96, 6, 9, 2, 256
    torch.manual_seed(0)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": False}, {"batch_size": n, "lr": 1e-3}, {"party_size": ps},
                 device=dev)
    assert tr._wide
    rng = np.random.default_rng(1)
    s = synth_packed(n, seed=21, party_size=ps, kind="full", tables=tables)
    s2 = synth_packed(n, seed=22, party_size=ps, kind="ragged", tables=tables)
    a = rng.integers(0, A, n).astype(np.int32)
    r = rng.uniform(-1, 1, n).astype(np.float32)
    t = (rng.random(n) < 0.1).astype(np.uint8)
    sd0 = {k: v.detach().cpu().clone() for k, v in tr.model.state_dict().items()}
    tsd0 = {k: v.detach().cpu().clone() for k, v in tr.target_model.state_dict().items()}
    f, _ = packed_oracle.encode(s, ps, tables)
    f2, m2 = packed_oracle.encode(s2, ps, tables)
    loss_o, grads_o = ref_port.dqn_update_grads(sd0, tsd0, f, a.astype(np.int64), r, f2, m2, t == 0, tr.gamma, True)
    loss = tr._update_from_arrays(s, s2, a, r, t)
    assert abs(loss - loss_o) < 2e-2, (loss, loss_o)
    # one Adam step from zero moments moves every parameter by ~lr * sign(grad): compare the direction where the
    # oracle gradient is clearly non-zero
    sd1 = {k: v.detach().cpu() for k, v in tr.model.state_dict().items()}
    for k in sd0:
        step = sd1[k] - sd0[k]
        g = grads_o[k]
        big = g.abs() > 0.05 * g.abs().max()
        agree = float((torch.sign(step[big]) == -torch.sign(g[big])).float().mean())
        assert agree > 0.97, f"{k}: update direction agrees with the oracle on only {agree:.3f} of the large gradients"
        assert float(step.abs().max()) <= 1.01e-3
    # the prepared bf16 weights follow the parameters: a fresh agent sees the updated network
    q_new, _ = tr.get_val_agent().q_and_action_batch(s[:8])
    with torch.no_grad():
        q_ref = ref_port.mlp_forward(sd1, torch.from_numpy(f[:8])).numpy()
    np.testing.assert_allclose(q_new, q_ref, rtol=0, atol=2e-2)


def test_wide_overlapped_exchange_two_gpus():
    """per-layer NCCL buckets started from inside the backward (pk_wide_train_backward_ev) against one allreduce after
    it, one process per GPU: tools/check_wide_overlap.py under torchrun.  Needs two GPUs; skipped on a one-GPU box."""
    import subprocess
    import sys
    from pathlib import Path
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = Path(__file__).resolve().parents[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29583", str(root / "tools" / "check_wide_overlap.py")],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "wide overlap ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    # the same run also checks the peer-memory route (pk_peer_allreduce) where peer mapping is available
    assert "wide peer allreduce ok" in r.stdout or "wide peer allreduce skipped" in r.stdout, r.stdout[-2000:]
    assert "wide trainer peer route ok" in r.stdout or "wide peer allreduce skipped" in r.stdout, r.stdout[-2000:]
