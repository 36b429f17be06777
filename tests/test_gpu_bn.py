"""GPU parity of the train-mode BatchNorm update (bn: true; pk_bn_* entry points, pokeai_b200/bn_train.py) against
golden vectors of the REAL reference's Trainer._update (tools/make_golden_bn.py -> tests/golden/dqn_update_bn.npz) and
against a plain torch fp32 statement of the same network on the CPU.

Tolerances (fp32; our reductions run in double and in another order than torch-CPU's): loss 1e-6 abs; pre-clamp
gradients rtol 1e-4 + atol 2e-7; running statistics rtol 1e-5 + atol 1e-6; parameters after Adam rtol 1e-5 + atol 2e-6
wherever the reference's gradient is not below 1e-6 in magnitude (Adam's first steps move a parameter by
lr * g / (|g| + 1e-8): for a gradient that is zero in exact arithmetic and 1e-9 of rounding noise in floating point the
step is noise in ANY implementation; those elements must only stay within 2 * lr per step)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
G_RTOL, G_ATOL = 1e-4, 2e-7
P_RTOL, P_ATOL = 1e-5, 2e-6
S_RTOL, S_ATOL = 1e-5, 1e-6


def _sd(z, prefix):
    return {k[len(prefix):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix)}


def _records(z, key):
    from pokeai_b200.packed import PK_OBS_DTYPE
    return np.frombuffer(z[key].tobytes(), dtype=PK_OBS_DTYPE)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_bn_update_matches_reference_golden(tag):
    """Two consecutive updates of the real reference Trainer with bn: true (a: 2 x 16 Double DQN, b: 3 x 24 plain DQN)."""
    from pokeai_b200.trainer import Trainer
    z = np.load(GOLDEN / "dqn_update_bn.npz")
    p = tag + "_"
    ps, nl, H = int(z[p + "party_size"]), int(z[p + "n_layers"]), int(z[p + "n_channels"])
    n = len(z[p + "action"])
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": True},
                 {"batch_size": n, "gamma": float(z[p + "gamma"]), "lr": float(z[p + "lr"]), "replay_buffer_size": n,
                  "double_dqn": bool(z[p + "double_dqn"])}, {"party_size": ps}, device=DEV)
    tr.model.load_state_dict(_sd(z, p + "sd0_"))
    tr.target_model.load_state_dict(_sd(z, p + "tgt0_"))
    rec_s, rec_n = _records(z, p + "packed_s"), _records(z, p + "packed_n")
    action, reward, terminal = z[p + "action"], z[p + "reward"], z[p + "terminal"]
    names = [k for k, _ in tr.model.named_parameters()]
    shaky = {k: np.zeros(z[f"{p}grad1_{k}"].shape, dtype=bool) for k in names}
    for step in (1, 2):
        loss = tr._update_from_arrays(rec_s, rec_n, action, reward, terminal)
        assert abs(loss - float(z[f"{p}loss{step}"])) < 1e-6
        off = 0
        for k, prm in tr.model.named_parameters():
            g = tr._grad[off:off + prm.numel()].view(prm.shape).cpu().numpy()
            off += prm.numel()
            want = z[f"{p}grad{step}_{k}"]
            np.testing.assert_allclose(g, want, rtol=G_RTOL, atol=G_ATOL, err_msg=f"grad{step} {k}")
            if k == "layers.0.weight":  # weights of features that never occur: exactly zero, so Adam leaves them alone
                assert np.all(g[want == 0] == 0)
            shaky[k] |= np.abs(want) < 1e-6
        sd = tr.model.state_dict()
        for k in names:
            got, want = sd[k].cpu().numpy(), z[f"{p}sd{step}_{k}"]
            ok = ~shaky[k] | (z[f"{p}grad1_{k}"] == 0)
            np.testing.assert_allclose(got[ok], want[ok], rtol=P_RTOL, atol=P_ATOL, err_msg=f"sd{step} {k}")
            assert np.abs(got - want).max() <= 2 * step * float(z[p + "lr"]) + 1e-6, k
        tsd = tr.target_model.state_dict()
        for l in range(nl):
            for which in ("running_mean", "running_var"):
                k = f"bn_layers.{l}.{which}"
                np.testing.assert_allclose(sd[k].cpu().numpy(), z[f"{p}sd{step}_{k}"], rtol=S_RTOL, atol=S_ATOL, err_msg=k)
                np.testing.assert_allclose(tsd[k].cpu().numpy(), z[f"{p}tgt{step}_{k}"], rtol=S_RTOL, atol=S_ATOL,
                                           err_msg="target " + k)
            k = f"bn_layers.{l}.num_batches_tracked"
            assert int(sd[k]) == int(z[f"{p}sd{step}_{k}"]) and int(tsd[k]) == int(z[f"{p}tgt{step}_{k}"])
    # a snapshot agent's network is in eval mode: BatchNorm folded with the running statistics of this moment
    from pokeai_b200 import ops
    model = tr.get_val_agent()._model
    assert not model.training
    feat, _ = ops.encode(ops.packed_to_device(rec_s, torch.device(DEV)), 4 + ps - 1)
    q = model(feat).cpu().numpy()
    np.testing.assert_allclose(q, z[p + "q_eval"], rtol=1e-4, atol=2e-4)  # parameters differ by <= P_ATOL-ish after 2 steps


def _torch_forward(model, x):
    """the network of mlp_model.py:21-30 from the module's own torch layers (CPU, fp32)"""
    h = x
    for conv, bn in zip(model.layers, model.bn_layers):
        h = F.relu(bn(conv(h)))
    return model.output(h).view(x.shape[0], -1)


@pytest.mark.parametrize("nl,H,B,ps,seed", [(2, 64, 700, 6, 3), (1, 40, 33, 3, 3), (3, 100, 257, 6, 4)])
def test_bn_forward_and_backward_match_torch(tables, nl, H, B, ps, seed):
    """Training-mode forward (batch statistics, running statistics, num_batches_tracked) and the gradient of a random
    linear functional of Q against torch in FLOAT64 on the CPU, at sizes that exercise several GEMM tiles, ragged edges
    and the split-K weight gradient.  torch's own fp32 path differs from its fp64 path by ~6e-7 (q) and ~2e-7 (gradient)
    on these cases; ours must stay within 5e-6 of the fp64 values, relative to the largest.  The seeds are chosen so that
    no pre-activation lies within 2e-6 of zero (checked below): a ReLU input of 7e-8, as seed 1 gives the
    3 x 100 case, is below fp32 rounding noise and flips the unit on or off in any fp32 implementation."""
    import copy
    from pokeai_b200 import bn_train, ops
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    A = 4 + ps - 1
    torch.manual_seed(seed)
    model = MLPModel((tables.n_channels, A), A, n_layers=nl, n_channels=H, bn=True)
    with torch.no_grad():
        for b in model.bn_layers:
            b.weight.uniform_(0.5, 1.5)
            b.bias.normal_(0, 0.1)
            b.running_mean.normal_(0, 0.2)
            b.running_var.uniform_(0.5, 1.5)
    ref = copy.deepcopy(model).double()
    dev = torch.device(DEV)
    packed = ops.packed_to_device(synth_packed(B, seed=3, party_size=ps, kind="ragged", tables=tables), dev)
    feat, mask = ops.encode(packed, A)
    x = feat.cpu().double()
    for p_ in ref.parameters():
        p_.requires_grad_(True)
    ref.train()
    with torch.no_grad():  # the fixture's precondition (on a throw-away copy: a training forward moves the buffers)
        h = x
        probe = copy.deepcopy(ref)
        for conv, bn in zip(probe.layers, probe.bn_layers):
            pre = bn(conv(h))
            assert float(pre.abs().min()) > 2e-6, "ill-conditioned fixture: a ReLU input within fp32 noise of zero"
            h = F.relu(pre)
    q_ref = _torch_forward(ref, x)
    # product: the module's forward in training mode, on the GPU
    model.train()
    q = model(feat)
    scale = max(float(q_ref.detach().abs().max()), 1.0)
    assert float((q.cpu().double() - q_ref.detach()).abs().max()) <= 5e-6 * scale
    for bm, br in zip(model.bn_layers, ref.bn_layers):
        np.testing.assert_allclose(bm.running_mean.cpu().numpy(), br.running_mean.numpy(), rtol=S_RTOL, atol=S_ATOL)
        np.testing.assert_allclose(bm.running_var.cpu().numpy(), br.running_var.numpy(), rtol=S_RTOL, atol=S_ATOL)
        assert int(bm.num_batches_tracked) == int(br.num_batches_tracked) == 1
    # gradient of sum(q * dq) for a dense random dq
    g = torch.Generator().manual_seed(1)
    dq = torch.randn(B, A, generator=g) / B
    (q_ref * dq.double()).sum().backward()
    want = torch.cat([p_.grad.reshape(-1) for p_ in ref.parameters()])
    flat = bn_train.flat_parameters(model, dev)  # the parameters the forward above used (the forward moves buffers only)
    running = bn_train.stacked_running(model, dev)
    q2, ws = bn_train.forward_train(flat, running, nl, H, feat)
    assert float((q2 - q).abs().max()) <= 1e-6 * scale  # same kernels; the double atomics may round a mean differently
    grad = torch.empty_like(flat)
    bn_train.backward(flat, nl, H, feat, ws, dq.to(dev).contiguous(), grad)
    gs = float(want.abs().max())
    assert float((grad.cpu().double() - want).abs().max()) <= 5e-6 * gs
    # the same straight from the packed records (what Trainer uses): layer 0 gathered, dW0's one-hot rows by atomics
    qp, maskp, wsp = bn_train.forward_train_packed(flat, None, nl, H, packed, A)
    assert torch.equal(maskp, mask)
    assert float((qp.cpu().double() - q_ref.detach()).abs().max()) <= 5e-6 * scale
    gradp = torch.full_like(flat, float("nan"))
    bn_train.backward_packed(flat, nl, H, packed, A, wsp, dq.to(dev).contiguous(), gradp)
    assert float((gradp.cpu().double() - want).abs().max()) <= 5e-6 * gs
    unused = (feat.abs().amax(dim=(0, 2)) == 0).cpu()  # feature rows that never occur in this batch
    gw0 = gradp[:H * tables.n_channels].view(H, tables.n_channels).cpu()
    assert bool((gw0[:, unused] == 0).all())  # exactly zero: Adam leaves those weights alone
    assert B > 100 or int(unused.sum()) > 100  # the small batch leaves most of the 617 rows unused
    # in eval mode the same module folds BatchNorm (running statistics) and runs the inference kernel
    model.eval()
    ref.eval()
    with torch.no_grad():  # give the torch copy the module's running statistics, then compare the eval-mode outputs
        for bm, br in zip(model.bn_layers, ref.bn_layers):
            br.running_mean.copy_(bm.running_mean.cpu())
            br.running_var.copy_(bm.running_var.cpu())
        q_eval_ref = _torch_forward(ref, x)
    assert float((model(feat).cpu().double() - q_eval_ref).abs().max()) <= 1e-5 * max(float(q_eval_ref.abs().max()), 1.0)


def test_clamp_adam_flat_matches_torch_adam():
    from pokeai_b200 import bn_train
    torch.manual_seed(0)
    n = 5000
    p0 = torch.randn(n)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1e-3)
    p = p0.clone().to(DEV)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in range(1, 4):
        g = torch.randn(n) * 2
        ref.grad = g.clamp(-1, 1)
        opt.step()
        bn_train.clamp_adam_flat(p, g.to(DEV), m, v, step)
        np.testing.assert_allclose(p.cpu().numpy(), ref.detach().numpy(), rtol=1e-6, atol=1e-7)


def test_bn_trainer_schedule_and_checkpoint(tables):
    """train() with target updates in between, then a save_state / load_state round trip that carries the BatchNorm
    buffers of both networks: the restored trainer takes the same next step."""
    import random
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    ps, A, n = 3, 6, 64
    torch.manual_seed(3)
    tr = Trainer({"n_layers": 2, "n_channels": 32, "bn": True},
                 {"batch_size": 16, "first_update_steps": 0, "target_update": 2, "replay_buffer_size": 1000},
                 {"party_size": ps}, device=DEV)
    rs = synth_packed(n, seed=1, party_size=ps, kind="ragged", tables=tables)
    rn = synth_packed(n, seed=2, party_size=ps, kind="ragged", tables=tables)
    rng = np.random.default_rng(0)
    action = (rng.random(n) * rs["n_act"]).astype(np.int32)
    reward = rng.uniform(-1, 1, n).astype(np.float32)
    terminal = (np.arange(n) % 7 == 0)
    random.seed(0)
    tr.replay_buffer.extend_arrays(rs, rn, action, reward, terminal)
    tr.total_steps = n
    tr.update_steps = n - 4
    tr.train()  # 4 updates, target updates in between
    assert len(tr.update_loss_history) == 4 and all(np.isfinite(tr.update_loss_history))
    sd = tr.model.state_dict()
    assert int(sd["bn_layers.0.num_batches_tracked"]) == 8  # two training forwards of the online net per update
    # agents get an eval-mode snapshot (BatchNorm folded with the running statistics): the decision kernel's Q equals the
    # snapshot module's own forward on the encoded features
    from pokeai_b200 import ops
    agent = tr.get_train_agent()
    q_agent, _ = agent.q_and_action_batch(rs[:16])
    feat, _ = ops.encode(ops.packed_to_device(rs[:16], torch.device(DEV)), A)
    assert not agent._model.training
    np.testing.assert_allclose(np.array(q_agent), agent._model(feat).cpu().numpy(), rtol=1e-5, atol=1e-5)
    state = tr.save_state(resume=True)
    tr2 = Trainer.load_state(state, resume=True, device=DEV)
    for k, v in tr.model.state_dict().items():
        assert torch.equal(v.cpu(), tr2.model.state_dict()[k].cpu()), k
    for k, v in tr.target_model.state_dict().items():
        assert torch.equal(v.cpu(), tr2.target_model.state_dict()[k].cpu()), k
    random.seed(1)
    st = random.getstate()
    tr._update()
    random.setstate(st)
    tr2._update()
    assert tr.update_loss_history[-1] == pytest.approx(tr2.update_loss_history[-1], abs=1e-7)


@pytest.mark.parametrize("n,term,double", [(1, [0], True), (2, [1, 0], True), (3, [0, 1, 0], False), (5, [0, 0, 0, 0, 1], True)])
def test_bn_update_tiny_batches_match_torch(tables, n, term, double):
    """The whole update at batch sizes below every tile size (one transition = six columns for the statistics; a single
    non-final next state) against the update of trainer.py:152-195 written out with torch autograd in float64."""
    import copy
    from pokeai_b200 import ops
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    ps, A, nl, H, gamma = 3, 6, 2, 16, 0.9
    torch.manual_seed(40 + n)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": True}, {"batch_size": n, "gamma": gamma, "double_dqn": double},
                 {"party_size": ps}, device=DEV)
    proto = MLPModel((tables.n_channels, A), A, n_layers=nl, n_channels=H, bn=True)
    tproto = MLPModel((tables.n_channels, A), A, n_layers=nl, n_channels=H, bn=True)
    with torch.no_grad():
        for m_ in (proto, tproto):
            for b in m_.bn_layers:
                b.weight.uniform_(0.5, 1.5)
                b.bias.normal_(0, 0.1)
                b.running_mean.normal_(0, 0.2)
                b.running_var.uniform_(0.5, 1.5)
    tr.model.load_state_dict(proto.state_dict())
    tr.target_model.load_state_dict(tproto.state_dict())
    ref, tgt = copy.deepcopy(proto).double(), copy.deepcopy(tproto).double()
    for p_ in ref.parameters():
        p_.requires_grad_(True)
    rs = synth_packed(n, seed=70 + n, party_size=ps, kind="ragged", tables=tables)
    rn = synth_packed(n, seed=80 + n, party_size=ps, kind="ragged", tables=tables)
    rng = np.random.default_rng(n)
    action = (rng.random(n) * rs["n_act"]).astype(np.int64)
    reward = rng.uniform(-2, 2, n).astype(np.float32)
    terminal = np.array(term, dtype=bool)
    dev = torch.device(DEV)
    fs = ops.encode(ops.packed_to_device(rs, dev), A)[0].cpu().double()
    fn, mn = ops.encode(ops.packed_to_device(rn, dev), A)
    fn, mn = fn.cpu().double(), mn.cpu().numpy()
    nf = ~terminal
    # trainer.py:163-187 (every forward in training mode, the next-state forwards on the non-final subset)
    q_sa = _torch_forward(ref, fs).gather(1, torch.from_numpy(action)[:, None])
    bias = torch.from_numpy(np.where(mn[nf] != 0, 0.0, -np.inf))
    nsv = torch.zeros(n, dtype=torch.float64)
    with torch.no_grad():
        if double:
            a_star = (_torch_forward(ref, fn[nf]) + bias).max(1)[1]
            nsv[torch.from_numpy(nf)] = _torch_forward(tgt, fn[nf]).gather(1, a_star[:, None]).squeeze(1)
        else:
            nsv[torch.from_numpy(nf)] = (_torch_forward(tgt, fn[nf]) + bias).max(1)[0]
    expected = nsv * gamma + torch.from_numpy(reward).double()
    loss_ref = F.smooth_l1_loss(q_sa, expected[:, None])
    loss_ref.backward()
    want = torch.cat([p_.grad.reshape(-1) for p_ in ref.parameters()])
    loss = tr._update_from_arrays(rs, rn, action, reward, terminal)
    assert abs(loss - float(loss_ref.detach())) < 2e-6
    gs = float(want.abs().max())
    assert float((tr._grad.cpu().double() - want).abs().max()) <= 2e-5 * gs
    for bm, br, bt, btr in zip(tr.model.bn_layers, ref.bn_layers, tr.target_model.bn_layers, tgt.bn_layers):
        np.testing.assert_allclose(bm.running_mean.cpu().numpy(), br.running_mean.numpy(), rtol=S_RTOL, atol=S_ATOL)
        np.testing.assert_allclose(bm.running_var.cpu().numpy(), br.running_var.numpy(), rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(bt.running_var.cpu().numpy(), btr.running_var.numpy(), rtol=1e-4, atol=1e-5)
        assert int(bm.num_batches_tracked) == int(br.num_batches_tracked) and int(bt.num_batches_tracked) == 1
    with pytest.raises(ValueError):  # all transitions terminal: np.stack([]) fails in the reference too (trainer.py:157)
        tr._update_from_arrays(rs, rn, action, reward, np.ones(n, dtype=bool))


@pytest.mark.parametrize("n,ps,nl,H,double", [(300, 6, 2, 64, True), (129, 2, 3, 32, False)])
def test_bn_update_matches_oracle(tables, n, ps, nl, H, double):
    """Trainer's bn update against the oracle's written-out restatement (oracle/ref_port.py::dqn_update_grads_bn, itself
    pinned to the reference's golden vectors by tests/test_oracle.py) evaluated in float64, on seeded synthetic transitions:
    loss, pre-clamp gradient, running statistics and counters of both networks."""
    from oracle import packed_oracle, ref_port
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    A = 4 + ps - 1
    torch.manual_seed(1000 + n)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": True}, {"batch_size": n, "gamma": 0.95, "double_dqn": double},
                 {"party_size": ps}, device=DEV)
    protos = [MLPModel((tables.n_channels, A), A, n_layers=nl, n_channels=H, bn=True) for _ in range(2)]
    with torch.no_grad():
        for m_ in protos:
            for b in m_.bn_layers:
                b.weight.uniform_(0.5, 1.5)
                b.bias.normal_(0, 0.1)
                b.running_mean.normal_(0, 0.2)
                b.running_var.uniform_(0.5, 1.5)
    tr.model.load_state_dict(protos[0].state_dict())
    tr.target_model.load_state_dict(protos[1].state_dict())
    sd = {k: v.detach().clone().double() if v.is_floating_point() else v.clone() for k, v in protos[0].state_dict().items()}
    tgt = {k: v.detach().clone().double() if v.is_floating_point() else v.clone() for k, v in protos[1].state_dict().items()}
    rs = synth_packed(n, seed=n, party_size=ps, kind="ragged", tables=tables)
    rn = synth_packed(n, seed=n + 1, party_size=ps, kind="ragged", tables=tables)
    rng = np.random.default_rng(n)
    action = (rng.random(n) * rs["n_act"]).astype(np.int64)
    reward = rng.uniform(-2, 2, n).astype(np.float32)
    terminal = rng.random(n) < 0.1
    fs, _ = packed_oracle.encode(rs, ps, tables)
    fn, mn = packed_oracle.encode(rn, ps, tables)
    loss_o, grads_o = ref_port.dqn_update_grads_bn(sd, tgt, fs, action, reward, fn, mn, ~terminal, 0.95, double_dqn=double,
                                                   dtype=torch.float64)
    loss = tr._update_from_arrays(rs, rn, action, reward, terminal)
    assert abs(loss - loss_o) < 2e-6
    want = torch.cat([grads_o[k].reshape(-1) for k, _ in tr.model.named_parameters()])
    assert float((tr._grad.cpu().double() - want).abs().max()) <= 2e-5 * float(want.abs().max())
    for which, mine, theirs in (("online", tr.model.state_dict(), sd), ("target", tr.target_model.state_dict(), tgt)):
        for k, v in theirs.items():
            if k.endswith("num_batches_tracked"):
                assert int(mine[k]) == int(v), (which, k)
            elif k.endswith(("running_mean", "running_var")):
                np.testing.assert_allclose(mine[k].cpu().numpy(), v.numpy(), rtol=1e-5, atol=1e-6, err_msg=f"{which} {k}")
