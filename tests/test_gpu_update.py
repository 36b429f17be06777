"""GPU parity of the DQN update path (pk_dqn_grad / pk_clamp_adam / Trainer) against the golden vectors of the
real reference's Trainer._update and against the oracle port.  Tolerances (fp32, different summation order
than torch-CPU autograd): loss 1e-6 abs; pre-clamp gradients rtol 1e-4 + atol 1e-7; parameters after Adam
rtol 1e-5 + atol 1e-6."""
import random

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

G_RTOL, G_ATOL = 1e-4, 1e-7
P_RTOL, P_ATOL = 1e-5, 1e-6
DEV = "cuda:0"


def _sd(z, prefix):
    return {k[len(prefix):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix)}


def _dev_batch(rec_s, rec_n, action, reward, terminal):
    from pokeai_b200 import ops
    dev = torch.device(DEV)
    return (ops.packed_to_device(rec_s, dev), ops.packed_to_device(rec_n, dev),
            torch.from_numpy(np.asarray(action, dtype=np.int32)).to(dev),
            torch.from_numpy(np.asarray(reward, dtype=np.float32)).to(dev),
            torch.from_numpy(np.asarray(terminal, dtype=np.uint8)).to(dev))


def test_update_matches_reference_golden(tables):
    """Two consecutive Trainer._update steps of the REAL reference (tools/make_golden.py) on a fixed buffer."""
    from pokeai_b200 import ops
    from pokeai_b200.packed import PK_OBS_DTYPE
    z = np.load(GOLDEN / "dqn_update.npz")
    ps, nl, H = int(z["party_size"]), int(z["n_layers"]), int(z["n_channels"])
    A = 4 + ps - 1
    rec_s = np.frombuffer(z["packed_s"].tobytes(), dtype=PK_OBS_DTYPE)
    rec_n = np.frombuffer(z["packed_n"].tobytes(), dtype=PK_OBS_DTYPE)
    batch = _dev_batch(rec_s, rec_n, z["action"], z["reward"], z["terminal"])
    prepared, nl2, H2, flat = ops.prepare_weights(_sd(z, "sd0_"), DEV)
    prepared_t = ops.prepare_weights(_sd(z, "tgt_"), DEV)[0]
    assert (nl2, H2) == (nl, H)
    m, v = torch.zeros_like(flat), torch.zeros_like(flat)
    for step in (1, 2):
        grad, loss = ops.dqn_grad(prepared, prepared_t, nl, H, A, *batch, gamma=float(z["gamma"]))
        assert abs(float(loss.item()) - float(z[f"loss{step}"])) < 1e-6
        for k, g in ops.unflatten(grad, nl, H, tables.n_channels).items():
            np.testing.assert_allclose(g.cpu().numpy(), z[f"grad{step}_{k}"], rtol=G_RTOL, atol=G_ATOL, err_msg=k)
        ops.clamp_adam(flat, grad, m, v, nl, H, step, lr=float(z["lr"]), prepared=prepared)
        for k, p in ops.unflatten(flat, nl, H, tables.n_channels).items():
            np.testing.assert_allclose(p.cpu().numpy(), z[f"sd{step}_{k}"], rtol=P_RTOL, atol=P_ATOL, err_msg=k)
        # the fused refresh of the prepared weights == preparing the new parameters from scratch
        fresh = ops.prepare_from_flat(flat, nl, H, tables.n_channels)
        assert torch.equal(fresh, prepared)


def _synth_transitions(n, ps, tables, seed, p_terminal=0.05):
    from pokeai_b200.packed import synth_packed
    rng = np.random.default_rng(seed)
    rec_s = synth_packed(n, seed=seed, party_size=ps, kind="ragged", tables=tables)
    rec_n = synth_packed(n, seed=seed + 1, party_size=ps, kind="ragged", tables=tables)
    action = (rng.random(n) * rec_s["n_act"]).astype(np.int32)
    reward = rng.uniform(-1, 1, n).astype(np.float32)
    reward[rng.random(n) < 0.1] *= 8.0  # some |delta| > 1: the linear branch of the Huber loss
    terminal = rng.random(n) < p_terminal
    return rec_s, rec_n, action, reward, terminal


@pytest.mark.parametrize("ps,nl,H,n,double", [(6, 2, 64, 777, True), (6, 2, 64, 300, False), (3, 3, 64, 500, True),
                                              (6, 3, 16, 1000, True), (2, 1, 32, 129, True), (6, 4, 64, 400, True),
                                              (6, 2, 64, 1, True), (6, 2, 64, 4096, True)])
def test_update_matches_oracle(tables, ps, nl, H, n, double):
    from oracle import packed_oracle, ref_port
    from pokeai_b200 import ops
    A = 4 + ps - 1
    rec_s, rec_n, action, reward, terminal = _synth_transitions(n, ps, tables, seed=100 + n)
    if n == 1:
        terminal[:] = False
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=3)
    tgt = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=4)
    fs, _ = packed_oracle.encode(rec_s, ps, tables)
    fn, mn = packed_oracle.encode(rec_n, ps, tables)
    loss_o, grads_o = ref_port.dqn_update_grads(sd, tgt, fs, action, reward, fn, mn, ~terminal, 0.95, double_dqn=double)
    prepared, _, _, flat = ops.prepare_weights(sd, DEV)
    prepared_t = ops.prepare_weights(tgt, DEV)[0]
    batch = _dev_batch(rec_s, rec_n, action, reward, terminal)
    grad, loss = ops.dqn_grad(prepared, prepared_t, nl, H, A, *batch, gamma=0.95, double_dqn=double)
    assert abs(float(loss.item()) - loss_o) < 2e-6 * max(1.0, abs(loss_o))
    scale = max(float(g.abs().max()) for g in grads_o.values())
    for k, g in ops.unflatten(grad, nl, H, tables.n_channels).items():
        np.testing.assert_allclose(g.cpu().numpy(), grads_o[k].numpy(), rtol=G_RTOL, atol=max(G_ATOL, 1e-6 * scale),
                                   err_msg=k)
    # Adam step against the oracle's restatement of clamp + torch.optim.Adam, both fed the ORACLE's gradient
    # (at step 1 Adam moves every weight by ~lr * sign(g), so feeding each side its own gradient would turn a
    # 1e-9 difference in a near-zero gradient into a 1e-3 * lr difference in the weight)
    keys = list(ops.unflatten(flat, nl, H, tables.n_channels).keys())
    g_flat = torch.cat([grads_o[k].reshape(-1) for k in keys]).to(DEV)
    m, v = torch.zeros_like(flat), torch.zeros_like(flat)
    sd_o = {k: t.clone() for k, t in sd.items()}
    adam_o = {}
    for step in (1, 2, 3):
        ops.clamp_adam(flat, g_flat, m, v, nl, H, step, lr=1e-3, prepared=prepared)
        ref_port.clamp_adam_step(sd_o, grads_o, adam_o, 1e-3, step)
    for k, p in ops.unflatten(flat, nl, H, tables.n_channels).items():
        np.testing.assert_allclose(p.cpu().numpy(), sd_o[k].numpy(), rtol=P_RTOL, atol=P_ATOL, err_msg=k)
    assert torch.equal(ops.prepare_from_flat(flat, nl, H, tables.n_channels), prepared)


def test_update_is_deterministic_and_shards_add_up(tables):
    """(1) bit-identical gradients run to run (fixed summation order); (2) the data-parallel invariant: the
    gradients of two shards computed with n_global = n sum to the full-batch gradient."""
    from oracle import ref_port
    from pokeai_b200 import ops
    n, ps, nl, H, A = 2048, 6, 2, 64, 9
    rec_s, rec_n, action, reward, terminal = _synth_transitions(n, ps, tables, seed=7)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=8)
    prepared = ops.prepare_weights(sd, DEV)[0]
    prepared_t = ops.prepare_weights(ref_port.init_mlp_params(tables.n_channels, nl, H, seed=9), DEV)[0]
    full = _dev_batch(rec_s, rec_n, action, reward, terminal)
    g1, l1 = ops.dqn_grad(prepared, prepared_t, nl, H, A, *full, gamma=0.95)
    g2, l2 = ops.dqn_grad(prepared, prepared_t, nl, H, A, *full, gamma=0.95)
    assert torch.equal(g1, g2) and torch.equal(l1, l2)
    cut = 900
    ga, la = ops.dqn_grad(prepared, prepared_t, nl, H, A, *[t[:cut].contiguous() for t in full], gamma=0.95, n_global=n)
    gb, lb = ops.dqn_grad(prepared, prepared_t, nl, H, A, *[t[cut:].contiguous() for t in full], gamma=0.95, n_global=n)
    np.testing.assert_allclose((ga + gb).cpu().numpy(), g1.cpu().numpy(), rtol=1e-4, atol=1e-7)
    assert abs(float((la + lb).item()) - float(l1.item())) < 1e-6


def test_update_full_size_properties(tables):
    """32768 transitions (BASELINE configs[3]): bit-identical run to run; eight contiguous shards with n_global =
    32768 sum to the full-batch gradient and loss (the data-parallel invariant at the benchmark's size); 256 sampled
    transitions agree with the oracle's per-sample gradient sum through linearity (their shard alone, n_global kept)"""
    from oracle import ref_port
    from pokeai_b200 import ops
    n, ps, nl, H, A = 32768, 6, 2, 64, 9
    rec_s, rec_n, action, reward, terminal = _synth_transitions(n, ps, tables, seed=21)
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=22)
    sd_t = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=23)
    prepared = ops.prepare_weights(sd, DEV)[0]
    prepared_t = ops.prepare_weights(sd_t, DEV)[0]
    full = _dev_batch(rec_s, rec_n, action, reward, terminal)
    g1, l1 = ops.dqn_grad(prepared, prepared_t, nl, H, A, *full, gamma=0.95)
    g2, l2 = ops.dqn_grad(prepared, prepared_t, nl, H, A, *full, gamma=0.95)
    assert torch.equal(g1, g2) and torch.equal(l1, l2)
    assert bool(torch.isfinite(g1).all()) and float(g1.abs().max()) > 0
    gs, ls = torch.zeros_like(g1), 0.0
    for r in range(8):
        lo, hi = n * r // 8, n * (r + 1) // 8
        g, l = ops.dqn_grad(prepared, prepared_t, nl, H, A, *[t[lo:hi].contiguous() for t in full], gamma=0.95, n_global=n)
        gs += g
        ls += float(l.item())
    np.testing.assert_allclose(gs.cpu().numpy(), g1.cpu().numpy(), rtol=2e-4, atol=2e-7)
    assert abs(ls - float(l1.item())) < 1e-5
    # a 256-transition slice against the oracle, scaled by its share of the global batch
    lo, hi = 5000, 5256
    g, l = ops.dqn_grad(prepared, prepared_t, nl, H, A, *[t[lo:hi].contiguous() for t in full], gamma=0.95, n_global=n)
    from oracle import packed_oracle
    fs, _ = packed_oracle.encode(rec_s[lo:hi], ps, tables)
    fn, mn = packed_oracle.encode(rec_n[lo:hi], ps, tables)
    loss_o, grads_o = ref_port.dqn_update_grads(sd, sd_t, fs, action[lo:hi], reward[lo:hi], fn, mn, ~terminal[lo:hi], 0.95,
                                                double_dqn=True)
    scale = (hi - lo) / n  # the oracle averages over its 256 transitions, the kernel divides by n_global
    gmax = max(float(t.abs().max()) for t in grads_o.values()) * scale
    for k, t in ops.unflatten(g, nl, H, tables.n_channels).items():
        np.testing.assert_allclose(t.cpu().numpy(), grads_o[k].numpy() * scale, rtol=G_RTOL, atol=max(1e-9, 1e-6 * gmax),
                                   err_msg=k)
    assert abs(float(l.item()) - loss_o * scale) < 1e-7


def test_update_bad_arguments(tables):
    from pokeai_b200 import _native as N
    from pokeai_b200 import ops
    from oracle import ref_port
    rec_s, rec_n, action, reward, terminal = _synth_transitions(8, 6, tables, seed=1)
    batch = _dev_batch(rec_s, rec_n, action, reward, terminal)
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 128, seed=3)
    prepared = ops.prepare_weights(sd, DEV)[0]
    with pytest.raises(N.NativeError) as e:
        ops.dqn_grad(prepared, prepared, 2, 128, 9, *batch, gamma=0.9)
    assert e.value.code == 4  # PK_ERR_UNSUPPORTED: loud, no fallback
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=3)
    prepared = ops.prepare_weights(sd, DEV)[0]
    with pytest.raises(N.NativeError):
        ops.dqn_grad(prepared, prepared, 2, 64, 9, *batch, gamma=0.9, n_global=4)  # n_global < n


# ---------------------------------------------------------------------------------------------------------
# Trainer (drop-in class)
# ---------------------------------------------------------------------------------------------------------
def _fill_by_self_play(trainer, tables, n_episodes, steps_per_episode, seed):
    """Drive AgentTrain the way RLPolicy does (rl_policy.py:95-112) on synthetic observations."""
    from pokeai_b200.packed import synth_packed, unpack_observation
    rec = synth_packed(n_episodes * steps_per_episode, seed=seed, party_size=trainer.feature_extractor.party_size,
                       kind="ragged", tables=tables)
    obs = [unpack_observation(r, tables) for r in rec]
    k = 0
    for _ in range(n_episodes):
        agent = trainer.get_train_agent()
        for s in range(steps_per_episode):
            a = agent.act(obs[k], 0.0 if s else 0.0)
            assert 0 <= a < len(obs[k].possible_actions)
            k += 1
        agent.stop_episode(1.0 if k % 2 else -1.0)
        trainer.extend_replay_buffer(agent._replay_buffer)
        trainer.total_battles += 1
        trainer.train()


def test_trainer_update_equals_oracle_on_the_same_minibatch(tables):
    from oracle import packed_oracle, ref_port
    from pokeai_b200.trainer import Trainer
    torch.manual_seed(0)
    np.random.seed(0)
    random.seed(0)
    tr = Trainer({"n_layers": 2, "n_channels": 64}, {"batch_size": 32, "first_update_steps": 40, "lr": 1e-3,
                                                      "target_update": 3}, {"party_size": 3})
    assert tr.model.state_dict().keys() == {"layers.0.weight", "layers.0.bias", "layers.1.weight", "layers.1.bias",
                                            "output.weight", "output.bias"}
    _fill_by_self_play(tr, tables, n_episodes=4, steps_per_episode=10, seed=5)
    assert tr.total_steps == 40 and tr.update_steps == 40 and tr.update_loss_history == []
    # one more episode: 10 new steps -> 10 updates (update_step >= first_update_steps), trainer.py:141-150
    sd_before = {k: v.cpu().clone() for k, v in tr.model.state_dict().items()}
    tgt_before = {k: v.cpu().clone() for k, v in tr.target_model.state_dict().items()}
    st = random.getstate()
    idx = tr.replay_buffer.sample_indices(32)  # what the first _update will draw
    random.setstate(st)
    s, nx, act, rew, term = tr.replay_buffer.gather(idx)
    fs, _ = packed_oracle.encode(s, 3, tables)
    nx = nx.copy()
    nx[term == 1] = s[term == 1]  # terminal rows are ignored; give the oracle a valid record to encode
    fn, mn = packed_oracle.encode(nx, 3, tables)
    loss_o, grads_o = ref_port.dqn_update_grads(sd_before, tgt_before, fs, act, rew, fn, mn, term == 0, tr.gamma)
    ref_port.clamp_adam_step(sd_before, grads_o, {}, 1e-3, 1)
    tr._update()
    assert abs(tr.update_loss_history[-1] - loss_o) < 2e-6
    for k, v in tr.model.state_dict().items():
        # first Adam step = lr * g / (|g| + eps): only well-conditioned where |g| >> the gradient tolerance
        well = np.abs(grads_o[k].numpy()) > 1e-5
        got, want = v.cpu().numpy(), sd_before[k].numpy()
        np.testing.assert_allclose(got[well], want[well], rtol=P_RTOL, atol=2e-5, err_msg=k)
        assert np.abs(got - want).max() <= 2.1e-3  # nobody moves further than both steps could differ
    # the prepared copy the kernels read follows the parameters
    from pokeai_b200 import ops
    assert torch.equal(ops.prepare_from_flat(tr._flat, 2, 64, 617), tr._prepared)


@pytest.mark.parametrize("nl,H", [(2, 64), (2, 256)])
def test_trainer_model_follows_in_place_updates(tables, nl, H):
    """the update kernels rewrite the parameters through raw pointers (torch's version counters do not move):
    trainer.model(x), an agent built on trainer.model and a DecisionPipeline on it must still see every step"""
    from oracle import packed_oracle, ref_port
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    torch.manual_seed(2)
    np.random.seed(2)
    random.seed(2)
    tr = Trainer({"n_layers": nl, "n_channels": H}, {"batch_size": 16, "first_update_steps": 20, "lr": 1e-2},
                 {"party_size": 3})
    _fill_by_self_play(tr, tables, n_episodes=3, steps_per_episode=8, seed=3)
    rec = synth_packed(40, seed=77, party_size=3, kind="ragged", tables=tables)
    fo, mo = packed_oracle.encode(rec, 3, tables)
    x = torch.from_numpy(fo).to(tr.device)
    agent = AgentVal(tr.model, tr.feature_extractor)
    tol = dict(rtol=1e-5, atol=1e-5) if H == 64 else dict(rtol=0, atol=3e-2)  # wide nets decide in bf16

    def check():
        sd = {k: v.detach().cpu().clone() for k, v in tr.model.state_dict().items()}
        with torch.no_grad():
            qo = ref_port.mlp_forward(sd, torch.from_numpy(fo)).numpy()
        if H == 64:
            np.testing.assert_allclose(tr.model(x).cpu().numpy(), qo, **tol)
        q, _ = agent.q_and_action_batch(rec)
        np.testing.assert_allclose(np.array(q), qo, **tol)
        return qo

    q0 = check()
    for _ in range(3):
        tr._update()
    q1 = check()
    assert np.abs(q1 - q0).max() > 1e-3  # the weights really moved
    tr._target_update()
    tr._update()
    check()


def test_update_step_replayed_as_cuda_graph(tables):
    """pk_update_ctl: Adam's step count lives in device memory, so a captured graph of the whole update step
    (advance, next-state pass, backward, reduce, clamp + Adam) can be replayed; same bits as the plain calls"""
    from oracle import ref_port
    from pokeai_b200 import ops
    from pokeai_b200.packed import synth_packed
    dev = torch.device("cuda:0")
    nl, H, A, n = 2, 64, 9, 1024
    sd = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=1)
    sd_t = ref_port.init_mlp_params(tables.n_channels, nl, H, seed=2)
    prep_t = ops.prepare_weights(sd_t, dev)[0]
    rng = np.random.default_rng(0)
    sets = []
    for i in range(2):
        s = synth_packed(n, seed=20 + i, party_size=6, kind="ragged", tables=tables)
        s2 = synth_packed(n, seed=40 + i, party_size=6, kind="ragged", tables=tables)
        sets.append((ops.packed_to_device(s, dev), ops.packed_to_device(s2, dev),
                     torch.from_numpy((rng.random(n) * s["n_act"]).astype(np.int32)).to(dev),
                     torch.from_numpy(rng.uniform(-1, 1, n).astype(np.float32)).to(dev),
                     torch.from_numpy((rng.random(n) < 0.05).astype(np.uint8)).to(dev)))

    def fresh():
        prep, _, _, flat = ops.prepare_weights(sd, dev)
        return prep, flat, torch.zeros_like(flat), torch.zeros_like(flat)

    # plain: launch arguments carry the step count
    prep_a, flat_a, ma, va = fresh()
    for k in range(1, 6):
        g, _ = ops.dqn_grad(prep_a, prep_t, nl, H, A, *sets[k % 2], gamma=0.95)
        ops.clamp_adam(flat_a, g, ma, va, nl, H, k, lr=1e-2, prepared=prep_a)
    # graphs: one per input set, the step count advances on the device
    prep_b, flat_b, mb, vb = fresh()
    ctl = ops.make_update_ctl(dev, lr=1e-2)
    grad, loss = torch.empty_like(flat_b), torch.zeros(1, device=dev)
    graphs = {}
    side = torch.cuda.Stream(device=dev)
    for p in (0, 1):
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=side):
            ops.update_ctl_advance(ctl)
            ops.dqn_grad(prep_b, prep_t, nl, H, A, *sets[p], gamma=0.95, out_grad=grad, out_loss=loss)
            ops.clamp_adam_ctl(flat_b, grad, mb, vb, nl, H, ctl, prepared=prep_b)
        graphs[p] = gph
    for k in range(1, 6):
        graphs[k % 2].replay()
    torch.cuda.synchronize()
    assert int(ctl.cpu()[1]) == 5  # the device-side step count
    assert float((flat_a - flat_b).abs().max()) <= 1e-7
    assert torch.equal(ops.prepare_from_flat(flat_b, nl, H, tables.n_channels), prep_b)


def test_resume_from_a_reference_style_checkpoint_with_dense_replay(synth_golden, tables):
    """trainer.py:73-117: a `resume=True` state whose replay buffer is the reference's (a deque of dense (617, A) items)
    loads: the items are converted to packed records that encode to the same bits, and training goes on"""
    from collections import deque, namedtuple
    from oracle import packed_oracle
    from pokeai_b200.trainer import Trainer
    rec, ps, feat, mask = synth_golden["ragged6"]
    Item = namedtuple("ReplayBufferItem", "state action_mask action next_state next_action_mask reward")

    class RefReplay:  # replay_buffer.py:18-32
        def __init__(self, n):
            self.buffer = deque(maxlen=n)

    torch.manual_seed(4)
    random.seed(4)
    tr = Trainer({"n_layers": 2, "n_channels": 64}, {"batch_size": 16, "first_update_steps": 0, "replay_buffer_size": 500},
                 {"party_size": 6})
    state = tr.save_state(resume=True)
    ref_rb = RefReplay(500)
    for i in range(0, 120, 2):
        n_act = int(mask[i].sum())
        last = i % 20 == 18
        ref_rb.buffer.append(Item(feat[i], mask[i], i % max(n_act, 1), None if last else feat[i + 1],
                                  None if last else mask[i + 1], float(i % 3) - 1.0))
    state["replay_buffer"] = ref_rb
    tr2 = Trainer.load_state(state, resume=True)
    assert len(tr2.replay_buffer) == 60 and tr2.replay_buffer.maxlen == 500
    s, nx, act, rew, term = tr2.replay_buffer.gather(np.arange(60))
    fs, ms = packed_oracle.encode(s, ps, tables)
    assert np.array_equal(fs.view(np.uint32), feat[0:120:2].view(np.uint32)) and np.array_equal(ms, mask[0:120:2])
    before = tr2._flat.clone()
    tr2._update()
    assert len(tr2.update_loss_history) == 1 and np.isfinite(tr2.update_loss_history[0])
    assert not torch.equal(before, tr2._flat)


def test_trainer_schedule_checkpoint_and_agents(tables):
    from pokeai_b200.trainer import Trainer, DQN_DEFAULT_PARAMS
    assert DQN_DEFAULT_PARAMS["batch_size"] == 32 and DQN_DEFAULT_PARAMS["target_update"] == 100
    torch.manual_seed(1)
    np.random.seed(1)
    random.seed(1)
    tr = Trainer({"n_layers": 3, "n_channels": 16}, {"batch_size": 16, "first_update_steps": 20, "target_update": 5,
                                                      "epsilon": 0.5, "epsilon_decay": 0.01, "epsilon_min": 0.1},
                 {"party_size": 6})
    _fill_by_self_play(tr, tables, n_episodes=6, steps_per_episode=8, seed=11)
    assert tr.total_steps == 48 and tr.update_steps == 48 and tr.total_battles == 6
    assert len(tr.update_loss_history) == 48 - 20 and all(np.isfinite(tr.update_loss_history))
    assert tr.optimizer.step_count == 28
    # epsilon schedule (trainer.py:127)
    ag = tr.get_train_agent()
    assert ag._epsilon == pytest.approx(max((1 - 0.01) ** 48 * 0.5, 0.1))
    # target network: last hard copy at update_step 45 (multiples of optimize_per_steps * target_update)
    assert not torch.equal(tr._flat, tr._flat_target)
    tr._target_update()
    assert torch.equal(tr._flat, tr._flat_target) and torch.equal(tr._prepared, tr._prepared_target)
    # checkpoint round trip (trainer.py:73-117), through pickle like the reference's GridFS blob
    import pickle
    state = pickle.loads(pickle.dumps(tr.save_state(resume=True)))
    assert set(state) == {"model", "target_model", "optimizer", "replay_buffer", "constructor_params", "total_steps",
                          "total_battles", "update_steps", "update_loss_history"}
    assert all(not v.is_cuda for v in state["model"].values())
    tr2 = Trainer.load_state(state, resume=True)
    assert tr2.total_steps == 48 and tr2.update_steps == 48 and len(tr2.replay_buffer) == 48
    assert torch.equal(tr2._flat, tr._flat) and torch.equal(tr2._prepared, tr._prepared)
    assert torch.equal(tr2.optimizer.exp_avg, tr.optimizer.exp_avg) and tr2.optimizer.step_count == 28
    # both continue identically from the checkpoint
    random.seed(77)
    tr._update()
    random.seed(77)
    tr2._update()
    assert torch.equal(tr2._flat, tr._flat)
    # the optimizer state loads into a real torch Adam over a reference-shaped model (checkpoint interchange)
    ref_like = torch.nn.ModuleDict()  # parameter shapes only
    params = [torch.nn.Parameter(v.clone()) for v in state["model"].values()]
    torch.optim.Adam(params, lr=1e-3).load_state_dict(state["optimizer"])
    light = tr.save_state(resume=False)
    assert set(light) == {"model", "constructor_params", "update_steps", "total_battles"}
    tr3 = Trainer.load_state(light)
    va, vb = tr.get_val_agent(), tr3.get_val_agent()
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(200, seed=3, party_size=6, kind="ragged", tables=tables)
    assert np.array_equal(va.act_batch(rec), vb.act_batch(rec))


def test_trainer_all_terminal_minibatch_raises_like_the_reference(tables):
    from pokeai_b200.replay_buffer import ReplayBufferItem
    from pokeai_b200.packed import synth_packed
    from pokeai_b200.trainer import Trainer
    tr = Trainer({}, {"batch_size": 4}, {"party_size": 6})
    rec = synth_packed(4, seed=1, party_size=6, kind="full", tables=tables)
    for r in rec:
        tr.replay_buffer.append(ReplayBufferItem(r, None, 0, None, None, 1.0))
    with pytest.raises(ValueError):  # np.stack([]) at trainer.py:157
        tr._update()
    with pytest.raises(IndexError):  # replay_buffer.py:30-31
        Trainer({}, {"batch_size": 64}, {"party_size": 6})._update()


def test_peer_gradient_exchange_two_gpus():
    """pk_peer_clamp_adam (gradient sum over NVLink peer memory fused with clamp + Adam) against the NCCL route, one
    process per GPU: tools/check_peer_update.py under torchrun.  Needs two GPUs; skipped on a one-GPU box."""
    import subprocess
    import sys
    from pathlib import Path
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = Path(__file__).resolve().parents[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29581", str(root / "tools" / "check_peer_update.py")],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "peer update ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
