"""Golden vectors of the REAL reference's Trainer._update with `bn: true` (train-mode BatchNorm in every forward of the
update; the reference never calls .eval() on Trainer.model / target_model).  Run in the build container, where
/root/reference is importable; writes tests/golden/dqn_update_bn.npz.  Companion of tools/make_golden.py (same
transition set-up as its section 4), kept separate so that the other fixtures keep their bytes.

  case a: 2 x 16, Double DQN, 32 transitions, party size 3
  case b: 3 x 24, plain DQN, 48 transitions, party size 3

Per case: the packed records, action / reward / terminal, the online and target state_dicts before the first update
(BatchNorm weights, biases and running statistics randomised so that none sits at its default), and after each of two
updates: loss, pre-clamp gradients, the online state_dict and the target's (its running statistics move too).
"""
import random
import sys
from pathlib import Path

sys.dont_write_bytecode = True
sys.path.insert(0, str(Path(__file__).resolve().parent))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from make_golden import GOLD, RefMLP, RefTrainer, ReplayBufferItem, T, ref_transform  # noqa: E402
from pokeai_b200.packed import synth_packed  # noqa: E402


def perturb_bn(model, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for b in model.bn_layers:
            b.weight.copy_(torch.rand(b.weight.shape, generator=g) + 0.5)
            b.bias.copy_(torch.randn(b.bias.shape, generator=g) * 0.1)
            b.running_mean.copy_(torch.randn(b.running_mean.shape, generator=g) * 0.2)
            b.running_var.copy_(torch.rand(b.running_var.shape, generator=g) + 0.5)


def case(tag, nl, nc, double_dqn, N, seed, out):
    ps, A = 3, 6
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    tr = RefTrainer({"n_layers": nl, "n_channels": nc, "bn": True},
                    {"batch_size": N, "gamma": 0.95, "lr": 1e-3, "replay_buffer_size": N, "double_dqn": double_dqn},
                    {"party_size": ps})
    perturb_bn(tr.model, seed + 100)
    torch.manual_seed(seed + 1)
    tgt = RefMLP((T.n_channels, A), A, n_layers=nl, n_channels=nc, bn=True)
    perturb_bn(tgt, seed + 200)
    tr.target_model.load_state_dict(tgt.state_dict())
    assert tr.model.training and tr.target_model.training  # the point of this fixture
    rec_s = synth_packed(N, seed=seed + 21, party_size=ps, kind="ragged", tables=T)
    rec_n = synth_packed(N, seed=seed + 22, party_size=ps, kind="ragged", tables=T)
    fs, ms = ref_transform(rec_s, ps)
    fn, mn = ref_transform(rec_n, ps)
    rng = np.random.default_rng(seed + 23)
    terminal = rng.random(N) < 0.2
    terminal[0] = True
    action = np.array([rng.choice(np.flatnonzero(ms[i])) for i in range(N)], dtype=np.int64)
    reward = (rng.random(N) * 4 - 2).astype(np.float32)
    for i in range(N):
        tr.replay_buffer.append(ReplayBufferItem(fs[i], ms[i], int(action[i]),
                                                 None if terminal[i] else fn[i],
                                                 None if terminal[i] else mn[i], float(reward[i])))
    p = tag + "_"
    out.update({p + "party_size": np.array(ps), p + "n_layers": np.array(nl), p + "n_channels": np.array(nc),
                p + "double_dqn": np.array(double_dqn), p + "gamma": np.array(0.95), p + "lr": np.array(1e-3),
                p + "packed_s": np.frombuffer(rec_s.tobytes(), dtype=np.uint8).reshape(N, 144),
                p + "packed_n": np.frombuffer(rec_n.tobytes(), dtype=np.uint8).reshape(N, 144),
                p + "terminal": terminal, p + "action": action, p + "reward": reward})
    for k, v in tr.model.state_dict().items():
        out[f"{p}sd0_{k}"] = v.numpy().copy()
    for k, v in tr.target_model.state_dict().items():
        out[f"{p}tgt0_{k}"] = v.numpy().copy()
    names = [k for k, _ in tr.model.named_parameters()]
    for step in (1, 2):
        captured = {}
        hooks = [q.register_hook(lambda g, k=k: captured.__setitem__(k, g.detach().clone()))
                 for k, q in tr.model.named_parameters()]
        tr._update()  # random.sample of all N items == a permutation of the buffer
        for h in hooks:
            h.remove()
        out[f"{p}loss{step}"] = np.array(tr.update_loss_history[-1], dtype=np.float64)
        for k in names:
            out[f"{p}grad{step}_{k}"] = captured[k].numpy().copy()
        for k, v in tr.model.state_dict().items():
            out[f"{p}sd{step}_{k}"] = v.numpy().copy()
        for k, v in tr.target_model.state_dict().items():
            if "bn_layers" in k:  # only its BatchNorm buffers move in an update
                out[f"{p}tgt{step}_{k}"] = v.numpy().copy()
    # the eval-mode Q of the updated network on the state batch (BatchNorm folded with the running statistics)
    tr.model.eval()
    with torch.no_grad():
        out[p + "q_eval"] = tr.model(torch.from_numpy(fs)).numpy()


def main():
    out = {}
    case("a", 2, 16, True, 32, 5, out)
    case("b", 3, 24, False, 48, 9, out)
    path = GOLD / "dqn_update_bn.npz"
    np.savez_compressed(path, **out)
    print(path.name, path.stat().st_size)


if __name__ == "__main__":
    main()
