"""Generate tests/golden/* by running the REAL reference (select766/pokeai, mounted
read-only at /root/reference) in the build container.  The GPU box has no reference,
so the outputs are committed as small fixtures; this script is how they were made.

    PYTHONDONTWRITEBYTECODE=1 python tools/make_golden.py

What is recorded (all produced by the reference's own classes):
  fixture_decisions.json.gz  220 logged real decisions (battle_status, possible_actions,
                             q_func, action) from
                             reproduce/ipynb_202104/rl_loop_210227_1_tops_battles.top10.zip
  fixture_features.npz       FeatureExtractor(3).transform of each of them (sparse form)
  synth_features.npz         seeded synthetic packed records (party sizes 6, 3, 1; full and
                             ragged) + FeatureExtractor.transform outputs (sparse form)
  mlp_forward.npz            MLPModel weights (2x64 default @A=9, 3x16 @A=6) + Q outputs on
                             a slice of the synthetic observations, torch CPU fp32
  dqn_update.npz             Trainer._update x2 on a fixed 32-transition buffer: loss,
                             pre-clamp grads, parameters after each Adam step
  dim_meanings.json          FeatureExtractor(6).get_dim_meanings()
"""
import gzip
import json
import os
import random
import sys
import zipfile
from pathlib import Path

sys.dont_write_bytecode = True
ROOT = Path(__file__).resolve().parents[1]
REF = Path(os.environ.get("POKEAI_REFERENCE", "/root/reference"))
sys.path.insert(0, str(ROOT))
sys.path.insert(1, str(REF))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from pokeai.ai.generic_move_model.feature_extractor import FeatureExtractor as RefFE  # noqa: E402
from pokeai.ai.generic_move_model.mlp_model import MLPModel as RefMLP  # noqa: E402
from pokeai.ai.generic_move_model.replay_buffer import ReplayBufferItem  # noqa: E402
from pokeai.ai.generic_move_model.trainer import Trainer as RefTrainer  # noqa: E402
from pokeai.ai.battle_status import BattleStatus as RefBS, SideStatus as RefSS, ActivePokeStatus as RefAP  # noqa: E402
from pokeai.ai.common import PossibleAction as RefPA  # noqa: E402
from pokeai.ai.rl_policy_observation import RLPolicyObservation as RefObs  # noqa: E402

from pokeai_b200.packed import synth_packed, unpack_observation  # noqa: E402
from pokeai_b200.tables import load_tables  # noqa: E402

GOLD = ROOT / "tests" / "golden"
GOLD.mkdir(parents=True, exist_ok=True)
T = load_tables()


def sparse(f: np.ndarray) -> dict:
    idx = np.flatnonzero(f)
    return {"shape": np.array(f.shape, dtype=np.int64), "idx": idx.astype(np.uint32),
            "val": f.ravel()[idx].astype(np.float32)}


def to_ref_obs(o):
    """our mirror objects -> the reference's own classes (so that nothing of ours is on the oracle side)"""
    b = o.battle_status
    rb = RefBS(b.side_friend, b.side_party)
    rb.turn = b.turn
    rb.weather = b.weather
    for side, ss in b.side_statuses.items():
        rs = RefSS()
        rs.total_pokes, rs.remaining_pokes = ss.total_pokes, ss.remaining_pokes
        a = ss.active
        ra = RefAP(a.pokemon, a.species, a.level, a.gender, a.hp_current, a.hp_max, a.status)
        ra.ranks = dict(a.ranks)
        rs.active = ra
        rb.side_statuses[side] = rs
    return RefObs(rb, o.request, [RefPA(*p) for p in o.possible_actions])


def ref_transform(rec, party_size):
    fe = RefFE(party_size)
    outs = [fe.transform(to_ref_obs(unpack_observation(r, T))) for r in rec]
    return np.stack([o[0] for o in outs]), np.stack([o[1] for o in outs])


def main():
    torch.set_num_threads(1)
    # ---- 1. the real battle-log fixture -------------------------------------------------
    z = zipfile.ZipFile(REF / "reproduce/ipynb_202104/rl_loop_210227_1_tops_battles.top10.zip")
    lines = z.read(z.infolist()[0].filename).decode().strip().split("\n")
    from pokeai_b200.observation import battle_status_from_json, possible_actions_from_json, RLPolicyObservation
    decisions, feats, masks = [], [], []
    fe3 = RefFE(3)
    for bi, line in enumerate(lines):
        for ev in json.loads(line)["events"]:
            if ev["type"] != "choice" or not ev["choice"].get("q_func"):
                continue
            c = ev["choice"]
            bs = dict(c["battle_status"])
            bs["side_party"] = None  # not read by the path; keeps the fixture small
            decisions.append({"battle": bi, "player": c["player"], "battle_status": bs,
                              "possible_actions": c["possible_actions"], "q_func": c["q_func"]["q_func"],
                              "action": c["q_func"]["action"], "choice": c["choice"]})
            obs = RLPolicyObservation(battle_status_from_json(bs), {}, possible_actions_from_json(c["possible_actions"]))
            f, m = fe3.transform(to_ref_obs(obs))
            feats.append(f)
            masks.append(m)
    with gzip.open(GOLD / "fixture_decisions.json.gz", "wt", encoding="utf-8") as f:
        json.dump(decisions, f, separators=(",", ":"))  # -Infinity is kept as the reference logs it
    np.savez_compressed(GOLD / "fixture_features.npz", mask=np.stack(masks), **sparse(np.stack(feats)))
    print("fixture decisions:", len(decisions))

    # ---- 2. synthetic packed records ------------------------------------------------------
    synth = {}
    for name, n, seed, ps, kind in [("full6", 192, 11, 6, "full"), ("ragged6", 192, 12, 6, "ragged"),
                                     ("ragged3", 96, 13, 3, "ragged"), ("ragged1", 48, 14, 1, "ragged"),
                                     ("full4", 48, 15, 4, "full")]:
        rec = synth_packed(n, seed=seed, party_size=ps, kind=kind, tables=T)
        f, m = ref_transform(rec, ps)
        synth[f"{name}_packed"] = np.frombuffer(rec.tobytes(), dtype=np.uint8).reshape(n, 144)
        synth[f"{name}_party_size"] = np.array(ps)
        synth[f"{name}_mask"] = m
        for k, v in sparse(f).items():
            synth[f"{name}_{k}"] = v
    np.savez_compressed(GOLD / "synth_features.npz", **synth)

    # ---- 3. MLP forward -------------------------------------------------------------------
    out = {}
    for name, ps, nl, nc, src in [("default", 6, 2, 64, "full6"), ("small", 3, 3, 16, "ragged3")]:
        torch.manual_seed(0)
        A = 4 + ps - 1
        model = RefMLP((T.n_channels, A), A, n_layers=nl, n_channels=nc, bn=False)
        model.eval()
        n = 64
        rec = np.frombuffer(synth[f"{src}_packed"].tobytes(), dtype=synth_packed(1).dtype)[:n]
        f, m = ref_transform(rec, ps)
        with torch.no_grad():
            q = model(torch.from_numpy(f)).numpy()
        out[f"{name}_q"] = q
        out[f"{name}_n"] = np.array(n)
        out[f"{name}_src"] = np.array(src)
        out[f"{name}_n_layers"] = np.array(nl)
        out[f"{name}_n_channels"] = np.array(nc)
        for k, v in model.state_dict().items():
            out[f"{name}_sd_{k}"] = v.numpy()
    np.savez_compressed(GOLD / "mlp_forward.npz", **out)

    # ---- 4. two DQN updates through the reference Trainer -------------------------------
    ps, A, N = 3, 6, 32
    random.seed(5)
    np.random.seed(5)
    torch.manual_seed(5)
    tr = RefTrainer({"n_layers": 2, "n_channels": 16, "bn": False},
                    {"batch_size": N, "gamma": 0.95, "lr": 1e-3, "replay_buffer_size": N, "double_dqn": True},
                    {"party_size": ps})
    # make the target network differ from the online one, as it does mid-training
    torch.manual_seed(6)
    tgt = RefMLP((T.n_channels, A), A, n_layers=2, n_channels=16, bn=False)
    tr.target_model.load_state_dict(tgt.state_dict())
    rec_s = synth_packed(N, seed=21, party_size=ps, kind="ragged", tables=T)
    rec_n = synth_packed(N, seed=22, party_size=ps, kind="ragged", tables=T)
    fs, ms = ref_transform(rec_s, ps)
    fn, mn = ref_transform(rec_n, ps)
    rng = np.random.default_rng(23)
    terminal = rng.random(N) < 0.2
    terminal[0] = True
    action = np.array([rng.choice(np.flatnonzero(ms[i])) for i in range(N)], dtype=np.int64)
    reward = (rng.random(N) * 4 - 2).astype(np.float32)  # spans both Huber branches
    for i in range(N):
        tr.replay_buffer.append(ReplayBufferItem(fs[i], ms[i], int(action[i]),
                                                 None if terminal[i] else fn[i],
                                                 None if terminal[i] else mn[i], float(reward[i])))
    up = {"party_size": np.array(ps), "n_layers": np.array(2), "n_channels": np.array(16),
          "gamma": np.array(0.95), "lr": np.array(1e-3),
          "packed_s": np.frombuffer(rec_s.tobytes(), dtype=np.uint8).reshape(N, 144),
          "packed_n": np.frombuffer(rec_n.tobytes(), dtype=np.uint8).reshape(N, 144),
          "terminal": terminal, "action": action, "reward": reward}
    for k, v in tr.model.state_dict().items():
        up[f"sd0_{k}"] = v.numpy().copy()
    for k, v in tr.target_model.state_dict().items():
        up[f"tgt_{k}"] = v.numpy().copy()
    names = [k for k, _ in tr.model.named_parameters()]
    for step in (1, 2):
        captured = {}
        hooks = [p.register_hook(lambda g, k=k: captured.__setitem__(k, g.detach().clone()))
                 for k, p in tr.model.named_parameters()]
        tr._update()  # random.sample of all 32 items == a permutation of the buffer
        for h in hooks:
            h.remove()
        up[f"loss{step}"] = np.array(tr.update_loss_history[-1], dtype=np.float64)
        for k in names:
            up[f"grad{step}_{k}"] = captured[k].numpy().copy()
        for k, v in tr.model.state_dict().items():
            up[f"sd{step}_{k}"] = v.numpy().copy()
    np.savez_compressed(GOLD / "dqn_update.npz", **up)

    # ---- 5. row meanings ------------------------------------------------------------------
    with open(GOLD / "dim_meanings.json", "w", encoding="utf-8") as f:
        json.dump({"party_size_6": RefFE(6).get_dim_meanings(), "input_shape_6": list(RefFE(6).input_shape),
                   "input_shape_3": list(RefFE(3).input_shape)}, f)
    for p in sorted(GOLD.iterdir()):
        print(p.name, p.stat().st_size)


if __name__ == "__main__":
    main()
