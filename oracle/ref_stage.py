"""ORACLE side (test infrastructure, not product code): the UNMODIFIED reference, staged for the GPU box.

The reference is pure Python.  `stage()` -- called by `__graft_entry__.build()` in the build
container, where /root/reference exists -- copies the hot-path modules of select766/pokeai and the four JSON tables
they read, byte for byte and with their relative paths, into the git-ignored `oracle/_ref/`, which travels to the GPU
box with the snapshot (nothing under `oracle/_ref/` enters the history; nothing here edits a reference file).
`load()` puts `oracle/_ref` on sys.path and returns the reference's own classes, or None when it was never staged.

Used by: `bench.py --impl reference` and `bench.py`'s `cpu_baseline` (kind "reference"; they fall back to the port
`oracle/ref_port.py`, kind "port", only when `oracle/_ref` is absent), and `tests/test_oracle_vs_reference.py`.
The product path (pokeai_b200/) never imports this module.

Reference call sites reproduced by `decide_step()`: feature_extractor.py:40-52 (one transform per observation, a
Python loop), agent.py:28-30 (`_calc_q_vector_batch`), agent.py:34-35 / party_evaluator_quick.py:101-104
(mask -> -inf, argmax).
"""
from __future__ import annotations

import os
import shutil
import sys
from pathlib import Path
from types import SimpleNamespace

HERE = Path(__file__).resolve().parent
REF_DST = HERE / "_ref"
REF_SRC = Path(os.environ.get("POKEAI_REFERENCE", "/root/reference"))

# everything `import pokeai.ai.generic_move_model.{feature_extractor,agent,trainer}` and
# `pokeai.ai.party_feature.party_evaluator_quick` pull in (probed in the build container), plus the tables
FILES = [
    "pokeai/__init__.py", "pokeai/util.py",
    "pokeai/ai/__init__.py", "pokeai/ai/battle_status.py", "pokeai/ai/common.py", "pokeai/ai/dex.py",
    "pokeai/ai/rl_policy_observation.py", "pokeai/ai/state_feature_extractor.py",
    "pokeai/ai/generic_move_model/__init__.py", "pokeai/ai/generic_move_model/agent.py",
    "pokeai/ai/generic_move_model/agent_train.py", "pokeai/ai/generic_move_model/agent_val.py",
    "pokeai/ai/generic_move_model/choice_to_vec.py", "pokeai/ai/generic_move_model/feature_extractor.py",
    "pokeai/ai/generic_move_model/mlp_model.py", "pokeai/ai/generic_move_model/replay_buffer.py",
    "pokeai/ai/generic_move_model/trainer.py",
    "pokeai/ai/party_feature/__init__.py", "pokeai/ai/party_feature/party_evaluator.py",
    "pokeai/ai/party_feature/party_evaluator_quick.py",
    "pokeai/sim/__init__.py", "pokeai/sim/party_generator.py",
    "data/dataset/all_pokemons.json", "data/dataset/all_moves_with_hiddenpower_type.json",
    "data/dataset/all_items.json", "data/dataset/pokedex.json",
]


def stage(verbose: bool = False) -> bool:
    """Copy the reference's hot-path files into oracle/_ref (build container only).  Returns True when staged."""
    if not (REF_SRC / "pokeai" / "ai").is_dir():
        return (REF_DST / "pokeai").is_dir()
    for rel in FILES:
        src, dst = REF_SRC / rel, REF_DST / rel
        if not src.exists():
            if rel.endswith("__init__.py"):  # namespace package in the reference: keep it one
                continue
            raise FileNotFoundError(f"reference file missing: {src}")
        dst.parent.mkdir(parents=True, exist_ok=True)
        if not dst.exists() or dst.read_bytes() != src.read_bytes():
            shutil.copyfile(src, dst)
    (REF_DST / "STAGED_FROM").write_text(f"{REF_SRC} (unmodified copies; see oracle/ref_stage.py)\n")
    if verbose:
        print(f"staged {len(FILES)} reference files into {REF_DST}")
    return True


_loaded = None


def load():
    """The reference's own classes (imported from oracle/_ref), or None when the reference was not staged."""
    global _loaded
    if _loaded is not None:
        return _loaded or None
    if not (REF_DST / "pokeai" / "ai" / "generic_move_model" / "feature_extractor.py").exists():
        _loaded = False
        return None
    sys.dont_write_bytecode = True
    if str(REF_DST) not in sys.path:
        sys.path.insert(0, str(REF_DST))
    from pokeai.ai.battle_status import ActivePokeStatus, BattleStatus, SideStatus
    from pokeai.ai.common import PossibleAction
    from pokeai.ai.generic_move_model.agent import Agent
    from pokeai.ai.generic_move_model.feature_extractor import FeatureExtractor
    from pokeai.ai.generic_move_model.mlp_model import MLPModel
    from pokeai.ai.rl_policy_observation import RLPolicyObservation
    _loaded = SimpleNamespace(FeatureExtractor=FeatureExtractor, MLPModel=MLPModel, Agent=Agent,
                              BattleStatus=BattleStatus, SideStatus=SideStatus, ActivePokeStatus=ActivePokeStatus,
                              PossibleAction=PossibleAction, RLPolicyObservation=RLPolicyObservation, root=REF_DST)
    return _loaded


def to_ref_obs(R, o):
    """mirror observation (pokeai_b200.observation) -> the reference's own classes"""
    b = o.battle_status
    rb = R.BattleStatus(b.side_friend, b.side_party)
    rb.turn = b.turn
    rb.weather = b.weather
    for side, ss in b.side_statuses.items():
        rs = R.SideStatus()
        rs.total_pokes, rs.remaining_pokes = ss.total_pokes, ss.remaining_pokes
        a = ss.active
        ra = R.ActivePokeStatus(a.pokemon, a.species, a.level, a.gender, a.hp_current, a.hp_max, a.status)
        ra.ranks = dict(a.ranks)
        rs.active = ra
        rb.side_statuses[side] = rs
    return R.RLPolicyObservation(rb, o.request, [R.PossibleAction(*p) for p in o.possible_actions])


def decide_step(R, observations, party_size: int, state_dict, n_layers: int, n_channels: int):
    """Returns step(n) -> actions of the first n observations, computed by the reference exactly as it runs:
    a Python loop of FeatureExtractor.transform, np.stack, Agent._calc_q_vector_batch (MLPModel on torch CPU),
    q[mask == 0] = -inf, argmax."""
    import numpy as np
    import torch
    fe = R.FeatureExtractor(party_size)
    model = R.MLPModel(fe.input_shape, fe.output_dim, n_layers=n_layers, n_channels=n_channels, bn=False)
    model.load_state_dict(state_dict)
    model.eval()
    agent = R.Agent(model, fe)

    def step(n):
        outs = [fe.transform(o) for o in observations[:n]]
        feats = np.stack([o[0] for o in outs])
        masks = np.stack([o[1] for o in outs])
        with torch.no_grad():
            q = agent._calc_q_vector_batch(feats)
        q[masks == 0] = -np.inf
        return np.argmax(q, axis=1)

    return step


if __name__ == "__main__":
    print("staged" if stage(verbose=True) else "reference not available; nothing staged")
