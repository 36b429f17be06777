"""Lock-step battle driver (pokeai_b200/lockstep.py).

CPU: the driver's bookkeeping with a stand-in agent that decides through the oracle -- batched lock-step decisions must
equal the per-battle, one-at-a-time decisions `RLPolicy._choice_by_model` makes (rl_policy.py:76-105), single-choice
bypass and forced switches included.
`reference` marker (build container only): the recorded battle log's raw protocol replayed through the REFERENCE's own
BattleStreamProcessor (battle_stream_processor.py:113-147) for N interleaved copies; the driver must make the same
choices as the reference's RLPolicy + AgentVal + FeatureExtractor + MLPModel running each battle alone.
GPU: 64 synthetic battles through the real kernels, and lock-step experience collection with training agents."""
import json
import os
import re
import sys
import zipfile
from pathlib import Path

import numpy as np
import pytest

REF = Path(os.environ.get("POKEAI_REFERENCE", "/root/reference"))


class OracleAgent:
    """AgentVal stand-in for machines without a GPU: act_batch through the oracle (test infrastructure)."""

    class _FE:
        def __init__(self, party_size):
            self.party_size = party_size
            self.output_dim = 4 + party_size - 1

    def __init__(self, party_size, sd, tables):
        self._model = object()
        self._feature_extractor = self._FE(party_size)
        self.sd, self.tables, self.calls, self.batch_sizes = sd, tables, 0, []

    def act_batch(self, rec):
        import torch
        from oracle import packed_oracle, ref_port
        f, m = packed_oracle.encode(rec, self._feature_extractor.party_size, self.tables)
        with torch.no_grad():
            q = ref_port.mlp_forward(self.sd, torch.from_numpy(f)).numpy()
        self.calls += 1
        self.batch_sizes.append(len(rec))
        return ref_port.masked_argmax(q, m)

    def act(self, rec, reward):
        return int(self.act_batch(rec)[0])

    def stop_episode(self, reward):
        pass


def _sequential_keys(agent, pairs, tables, max_actions):
    """what RLPolicy._choice_by_model returns for each (battle_status, request), one call at a time"""
    from pokeai_b200.protocol import pack_request
    out = []
    for bs, request in pairs:
        rec, keys = pack_request(bs, request, tables, max_actions)
        out.append(keys[0] if len(keys) == 1 else keys[int(agent.act_batch(rec)[0])])
    return out


@pytest.mark.parametrize("party_size,n_battles", [(6, 16), (3, 5), (1, 3)])
def test_lockstep_equals_per_battle_decisions(tables, party_size, n_battles):
    from oracle import ref_port
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=party_size)
    agent = OracleAgent(party_size, sd, tables)
    battles = [SyntheticBattle(seed=40 + i, party_size=party_size, tables=tables) for i in range(n_battles)]
    for b in battles:
        b.keep_history = True
    drv = LockStepDriver(agent, battles, tables=tables, use_pipeline=False)
    rounds = 12
    made = sum(drv.step() for _ in range(rounds))
    assert made == rounds * n_battles                      # every request got an answer
    assert agent.calls <= rounds                           # at most ONE model call per round
    assert drv.decisions == sum(agent.batch_sizes)
    n_bypass = 0
    for b in battles:
        assert len(b.choices) == rounds
        ref_agent = OracleAgent(party_size, sd, tables)
        assert b.choices == _sequential_keys(ref_agent, b.history, tables, 4 + party_size - 1)
        n_bypass += rounds - ref_agent.calls
        for (bs, request), key in zip(b.history, b.choices):
            assert key.startswith("switch") if request.get("forceSwitch") else key.split()[0] in ("move", "switch")
    if party_size == 1:
        assert n_bypass > 0                                # single legal action: the model is not asked


def test_lockstep_keeps_surrogate_reward_state(tables):
    """potentials are recorded exactly when the reference records them (after a model decision, not on a bypass)"""
    from collections import namedtuple
    from oracle import ref_port
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    from pokeai_b200.protocol import PackedRLPolicy
    Cfg = namedtuple("Cfg", "hp_ratio alive_ratio only_opponent offset_at_end")
    cfg = Cfg(0.25, 0.25, False, True)
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=3)
    agent = OracleAgent(2, sd, tables)
    battles = [SyntheticBattle(seed=7 + i, party_size=2, tables=tables) for i in range(4)]
    for b in battles:
        b.keep_history = True
    drv = LockStepDriver(agent, battles, surrogate_reward_config=cfg, tables=tables, use_pipeline=False)
    for _ in range(8):
        drv.step()
    for i, b in enumerate(battles):
        solo = PackedRLPolicy(OracleAgent(2, sd, tables), cfg, tables)
        solo.game_start()
        for (bs, request), key in zip(b.history, b.choices):
            fn = solo.choice_force_switch if request.get("forceSwitch") else solo.choice_turn_start
            assert fn(bs, request) == key
        assert drv.policies[i].inner.last_reward_potential == solo.last_reward_potential


def test_lockstep_rejects_agents_with_different_models(tables):
    from oracle import ref_port
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    sd = ref_port.init_mlp_params(tables.n_channels, 2, 64, seed=3)
    a, b = OracleAgent(3, sd, tables), OracleAgent(3, sd, tables)
    with pytest.raises(ValueError):
        LockStepDriver([a, b], [SyntheticBattle(1, 3, tables), SyntheticBattle(2, 3, tables)], tables=tables,
                       use_pipeline=False)


# ---- the recorded protocol through the reference's own parser ---------------------------------------------------------
def _extract_update_for_side(side, chunk_data):
    """pokeai/sim/sim.py:101-110 (a port of Pokemon-Showdown/sim/battle.ts): drop the other side's secret lines"""
    chunk_data = re.sub('\n\\|split\\|' + side + '\n([^\n]*)\n(?:[^\n]*)', '\n\\1', chunk_data)
    return re.sub('\n\\|split\\|(?:[^\n]*)\n(?:[^\n]*)\n\n?', '\n', chunk_data)


def _recorded_battles():
    z = zipfile.ZipFile(REF / "reproduce/ipynb_202104/rl_loop_210227_1_tops_battles.top10.zip")
    lines = z.read(z.infolist()[0].filename).decode().strip().split("\n")
    battles = []
    for line in lines:
        ev = json.loads(line)["events"]
        steps = []  # (update chunk, {player: request}) in protocol order
        for i, e in enumerate(ev):
            if e["type"] != "update":
                continue
            reqs = {}
            for f in ev[i + 1:]:
                if f["type"] != "choice":
                    break
                reqs[f["choice"]["player"]] = f["choice"]["request"]
            steps.append(("\n".join(e["update"]), reqs))
        battles.append(steps)
    return battles


@pytest.mark.reference
@pytest.mark.skipif(not (REF / "pokeai" / "sim" / "battle_stream_processor.py").exists(), reason="no reference checkout")
def test_recorded_protocol_through_reference_parser(tables):
    sys.dont_write_bytecode = True
    if str(REF) not in sys.path:
        sys.path.insert(1, str(REF))
    import torch
    from pokeai.ai.generic_move_model.agent_val import AgentVal as RefAgentVal
    from pokeai.ai.generic_move_model.feature_extractor import FeatureExtractor as RefFE
    from pokeai.ai.generic_move_model.mlp_model import MLPModel as RefMLP
    from pokeai.ai.rl_policy import RLPolicy as RefRLPolicy
    from pokeai.ai.surrogate_reward_config import SurrogateRewardConfig
    from pokeai.sim.battle_stream_processor import BattleStreamProcessor
    from oracle import ref_port
    from pokeai_b200.lockstep import LockStepDriver

    battles = _recorded_battles()
    assert len(battles) == 10 and sum(len(s) for s in battles) == 137
    sd = ref_port.init_mlp_params(tables.n_channels, 3, 64, seed=5)
    cfg = SurrogateRewardConfig(0.25, 0.25, False, True)

    def feed_all(processors, on_choice):
        """interleave the battles chunk by chunk, as N simulators running side by side would deliver them"""
        for step in range(max(len(s) for s in battles)):
            for bi, steps in enumerate(battles):
                if step >= len(steps):
                    continue
                update, reqs = steps[step]
                for side in ("p1", "p2"):
                    slot = 2 * bi + (side == "p2")
                    if side in reqs:
                        c = processors[slot].process_chunk("sideupdate", "|request|" + json.dumps(reqs[side]))
                        assert c is None
                    on_choice(slot, processors[slot].process_chunk("update", _extract_update_for_side(side, update)))
            yield step

    # (a) the reference alone: its parser, RLPolicy, AgentVal, FeatureExtractor(3), MLPModel -- one decision at a time
    fe = RefFE(3)
    model = RefMLP(fe.input_shape, fe.output_dim, n_layers=3, n_channels=64, bn=False)
    model.load_state_dict(sd)
    model.eval()
    ref_choices = {}
    procs = []
    for bi in range(len(battles)):
        for side in ("p1", "p2"):
            p = BattleStreamProcessor()
            p.set_policy(RefRLPolicy(RefAgentVal(model, fe), cfg))
            p.start_battle(side, None)
            procs.append(p)
    with torch.no_grad():
        for _ in feed_all(procs, lambda slot, c: ref_choices.setdefault(slot, []).append(c) if c is not None else None):
            pass
    n_ref = sum(len(v) for v in ref_choices.values())
    assert n_ref == 231  # the decisions the log holds (220 through the model + 11 single-choice)

    # (b) the same chunks through the same parser, decided in lock step by the driver
    procs2 = [BattleStreamProcessor() for _ in range(2 * len(battles))]
    agent = OracleAgent(3, sd, tables)
    drv = LockStepDriver(agent, procs2, surrogate_reward_config=cfg, tables=tables, use_pipeline=False)
    for bi in range(len(battles)):
        for j, side in enumerate(("p1", "p2")):
            procs2[2 * bi + j].start_battle(side, None)
    got = {}
    waiting = []

    def on_choice(slot, c):
        if c == LockStepDriver.PENDING:
            waiting.append(slot)
        elif c is not None:
            got.setdefault(slot, []).append(c)

    for _ in feed_all(procs2, on_choice):
        for slot, key in drv.flush().items():   # ONE launch for all the battles that asked in this step
            got.setdefault(slot, []).append(key)
        waiting.clear()
    assert got == ref_choices
    assert agent.calls <= 23 and max(agent.batch_sizes) > 10   # batched: far fewer launches than decisions
    for slot in range(len(procs2)):
        assert drv.policies[slot].inner.last_reward_potential == procs[slot].policy.last_reward_potential


# ---- GPU ----------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_lockstep_64_battles_on_the_kernels(tables):
    import torch
    from pokeai_b200.agent import AgentVal
    from pokeai_b200.feature_extractor import FeatureExtractor
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    from pokeai_b200.mlp_model import MLPModel
    from pokeai_b200.protocol import PackedRLPolicy
    dev = torch.device("cuda:0")
    fe = FeatureExtractor(6, device=dev)
    torch.manual_seed(0)
    model = MLPModel(fe.input_shape, fe.output_dim).eval()
    agent = AgentVal(model, fe)
    battles = [SyntheticBattle(seed=100 + i, party_size=6, tables=tables) for i in range(64)]
    for b in battles:
        b.keep_history = True
    drv = LockStepDriver(agent, battles, tables=tables)
    assert drv._pipe is not None  # pinned records, read in place by the kernel
    rounds = 10
    assert sum(drv.step() for _ in range(rounds)) == 64 * rounds
    assert drv.rounds == rounds
    drv.close()
    from collections import namedtuple
    Cfg = namedtuple("Cfg", "hp_ratio alive_ratio only_opponent offset_at_end")
    solo = PackedRLPolicy(AgentVal(model, fe), Cfg(0.0, 0.0, False, False), tables)
    for b in battles:
        for (bs, request), key in zip(b.history, b.choices):
            fn = solo.choice_force_switch if request.get("forceSwitch") else solo.choice_turn_start
            assert fn(bs, request) == key


@pytest.mark.gpu
@pytest.mark.parametrize("nl,H", [(2, 64), (2, 256)])
def test_lockstep_experience_collection_equals_one_at_a_time(tables, nl, H):
    """training agents (replay append + epsilon draw per battle, greedy decisions batched) against the same agents
    stepped one decision at a time in the same order; small net and a wide (tensor-core path) net"""
    import torch
    from pokeai_b200.lockstep import LockStepDriver, SyntheticBattle
    from pokeai_b200.protocol import PackedRLPolicy
    from pokeai_b200.trainer import Trainer
    from collections import namedtuple
    Cfg = namedtuple("Cfg", "hp_ratio alive_ratio only_opponent offset_at_end")
    cfg = Cfg(0.25, 0.25, False, True)
    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    tr = Trainer({"n_layers": nl, "n_channels": H, "bn": False}, {"epsilon": 0.3}, {"party_size": 3}, device=dev)
    n, rounds = 12, 9
    np.random.seed(5)
    agents = tr.get_train_agents(n)
    battles = [SyntheticBattle(seed=300 + i, party_size=3, tables=tables) for i in range(n)]
    for b in battles:
        b.keep_history = True
    drv = LockStepDriver(agents, battles, surrogate_reward_config=cfg, tables=tables)
    for _ in range(rounds):
        drv.step()
    for i, a in enumerate(agents):
        drv.policies[i].game_end(1.0 if i % 2 else -1.0)
    drv.close()
    # the same, one act() at a time in the same order (so the global numpy RNG makes the same draws)
    np.random.seed(5)
    agents2 = tr.get_train_agents(n)
    pols = [PackedRLPolicy(a, cfg, tables) for a in agents2]
    for p in pols:
        p.game_start()
    for r in range(rounds):
        for i, b in enumerate(battles):
            bs, request = b.history[r]
            fn = pols[i].choice_force_switch if request.get("forceSwitch") else pols[i].choice_turn_start
            assert fn(bs, request) == b.choices[r], (r, i)
    for i, p in enumerate(pols):
        p.game_end(1.0 if i % 2 else -1.0)
    for a, a2 in zip(agents, agents2):
        b1, b2 = a._replay_buffer.buffer, a2._replay_buffer.buffer
        assert len(b1) == len(b2) > 0
        for x, y in zip(b1, b2):
            assert x.action == y.action and x.reward == pytest.approx(y.reward)
            assert x.state.tobytes() == y.state.tobytes()
            assert (x.next_state is None) == (y.next_state is None)
            if x.next_state is not None:
                assert x.next_state.tobytes() == y.next_state.tobytes()
