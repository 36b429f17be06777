"""Lock-step battle driver: N concurrent battles, ONE decision launch per round.

The reference steps one battle per process (`Sim.run`, pokeai/sim/sim.py:54-99: read a chunk from the simulator, hand
it to each side's `BattleStreamProcessor.process_chunk`, write the returned choice back) and `rl_train.train_episode`
(pokeai/ai/generic_move_model/rl_train.py:51-68) runs one episode at a time, so the decision path sees batch 1.  Here
a driver owns the processors of many battles.  Each processor gets an `ActionPolicy` (action_policy.py:7-44) whose
`choice_turn_start` / `choice_force_switch` do what `RLPolicy._choice_by_model` does (rl_policy.py:76-105) up to the
model call -- surrogate-reward potential, request walked straight into a packed record in PINNED host memory
(protocol.pack_request_into), single-choice bypass (rl_policy.py:87-91) -- and then defer: the processor returns the
sentinel `PENDING`, the driver collects every waiting battle, runs ONE `pk_decide` launch over the pinned records
(read by the kernel over PCIe, `DecisionPipeline(..., pinned=True)`), and hands the simulator keys back in order.

    drv = LockStepDriver(agent, processors)             # processors: BattleStreamProcessor-like, one per battle side
    for i, (chunk_type, data) in incoming:              # chunks as they arrive from the N simulators
        choice = drv.feed(i, chunk_type, data)          # None, a key ("move 2"), or LockStepDriver.PENDING
    for i, key in drv.flush().items(): ...              # one launch; write f">{side} {key}" to simulator i

`agent` is an `AgentVal` shared by all battles (greedy, rating battles: rl_rating_battle.py:199), or a list of
`AgentTrain` -- one per processor, sharing ONE model snapshot (`Trainer.get_train_agents`) -- for experience
collection: the replay append and the exploration draw stay per battle (`AgentTrain.begin_act`), only the greedy
decisions are batched.  The protocol parser stays the reference's (`pokeai/sim/battle_stream_processor.py`); anything
with `set_policy(policy)` and `process_chunk(chunk_type, data) -> Optional[str]` works, e.g. `SyntheticBattle` below,
which stands in for simulator + parser where Node is absent (tests on the GPU box, bench.py's configs[0] leg).
"""
from __future__ import annotations

import random
import time
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

from .observation import ActivePokeStatus, BattleStatus, RANK_KEYS
from .packed import PK_OBS_BYTES, PK_OBS_DTYPE
from .protocol import PackedRLPolicy, pack_request_into
from .tables import NV_CONDITIONS, WEATHERS, Tables, load_tables


class _NoSurrogate:
    """surrogate_reward_config.py:4-15 with every term off"""
    hp_ratio = 0.0
    alive_ratio = 0.0
    only_opponent = False
    offset_at_end = False


class _SlotPolicy:
    """The ActionPolicy (action_policy.py:7-44) a processor of the driver holds."""

    train = False

    def __init__(self, driver: "LockStepDriver", slot: int, inner: PackedRLPolicy):
        self.driver, self.slot, self.inner = driver, slot, inner

    def game_start(self) -> None:
        self.inner.game_start()

    def choice_turn_start(self, battle_status, request: dict) -> str:
        return self.driver._request(self.slot, battle_status, request)

    choice_force_switch = choice_turn_start

    def game_end(self, reward: float) -> None:
        self.inner.game_end(reward)


class LockStepDriver:
    PENDING = "\x00pending"  # what a processor returns while its decision waits for the batched launch

    def __init__(self, agent, processors: Sequence, surrogate_reward_config=None, tables: Optional[Tables] = None,
                 use_pipeline: Optional[bool] = None, materialise_features: bool = False):
        self.tables = tables or load_tables()
        n = len(processors)
        agents = list(agent) if isinstance(agent, (list, tuple)) else [agent] * n
        if len(agents) != n:
            raise ValueError("one agent per processor (or one shared agent)")
        if any(a._model is not agents[0]._model for a in agents):
            raise ValueError("lock-step batching needs agents that share one model")
        self.processors = list(processors)
        cfg = surrogate_reward_config if surrogate_reward_config is not None else _NoSurrogate
        self.policies = [_SlotPolicy(self, i, PackedRLPolicy(a, cfg, self.tables)) for i, a in enumerate(agents)]
        for p, pol in zip(self.processors, self.policies):
            p.set_policy(pol)
        self._max_actions = agents[0]._feature_extractor.output_dim
        self._pipe = None
        if use_pipeline is None:
            import torch
            use_pipeline = torch.cuda.is_available() and not getattr(agents[0], "_is_wide", False) \
                and hasattr(agents[0]._feature_extractor, "session")
        if use_pipeline:
            from .session import DecisionPipeline, pinned_records
            from .wide import is_wide
            m = agents[0]._model
            if not (hasattr(m, "n_channels") and is_wide(m.n_channels)):
                self._pipe = DecisionPipeline(agents[0], max(n, 1), depth=1, materialise_features=materialise_features)
                self._rec = pinned_records(max(n, 1))
        if self._pipe is None:
            self._rec = np.zeros(max(n, 1), dtype=PK_OBS_DTYPE)
        self._buf = memoryview(self._rec.view(np.uint8).reshape(-1))  # the records' bytes: struct.pack_into target
        self._pending: List[Tuple[int, List[str]]] = []
        self.rounds = 0
        self.decisions = 0
        self.time_pack = 0.0
        self.time_gpu = 0.0

    # ---- what a processor's policy calls ------------------------------------------------------------------------
    def _request(self, slot: int, battle_status, request: dict) -> str:
        t0 = time.perf_counter()
        inner = self.policies[slot].inner
        potential = inner._calc_reward_potential(battle_status)      # rl_policy.py:83
        k = len(self._pending)
        if k >= len(self._rec):
            raise RuntimeError("more pending decisions than processors: flush() first")
        keys = pack_request_into(self._buf, k * PK_OBS_BYTES, battle_status, request, self.tables, self._max_actions)
        if len(keys) == 1:                                            # rl_policy.py:87-91: the model is not asked
            self.time_pack += time.perf_counter() - t0
            return keys[0]
        agent = inner.agent
        if hasattr(agent, "begin_act"):                               # AgentTrain: replay append + exploration draw
            a = agent.begin_act(self._rec[k:k + 1], inner._surrogate(potential))
            inner.last_reward_potential = potential
            if a is not None:
                self.time_pack += time.perf_counter() - t0
                return keys[agent.finish_act(a)]
        else:
            inner.last_reward_potential = potential
        self._pending.append((slot, keys))
        self.time_pack += time.perf_counter() - t0
        return self.PENDING

    # ---- the driver's own interface -------------------------------------------------------------------------------
    def feed(self, slot: int, chunk_type: str, data: str) -> Optional[str]:
        """Hand one simulator chunk to processor `slot` (battle_stream_processor.py:113-147).  Returns None (no choice
        asked), a simulator key (single legal action, or an exploring train agent), or PENDING."""
        return self.processors[slot].process_chunk(chunk_type, data)

    def flush(self) -> Dict[int, str]:
        """ONE decision launch for every waiting battle; {slot: simulator key} in the order the requests arrived."""
        k = len(self._pending)
        if k == 0:
            return {}
        t0 = time.perf_counter()
        if self._pipe is not None:
            self._pipe.submit(self._rec[:k], pinned=True)
            actions = self._pipe.result()
        else:
            actions = self.policies[self._pending[0][0]].inner.agent.act_batch(np.array(self._rec[:k]))
        self.time_gpu += time.perf_counter() - t0
        out: Dict[int, str] = {}
        for (slot, keys), a in zip(self._pending, actions):
            agent = self.policies[slot].inner.agent
            a = int(a)
            if hasattr(agent, "finish_act"):
                agent.finish_act(a)
            out[slot] = keys[a]
        self._pending = []
        self.rounds += 1
        self.decisions += k
        return out

    def step(self, chunks: Optional[Sequence[Tuple[int, str, str]]] = None) -> int:
        """One lock-step round: feed a chunk to every processor (default: an empty `update` to each, which is all a
        `SyntheticBattle` needs), flush, deliver.  Returns the number of decisions made (bypasses included)."""
        if chunks is None:
            chunks = [(i, "update", "") for i in range(len(self.processors))]
        made = 0
        waiting = []
        for slot, chunk_type, data in chunks:
            c = self.feed(slot, chunk_type, data)
            if c == self.PENDING:
                waiting.append(slot)
            elif c is not None:
                made += 1
                self._deliver(slot, c)
        for slot, key in self.flush().items():
            made += 1
            self._deliver(slot, key)
        return made

    def _deliver(self, slot: int, key: str) -> None:
        p = self.processors[slot]
        if hasattr(p, "deliver"):
            p.deliver(key)

    def close(self) -> None:
        if self._pipe is not None:
            self._pipe.close()
            self._pipe = None


# ---- a stand-in for simulator + protocol parser (no Node on the GPU box) ------------------------------------------
def synthetic_request(rng: random.Random, tables: Tables, party_size: int) -> dict:
    """A `|request|` object in the simulator's format (turn start, forced switch, trapped, disabled moves, fainted
    bench members, Hidden Power / Return / Frustration spellings), as tools/make_golden_protocol.py draws them."""
    names = tables.species_names
    move_ids = [m for m in tables.moves if m != "hiddenpower"]
    species = rng.sample(range(len(names)), party_size)
    kind = rng.choice(["turn", "turn", "turn", "force", "trapped"]) if party_size > 1 else "turn"
    members = []
    for i, sp in enumerate(species):
        moves = rng.sample(move_ids, rng.randint(1, 4))
        hp_max = rng.randint(100, 400)
        fainted = (i > 1 and rng.random() < 0.25) or (kind == "force" and i == 0)  # member 1 stays able to switch in
        cond = "0 fnt" if fainted else f"{rng.randint(1, hp_max)}/{hp_max}" + rng.choice(["", "", " par", " slp"])
        members.append({"ident": f"p1: {names[sp]}", "details": f"{names[sp]}, L{rng.randint(50, 55)}" +
                        rng.choice(["", ", M", ", F"]), "condition": cond, "active": i == 0,
                        "moves": ["return102" if m == "return" else "frustration1" if m == "frustration" else m
                                  for m in moves], "item": rng.choice(tables.items), "_moves": moves})
    req = {"side": {"name": "p1", "id": "p1", "pokemon": members}, "rqid": rng.randint(1, 99)}
    if kind == "force":
        req["forceSwitch"] = [True]
    else:
        am = []
        for m in members[0]["_moves"]:
            if m.startswith("hiddenpower"):
                am.append({"move": f"Hidden Power {m[len('hiddenpower'):].capitalize()} 70", "id": "hiddenpower",
                           "disabled": False})
            else:
                am.append({"move": m.capitalize(), "id": m, "disabled": False})
        for m in am:
            if len(am) > 1 and rng.random() < 0.15:
                m["disabled"] = True
        if all(m["disabled"] for m in am):
            am[0]["disabled"] = False
        act = {"moves": am}
        if kind == "trapped":
            act["trapped"] = True
        req["active"] = [act]
    for m in members:
        del m["_moves"]
    return req


def synthetic_battle_status(rng: random.Random, tables: Tables, party_size: int) -> BattleStatus:
    """SURVEY.md 8d's value distribution as a BattleStatus object."""
    bs = BattleStatus("p1", None)
    for side in ("p1", "p2"):
        ss = bs.side_statuses[side]
        ss.total_pokes = party_size
        ss.remaining_pokes = rng.randint(1, party_size)
        name = tables.species_names[rng.randrange(len(tables.species_names))]
        hp_max = rng.randint(100, 400)
        status = "" if rng.random() < 1 / 3 else rng.choice(NV_CONDITIONS)
        ap = ActivePokeStatus(f"{side}a: {name}", name, 55, "M", rng.randint(1, hp_max), hp_max, status)
        ap.ranks = {k: rng.randint(-6, 6) for k in RANK_KEYS}
        ss.active = ap
    bs.weather = rng.choice([BattleStatus.WEATHER_NONE] + list(WEATHERS))
    return bs


class SyntheticBattle:
    """Processor-like source of decisions: every `update` chunk produces one seeded (battle_status, request) pair and
    asks the policy, as `BattleStreamProcessor.process_chunk` does at battle_stream_processor.py:134-138."""

    def __init__(self, seed: int, party_size: int = 6, tables: Optional[Tables] = None):
        self.rng = random.Random(seed)
        self.party_size = party_size
        self.tables = tables or load_tables()
        self.policy = None
        self.history: List[Tuple[BattleStatus, dict]] = []
        self.choices: List[str] = []
        self.keep_history = False
        self._queue: List[Tuple[BattleStatus, dict]] = []

    def pregenerate(self, n: int) -> None:
        """draw the next `n` (battle_status, request) pairs now, so that a timed loop measures the decision path and not
        this stand-in for the simulator"""
        self._queue.extend(self._draw() for _ in range(n))

    def set_policy(self, policy) -> None:
        self.policy = policy

    def _draw(self) -> Tuple[BattleStatus, dict]:
        return (synthetic_battle_status(self.rng, self.tables, self.party_size),
                synthetic_request(self.rng, self.tables, self.party_size))

    def next_pair(self) -> Tuple[BattleStatus, dict]:
        return self._queue.pop(0) if self._queue else self._draw()

    def process_chunk(self, chunk_type: str, data: str) -> Optional[str]:
        if chunk_type != "update":
            return None
        bs, request = self.next_pair()
        if self.keep_history:
            self.history.append((bs, request))
        if request.get("forceSwitch"):
            return self.policy.choice_force_switch(bs, request)
        return self.policy.choice_turn_start(bs, request)

    def deliver(self, key: str) -> None:
        if self.keep_history:
            self.choices.append(key)
