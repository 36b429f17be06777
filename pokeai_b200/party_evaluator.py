"""PartyEvaluator / PartyEvaluatorQuick on the fused decision kernel.

The reference evaluates a party by the Q values of its lead against an opponent species
(pokeai/ai/party_feature/party_evaluator.py:17-105) and speeds that up by synthesising the dense feature batch
by hand (party_evaluator_quick.py:30-104: cached state vector per opponent + one choice block per party, then
`model(torch.from_numpy(batch))` at batch ~129).  The observation it builds is fully determined by
(party, opponent species): both sides just switched in at 100/100 HP, no status, no boosts, no weather,
remaining = total = party_size, the lead's moves in order followed by switches to the other members
(party_evaluator.py:40-72).  That is exactly one 144-byte packed record, so here the whole evaluation is
"pack -> pk_decide (features never materialised) -> max over legal slots": no dense tensor on either side of the
PCIe bus.  Same method names and results as the reference classes.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

from .packed import FLAG_SWITCH, PK_OBS_DTYPE
from .tables import load_tables


class PartyEvaluatorQuick:
    def __init__(self, agent, party_size: int, device: Optional[str] = None):
        self.agent = agent
        self.party_size = party_size
        self.n_choices = 4 + party_size - 1  # party_evaluator.py:26
        self.tables = load_tables()
        if agent._feature_extractor.output_dim != self.n_choices:
            raise ValueError("the agent's feature extractor was built for another party size")

    # -- (party, opponents) -> packed records (party_evaluator.py:40-72) ---------------------------------
    def pack(self, party: Sequence[dict], opponent_pokes: Sequence[str]) -> np.ndarray:
        T = self.tables
        if len(party) != self.party_size:
            raise AssertionError("party size mismatch")  # party_evaluator.py:41
        n = len(opponent_pokes)
        one = np.zeros(1, dtype=PK_OBS_DTYPE)
        r = one[0]
        r["remaining"] = (self.party_size, self.party_size)  # :44-46
        r["total"] = (self.party_size, self.party_size)
        r["hp_cur"] = (100, 100)  # switched in at "100/100" (:48-50)
        r["hp_max"] = (100, 100)
        lead = party[0]
        acts = []
        for mv in lead["moves"]:  # :53-61 one action per move of the lead
            acts.append((0, lead["species"], lead["item"], [mv]))
        for member in party[1:]:  # :62-70 switches carry all the target's moves
            acts.append((FLAG_SWITCH, member["species"], member["item"], list(member["moves"])))
        if len(acts) > self.n_choices:
            raise ValueError(f"{len(acts)} possible actions exceed output_dim {self.n_choices}")
        r["n_act"] = len(acts)
        for i, (flags, species, item, moves) in enumerate(acts):
            ar = r["act"][i]
            ar["flags"] = flags
            ar["species"] = T.species_id_to_idx[species]
            ar["item"] = T.item_to_idx[item]
            ar["n_moves"] = len(moves)
            for j, mv in enumerate(moves):
                ar["move"][j] = T.move_to_idx[mv]
        rec = np.repeat(one, n)
        rec["opp_species"] = [T.species_id_to_idx[p] for p in opponent_pokes]  # KeyError for unknown ids
        return rec

    def _q(self, rec: np.ndarray) -> np.ndarray:
        q, _ = self.agent.q_and_action_batch(rec)
        q = np.array(q, dtype=np.float32, copy=True)
        legal = np.arange(self.n_choices)[None, :] < rec["n_act"][:, None].astype(np.int64)
        q[~legal] = -np.inf  # q_vector[action_mask == 0] = -np.inf
        return q

    # -- reference surface --------------------------------------------------------------------------------------
    def calc_q_func(self, party: Sequence[dict], opponent_poke: str) -> np.ndarray:
        """Q of every choice of `party`'s lead against `opponent_poke`; -inf on illegal slots."""
        return self._q(self.pack(party, [opponent_poke]))[0]

    def gather_best_q(self, party: Sequence[dict], opponent_pokes: List[str]) -> np.ndarray:
        """max over legal choices, one value per opponent (party_evaluator_quick.py:97-104)."""
        return self._q(self.pack(party, opponent_pokes)).max(axis=1)

    # -- batched over parties: what hillclimb_party.py loops over ------------------------------------------------
    def gather_best_q_many(self, parties: Sequence[Sequence[dict]], opponent_pokes: List[str]) -> np.ndarray:
        """(len(parties), len(opponent_pokes)) best Q: all parties x opponents in ONE kernel launch."""
        if len(parties) == 0:
            return np.zeros((0, len(opponent_pokes)), dtype=np.float32)
        rec = np.concatenate([self.pack(p, opponent_pokes) for p in parties])
        return self._q(rec).max(axis=1).reshape(len(parties), len(opponent_pokes))


PartyEvaluator = PartyEvaluatorQuick  # the slow and the quick class of the reference give the same numbers
