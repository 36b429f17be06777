"""Agent / AgentVal / AgentTrain with the reference's interface
(pokeai/ai/generic_move_model/agent.py:13-45, agent_val.py:4-14, agent_train.py:9-44),
running on the fused B200 decision kernel.

``act(obs, reward) -> int`` is the call RLPolicy makes (rl_policy.py:95-105); ``act_batch`` is
the same decision for many battles in one launch.  The host side only packs the observation
into the pinned 144-byte record; encode, MLP and the masked choice happen in one kernel.
"""
from __future__ import annotations

import json
import logging
from logging import getLogger
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _native as N
from .feature_extractor import FeatureExtractor
from .packed import PK_OBS_DTYPE
from .replay_buffer import ReplayBuffer, ReplayBufferItem

logger = getLogger(__name__)


class Agent:
    def __init__(self, model: torch.nn.Module, feature_extractor: FeatureExtractor):
        self._model = model
        self._feature_extractor = feature_extractor

    def act(self, obs: object, reward: float) -> int:
        raise NotImplementedError

    def stop_episode(self, reward: float) -> None:
        raise NotImplementedError

    # -- device plumbing ---------------------------------------------------------------------------
    def _session(self, n: int):
        s = self._feature_extractor.session(n)
        m = self._model
        # cheap change detector (the full key lives in MLPModel.prepared): generation, parameter versions, storage
        # pointers -- and which agent pointed the (shared) session last: two agents whose models have no `.prepared`
        # (a reference nn.Module) would otherwise both see "nothing changed" while the session holds the other's weights
        plist = self.__dict__.get("_plist")
        if plist is None:
            plist = self._plist = list(m.parameters()) + list(m.buffers())
        fast = (id(s), getattr(m, "_generation", 0)) + tuple(p._version for p in plist) + tuple(p.data_ptr() for p in plist)
        if self.__dict__.get("_fast_key") == fast and getattr(s, "_weights_owner", None) is m:
            return s  # same parameters, and no other model has been loaded into the shared session since
        if hasattr(m, "prepared"):
            prep = m.prepared(s.device)  # rebuilt (a NEW tensor) only when a parameter changed
            if getattr(s, "_prepared_ref", None) is not prep:  # (re)pointing synchronises: do it only on a change
                s.use_prepared(prep, m.n_layers, m.n_channels)
        else:
            s.load_model(m)
        s._weights_owner = m
        self._fast_key = fast
        return s

    # -- dense-input Q (agent.py:23-30; used by party_evaluator.py:102, party_evaluator_quick.py:101) ---
    def _calc_q_vector(self, obs_vector) -> np.ndarray:
        return self._calc_q_vector_batch(obs_vector[np.newaxis, ...])[0]

    def _calc_q_vector_batch(self, obs_vector_batch) -> np.ndarray:
        s = self._feature_extractor.session(1)
        x = torch.from_numpy(np.ascontiguousarray(obs_vector_batch, dtype=np.float32)).to(s.device)
        q = self._model(x)
        return q.cpu().numpy()

    def _act_by_model(self, obs_vector, action_mask) -> int:
        q_vector = self._calc_q_vector(obs_vector)
        q_vector[action_mask == 0] = -np.inf
        action = int(np.argmax(q_vector))
        if logger.isEnabledFor(logging.DEBUG):
            logger.debug("q_func: " + json.dumps({"q_func": q_vector.tolist(), "action": action}))
        return action

    def _act_random(self, obs_vector, action_mask) -> int:
        actions = np.flatnonzero(action_mask)
        return int(np.random.choice(actions))

    # -- fused path -----------------------------------------------------------------------------------
    def q_and_action_batch(self, obs_list, mode: int = N.PK_SELECT_ARGMAX, eps: float = 0.0,
                           temperature: float = 1.0, u_explore: Optional[np.ndarray] = None,
                           u_pick: Optional[np.ndarray] = None):
        """Observations (objects or packed records) -> (q (B, A) float32 with -inf NOT applied,
        action (B,) int32).  Views of pinned buffers, valid until the next call."""
        n = len(obs_list)
        m = self._model
        wide_net = self.__dict__.get("_is_wide")
        if wide_net is None:
            from .wide import is_wide
            wide_net = self._is_wide = bool(hasattr(m, "n_channels") and is_wide(m.n_channels))
        if wide_net:
            from .wide import WideRunner
            # wide network (e.g. 4 x 2048): tensor-core path (pokeai_b200/wide.py), same semantics
            from .ops import SelectSpec
            fe = self._feature_extractor
            if getattr(self, "_wide", None) is None or self._wide.model is not m:
                self._wide = WideRunner(m, fe.output_dim, fe.session(1).device)
            rec = fe.pack(obs_list)
            spec = None
            if mode != N.PK_SELECT_ARGMAX:
                dev = self._wide.device
                ue = torch.from_numpy(np.asarray(u_explore, dtype=np.float32)).to(dev) if u_explore is not None else None
                up = torch.from_numpy(np.asarray(u_pick, dtype=np.float32)).to(dev)
                spec = SelectSpec(mode, eps, temperature, ue, up)
            return self._wide.decide(rec, spec)
        s = self._session(n)
        direct = None
        if isinstance(obs_list, np.ndarray) and obs_list.dtype == PK_OBS_DTYPE and obs_list.flags.c_contiguous:
            direct = obs_list  # already packed: read in place (no 144 B x n host copy into the staging buffer)
        else:
            self._feature_extractor.pack(obs_list, out=s.packed)
        if mode == N.PK_SELECT_EPSGREEDY:
            s.u_explore[:n] = u_explore
        if mode != N.PK_SELECT_ARGMAX:
            s.u_pick[:n] = u_pick
        return s.decide(n, mode, eps, temperature, packed=direct)

    def act_batch(self, obs_list) -> np.ndarray:
        """Greedy decisions for a batch: q[mask == 0] = -inf; argmax (agent.py:32-35)."""
        return self.q_and_action_batch(obs_list)[1].copy()

    def _log_q(self, q: np.ndarray, n_act: int, action: int) -> None:
        if logger.isEnabledFor(logging.DEBUG):
            qq = q.astype(np.float32).copy()
            qq[n_act:] = -np.inf
            logger.debug("q_func: " + json.dumps({"q_func": qq.tolist(), "action": action}))


class AgentVal(Agent):
    def __init__(self, model, feature_extractor):
        super().__init__(model, feature_extractor)

    def act(self, obs: object, reward: float) -> int:
        # a (1,) packed record (protocol.pack_request) is taken as it is
        q, a = self.q_and_action_batch(obs if isinstance(obs, np.ndarray) else [obs])
        action = int(a[0])
        if logger.isEnabledFor(logging.DEBUG):
            n_act = int(obs["n_act"][0]) if isinstance(obs, np.ndarray) else len(obs.possible_actions)
            self._log_q(q[0], n_act, action)
        return action

    def stop_episode(self, reward: float) -> None:
        pass


class AgentTrain(Agent):
    def __init__(self, model, feature_extractor: FeatureExtractor, epsilon: float):
        super().__init__(model, feature_extractor)
        self._replay_buffer = ReplayBuffer(None)
        self._epsilon = epsilon
        self._last_state = None
        self._last_action_mask = None
        self._last_action = 0
        self.pending_record = None

    def begin_act(self, obs, reward: float) -> Optional[int]:
        """First half of `act` (agent_train.py:24-31): pack the observation, append the transition that ended with it
        to the replay buffer, and make the exploration draw.  Returns the random action when exploring, else None:
        the caller obtains the greedy action for `self.pending_record` from the model (alone, or batched with other
        battles by `LockStepDriver`) and passes it to `finish_act`.
        The two numpy global-RNG draws happen on the host in the reference's order (np.random.random(), then
        np.random.choice over the legal slots when exploring), so a seeded run makes the same explore/exploit choices
        as the reference."""
        rec = self._feature_extractor.pack(obs if isinstance(obs, np.ndarray) else [obs])
        packed = rec[0].copy()
        n_act = int(packed["n_act"])
        action_mask = np.zeros((self._feature_extractor.output_dim,), dtype=np.int32)
        action_mask[:n_act] = 1
        if self._last_state is not None:
            self._replay_buffer.append(
                ReplayBufferItem(self._last_state, self._last_action_mask, self._last_action, packed, action_mask,
                                 reward))
        self._last_state = packed
        self._last_action_mask = action_mask
        self.pending_record = packed
        if np.random.random() < self._epsilon:
            return self._act_random(None, action_mask)
        return None

    def finish_act(self, action: int) -> int:
        """Second half of `act` (agent_train.py:32-36): remember the action taken in the pending state."""
        self._last_action = int(action)
        self.pending_record = None
        return self._last_action

    def act(self, obs, reward: float) -> int:
        """agent_train.py:24-37.  The greedy branch goes through `q_and_action_batch`, i.e. the fused small-net kernel
        or, for a wide model, the tensor-core path."""
        action = self.begin_act(obs, reward)
        if action is None:
            rec = np.zeros(1, dtype=self.pending_record.dtype)
            rec[0] = self.pending_record
            q, a = self.q_and_action_batch(rec)
            action = int(a[0])
            self._log_q(q[0], int(rec["n_act"][0]), action)
        return self.finish_act(action)

    def stop_episode(self, reward: float) -> None:
        self._replay_buffer.append(
            ReplayBufferItem(self._last_state, self._last_action_mask, self._last_action, None, None, reward))
        self._last_state = None
        self._last_action_mask = None
        self._last_action = 0
