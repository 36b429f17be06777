"""ORACLE (test infrastructure, not product code) -- object-level CPU restatement of
the reference's decision path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg may import this module.  The product path (pokeai_b200/) never does.

Parity pinning: PINNED.  tools/make_golden.py runs the *real* reference
(/root/reference, imported read-only in the build container) on the recorded battle
log fixture and on seeded synthetic observations and commits its outputs under
tests/golden/; tests/test_oracle.py checks this port against those vectors bit for
bit (features, masks, argmax actions) and within fp32 tolerance (Q values, update).
tests/test_oracle_vs_reference.py additionally compares against the live reference
wherever it was staged into oracle/_ref (oracle/ref_stage.py; the build container and,
because the snapshot carries it, the GPU box).

Third-party arithmetic: like the reference, the MLP / Huber / autograd / Adam are
torch CPU fp32 (reference pins torch 2.6.0, poetry.lock:3316; this image has 2.11).

Each function cites the reference lines it restates.  The code is deliberately
written the way the reference executes (a Python loop over observations, string
lookups, numpy scatter), because it is also the timed CPU baseline.
"""
from __future__ import annotations

import math
import random
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from pokeai_b200.tables import (NV_CONDITIONS, N_STATE, POKE_TYPES, RANKS, WEATHERS, Tables, load_tables)


# ---------------------------------------------------------------------------------------
# a3: StateFeatureExtractor.transform -- pokeai/ai/state_feature_extractor.py:92-197
# ---------------------------------------------------------------------------------------
def state_features(obs, tables: Tables) -> np.ndarray:
    bs = obs.battle_status
    friend = bs.side_statuses[bs.side_friend]
    opponent = bs.side_statuses[bs.side_opponent]
    out = np.zeros((N_STATE,), dtype=np.float32)
    pos = 0
    # :113-123 remaining / total, python float (double) division stored into float32
    for side in (friend, opponent):
        out[pos] = side.remaining_pokes / side.total_pokes
        pos += 1
    # :125-138 opponent active's types, multi-hot over the 17-type list
    assert opponent.active is not None
    for t in tables.species_types[tables.species_name_to_idx[opponent.active.species]]:
        out[pos + t] = 1.0
    pos += len(POKE_TYPES)
    # :140-151 hp_current / hp_max
    for side in (friend, opponent):
        assert side.active is not None
        out[pos] = side.active.hp_current / side.active.hp_max
        pos += 1
    # :153-167 non-volatile condition one-hot ('' and 'fnt' match nothing)
    for side in (friend, opponent):
        for i, cond in enumerate(NV_CONDITIONS):
            if cond == side.active.status:
                out[pos + i] = 1.0
                break
        pos += len(NV_CONDITIONS)
    # :169-182 (rank + 6) / 12
    for side in (friend, opponent):
        for i, key in enumerate(RANKS):
            out[pos + i] = (side.active.ranks[key] + 6.0) / 12.0
        pos += len(RANKS)
    # :184-197 weather one-hot ('none' matches nothing)
    for i, w in enumerate(WEATHERS):
        if w == bs.weather:
            out[pos + i] = 1.0
            break
    pos += len(WEATHERS)
    assert pos == N_STATE
    return out


# ---------------------------------------------------------------------------------------
# a4: ChoiceToVec.transform -- pokeai/ai/generic_move_model/choice_to_vec.py:40-61
# ---------------------------------------------------------------------------------------
def choice_features(obs, tables: Tables) -> np.ndarray:
    acts = obs.possible_actions
    out = np.zeros((tables.n_choice, len(acts)), dtype=np.float32)
    for col, pa in enumerate(acts):
        if pa.force_switch:
            out[tables.CH_FORCE, col] = 1
        if pa.switch:
            out[tables.CH_SWITCH, col] = 1
        out[tables.CH_SPECIES + tables.species_id_to_idx[pa.poke], col] = 1
        if pa.switch:
            for mv in pa.allMoves:  # all of the switch target's moves
                out[tables.CH_MOVE + tables.move_to_idx[mv], col] = 1
        else:
            out[tables.CH_MOVE + tables.move_to_idx[pa.move], col] = 1
        out[tables.CH_ITEM + tables.item_to_idx[pa.item], col] = 1
    return out


# ---------------------------------------------------------------------------------------
# a5: FeatureExtractor.transform -- pokeai/ai/generic_move_model/feature_extractor.py:40-52
# ---------------------------------------------------------------------------------------
def output_dim(party_size: int) -> int:
    return 4 + party_size - 1  # feature_extractor.py:30-38


def transform(obs, party_size: int, tables: Optional[Tables] = None) -> Tuple[np.ndarray, np.ndarray]:
    tables = tables or load_tables()
    A = output_dim(party_size)
    feat = np.zeros((tables.n_channels, A), dtype=np.float32)
    st = state_features(obs, tables)
    feat[:N_STATE, :] = st[:, None]  # every column, legal or not (:48)
    n_act = len(obs.possible_actions)
    feat[N_STATE:, :n_act] = choice_features(obs, tables)  # (:49)
    mask = np.zeros((A,), dtype=np.int32)
    mask[:n_act] = 1
    return feat, mask


def transform_batch(obs_list: Sequence, party_size: int, tables: Optional[Tables] = None):
    tables = tables or load_tables()
    pairs = [transform(o, party_size, tables) for o in obs_list]
    return np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])


# ---------------------------------------------------------------------------------------
# a6: MLPModel -- pokeai/ai/generic_move_model/mlp_model.py:5-30 (bn=False path and eval-mode bn)
# ---------------------------------------------------------------------------------------
def init_mlp_params(in_channels: int, n_layers: int = 2, n_channels: int = 64, seed: int = 0) -> dict:
    """Random-init parameters with the reference's state_dict keys/shapes
    (Conv1d default init, same bounds as torch: U(-1/sqrt(fan_in), 1/sqrt(fan_in)))."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    c = in_channels
    for i in range(n_layers):
        bound = 1.0 / math.sqrt(c)
        sd[f"layers.{i}.weight"] = (torch.rand((n_channels, c, 1), generator=g) * 2 - 1) * bound
        sd[f"layers.{i}.bias"] = (torch.rand((n_channels,), generator=g) * 2 - 1) * bound
        c = n_channels
    bound = 1.0 / math.sqrt(c)
    sd["output.weight"] = (torch.rand((1, c, 1), generator=g) * 2 - 1) * bound
    sd["output.bias"] = (torch.rand((1,), generator=g) * 2 - 1) * bound
    return sd


def mlp_forward(sd: dict, x: torch.Tensor) -> torch.Tensor:
    """x (B, C, A) fp32 -> (B, A).  mlp_model.py:21-30."""
    h = x
    i = 0
    while f"layers.{i}.weight" in sd:
        h = F.relu(F.conv1d(h, sd[f"layers.{i}.weight"], sd.get(f"layers.{i}.bias")))
        i += 1
    h = F.conv1d(h, sd["output.weight"], sd["output.bias"])
    return h.view(h.shape[0], -1)


# ---------------------------------------------------------------------------------------
# a7: action selection -- agent.py:32-45, agent_train.py:30-33
# ---------------------------------------------------------------------------------------
def masked_argmax(q: np.ndarray, mask: np.ndarray) -> np.ndarray:
    """q (B, A), mask (B, A) -> int32 (B,).  q[mask==0] = -inf; np.argmax (first max wins)."""
    qq = np.array(q, dtype=np.float32, copy=True)
    qq[mask == 0] = -np.inf
    return np.argmax(qq, axis=1).astype(np.int32)


def epsilon_greedy(q: np.ndarray, mask: np.ndarray, eps: float, u_explore: np.ndarray,
                   u_pick: np.ndarray) -> np.ndarray:
    """Batched restatement of AgentTrain.act's branch (agent_train.py:30-33) with the two
    uniforms supplied by the caller instead of numpy's global RNG:
      explore  <=> u_explore < eps                      (np.random.random() < eps)
      pick     =   flatnonzero(mask)[floor(u_pick * n)] (np.random.choice(actions), agent.py:43-45)
    """
    greedy = masked_argmax(q, mask)
    out = greedy.copy()
    for b in range(q.shape[0]):
        if u_explore[b] < eps:
            legal = np.flatnonzero(mask[b])
            if len(legal):
                k = min(int(np.float32(u_pick[b]) * np.float32(len(legal))), len(legal) - 1)
                out[b] = legal[k]
    return out


def softmax_sample(q: np.ndarray, mask: np.ndarray, temperature: float, u: np.ndarray):
    """NO REFERENCE COUNTERPART (the reference is a Q-network with argmax).
    North-star extension: masked softmax over legal slots and inverse-CDF sampling.
    p_k = exp((q_k - max_legal q) / T) / sum ; action = first legal k with cdf_k > u * sum.
    Returns (actions int32 (B,), probs float32 (B, A))."""
    B, A = q.shape
    acts = np.zeros((B,), dtype=np.int32)
    probs = np.zeros((B, A), dtype=np.float32)
    for b in range(B):
        legal = np.flatnonzero(mask[b])
        if len(legal) == 0:
            continue
        z = (q[b, legal].astype(np.float32) - np.max(q[b, legal])) / np.float32(temperature)
        e = np.exp(z.astype(np.float32)).astype(np.float32)
        total = np.float32(0)
        for v in e:  # sequential fp32 sum, slot order (the kernel sums in the same order)
            total = np.float32(total + v)
        probs[b, legal] = e / total
        thr = np.float32(u[b]) * total
        run = np.float32(0)
        pick = legal[-1]
        for k, v in zip(legal, e):
            run = np.float32(run + v)
            if run > thr:
                pick = k
                break
        acts[b] = pick
    return acts, probs


# ---------------------------------------------------------------------------------------
# a8: Trainer._update -- pokeai/ai/generic_move_model/trainer.py:152-195
# ---------------------------------------------------------------------------------------
def dqn_update_grads(sd: dict, target_sd: dict, state: np.ndarray, action: np.ndarray, reward: np.ndarray,
                     next_state: np.ndarray, next_mask: np.ndarray, non_final: np.ndarray, gamma: float,
                     double_dqn: bool = True, global_batch: Optional[int] = None):
    """One Double-DQN loss/backward on explicit arrays.

    state (N,C,A) f32, action (N,) int, reward (N,) f32, next_state (N,C,A) f32 and
    next_mask (N,A) for *all* N (rows of terminal transitions are ignored),
    non_final (N,) bool.  Returns (loss, grads dict *before* the [-1,1] clamp).
    ``global_batch``: divide the Huber sum by this instead of N (data-parallel shard,
    SURVEY.md §8 e); default = N which is F.smooth_l1_loss's mean (trainer.py:187).
    """
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    N = state.shape[0]
    nf = torch.from_numpy(np.asarray(non_final, dtype=bool))
    s = torch.from_numpy(state)
    a = torch.from_numpy(np.asarray(action, dtype=np.int64)[:, None])
    r = torch.from_numpy(np.asarray(reward, dtype=np.float32))
    q_sa = mlp_forward(params, s).gather(1, a)  # :163
    v_next = torch.zeros(N)
    if bool(nf.any()):
        ns = torch.from_numpy(next_state[non_final])
        bias = torch.from_numpy(np.where(next_mask[non_final], 0.0, -np.inf).astype(np.float32))  # :169-170
        with torch.no_grad():
            if double_dqn:  # :171-177
                a_star = (mlp_forward(sd, ns) + bias).max(1)[1]
                v = mlp_forward(target_sd, ns).gather(1, a_star.unsqueeze(1)).squeeze(1)
            else:  # :178-182
                v = (mlp_forward(target_sd, ns) + bias).max(1)[0]
        v_next[nf] = v
    y = v_next * gamma + r  # :184
    if global_batch is None:
        loss = F.smooth_l1_loss(q_sa, y.unsqueeze(1))  # :187  (beta=1, mean)
    else:
        loss = F.smooth_l1_loss(q_sa, y.unsqueeze(1), reduction="sum") / float(global_batch)
    loss.backward()  # :192
    grads = {k: p.grad.detach().clone() for k, p in params.items()}
    return float(loss.detach()), grads


def clamp_adam_step(sd: dict, grads: dict, adam_state: dict, lr: float, step: int, betas=(0.9, 0.999),
                    eps: float = 1e-8) -> None:
    """trainer.py:193-195: clamp every grad element to [-1,1] then torch.optim.Adam defaults
    (no weight decay, no amsgrad).  ``adam_state[k] = (exp_avg, exp_avg_sq)``; ``step`` is 1-based."""
    b1, b2 = betas
    for k, p in sd.items():
        g = grads[k].clamp(-1, 1)
        m, v = adam_state.setdefault(k, (torch.zeros_like(p), torch.zeros_like(p)))
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1 = 1 - b1 ** step
        bc2 = 1 - b2 ** step
        denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
        p.addcdiv_(m, denom, value=-(lr / bc1))


# ---------------------------------------------------------------------------------------
# a6 / a8 with bn=True: MLPModel.forward in TRAINING mode (mlp_model.py:12-15,21-30) under Trainer._update.
# The reference's Trainer never calls .eval() on model / target_model (trainer.py:41-43), so every forward of the update
# normalises with the statistics of its own batch and moves the running statistics.  BatchNorm1d is written out here
# (torch/nn/modules/batchnorm.py + F.batch_norm semantics): per channel over (batch, action columns), biased variance
# for the normalisation, eps 1e-5; running_mean / running_var move by momentum 0.1 towards the batch mean / the
# UNBIASED batch variance; num_batches_tracked += 1.
# ---------------------------------------------------------------------------------------
def mlp_forward_bn_train(sd: dict, x: torch.Tensor, buffers: Optional[dict] = None, momentum: float = 0.1,
                         eps: float = 1e-5) -> torch.Tensor:
    """sd: parameters by state_dict key (layers.i.weight (H, C, 1) without bias, bn_layers.i.weight / .bias, output.*);
    `buffers` (state_dict keys bn_layers.i.running_mean / running_var / num_batches_tracked) is updated in place."""
    h = x
    i = 0
    while f"layers.{i}.weight" in sd:
        z = torch.einsum("oc,bca->boa", sd[f"layers.{i}.weight"][:, :, 0], h)  # mlp_model.py:23, bias=False (:12)
        n = z.shape[0] * z.shape[2]
        mean = z.mean(dim=(0, 2))
        var = ((z - mean[None, :, None]) ** 2).mean(dim=(0, 2))  # biased
        if buffers is not None:
            with torch.no_grad():
                rm, rv = buffers[f"bn_layers.{i}.running_mean"], buffers[f"bn_layers.{i}.running_var"]
                rm.mul_(1 - momentum).add_(mean.detach().to(rm.dtype), alpha=momentum)
                rv.mul_(1 - momentum).add_((var.detach() * (n / max(n - 1, 1))).to(rv.dtype), alpha=momentum)
                buffers[f"bn_layers.{i}.num_batches_tracked"] += 1
        xhat = (z - mean[None, :, None]) / torch.sqrt(var[None, :, None] + eps)
        h = torch.relu(xhat * sd[f"bn_layers.{i}.weight"][None, :, None] + sd[f"bn_layers.{i}.bias"][None, :, None])  # :24-27
        i += 1
    out = torch.einsum("oc,bca->boa", sd["output.weight"][:, :, 0], h) + sd["output.bias"][None, :, None]  # :28
    return out.reshape(out.shape[0], -1)  # :29


def dqn_update_grads_bn(sd: dict, target_sd: dict, state: np.ndarray, action: np.ndarray, reward: np.ndarray,
                        next_state: np.ndarray, next_mask: np.ndarray, non_final: np.ndarray, gamma: float,
                        double_dqn: bool = True, dtype=torch.float32):
    """dqn_update_grads for a bn=True network.  sd / target_sd: FULL state_dicts (parameters and BatchNorm buffers);
    their buffers move as the reference's do: the online network's once for model(state_batch) (trainer.py:163) and once
    more for model(non_final_next_states) under Double DQN (:174), the target's once (:175 / :182).
    Returns (loss, grads of the parameters before the clamp)."""
    is_param = lambda k: not (k.endswith("running_mean") or k.endswith("running_var") or k.endswith("num_batches_tracked"))
    params = {k: v.detach().clone().to(dtype).requires_grad_(True) for k, v in sd.items() if is_param(k)}
    tparams = {k: v.detach().to(dtype) for k, v in target_sd.items() if is_param(k)}
    nf = np.asarray(non_final, dtype=bool)
    s = torch.from_numpy(state).to(dtype)
    a = torch.from_numpy(np.asarray(action, dtype=np.int64)[:, None])
    r = torch.from_numpy(np.asarray(reward, dtype=np.float32)).to(dtype)
    q_sa = mlp_forward_bn_train(params, s, sd).gather(1, a)  # :163
    ns = torch.from_numpy(next_state[nf]).to(dtype)  # :157-158: np.stack([]) raises when every transition is terminal
    if ns.shape[0] == 0:
        raise ValueError("need at least one array to stack")
    bias = torch.from_numpy(np.where(next_mask[nf], 0.0, -np.inf)).to(dtype)  # :169-170
    v_next = torch.zeros(state.shape[0], dtype=dtype)
    with torch.no_grad():
        if double_dqn:  # :171-177
            a_star = (mlp_forward_bn_train({k: p.detach() for k, p in params.items()}, ns, sd) + bias).max(1)[1]
            v = mlp_forward_bn_train(tparams, ns, target_sd).gather(1, a_star.unsqueeze(1)).squeeze(1)
        else:  # :178-182
            v = (mlp_forward_bn_train(tparams, ns, target_sd) + bias).max(1)[0]
    v_next[torch.from_numpy(nf)] = v
    y = v_next * gamma + r  # :184
    loss = F.smooth_l1_loss(q_sa, y.unsqueeze(1))  # :187
    loss.backward()  # :192
    return float(loss.detach()), {k: p.grad.detach().clone() for k, p in params.items()}
