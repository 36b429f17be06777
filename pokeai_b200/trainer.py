"""Trainer: the reference's DQN trainer interface (pokeai/ai/generic_move_model/trainer.py:15-198) on the
B200 update kernels.

Same constructor, attributes, schedule (``train``), checkpoint dictionaries (``save_state`` / ``load_state``)
and agent factories as the reference.  What changes underneath:
  * parameters, gradients and Adam moments live in flat fp32 CUDA vectors (state_dict order); ``self.model`` /
    ``self.target_model`` are MLPModel modules whose parameters are views of those vectors, so
    ``state_dict()`` / ``load_state_dict()`` keep working;
  * ``_update`` ships the sampled packed transitions to the device and runs pk_dqn_grad -> [allreduce] ->
    pk_clamp_adam (Double-DQN target, Huber, backward, clamp, Adam: trainer.py:152-195);
  * with ``torch.distributed`` initialised (one process per GPU), every rank draws the same minibatch, takes its
    contiguous shard, and the flat gradient is summed with one allreduce BEFORE the clamp.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np
import torch

from . import _native as N
from .agent import AgentTrain, AgentVal
from .feature_extractor import FeatureExtractor
from .mlp_model import MLPModel
from .packed import PK_OBS_BYTES
from .replay_buffer import ReplayBuffer

DQN_DEFAULT_PARAMS = {  # trainer.py:15-27
    "epsilon": 0.3,
    "epsilon_decay": 0.0,
    "epsilon_min": 0.0,
    "gamma": 0.95,
    "batch_size": 32,
    "first_update_steps": 500,
    "optimize_per_steps": 1,
    "target_update": 100,
    "double_dqn": True,
    "replay_buffer_size": 100000,
    "lr": 1e-3,
}

_PARAM_KEYS_CACHE = {}


def _param_keys(n_layers: int):
    keys = []
    for i in range(n_layers):
        keys += [f"layers.{i}.weight", f"layers.{i}.bias"]
    return keys + ["output.weight", "output.bias"]


class FlatAdam:
    """Adam state of the flat parameter vector, exposing ``state_dict`` / ``load_state_dict`` in
    torch.optim.Adam's format (per-parameter ``step`` / ``exp_avg`` / ``exp_avg_sq``) so that resume
    checkpoints interchange with the reference's ``optim.Adam`` (trainer.py:54,84,114)."""

    def __init__(self, model: MLPModel, flat: torch.Tensor, lr: float):
        self.lr, self.betas, self.eps = float(lr), (0.9, 0.999), 1e-8
        self.step_count = 0
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self._shapes = [tuple(p.shape) for p in model.parameters()]

    def state_dict(self) -> dict:
        state = {}
        if self.step_count > 0:
            off = 0
            for i, shp in enumerate(self._shapes):
                n = int(np.prod(shp))
                state[i] = {"step": torch.tensor(float(self.step_count)),
                            "exp_avg": self.exp_avg[off:off + n].view(shp).cpu().clone(),
                            "exp_avg_sq": self.exp_avg_sq[off:off + n].view(shp).cpu().clone()}
                off += n
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": 0, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": False, "params": list(range(len(self._shapes)))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd: dict) -> None:
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps = float(g["lr"]), tuple(g["betas"]), float(g["eps"])
        if g.get("weight_decay", 0) or g.get("amsgrad", False):
            raise NotImplementedError("weight decay / amsgrad have no kernel (the reference uses neither)")
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        self.step_count = 0
        off = 0
        for i, shp in enumerate(self._shapes):
            n = int(np.prod(shp))
            st = sd["state"].get(i)
            if st is not None:
                self.exp_avg[off:off + n].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
                self.step_count = int(st["step"])
            off += n

    def zero_grad(self) -> None:  # kept for API familiarity; gradients are overwritten by every update
        pass


class Trainer:
    def __init__(self, model_params: dict, dqn_params: dict, feature_params: dict, device=None):
        self.constructor_params = {
            "model_params": model_params,
            "dqn_params": dqn_params,
            "feature_params": feature_params,
        }
        if not torch.cuda.is_available():
            raise RuntimeError("pokeai_b200.Trainer needs a CUDA device (there is no CPU training path)")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        N.ensure_init(self.device.index if self.device.index is not None else torch.cuda.current_device())
        self.feature_extractor = FeatureExtractor(**feature_params)
        self.feature_extractor.device = self.device
        self.model_params = model_params.copy()
        self.model_params["input_shape"] = self.feature_extractor.input_shape
        self.model_params["output_dim"] = self.feature_extractor.output_dim
        # bn: true -> train-mode BatchNorm in every forward of the update, as in the reference, whose Trainer never calls
        # .eval() on model / target_model (trainer.py:41-43): pokeai_b200/bn_train.py, csrc/pk_bn.cuh
        self._bn = bool(self.model_params.get("bn", False))
        # same construction order as the reference (two random inits, then copy): identical torch RNG use
        self.model, self._flat = self._construct_flat_model()
        self.target_model, self._flat_target = self._construct_flat_model()
        self._flat_target.copy_(self._flat)
        self.total_steps = 0
        self.total_battles = 0
        self.update_steps = 0
        p = DQN_DEFAULT_PARAMS.copy()
        p.update(dqn_params)
        self.optimizer = FlatAdam(self.model, self._flat, p["lr"])
        self.replay_buffer = ReplayBuffer(p["replay_buffer_size"], self.feature_extractor.output_dim)
        self.epsilon = p["epsilon"]
        self.epsilon_decay = p["epsilon_decay"]
        self.epsilon_min = p["epsilon_min"]
        self.gamma = p["gamma"]
        self.batch_size = p["batch_size"]
        self.first_update_steps = p["first_update_steps"]
        self.optimize_per_steps = p["optimize_per_steps"]
        self.target_update = p["target_update"]
        self.double_dqn = p["double_dqn"]
        self.update_loss_history = []

        m = self.model
        self._nl, self._H, self._C = m.n_layers, m.n_channels, m.in_channels
        lib = N.lib()
        from .wide import is_wide
        self._wide = is_wide(self._H) and not self._bn  # 4 x 2048 and the like: tensor-core path (pokeai_b200/wide.py)
        if self._bn:  # no folded / prepared copies while training: snapshots fold the running statistics (eval mode)
            self._prepared = self._prepared_target = None
        elif self._wide:
            n_prep = lib.pk_wide_prepared_bytes(self._nl, self._H, self._C)
            self._prepared = torch.zeros(n_prep, dtype=torch.uint8, device=self.device)
            self._prepared_target = torch.zeros(n_prep, dtype=torch.uint8, device=self.device)
        else:
            n_prep = lib.pk_mlp_prepared_floats(self._nl, self._H, self._C)
            self._prepared = torch.zeros(n_prep, dtype=torch.float32, device=self.device)
            self._prepared_target = torch.zeros(n_prep, dtype=torch.float32, device=self.device)
        self._grad = torch.zeros_like(self._flat)
        self._loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._staging = None
        if self._bn:
            self._bn_ws = {}
            return
        self._refresh_prepared()
        self._prepared_target.copy_(self._prepared)

    # -- model plumbing -------------------------------------------------------------------------------
    def _construct_model(self):
        return MLPModel(**self.model_params)  # trainer.py:69-71

    def _construct_flat_model(self, flat: Optional[torch.Tensor] = None):
        """An MLPModel whose parameters are views of one flat fp32 CUDA vector (state_dict order)."""
        model = self._construct_model()
        params = list(model.parameters())
        if flat is None:
            flat = torch.cat([p.detach().reshape(-1) for p in params]).to(self.device, torch.float32)
        if model.bn:
            assert flat.numel() == N.lib().pk_bn_param_floats(model.n_layers, model.n_channels, model.in_channels)
        else:
            assert flat.numel() == N.lib().pk_mlp_param_floats(model.n_layers, model.n_channels, model.in_channels)
        off = 0
        for p in params:
            n = p.numel()
            p.data = flat[off:off + n].view(p.shape)
            p.requires_grad_(False)
            off += n
        if model.bn:
            # running_mean / running_var of every layer are views of ONE (n_layers, 2, H) CUDA tensor the kernels move
            running = torch.stack([torch.stack([b.running_mean, b.running_var]) for b in model.bn_layers]) \
                .to(self.device, torch.float32).contiguous()
            for l, b in enumerate(model.bn_layers):
                b.running_mean = running[l, 0]
                b.running_var = running[l, 1]
            model._running = running
        return model, flat

    def _refresh_prepared(self) -> None:
        if self._bn:
            self.model.mark_updated()
            return
        if self._wide:
            from . import wide
            wide.prepare_wide(self.model.state_dict(), self.device, out=self._prepared)  # parameters are views of _flat
            return
        from . import ops
        ops.prepare_from_flat(self._flat, self._nl, self._H, self._C, out=self._prepared)

    def _snapshot_model(self) -> MLPModel:
        """A frozen copy of the online network (get_train_agent / get_val_agent, trainer.py:123-134)."""
        if self._bn:  # eval mode: BatchNorm folds into the convs with the running statistics of this moment
            model = self._construct_model()
            model.load_state_dict(self._cpu_sd(self.model))
            model.eval()
            return model
        model, flat = self._construct_flat_model(self._flat.clone())
        model.eval()
        if not self._wide:  # wide models are prepared by the agent's WideRunner
            model._prepared = self._prepared.clone()
            model._prepared_key = (str(self.device),) + model.weights_key()
        return model

    # -- checkpoints (trainer.py:73-121) ---------------------------------------------------------------
    @staticmethod
    def _cpu_sd(model) -> dict:
        return {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}

    def save_state(self, resume=False):
        if resume:
            return {
                "model": self._cpu_sd(self.model),
                "target_model": self._cpu_sd(self.target_model),
                "optimizer": self.optimizer.state_dict(),
                "replay_buffer": self.replay_buffer,
                "constructor_params": self.constructor_params,
                "total_steps": self.total_steps,
                "total_battles": self.total_battles,
                "update_steps": self.update_steps,
                "update_loss_history": self.update_loss_history,
            }
        return {
            "model": self._cpu_sd(self.model),
            "constructor_params": self.constructor_params,
            "update_steps": self.update_steps,
            "total_battles": self.total_battles,
        }

    @classmethod
    def load_state(cls, state, resume=False, **kwargs) -> "Trainer":
        trainer = cls(**state["constructor_params"], **kwargs)
        trainer.model.load_state_dict(state["model"])
        trainer._refresh_prepared()
        trainer.update_steps = state["update_steps"]
        if resume:
            trainer.total_steps = state["total_steps"]
            trainer.total_battles = state["total_battles"]
            trainer.target_model.load_state_dict(state["target_model"])
            if trainer._bn:
                pass
            elif trainer._wide:
                from . import wide
                wide.prepare_wide(trainer.target_model.state_dict(), trainer.device, out=trainer._prepared_target)
            else:
                from . import ops
                ops.prepare_from_flat(trainer._flat_target, trainer._nl, trainer._H, trainer._C,
                                      out=trainer._prepared_target)
            trainer.optimizer.load_state_dict(state["optimizer"])
            # a checkpoint written by the reference holds ITS ReplayBuffer (dense (617, A) arrays): converted to packed
            # records that encode to the same bits (packed.record_from_dense)
            from .replay_buffer import from_reference
            trainer.replay_buffer = from_reference(state["replay_buffer"])
            trainer.update_loss_history = state["update_loss_history"]
        return trainer

    def load_initial_model(self, state_dict):
        self.model.load_state_dict(state_dict)
        self.target_model.load_state_dict(state_dict)
        self._refresh_prepared()
        if not self._bn:
            self._prepared_target.copy_(self._prepared)

    # -- agents (trainer.py:123-134) ----------------------------------------------------------------------
    def get_train_agent(self):
        epsilon = max(math.pow(1.0 - self.epsilon_decay, self.total_steps) * self.epsilon, self.epsilon_min)
        return AgentTrain(self._snapshot_model(), self.feature_extractor, epsilon)

    def get_val_agent(self):
        return AgentVal(self._snapshot_model(), self.feature_extractor)

    def get_train_agents(self, n: int):
        """`n` training agents for `n` concurrent battles that share ONE snapshot of the online network (each keeps its
        own replay buffer and episode state), so that `LockStepDriver` can decide for all of them in one launch.
        Equivalent to `n` calls of `get_train_agent()` made at the same `total_steps` (trainer.py:123-128)."""
        epsilon = max(math.pow(1.0 - self.epsilon_decay, self.total_steps) * self.epsilon, self.epsilon_min)
        model = self._snapshot_model()
        return [AgentTrain(model, self.feature_extractor, epsilon) for _ in range(n)]

    def extend_replay_buffer(self, buffer: ReplayBuffer):
        steps = len(buffer)
        self.total_steps += steps
        self.replay_buffer.extend(buffer.buffer)

    # -- schedule (trainer.py:141-150) ------------------------------------------------------------------------
    def train(self):
        while self.update_steps < self.total_steps:
            update_step = self.update_steps
            if update_step % self.optimize_per_steps == 0 and update_step >= self.first_update_steps:
                self._update()
            if update_step % (self.optimize_per_steps * self.target_update) == 0:
                self._target_update()
            self.update_steps += 1

    # -- the update (trainer.py:152-195) ------------------------------------------------------------------------
    def _ensure_staging(self, n: int) -> None:
        if self._staging is not None and self._staging["n"] >= n:
            return
        # one pinned host block and one device block: [state n*144 | next n*144 | action n*4 | reward n*4 | terminal n]
        off_next = n * PK_OBS_BYTES
        off_act = 2 * n * PK_OBS_BYTES
        off_rew = off_act + 4 * n
        off_term = off_rew + 4 * n
        total = (off_term + n + 15) & ~15
        host = torch.empty(total, dtype=torch.uint8, pin_memory=True)
        dev = torch.empty(total, dtype=torch.uint8, device=self.device)
        hn = host.numpy()
        n_scr = N.lib().pk_dqn_scratch_floats(self._nl, self._H, self._C, n)
        self._staging = {
            "n": n, "host": host, "dev": dev, "off": (0, off_next, off_act, off_rew, off_term),
            "h_state": hn[0:off_next], "h_next": hn[off_next:off_act],
            "h_action": hn[off_act:off_rew].view(np.int32), "h_reward": hn[off_rew:off_term].view(np.float32),
            "h_terminal": hn[off_term:off_term + n],
            "scratch": torch.empty(n_scr, dtype=torch.float32, device=self.device),
        }

    def _dist(self):
        from .dist import active_group
        return active_group()

    def _update_from_arrays(self, state, next_state, action, reward, terminal, n_global: Optional[int] = None) -> float:
        """One optimiser step on explicit (already sharded) transition arrays.  Returns the loss
        (sum of this shard's Huber terms / n_global; the all-reduced mean when distributed)."""
        n = len(state)
        n_global = n if n_global is None else n_global
        if self._bn:
            return self._update_bn(state, next_state, action, reward, terminal, n_global)
        if self._wide:
            return self._update_wide(state, next_state, action, reward, terminal, n_global)
        self._ensure_staging(max(n, 1))
        S = self._staging
        S["h_state"][:n * PK_OBS_BYTES] = np.ascontiguousarray(state).view(np.uint8).reshape(-1)
        S["h_next"][:n * PK_OBS_BYTES] = np.ascontiguousarray(next_state).view(np.uint8).reshape(-1)
        S["h_action"][:n] = action
        S["h_reward"][:n] = reward
        S["h_terminal"][:n] = terminal
        S["dev"].copy_(S["host"], non_blocking=True)
        base = S["dev"].data_ptr()
        o = S["off"]
        b = N.pk_batch()
        b.n, b.n_global = n, n_global
        b.state, b.next_state = base + o[0], base + o[1]
        b.action, b.reward, b.terminal = base + o[2], base + o[3], base + o[4]
        stream = torch.cuda.current_stream(self.device).cuda_stream
        lib = N.lib()
        dist = self._dist()
        px = self._peer_exchange(dist)
        grad, loss = (self._grad, self._loss) if px is None else px.next_slot()
        if n == 0:  # a rank whose shard of a tiny minibatch is empty still takes part in the exchange
            grad.zero_()
            loss.zero_()
        else:
            N.check(lib.pk_dqn_grad(self._prepared.data_ptr(), self._prepared_target.data_ptr(), self._nl, self._H,
                                    self.feature_extractor.output_dim, C.byref(b), float(self.gamma),
                                    1 if self.double_dqn else 0, grad.data_ptr(), loss.data_ptr(),
                                    S["scratch"].data_ptr(), stream))
        opt = self.optimizer
        opt.step_count += 1
        if px is not None:  # ONE kernel over NVLink peer memory: sum over ranks (rank order) + clamp + Adam
            px.reduce_clamp_adam(self._flat, opt.exp_avg, opt.exp_avg_sq, self._nl, self._H, self._C, opt.step_count,
                                 opt.lr, opt.betas, opt.eps, prepared=self._prepared, grad_sum=self._grad)
            ls = px.loss_status.tolist()
            if ls[1] != 0.0:
                # the kernel gave up without writing parameters or Adam moments: undo the step count and say so
                opt.step_count -= 1
                raise RuntimeError("peer gradient exchange: a rank did not arrive within the time limit "
                                   "(PK_PEER_TIMEOUT_MS); this update was NOT applied")
            self.model.mark_updated(self._prepared)
            return float(ls[0])
        if dist is not None:  # sum of shard gradients == gradient of the global mean loss; clamp comes after
            from .dist import allreduce_sum_
            allreduce_sum_(self._grad, self._loss, dist)
        N.check(lib.pk_clamp_adam(self._flat.data_ptr(), self._grad.data_ptr(), opt.exp_avg.data_ptr(),
                                  opt.exp_avg_sq.data_ptr(), self._nl, self._H, self._C, opt.lr, opt.betas[0],
                                  opt.betas[1], opt.eps, opt.step_count, self._prepared.data_ptr(), stream))
        self.model.mark_updated(self._prepared)  # trainer.model(x) / Agent(trainer.model, fe) see the new weights
        return float(self._loss.item())

    def _peer_exchange_wide(self, dist):
        """PeerGradExchange for the wide network's 55 MB gradient (sliced pk_peer_allreduce), created collectively on
        the first distributed update; None when unavailable or switched off (PK_PEER_ALLREDUCE=0): per-layer NCCL buckets."""
        if dist is None or not self._wide:
            return None
        if "_pxw" not in self.__dict__:
            from .dist import PeerGradExchange
            self._pxw = PeerGradExchange.create(self._flat.numel(), self.device)
        return self._pxw

    def _peer_exchange(self, dist):
        """PeerGradExchange of the default (small) network when a NCCL group is active and the box maps peer memory;
        created collectively on the first distributed update, None otherwise (then: NCCL allreduce)."""
        if dist is None or self._wide:
            return None
        if "_px" not in self.__dict__:
            from .dist import PeerGradExchange
            self._px = PeerGradExchange.create(self._flat.numel(), self.device)
        return self._px

    def _wide_grad_device(self, s_d, s2_d, a_d, r_d, t_d, n_global: int, dist=None, out_grad=None):
        """Device-resident part of the wide update: fills self._grad, returns the loss tensor (1,).
        With `dist` (an active NCCL group) the gradient is also summed over ranks, one bucket per layer, each started as
        soon as that layer's slice is complete, so the 55 MB exchange overlaps the backward of the earlier layers
        (SURVEY.md 8e); the caller must not all-reduce self._grad again."""
        from . import wide
        n, A = s_d.shape[0], self.feature_extractor.output_dim
        if getattr(self, "_wide_ws", None) is None or self._wide_ws_n < n:
            self._wide_ws = wide.train_workspace(self._nl, self._H, n, A, self.device)
            self._wide_fw = wide.workspace(self._nl, self._H, n, A, self.device)
            self._wide_ws_n = n
        ws = self._wide_ws
        m2, q_no, _ = wide.wide_forward(s2_d, A, self._prepared, self._nl, self._H, ws=self._wide_fw)
        _, q_nt, _ = wide.wide_forward(s2_d, A, self._prepared_target, self._nl, self._H, ws=self._wide_fw)
        _, q_sa = wide.train_forward(s_d, A, self._prepared, self._nl, self._H, ws, col_select=a_d)
        dq, loss = wide.dqn_dq(q_sa, q_no if self.double_dqn else None, q_nt, m2, r_d, t_d, n_global, self.gamma,
                               self.double_dqn)
        if dist is None:
            wide.train_backward(self._prepared, self._nl, self._H, n, A, ws, dq,
                                out_grad=self._grad if out_grad is None else out_grad, chosen_only=True)
            return loss
        if getattr(self, "_wide_events", None) is None:
            self._wide_events = [torch.cuda.Event() for _ in range(self._nl)]
            for e in self._wide_events:
                e.record()  # materialises the CUDA event so that the library can record it
            self._wide_comm = torch.cuda.Stream(self.device)
            self._wide_buckets = wide.layer_grad_buckets(self._nl, self._H)
        main = torch.cuda.current_stream(self.device)
        wide.train_backward(self._prepared, self._nl, self._H, n, A, ws, dq, out_grad=self._grad, chosen_only=True,
                            layer_events=self._wide_events)
        comm = self._wide_comm
        with torch.cuda.stream(comm):
            for l in range(self._nl - 1, -1, -1):  # the order the backward finishes them in
                comm.wait_event(self._wide_events[l])
                lo, hi = self._wide_buckets[l]
                dist.all_reduce(self._grad[lo:hi], op=dist.ReduceOp.SUM)
        main.wait_stream(comm)
        return loss

    def _wide_apply(self, grad=None) -> None:
        """clamp + Adam on the flat parameters, then the bf16 / transposed weight copies are re-derived"""
        opt = self.optimizer
        opt.step_count += 1
        grad = self._grad if grad is None else grad
        N.check(N.lib().pk_clamp_adam(self._flat.data_ptr(), grad.data_ptr(), opt.exp_avg.data_ptr(),
                                      opt.exp_avg_sq.data_ptr(), self._nl, self._H, self._C, opt.lr, opt.betas[0],
                                      opt.betas[1], opt.eps, opt.step_count, None,
                                      torch.cuda.current_stream(self.device).cuda_stream))
        self._refresh_prepared()
        self.model.mark_updated()  # wide: agents re-derive their bf16 copies from the new parameters

    def _update_wide(self, state, next_state, action, reward, terminal, n_global: int) -> float:
        """The same Double-DQN step for a wide network (tensor-core path): Q(s, a) of the CHOSEN column with the
        activations kept (trainer.py:163), online + target forward over next_state (:171-177), Huber dq (:184-187),
        backward to the flat gradient (:192), [allreduce], clamp + Adam (:193-195), bf16 weights re-derived."""
        from . import ops, wide
        n = len(state)
        dev, A = self.device, self.feature_extractor.output_dim
        lib = N.lib()
        stream = torch.cuda.current_stream(dev).cuda_stream
        if n == 0:
            self._grad.zero_()
            self._loss.zero_()
        else:
            s_d, s2_d = ops.packed_to_device(state, dev), ops.packed_to_device(next_state, dev)
            a_d = torch.from_numpy(np.ascontiguousarray(action, dtype=np.int32)).to(dev)
            r_d = torch.from_numpy(np.ascontiguousarray(reward, dtype=np.float32)).to(dev)
            t_d = torch.from_numpy(np.ascontiguousarray(terminal, dtype=np.uint8)).to(dev)
        dist = self._dist()
        nccl = dist is not None and dist.get_backend() == "nccl"
        px = self._peer_exchange_wide(dist) if nccl else None
        if px is not None:
            # peer-memory route: the shard gradient is written straight into the exchange buffer of this epoch, ONE kernel
            # sums it over the ranks in place (pk_peer_allreduce: every element summed by one rank, in rank order), and
            # clamp + Adam run on the summed buffer; no NCCL call in the step
            grad, loss_slot = px.next_slot()
            if n > 0:
                loss_slot.copy_(self._wide_grad_device(s_d, s2_d, a_d, r_d, t_d, n_global, None, out_grad=grad))
            else:
                grad.zero_()
                loss_slot.zero_()
            px.allreduce()
            self._wide_apply(grad)
            if float(px.loss_status[1].item()) != 0.0:
                raise RuntimeError("peer gradient exchange: a rank did not arrive within the time limit "
                                   "(PK_PEER_TIMEOUT_MS); the parameters of this rank are no longer in step")
            return float(loss_slot.item())
        if n > 0:
            # NCCL group: the gradient is reduced bucket by bucket inside, overlapped with the backward
            loss = self._wide_grad_device(s_d, s2_d, a_d, r_d, t_d, n_global, dist if nccl else None)
            self._loss.copy_(loss)
        if dist is not None:
            if nccl:
                if n == 0:  # an empty shard still takes part in the same per-layer collectives, in the same order
                    from . import wide as _w
                    for lo, hi in reversed(_w.layer_grad_buckets(self._nl, self._H)):
                        dist.all_reduce(self._grad[lo:hi], op=dist.ReduceOp.SUM)
                dist.all_reduce(self._loss, op=dist.ReduceOp.SUM)
            else:
                from .dist import allreduce_sum_
                allreduce_sum_(self._grad, self._loss, dist)
        self._wide_apply()
        return float(self._loss.item())

    def _update_bn(self, state, next_state, action, reward, terminal, n_global: int) -> float:
        """The update of a bn=True network: every forward in TRAINING mode as in the reference (model(state_batch)
        trainer.py:163, model / target_model on the non-final next states :157-158,174-175,181-182: batch statistics,
        running statistics moved, in that order), Huber dq (:184-187), backward through the batch statistics (:192),
        clamp + Adam (:193-195)."""
        from . import bn_train, ops
        if self._dist() is not None:
            raise NotImplementedError("bn=True under data parallelism needs batch statistics synchronised across ranks; "
                                      "run the bn=True update on one GPU (the reference is single-process)")
        n = len(state)
        if n < 1:
            raise ValueError("empty minibatch")
        dev, A = self.device, self.feature_extractor.output_dim
        nl, H = self._nl, self._H
        terminal = np.asarray(terminal).astype(bool)
        idx_nf = np.flatnonzero(~terminal)
        if len(idx_nf) == 0:
            raise ValueError("need at least one array to stack")  # np.stack([]) at trainer.py:157
        slot = np.full(n, -1, dtype=np.int32)
        slot[idx_nf] = np.arange(len(idx_nf), dtype=np.int32)
        # one pinned block, one asynchronous copy: [state n*144 | non-final next m*144 | action | reward | slot]
        m = len(idx_nf)
        off_n, off_a = n * PK_OBS_BYTES, (n + m) * PK_OBS_BYTES
        off_r, off_s, total = off_a + 4 * n, off_a + 8 * n, off_a + 12 * n
        st = self._bn_ws.get("stage")
        if st is None or st[0].numel() < total:
            cap = 2 * total
            st = self._bn_ws["stage"] = (torch.empty(cap, dtype=torch.uint8, pin_memory=True),
                                         torch.empty(cap, dtype=torch.uint8, device=dev))
        host, dblk = st
        hn = host.numpy()
        hn[:off_n] = np.ascontiguousarray(state).view(np.uint8).reshape(-1)
        hn[off_n:off_a] = np.ascontiguousarray(next_state[idx_nf]).view(np.uint8).reshape(-1)
        hn[off_a:off_r].view(np.int32)[:] = action
        hn[off_r:off_s].view(np.float32)[:] = reward
        hn[off_s:total].view(np.int32)[:] = slot
        dblk[:total].copy_(host[:total], non_blocking=True)
        s_d, s2_d = dblk[:off_n].view(n, PK_OBS_BYTES), dblk[off_n:off_a].view(m, PK_OBS_BYTES)
        a_d, r_d = dblk[off_a:off_r].view(torch.int32), dblk[off_r:off_s].view(torch.float32)
        slot_d = dblk[off_s:total].view(torch.int32)
        W = self._bn_ws
        fwd = bn_train.forward_train_packed  # layer 0 gathered from the records: no dense (B, 617, A) block
        q, _, W["s"] = fwd(self._flat, self.model._running, nl, H, s_d, A, W.get("s"))
        q_no = None
        if self.double_dqn:
            q_no, _, W["n"] = fwd(self._flat, self.model._running, nl, H, s2_d, A, W.get("n"))
        q_nt, mask_n, W["n"] = fwd(self._flat_target, self.target_model._running, nl, H, s2_d, A, W.get("n"))
        dq, loss = bn_train.dqn_dq(q, a_d, r_d, slot_d, q_no, q_nt, mask_n, n_global, self.gamma, self.double_dqn)
        W["b"] = bn_train.backward_packed(self._flat, nl, H, s_d, A, W["s"], dq, self._grad, W.get("b"))
        opt = self.optimizer
        opt.step_count += 1
        bn_train.clamp_adam_flat(self._flat, self._grad, opt.exp_avg, opt.exp_avg_sq, opt.step_count, opt.lr, opt.betas,
                                 opt.eps)
        for b in self.model.bn_layers:
            b.num_batches_tracked += 2 if self.double_dqn else 1
        for b in self.target_model.bn_layers:
            b.num_batches_tracked += 1
        self.model.mark_updated()
        self.target_model.mark_updated()
        return float(loss.item())

    def _update(self):
        idx = self.replay_buffer.sample_indices(self.batch_size)
        state, next_state, action, reward, terminal = self.replay_buffer.gather(idx)
        if bool(terminal.all()):
            # the reference fails here too: np.stack([]) at trainer.py:157
            raise ValueError("need at least one array to stack")
        n_global = len(idx)
        dist = self._dist()
        if dist is not None:  # every rank drew the same minibatch (same `random` state); keep a contiguous shard
            from .dist import shard_bounds
            lo, hi = shard_bounds(n_global, dist.get_world_size(), dist.get_rank())
            state, next_state, action, reward, terminal = (x[lo:hi] for x in (state, next_state, action, reward, terminal))
        loss = self._update_from_arrays(state, next_state, action, reward, terminal, n_global)
        self.update_loss_history.append(loss)

    def _target_update(self):  # trainer.py:197-198
        self._flat_target.copy_(self._flat)
        if self._bn:  # load_state_dict copies the buffers too: running statistics and num_batches_tracked
            self.target_model._running.copy_(self.model._running)
            for bt, bm in zip(self.target_model.bn_layers, self.model.bn_layers):
                bt.num_batches_tracked.copy_(bm.num_batches_tracked)
            self.target_model.mark_updated()
            return
        self._prepared_target.copy_(self._prepared)
        self.target_model.mark_updated(None if self._wide else self._prepared_target)
