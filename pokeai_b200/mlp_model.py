"""MLPModel: the Q-network of the decision path, with the reference's constructor, parameter
names and shapes (pokeai/ai/generic_move_model/mlp_model.py:5-30) so that reference
checkpoints load with ``load_state_dict`` and ours load in the reference.

The module only *holds* parameters (torch is the memory manager).  ``forward`` runs the
hand-written sm_100a kernel behind ``pk_mlp_forward`` on dense CUDA input and raises on
anything else: there is no eager / CPU path in this package.  Inference from packed
observations does not go through ``forward`` at all (Agent.act_batch -> pk_decide).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn


class MLPModel(nn.Module):
    def __init__(self, input_shape, output_dim, n_layers: int = 2, n_channels: int = 64, bn: bool = False):
        super().__init__()
        layers = []
        bn_layers = []
        cur = input_shape[0]
        for _ in range(n_layers):
            layers.append(nn.Conv1d(cur, n_channels, 1, bias=not bn))  # mlp_model.py:12
            cur = n_channels
            if bn:
                bn_layers.append(nn.BatchNorm1d(n_channels))  # mlp_model.py:14-15
        self.layers = nn.ModuleList(layers)
        self.bn_layers = nn.ModuleList(bn_layers)
        self.output = nn.Conv1d(cur, 1, 1)  # mlp_model.py:18
        self.bn = bn
        self.in_channels = int(input_shape[0])
        self.output_dim = int(output_dim)
        self.n_layers = int(n_layers)
        self.n_channels = int(n_channels)
        self._prepared: Optional[torch.Tensor] = None
        self._prepared_key: Optional[Tuple] = None
        # bumped by whoever rewrites the parameters through raw pointers (Trainer's update kernels): torch's version
        # counters do not see such writes, and every cache of derived weights keys on `weights_key()`
        self._generation = 0

    # -- folded parameters --------------------------------------------------------------------
    def folded_state_dict(self) -> dict:
        """Parameters with eval-mode BatchNorm folded into the preceding conv (mlp_model.py:24-26):
        y = gamma * (W x - mean) / sqrt(var + eps) + beta  ==  (s W) x + (beta - s mean)."""
        sd = {}
        for i, conv in enumerate(self.layers):
            w = conv.weight.detach()
            b = conv.bias.detach() if conv.bias is not None else torch.zeros(w.shape[0], device=w.device)
            if self.bn:
                bnl = self.bn_layers[i]
                s = bnl.weight.detach() / torch.sqrt(bnl.running_var.detach() + bnl.eps)
                w = w * s[:, None, None]
                b = (b - bnl.running_mean.detach()) * s + bnl.bias.detach()
            sd[f"layers.{i}.weight"] = w
            sd[f"layers.{i}.bias"] = b
        sd["output.weight"] = self.output.weight.detach()
        sd["output.bias"] = self.output.bias.detach()
        return sd

    def weights_key(self) -> Tuple:
        """Changes whenever the parameters may have: storage pointers, torch version counters, and the generation
        counter that in-place kernel updates bump (`mark_updated`)."""
        return (self._generation,) + tuple((p.data_ptr(), p._version) for p in self.parameters()) + \
            tuple((b.data_ptr(), b._version) for b in self.buffers())

    def mark_updated(self, prepared: Optional[torch.Tensor] = None) -> None:
        """The parameters were rewritten behind torch's back (pk_clamp_adam / pk_peer_clamp_adam through raw
        pointers).  `prepared`: the up-to-date prepared copy the same kernel refreshed, adopted as this model's."""
        self._generation += 1
        if prepared is not None:
            self._prepared = prepared
            self._prepared_key = (str(prepared.device),) + self.weights_key()
        else:
            self._prepared = None
            self._prepared_key = None

    def prepared(self, device) -> torch.Tensor:
        """The kernels' transposed / padded copy of the weights on `device`, rebuilt when any
        parameter changes (`weights_key`)."""
        from . import ops
        device = torch.device(device)
        key = (str(device),) + self.weights_key()
        if self._prepared is None or self._prepared_key != key:
            if self.bn and self.training:
                raise RuntimeError("a bn=True model in training mode has no folded weights: call .eval() for the "
                                   "inference kernels (forward() in training mode uses batch statistics)")
            self._prepared = ops.prepare_weights(self.folded_state_dict(), device)[0]
            self._prepared_key = key
        return self._prepared

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x (B, C, A) fp32 CUDA -> (B, A).  mlp_model.py:21-30."""
        from . import ops
        if not x.is_cuda:
            raise RuntimeError("pokeai_b200.MLPModel.forward runs on CUDA tensors only (no CPU fallback); "
                               "use Agent.act_batch / _calc_q_vector_batch for host arrays")
        if x.requires_grad:
            raise RuntimeError("autograd does not flow through the CUDA forward; the update is Trainer._update")
        if self.bn and self.training:  # batch statistics, running statistics moved (pk_bn_forward_train)
            from . import bn_train
            return bn_train.model_forward_train(self, x)
        return ops.mlp_forward(self.prepared(x.device), self.n_layers, self.n_channels, x.contiguous().float())
