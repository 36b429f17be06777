"""CPU, world_size 2, gloo: the host-side logic of the data-parallel update (pokeai_b200/dist.py) --
sharding, sum-allreduce of shard gradients scaled by the GLOBAL batch, clamp AFTER the reduce.  The per-shard
gradient comes from the oracle here (no GPU in this container); on the GPU box the same invariant is checked on
the CUDA kernel by tests/test_gpu_update.py::test_update_is_deterministic_and_shards_add_up."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_shard_bounds_cover_and_balance():
    from pokeai_b200.dist import shard_bounds
    for n in (0, 1, 5, 32, 32768, 65537):
        for w in (1, 2, 3, 4, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import packed_oracle, ref_port
        from pokeai_b200.dist import active_group, allreduce_sum_, shard_bounds
        from pokeai_b200.packed import synth_packed
        from pokeai_b200.tables import load_tables
        T = load_tables()
        n, ps = 48, 3
        rng = np.random.default_rng(5)  # every rank builds the same global minibatch
        rec_s = synth_packed(n, seed=1, party_size=ps, kind="ragged", tables=T)
        rec_n = synth_packed(n, seed=2, party_size=ps, kind="ragged", tables=T)
        action = (rng.random(n) * rec_s["n_act"]).astype(np.int64)
        reward = (rng.uniform(-1, 1, n) * 40).astype(np.float32)
        terminal = rng.random(n) < 0.1
        sd = {k: v * 12.0 for k, v in ref_port.init_mlp_params(T.n_channels, 2, 16, seed=3).items()}  # big weights ->
        # gradients beyond the [-1, 1] clamp (the Huber loss alone bounds dL/dq by 1/N)
        tgt = ref_port.init_mlp_params(T.n_channels, 2, 16, seed=4)
        fs, _ = packed_oracle.encode(rec_s, ps, T)
        fn, mn = packed_oracle.encode(rec_n, ps, T)
        lo, hi = shard_bounds(n, world, rank)
        loss, grads = ref_port.dqn_update_grads(sd, tgt, fs[lo:hi], action[lo:hi], reward[lo:hi], fn[lo:hi], mn[lo:hi],
                                                ~terminal[lo:hi], 0.95, global_batch=n)
        keys = list(sd.keys())
        flat = torch.cat([grads[k].reshape(-1) for k in keys])
        local = flat.clone()
        loss_t = torch.tensor([loss], dtype=torch.float32)
        assert active_group() is not None
        allreduce_sum_(flat, loss_t)
        # the alternative order (clamp each shard, then sum) -- must NOT be what we do
        wrong = local.clamp(-1, 1)
        dist.all_reduce(wrong)
        if rank == 0:
            full_loss, full = ref_port.dqn_update_grads(sd, tgt, fs, action, reward, fn, mn, ~terminal, 0.95)
            full_flat = torch.cat([full[k].reshape(-1) for k in keys])
            np.savez(os.path.join(out_dir, "r0.npz"), reduced=flat.numpy(), full=full_flat.numpy(),
                     loss=np.float32(loss_t.item()), full_loss=np.float32(full_loss), wrong=wrong.numpy())
        np.save(os.path.join(out_dir, f"reduced_rank{rank}.npy"), flat.numpy())
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_allreduce_matches_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    z = np.load(tmp_path / "r0.npz")
    np.testing.assert_allclose(z["reduced"], z["full"], rtol=1e-5, atol=1e-7)
    assert abs(float(z["loss"]) - float(z["full_loss"])) < 1e-5 * max(1.0, abs(float(z["full_loss"])))
    # every rank holds the same reduced gradient bit for bit -> identical Adam steps, replicas stay in sync
    assert np.array_equal(np.load(tmp_path / "reduced_rank0.npy"), np.load(tmp_path / "reduced_rank1.npy"))
    # clamp-after-reduce is what the reference's semantics need: clamping first gives a different update
    assert np.abs(z["full"]).max() > 1.0
    right = np.clip(z["reduced"], -1, 1)
    assert np.abs(z["wrong"] - right).max() > 1e-3
