"""The oracle port (oracle/ref_port.py, oracle/packed_oracle.py) against the LIVE reference: the unmodified reference
modules staged in oracle/_ref by `__graft_entry__.build()` (oracle/ref_stage.py).  Skipped where nothing was staged
(the committed golden vectors of tests/test_oracle.py pin the same things there)."""
import numpy as np
import pytest

from conftest import bits

pytestmark = pytest.mark.reference


@pytest.fixture(scope="module")
def R():
    from oracle import ref_stage
    ref_stage.stage()  # no-op outside the build container
    r = ref_stage.load()
    if r is None:
        pytest.skip("reference not staged in oracle/_ref")
    return r


@pytest.mark.parametrize("party_size,kind,n", [(6, "full", 96), (6, "ragged", 96), (3, "ragged", 64), (1, "ragged", 32)])
def test_features_masks_q_and_actions_equal_the_reference(R, tables, party_size, kind, n):
    import torch
    from oracle import packed_oracle, ref_port, ref_stage
    from pokeai_b200.packed import synth_packed, unpack_observation
    rec = synth_packed(n, seed=900 + party_size, party_size=party_size, kind=kind, tables=tables)
    obs = [unpack_observation(r, tables) for r in rec]
    fe = R.FeatureExtractor(party_size)
    outs = [fe.transform(ref_stage.to_ref_obs(R, o)) for o in obs]
    f_ref, m_ref = np.stack([o[0] for o in outs]), np.stack([o[1] for o in outs])
    # object-level port and vectorised packed oracle: bit for bit
    f_port, m_port = ref_port.transform_batch(obs, party_size, tables)
    f_vec, m_vec = packed_oracle.encode(rec, party_size, tables)
    assert np.array_equal(bits(f_port), bits(f_ref)) and np.array_equal(m_port, m_ref)
    assert np.array_equal(bits(f_vec), bits(f_ref)) and np.array_equal(m_vec, m_ref)
    assert fe.get_dim_meanings()[:3] == ["state/" + s for s in tables.state_dim_meanings()[:3]]
    # Q and the masked argmax: the reference's MLPModel / Agent against the port on the same weights
    sd = ref_port.init_mlp_params(tables.n_channels, 3, 64, seed=party_size)
    model = R.MLPModel(fe.input_shape, fe.output_dim, n_layers=3, n_channels=64, bn=False)
    model.load_state_dict(sd)
    model.eval()
    agent = R.Agent(model, fe)
    with torch.no_grad():
        q_ref = agent._calc_q_vector_batch(f_ref)
        q_port = ref_port.mlp_forward(sd, torch.from_numpy(f_ref)).numpy()
    np.testing.assert_allclose(q_port, q_ref, rtol=1e-6, atol=1e-6)
    with torch.no_grad():  # the reference's callers run under no_grad too (rl_rating_battle.py)
        a_ref = [agent._act_by_model(f.copy(), m) for f, m in zip(f_ref, m_ref) if m.sum() > 0]
    a_port = ref_port.masked_argmax(q_ref, m_ref)[m_ref.sum(1) > 0]
    assert list(a_port) == a_ref


def test_reference_arm_of_the_bench_runs_the_reference(R, tables):
    """bench.py's CPU legs: kind == "reference" when staged, and the step really is the reference's classes"""
    import bench
    from pokeai_b200.packed import synth_packed
    rec = synth_packed(64, seed=0, party_size=6, kind="full", tables=tables)
    step, kind = bench.cpu_decide_factory(rec, 2, 64, 1)
    assert kind == "reference" and step(64) == 64


def test_reference_arm_under_torchrun_prints_one_line_from_rank_0(R):
    """`bench.py --impl reference --gpus 2` launched as the driver launches it (one process per GPU): rank 0 alone
    runs the CPU arm and prints ONE JSON line, the other rank exits 0 without work; no GPU is touched"""
    import json
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29571", "bench.py", "--impl", "reference",
                        "--gpus", "2", "--steps", "1", "--warmup", "1"], cwd=root, env=env, capture_output=True,
                       text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [json.loads(x) for x in p.stdout.splitlines() if x.startswith("{")]
    assert len(lines) == 1
    d = lines[0]
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["value"] > 0 and d["higher_is_better"] is True
