"""End-to-end timing of AgentVal.act_batch (host records -> actions) at a few batch sizes, with the parts timed."""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from pokeai_b200 import AgentVal, FeatureExtractor, MLPModel
from pokeai_b200.packed import synth_packed
dev = torch.device("cuda:0")
fe = FeatureExtractor(6, device=dev)
torch.manual_seed(0)
model = MLPModel(fe.input_shape, fe.output_dim).eval()
agent = AgentVal(model, fe)
for B in (64, 4096, 65536):
    recs = [synth_packed(B, seed=i, party_size=6, kind="full") for i in range(4)]
    for i in range(5):
        a = agent.act_batch(recs[i % 4])
    t0 = time.perf_counter(); n = 200 if B <= 4096 else 30
    for i in range(n):
        a = agent.act_batch(recs[i % 4])
    dt = (time.perf_counter() - t0) / n
    s = agent._session(B)
    t1 = time.perf_counter()
    for i in range(n):
        fe.pack(recs[i % 4], out=s.packed)
    tp = (time.perf_counter() - t1) / n
    t2 = time.perf_counter()
    for i in range(n):
        s.decide(B)
    td = (time.perf_counter() - t2) / n
    print(f"B={B}: act_batch {dt*1e6:.1f} us ({B/dt/1e6:.2f} M decisions/s); staging copy {tp*1e6:.1f} us; session.decide {td*1e6:.1f} us; env PK_ZERO_COPY={os.environ.get('PK_ZERO_COPY','1')}")
    from pokeai_b200 import DecisionPipeline
    for depth in (1, 2, 3, 4):
        pipe = DecisionPipeline(agent, B, depth=depth)
        for _ in pipe.map(recs[i % 4] for i in range(8)):
            pass
        t3 = time.perf_counter()
        for _ in pipe.map(recs[i % 4] for i in range(n)):
            pass
        tq = (time.perf_counter() - t3) / n
        pipe.close()
        print(f"   pipeline depth {depth}: {tq*1e6:.1f} us per batch ({B/tq/1e6:.2f} M decisions/s)")
    t4 = time.perf_counter()
    for i in range(n):
        s.submit(B); s.wait()
    print(f"   submit+wait without the staging copy: {(time.perf_counter()-t4)/n*1e6:.1f} us; result copy {B*40/1e3:.0f} KB")
