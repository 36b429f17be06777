"""Host-side cost of each part of one pipelined decision step (B = 4096 by default)."""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from pokeai_b200 import AgentVal, FeatureExtractor, MLPModel, DecisionPipeline
from pokeai_b200.packed import synth_packed
B = int(os.environ.get("B", 4096))
fe = FeatureExtractor(6, device=torch.device("cuda:0"))
torch.manual_seed(0)
agent = AgentVal(MLPModel(fe.input_shape, fe.output_dim).eval(), fe)
recs = [synth_packed(B, seed=i, party_size=6, kind="full") for i in range(4)]
pipe = DecisionPipeline(agent, B, depth=2)
for _ in pipe.map(recs[i % 4] for i in range(8)):
    pass
s = pipe._sessions[0]
n = 300
def t(fn):
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    return (time.perf_counter() - t0) / n * 1e6
print(f"B={B}")
print(f"  pack copy into pinned staging      {t(lambda i: fe.pack(recs[i % 4], out=s.packed)):7.1f} us")
print(f"  _point_weights (unchanged)         {t(lambda i: pipe._point_weights()):7.1f} us")
def sub(i):
    s.submit(B)
ts = []
for i in range(n):
    t0 = time.perf_counter(); s.submit(B); t1 = time.perf_counter(); s.wait(); t2 = time.perf_counter()
    ts.append((t1 - t0, t2 - t1))
print(f"  submit call returns after          {np.median([a for a, _ in ts])*1e6:7.1f} us; wait then takes {np.median([b for _, b in ts])*1e6:7.1f} us")
for i in range(n):
    s.submit(B); time.sleep(0.0005); t1 = time.perf_counter(); s.wait(); ts[i] = time.perf_counter() - t1
print(f"  wait on a finished batch           {np.median(ts)*1e6:7.1f} us")
print(f"  result copies (q, action)          {t(lambda i: (s.q[:B].copy(), s.action[:B].copy())):7.1f} us;  action only {t(lambda i: s.action[:B].copy()):7.1f} us")
# the same parts timed INSIDE a depth-3 pipeline loop
pipe3 = DecisionPipeline(agent, B, depth=3)
pipe3._point_weights()
acc = {"pack": 0.0, "submit": 0.0, "wait": 0.0, "copy": 0.0}
N_ = 600
pc = time.perf_counter
t_all = pc()
for i in range(N_):
    if pipe3._inflight == 3:
        s_ = pipe3._sessions[(pipe3._head - pipe3._inflight) % 3]
        t0 = pc(); q_, a_ = s_.wait(); t1 = pc(); a_ = a_.copy(); t2 = pc()
        pipe3._inflight -= 1
        acc["wait"] += t1 - t0; acc["copy"] += t2 - t1
    s_ = pipe3._sessions[pipe3._head]
    t0 = pc(); fe.pack(recs[i % 4], out=s_.packed); t1 = pc(); s_.submit(B); t2 = pc()
    acc["pack"] += t1 - t0; acc["submit"] += t2 - t1
    pipe3._head = (pipe3._head + 1) % 3; pipe3._inflight += 1
t_all = pc() - t_all
print(f"depth-3 loop: {t_all / N_ * 1e6:.1f} us per batch; " + ", ".join(f"{k} {v / N_ * 1e6:.1f}" for k, v in acc.items()))
