"""vsp_groth16_generate on the bench-sized circuit, wall clock per call plus device time per kernel class (development aid)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
vsp = importlib.import_module("vote-saver-protocol_b200")
import workload as wl
ctx = vsp.Context(0)
cs = wl.ballot_r1cs(65505, 30, seed=1)
tox = [int.from_bytes(x.tobytes(), "little") | 1 for x in wl.random_fr(5, 77)]
for it in range(4):
    if it == 2: ctx.profile_enable(True)
    t0 = time.perf_counter()
    pk, vk = vsp.generate(ctx, cs, tox)
    print("generate %.1f ms" % ((time.perf_counter() - t0) * 1e3))
print({k: (v["launches"], round(v["ms"], 2)) for k, v in ctx.profile_read().items() if v["launches"]})
