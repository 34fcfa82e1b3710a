"""One Groth16 proof of a 2^log_m-constraint synthetic R1CS (BASELINE configs[4] shape) on ONE GPU through
sharded.ShardedProver at world 1: times the load (window tables included) and the proof.  Development aid."""
import importlib
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402

vsp = importlib.import_module("vote-saver-protocol_b200")
log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
stream = torch.cuda.Stream()
ctx = vsp.Context(0, stream.cuda_stream)
ctx.profile_enable(True)
t0 = time.perf_counter()
out = bench.run_sharded_prove({"vsp": vsp, "ctx": ctx, "world": 1, "rank": 0}, log_m)
out["total_s_including_key_synthesis"] = time.perf_counter() - t0
free, total = torch.cuda.mem_get_info()
out["hbm_used_gb_after"] = (total - free) / 1e9
out["kernel_ms_over_4_proofs"] = {k: round(v["ms"], 2) for k, v in ctx.profile_read().items() if v["launches"]}
print(json.dumps(out))
