"""G1 multiexp 2^log_n on resident bases, a few runs (development aid for ncu launch lists; bench.py is the contract)."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
vsp = importlib.import_module("vote-saver-protocol_b200")
import workload as wl  # noqa: E402

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = 1 << log_n
stream = torch.cuda.Stream()
ctx = vsp.Context(0, stream.cuda_stream)
h = vsp.G1Bases.generate(ctx, n, 1234)
sc = torch.from_numpy(wl.random_fr(n, 1).view(np.int64)).cuda()
out = torch.zeros(12, dtype=torch.int64, device="cuda")
with torch.cuda.stream(stream):
    for _ in range(3):
        h.multiexp_dev(n, sc.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(5):
        h.multiexp_dev(n, sc.data_ptr(), out.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
print("msm 2^%d: %.3f ms" % (log_n, e0.elapsed_time(e1) / 5))
