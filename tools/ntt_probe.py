"""Fr NTT 2^log_n on resident vectors (8 buffers round-robin so no launch finds its input in L2); a development aid for
A/B runs of the pass kernel (VSP_NTT_RADIX2, VSP_NTT_BLOCK_LOG) — bench.py is the contract."""
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

stream = torch.cuda.Stream()
ctx = vsp.Context(0, stream.cuda_stream)
for log_n in [int(a) for a in sys.argv[1:]] or [16, 20, 24]:
    n = 1 << log_n
    nb = max(2, min(64, (1 << 23) // n))
    bufs = [torch.from_numpy(wl.random_fr(n, 2 + i).view(np.int64)).cuda() for i in range(nb)]
    dom = vsp.make_evaluation_domain(ctx, n)
    reps = 40
    with torch.cuda.stream(stream):
        for i in range(8):
            dom.fft_dev(bufs[i % nb].data_ptr(), False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(reps):
            dom.fft_dev(bufs[i % nb].data_ptr(), False)
        e1.record(stream)
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("ntt 2^%d: %.4f ms  %.2f TMAD/s" % (log_n, ms, (n / 2) * log_n * 136.0 / ms / 1e9))
