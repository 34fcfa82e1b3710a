"""Quick device-side timings of the hot kernels (development aid; bench.py is the contract)."""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
vsp = importlib.import_module("vote-saver-protocol_b200")
import workload as fx  # noqa: E402  (seeded input generation)


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    which = sys.argv[1:] or ["imad", "ntt", "msm"]
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    ctx = vsp.Context(0, stream if stream else None)
    if stream == 0:
        print("note: torch default stream is 0; ctx uses a private stream, timing by wall clock + sync")
    if "imad" in which:
        for it in (1024, 8192):
            print("imad rate iters=%d: %.3f TMAD/s" % (it, ctx.measure_imad_rate(it) / 1e12))

    def wall(fn, iters=5, warm=2):
        for _ in range(warm):
            fn()
        ctx.synchronize()
        t = time.perf_counter()
        for _ in range(iters):
            fn()
        ctx.synchronize()
        return (time.perf_counter() - t) / iters * 1e3

    if "ntt" in which:
        for lg in (14, 16, 18, 20, 22, 24):
            n = 1 << lg
            a = torch.from_numpy(fx.random_fr(n, 2).view(np.int64)).cuda()
            dom = vsp.make_evaluation_domain(ctx, n)
            ms = wall(lambda: dom.fft_dev(a.data_ptr(), False))
            print("ntt 2^%d: %.3f ms  (%.1f GB/s algorithmic 64 B/elt)" % (lg, ms, n * 64 / ms / 1e6))
    if "msm" in which or "msm20" in which or "msm16" in which:
        for lg in ((16, 18, 20, 22) if "msm" in which else ((20,) if "msm20" in which else (16,))):
            n = 1 << lg
            h = vsp.G1Bases.generate(ctx, n, 1234)
            s = torch.from_numpy(fx.random_fr(n, 1).view(np.int64)).cuda()
            out = torch.zeros(12, dtype=torch.int64, device="cuda")
            ms = wall(lambda: h.multiexp_dev(n, s.data_ptr(), out.data_ptr()), iters=3, warm=1)
            ctx.profile_enable(True)
            h.multiexp_dev(n, s.data_ptr(), out.data_ptr())
            prof = ctx.profile_read()
            ctx.profile_enable(False)
            print("g1 msm 2^%d (VSP_MSM_C=%s): %.3f ms  | %s" % (lg, os.environ.get("VSP_MSM_C", "auto"), ms, "  ".join(
                "%s %.2f" % (k, v["ms"]) for k, v in prof.items() if v["launches"])))
            h.close()


if __name__ == "__main__":
    main()
