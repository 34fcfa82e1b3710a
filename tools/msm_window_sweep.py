"""Window width / segment length sweep of one G1 multiexp with a resident window table (B200 box; tuning aid).

python tools/msm_window_sweep.py [log_n ...]   -> one line per (log_n, c, seg_log): median ms of 10 calls after 3 warm-ups
Uses the library's own overrides (VSP_MSM_C at table build, VSP_MSM_SEG_LOG per call); results are bit-identical under all
of them (tests/test_gpu_prover.py::test_tuning_switches_do_not_change_results)."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import workload as wl  # noqa: E402

vsp = importlib.import_module("vote-saver-protocol_b200")


def main():
    logs = [int(x) for x in sys.argv[1:]] or [20]
    stream = torch.cuda.Stream()
    ctx = vsp.Context(0, stream.cuda_stream)
    for log_n in logs:
        n = 1 << log_n
        sc = torch.from_numpy(wl.random_fr(n, seed=1).view(np.int64)).cuda()
        out = torch.zeros(12, dtype=torch.int64, device="cuda")
        ref = None
        for c in ([None] + list(range(12, 20))):
            if c is None:
                os.environ.pop("VSP_MSM_C", None)
            else:
                os.environ["VSP_MSM_C"] = str(c)
            bases = vsp.G1Bases.generate(ctx, n, 4321, table=True)
            for seg in ([None] if c is None else [2, 3, 4, 5, 6]):
                if seg is None:
                    os.environ.pop("VSP_MSM_SEG_LOG", None)
                else:
                    os.environ["VSP_MSM_SEG_LOG"] = str(seg)
                ms = []
                with torch.cuda.stream(stream):
                    for it in range(13):
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        bases.multiexp_dev(n, sc.data_ptr(), out.data_ptr())
                        e1.record(stream)
                        torch.cuda.synchronize()
                        if it >= 3:
                            ms.append(e0.elapsed_time(e1))
                res = out.cpu().numpy().copy()
                if ref is None:
                    ref = res
                assert np.array_equal(res, ref), "tuning switches changed the result"
                print("log_n %d c %s seg_log %s: %.3f ms (min %.3f)" % (log_n, c, seg, float(np.median(ms)), min(ms)), flush=True)
            bases.close()
            ctx.trim()


if __name__ == "__main__":
    main()
