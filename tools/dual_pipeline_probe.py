"""Development probe: do two independent stream sets (two contexts, two host threads, half of the ballots each) prove a
step faster than one?  Tests whether de-phasing the per-chunk sort -> accumulate -> reduce pipelines pays."""
import importlib
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import workload as wl  # noqa: E402

vsp = importlib.import_module("vote-saver-protocol_b200")
nctx = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B = 256
cs = wl.ballot_r1cs(32000, 30, seed=0)  # XOR stand-in of the ballot circuit's size, but with wide values below
ni, na = cs["num_inputs"], cs["num_aux"]
_, m = vsp.domain_plan(cs["num_constraints"] + ni + 1)
ctxs, pks, keys, streams = [], [], [], []
for i in range(nctx):
    st = torch.cuda.Stream()
    ctx = vsp.Context(0, st.cuda_stream)
    pk = vsp.ProvingKey(ctx, wl.proving_key_arrays(vsp, ctx, cs, m, seed=7))
    pub_arr, gabc = wl.saver_pubkey_arrays(vsp, ctx, 25, seed=11)
    pub = vsp.SaverPublicKey(pub_arr)
    keys.append(pub.resident(ctx, np.ascontiguousarray(gabc[:25])))
    ctxs.append(ctx); pks.append(pk); streams.append(st)
# witnesses need not satisfy the circuit for timing: a third of the wires full-width, the rest bits, messages bits
rng = np.random.default_rng(1)
assign = np.zeros((B, ni + na + 1, 4), dtype=np.uint64)
assign[:, :, 0] = rng.integers(0, 2, size=(B, ni + na + 1), dtype=np.uint64)
wide = rng.random(ni + na + 1) < 0.3
wide[:31] = False
assign[:, wide] = wl.random_fr(int(wide.sum()) * B, 5).reshape(B, -1, 4)
assign[:, 0] = 0
assign[:, 0, 0] = 1
d_assign = torch.from_numpy(assign.view(np.int64)).cuda()
rs = wl.random_fr(3 * B, seed=99).reshape(B, 3, 4)
d_pr = torch.zeros((B, 48), dtype=torch.int64, device="cuda")
d_ct = torch.zeros((B, 27, 12), dtype=torch.int64, device="cuda")
per = B // nctx


def work(i):
    lo = i * per
    pks[i].encrypt_batch_dev(keys[i], per, d_assign[lo:lo + per].data_ptr(), rs[lo:lo + per], d_ct[lo:lo + per].data_ptr(),
                             d_pr[lo:lo + per].data_ptr())


def step():
    th = [threading.Thread(target=work, args=(i,)) for i in range(nctx)]
    for t in th:
        t.start()
    for t in th:
        t.join()


for _ in range(3):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
K = 5
for _ in range(K):
    step()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / K
print("contexts %d: %.1f ms per %d ballots = %.0f ballots/s" % (nctx, dt * 1e3, B, B / dt))
