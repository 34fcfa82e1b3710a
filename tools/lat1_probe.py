import sys, time, importlib, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tools')
import workload as wl
vsp = importlib.import_module("vote-saver-protocol_b200")
ctx = vsp.Context(0)
cs = wl.ballot_r1cs(65505, 30, seed=0)
_, m = vsp.domain_plan(65505+31)
pk = vsp.ProvingKey(ctx, wl.proving_key_arrays(vsp, ctx, cs, m, seed=7))
pub_arr, gabc = wl.saver_pubkey_arrays(vsp, ctx, 25, seed=11)
pub = vsp.SaverPublicKey(pub_arr); key = pub.resident(ctx, np.ascontiguousarray(gabc[:25]))
p,a = wl.ballot_witness(cs, seed=1)
rs = wl.random_fr(3, seed=9).reshape(1,3,4)
for i in range(6):
    t0=time.perf_counter(); pk.encrypt_batch(key, p[None], a[None], rs); print("ms", (time.perf_counter()-t0)*1e3)
