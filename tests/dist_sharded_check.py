"""torchrun worker of tests/test_gpu_sharded_prover.py::test_sharded_prover_over_nccl_two_gpus (and usable by hand with any
world size): one circuit, every query sharded by point range, proved through sharded.ShardedProver.prove_dev over NCCL;
rank 0 compares the proof with the oracle's single prover and a sharded multiexp with the oracle's multiexp."""
import importlib
import os
import random
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    vsp = importlib.import_module("vote-saver-protocol_b200")
    sh = importlib.import_module("vote-saver-protocol_b200.sharded")
    from oracle import cref, fixtures as fx, pyref as o
    import helpers as h
    stream = torch.cuda.Stream()
    ctx = vsp.Context(local, stream.cuda_stream)
    nc, ni = 1057, 30
    cs, primary, aux = h.synthetic_r1cs(nc, ni, seed=3)
    arr = h.synthetic_pk_arrays(cs, 8)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    prover = sh.ShardedProver(vsp, ctx, arr)
    assign = np.concatenate([cref.ints_to_limbs([1], 4), P, A])
    rng = random.Random(4)
    r, s = rng.randrange(o.R), rng.randrange(o.R)
    with torch.cuda.stream(stream):
        d_assign = torch.from_numpy(assign.view(np.int64)).cuda()
        proof = prover.prove_dev(d_assign, r, s)
        torch.cuda.synchronize()
        got = proof.cpu().numpy().view(np.uint64)
    host = prover.prove(P, A, r, s)  # the host-array path over the same NCCL group
    ok = True
    if rank == 0:
        want = cref.groth16_prove(fx.fill_pk(cref.ProvingKeyDesc(), arr), P, A, r, s)
        ok = bool(np.array_equal(got, want) and np.array_equal(host, want))
    # a sharded multiexp against the oracle
    n = 5000
    bases, scalars = h.random_g1(n, seed=5), fx.random_fr(n, seed=6)
    lo, hi = sh.shard_range(n, rank, world)
    sb = sh.ShardedBases(vsp, ctx, n, bases_local=bases[lo:hi])
    res = sb.multiexp(scalars[lo:hi])
    if rank == 0:
        ok = ok and bool(np.array_equal(res, cref.msm_g1(bases, scalars)))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    sb.close()
    prover.close()
    if rank == 0:
        print("SHARDED_OK" if int(flag.item()) == 1 else "SHARDED_MISMATCH", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
