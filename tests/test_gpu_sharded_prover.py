"""Point-range sharded proving through the C ABI (vsp_pk_load_slice / vsp_prove_partial / vsp_prove_assemble): the G
ranks of BASELINE configs[4] are emulated one after another on the one GPU the test tier has; the exchange is the
all_gather's identity (stacking).  Proofs must equal the oracle's single-prover proof bit for bit."""
import importlib
import random

import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import synthetic_pk_arrays, synthetic_r1cs, synthetic_saver_pub

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nc,ni,world", [(97, 30, 2), (1057, 30, 3), (18, 5, 8)])
def test_sharded_proof_equals_single_prover(vsp, ctx, nc, ni, world):
    sh = importlib.import_module("vote-saver-protocol_b200.sharded")
    cs, primary, aux = synthetic_r1cs(nc, ni, seed=nc)
    arr = synthetic_pk_arrays(cs, nc + 1)
    pub_arr, _ = synthetic_saver_pub(5, seed=9)
    pub = vsp.SaverPublicKey(pub_arr)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    _, m = vsp.domain_plan(nc + ni + 1)
    parts, keys = [], []
    for rank in range(world):
        sliced, rng = sh.slice_proving_key(arr, m, rank, world)
        pk = vsp.ProvingKey(ctx, sliced, slice=rng)
        parts.append(pk.prove_partial(P, A))
        keys.append(pk)
    sums = sh.sum_partials(np.stack(parts), lambda p: vsp.points_sum(ctx, p), lambda p: vsp.points_sum(ctx, p))
    rng_ = random.Random(5)
    r, s, rsv = (rng_.randrange(o.R) for _ in range(3))
    opk = fx.fill_pk(cref.ProvingKeyDesc(), arr)
    assert np.array_equal(keys[0].prove_assemble(sums, r, s), cref.groth16_prove(opk, P, A, r, s))
    assert np.array_equal(keys[-1].prove_assemble(sums, r, s, pub, rsv),
                          cref.groth16_prove(opk, P, A, r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"]))
    # a sliced key refuses the whole-proof entry point
    with pytest.raises(ValueError):
        keys[0].prove(P, A, r, s)
    # a full key's partial is the whole multiexp: assembling it alone gives the same proof
    full = vsp.ProvingKey(ctx, arr)
    assert np.array_equal(full.prove_assemble(full.prove_partial(P, A), r, s), cref.groth16_prove(opk, P, A, r, s))
    full.close()
    for k in keys:
        k.close()


@pytest.mark.parametrize("nc,ni,world,binary", [(97, 30, 2, True), (4500, 6, 3, False)])
def test_device_resident_sharded_proof(vsp, ctx, nc, ni, world, binary):
    """the path bench.py times (sharded.ShardedProver.prove_dev): vsp_prove_partial_dev leaves raw XYZZ shares on the
    device, the all_gather is emulated by writing every rank's 144 limbs into one device tensor, vsp_prove_assemble_dev
    sums them on the device.  Slices generated per rank exactly as the bench does.  Equal to the oracle's single prover."""
    import torch
    sh = importlib.import_module("vote-saver-protocol_b200.sharded")
    cs, primary, aux = synthetic_r1cs(nc, ni, seed=nc, binary=binary)
    arr = synthetic_pk_arrays(cs, nc + 2)
    pub_arr, _ = synthetic_saver_pub(5, seed=9)
    pub = vsp.SaverPublicKey(pub_arr)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    _, m = vsp.domain_plan(nc + ni + 1)
    assign = np.concatenate([cref.ints_to_limbs([1], 4), P, A])
    d_assign = torch.from_numpy(assign.view(np.int64)).cuda()
    gathered = torch.zeros((world, sh.PARTIAL_LIMBS), dtype=torch.int64, device="cuda")
    d_proof = torch.zeros(48, dtype=torch.int64, device="cuda")
    keys = []
    for rank in range(world):
        sliced, rng = sh.slice_proving_key(arr, m, rank, world)
        assert rng == sh.key_ranges(arr["cs"], m, int(arr["B_count"]), rank, world)
        pk = vsp.ProvingKey(ctx, sliced, slice=rng)
        pk.prove_partial_dev(d_assign.data_ptr(), gathered[rank].data_ptr())
        keys.append(pk)
    rng_ = random.Random(6)
    r, s, rsv = (rng_.randrange(o.R) for _ in range(3))
    opk = fx.fill_pk(cref.ProvingKeyDesc(), arr)
    keys[0].prove_assemble_dev(world, gathered.data_ptr(), r, s, d_proof.data_ptr())
    ctx.synchronize()
    assert np.array_equal(d_proof.cpu().numpy().view(np.uint64), cref.groth16_prove(opk, P, A, r, s))
    keys[-1].prove_assemble_dev(world, gathered.data_ptr(), r, s, d_proof.data_ptr(), pub, rsv)
    ctx.synchronize()
    assert np.array_equal(d_proof.cpu().numpy().view(np.uint64),
                          cref.groth16_prove(opk, P, A, r, s, rsv, pub_arr["gamma_inverse_sum_s_g1"]))
    for k in keys:
        k.close()


def test_sharded_prover_over_nccl_two_gpus():
    """real NCCL, two processes, two GPUs (skipped on a one-GPU box): tests/dist_sharded_check.py proves one circuit with
    sharded.ShardedProver.prove_dev on both ranks and compares with the oracle's single prover on rank 0"""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", os.path.join(here, "dist_sharded_check.py")],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "SHARDED_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_split_witness_map_begin_finish(vsp, ctx):
    """vsp_prove_partial_begin_dev / _finish_dev: three emulated ranks each extend one of aA, aB, aC to the coset, the
    'broadcast' is a device copy of the owner's vector into the others' buffers, and the finished shares assemble to the
    oracle's proof (basic and step domains); begin with mask 7 + finish equals prove_partial_dev"""
    import torch
    sh = importlib.import_module("vote-saver-protocol_b200.sharded")
    for nc, ni in ((97, 30), (1057, 30)):
        world = 3
        cs, primary, aux = synthetic_r1cs(nc, ni, seed=nc + 5)
        arr = synthetic_pk_arrays(cs, nc + 3)
        P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
        _, m = vsp.domain_plan(nc + ni + 1)
        d_assign = torch.from_numpy(np.concatenate([cref.ints_to_limbs([1], 4), P, A]).view(np.int64)).cuda()
        gathered = torch.zeros((world, sh.PARTIAL_LIMBS), dtype=torch.int64, device="cuda")
        abc = [torch.zeros((3, m * 4), dtype=torch.int64, device="cuda") for _ in range(world)]
        keys = []
        owners = sh.vector_owners(world)
        for rank in range(world):
            sliced, rng = sh.slice_proving_key(arr, m, rank, world)
            pk = vsp.ProvingKey(ctx, sliced, slice=rng)
            mask = sum(1 << v for v in range(3) if owners[v] == rank)
            pk.prove_partial_begin_dev(d_assign.data_ptr(), mask, abc[rank].data_ptr(), gathered[rank].data_ptr())
            keys.append(pk)
        ctx.synchronize()
        for v in range(3):  # the exchange
            for rank in range(world):
                if rank != owners[v]:
                    abc[rank][v].copy_(abc[owners[v]][v])
        torch.cuda.synchronize()
        for rank in range(world):
            keys[rank].prove_partial_finish_dev(abc[rank].data_ptr(), gathered[rank].data_ptr())
        d_proof = torch.zeros(48, dtype=torch.int64, device="cuda")
        r, s = 0x1234, o.R - 99
        keys[1].prove_assemble_dev(world, gathered.data_ptr(), r, s, d_proof.data_ptr())
        ctx.synchronize()
        opk = fx.fill_pk(cref.ProvingKeyDesc(), arr)
        assert np.array_equal(d_proof.cpu().numpy().view(np.uint64), cref.groth16_prove(opk, P, A, r, s))
        # mask 7 on one rank = the undivided call
        one = torch.zeros(sh.PARTIAL_LIMBS, dtype=torch.int64, device="cuda")
        two = torch.zeros(sh.PARTIAL_LIMBS, dtype=torch.int64, device="cuda")
        keys[0].prove_partial_dev(d_assign.data_ptr(), one.data_ptr())
        keys[0].prove_partial_begin_dev(d_assign.data_ptr(), 7, abc[0].data_ptr(), two.data_ptr())
        keys[0].prove_partial_finish_dev(abc[0].data_ptr(), two.data_ptr())
        # raw XYZZ shares are projective (their limbs depend on the order the sort left inside each bucket): compare the points
        pa, pb = torch.zeros(48, dtype=torch.int64, device="cuda"), torch.zeros(48, dtype=torch.int64, device="cuda")
        keys[0].prove_assemble_dev(1, one.data_ptr(), r, s, pa.data_ptr())
        keys[0].prove_assemble_dev(1, two.data_ptr(), r, s, pb.data_ptr())
        ctx.synchronize()
        assert torch.equal(pa, pb)
        for k in keys:
            k.close()
