"""CPU tier: the synthetic workloads bench.py runs (tools/workload.py, numpy only) against the oracle's restatement of the
reference's generators, and the integer helpers of the sharded legs' checksums."""
import os
import random
import sys

import numpy as np

from oracle import cref, fixtures as fx, pyref as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import workload as wl  # noqa: E402


def _csr_to_r1cs(cs: dict) -> o.R1CS:
    out = o.R1CS(cs["num_inputs"], cs["num_aux"])
    for i in range(cs["num_constraints"]):
        lcs = []
        for m in "abc":
            lo, hi = int(cs[m + "_row"][i]), int(cs[m + "_row"][i + 1])
            lcs.append([(int(cs[m + "_col"][k]), cref.limbs_to_ints(cs[m + "_val"][k:k + 1])[0]) for k in range(lo, hi)])
        out.constraints.append(tuple(lcs))
    return out


def test_field_input_system_is_the_reference_generator_shape():
    """workload.field_r1cs / field_witness (BASELINE configs[4]) == generate_r1cs_example_with_field_input
    (r1cs_examples.hpp:75-144) as the oracle restates it: same constraints term for term, and the witness satisfies it"""
    for nc, ni in ((7, 3), (20, 4), (33, 30)):
        cs = wl.field_r1cs(nc, ni)
        want, _p, _a = o.r1cs_example_with_field_input(nc, ni, random.Random(1))
        got = _csr_to_r1cs(cs)
        assert got.primary_input_size == want.primary_input_size and got.auxiliary_input_size == want.auxiliary_input_size
        norm = lambda lc: sorted((i, c % o.R) for i, c in lc)
        assert [tuple(norm(x) for x in c) for c in got.constraints] == [tuple(norm(x) for x in c) for c in want.constraints]
        prim, aux = wl.field_witness(cs, seed=nc)
        assert got.is_satisfied(cref.limbs_to_ints(prim), cref.limbs_to_ints(aux))
        lay = wl.key_layout(cs, 64)
        assert lay["nb"] == len(wl.b_query_support(cs)) and lay["A"][1] == nc + 3


def test_sharded_checksum_helpers():
    sys.path.insert(0, ROOT)
    import bench
    sc, k = bench.indexed_scalars(10, 300, 77), bench.base_multipliers(10, 300, 4321)
    ints = cref.limbs_to_ints(sc)
    assert all(v < o.R for v in ints)
    assert bench.dot_mod_r(sc, k) == sum(a * int(b) for a, b in zip(ints, k)) % o.R
    assert np.array_equal(bench.indexed_scalars(0, 20, 77)[10:], bench.indexed_scalars(10, 10, 77))  # a function of the index
    from helpers import generated_base_scalars
    assert np.array_equal(k, generated_base_scalars(300, 4321, first=10))
