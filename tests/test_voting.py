"""CPU tier: the ballot circuit (vote-saver-protocol_b200/voting.py).  Parity with crypto3-blueprint's wire
values is unpinned (its sources are absent); what is checked here is that the restated circuit is a correct circuit for
the statement the reference proves: every honest witness satisfies the R1CS under the ORACLE's evaluator
(oracle/pyref.py R1CS.is_satisfied), the public inputs it exposes (eid, sn, rt) equal an independent native computation
(Edwards scalar multiples, not the circuit's Montgomery window walk), dishonest witnesses do not satisfy it, and the
interface matches the reference's (30 primary inputs: 25 message slots, packed eid, sn, rt: common.hpp:861-876)."""
import importlib
import random

import pytest

from oracle import pyref as o

vt = importlib.import_module("vote-saver-protocol_b200.voting")


def test_jubjub_constants_and_generators():
    assert vt.Q == o.R
    assert vt.MONT_SCALE * vt.MONT_SCALE % vt.Q == -(vt.MONT_A + 2) % vt.Q
    for i in range(3):
        u, v = vt.generator(i)
        assert (-u * u + v * v) % vt.Q == (1 + vt.JUBJUB_D * u * u * v * v) % vt.Q
        assert vt.ed_mul(vt.JUBJUB_ORDER, (u, v)) == (0, 1) and (u, v) != (0, 1)
        x, y = vt.ed_to_mont((u, v))
        assert y * y % vt.Q == (x ** 3 + vt.MONT_A * x * x + x) % vt.Q
    assert len({vt.generator(i) for i in range(3)}) == 3
    # the group law used by the native hash: associativity and the order of a sum
    a, b, c = vt.generator(0), vt.generator(1), vt.generator(2)
    assert vt.ed_add(vt.ed_add(a, b), c) == vt.ed_add(a, vt.ed_add(b, c))
    assert vt.ed_add(a, vt.ed_neg(a)) == (0, 1)


@pytest.fixture(scope="module")
def small():
    rng = random.Random(1)
    c = vt.VotingCircuit(depth=2)
    sks = [rng.getrandbits(255) for _ in range(4)]
    pks = [vt.pedersen_hash_native(vt.digest_bits(s)) for s in sks]
    return c, rng, sks, pks


def test_interface_matches_the_reference(small):
    c = small[0]
    assert c.num_inputs == 30 and c.msg_size == 25  # common.hpp:163, 861-876
    assert (c.eid_packed, c.sn_packed, c.rt_packed) == (26, 27, 29)
    arr = c.r1cs_arrays()
    assert arr["num_constraints"] == c.num_constraints == len(arr["a_row"]) - 1
    assert arr["num_inputs"] + arr["num_aux"] == c.num_vars
    # size formula of the module docstring: 5 W + 5 S + 250 per hash
    per_hash = lambda n: 5 * ((n + 2) // 3) + 5 * (((n + 2) // 3 + 62) // 63) + 250
    want = 25 + 1 + 64 + 255 + c.depth + per_hash(255) + per_hash(64 + 255) + c.depth * (255 + 510 + per_hash(510)) + 1 + 2 + 2
    assert c.num_constraints == want
    big = vt.VotingCircuit(depth=16)
    assert big.num_constraints + big.num_inputs + 1 <= 1 << 15  # the depth-16 ballot circuit fits a 2^15 domain


@pytest.mark.parametrize("idx,vote", [(0, 0), (2, 3), (3, 24)])
def test_honest_witness_satisfies_and_exposes_the_right_public_inputs(small, idx, vote):
    c, rng, sks, pks = small
    eid = rng.getrandbits(64)
    prim, aux = c.witness(vote, eid, sks[idx], idx, vt.merkle_path_native(pks, idx))
    cs = o.R1CS(c.num_inputs, c.num_aux, c.constraints)
    assert cs.is_satisfied(prim, aux)
    assert prim[:25] == [1 if i == vote else 0 for i in range(25)]
    sn = vt.pedersen_hash_native([(eid >> i) & 1 for i in range(64)] + vt.digest_bits(sks[idx]))
    assert prim[25] == eid
    assert prim[26] + (prim[27] << 254) == sn
    assert prim[28] + (prim[29] << 254) == vt.merkle_root_native(pks)


def test_dishonest_witnesses_do_not_satisfy(small):
    c, rng, sks, pks = small
    cs = o.R1CS(c.num_inputs, c.num_aux, c.constraints)
    eid, idx = rng.getrandbits(64), 1
    path = vt.merkle_path_native(pks, idx)
    prim, aux = c.witness(5, eid, sks[idx], idx, path)
    assert cs.is_satisfied(prim, aux)
    two_votes = list(prim)
    two_votes[6] = 1
    assert not cs.is_satisfied(two_votes, aux)
    # a secret key that is not in the tree: the circuit is satisfiable only with a different root
    outsider = rng.getrandbits(255)
    p2, a2 = c.witness(5, eid, outsider, idx, path)
    assert cs.is_satisfied(p2, a2) and p2[28:30] != prim[28:30]
    forged = list(p2)
    forged[28:30] = prim[28:30]
    assert not cs.is_satisfied(forged, a2)
    # a serial number for another election
    other = list(prim)
    other[25] ^= 1
    assert not cs.is_satisfied(other, aux)
    # one wire flipped inside a hash
    bad = list(aux)
    bad[c.h_pk.var_base - c.num_inputs + 7] = (bad[c.h_pk.var_base - c.num_inputs + 7] + 1) % o.R
    assert not cs.is_satisfied(prim, bad)


def test_csr_arrays_equal_the_constraint_list(small):
    c = small[0]
    arr = c.r1cs_arrays()
    from oracle import cref
    for k, name in enumerate("abc"):
        for i in (0, 26, 400, c.num_constraints - 1):
            lo, hi = int(arr[name + "_row"][i]), int(arr[name + "_row"][i + 1])
            got = {int(arr[name + "_col"][q]): cref.limbs_to_ints(arr[name + "_val"][q:q + 1])[0] for q in range(lo, hi)}
            want = {}
            for idx, co in c.constraints[i][k]:
                want[idx] = (want.get(idx, 0) + co) % o.R
            assert got == {i_: v for i_, v in want.items() if v}
