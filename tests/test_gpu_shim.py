"""The C++ host side: include/vsp_crypto3_shim.hpp (crypto3-named entry points over the C ABI) driven by
tests/cpp/shim_check the way bin/cli's vote phase calls them (common.hpp:1132-1145), every result compared with the
oracle through a fixture file."""
import os
import random
import struct
import subprocess

import numpy as np
import pytest

from oracle import cref, fixtures as fx, pyref as o

from helpers import random_g1, random_g2, synthetic_pk_arrays, synthetic_r1cs, synthetic_saver_pub

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")


def build_shim_check() -> str:
    subprocess.check_call(["make", "-C", CPP], stdout=subprocess.DEVNULL)
    return os.path.join(CPP, "shim_check")


def write_fixture(path: str):
    rng = random.Random(77)
    rec = {}
    # domains: m_req = 21 -> step_radix2 of size 24 (16 + 8)
    _, m = cref.domain_plan(21)
    assert m == 24
    a = fx.random_fr(m, seed=3)
    rec["fft_m"], rec["fft_m_out"] = [21], [m]
    rec["fft_in"], rec["fft_out"] = a, cref.domain_fft(m, a)
    g = rng.randrange(2, o.R)
    rec["coset_g"], rec["coset_out"] = cref.ints_to_limbs([g], 4), cref.multiply_by_coset(a, g)
    # multiexp (zero / one scalars and a point at infinity included)
    b1, s1 = random_g1(300, 5), fx.random_fr(300, seed=6)
    s1[0], s1[1] = 0, 0
    s1[1, 0] = 1
    b1[2] = 0
    rec["msm_bases"], rec["msm_scalars"], rec["msm_out"] = b1, s1, cref.msm_g1(b1, s1)
    b2, s2 = random_g2(40, 7), fx.random_fr(40, seed=8)
    rec["msm2_bases"], rec["msm2_scalars"], rec["msm2_out"] = b2, s2, cref.msm_g2(b2, s2)
    # keys and one ballot
    nc, ni, msg_size = 97, 30, 25
    cs, primary, aux = synthetic_r1cs(nc, ni, seed=21, binary=True)
    arr = synthetic_pk_arrays(cs, 22)
    pub_arr, _ = synthetic_saver_pub(msg_size, seed=23)
    gabc = random_g1(ni, 24)
    csa = arr["cs"]
    rec["num_inputs"], rec["num_aux"], rec["num_constraints"] = [csa["num_inputs"]], [csa["num_aux"]], [csa["num_constraints"]]
    for k in "abc":
        rec[k + "_row"], rec[k + "_col"], rec[k + "_val"] = csa[k + "_row"], csa[k + "_col"].astype(np.uint64), csa[k + "_val"]
    for k in ("alpha_g1", "beta_g1", "beta_g2", "delta_g1", "delta_g2", "A_query", "H_query", "L_query", "B_g1", "B_g2"):
        rec[k] = arr[k]
    rec["B_count"], rec["B_idx"] = [arr["B_count"]], arr["B_idx"].astype(np.uint64)
    rec["gamma_abc_rest"] = gabc
    for k in ("delta_g1", "delta_s_g1", "t_g1", "t_g2", "delta_sum_s_g1", "gamma_inverse_sum_s_g1"):
        rec["pub_" + k] = pub_arr[k]
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    r_saver, r, s = (rng.randrange(o.R) for _ in range(3))
    rec["primary"], rec["aux"] = P, A
    rec["r"], rec["s"], rec["r_saver"] = (cref.ints_to_limbs([x], 4) for x in (r, s, r_saver))
    opk = fx.fill_pk(cref.ProvingKeyDesc(), arr)
    opub = fx.fill_struct(cref.SaverPubkeyDesc(), pub_arr)
    rec["proof_basic"] = cref.groth16_prove(opk, P, A, r, s)
    rec["proof_enc"] = cref.groth16_prove(opk, P, A, r, s, r_saver, pub_arr["gamma_inverse_sum_s_g1"])
    ct = cref.saver_ciphertext(opub, np.ascontiguousarray(gabc[:msg_size]), np.ascontiguousarray(P[:msg_size]), r_saver)
    rec["ct"] = ct
    rnd = [rng.randrange(1, o.R) for _ in range(3)]
    rec["rnd"] = cref.ints_to_limbs(rnd, 4)
    rec["rerand_ct"], rec["rerand_proof"] = cref.saver_rerandomize(opub, arr["delta_g2"], rnd, ct, rec["proof_enc"])
    # key generation (zk::generate, pubkey::generate_keypair) on a small circuit, expected values from the Python oracle
    kcs, _kp, _ka = synthetic_r1cs(12, 5, seed=31, binary=True)  # bit witness: message slots within decrypt's search range
    rec["kg_primary"], rec["kg_aux"] = cref.ints_to_limbs(_kp, 4), cref.ints_to_limbs(_ka, 4)
    tox = [rng.randrange(1, o.R) for _ in range(5)]
    opk, ovk = o.groth16_generate(kcs, None, toxic=tox)
    karr = fx.pk_arrays(opk)
    kcsa = karr["cs"]
    rec["kg_num_inputs"], rec["kg_num_aux"], rec["kg_num_constraints"] = ([kcsa["num_inputs"]], [kcsa["num_aux"]],
                                                                         [kcsa["num_constraints"]])
    for k in "abc":
        rec["kg_" + k + "_row"], rec["kg_" + k + "_col"], rec["kg_" + k + "_val"] = (
            kcsa[k + "_row"], kcsa[k + "_col"].astype(np.uint64), kcsa[k + "_val"])
    rec["kg_toxic"] = cref.ints_to_limbs(tox, 4)
    for k in ("alpha_g1", "beta_g1", "beta_g2", "delta_g1", "delta_g2", "A_query", "H_query", "L_query", "B_g1", "B_g2"):
        rec["kg_" + k] = karr[k]
    rec["kg_B_count"], rec["kg_B_idx"] = [karr["B_count"]], karr["B_idx"].astype(np.uint64)
    A1 = lambda pts: cref.g1_to_limbs([o.G1.to_affine(p) for p in pts])
    A2 = lambda pts: cref.g2_to_limbs([o.G2.to_affine(p) for p in pts])
    rec["kg_gamma_ABC_g1"] = A1([ovk.gamma_ABC_first] + ovk.gamma_ABC_rest)
    rec["kg_gamma_g2"], rec["kg_gamma_g1"] = A2([ovk.gamma_g2]), A1([ovk.gamma_g1])
    kmsg = 3
    krnd = [rng.randrange(1, o.R) for _ in range(3 * kmsg + 2)]
    kpub, _rho, ksvk = o.saver_keygen(ovk, kmsg, krnd)
    rec["kg_msg_size"], rec["kg_rnd"] = [kmsg], cref.ints_to_limbs(krnd, 4)
    kpa = fx.saver_pub_arrays(kpub)
    for k in ("delta_s_g1", "t_g1", "t_g2", "delta_sum_s_g1", "gamma_inverse_sum_s_g1"):
        rec["kg_pub_" + k] = kpa[k]
    rec["kg_rho_g2"], rec["kg_rho_sv_g2"], rec["kg_rho_rhov_g2"] = A2([ksvk.rho_g2]), A2(ksvk.rho_sv_g2), A2(ksvk.rho_rhov_g2)
    with open(path, "wb") as f:
        for name, v in rec.items():
            data = np.ascontiguousarray(np.asarray(v, dtype=np.uint64)).reshape(-1)
            nb = name.encode()
            f.write(struct.pack("<I", len(nb)) + nb + struct.pack("<Q", data.size) + data.tobytes())


def test_shim_check_builds_and_refuses_to_run_without_a_gpu(tmp_path):
    """CPU tier: the shim compiles as plain C++17 against the header + .so; with no CUDA device it fails loudly
    (exit code 3, 'a CUDA device is required') instead of computing anything on the host."""
    import torch
    exe = build_shim_check()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test below")
    fixture = str(tmp_path / "fx.bin")
    with open(fixture, "wb") as f:
        for name, v in (("fft_m", [8]), ("fft_m_out", [8]), ("fft_in", [0] * 32)):
            nb = name.encode()
            d = np.asarray(v, dtype=np.uint64)
            f.write(struct.pack("<I", len(nb)) + nb + struct.pack("<Q", d.size) + d.tobytes())
    p = subprocess.run([exe, fixture], capture_output=True, text=True)
    assert p.returncode == 3 and "CUDA device is required" in p.stderr


@pytest.mark.gpu
def test_crypto3_shim_matches_oracle(tmp_path):
    exe = build_shim_check()
    fixture = str(tmp_path / "shim_fixture.bin")
    write_fixture(fixture)
    p = subprocess.run([exe, fixture], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "all shim checks passed" in p.stdout


def _build_blob_prove_check(tmp_path) -> str:
    import importlib
    import shutil
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    lib_dir = os.path.dirname(importlib.import_module("vote-saver-protocol_b200").library_path())
    exe = str(tmp_path / "blob_prove_check")
    subprocess.check_call([gxx, "-O1", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(CPP, "blob_prove_check.cpp"), "-o", exe, "-L", lib_dir, "-lvsp_b200",
                           "-Wl,-rpath," + lib_dir])
    return exe


def _write_prove_blobs(tmp_path, seed=41):
    import importlib
    codec = importlib.import_module("vote-saver-protocol_b200.codec")
    cs, primary, aux = synthetic_r1cs(97, 30, seed=seed, binary=True)
    arr = synthetic_pk_arrays(cs, seed + 1)
    P, A = cref.ints_to_limbs(primary, 4), cref.ints_to_limbs(aux, 4)
    r, s = 0x1234567890abcdef, o.R - 5
    (tmp_path / "pk.bin").write_bytes(codec.write_fast_proving_key(arr))
    (tmp_path / "primary.bin").write_bytes(codec.primary_input_bytes(P))
    (tmp_path / "aux.bin").write_bytes(codec.primary_input_bytes(A))
    (tmp_path / "rs.bin").write_bytes(codec.primary_input_bytes(cref.ints_to_limbs([r, s], 4)))
    return codec, arr, P, A, r, s


def test_blob_prove_check_builds_and_refuses_to_run_without_a_gpu(tmp_path):
    """CPU tier: the marshalling header + shim compile together under -Werror; the blobs are read on the host; without a CUDA
    device the program stops with exit code 3 before proving anything"""
    import torch
    exe = _build_blob_prove_check(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test below")
    _write_prove_blobs(tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 3 and "CUDA device is required" in p.stderr, p.stdout + p.stderr
    (tmp_path / "pk.bin").write_bytes((tmp_path / "pk.bin").read_bytes()[:-3])
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "malformed blob" in p.stderr


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="tests/cpp/blob_prove_check.cpp was written after the round's GPU minutes were spent and "
                                        "has not run on hardware yet")
def test_cpp_host_proves_from_the_reference_blobs(tmp_path):
    """r1cs_proving_key.bin -> proof.bin through C++ alone (vsp_crypto3_marshalling.hpp + vsp_crypto3_shim.hpp): the bytes equal
    the oracle's proof in the reference's compressed encoding"""
    exe = _build_blob_prove_check(tmp_path)
    codec, arr, P, A, r, s = _write_prove_blobs(tmp_path)
    p = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    want = cref.groth16_prove(fx.fill_pk(cref.ProvingKeyDesc(), arr), P, A, r, s)
    assert p.stdout.strip() == codec.proof_bytes(want).hex()
