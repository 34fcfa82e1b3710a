#!/usr/bin/env python3
"""Extract the reference's own known-answer vectors for the hot path into small JSON fixtures.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py
The fixtures it writes are committed; tests never read /root/reference at run time.

Sources (paths relative to /root/reference):
  libs/math/test/polynomial_dfs.cpp:50-113    size-8 Fr FFT / iFFT known answers
  libs/math/test/polynomial_dfs.cpp:387-581   polynomial_dfs multiplication (iFFT(k) -> pad -> FFT(2k) -> pointwise)
  libs/pubkey/test/bls.cpp:111-303            G1 / G2 points with blst-generated encodings, non-affine Jacobian points
  libs/pubkey/test/bls.cpp:669-998            ten G2 / ten G1 signatures and their aggregate (= all-ones MSM)
"""
import json
import os
import re
import sys

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def _strip_comments(s):
    return re.sub(r"//[^\n]*", "", s)


def _case_body(src, name):
    i = src.index("BOOST_AUTO_TEST_CASE(%s)" % name)
    j = src.find("BOOST_AUTO_TEST_CASE(", i + 10)
    return src[i: j if j > 0 else len(src)]


def _ints(text):
    """All integer literals in a brace list: 0x.._cppui253, 0x.., decimal."""
    out = []
    for tok in re.findall(r"0x[0-9a-fA-F]+|\b\d+\b", re.sub(r"_cppui\d+", "", text)):
        out.append(int(tok, 16) if tok.startswith("0x") else int(tok))
    return out


def _brace(text, start):
    """Return (content, end) of the balanced {...} starting at text[start] == '{'."""
    depth = 0
    for k in range(start, len(text)):
        if text[k] == "{":
            depth += 1
        elif text[k] == "}":
            depth -= 1
            if depth == 0:
                return text[start + 1: k], k + 1
    raise ValueError("unbalanced")


def _dfs_objects(body):
    """polynomial_dfs<...> NAME = { d, { v... } };  ->  {NAME: (d, [v...])}"""
    res = {}
    for m in re.finditer(r"polynomial_dfs<[^;]*?>\s+(\w+)\s*=\s*\{", body):
        content, _ = _brace(body, m.end() - 1)
        k = content.index("{")
        d = _ints(content[:k])[0]
        inner, _ = _brace(content, k)
        res[m.group(1)] = (d, _ints(inner))
    return res


def _vec_objects(body):
    res = {}
    for m in re.finditer(r"std::vector<[^;]*?>\s+(\w+)\s*=\s*\{", body):
        content, _ = _brace(body, m.end() - 1)
        res[m.group(1)] = _ints(content)
    return res


def fr_vectors():
    src = _strip_comments(open(f"{REF}/libs/math/test/polynomial_dfs.cpp").read())
    out = {"source": "libs/math/test/polynomial_dfs.cpp", "field": "bls12_fr<381>", "cases": []}

    b = _case_body(src, "polynomial_dfs_from_coefficients_less_degree")
    v = _vec_objects(b)
    out["cases"].append({"name": "from_coefficients_less_degree", "lines": "50-71", "op": "fft",
                         "n": 8, "in": [hex(x) for x in v["polynomial"]], "out": [hex(x) for x in v["c_res"]]})
    for nm, ln in (("polynomial_dfs_coefficients_less_degree", "77-93"),
                   ("polynomial_dfs_coefficients_same_degree", "95-111")):
        b = _case_body(src, nm)
        d = _dfs_objects(b)
        v = _vec_objects(b)
        out["cases"].append({"name": nm[len("polynomial_dfs_"):], "lines": ln, "op": "ifft", "n": len(d["a"][1]),
                             "in": [hex(x) for x in d["a"][1]], "out": [hex(x) for x in v["c_res"]]})
    for nm, ln in (("polynomial_dfs_multiplication_without_resize", "387-432"),
                   ("polynomial_dfs_multiplication_resize_a", "434-483"),
                   ("polynomial_dfs_multiplication_resize_b", "485-534"),
                   ("polynomial_dfs_multiplication_resize_both", "536-581")):
        b = _case_body(src, nm)
        d = _dfs_objects(b)
        out["cases"].append({"name": nm[len("polynomial_dfs_"):], "lines": ln, "op": "dfs_mul",
                             "a_deg": d["a"][0], "a": [hex(x) for x in d["a"][1]],
                             "b_deg": d["b"][0], "b": [hex(x) for x in d["b"][1]],
                             "c_deg": d["c_res"][0], "c": [hex(x) for x in d["c_res"][1]]})
    return out


def _dec_strings(text):
    """integral_type("123" "456") -> 123456, in order of appearance."""
    vals = []
    for m in re.finditer(r"integral_type\(((?:\s*\"\d+\")+)\s*\)", text):
        vals.append(int("".join(re.findall(r"\"(\d+)\"", m.group(1)))))
    return vals


def _u8_array(body, name):
    m = re.search(r"std::array<std::uint8_t,[^>]*>\s+%s\s*=\s*\{" % name, body)
    content, _ = _brace(body, m.end() - 1)
    return bytes(_ints(content)).hex()


def _signature_objects(body, g2):
    res = {}
    for m in re.finditer(r"signature_type\s+(etalon_\w+)\s*=\s*signature_type\(", body):
        j = body.index(";", m.end())
        vals = _dec_strings(body[m.end(): j])
        if g2:
            assert len(vals) == 4, (m.group(1), len(vals))
            res[m.group(1)] = {"x": [hex(vals[0]), hex(vals[1])], "y": [hex(vals[2]), hex(vals[3])]}
        else:
            assert len(vals) == 2, (m.group(1), len(vals))
            res[m.group(1)] = {"x": hex(vals[0]), "y": hex(vals[1])}
    return res


def curve_vectors():
    src = _strip_comments(open(f"{REF}/libs/pubkey/test/bls.cpp").read())
    out = {"source": "libs/pubkey/test/bls.cpp"}

    b = _case_body(src, "g1_serialization_test")
    pts = {}
    for m in re.finditer(r"group_value_type\s+(p\d)\s*=\s*group_value_type\(", b):
        j = b.index(";", m.end())
        chunk = b[m.end(): j]
        vals = _dec_strings(chunk)
        if len(vals) == 2:
            vals.append(1)
        pts[m.group(1)] = [hex(v) for v in vals]
    out["g1_serialization"] = {"lines": "111-196", "p1_affine": pts["p1"], "p3_jacobian": pts["p3"],
                               "p4_affine": pts["p4"],
                               "p4_octets": _u8_array(b, "etalon_serialization"),
                               "p4_octets_compressed": _u8_array(b, "etalon_serialization_comp")}

    b = _case_body(src, "g2_serialization_test")
    pts = {}
    for m in re.finditer(r"group_value_type\s+(p\d)\s*=\s*group_value_type\(", b):
        j = b.index(";", m.end())
        vals = _dec_strings(b[m.end(): j])
        if len(vals) == 4:
            vals += [1, 0]
        pts[m.group(1)] = [hex(v) for v in vals]  # x.c0 x.c1 y.c0 y.c1 z.c0 z.c1
    out["g2_serialization"] = {"lines": "198-303", "p1_affine": pts["p1"], "p3_jacobian": pts["p3"],
                               "p4_affine": pts["p4"],
                               "p4_octets": _u8_array(b, "etalon_serialization"),
                               "p4_octets_compressed": _u8_array(b, "etalon_serialization_comp")}

    b = _case_body(src, "bls_basic_mss")
    sigs = _signature_objects(b.replace("base_integral_type", "integral_type"), g2=False)
    out["g1_aggregate"] = {"lines": "858-998", "note": "etalon_agg_sig = sum(etalon_sig0..etalon_sig9)",
                           "standalone": sigs["etalon_sig"],
                           "points": [sigs["etalon_sig%d" % i] for i in range(10)],
                           "sum": sigs["etalon_agg_sig"]}
    b = _case_body(src, "bls_basic_mps")
    sigs = _signature_objects(b.replace("base_integral_type", "integral_type"), g2=True)
    out["g2_aggregate"] = {"lines": "669-856", "note": "etalon_agg_sig = sum(etalon_sig0..etalon_sig9); Fq2 as [c0, c1]",
                           "standalone": sigs["etalon_sig"],
                           "points": [sigs["etalon_sig%d" % i] for i in range(10)],
                           "sum": sigs["etalon_agg_sig"]}
    return out


def databin():
    """bin/cli/src/data.bin (1,072 bytes, the only crypto3-made proof in the tree): the blob share/tvm/voting_voter.sol:94
    hands to tvm.vergrth16 = proof | primary-input count | verification key.  Layout found by decoding it:
      [0, 192)      proof, compressed g_A | g_B | g_C           (marshalling .../r1cs_gg_ppzksnark/proof.hpp:63-71)
      [192, 196)    00 00 00 00: number of primary inputs (none)
      [196, 772)    alpha_g1_beta_g2: 12 Fq coefficients, 48 bytes each, LITTLE-endian, tower order
                    c0.c0.c0 | c0.c0.c1 | c0.c1.c0 | ... | c1.c2.c1   (verification_key.hpp:69)
      [772, 964)    gamma_g2 | delta_g2 compressed               (verification_key.hpp:71-73)
      [964, 1012)   gamma_ABC_g1.first compressed; [1012, 1072) zero: empty rest, padding
    Every point decodes on its curve and lies in the order-r subgroup, and (tests/test_oracle.py)
      e(g_A, g_B) = alpha_g1_beta_g2 * e(gamma_ABC.first, gamma_g2) * e(g_C, delta_g2)
    holds coefficient for coefficient when the reduced pairing is f^(3 (p^12 - 1) / r)."""
    with open(os.path.join(REF, "bin/cli/src/data.bin"), "rb") as f:
        d = f.read()
    assert len(d) == 1072 and d[192:196] == bytes(4) and not any(d[1012:])
    return {"source": "bin/cli/src/data.bin", "proof": d[0:192].hex(), "n_primary": 0,
            "alpha_g1_beta_g2": ["%096x" % int.from_bytes(d[196 + 48 * i:196 + 48 * (i + 1)], "little") for i in range(12)],
            "gamma_g2": d[772:868].hex(), "delta_g2": d[868:964].hex(), "gamma_ABC_first": d[964:1012].hex()}


def main():
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    with open(os.path.join(OUT, "databin_groth16.json"), "w") as f:
        json.dump(databin(), f, indent=1)
    with open(os.path.join(OUT, "fr_fft_kat.json"), "w") as f:
        json.dump(fr_vectors(), f, indent=1)
    with open(os.path.join(OUT, "curve_kat.json"), "w") as f:
        json.dump(curve_vectors(), f, indent=1)
    print("wrote fr_fft_kat.json, curve_kat.json, databin_groth16.json")


if __name__ == "__main__":
    main()
