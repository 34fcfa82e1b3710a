// CPU tier (built and run by tests/test_codec.py): include/vsp_crypto3_marshalling.hpp on blobs written by the Python codec —
// every deserialiser of bin/cli's marshaling_policy into the shim's types, every serialiser back, byte for byte; the
// structure of what was read (sizes, the GT flag, the sparse B_query) checked on the way.  No GPU call is made.
// usage: marshalling_check <directory with pk.bin vk.bin pub.bin svk.bin sk.bin ct.bin pi.bin proof.bin tree.bin>
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>

#include "vsp_crypto3_marshalling.hpp"

using namespace nil::crypto3;
typedef b200::marshaling_policy mp;

static int failures = 0;
static void expect(bool ok, const char *what) {
    if (!ok) {
        std::printf("FAIL %s\n", what);
        ++failures;
    }
}
static mp::blob_type load(const std::string &dir, const char *name) {
    std::ifstream f(dir + "/" + name, std::ios::binary);
    if (!f) {
        std::printf("FAIL cannot open %s\n", name);
        ++failures;
    }
    return mp::blob_type(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
}

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    const std::string dir = argv[1];
    try {
        const auto pk_blob = load(dir, "pk.bin");
        const auto pk = mp::deserialize_pk_crs(pk_blob);
        expect(pk.A_query.size() == pk.constraint_system.num_variables() + 1, "A_query covers 1 || primary || auxiliary");
        expect(pk.L_query.size() == pk.constraint_system.auxiliary_input_size, "L_query covers the auxiliary input");
        expect(pk.B_query.indices.size() == pk.B_query.values.size() && pk.B_query.domain_size_ == pk.A_query.size(), "B_query");
        expect(pk.constraint_system.num_constraints() == 12 && pk.constraint_system.primary_input_size == 3, "constraint system sizes");
        expect(mp::serialize_pk_crs(pk) == pk_blob, "proving key round trip");

        const auto vk_blob = load(dir, "vk.bin");
        const auto vk = mp::deserialize_vk_crs(vk_blob);
        expect(vk.has_alpha_g1_beta_g2 && vk.gamma_ABC_g1.rest.size() == 7, "verification key: GT form, 7 rest points");
        expect(mp::serialize_vk_crs(vk) == vk_blob, "verification key round trip");
        zk::snark::r1cs_gg_ppzksnark_verification_key<algebra::curves::bls12<381>> no_gt = vk;
        no_gt.has_alpha_g1_beta_g2 = false;
        bool threw = false;
        try {
            mp::serialize_vk_crs(no_gt);
        } catch (const std::invalid_argument &) {
            threw = true;
        }
        expect(threw, "a key without its GT element cannot be serialised");

        const auto pub_blob = load(dir, "pub.bin");
        const auto pub = mp::deserialize_pk_eid(pub_blob);
        expect(pub.delta_s_g1.size() == 5 && pub.t_g1.size() == 5 && pub.t_g2.size() == 6, "public key list sizes");
        expect(mp::serialize_pk_eid(pub) == pub_blob, "public key round trip");
        const auto svk_blob = load(dir, "svk.bin");
        const auto svk = mp::deserialize_vk_eid(svk_blob);
        expect(svk.rho_sv_g2.size() == 5 && svk.rho_rhov_g2.size() == 5, "SAVER verification key list sizes");
        expect(mp::serialize_vk_eid(svk) == svk_blob, "SAVER verification key round trip");
        const auto sk_blob = load(dir, "sk.bin");
        expect(mp::serialize_sk_eid(mp::deserialize_sk_eid(sk_blob)) == sk_blob, "secret key round trip");

        const auto ct_blob = load(dir, "ct.bin");
        const auto ct = mp::deserialize_ct(ct_blob);
        expect(ct.size() == 7 && mp::serialize_ct(ct) == ct_blob, "cipher text round trip");
        const auto pi_blob = load(dir, "pi.bin");
        const auto pi = mp::deserialize_scalar_vector(pi_blob);
        expect(pi.size() == 3 && pi[0].l[0] == 1 && pi[2].l[0] == 3 && mp::serialize_scalar_vector(pi) == pi_blob, "scalar vector round trip");
        const auto proof_blob = load(dir, "proof.bin");
        const auto proof = mp::deserialize_proof(proof_blob);
        expect(mp::serialize_proof(proof) == proof_blob, "proof round trip");

        // the bundles: serialize_data and serialize_tally_phase_data give the single blobs' bytes
        mp::blob_type a, b, c, d;
        mp::serialize_data(0, proof, pi, ct, pi, a, b, c, d);
        expect(a == proof_blob && b == pi_blob && c == ct_blob && d == pi_blob, "serialize_data");
        mp::serialize_tally_phase_data({pi, ct[0]}, a, b);
        expect(b == pi_blob && a.size() == 48 && mp::deserialize_decryption_proof(a) == ct[0], "serialize_tally_phase_data");
        mp::blob_type k1, k2, k3, k4, k5;
        mp::serialize_initial_phase_admin_keys(pk, vk, pub, mp::deserialize_sk_eid(sk_blob), svk, k1, k2, k3, k4, k5);
        expect(k1 == pk_blob && k2 == vk_blob && k3 == pub_blob && k4 == sk_blob && k5 == svk_blob, "serialize_initial_phase_admin_keys");

        // bit strings, packing, the voters' tree and a copath
        std::array<bool, 12> bits = {true, false, false, false, false, false, false, true, false, true, false, false};
        const auto bb = mp::serialize_bitarray<12>(bits);
        expect(bb.size() == 2 && bb[0] == 0x81 && bb[1] == 0x40, "serialize_bitarray");
        expect(mp::deserialize_bitarray<12>(bb.begin(), bb.end()) == bits, "deserialize_bitarray");
        std::vector<bool> many(255, false);
        many[0] = many[253] = many[254] = true;
        const auto packed = mp::get_multi_field_element_from_bits(many);
        expect(packed.size() == 2 && packed[0].l[0] == 1 && packed[0].l[3] == (std::uint64_t(1) << 61) && packed[1].l[0] == 1, "254-bit packing");
        const auto tree_blob = load(dir, "tree.bin");
        const auto tree = mp::deserialize_merkle_tree(2, tree_blob);
        const auto path = mp::merkle_copath(tree, 2, 2);
        expect(tree.size() == 7 && path.size() == 2 && path[0] == tree[3] && path[1] == tree[4], "tree and copath of leaf 2");

        // malformed input is an exception, not a crash
        threw = false;
        try {
            mp::deserialize_pk_eid(mp::blob_type(pub_blob.begin(), pub_blob.end() - 1));
        } catch (const std::invalid_argument &) {
            threw = true;
        }
        expect(threw, "a truncated public key throws");
    } catch (const std::exception &e) {
        std::printf("FAIL exception: %s\n", e.what());
        return 1;
    }
    std::printf(failures ? "FAILED %d\n" : "OK%.0d\n", failures);
    return failures ? 1 : 0;
}
