// shim_check — drives the crypto3-named entry points of include/vsp_crypto3_shim.hpp the way
// process_encrypted_input_mode_vote_phase does (bin/cli/include/nil/vote_saver/common.hpp:1132-1145) and compares
// every result with the expected values in a fixture written by tests/test_gpu_shim.py from the CPU oracle.
//
// fixture: records of  u32 name_len | name | u64 count | count x u64
#include <cstdio>
#include <fstream>
#include <iostream>
#include <map>
#include <string>

#include "vsp_crypto3_shim.hpp"

using namespace nil::crypto3;
typedef algebra::curves::bls12<381> curve_type;
typedef curve_type::scalar_field_type fr_type;
typedef b200::fr_value fr;
typedef b200::g1_value g1;
typedef b200::g2_value g2;
typedef pubkey::elgamal_verifiable<curve_type> scheme_type;
typedef scheme_type::proof_system_type proof_system_type;

static std::map<std::string, std::vector<std::uint64_t>> F;
static int failures = 0;

static void load(const char *path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    for (;;) {
        std::uint32_t nl;
        if (!f.read(reinterpret_cast<char *>(&nl), 4)) break;
        std::string name(nl, '\0');
        f.read(&name[0], nl);
        std::uint64_t cnt;
        f.read(reinterpret_cast<char *>(&cnt), 8);
        std::vector<std::uint64_t> v(cnt);
        f.read(reinterpret_cast<char *>(v.data()), cnt * 8);
        F[name] = std::move(v);
    }
}
template<typename T>
static std::vector<T> vec(const std::string &name) {
    const auto &v = F.at(name);
    const std::size_t L = sizeof(T) / 8;
    std::vector<T> out(v.size() / L);
    for (std::size_t i = 0; i < out.size(); ++i) std::memcpy(out[i].l.data(), &v[i * L], sizeof(T));
    return out;
}
template<typename T>
static T one(const std::string &name) {
    return vec<T>(name).at(0);
}
static void expect(bool ok, const char *what) {
    std::printf("%-58s %s\n", what, ok ? "ok" : "MISMATCH");
    if (!ok) ++failures;
}
template<typename F2>
static void expect_invalid(F2 fn, const char *what) {
    bool threw = false;
    try {
        fn();
    } catch (const std::invalid_argument &) {
        threw = true;
    }
    expect(threw, what);
}

int main(int argc, char **argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: shim_check fixture.bin\n");
        return 2;
    }
    try {
        load(argv[1]);
        // ---- math::make_evaluation_domain / fft / inverse_fft / multiply_by_coset ----------------------------------
        {
            auto a = vec<fr>("fft_in");
            auto dom = math::make_evaluation_domain<fr_type>(F.at("fft_m")[0]);
            expect(dom->size() == F.at("fft_m_out")[0], "make_evaluation_domain size");
            auto b = a;
            dom->fft(b);
            expect(b == vec<fr>("fft_out"), "evaluation_domain::fft");
            dom->inverse_fft(b);
            auto padded = a;
            padded.resize(dom->size());
            expect(b == padded, "inverse_fft(fft(a)) == a");
            auto c = a;
            math::multiply_by_coset(c, one<fr>("coset_g"));
            expect(c == vec<fr>("coset_out"), "multiply_by_coset");
            std::vector<fr> too_long(dom->size() + 1);
            expect_invalid([&] { dom->fft(too_long); }, "fft on a.size() > m throws invalid_argument");
            expect_invalid([&] { math::make_evaluation_domain<fr_type>(1); }, "make_evaluation_domain(1) throws");
        }
        // ---- algebra::multiexp ------------------------------------------------------------------------------------------
        {
            auto bases = vec<g1>("msm_bases");
            auto scalars = vec<fr>("msm_scalars");
            g1 r = algebra::multiexp<algebra::policies::multiexp_method_BDLO12>(bases.begin(), bases.end(), scalars.begin(),
                                                                                scalars.end(), 1);
            expect(r == one<g1>("msm_out"), "algebra::multiexp<BDLO12> (G1)");
            auto bases2 = vec<g2>("msm2_bases");
            auto scalars2 = vec<fr>("msm2_scalars");
            g2 r2 = algebra::multiexp_with_mixed_addition<algebra::policies::multiexp_method_BDLO12>(
                bases2.begin(), bases2.end(), scalars2.begin(), scalars2.end(), 4);
            expect(r2 == one<g2>("msm2_out"), "algebra::multiexp_with_mixed_addition<BDLO12> (G2)");
        }
        // ---- keys -------------------------------------------------------------------------------------------------------
        proof_system_type::keypair_type keypair;
        auto &pk = keypair.first;
        auto &vk = keypair.second;
        {
            auto &cs = pk.constraint_system;
            cs.primary_input_size = F.at("num_inputs")[0];
            cs.auxiliary_input_size = F.at("num_aux")[0];
            const std::size_t nc = F.at("num_constraints")[0];
            cs.constraints.resize(nc);
            const char *mn[3] = {"a", "b", "c"};
            for (int k = 0; k < 3; ++k) {
                const auto &row = F.at(std::string(mn[k]) + "_row"), &col = F.at(std::string(mn[k]) + "_col");
                auto val = vec<fr>(std::string(mn[k]) + "_val");
                for (std::size_t i = 0; i < nc; ++i)
                    for (std::uint64_t q = row[i]; q < row[i + 1]; ++q) {
                        zk::snark::linear_term<fr_type> t {static_cast<std::size_t>(col[q]), val[q]};
                        (k == 0 ? cs.constraints[i].a : k == 1 ? cs.constraints[i].b : cs.constraints[i].c).push_back(t);
                    }
            }
            pk.alpha_g1 = one<g1>("alpha_g1");
            pk.beta_g1 = one<g1>("beta_g1");
            pk.beta_g2 = one<g2>("beta_g2");
            pk.delta_g1 = one<g1>("delta_g1");
            pk.delta_g2 = one<g2>("delta_g2");
            pk.A_query = vec<g1>("A_query");
            pk.H_query = vec<g1>("H_query");
            pk.L_query = vec<g1>("L_query");
            auto bg1 = vec<g1>("B_g1");
            auto bg2 = vec<g2>("B_g2");
            const auto &bidx = F.at("B_idx");
            for (std::size_t i = 0; i < F.at("B_count")[0]; ++i) {
                pk.B_query.indices.push_back(bidx[i]);
                pk.B_query.values.push_back({bg2[i], bg1[i]});
            }
            vk.delta_g2 = pk.delta_g2;
            vk.gamma_ABC_g1.rest = vec<g1>("gamma_abc_rest");
        }
        pubkey::public_key<scheme_type> pub;
        pub.delta_g1 = one<g1>("pub_delta_g1");
        pub.delta_s_g1 = vec<g1>("pub_delta_s_g1");
        pub.t_g1 = vec<g1>("pub_t_g1");
        pub.t_g2 = vec<g2>("pub_t_g2");
        pub.delta_sum_s_g1 = one<g1>("pub_delta_sum_s_g1");
        pub.gamma_inverse_sum_s_g1 = one<g1>("pub_gamma_inverse_sum_s_g1");
        auto primary = vec<fr>("primary"), aux = vec<fr>("aux");
        const fr r = one<fr>("r"), s = one<fr>("s"), r_saver = one<fr>("r_saver");
        // ---- zk::prove, basic and encrypted_input ---------------------------------------------------------------------------
        {
            typedef zk::snark::r1cs_gg_ppzksnark<curve_type> basic_system;
            auto p = zk::prove<basic_system>(pk, primary, aux, r, s);
            auto want = F.at("proof_basic");
            expect(!std::memcmp(p.g_A.l.data(), &want[0], 96) && !std::memcmp(p.g_B.l.data(), &want[12], 192) &&
                       !std::memcmp(p.g_C.l.data(), &want[36], 96),
                   "zk::prove<r1cs_gg_ppzksnark> (basic mode)");
            auto q = zk::prove<proof_system_type>(pk, pub, primary, aux, r_saver, r, s);
            auto want2 = F.at("proof_enc");
            expect(!std::memcmp(q.g_A.l.data(), &want2[0], 96) && !std::memcmp(q.g_B.l.data(), &want2[12], 192) &&
                       !std::memcmp(q.g_C.l.data(), &want2[36], 96),
                   "zk::prove<..., encrypted_input>(pk, pubkey, ..., r)");
            auto p2 = zk::prove<basic_system>(pk, primary, aux);    // internally drawn r, s: g_A, g_B, g_C all change
            expect(!(p2.g_A == p.g_A) && !p2.g_A.is_zero(), "zk::prove with internally drawn blinding differs");
        }
        // ---- pubkey::encrypt / rerandomize, as common.hpp:1132-1145 calls them -----------------------------------------------
        {
            std::vector<fr> m_field(primary.begin(), primary.begin() + pub.delta_s_g1.size());
            scheme_type::cipher_type ct = pubkey::encrypt<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>>(
                m_field, {r_saver, pub, keypair, primary, aux, &r, &s});
            expect(ct.first == vec<g1>("ct"), "pubkey::encrypt cipher text (msg_size + 2 G1)");
            auto want = F.at("proof_enc");
            expect(!std::memcmp(ct.second.g_A.l.data(), &want[0], 96) && !std::memcmp(ct.second.g_C.l.data(), &want[36], 96),
                   "pubkey::encrypt proof");
            auto rnd = vec<fr>("rnd");
            scheme_type::cipher_type rr = pubkey::rerandomize<scheme_type>(rnd, ct.first, {pub, keypair, ct.second});
            auto wantp = F.at("rerand_proof");
            expect(rr.first == vec<g1>("rerand_ct"), "pubkey::rerandomize cipher text");
            expect(!std::memcmp(rr.second.g_A.l.data(), &wantp[0], 96) && !std::memcmp(rr.second.g_B.l.data(), &wantp[12], 192) &&
                       !std::memcmp(rr.second.g_C.l.data(), &wantp[36], 96),
                   "pubkey::rerandomize proof");
            auto bad = m_field;
            bad[0].l[0] ^= 1;
            expect_invalid(
                [&] {
                    pubkey::encrypt<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>>(
                        bad, {r_saver, pub, keypair, primary, aux, &r, &s});
                },
                "encrypt: plain text not a prefix of primary input throws");
        }
        // ---- zk::generate / pubkey::generate_keypair, as common.hpp:916-931 calls them ------------------------------------------
        if (F.count("kg_num_constraints")) {
            zk::snark::r1cs_constraint_system<fr_type> cs;
            cs.primary_input_size = F.at("kg_num_inputs")[0];
            cs.auxiliary_input_size = F.at("kg_num_aux")[0];
            const std::size_t nc = F.at("kg_num_constraints")[0];
            cs.constraints.resize(nc);
            const char *mn[3] = {"a", "b", "c"};
            for (int k = 0; k < 3; ++k) {
                const auto &row = F.at(std::string("kg_") + mn[k] + "_row"), &col = F.at(std::string("kg_") + mn[k] + "_col");
                auto val = vec<fr>(std::string("kg_") + mn[k] + "_val");
                for (std::size_t i = 0; i < nc; ++i)
                    for (std::uint64_t q = row[i]; q < row[i + 1]; ++q) {
                        zk::snark::linear_term<fr_type> t {static_cast<std::size_t>(col[q]), val[q]};
                        (k == 0 ? cs.constraints[i].a : k == 1 ? cs.constraints[i].b : cs.constraints[i].c).push_back(t);
                    }
            }
            auto tox = vec<fr>("kg_toxic");
            auto kp = zk::generate<proof_system_type>(cs, {tox[0], tox[1], tox[2], tox[3], tox[4]});
            expect(kp.first.alpha_g1 == one<g1>("kg_alpha_g1") && kp.first.beta_g1 == one<g1>("kg_beta_g1") &&
                       kp.first.beta_g2 == one<g2>("kg_beta_g2") && kp.first.delta_g1 == one<g1>("kg_delta_g1") &&
                       kp.first.delta_g2 == one<g2>("kg_delta_g2"),
                   "zk::generate alpha / beta / delta");
            expect(kp.first.A_query == vec<g1>("kg_A_query"), "zk::generate A_query");
            expect(kp.first.H_query == vec<g1>("kg_H_query"), "zk::generate H_query");
            expect(kp.first.L_query == vec<g1>("kg_L_query"), "zk::generate L_query");
            bool b_ok = kp.first.B_query.indices.size() == F.at("kg_B_count")[0];
            auto bg1 = vec<g1>("kg_B_g1");
            auto bg2 = vec<g2>("kg_B_g2");
            for (std::size_t i = 0; b_ok && i < kp.first.B_query.indices.size(); ++i)
                b_ok = kp.first.B_query.indices[i] == F.at("kg_B_idx")[i] && kp.first.B_query.values[i].g == bg2[i] &&
                       kp.first.B_query.values[i].h == bg1[i];
            expect(b_ok, "zk::generate B_query (sparse)");
            auto gabc = vec<g1>("kg_gamma_ABC_g1");
            expect(kp.second.gamma_ABC_g1.first == gabc[0] &&
                       kp.second.gamma_ABC_g1.rest == std::vector<g1>(gabc.begin() + 1, gabc.end()) &&
                       kp.second.gamma_g2 == one<g2>("kg_gamma_g2") && kp.second.gamma_g1 == one<g1>("kg_gamma_g1"),
                   "zk::generate verification key");
            auto rnd = vec<fr>("kg_rnd");
            const std::size_t msg = F.at("kg_msg_size")[0];
            auto keys = pubkey::generate_keypair<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>>(rnd, {kp, msg});
            const auto &spub = std::get<0>(keys);
            expect(spub.delta_g1 == kp.second.delta_g1 && spub.delta_s_g1 == vec<g1>("kg_pub_delta_s_g1") &&
                       spub.t_g1 == vec<g1>("kg_pub_t_g1") && spub.t_g2 == vec<g2>("kg_pub_t_g2") &&
                       spub.delta_sum_s_g1 == one<g1>("kg_pub_delta_sum_s_g1") &&
                       spub.gamma_inverse_sum_s_g1 == one<g1>("kg_pub_gamma_inverse_sum_s_g1"),
                   "pubkey::generate_keypair public key");
            expect(std::get<1>(keys).rho == rnd[0], "pubkey::generate_keypair private key");
            expect(std::get<2>(keys).rho_g2 == one<g2>("kg_rho_g2") && std::get<2>(keys).rho_sv_g2 == vec<g2>("kg_rho_sv_g2") &&
                       std::get<2>(keys).rho_rhov_g2 == vec<g2>("kg_rho_rhov_g2"),
                   "pubkey::generate_keypair verification key");
            // the rest of the protocol on those keys (common.hpp:1132-1168, 1221-1223, 1282-1283): encrypt, verify_encryption,
            // decrypt, verify_decryption
            {
                auto kprim = vec<fr>("kg_primary"), kaux = vec<fr>("kg_aux");
                std::vector<fr> m_field(kprim.begin(), kprim.begin() + msg);
                const fr rs = one<fr>("r_saver");
                auto ct = pubkey::encrypt<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>>(
                    m_field, {rs, spub, kp, kprim, kaux, nullptr, nullptr});
                proof_system_type::primary_input_type rest(kprim.begin() + msg, kprim.end());
                expect(pubkey::verify_encryption<scheme_type>(ct.first, {spub, kp.second, ct.second, rest}),
                       "pubkey::verify_encryption accepts the ballot");
                auto bad_ct = ct.first;
                bad_ct[1] = bad_ct[2];
                expect(!pubkey::verify_encryption<scheme_type>(bad_ct, {spub, kp.second, ct.second, rest}),
                       "pubkey::verify_encryption rejects a tampered cipher text");
                typedef zk::snark::r1cs_gg_ppzksnark<curve_type> basic_system;
                auto bp = zk::prove<basic_system>(kp.first, kprim, kaux);
                expect(zk::verify<basic_system>(kp.second, kprim, bp), "zk::verify accepts a basic-mode proof");
                auto dec = pubkey::decrypt<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>, 4>(
                    ct.first, {std::get<1>(keys), std::get<2>(keys), kp});
                expect(dec.first == m_field, "pubkey::decrypt recovers the plain text");
                expect(pubkey::verify_decryption<scheme_type>(ct.first, dec.first, {std::get<2>(keys), kp, dec.second}),
                       "pubkey::verify_decryption accepts");
                auto wrong = dec.first;
                wrong[0].l[0] ^= 1;
                expect(!pubkey::verify_decryption<scheme_type>(ct.first, wrong, {std::get<2>(keys), kp, dec.second}),
                       "pubkey::verify_decryption rejects a wrong plain text");
            }
            rnd.pop_back();
            expect_invalid(
                [&] {
                    pubkey::generate_keypair<scheme_type, pubkey::modes::verifiable_encryption<scheme_type>>(rnd, {kp, msg});
                },
                "generate_keypair: too few random numbers throws");
        }
    } catch (const std::exception &e) {
        std::fprintf(stderr, "shim_check: exception: %s\n", e.what());
        return 3;
    }
    std::printf("%s (%d mismatches)\n", failures ? "FAILED" : "all shim checks passed", failures);
    return failures ? 1 : 0;
}
