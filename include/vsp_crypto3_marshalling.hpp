// vsp_crypto3_marshalling.hpp — the (de)serialisers of bin/cli's `marshaling_policy`
// (/root/reference/bin/cli/include/nil/vote_saver/common.hpp:168-800) over the shim's value types
// (vsp_crypto3_shim.hpp), by the reference's names and argument order, for a C++ host that does not carry crypto3's
// marshalling library: every blob the process_encrypted_input_mode_* phases read or write, decoded straight into the
// types zk::prove / pubkey::encrypt / rerandomize / verify_encryption / decrypt / verify_decryption of the shim take.
// Header-only over the C ABI's vsp_codec_* entry points (host-only code, no GPU): tests/cpp/marshalling_check.cpp runs on
// the CPU tier.  A maintainer who links the shim under the real bin/cli keeps crypto3's own marshalling and does not need
// this file.
//
// Formats: include/vsp_b200.h, "wire codecs" (big-endian fixed-width integers, 8-byte count prefixes, compressed points;
// libs/marshalling/zk/include/nil/crypto3/marshalling/zk/types/r1cs_gg_ppzksnark/*.hpp, .../pubkey/types/elgamal_verifiable.hpp).
#ifndef VSP_CRYPTO3_MARSHALLING_HPP
#define VSP_CRYPTO3_MARSHALLING_HPP

#include <array>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <utility>
#include <vector>

#include "vsp_crypto3_shim.hpp"

namespace nil {
namespace crypto3 {
namespace b200 {

    struct marshaling_policy {
        typedef algebra::curves::bls12<381> curve_type;
        typedef pubkey::elgamal_verifiable<curve_type> encryption_scheme_type;
        typedef encryption_scheme_type::proof_system_type proof_system;
        typedef fr_value scalar_field_value_type;
        typedef proof_system::proof_type proof_type;
        typedef proof_system::verification_key_type verification_key_type;
        typedef proof_system::proving_key_type proving_key_type;
        typedef proof_system::primary_input_type primary_input_type;
        typedef pubkey::public_key<encryption_scheme_type> elgamal_public_key_type;
        typedef pubkey::private_key<encryption_scheme_type> elgamal_private_key_type;
        typedef pubkey::verification_key<encryption_scheme_type> elgamal_verification_key_type;
        typedef std::vector<std::uint8_t> blob_type;

        static void check(int rc, const char *what) {
            if (rc != VSP_OK) throw std::invalid_argument(std::string(what) + ": malformed blob");
        }

        // ---- scalars, bit strings ------------------------------------------------------------------------------------
        // deserialize_scalar_vector (common.hpp:526-532): count | Fr...
        static std::vector<scalar_field_value_type> deserialize_scalar_vector(const blob_type &blob) {
            std::uint64_t n = 0;
            check(vsp_codec_primary_input_read(blob.data(), blob.size(), &n, nullptr, 0), "deserialize_scalar_vector");
            std::vector<scalar_field_value_type> out(n);
            check(vsp_codec_primary_input_read(blob.data(), blob.size(), &n, n ? out[0].l.data() : nullptr, n),
                  "deserialize_scalar_vector");
            return out;
        }
        static blob_type serialize_scalar_vector(const std::vector<scalar_field_value_type> &v) {
            blob_type out(8 + 32 * v.size());
            check(vsp_codec_primary_input_write(v.empty() ? nullptr : v[0].l.data(), v.size(), out.data()), "serialize_scalar_vector");
            return out;
        }
        // get_multi_field_element_from_bits (common.hpp:557-570)
        static std::vector<scalar_field_value_type> get_multi_field_element_from_bits(const std::vector<bool> &bits) {
            std::vector<std::uint8_t> b(bits.begin(), bits.end());
            std::vector<scalar_field_value_type> out((bits.size() + 253) / 254);
            check(vsp_codec_bits_to_field_elements(b.data(), b.size(), out.empty() ? nullptr : out[0].l.data()),
                  "get_multi_field_element_from_bits");
            return out;
        }
        // serialize_bitarray / deserialize_bitarray (common.hpp:574-612)
        template<std::size_t Bits>
        static blob_type serialize_bitarray(const std::array<bool, Bits> &bitarray) {
            std::uint8_t b[Bits ? Bits : 1];
            for (std::size_t i = 0; i < Bits; ++i) b[i] = bitarray[i];
            blob_type out((Bits + 7) / 8);
            check(vsp_codec_bitarray_write(b, Bits, out.data()), "serialize_bitarray");
            return out;
        }
        template<std::size_t Bits, typename InputIterator>
        static std::array<bool, Bits> deserialize_bitarray(InputIterator begin, InputIterator end) {
            const blob_type in(begin, end);
            std::uint8_t b[Bits ? Bits : 1];
            check(vsp_codec_bitarray_read(in.data(), in.size(), Bits, b), "deserialize_bitarray");
            std::array<bool, Bits> out {};
            for (std::size_t i = 0; i < Bits; ++i) out[i] = b[i] != 0;
            return out;
        }
        // deserialize_merkle_tree (common.hpp:614-636): digests as 4-limb integers (bit i = digest bit i), leaves first, root
        // last; merkle_copath = containers::merkle_proof(tree, idx) for arity 2 (merkle/proof.hpp:65-91), leaf level first
        static std::vector<std::array<std::uint64_t, 4>> deserialize_merkle_tree(std::size_t tree_depth, const blob_type &blob) {
            if (tree_depth > 40 || blob.size() != 32 * ((std::size_t(2) << tree_depth) - 1))
                throw std::invalid_argument("deserialize_merkle_tree: the blob is not 2^(depth + 1) - 1 digests of 32 bytes");
            std::vector<std::array<std::uint64_t, 4>> out((std::size_t(2) << tree_depth) - 1);
            check(vsp_codec_merkle_tree_read(blob.data(), blob.size(), tree_depth, out[0].data()), "deserialize_merkle_tree");
            return out;
        }
        static std::vector<std::array<std::uint64_t, 4>> merkle_copath(const std::vector<std::array<std::uint64_t, 4>> &tree,
                                                                       std::size_t tree_depth, std::size_t leaf_idx) {
            if (tree_depth > 40 || tree.size() != (std::size_t(2) << tree_depth) - 1)
                throw std::invalid_argument("merkle_copath: tree size");
            std::vector<std::array<std::uint64_t, 4>> out(tree_depth);
            check(vsp_codec_merkle_copath(tree[0].data(), tree_depth, leaf_idx, tree_depth ? out[0].data() : nullptr), "merkle_copath");
            return out;
        }

        // ---- Groth16 ------------------------------------------------------------------------------------------------
        // deserialize_pk_crs (common.hpp:749-753): r1cs_gg_ppzksnark_fast_proving_key
        static proving_key_type deserialize_pk_crs(const blob_type &blob) {
            vsp_fast_pk_sizes_t sz;
            check(vsp_codec_fast_pk_sizes(blob.data(), blob.size(), &sz), "deserialize_pk_crs");
            const std::size_t nc = sz.num_constraints;
            std::vector<std::uint64_t> row[3], val[3];
            std::vector<std::uint32_t> col[3];
            const std::uint64_t nnz[3] = {sz.nnz_a, sz.nnz_b, sz.nnz_c};
            for (int k = 0; k < 3; ++k) {
                row[k].assign(nc + 1, 0);
                col[k].assign(nnz[k] + 1, 0);
                val[k].assign(4 * nnz[k] + 4, 0);
            }
            proving_key_type pk;
            pk.A_query.resize(sz.A_count);
            pk.H_query.resize(sz.H_count);
            pk.L_query.resize(sz.L_count);
            std::vector<std::uint32_t> b_idx(sz.B_count + 1);
            std::vector<g2_value> b_g2(sz.B_count + 1);
            std::vector<g1_value> b_g1(sz.B_count + 1);
            g1_value spare;  // target of an empty list's pointer
            vsp_proving_key_t d;
            std::memset(&d, 0, sizeof(d));
            d.cs.a_row = row[0].data(); d.cs.b_row = row[1].data(); d.cs.c_row = row[2].data();
            d.cs.a_col = col[0].data(); d.cs.b_col = col[1].data(); d.cs.c_col = col[2].data();
            d.cs.a_val = val[0].data(); d.cs.b_val = val[1].data(); d.cs.c_val = val[2].data();
            d.alpha_g1 = pk.alpha_g1.l.data(); d.beta_g1 = pk.beta_g1.l.data(); d.beta_g2 = pk.beta_g2.l.data();
            d.delta_g1 = pk.delta_g1.l.data(); d.delta_g2 = pk.delta_g2.l.data();
            d.A_query = pk.A_query.empty() ? spare.l.data() : pk.A_query[0].l.data();
            d.B_idx = b_idx.data();
            d.B_g2 = b_g2[0].l.data();
            d.B_g1 = b_g1[0].l.data();
            d.H_query = pk.H_query.empty() ? spare.l.data() : pk.H_query[0].l.data();
            d.L_query = pk.L_query.empty() ? spare.l.data() : pk.L_query[0].l.data();
            check(vsp_codec_fast_pk_read(blob.data(), blob.size(), &d), "deserialize_pk_crs");
            pk.B_query.indices.assign(b_idx.begin(), b_idx.begin() + sz.B_count);
            pk.B_query.values.resize(sz.B_count);
            for (std::size_t i = 0; i < sz.B_count; ++i) {
                pk.B_query.values[i].g = b_g2[i];
                pk.B_query.values[i].h = b_g1[i];
            }
            pk.B_query.domain_size_ = sz.B_domain_size;
            auto &cs = pk.constraint_system;
            cs.primary_input_size = sz.num_inputs;
            cs.auxiliary_input_size = sz.num_aux;
            cs.constraints.resize(nc);
            for (std::size_t i = 0; i < nc; ++i) {
                std::vector<zk::snark::linear_term<curve_type::scalar_field_type>> *lc[3] = {&cs.constraints[i].a, &cs.constraints[i].b,
                                                                                              &cs.constraints[i].c};
                for (int k = 0; k < 3; ++k)
                    for (std::uint64_t q = row[k][i]; q < row[k][i + 1]; ++q) {
                        zk::snark::linear_term<curve_type::scalar_field_type> t;
                        t.index = col[k][q];
                        std::memcpy(t.coeff.l.data(), &val[k][4 * q], 32);
                        lc[k]->push_back(t);
                    }
            }
            return pk;
        }
        static blob_type serialize_pk_crs(const proving_key_type &pk) {
            const auto &cs = pk.constraint_system;
            const std::size_t nc = cs.num_constraints();
            std::vector<std::uint64_t> row[3], val[3];
            std::vector<std::uint32_t> col[3];
            for (int k = 0; k < 3; ++k) row[k].assign(nc + 1, 0);
            for (std::size_t i = 0; i < nc; ++i) {
                const std::vector<zk::snark::linear_term<curve_type::scalar_field_type>> *lc[3] = {
                    &cs.constraints[i].a, &cs.constraints[i].b, &cs.constraints[i].c};
                for (int k = 0; k < 3; ++k) {
                    for (const auto &t : *lc[k]) {
                        col[k].push_back(static_cast<std::uint32_t>(t.index));
                        val[k].insert(val[k].end(), t.coeff.l.begin(), t.coeff.l.end());
                    }
                    row[k][i + 1] = col[k].size();
                }
            }
            if (pk.A_query.size() != cs.num_variables() + 1 || pk.L_query.size() != cs.auxiliary_input_size)
                throw std::invalid_argument("serialize_pk_crs: A_query / L_query sizes do not match the constraint system");
            std::vector<std::uint32_t> b_idx(pk.B_query.indices.begin(), pk.B_query.indices.end());
            std::vector<g2_value> b_g2(pk.B_query.values.size());
            std::vector<g1_value> b_g1(pk.B_query.values.size());
            for (std::size_t i = 0; i < b_g2.size(); ++i) {
                b_g2[i] = pk.B_query.values[i].g;
                b_g1[i] = pk.B_query.values[i].h;
            }
            vsp_proving_key_t d;
            std::memset(&d, 0, sizeof(d));
            d.cs.num_inputs = cs.primary_input_size;
            d.cs.num_aux = cs.auxiliary_input_size;
            d.cs.num_constraints = nc;
            d.cs.a_row = row[0].data(); d.cs.b_row = row[1].data(); d.cs.c_row = row[2].data();
            d.cs.a_col = col[0].data(); d.cs.b_col = col[1].data(); d.cs.c_col = col[2].data();
            d.cs.a_val = val[0].data(); d.cs.b_val = val[1].data(); d.cs.c_val = val[2].data();
            d.alpha_g1 = pk.alpha_g1.l.data(); d.beta_g1 = pk.beta_g1.l.data(); d.beta_g2 = pk.beta_g2.l.data();
            d.delta_g1 = pk.delta_g1.l.data(); d.delta_g2 = pk.delta_g2.l.data();
            d.A_query = pk.A_query[0].l.data();
            d.B_count = b_idx.size();
            d.B_idx = b_idx.data();
            d.B_g2 = b_g2.empty() ? nullptr : b_g2[0].l.data();
            d.B_g1 = b_g1.empty() ? nullptr : b_g1[0].l.data();
            d.H_count = pk.H_query.size();
            d.H_query = pk.H_query.empty() ? nullptr : pk.H_query[0].l.data();
            d.L_query = pk.L_query.empty() ? nullptr : pk.L_query[0].l.data();
            std::size_t n = 0;
            check(vsp_codec_fast_pk_write(&d, nullptr, 0, &n), "serialize_pk_crs");
            blob_type out(n);
            check(vsp_codec_fast_pk_write(&d, out.data(), out.size(), &n), "serialize_pk_crs");
            return out;
        }
        // deserialize_vk_crs (common.hpp:734-738): the extended verification key; alpha_g1_beta_g2 stays in GT form
        static verification_key_type deserialize_vk_crs(const blob_type &blob) {
            std::uint64_t n = 0;
            check(vsp_codec_groth16_vk_read(blob.data(), blob.size(), &n, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0),
                  "deserialize_vk_crs");
            std::vector<g1_value> gabc(n);
            verification_key_type vk;
            check(vsp_codec_groth16_vk_read(blob.data(), blob.size(), nullptr, vk.alpha_g1_beta_g2.data(), vk.gamma_g2.l.data(),
                                            vk.delta_g2.l.data(), vk.delta_g1.l.data(), vk.gamma_g1.l.data(), gabc[0].l.data(), n),
                  "deserialize_vk_crs");
            vk.has_alpha_g1_beta_g2 = true;
            vk.gamma_ABC_g1.first = gabc[0];
            vk.gamma_ABC_g1.rest.assign(gabc.begin() + 1, gabc.end());
            return vk;
        }
        // needs the GT form (a key deserialised here, or one whose alpha_g1_beta_g2 the caller computed with vsp_pairing_products)
        static blob_type serialize_vk_crs(const verification_key_type &vk) {
            if (!vk.has_alpha_g1_beta_g2) throw std::invalid_argument("serialize_vk_crs: alpha_g1_beta_g2 is not set");
            std::vector<g1_value> gabc;
            gabc.push_back(vk.gamma_ABC_g1.first);
            gabc.insert(gabc.end(), vk.gamma_ABC_g1.rest.begin(), vk.gamma_ABC_g1.rest.end());
            std::size_t n = 0;
            check(vsp_codec_groth16_vk_write(vk.alpha_g1_beta_g2.data(), vk.gamma_g2.l.data(), vk.delta_g2.l.data(), vk.delta_g1.l.data(),
                                             vk.gamma_g1.l.data(), gabc.size(), gabc[0].l.data(), nullptr, 0, &n),
                  "serialize_vk_crs");
            blob_type out(n);
            check(vsp_codec_groth16_vk_write(vk.alpha_g1_beta_g2.data(), vk.gamma_g2.l.data(), vk.delta_g2.l.data(), vk.delta_g1.l.data(),
                                             vk.gamma_g1.l.data(), gabc.size(), gabc[0].l.data(), out.data(), out.size(), &n),
                  "serialize_vk_crs");
            return out;
        }
        // deserialize_proof / proof half of serialize_data (common.hpp:462-470, 761-764)
        static proof_type deserialize_proof(const blob_type &blob) {
            if (blob.size() != 192) throw std::invalid_argument("deserialize_proof: a proof is 192 bytes");
            std::uint64_t l[48];
            check(vsp_codec_proof_read(blob.data(), l), "deserialize_proof");
            proof_type p;
            std::memcpy(p.g_A.l.data(), l, 96);
            std::memcpy(p.g_B.l.data(), l + 12, 192);
            std::memcpy(p.g_C.l.data(), l + 36, 96);
            return p;
        }
        static blob_type serialize_proof(const proof_type &p) {
            std::uint64_t l[48];
            std::memcpy(l, p.g_A.l.data(), 96);
            std::memcpy(l + 12, p.g_B.l.data(), 192);
            std::memcpy(l + 36, p.g_C.l.data(), 96);
            blob_type out(192);
            check(vsp_codec_proof_write(l, out.data()), "serialize_proof");
            return out;
        }

        // ---- SAVER keys and cipher texts ---------------------------------------------------------------------------------
        // deserialize_pk_eid (common.hpp:691-695)
        static elgamal_public_key_type deserialize_pk_eid(const blob_type &blob) {
            std::uint64_t n = 0;
            check(vsp_codec_saver_pubkey_read(blob.data(), blob.size(), &n, nullptr), "deserialize_pk_eid");
            elgamal_public_key_type pk;
            pk.delta_s_g1.resize(n);
            pk.t_g1.resize(n);
            pk.t_g2.resize(n + 1);
            g1_value spare;
            vsp_saver_pubkey_t d = pk.descriptor();
            if (n == 0) d.delta_s_g1 = d.t_g1 = spare.l.data();
            check(vsp_codec_saver_pubkey_read(blob.data(), blob.size(), nullptr, &d), "deserialize_pk_eid");
            return pk;
        }
        static blob_type serialize_pk_eid(const elgamal_public_key_type &pk) {
            if (pk.t_g1.size() != pk.delta_s_g1.size() || pk.t_g2.size() != pk.delta_s_g1.size() + 1)
                throw std::invalid_argument("serialize_pk_eid: list sizes");
            const vsp_saver_pubkey_t d = pk.descriptor();
            std::size_t n = 0;
            check(vsp_codec_saver_pubkey_write(&d, nullptr, 0, &n), "serialize_pk_eid");
            blob_type out(n);
            check(vsp_codec_saver_pubkey_write(&d, out.data(), out.size(), &n), "serialize_pk_eid");
            return out;
        }
        // deserialize_vk_eid (common.hpp:706-710)
        static elgamal_verification_key_type deserialize_vk_eid(const blob_type &blob) {
            std::uint64_t n = 0;
            check(vsp_codec_saver_vk_read(blob.data(), blob.size(), &n, nullptr, nullptr, nullptr, 0), "deserialize_vk_eid");
            elgamal_verification_key_type vk;
            vk.rho_sv_g2.resize(n);
            vk.rho_rhov_g2.resize(n);
            g2_value spare;
            check(vsp_codec_saver_vk_read(blob.data(), blob.size(), nullptr, vk.rho_g2.l.data(), n ? vk.rho_sv_g2[0].l.data() : spare.l.data(),
                                          n ? vk.rho_rhov_g2[0].l.data() : spare.l.data(), n),
                  "deserialize_vk_eid");
            return vk;
        }
        static blob_type serialize_vk_eid(const elgamal_verification_key_type &vk) {
            if (vk.rho_sv_g2.size() != vk.rho_rhov_g2.size()) throw std::invalid_argument("serialize_vk_eid: list sizes");
            const std::uint64_t n = vk.rho_sv_g2.size();
            g2_value spare;
            const std::uint64_t *a = n ? vk.rho_sv_g2[0].l.data() : spare.l.data(), *b = n ? vk.rho_rhov_g2[0].l.data() : spare.l.data();
            std::size_t len = 0;
            check(vsp_codec_saver_vk_write(n, vk.rho_g2.l.data(), a, b, nullptr, 0, &len), "serialize_vk_eid");
            blob_type out(len);
            check(vsp_codec_saver_vk_write(n, vk.rho_g2.l.data(), a, b, out.data(), out.size(), &len), "serialize_vk_eid");
            return out;
        }
        // deserialize_sk_eid (common.hpp:721-725)
        static elgamal_private_key_type deserialize_sk_eid(const blob_type &blob) {
            elgamal_private_key_type sk;
            check(vsp_codec_saver_sk_read(blob.data(), blob.size(), sk.rho.l.data()), "deserialize_sk_eid");
            return sk;
        }
        static blob_type serialize_sk_eid(const elgamal_private_key_type &sk) {
            blob_type out(32);
            check(vsp_codec_saver_sk_write(sk.rho.l.data(), out.data()), "serialize_sk_eid");
            return out;
        }
        // deserialize_ct (common.hpp:775-783)
        static std::vector<g1_value> deserialize_ct(const blob_type &blob) {
            std::uint64_t n = 0;
            check(vsp_codec_ciphertext_read(blob.data(), blob.size(), &n, nullptr, 0), "deserialize_ct");
            std::vector<g1_value> ct(n);
            check(vsp_codec_ciphertext_read(blob.data(), blob.size(), &n, n ? ct[0].l.data() : nullptr, n), "deserialize_ct");
            return ct;
        }
        static blob_type serialize_ct(const std::vector<g1_value> &ct) {
            blob_type out(8 + 48 * ct.size());
            check(vsp_codec_ciphertext_write(ct.empty() ? nullptr : ct[0].l.data(), ct.size(), out.data()), "serialize_ct");
            return out;
        }
        // deserialize_decryption_proof (common.hpp:793-798): one compressed G1 point
        static g1_value deserialize_decryption_proof(const blob_type &blob) {
            if (blob.size() != 48) throw std::invalid_argument("deserialize_decryption_proof: 48 bytes");
            g1_value p;
            check(vsp_codec_g1_decompress(blob.data(), p.l.data()), "deserialize_decryption_proof");
            return p;
        }

        // ---- the phases' output bundles ------------------------------------------------------------------------------
        // serialize_data (common.hpp:462-485)
        static void serialize_data(std::size_t /*proof_idx*/, const proof_type &proof, const primary_input_type &pinput,
                                   const std::vector<g1_value> &ct, const primary_input_type &sn, blob_type &proof_blob,
                                   blob_type &pinput_blob, blob_type &ct_blob, blob_type &sn_blob) {
            proof_blob = serialize_proof(proof);
            pinput_blob = serialize_scalar_vector(pinput);
            ct_blob = serialize_ct(ct);
            sn_blob = serialize_scalar_vector(sn);
        }
        // serialize_tally_phase_data (common.hpp:511-522): dec = (plain text, decryption proof)
        static void serialize_tally_phase_data(const std::pair<std::vector<scalar_field_value_type>, g1_value> &dec,
                                               blob_type &dec_proof_blob, blob_type &voting_res_blob) {
            dec_proof_blob.assign(48, 0);
            check(vsp_codec_g1_compress(dec.second.l.data(), dec_proof_blob.data()), "serialize_tally_phase_data");
            voting_res_blob = serialize_scalar_vector(dec.first);
        }
        // serialize_initial_phase_admin_keys (common.hpp:330-366)
        static void serialize_initial_phase_admin_keys(const proving_key_type &pk_crs, const verification_key_type &vk_crs,
                                                       const elgamal_public_key_type &pk_eid, const elgamal_private_key_type &sk_eid,
                                                       const elgamal_verification_key_type &vk_eid, blob_type &r1cs_proving_key_out,
                                                       blob_type &r1cs_verification_key_out, blob_type &public_key_output,
                                                       blob_type &secret_key_output, blob_type &verification_key_output) {
            r1cs_proving_key_out = serialize_pk_crs(pk_crs);
            r1cs_verification_key_out = serialize_vk_crs(vk_crs);
            public_key_output = serialize_pk_eid(pk_eid);
            secret_key_output = serialize_sk_eid(sk_eid);
            verification_key_output = serialize_vk_eid(vk_eid);
        }
    };

}    // namespace b200
}    // namespace crypto3
}    // namespace nil

#endif    // VSP_CRYPTO3_MARSHALLING_HPP
