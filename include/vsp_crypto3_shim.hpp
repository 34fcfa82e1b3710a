// vsp_crypto3_shim.hpp — the crypto3 entry points of VoteSAVER's ballot-proof hot path, restated with their reference
// names, argument order and error behaviour, forwarding to the C ABI of libvsp_b200.so (include/vsp_b200.h).
//
// Header-only C++17, no Boost, no crypto3: this is the host side a maintainer puts under bin/cli in place of the
// templated CPU code (INTEGRATION.md shows the three conversions needed in the real tree).  Reference interfaces
// (paths relative to NilFoundation/vote-saver-protocol @ ae4f646c):
//   math::make_evaluation_domain<FieldType>(m)          libs/math/include/nil/crypto3/math/algorithms/make_evaluation_domain.hpp:51-106
//   evaluation_domain<FieldType>::fft / inverse_fft /   libs/math/include/nil/crypto3/math/domains/evaluation_domain.hpp:41-118
//       divide_by_z_on_coset / size
//   math::multiply_by_coset(a, g)                       libs/math/include/nil/crypto3/math/coset.hpp:37-43
//   algebra::multiexp<Method>(vec_start, vec_end,       call site libs/containers/include/nil/crypto3/container/sparse_vector.hpp:258-261
//       scalar_start, scalar_end, chunks)
//   zk::prove<ProofSystem>(pk, pubkey, primary, aux, r) call site libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:353-354
//   pubkey::encrypt<Scheme, Mode>(range, init_params)   libs/pubkey/include/nil/crypto3/pubkey/elgamal_verifiable.hpp:283-357,
//                                                       call site bin/cli/include/nil/vote_saver/common.hpp:1132-1135
//   pubkey::rerandomize<Scheme>(rnd, ct, init_params)   elgamal_verifiable.hpp:600-667, call site common.hpp:1143-1145
//
// Value types are plain limb arrays in the C ABI's format (canonical little-endian 64-bit limbs; infinity = all zero).
// What throws std::invalid_argument or trips BOOST_ASSERT in the reference throws std::invalid_argument here; device
// and library failures throw std::runtime_error.  There is no CPU path: without a CUDA device every call throws.
#ifndef VSP_CRYPTO3_SHIM_HPP
#define VSP_CRYPTO3_SHIM_HPP

#include <array>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#if defined(__linux__)
#include <cerrno>
#include <sys/random.h>
#include <sys/types.h>
#endif

#include <tuple>

#include "vsp_b200.h"

namespace nil {
namespace crypto3 {

namespace b200 {
    // ---- value types ---------------------------------------------------------------------------------------------
    struct fr_value {  // bls12_fr<381>::value_type, canonical
        std::array<std::uint64_t, 4> l {};
        fr_value() = default;
        explicit fr_value(std::uint64_t x) { l[0] = x; }
        bool operator==(const fr_value &o) const { return l == o.l; }
        bool operator!=(const fr_value &o) const { return l != o.l; }
        bool is_zero() const { return (l[0] | l[1] | l[2] | l[3]) == 0; }
    };
    struct g1_value {  // bls12<381>::g1_type<>::value_type as a canonical affine point
        std::array<std::uint64_t, 12> l {};
        bool operator==(const g1_value &o) const { return l == o.l; }
        bool is_zero() const {
            for (auto v : l) if (v) return false;
            return true;
        }
    };
    struct g2_value {
        std::array<std::uint64_t, 24> l {};
        bool operator==(const g2_value &o) const { return l == o.l; }
    };
    static_assert(sizeof(fr_value) == 32 && sizeof(g1_value) == 96 && sizeof(g2_value) == 192, "packed limb arrays");

    // Fr modulus
    inline const std::array<std::uint64_t, 4> &fr_modulus() {
        static const std::array<std::uint64_t, 4> m = {0xffffffff00000001ull, 0x53bda402fffe5bfeull, 0x3339d80809a1d805ull,
                                                       0x73eda753299d7d48ull};
        return m;
    }
    inline bool fr_is_reduced(const fr_value &x) {
        const auto &m = fr_modulus();
        for (int i = 3; i >= 0; --i) {
            if (x.l[i] < m[i]) return true;
            if (x.l[i] > m[i]) return false;
        }
        return false;
    }
    // uniform element of Fr by rejection sampling (stands in for random::algebraic_random_device, common.hpp:1131)
    template<typename Engine>
    fr_value random_fr(Engine &eng) {
        static_assert(Engine::min() == 0 && Engine::max() == ~std::uint64_t(0), "64 uniform bits per draw expected");
        for (;;) {
            fr_value x;
            for (auto &v : x.l) v = eng();
            x.l[3] &= 0x7fffffffffffffffull;
            if (fr_is_reduced(x)) return x;
        }
    }

    // ---- the GPU context behind every entry point -----------------------------------------------------------------
    class device {
    public:
        // one context per process and GPU; VSP_DEVICE selects the GPU (default 0)
        static device &instance() {
            static device d;
            return d;
        }
        vsp_ctx *ctx() const { return ctx_; }
        void check(int rc, const char *what) const {
            if (rc == VSP_OK) return;
            const std::string text = std::string(what) + ": " + (ctx_ ? vsp_last_error(ctx_) : "no context");
            if (rc == VSP_E_INVALID || rc == VSP_E_DOMAIN) throw std::invalid_argument(text);
            throw std::runtime_error(text);
        }
        device(const device &) = delete;
        device &operator=(const device &) = delete;

    private:
        device() {
            const char *e = std::getenv("VSP_DEVICE");
            const int dev = e ? std::atoi(e) : 0;
            const int rc = vsp_ctx_create(dev, nullptr, &ctx_);
            if (rc != VSP_OK)
                throw std::runtime_error("vsp_ctx_create failed (" + std::to_string(rc) +
                                         "): a CUDA device is required, there is no CPU path");
        }
        ~device() {
            if (ctx_) vsp_ctx_destroy(ctx_);
        }
        vsp_ctx *ctx_ = nullptr;
    };
}    // namespace b200

namespace algebra {
    namespace fields {
        template<std::size_t Version>
        struct bls12_fr;
        template<>
        struct bls12_fr<381> {
            typedef b200::fr_value value_type;
            constexpr static const std::size_t value_bits = 255;
        };
    }    // namespace fields
    namespace curves {
        template<std::size_t Version>
        struct bls12;
        template<>
        struct bls12<381> {
            typedef fields::bls12_fr<381> scalar_field_type;
            template<typename = void>
            struct g1_type {
                typedef b200::g1_value value_type;
            };
            template<typename = void>
            struct g2_type {
                typedef b200::g2_value value_type;
            };
        };
    }    // namespace curves
}    // namespace algebra

// =====================================================================================================================
// math: evaluation domains and the coset shift
// =====================================================================================================================
namespace math {
    template<typename FieldType>
    class evaluation_domain {
    public:
        typedef FieldType field_type;
        typedef typename FieldType::value_type value_type;

        std::size_t m;
        int kind;    // VSP_DOMAIN_BASIC_RADIX2 / VSP_DOMAIN_STEP_RADIX2

        evaluation_domain(std::size_t m, int kind) : m(m), kind(kind) {}
        virtual ~evaluation_domain() = default;

        std::size_t size() const { return m; }

        // basic_radix2_domain zero-pads short input; longer input, and step_radix2::inverse_fft on a.size() != m,
        // throw std::invalid_argument (basic_radix2_domain.hpp:70-88, step_radix2_domain.hpp:112-114)
        virtual void fft(std::vector<value_type> &a) { run(a, 0); }
        virtual void inverse_fft(std::vector<value_type> &a) { run(a, 1); }

        virtual void divide_by_z_on_coset(std::vector<value_type> &P) {
            if (P.size() != m) throw std::invalid_argument("expected P.size() == this->m");
            auto &d = b200::device::instance();
            d.check(vsp_fr_divide_by_z_on_coset(d.ctx(), m, P[0].l.data()), "divide_by_z_on_coset");
        }

    private:
        void run(std::vector<value_type> &a, int inverse) {
            auto &d = b200::device::instance();
            std::vector<value_type> out(m);
            d.check(vsp_fr_fft(d.ctx(), m, inverse, a.empty() ? out[0].l.data() : a[0].l.data(), a.size(), out[0].l.data()),
                    inverse ? "inverse_fft" : "fft");
            a.swap(out);
        }
    };

    // make_evaluation_domain.hpp:51-106: basic_radix2 / step_radix2 by the reference's own selection rules; a size
    // with no radix-2 family domain throws std::invalid_argument as there
    template<typename FieldType>
    std::shared_ptr<evaluation_domain<FieldType>> make_evaluation_domain(std::size_t m) {
        int kind = 0;
        std::uint64_t size = 0;
        if (vsp_domain_plan(m, &kind, &size) != VSP_OK || kind == VSP_DOMAIN_EXTENDED_RADIX2)
            throw std::invalid_argument("make_evaluation_domain: no radix-2 family domain of that size");
        return std::make_shared<evaluation_domain<FieldType>>(static_cast<std::size_t>(size), kind);
    }

    // coset.hpp:37-43: a[i] *= g^i
    template<typename Range, typename FieldValueType>
    void multiply_by_coset(Range &a, const FieldValueType &g) {
        if (a.size() == 0) return;
        auto &d = b200::device::instance();
        d.check(vsp_fr_multiply_by_coset(d.ctx(), a.size(), g.l.data(), a[0].l.data()), "multiply_by_coset");
    }
}    // namespace math

// =====================================================================================================================
// algebra::multiexp
// =====================================================================================================================
namespace algebra {
    namespace policies {
        struct multiexp_method_naive_plain {};
        struct multiexp_method_bos_coster {};
        struct multiexp_method_BDLO12 {};
    }    // namespace policies

    namespace detail {
        inline void multiexp_run(const b200::g1_value *b, std::size_t n, const b200::fr_value *s, b200::g1_value &out) {
            auto &d = b200::device::instance();
            d.check(vsp_g1_msm(d.ctx(), n, n ? b->l.data() : nullptr, n ? s->l.data() : nullptr, out.l.data()), "multiexp");
        }
        inline void multiexp_run(const b200::g2_value *b, std::size_t n, const b200::fr_value *s, b200::g2_value &out) {
            auto &d = b200::device::instance();
            d.check(vsp_g2_msm(d.ctx(), n, n ? b->l.data() : nullptr, n ? s->l.data() : nullptr, out.l.data()), "multiexp");
        }
    }    // namespace detail

    // Every Method computes the same group element; all of them run the GPU bucket multiexp.  `chunks` (the reference's
    // OpenMP split) is accepted and ignored.  Iterators must be contiguous (std::vector / pointers), as they are at the
    // reference's call sites.
    template<typename MultiexpMethod, typename InputBaseIterator, typename InputFieldIterator>
    typename std::iterator_traits<InputBaseIterator>::value_type
        multiexp(InputBaseIterator vec_start, InputBaseIterator vec_end, InputFieldIterator scalar_start,
                 InputFieldIterator scalar_end, std::size_t chunks) {
        (void)chunks;
        const std::size_t n = static_cast<std::size_t>(vec_end - vec_start);
        if (static_cast<std::size_t>(scalar_end - scalar_start) != n)
            throw std::invalid_argument("multiexp: bases and scalars differ in length");
        typename std::iterator_traits<InputBaseIterator>::value_type out;
        detail::multiexp_run(n ? &*vec_start : nullptr, n, n ? &*scalar_start : nullptr, out);
        return out;
    }
    // zeros and ones need no pre-pass on the GPU: same entry
    template<typename MultiexpMethod, typename InputBaseIterator, typename InputFieldIterator>
    typename std::iterator_traits<InputBaseIterator>::value_type
        multiexp_with_mixed_addition(InputBaseIterator vec_start, InputBaseIterator vec_end, InputFieldIterator scalar_start,
                                     InputFieldIterator scalar_end, std::size_t chunks) {
        return multiexp<MultiexpMethod>(vec_start, vec_end, scalar_start, scalar_end, chunks);
    }
}    // namespace algebra

// =====================================================================================================================
// zk: r1cs_gg_ppzksnark data types (field order of fast_proving_key.hpp:66-103, r1cs.hpp:57-113, proof.hpp:63-71)
// =====================================================================================================================
namespace zk {
    namespace snark {
        enum class proving_mode { basic, encrypted_input };

        template<typename FieldType>
        struct linear_term {
            std::size_t index;    // variable 0 is the constant one
            typename FieldType::value_type coeff;
        };
        template<typename FieldType>
        struct r1cs_constraint {
            std::vector<linear_term<FieldType>> a, b, c;
        };
        template<typename FieldType>
        struct r1cs_constraint_system {
            std::size_t primary_input_size = 0, auxiliary_input_size = 0;
            std::vector<r1cs_constraint<FieldType>> constraints;
            std::size_t num_inputs() const { return primary_input_size; }
            std::size_t num_variables() const { return primary_input_size + auxiliary_input_size; }
            std::size_t num_constraints() const { return constraints.size(); }
        };

        template<typename CurveType>
        struct knowledge_commitment {
            typename CurveType::template g2_type<>::value_type g;
            typename CurveType::template g1_type<>::value_type h;
        };
        template<typename T>
        struct sparse_vector {    // containers/sparse_vector.hpp: sorted indices + values
            std::vector<std::size_t> indices;
            std::vector<T> values;
            std::size_t domain_size_ = 0;
        };

        template<typename CurveType>
        struct r1cs_gg_ppzksnark_proof {
            typename CurveType::template g1_type<>::value_type g_A;
            typename CurveType::template g2_type<>::value_type g_B;
            typename CurveType::template g1_type<>::value_type g_C;
        };

        template<typename CurveType>
        class r1cs_gg_ppzksnark_proving_key {
        public:
            typedef typename CurveType::scalar_field_type field_type;
            typedef typename CurveType::template g1_type<>::value_type g1_value_type;
            typedef typename CurveType::template g2_type<>::value_type g2_value_type;

            g1_value_type alpha_g1, beta_g1;
            g2_value_type beta_g2;
            g1_value_type delta_g1;
            g2_value_type delta_g2;
            std::vector<g1_value_type> A_query;
            sparse_vector<knowledge_commitment<CurveType>> B_query;
            std::vector<g1_value_type> H_query, L_query;
            r1cs_constraint_system<field_type> constraint_system;

            r1cs_gg_ppzksnark_proving_key() = default;
            r1cs_gg_ppzksnark_proving_key(const r1cs_gg_ppzksnark_proving_key &) = delete;
            r1cs_gg_ppzksnark_proving_key &operator=(const r1cs_gg_ppzksnark_proving_key &) = delete;
            r1cs_gg_ppzksnark_proving_key(r1cs_gg_ppzksnark_proving_key &&o) noexcept :
                alpha_g1(o.alpha_g1), beta_g1(o.beta_g1), beta_g2(o.beta_g2), delta_g1(o.delta_g1), delta_g2(o.delta_g2),
                A_query(std::move(o.A_query)), B_query(std::move(o.B_query)), H_query(std::move(o.H_query)),
                L_query(std::move(o.L_query)), constraint_system(std::move(o.constraint_system)), resident_(o.resident_) {
                o.resident_ = nullptr;
            }
            ~r1cs_gg_ppzksnark_proving_key() {
                if (resident_) vsp_pk_destroy(resident_);
            }

            // the key's GPU residency (query window tables, CSR, coset tables): built on first use, then reused by
            // every proof.  Call again by hand (reload = true) after editing the fields.
            vsp_pk *resident(bool reload = false) const {
                if (resident_ && !reload) return resident_;
                if (resident_) vsp_pk_destroy(resident_);
                resident_ = nullptr;
                const auto &cs = constraint_system;
                const std::size_t nc = cs.num_constraints();
                std::vector<std::uint64_t> row[3], val[3];
                std::vector<std::uint32_t> col[3];
                for (int k = 0; k < 3; ++k) row[k].assign(nc + 1, 0);
                for (std::size_t i = 0; i < nc; ++i) {
                    const std::vector<linear_term<field_type>> *lc[3] = {&cs.constraints[i].a, &cs.constraints[i].b,
                                                                         &cs.constraints[i].c};
                    for (int k = 0; k < 3; ++k) {
                        for (const auto &t : *lc[k]) {
                            col[k].push_back(static_cast<std::uint32_t>(t.index));
                            val[k].insert(val[k].end(), t.coeff.l.begin(), t.coeff.l.end());
                        }
                        row[k][i + 1] = col[k].size();
                    }
                }
                std::vector<std::uint32_t> b_idx(B_query.indices.begin(), B_query.indices.end());
                std::vector<g2_value_type> b_g2(B_query.values.size());
                std::vector<g1_value_type> b_g1(B_query.values.size());
                for (std::size_t i = 0; i < B_query.values.size(); ++i) {
                    b_g2[i] = B_query.values[i].g;
                    b_g1[i] = B_query.values[i].h;
                }
                vsp_proving_key_t d;
                std::memset(&d, 0, sizeof(d));
                d.cs.num_inputs = cs.primary_input_size;
                d.cs.num_aux = cs.auxiliary_input_size;
                d.cs.num_constraints = nc;
                d.cs.a_row = row[0].data(); d.cs.b_row = row[1].data(); d.cs.c_row = row[2].data();
                d.cs.a_col = col[0].data(); d.cs.b_col = col[1].data(); d.cs.c_col = col[2].data();
                d.cs.a_val = val[0].data(); d.cs.b_val = val[1].data(); d.cs.c_val = val[2].data();
                d.alpha_g1 = alpha_g1.l.data(); d.beta_g1 = beta_g1.l.data(); d.beta_g2 = beta_g2.l.data();
                d.delta_g1 = delta_g1.l.data(); d.delta_g2 = delta_g2.l.data();
                d.A_query = A_query.empty() ? nullptr : A_query[0].l.data();
                d.B_count = b_idx.size();
                d.B_idx = b_idx.data();
                d.B_g2 = b_g2.empty() ? nullptr : b_g2[0].l.data();
                d.B_g1 = b_g1.empty() ? nullptr : b_g1[0].l.data();
                d.H_count = H_query.size();
                d.H_query = H_query.empty() ? nullptr : H_query[0].l.data();
                d.L_query = L_query.empty() ? nullptr : L_query[0].l.data();
                if (A_query.size() != cs.num_variables() + 1 || L_query.size() != cs.auxiliary_input_size)
                    throw std::invalid_argument("proving key: A_query / L_query sizes do not match the constraint system");
                auto &dev = b200::device::instance();
                dev.check(vsp_pk_load(dev.ctx(), &d, &resident_), "vsp_pk_load");
                return resident_;
            }

        private:
            mutable vsp_pk *resident_ = nullptr;
        };

        template<typename T>
        struct accumulation_vector {    // containers/accumulation_vector.hpp: first + rest
            T first;
            std::vector<T> rest;
        };
        template<typename CurveType>
        struct r1cs_gg_ppzksnark_verification_key {    // the extended (encrypted_input) key's group members; the GT element
                                                       // alpha_g1_beta_g2 = e(alpha_g1, beta_g2) belongs to the pairing-side
                                                       // verifier and is kept as its two preimages here
            typename CurveType::template g1_type<>::value_type alpha_g1;
            typename CurveType::template g2_type<>::value_type beta_g2;
            typename CurveType::template g2_type<>::value_type gamma_g2, delta_g2;
            typename CurveType::template g1_type<>::value_type delta_g1, gamma_g1;
            accumulation_vector<typename CurveType::template g1_type<>::value_type> gamma_ABC_g1;
            // A key that was deserialised (verification_key.hpp:128-152, vsp_codec_groth16_vk_read) has no alpha_g1 / beta_g2,
            // only their pairing: 12 Fq coefficients in tower order, 72 limbs.  When has_alpha_g1_beta_g2 is set the verifiers
            // below use it and ignore alpha_g1 / beta_g2.
            std::array<std::uint64_t, 72> alpha_g1_beta_g2 {};
            bool has_alpha_g1_beta_g2 = false;
        };

        template<typename CurveType, proving_mode Mode = proving_mode::basic>
        struct r1cs_gg_ppzksnark {
            typedef CurveType curve_type;
            constexpr static const proving_mode mode = Mode;
            typedef r1cs_gg_ppzksnark_proving_key<CurveType> proving_key_type;
            typedef r1cs_gg_ppzksnark_verification_key<CurveType> verification_key_type;
            typedef std::pair<proving_key_type, verification_key_type> keypair_type;
            typedef r1cs_gg_ppzksnark_proof<CurveType> proof_type;
            typedef std::vector<typename CurveType::scalar_field_type::value_type> primary_input_type;
            typedef std::vector<typename CurveType::scalar_field_type::value_type> auxiliary_input_type;
        };
    }    // namespace snark
}    // namespace zk

// =====================================================================================================================
// pubkey: elgamal_verifiable (SAVER) types — elgamal_verifiable.hpp:60-180
// =====================================================================================================================
namespace pubkey {
    template<typename CurveType, std::size_t BlockBits = 4>
    struct elgamal_verifiable {
        typedef CurveType curve_type;
        typedef zk::snark::r1cs_gg_ppzksnark<CurveType, zk::snark::proving_mode::encrypted_input> proof_system_type;
        typedef typename CurveType::scalar_field_type scalar_field_type;
        typedef typename CurveType::template g1_type<>::value_type g1_value_type;
        typedef typename CurveType::template g2_type<>::value_type g2_value_type;
        // (cipher text, proof)  — elgamal_verifiable.hpp:77-78
        typedef std::pair<std::vector<g1_value_type>, typename proof_system_type::proof_type> cipher_type;
    };

    template<typename Scheme>
    struct public_key;
    template<typename CurveType, std::size_t BlockBits>
    struct public_key<elgamal_verifiable<CurveType, BlockBits>> {    // elgamal_verifiable.hpp:152-157
        typedef elgamal_verifiable<CurveType, BlockBits> scheme_type;
        typename scheme_type::g1_value_type delta_g1;
        std::vector<typename scheme_type::g1_value_type> delta_s_g1;
        std::vector<typename scheme_type::g1_value_type> t_g1;
        std::vector<typename scheme_type::g2_value_type> t_g2;
        typename scheme_type::g1_value_type delta_sum_s_g1;
        typename scheme_type::g1_value_type gamma_inverse_sum_s_g1;

        vsp_saver_pubkey_t descriptor() const {
            vsp_saver_pubkey_t d;
            d.msg_size = delta_s_g1.size();
            d.delta_g1 = delta_g1.l.data();
            d.delta_s_g1 = delta_s_g1.empty() ? nullptr : delta_s_g1[0].l.data();
            d.t_g1 = t_g1.empty() ? nullptr : t_g1[0].l.data();
            d.t_g2 = t_g2.empty() ? nullptr : t_g2[0].l.data();
            d.delta_sum_s_g1 = delta_sum_s_g1.l.data();
            d.gamma_inverse_sum_s_g1 = gamma_inverse_sum_s_g1.l.data();
            return d;
        }
    };

    template<typename Scheme>
    struct private_key;
    template<typename CurveType, std::size_t BlockBits>
    struct private_key<elgamal_verifiable<CurveType, BlockBits>> {    // elgamal_verifiable.hpp:131-139
        typename CurveType::scalar_field_type::value_type rho;
    };
    template<typename Scheme>
    struct verification_key;
    template<typename CurveType, std::size_t BlockBits>
    struct verification_key<elgamal_verifiable<CurveType, BlockBits>> {    // elgamal_verifiable.hpp:85-113
        typedef elgamal_verifiable<CurveType, BlockBits> scheme_type;
        typename scheme_type::g2_value_type rho_g2;
        std::vector<typename scheme_type::g2_value_type> rho_sv_g2;
        std::vector<typename scheme_type::g2_value_type> rho_rhov_g2;
    };

    namespace modes {
        template<typename Scheme>
        struct verifiable_encryption {};
    }    // namespace modes
}    // namespace pubkey

// =====================================================================================================================
// zk::prove — r1cs_gg_ppzksnark, basic and encrypted_input modes
// =====================================================================================================================
namespace zk {
    namespace detail {
        template<typename CurveType>
        snark::r1cs_gg_ppzksnark_proof<CurveType> prove_run(const snark::r1cs_gg_ppzksnark_proving_key<CurveType> &pk,
                                                            const std::vector<b200::fr_value> &primary_input,
                                                            const std::vector<b200::fr_value> &auxiliary_input,
                                                            const b200::fr_value &r, const b200::fr_value &s,
                                                            const b200::fr_value *r_saver, const b200::g1_value *gamma_inv) {
            const auto &cs = pk.constraint_system;
            if (primary_input.size() != cs.primary_input_size || auxiliary_input.size() != cs.auxiliary_input_size)
                throw std::invalid_argument("prove: assignment does not match the constraint system");
            auto &d = b200::device::instance();
            std::uint64_t out[48];
            d.check(vsp_prove(d.ctx(), pk.resident(), primary_input.empty() ? nullptr : primary_input[0].l.data(),
                              auxiliary_input.empty() ? nullptr : auxiliary_input[0].l.data(), r.l.data(), s.l.data(),
                              r_saver ? r_saver->l.data() : nullptr, gamma_inv ? gamma_inv->l.data() : nullptr, out),
                    "prove");
            snark::r1cs_gg_ppzksnark_proof<CurveType> proof;
            std::memcpy(proof.g_A.l.data(), out, 96);
            std::memcpy(proof.g_B.l.data(), out + 12, 192);
            std::memcpy(proof.g_C.l.data(), out + 36, 96);
            return proof;
        }
        // Every secret scalar (the generator's toxic waste, Groth16 r and s, SAVER's encryption and rerandomisation
        // randomness) comes from the operating system's CSPRNG, 64 bits per call, as the reference draws them from
        // random::algebraic_random_device (common.hpp:923, 1131): getrandom(2) on Linux, std::random_device (one call per
        // 32-bit half) elsewhere.  No seeded PRNG sits between the OS and a secret; a short read or an error throws.
        struct os_entropy_engine {
            typedef std::uint64_t result_type;
            static constexpr result_type min() { return 0; }
            static constexpr result_type max() { return ~result_type(0); }
            result_type operator()() {
                result_type v = 0;
#if defined(__linux__)
                unsigned char *p = reinterpret_cast<unsigned char *>(&v);
                std::size_t got = 0;
                while (got < sizeof(v)) {
                    const ssize_t n = ::getrandom(p + got, sizeof(v) - got, 0);
                    if (n < 0) {
                        if (errno == EINTR) continue;
                        throw std::runtime_error("getrandom failed: no OS entropy for a secret scalar");
                    }
                    got += static_cast<std::size_t>(n);
                }
#else
                std::random_device rd;
                v = (result_type(rd()) << 32) | result_type(rd() & 0xffffffffu);
#endif
                return v;
            }
        };
        inline os_entropy_engine &blinding_engine() {
            static os_entropy_engine eng;
            return eng;
        }
    }    // namespace detail

    // generate<ProofSystem>(constraint_system) -> keypair   (bin/cli/include/nil/vote_saver/common.hpp:916-917).
    // The reference draws t, alpha, beta, gamma, delta inside; the overload taking them makes key pairs reproducible.
    struct generator_randomness {
        b200::fr_value t, alpha, beta, gamma, delta;
    };
    template<typename ProofSystem>
    typename ProofSystem::keypair_type
        generate(const snark::r1cs_constraint_system<typename ProofSystem::curve_type::scalar_field_type> &cs,
                 const generator_randomness &rnd) {
        typedef typename ProofSystem::curve_type curve_type;
        typedef typename curve_type::scalar_field_type field_type;
        typedef typename curve_type::template g1_type<>::value_type g1;
        typedef typename curve_type::template g2_type<>::value_type g2;
        const std::size_t nc = cs.num_constraints(), nv = cs.num_variables(), ni = cs.primary_input_size;
        std::vector<std::uint64_t> row[3], val[3];
        std::vector<std::uint32_t> col[3];
        for (int k = 0; k < 3; ++k) row[k].assign(nc + 1, 0);
        for (std::size_t i = 0; i < nc; ++i) {
            const std::vector<snark::linear_term<field_type>> *lc[3] = {&cs.constraints[i].a, &cs.constraints[i].b,
                                                                        &cs.constraints[i].c};
            for (int k = 0; k < 3; ++k) {
                for (const auto &t : *lc[k]) {
                    col[k].push_back(static_cast<std::uint32_t>(t.index));
                    val[k].insert(val[k].end(), t.coeff.l.begin(), t.coeff.l.end());
                }
                row[k][i + 1] = col[k].size();
            }
        }
        vsp_r1cs_t d;
        std::memset(&d, 0, sizeof(d));
        d.num_inputs = ni;
        d.num_aux = cs.auxiliary_input_size;
        d.num_constraints = nc;
        d.a_row = row[0].data(); d.b_row = row[1].data(); d.c_row = row[2].data();
        d.a_col = col[0].data(); d.b_col = col[1].data(); d.c_col = col[2].data();
        d.a_val = val[0].data(); d.b_val = val[1].data(); d.c_val = val[2].data();
        int kind = 0;
        std::uint64_t m = 0;
        if (vsp_domain_plan(nc + ni + 1, &kind, &m) != VSP_OK)
            throw std::invalid_argument("generate: no evaluation domain for this constraint system");
        typename ProofSystem::keypair_type kp;
        auto &pk = kp.first;
        auto &vk = kp.second;
        pk.A_query.resize(nv + 1);
        pk.H_query.resize(m - 1);
        pk.L_query.resize(cs.auxiliary_input_size);
        std::vector<g1> b_g1(nv + 1), gabc(ni + 1);
        std::vector<g2> b_g2(nv + 1);
        std::vector<std::uint32_t> b_idx(nv + 1);
        g1 scratch_g1;
        vsp_keypair_t out;
        std::memset(&out, 0, sizeof(out));
        out.alpha_g1 = pk.alpha_g1.l.data(); out.beta_g1 = pk.beta_g1.l.data(); out.beta_g2 = pk.beta_g2.l.data();
        out.delta_g1 = pk.delta_g1.l.data(); out.delta_g2 = pk.delta_g2.l.data();
        out.A_query = pk.A_query[0].l.data();
        out.B_idx = b_idx.data(); out.B_g2 = b_g2[0].l.data(); out.B_g1 = b_g1[0].l.data();
        out.H_query = pk.H_query.empty() ? scratch_g1.l.data() : pk.H_query[0].l.data();
        out.L_query = pk.L_query.empty() ? nullptr : pk.L_query[0].l.data();
        out.gamma_g2 = vk.gamma_g2.l.data(); out.gamma_g1 = vk.gamma_g1.l.data();
        out.gamma_ABC_g1 = gabc[0].l.data();
        const std::uint64_t toxic[20] = {rnd.t.l[0], rnd.t.l[1], rnd.t.l[2], rnd.t.l[3],
                                         rnd.alpha.l[0], rnd.alpha.l[1], rnd.alpha.l[2], rnd.alpha.l[3],
                                         rnd.beta.l[0], rnd.beta.l[1], rnd.beta.l[2], rnd.beta.l[3],
                                         rnd.gamma.l[0], rnd.gamma.l[1], rnd.gamma.l[2], rnd.gamma.l[3],
                                         rnd.delta.l[0], rnd.delta.l[1], rnd.delta.l[2], rnd.delta.l[3]};
        auto &dev = b200::device::instance();
        dev.check(vsp_groth16_generate(dev.ctx(), &d, toxic, &out), "generate");
        for (std::uint64_t i = 0; i < out.B_count; ++i) {
            pk.B_query.indices.push_back(b_idx[i]);
            pk.B_query.values.push_back({b_g2[i], b_g1[i]});
        }
        pk.constraint_system = cs;
        vk.alpha_g1 = pk.alpha_g1;
        vk.beta_g2 = pk.beta_g2;
        vk.delta_g2 = pk.delta_g2;
        vk.delta_g1 = pk.delta_g1;
        vk.gamma_ABC_g1.first = gabc[0];
        vk.gamma_ABC_g1.rest.assign(gabc.begin() + 1, gabc.end());
        return kp;
    }
    template<typename ProofSystem>
    typename ProofSystem::keypair_type
        generate(const snark::r1cs_constraint_system<typename ProofSystem::curve_type::scalar_field_type> &cs) {
        auto &e = detail::blinding_engine();
        generator_randomness rnd;
        for (b200::fr_value *x : {&rnd.t, &rnd.alpha, &rnd.beta, &rnd.gamma, &rnd.delta})
            do *x = b200::random_fr(e);
            while (x->is_zero());
        return generate<ProofSystem>(cs, rnd);
    }

    // basic mode: prove<ProofSystem>(pk, primary, auxiliary)  (libs/marshalling/zk/test/r1cs_gg_ppzksnark.cpp:84).
    // The reference draws Groth16's blinding scalars r, s inside; the overloads with explicit (r, s) make proofs
    // reproducible and are what the parity tests use.
    template<typename ProofSystem>
    typename ProofSystem::proof_type prove(const typename ProofSystem::proving_key_type &pk,
                                           const typename ProofSystem::primary_input_type &primary_input,
                                           const typename ProofSystem::auxiliary_input_type &auxiliary_input,
                                           const b200::fr_value &r, const b200::fr_value &s) {
        return detail::prove_run<typename ProofSystem::curve_type>(pk, primary_input, auxiliary_input, r, s, nullptr, nullptr);
    }
    template<typename ProofSystem>
    typename ProofSystem::proof_type prove(const typename ProofSystem::proving_key_type &pk,
                                           const typename ProofSystem::primary_input_type &primary_input,
                                           const typename ProofSystem::auxiliary_input_type &auxiliary_input) {
        auto &e = detail::blinding_engine();
        const b200::fr_value r = b200::random_fr(e), s = b200::random_fr(e);
        return prove<ProofSystem>(pk, primary_input, auxiliary_input, r, s);
    }
    // encrypted_input mode: prove<ProofSystem>(pk, pubkey, primary, auxiliary, r_saver)  (elgamal_verifiable.hpp:353-354):
    // g_C additionally carries r_saver * pubkey.gamma_inverse_sum_s_g1
    template<typename ProofSystem, typename PublicKey>
    typename ProofSystem::proof_type prove(const typename ProofSystem::proving_key_type &pk, const PublicKey &pubkey,
                                           const typename ProofSystem::primary_input_type &primary_input,
                                           const typename ProofSystem::auxiliary_input_type &auxiliary_input,
                                           const b200::fr_value &r_saver, const b200::fr_value &r, const b200::fr_value &s) {
        static_assert(ProofSystem::mode == snark::proving_mode::encrypted_input, "pubkey overload is encrypted_input mode");
        return detail::prove_run<typename ProofSystem::curve_type>(pk, primary_input, auxiliary_input, r, s, &r_saver,
                                                                   &pubkey.gamma_inverse_sum_s_g1);
    }
    template<typename ProofSystem, typename PublicKey>
    typename ProofSystem::proof_type prove(const typename ProofSystem::proving_key_type &pk, const PublicKey &pubkey,
                                           const typename ProofSystem::primary_input_type &primary_input,
                                           const typename ProofSystem::auxiliary_input_type &auxiliary_input,
                                           const b200::fr_value &r_saver) {
        auto &e = detail::blinding_engine();
        const b200::fr_value r = b200::random_fr(e), s = b200::random_fr(e);
        return prove<ProofSystem>(pk, pubkey, primary_input, auxiliary_input, r_saver, r, s);
    }

    // verify<ProofSystem>(vk, primary_input, proof) -> bool   (libs/marshalling/zk/test/r1cs_gg_ppzksnark.cpp:160-163)
    template<typename ProofSystem>
    bool verify(const typename ProofSystem::verification_key_type &vk,
                const typename ProofSystem::primary_input_type &primary_input, const typename ProofSystem::proof_type &proof) {
        if (vk.gamma_ABC_g1.rest.size() != primary_input.size())
            throw std::invalid_argument("verify: primary input size does not match the verification key");
        std::vector<b200::g1_value> gabc;
        gabc.push_back(vk.gamma_ABC_g1.first);
        gabc.insert(gabc.end(), vk.gamma_ABC_g1.rest.begin(), vk.gamma_ABC_g1.rest.end());
        vsp_verification_key_t d;
        d.alpha_g1 = vk.alpha_g1.l.data();
        d.beta_g2 = vk.beta_g2.l.data();
        d.gamma_g2 = vk.gamma_g2.l.data();
        d.delta_g2 = vk.delta_g2.l.data();
        d.gamma_ABC_count = gabc.size();
        d.gamma_ABC_g1 = gabc[0].l.data();
        std::uint64_t pr[48];
        std::memcpy(pr, proof.g_A.l.data(), 96);
        std::memcpy(pr + 12, proof.g_B.l.data(), 192);
        std::memcpy(pr + 36, proof.g_C.l.data(), 96);
        std::uint8_t ok = 0;
        auto &dev = b200::device::instance();
        const std::uint64_t *x = primary_input.empty() ? nullptr : primary_input[0].l.data();
        if (vk.has_alpha_g1_beta_g2)  // the key as deserialised: alpha_g1_beta_g2 in GT form
            dev.check(vsp_groth16_verify_gt_batch(dev.ctx(), vk.alpha_g1_beta_g2.data(), d.gamma_g2, d.delta_g2, d.gamma_ABC_count,
                                                  d.gamma_ABC_g1, 1, x, primary_input.size(), pr, &ok),
                      "verify");
        else
            dev.check(vsp_groth16_verify_batch(dev.ctx(), &d, 1, x, primary_input.size(), pr, &ok), "verify");
        return ok != 0;
    }
}    // namespace zk

// =====================================================================================================================
// pubkey::encrypt / pubkey::rerandomize for elgamal_verifiable
// =====================================================================================================================
namespace pubkey {
    template<typename Scheme>
    struct generate_keypair_init_params {    // generate_keypair_op::init_params_type, elgamal_verifiable.hpp:189-192
        const typename Scheme::proof_system_type::keypair_type &gg_keypair;
        std::size_t msg_size;
    };
    // generate_keypair<Scheme, modes::verifiable_encryption<Scheme>>(rnd, {gg_keypair, msg_size})
    //   -> (public key, private key, verification key)     generate_keypair_op::process, elgamal_verifiable.hpp:218-271
    // as bin/cli/include/nil/vote_saver/common.hpp:928-931 calls it
    template<typename Scheme, typename Mode, typename RandomRange>
    std::tuple<public_key<Scheme>, private_key<Scheme>, verification_key<Scheme>>
        generate_keypair(const RandomRange &rnd, const generate_keypair_init_params<Scheme> &p) {
        typedef typename Scheme::scalar_field_type::value_type fr;
        const std::vector<fr> rr(std::begin(rnd), std::end(rnd));
        const auto &vk = p.gg_keypair.second;
        const std::size_t n = p.msg_size;
        // the reference's BOOST_ASSERT_MSG checks, elgamal_verifiable.hpp:199-201, 220-221
        if (!(vk.gamma_ABC_g1.rest.size() > n))
            throw std::invalid_argument("Array of gammas in vk should be longer than the message.");
        if (rr.size() < 3 * n + 2) throw std::invalid_argument("Too few numbers in the source of randomness.");
        std::tuple<public_key<Scheme>, private_key<Scheme>, verification_key<Scheme>> out;
        auto &pub = std::get<0>(out);
        auto &svk = std::get<2>(out);
        pub.delta_g1 = vk.delta_g1;
        pub.delta_s_g1.resize(n);
        pub.t_g1.resize(n);
        pub.t_g2.resize(n + 1);
        svk.rho_sv_g2.resize(n);
        svk.rho_rhov_g2.resize(n);
        vsp_saver_keypair_t d;
        d.delta_s_g1 = n ? pub.delta_s_g1[0].l.data() : nullptr;
        d.t_g1 = n ? pub.t_g1[0].l.data() : nullptr;
        d.t_g2 = pub.t_g2[0].l.data();
        d.delta_sum_s_g1 = pub.delta_sum_s_g1.l.data();
        d.gamma_inverse_sum_s_g1 = pub.gamma_inverse_sum_s_g1.l.data();
        d.rho_g2 = svk.rho_g2.l.data();
        d.rho_sv_g2 = n ? svk.rho_sv_g2[0].l.data() : nullptr;
        d.rho_rhov_g2 = n ? svk.rho_rhov_g2[0].l.data() : nullptr;
        d.rho = std::get<1>(out).rho.l.data();
        auto &dev = b200::device::instance();
        dev.check(vsp_saver_keygen(dev.ctx(), n, vk.delta_g1.l.data(), vk.gamma_g1.l.data(), vk.gamma_ABC_g1.rest[0].l.data(),
                                   rr[0].l.data(), &d),
                  "generate_keypair");
        return out;
    }

    template<typename Scheme>
    struct encrypt_init_params {    // encrypt_op::init_params_type, elgamal_verifiable.hpp:283-290
        typename Scheme::scalar_field_type::value_type r;
        const public_key<Scheme> &pubkey;
        const typename Scheme::proof_system_type::keypair_type &gg_keypair;
        const typename Scheme::proof_system_type::primary_input_type &primary_input;
        const typename Scheme::proof_system_type::auxiliary_input_type &auxiliary_input;
        // Groth16 blinding scalars; left null they are drawn as the reference's prover does
        const typename Scheme::scalar_field_type::value_type *groth16_r = nullptr;
        const typename Scheme::scalar_field_type::value_type *groth16_s = nullptr;
    };

    // encrypt<Scheme, modes::verifiable_encryption<Scheme>>(plain_text, {r, pubkey, gg_keypair, primary, auxiliary})
    //   -> (ct_g1[msg_size + 2], proof)       encrypt_op::process, elgamal_verifiable.hpp:322-357
    template<typename Scheme, typename Mode, typename SinglePassRange>
    typename Scheme::cipher_type encrypt(const SinglePassRange &plain_text, const encrypt_init_params<Scheme> &p) {
        typedef typename Scheme::scalar_field_type::value_type fr;
        const std::vector<fr> m(std::begin(plain_text), std::end(plain_text));
        const auto &vk = p.gg_keypair.second;
        // the reference's BOOST_ASSERT_MSG checks, elgamal_verifiable.hpp:323-339
        if (!(vk.gamma_ABC_g1.rest.size() > m.size()))
            throw std::invalid_argument("Array of gammas in vk should be longer than the plain text.");
        if (!(p.primary_input.size() > m.size()))
            throw std::invalid_argument("Primary input should be longer than plain text.");
        if (vk.gamma_ABC_g1.rest.size() != p.primary_input.size())
            throw std::invalid_argument("Number of gammas should be equal to the primary input size.");
        if (m.size() != p.pubkey.delta_s_g1.size() || m.size() != p.pubkey.t_g1.size() || m.size() + 1 != p.pubkey.t_g2.size())
            throw std::invalid_argument("Plain text size should be equal to the delta_s / t_g1 / t_g2 array sizes from pk.");
        for (std::size_t i = 0; i < m.size(); ++i)
            if (p.primary_input[i] != m[i]) throw std::invalid_argument("Plain text should be a prefix of primary input.");

        auto &e = zk::detail::blinding_engine();
        const fr r = p.groth16_r ? *p.groth16_r : b200::random_fr(e), s = p.groth16_s ? *p.groth16_s : b200::random_fr(e);
        auto &d = b200::device::instance();
        const vsp_saver_pubkey_t desc = p.pubkey.descriptor();
        typename Scheme::cipher_type out;
        out.first.resize(m.size() + 2);
        std::uint64_t proof[48];
        d.check(vsp_saver_encrypt(d.ctx(), p.gg_keypair.first.resident(), &desc, vk.gamma_ABC_g1.rest[0].l.data(),
                                  m.empty() ? nullptr : m[0].l.data(), p.primary_input[0].l.data(),
                                  p.auxiliary_input.empty() ? nullptr : p.auxiliary_input[0].l.data(), p.r.l.data(),
                                  r.l.data(), s.l.data(), out.first[0].l.data(), proof),
                "encrypt");
        std::memcpy(out.second.g_A.l.data(), proof, 96);
        std::memcpy(out.second.g_B.l.data(), proof + 12, 192);
        std::memcpy(out.second.g_C.l.data(), proof + 36, 96);
        return out;
    }

    template<typename Scheme>
    struct rerandomize_init_params {    // rerandomize_op::init_params_type, elgamal_verifiable.hpp:600-606
        const public_key<Scheme> &pubkey;
        const typename Scheme::proof_system_type::keypair_type &gg_keypair;
        const typename Scheme::proof_system_type::proof_type &proof;
    };

    // rerandomize<Scheme>(rnd = (r, z1, z2, ...), cipher_text, {pubkey, gg_keypair, proof}) -> (ct', proof')
    template<typename Scheme, typename RandomRange, typename CipherRange>
    typename Scheme::cipher_type rerandomize(const RandomRange &rnd, const CipherRange &cipher_text,
                                             const rerandomize_init_params<Scheme> &p) {
        typedef typename Scheme::scalar_field_type::value_type fr;
        typedef typename Scheme::g1_value_type g1;
        const std::vector<fr> rr(std::begin(rnd), std::end(rnd));
        const std::vector<g1> ct(std::begin(cipher_text), std::end(cipher_text));
        if (rr.size() < 3) throw std::invalid_argument("Too few numbers in the source of randomness (at least 3).");
        if (ct.size() < 2 || p.pubkey.delta_s_g1.size() != ct.size() - 2 || p.pubkey.t_g1.size() != ct.size() - 2 ||
            p.pubkey.t_g2.size() != ct.size() - 1)
            throw std::invalid_argument("Cipher text size should match the public key (exclusive of 2 elements in CT).");
        auto &d = b200::device::instance();
        const vsp_saver_pubkey_t desc = p.pubkey.descriptor();
        std::uint64_t proof_in[48], proof_out[48];
        std::memcpy(proof_in, p.proof.g_A.l.data(), 96);
        std::memcpy(proof_in + 12, p.proof.g_B.l.data(), 192);
        std::memcpy(proof_in + 36, p.proof.g_C.l.data(), 96);
        typename Scheme::cipher_type out;
        out.first.resize(ct.size());
        d.check(vsp_saver_rerandomize(d.ctx(), &desc, p.gg_keypair.second.delta_g2.l.data(), rr[0].l.data(), ct[0].l.data(),
                                      proof_in, out.first[0].l.data(), proof_out),
                "rerandomize");
        std::memcpy(out.second.g_A.l.data(), proof_out, 96);
        std::memcpy(out.second.g_B.l.data(), proof_out + 12, 192);
        std::memcpy(out.second.g_C.l.data(), proof_out + 36, 96);
        return out;
    }

    // ---- pairing side: verify_encryption / decrypt / verify_decryption ---------------------------------
    namespace detail {
        template<typename VK>
        inline vsp_verification_key_t vk_descriptor(const VK &vk, std::vector<b200::g1_value> &gabc) {
            gabc.clear();
            gabc.push_back(vk.gamma_ABC_g1.first);
            gabc.insert(gabc.end(), vk.gamma_ABC_g1.rest.begin(), vk.gamma_ABC_g1.rest.end());
            vsp_verification_key_t d;
            d.alpha_g1 = vk.alpha_g1.l.data();
            d.beta_g2 = vk.beta_g2.l.data();
            d.gamma_g2 = vk.gamma_g2.l.data();
            d.delta_g2 = vk.delta_g2.l.data();
            d.gamma_ABC_count = gabc.size();
            d.gamma_ABC_g1 = gabc[0].l.data();
            return d;
        }
        template<typename Proof>
        inline void proof_limbs(const Proof &p, std::uint64_t out[48]) {
            std::memcpy(out, p.g_A.l.data(), 96);
            std::memcpy(out + 12, p.g_B.l.data(), 192);
            std::memcpy(out + 36, p.g_C.l.data(), 96);
        }
    }    // namespace detail

    template<typename Scheme>
    struct verify_encryption_init_params {    // verify_encryption_op::init_params_type, elgamal_verifiable.hpp:453-458
        const public_key<Scheme> &pubkey;
        const typename Scheme::proof_system_type::verification_key_type &gg_vk;
        const typename Scheme::proof_system_type::proof_type &proof;
        const typename Scheme::proof_system_type::primary_input_type &unencrypted_primary_input;
    };
    // verify_encryption<Scheme>(cipher_text, {pubkey, gg_vk, proof, unencrypted_primary_input}) -> bool
    //   verify_encryption_op::process, elgamal_verifiable.hpp:486-490; called at bin/cli common.hpp:1164-1168
    template<typename Scheme, typename CipherRange>
    bool verify_encryption(const CipherRange &cipher_text, const verify_encryption_init_params<Scheme> &p) {
        const std::vector<typename Scheme::g1_value_type> ct(std::begin(cipher_text), std::end(cipher_text));
        if (ct.size() != p.pubkey.delta_s_g1.size() + 2 || p.pubkey.t_g2.size() + 1 != ct.size())
            throw std::invalid_argument("Cipher text size should match the public key (exclusive of 2 elements in CT).");
        std::vector<b200::g1_value> gabc;
        const vsp_verification_key_t vk = detail::vk_descriptor(p.gg_vk, gabc);
        const vsp_saver_pubkey_t desc = p.pubkey.descriptor();
        std::uint64_t proof[48];
        detail::proof_limbs(p.proof, proof);
        std::uint8_t ok = 0;
        auto &d = b200::device::instance();
        const auto &x = p.unencrypted_primary_input;
        if (p.gg_vk.has_alpha_g1_beta_g2)
            d.check(vsp_saver_verify_encryption_gt_batch(d.ctx(), p.gg_vk.alpha_g1_beta_g2.data(), vk.gamma_g2, vk.delta_g2,
                                                         vk.gamma_ABC_count, vk.gamma_ABC_g1, &desc, 1, ct[0].l.data(),
                                                         x.empty() ? nullptr : x[0].l.data(), x.size(), proof, &ok),
                    "verify_encryption");
        else
            d.check(vsp_saver_verify_encryption_batch(d.ctx(), &vk, &desc, 1, ct[0].l.data(),
                                                      x.empty() ? nullptr : x[0].l.data(), x.size(), proof, &ok),
                    "verify_encryption");
        return ok != 0;
    }

    template<typename Scheme>
    struct decrypt_init_params {    // decrypt_op::init_params_type, elgamal_verifiable.hpp:371-375
        const private_key<Scheme> &privkey;
        const verification_key<Scheme> &vk;
        const typename Scheme::proof_system_type::keypair_type &gg_keypair;
    };
    // decrypt<Scheme, modes::verifiable_encryption<Scheme>>(cipher_text, {privkey, vk, gg_keypair})
    //   -> (plain text, decryption proof = rho * ct[0])    decrypt_op::process, elgamal_verifiable.hpp:401-441;
    //   called on the tally's aggregated cipher text at bin/cli common.hpp:1221-1223.  BlockBits: scheme_type::block_bits.
    template<typename Scheme, typename Mode, std::size_t BlockBits = 4, typename CipherRange>
    std::pair<std::vector<typename Scheme::scalar_field_type::value_type>, typename Scheme::g1_value_type>
        decrypt(const CipherRange &cipher_text, const decrypt_init_params<Scheme> &p) {
        typedef typename Scheme::scalar_field_type::value_type fr;
        const std::vector<typename Scheme::g1_value_type> ct(std::begin(cipher_text), std::end(cipher_text));
        const auto &gabc = p.gg_keypair.second.gamma_ABC_g1.rest;
        // the reference's BOOST_ASSERT_MSG checks, elgamal_verifiable.hpp:403-412
        if (ct.size() < 2 || !(gabc.size() > ct.size() - 2))
            throw std::invalid_argument("Array of gammas in vk should be longer than the cipher text (exclusive of 2 element in CT).");
        if (ct.size() - 2 != p.vk.rho_sv_g2.size() || ct.size() - 2 != p.vk.rho_rhov_g2.size())
            throw std::invalid_argument("Cipher text size should be equal to the rho_sv_g2 / rho_rhov_g2 array size from vk.");
        const std::size_t n = ct.size() - 2;
        std::pair<std::vector<fr>, typename Scheme::g1_value_type> out;
        out.first.resize(n);
        auto &d = b200::device::instance();
        d.check(vsp_saver_decrypt(d.ctx(), n, ct[0].l.data(), p.privkey.rho.l.data(), n ? p.vk.rho_sv_g2[0].l.data() : nullptr,
                                  n ? p.vk.rho_rhov_g2[0].l.data() : nullptr, n ? gabc[0].l.data() : nullptr,
                                  static_cast<unsigned>(BlockBits), n ? out.first[0].l.data() : nullptr, out.second.l.data()),
                "decrypt");
        return out;
    }

    template<typename Scheme>
    struct verify_decryption_init_params {    // verify_decryption_op::init_params_type, elgamal_verifiable.hpp:506-510
        const verification_key<Scheme> &vk;
        const typename Scheme::proof_system_type::keypair_type &gg_keypair;
        const typename Scheme::g1_value_type &proof;
    };
    // verify_decryption<Scheme>(cipher_text, plain_text, {vk, gg_keypair, proof}) -> bool
    //   verify_decryption_op::process, elgamal_verifiable.hpp:548-575; called at bin/cli common.hpp:1282-1283
    template<typename Scheme, typename CipherRange, typename PlainRange>
    bool verify_decryption(const CipherRange &cipher_text, const PlainRange &plain_text,
                           const verify_decryption_init_params<Scheme> &p) {
        typedef typename Scheme::scalar_field_type::value_type fr;
        const std::vector<typename Scheme::g1_value_type> ct(std::begin(cipher_text), std::end(cipher_text));
        const std::vector<fr> m(std::begin(plain_text), std::end(plain_text));
        const auto &gabc = p.gg_keypair.second.gamma_ABC_g1.rest;
        if (m.size() + 2 != ct.size())
            throw std::invalid_argument("Cipher text size should be equal to the plain text size (exclusive of 2 element in CT).");
        if (!(gabc.size() > m.size())) throw std::invalid_argument("Array of gammas in vk should be longer than the plain text.");
        if (p.vk.rho_sv_g2.size() < m.size() || p.vk.rho_rhov_g2.size() < m.size())
            throw std::invalid_argument("verification key shorter than the plain text");
        std::uint8_t ok = 0;
        auto &d = b200::device::instance();
        const std::size_t n = m.size();
        d.check(vsp_saver_verify_decryption(d.ctx(), n, ct[0].l.data(), n ? m[0].l.data() : nullptr, p.vk.rho_g2.l.data(),
                                            n ? p.vk.rho_sv_g2[0].l.data() : nullptr, n ? p.vk.rho_rhov_g2[0].l.data() : nullptr,
                                            n ? gabc[0].l.data() : nullptr, p.proof.l.data(), &ok),
                "verify_decryption");
        return ok != 0;
    }
}    // namespace pubkey

}    // namespace crypto3
}    // namespace nil
#endif    // VSP_CRYPTO3_SHIM_HPP
