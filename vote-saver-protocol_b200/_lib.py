"""ctypes loader for libvsp_b200.so (the C-ABI product library, include/vsp_b200.h).

There is no CPU fallback: if the library is missing, or no CUDA device is usable, the caller gets an exception.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvsp_b200.so")

u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)

VSP_OK = 0
VSP_E_INVALID = -1
VSP_E_NO_DEVICE = -2
VSP_E_CUDA = -3
VSP_E_DOMAIN = -4
VSP_E_NOMEM = -5

DOMAIN_BASIC_RADIX2 = 1
DOMAIN_EXTENDED_RADIX2 = 2
DOMAIN_STEP_RADIX2 = 3


class R1CSDesc(C.Structure):
    """vsp_r1cs_t"""
    _fields_ = [("num_inputs", C.c_uint64), ("num_aux", C.c_uint64), ("num_constraints", C.c_uint64),
                ("a_row", u64p), ("b_row", u64p), ("c_row", u64p),
                ("a_col", u32p), ("b_col", u32p), ("c_col", u32p),
                ("a_val", u64p), ("b_val", u64p), ("c_val", u64p)]


class ProvingKeyDesc(C.Structure):
    """vsp_proving_key_t"""
    _fields_ = [("cs", R1CSDesc),
                ("alpha_g1", u64p), ("beta_g1", u64p), ("beta_g2", u64p), ("delta_g1", u64p), ("delta_g2", u64p),
                ("A_query", u64p), ("B_count", C.c_uint64), ("B_idx", u32p), ("B_g2", u64p), ("B_g1", u64p),
                ("H_count", C.c_uint64), ("H_query", u64p), ("L_query", u64p)]


class PkSliceDesc(C.Structure):
    """vsp_pk_slice_t"""
    _fields_ = [(n, C.c_uint64) for n in ("A_lo", "A_hi", "B_lo", "B_hi", "H_lo", "H_hi", "L_lo", "L_hi")]


class FastPkSizes(C.Structure):
    """vsp_fast_pk_sizes_t"""
    _fields_ = [(n, C.c_uint64) for n in ("num_inputs", "num_aux", "num_constraints", "nnz_a", "nnz_b", "nnz_c", "A_count",
                                          "B_count", "B_domain_size", "H_count", "L_count", "bytes")]


class SaverPubkeyDesc(C.Structure):
    """vsp_saver_pubkey_t"""
    _fields_ = [("msg_size", C.c_uint64), ("delta_g1", u64p), ("delta_s_g1", u64p), ("t_g1", u64p), ("t_g2", u64p),
                ("delta_sum_s_g1", u64p), ("gamma_inverse_sum_s_g1", u64p)]


class KeypairDesc(C.Structure):
    """vsp_keypair_t"""
    _fields_ = [("alpha_g1", u64p), ("beta_g1", u64p), ("beta_g2", u64p), ("delta_g1", u64p), ("delta_g2", u64p),
                ("A_query", u64p), ("B_count", C.c_uint64), ("B_idx", u32p), ("B_g2", u64p), ("B_g1", u64p),
                ("H_count", C.c_uint64), ("H_query", u64p), ("L_query", u64p),
                ("gamma_g2", u64p), ("gamma_g1", u64p), ("gamma_ABC_g1", u64p)]


class SaverKeypairDesc(C.Structure):
    """vsp_saver_keypair_t"""
    _fields_ = [(n, u64p) for n in ("delta_s_g1", "t_g1", "t_g2", "delta_sum_s_g1", "gamma_inverse_sum_s_g1", "rho_g2",
                                    "rho_sv_g2", "rho_rhov_g2", "rho")]


class VerificationKeyDesc(C.Structure):
    """vsp_verification_key_t"""
    _fields_ = [("alpha_g1", u64p), ("beta_g2", u64p), ("gamma_g2", u64p), ("delta_g2", u64p),
                ("gamma_ABC_count", C.c_uint64), ("gamma_ABC_g1", u64p)]


# every symbol include/vsp_b200.h declares: (name, restype, argtypes)
_vp = C.c_void_p
_u64 = C.c_uint64
_int = C.c_int
SYMBOLS = [
    ("vsp_ctx_create", _int, [_int, _vp, C.POINTER(_vp)]),
    ("vsp_ctx_destroy", None, [_vp]),
    ("vsp_last_error", C.c_char_p, [_vp]),
    ("vsp_last_warning", C.c_char_p, [_vp]),
    ("vsp_ctx_set_batch_chunk", _int, [_vp, _int]),
    ("vsp_ctx_synchronize", _int, [_vp]),
    ("vsp_ctx_trim", _int, [_vp]),
    ("vsp_ctx_scratch_stats", _int, [_vp, C.POINTER(_u64), C.POINTER(_u64), C.POINTER(_u64)]),
    ("vsp_ctx_launch_count", _u64, [_vp]),
    ("vsp_version", C.c_char_p, []),
    ("vsp_ctx_profile_enable", _int, [_vp, _int]),
    ("vsp_ctx_profile_read", _int, [_vp, _int, C.POINTER(_u64), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    ("vsp_domain_plan", _int, [_u64, C.POINTER(_int), C.POINTER(_u64)]),
    ("vsp_fr_fft", _int, [_vp, _u64, _int, _vp, _u64, _vp]),
    ("vsp_fr_fft_dev", _int, [_vp, _u64, _int, _vp]),
    ("vsp_fr_fft_noalloc_dev", _int, [_vp, _u64, _int, _vp]),
    ("vsp_fr_multiply_by_coset", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_fr_multiply_by_coset_dev", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_fr_divide_by_z_on_coset", _int, [_vp, _u64, _vp]),
    ("vsp_fr_divide_by_z_on_coset_dev", _int, [_vp, _u64, _vp]),
    ("vsp_g1_msm", _int, [_vp, _u64, _vp, _vp, _vp]),
    ("vsp_g2_msm", _int, [_vp, _u64, _vp, _vp, _vp]),
    ("vsp_g1_bases_create", _int, [_vp, _u64, _vp, C.POINTER(_vp)]),
    ("vsp_g1_bases_destroy", None, [_vp]),
    ("vsp_g2_bases_create", _int, [_vp, _u64, _vp, C.POINTER(_vp)]),
    ("vsp_g1_bases_create_ex", _int, [_vp, _u64, _vp, _int, C.POINTER(_vp)]),
    ("vsp_g2_bases_create_ex", _int, [_vp, _u64, _vp, _int, C.POINTER(_vp)]),
    ("vsp_g2_bases_destroy", None, [_vp]),
    ("vsp_g1_bases_generate", _int, [_vp, _u64, _u64, C.POINTER(_vp)]),
    ("vsp_g1_bases_generate_ex", _int, [_vp, _u64, _u64, _int, C.POINTER(_vp)]),
    ("vsp_g1_bases_read", _int, [_vp, _vp, _u64, _u64, _vp]),
    ("vsp_g2_points_generate", _int, [_vp, _u64, _u64, _vp]),
    ("vsp_g1_msm_bases", _int, [_vp, _vp, _u64, _vp, _int, _vp]),
    ("vsp_g2_msm_bases", _int, [_vp, _vp, _u64, _vp, _int, _vp]),
    ("vsp_g1_msm_bases_dev", _int, [_vp, _vp, _u64, _vp, _vp]),
    ("vsp_g2_msm_bases_dev", _int, [_vp, _vp, _u64, _vp, _vp]),
    ("vsp_g1_sum", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_g2_sum", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_g1_sum_dev", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_g2_sum_dev", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_pk_load", _int, [_vp, C.POINTER(ProvingKeyDesc), C.POINTER(_vp)]),
    ("vsp_pk_load_slice", _int, [_vp, C.POINTER(ProvingKeyDesc), C.POINTER(PkSliceDesc), C.POINTER(_vp)]),
    ("vsp_prove_partial", _int, [_vp, _vp, _vp, _vp, _vp]),
    ("vsp_prove_assemble", _int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_prove_partial_dev", _int, [_vp, _vp, _vp, _vp]),
    ("vsp_prove_partial_begin_dev", _int, [_vp, _vp, _vp, C.c_uint, _vp, _vp]),
    ("vsp_prove_partial_finish_dev", _int, [_vp, _vp, _vp, _vp]),
    ("vsp_prove_assemble_dev", _int, [_vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_pk_destroy", None, [_vp]),
    ("vsp_pk_domain_size", _u64, [_vp]),
    ("vsp_witness_map", _int, [_vp, _vp, _vp, _vp, _vp]),
    ("vsp_prove", _int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_prove_batch", _int, [_vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_prove_batch_dev", _int, [_vp, _vp, _u64, _vp, _vp, _vp, _vp]),
    ("vsp_saver_encrypt", _int, [_vp, _vp, C.POINTER(SaverPubkeyDesc), _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_saver_ciphertext", _int, [_vp, C.POINTER(SaverPubkeyDesc), _vp, _vp, _vp, _vp]),
    ("vsp_saver_pub_load", _int, [_vp, C.POINTER(SaverPubkeyDesc), _vp, C.POINTER(_vp)]),
    ("vsp_saver_pub_destroy", None, [_vp]),
    ("vsp_saver_encrypt_batch", _int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_saver_encrypt_batch_dev", _int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp]),
    ("vsp_saver_rerandomize", _int, [_vp, C.POINTER(SaverPubkeyDesc), _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_saver_rerandomize_batch", _int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_saver_rerandomize_batch_dev", _int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_groth16_generate", _int, [_vp, C.POINTER(R1CSDesc), _vp, C.POINTER(KeypairDesc)]),
    ("vsp_saver_keygen", _int, [_vp, _u64, _vp, _vp, _vp, _vp, C.POINTER(SaverKeypairDesc)]),
    ("vsp_g1_mul_batch", _int, [_vp, _u64, _vp, _vp, _vp]),
    ("vsp_g1_generator_mul_batch", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_g2_generator_mul_batch", _int, [_vp, _u64, _vp, _vp]),
    ("vsp_pairing_products", _int, [_vp, _u64, _u64, _vp, _vp, _vp, _vp]),
    ("vsp_groth16_verify_batch", _int, [_vp, C.POINTER(VerificationKeyDesc), _u64, _vp, _u64, _vp, _vp]),
    ("vsp_groth16_verify_gt_batch", _int, [_vp, _vp, _vp, _vp, _u64, _vp, _u64, _vp, _u64, _vp, _vp]),
    ("vsp_saver_verify_encryption_batch", _int, [_vp, C.POINTER(VerificationKeyDesc), C.POINTER(SaverPubkeyDesc), _u64, _vp,
                                                 _vp, _u64, _vp, _vp]),
    ("vsp_saver_verify_encryption_gt_batch", _int, [_vp, _vp, _vp, _vp, _u64, _vp, C.POINTER(SaverPubkeyDesc), _u64, _vp, _vp,
                                                    _u64, _vp, _vp]),
    ("vsp_saver_decrypt", _int, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, C.c_uint, _vp, _vp]),
    ("vsp_saver_verify_decryption", _int, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("vsp_voting_program_create", _int, [_vp, _vp, _u64, _vp, _u64, _vp, _u64, _vp, _u64, _vp, _u64, C.POINTER(_vp)]),
    ("vsp_voting_program_destroy", None, [_vp]),
    ("vsp_voting_program_num_vars", _u64, [_vp]),
    ("vsp_voting_witness_batch_dev", _int, [_vp, _vp, _u64, _vp, _u64, _vp, _vp, _vp, _vp]),
    ("vsp_codec_g1_compress", _int, [_vp, _vp]),
    ("vsp_codec_g2_compress", _int, [_vp, _vp]),
    ("vsp_codec_proof_write", _int, [_vp, _vp]),
    ("vsp_codec_g1_decompress", _int, [_vp, _vp]),
    ("vsp_codec_g2_decompress", _int, [_vp, _vp]),
    ("vsp_codec_proof_read", _int, [_vp, _vp]),
    ("vsp_codec_ciphertext_write", _int, [_vp, _u64, _vp]),
    ("vsp_codec_primary_input_write", _int, [_vp, _u64, _vp]),
    ("vsp_codec_primary_input_read", _int, [_vp, C.c_size_t, C.POINTER(_u64), _vp, _u64]),
    ("vsp_codec_fast_pk_sizes", _int, [_vp, C.c_size_t, C.POINTER(FastPkSizes)]),
    ("vsp_codec_fast_pk_read", _int, [_vp, C.c_size_t, C.POINTER(ProvingKeyDesc)]),
    ("vsp_codec_fast_pk_write", _int, [C.POINTER(ProvingKeyDesc), _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    ("vsp_codec_ciphertext_read", _int, [_vp, C.c_size_t, C.POINTER(_u64), _vp, _u64]),
    ("vsp_codec_saver_pubkey_write", _int, [C.POINTER(SaverPubkeyDesc), _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    ("vsp_codec_saver_pubkey_read", _int, [_vp, C.c_size_t, C.POINTER(_u64), C.POINTER(SaverPubkeyDesc)]),
    ("vsp_codec_saver_vk_write", _int, [_u64, _vp, _vp, _vp, _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    ("vsp_codec_saver_vk_read", _int, [_vp, C.c_size_t, C.POINTER(_u64), _vp, _vp, _vp, _u64]),
    ("vsp_codec_saver_sk_write", _int, [_vp, _vp]),
    ("vsp_codec_saver_sk_read", _int, [_vp, C.c_size_t, _vp]),
    ("vsp_codec_groth16_vk_write", _int, [_vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    ("vsp_codec_groth16_vk_read", _int, [_vp, C.c_size_t, C.POINTER(_u64), _vp, _vp, _vp, _vp, _vp, _vp, _u64]),
    ("vsp_codec_bitarray_write", _int, [_vp, _u64, _vp]),
    ("vsp_codec_bitarray_read", _int, [_vp, C.c_size_t, _u64, _vp]),
    ("vsp_codec_bits_to_field_elements", _int, [_vp, _u64, _vp]),
    ("vsp_codec_merkle_tree_write", _int, [_vp, _u64, _vp]),
    ("vsp_codec_merkle_tree_read", _int, [_vp, C.c_size_t, _u64, _vp]),
    ("vsp_codec_merkle_copath", _int, [_vp, _u64, _u64, _vp]),
    ("vsp_measure_imad_rate", _int, [_vp, _u64, C.POINTER(C.c_double)]),
    ("vsp_measure_dfma_rate", _int, [_vp, _u64, C.POINTER(C.c_double)]),
    ("vsp_accumulate_mads_per_add", _int, [_vp, C.POINTER(C.c_double)]),
]

_LIB = None


class VspError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__("vsp error %d: %s" % (code, text))
        self.code = code


def load():
    """dlopen libvsp_b200.so and type every export.  Raises if the library has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (or make -C "
                              "vote-saver-protocol_b200/csrc).  There is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB
