"""ctypes loader for libbiscotti_b200.so (C-ABI declared in include/biscotti_b200.h).

The library is built in-tree by ``biscotti_b200.build``; it is loaded lazily so that importing the
package never compiles anything.  On a machine with a GPU the CUDA entry points are mandatory: ops
raise if the library is missing instead of silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

_LIB = None
_PATH = Path(__file__).resolve().parent / "lib" / "libbiscotti_b200.so"


def lib_path() -> Path:
    return _PATH


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _PATH.exists():
        if not build_if_missing or os.environ.get("BISCOTTI_NO_BUILD"):
            raise RuntimeError(f"{_PATH} is missing; run `python -m biscotti_b200.build`")
        from . import build as _b

        _b.build()
    _LIB = ctypes.CDLL(str(_PATH))
    _declare(_LIB)
    return _LIB


def available() -> bool:
    try:
        load()
        return True
    except Exception:
        return False


c_i = ctypes.c_int
c_ll = ctypes.c_longlong
c_ull = ctypes.c_ulonglong
c_f = ctypes.c_float
c_d = ctypes.c_double
c_p = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_cp = ctypes.c_char_p


def _declare(L: ctypes.CDLL) -> None:
    def sig(name, res, *args):
        f = getattr(L, name, None)
        if f is None:
            return
        f.restype = res
        f.argtypes = list(args)

    sig("bsc_launch_count", c_ull)
    sig("bsc_launch_count_reset", None)
    # ML kernels
    sig("bsc_local_softmax_step", c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_f, c_p)
    sig("bsc_local_logistic_step", c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_f, c_f, c_f, c_ull, c_p, c_i, c_p)
    sig("bsc_noise_mix", c_i, c_p, c_p, c_i, c_ull, c_i, c_f, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p)
    sig("bsc_noise_mix_rows", c_i, c_p, c_p, c_i, c_ull, c_i, c_f, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p)
    sig("bsc_sample_minibatch", c_i, c_ull, c_i, c_p, c_i, c_i, c_i, c_p, c_p)
    sig("bsc_noise_row", c_i, c_p, c_i, c_ull, c_i, c_i, c_f, c_p)
    sig("bsc_noise_row_host", None, c_p, c_i, c_ull, c_i, c_i, c_f)
    sig("bsc_gram_simt", c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_p)
    sig("bsc_krum_select", c_i, c_p, c_i, c_i, c_p, c_i, c_i, c_p, c_p, c_p, c_p)
    sig("bsc_krum_select_mc", c_i, c_p, c_i, c_i, c_p, c_p, c_i, c_i, c_p, c_p, c_p, c_p)
    sig("bsc_eval_softmax", c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_p)
    sig("bsc_eval_logistic", c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_p)
    sig("bsc_cnn1_scratch_floats", c_ll, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i)
    sig("bsc_cnn1_local_step", c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_f, c_p)
    sig("bsc_cnn2_scratch_floats", c_ll, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i)
    sig("bsc_cnn2_local_step", c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i,
        c_i, c_f, c_p)
    # secure aggregation
    sig("bsc_quantize", c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_d, c_p)
    sig("bsc_shares_eval", c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_p)
    sig("bsc_share_sum", c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_d, c_p, c_p)
    sig("bsc_reconstruct_step", c_i, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_d, c_p, c_p, c_p, c_p, c_p)
    sig("bsc_plain_sum_step", c_i, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p)
    sig("bsc_quantize_host", None, c_p, c_p, c_i, c_i, c_i, c_i, c_d)
    sig("bsc_shares_eval_host", None, c_p, c_p, c_i, c_i, c_i, c_i)
    sig("bsc_shamir_recover_host", c_i, c_p, c_p, c_p, c_i, c_i, c_p)
    # MSM
    sig("bsc_msm_groups", c_i, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_ll, c_p)
    sig("bsc_g1_fixed_base_mul", c_i, c_p, c_i, c_p, c_p, c_p)
    sig("bsc_g1_normalize", c_i, c_p, c_p, c_ll, c_p)
    sig("bsc_g1_affine_to_bytes", c_i, c_p, c_p, c_ll, c_p)
    sig("bsc_g1_bytes_to_affine", c_i, c_p, c_p, c_ll, c_p)
    sig("bsc_g1_sum_rows", c_i, c_p, c_p, c_i, c_ll, c_p, c_p)
    sig("bsc_g1_sum_cols", c_i, c_p, c_i, c_i, c_p, c_p)
    sig("bsc_g1_sum_rows_jac", c_i, c_p, c_p, c_i, c_ll, c_p, c_p)
    sig("bsc_g1_sum_cols_jac", c_i, c_p, c_i, c_i, c_p, c_p)
    sig("bsc_g1_sum_rows_jac_host", None, c_p, c_p, c_i, c_ll, c_p)
    sig("bsc_g1_sum_cols_jac_host", None, c_p, c_i, c_i, c_p)
    sig("bsc_msm_table_build", c_i, c_p, c_ll, c_i, c_i, c_p, c_p)
    sig("bsc_msm_table_build_host", None, c_p, c_ll, c_i, c_i, c_p)
    sig("bsc_msm_table", c_i, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_p, c_i, c_i, c_p, c_ll, c_p)
    sig("bsc_msm_table_host", None, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_p, c_i, c_i, c_p, c_ll)
    sig("bsc_msm_groups_host", None, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_ll)
    sig("bsc_g1_normalize_host", None, c_p, c_p, c_ll)
    sig("bsc_g1_sum_rows_host", None, c_p, c_p, c_i, c_ll, c_p)
    sig("bsc_g1_sum_cols_host", None, c_p, c_i, c_i, c_p)
    sig("bsc_g1_fixed_base_mul_host", None, c_p, c_i, c_p, c_p)
    # VRF / hashing
    sig("bsc_sha512", None, c_cp, c_sz, c_p)
    sig("bsc_vrf_pubkey", None, c_cp, c_p)
    sig("bsc_vrf_secret_scalar", None, c_cp, c_p)
    sig("bsc_vrf_hash_to_curve", None, c_cp, c_sz, c_p, c_p)
    sig("bsc_vrf_output_from_gamma", None, c_cp, c_cp, c_sz, c_p)
    sig("bsc_vrf_prove", None, c_cp, c_cp, c_sz, c_p, c_p)
    sig("bsc_vrf_verify", c_i, c_cp, c_cp, c_sz, c_cp, c_cp)
    sig("bsc_ed25519_scalarmult_batch", c_i, c_p, c_p, c_i, c_i, c_p, c_p)
    sig("bsc_ed25519_scalarmult_batch_host", None, c_p, c_p, c_i, c_i, c_p)
    sig("bsc_vrf_hash_to_curve_xy", None, c_cp, c_sz, c_p)
    sig("bsc_vrf_launch", c_i, c_cp, c_i, c_p, c_p, c_p, c_i, c_p, c_i, c_p, c_p)
    sig("bsc_ed25519_scalarmult_r25", c_i, c_p, c_p, c_i, c_p, c_p)
    sig("bsc_ed25519_scalarmult_quad", c_i, c_p, c_p, c_i, c_p, c_p)
    sig("bsc_ed25519_scalarmult_quad_xyz", c_i, c_p, c_p, c_i, c_p, c_p)
    sig("bsc_ed25519_scalarmult_wide_xyz", c_i, c_p, c_p, c_i, c_p, c_p)
    sig("bsc_ed25519_compress_batch", None, c_p, c_i, c_p)
    sig("bsc_ed25519_scalarmult_r25_host", None, c_p, c_p, c_i, c_p)
    sig("bsc_sha256_batch", c_i, c_p, c_i, c_ll, c_i, c_p, c_p)
    sig("bsc_noisers_batch", None, c_p, c_cp, c_sz, c_p, c_i, c_p, c_i, c_p, c_p)
    sig("bsc_gram_tc", c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_i, c_p)
    sig("bsc_split_planes", c_i, c_p, c_p, c_i, c_i, c_i, c_ll, c_i, c_i, c_p, c_p, c_i, c_i, c_p, c_i, c_p)
    sig("bsc_eval_tc", c_i, c_p, c_p, c_i, c_p, c_p, c_i, c_i, c_p, c_p, c_p, c_i, c_i, c_p)
    sig("bsc_lottery_roles", c_i, c_cp, c_i, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p)
    sig("bsc_lottery_noisers", c_i, c_cp, c_i, c_p, c_i, c_i, c_i, c_p)
    # host crypto
    sig("bsc_sha256", None, c_cp, c_sz, c_p)
    sig("bsc_blake2b", None, c_cp, c_sz, c_cp, c_sz, c_p, c_sz)
    sig("bsc_blake2xb", None, c_cp, c_sz, c_cp, c_sz, c_p, c_sz)
    sig("bsc_shake256", None, c_cp, c_sz, c_p, c_sz)
    sig("bsc_g1_generator", None, c_p)
    sig("bsc_g1_on_curve", c_i, c_cp)
    sig("bsc_g1_add", None, c_p, c_cp, c_cp)
    sig("bsc_g1_neg", None, c_p, c_cp)
    sig("bsc_g1_mul", None, c_p, c_cp, c_cp)
    sig("bsc_g1_msm_i64_host", None, c_p, c_cp, c_p, c_i)
    sig("bsc_g1_bytes_to_affine_host", None, c_cp, c_p, c_ll)
    sig("bsc_g1_affine_to_bytes_host", None, c_p, c_p, c_ll)
    sig("bsc_g1_jac_to_bytes_host", None, c_p, c_p, c_ll)
    sig("bsc_g1_jac_to_bytes_batch_host", None, c_p, c_p, c_i, c_p)
    sig("bsc_g1_straus_host", None, c_p, c_p, c_p, c_i)
    # G2 / GT / pairing (host)
    sig("bsc_g2_generator", None, c_p)
    sig("bsc_g2_on_curve", c_i, c_cp)
    sig("bsc_g2_add", None, c_p, c_cp, c_cp)
    sig("bsc_g2_neg", None, c_p, c_cp)
    sig("bsc_g2_mul", None, c_p, c_cp, c_cp)
    sig("bsc_srs_g2_host", None, c_cp, c_i, c_p)
    sig("bsc_pairing", None, c_cp, c_cp, c_p)
    sig("bsc_pairing_check", c_i, c_cp, c_cp, c_i)
    sig("bsc_gt_mul", None, c_p, c_cp, c_cp)
    sig("bsc_gt_pow", None, c_p, c_cp, c_cp)
    sig("bsc_kzg_verify_share", c_i, c_cp, c_cp, c_cp, c_cp, c_cp, c_i, c_ll)
    sig("bsc_srs_g1_host", None, c_cp, c_i, c_p)
    sig("bsc_fr_powers", None, c_cp, c_i, c_p)
    sig("bsc_schnorr_pubkey", None, c_cp, c_p)
    sig("bsc_schnorr_sign", None, c_cp, c_sz, c_cp, c_cp, c_p)
    sig("bsc_schnorr_nonces", None, c_i, c_p, c_i, c_p, c_p, c_p)
    sig("bsc_schnorr_finish_batch", None, c_i, c_p, c_i, c_p, c_p, c_p, c_p)
    sig("bsc_g1_comb_mul", c_i, c_p, c_i, c_p, c_i, c_i, c_p, c_p)
    sig("bsc_schnorr_verify", c_i, c_cp, c_sz, c_cp, c_cp)
    # device-resident block chain (kernels_chain.cu) and the device-count variants of the aggregation kernels
    sig("bsc_vrf_noisers", c_i, c_p, c_p, c_i, c_p, c_i, c_p, c_p, c_i, c_i, c_p, c_p, c_p)
    sig("bsc_block_plan", c_i, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_p)
    sig("bsc_schnorr_nonces_dev", c_i, c_p, c_cp, c_ll, c_p, c_i, c_p, c_i, c_p, c_p)
    sig("bsc_schnorr_finish_dev", c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_i, c_p, c_i, c_p)
    sig("bsc_pack_commitments", c_i, c_p, c_i, c_p, c_i, c_p)
    sig("bsc_schnorr_finish64_host", None, c_cp, c_p, c_cp, c_cp, c_p)
    sig("bsc_share_sum_dc", c_i, c_p, c_p, c_i, c_p, c_i, c_i, c_i, c_i, c_i, c_d, c_p, c_p)
    sig("bsc_reconstruct_step_gated", c_i, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_d, c_p, c_p, c_p, c_p, c_p, c_i, c_p)
    sig("bsc_reconstruct_mat", c_i, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_d, c_p, c_p, c_p, c_p, c_p, c_i, c_p)
    sig("bsc_recover_mat_host", c_i, c_p, c_p, c_i, c_i, c_p)
    sig("bsc_recover_mat_sizeof", c_i)
    sig("bsc_plain_sum_step_dc", c_i, c_p, c_p, c_i, c_p, c_i, c_i, c_p, c_p, c_p, c_p, c_i, c_p)
    sig("bsc_msm_table_dc", c_i, c_p, c_i, c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_i, c_i, c_p, c_ll, c_p)
    sig("bsc_g1_sum_rows_dc", c_i, c_p, c_p, c_i, c_p, c_ll, c_p, c_p)
    sig("bsc_g1_sum_rows_jac_dc", c_i, c_p, c_p, c_i, c_p, c_ll, c_p, c_p)
    sig("bsc_seeded_perm", None, c_ull, c_i, c_p)
    sig("bsc_sum_rows_i64_dc", c_i, c_p, c_i, c_p, c_i, c_p, c_i, c_p, c_p)
    sig("bsc_plan_block", c_i, c_cp, c_i, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_ull, c_ull, c_ull, c_d, c_i, c_i,
        c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p)
    sig("bsc_chain_append_packed", c_i, c_p, c_ll, c_p, c_i, c_p, c_i, c_p, c_ll, c_i, c_p, c_i, c_i, c_ll, c_p)
    sig("bsc_symm_error", c_i)
    sig("bsc_fe_inv_both", None, c_i, c_cp, c_p, c_p)
    # native block executor (csrc/exec.cu)
    sig("bsc_exec", c_i, c_p, c_i, c_p, c_i)
    sig("bsc_cmd_sizeof", c_i)
    sig("bsc_event_create", c_p)
    sig("bsc_event_destroy", None, c_p)
    sig("bsc_event_record", c_i, c_p, c_p)
    sig("bsc_event_synchronize", c_i, c_p)
    sig("bsc_event_query", c_i, c_p)
    sig("bsc_stream_wait_event", c_i, c_p, c_p)
