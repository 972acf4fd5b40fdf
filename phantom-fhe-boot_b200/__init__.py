"""phantom-fhe-boot_b200 — B200-native RNS polynomial hot path for CKKS (NTT, hybrid key switching, HMult,
rotation, rescale), exposed through the C ABI in include/b200fhe.h (libb200fhe.so, hand-written sm_100a CUDA).

This Python module is only plumbing for tests and bench.py: ctypes over the C ABI, torch for device memory and
streams.  It never imports anything from oracle/ and has no CPU fallback: compute calls raise if the CUDA
library or a GPU is missing.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.environ.get("B200FHE_LIB") or os.path.join(_HERE, "libb200fhe.so")   # override: A/B builds of the same ABI
_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_lib = None


class B200FheError(RuntimeError):
    pass


def build(verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"] + ([] if verbose else ["-s"]))
    return _SO


_SIG = {
    "b200fhe_coeff_modulus_create": [C.c_size_t, C.POINTER(C.c_int), C.c_size_t, _u64p],
    "b200fhe_ctx_create": [C.c_int, _u64p, C.c_size_t, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)],
    "b200fhe_ctx_create_host": [C.c_int, _u64p, C.c_size_t, C.c_size_t, C.POINTER(C.c_void_p)],
    "b200fhe_ctx_destroy": [C.c_void_p],
    "b200fhe_ctx_export_table": [C.c_void_p, C.c_int, _u64p],
    "b200fhe_level_export": [C.c_void_p, C.c_size_t, C.c_int, _u64p, C.c_size_t, C.POINTER(C.c_size_t)],
    "b200fhe_level_beta": [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)],
    "b200fhe_ntt_forward": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_inverse": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_inverse_scale": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_ntt_forward_fuse_moddown": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_bconv_P_to_Ql": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_modup": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_inner_prod": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_moddown_from_ntt": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_keyswitch": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_tensor": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_hmult_relin": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_rescale": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p],
    "b200fhe_rescale_noalloc": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p],
    "b200fhe_automap_table": [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p],
    "b200fhe_galois_table_host": [C.c_int, C.c_uint32, _u32p],
    "b200fhe_apply_galois_ntt": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_rotate_fused": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_eltwise": [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_forward_host": [C.c_void_p, _u64p, C.c_size_t, C.c_size_t, C.c_size_t],
    "b200fhe_ntt_inverse_host": [C.c_void_p, _u64p, C.c_size_t, C.c_size_t, C.c_size_t],
    "b200fhe_ntt_host_batch": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int],
    "b200fhe_ctx_set_lanes": [C.c_void_p, C.c_int],
    "b200fhe_tune_cluster_ntt": [C.c_int],
    "b200fhe_hmult_relin_host": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_rotate_fused_host": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p],
    "b200fhe_boot_eval_host": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p],
}
_vp, _vpp, _sz, _dbl = C.c_void_p, C.POINTER(C.c_void_p), C.c_size_t, C.c_double
_SIG.update({
    "b200fhe_eval_create": [_vp, _vp, _vpp],
    "b200fhe_eval_destroy": [_vp],
    "b200fhe_eval_scaling_factors": [_vp, C.c_int, C.POINTER(_dbl), _sz, C.POINTER(_sz)],
    "b200fhe_eval_set_relin_key": [_vp, _vpp, _sz],
    "b200fhe_eval_set_galois_key": [_vp, C.c_uint32, _vpp, _sz],
    "b200fhe_ct_import": [_vp, _vp, _sz, C.c_uint32, C.c_int, _dbl, C.c_uint32, _vpp],
    "b200fhe_ct_export": [_vp, _vp, _vp, _sz],
    "b200fhe_ct_describe": [_vp, _vp],
    "b200fhe_ct_clone": [_vp, _vp, _vpp],
    "b200fhe_ct_free": [_vp],
    "b200fhe_eval_mod_reduce": [_vp, _vp, _sz],
    "b200fhe_eval_mod_switch_level": [_vp, _vp, _sz],
    "b200fhe_eval_mult_const": [_vp, _vp, _dbl],
    "b200fhe_eval_add_const": [_vp, _vp, _dbl],
    "b200fhe_eval_mult_by_integer": [_vp, _vp, C.c_uint64],
    "b200fhe_eval_mult_by_monomial": [_vp, _vp, C.c_uint32],
    "b200fhe_eval_add_auto": [_vp, _vp, _vp],
    "b200fhe_eval_sub_auto": [_vp, _vp, _vp],
    "b200fhe_eval_mult_auto": [_vp, _vp, _vp, _vpp],
    "b200fhe_eval_square": [_vp, _vp, _vpp],
    "b200fhe_eval_raise_mod": [_vp, _vp, _vpp],
    "b200fhe_eval_rotate_fused": [_vp, _vp, C.c_int32, _vpp],
    "b200fhe_eval_conj_fused": [_vp, _vp, _vpp],
    "b200fhe_eval_chebyshev": [_vp, _vp, C.POINTER(_dbl), _sz, _dbl, _dbl, _vpp],
    "b200fhe_eval_fast_rotations_ext": [_vp, _vp, C.POINTER(C.c_int32), _sz, C.c_int, _vpp],
    "b200fhe_eval_keyswitch_ext": [_vp, _vp, _vpp],
    "b200fhe_eval_keyswitch_down": [_vp, _vp, C.c_int, _vpp],
    "b200fhe_eval_mult_ext": [_vp, _vp, _vp, _dbl],
    "b200fhe_eval_add_ext": [_vp, _vp, _vp],
    "b200fhe_boot_create": [_vp, _vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _vpp],
    "b200fhe_boot_destroy": [_vp],
    "b200fhe_boot_params": [_vp, C.c_int, C.POINTER(C.c_int32)],
    "b200fhe_boot_rotation_indices": [_vp, C.POINTER(C.c_int32), _sz, C.POINTER(_sz)],
    "b200fhe_boot_stage_info": [_vp, C.c_int, _sz, C.POINTER(_sz), C.POINTER(_sz), C.POINTER(_sz)],
    "b200fhe_boot_set_diagonal": [_vp, C.c_int, _sz, _sz, _vp, _sz, _dbl],
    "b200fhe_boot_eval": [_vp, _vp, _vpp],
    "b200fhe_boot_set_peer_shard": [_vp, C.c_uint32, C.c_uint32, _vpp, _sz, _vp, C.c_int],
    "b200fhe_dnn_set_peer_ranks": [_vp, C.c_uint32, C.c_uint32, _vpp, _sz, _vp],
    "b200fhe_boot_graph_capture": [_vp, _vp, _vpp],
    "b200fhe_boot_graph_launch": [_vp],
    "b200fhe_boot_graph_io": [_vp, _vpp, _vpp],
    "b200fhe_boot_graph_destroy": [_vp],
    "b200fhe_boot_eval_iter": [_vp, _vp, C.c_uint32, C.c_uint32, _vpp],
    "b200fhe_boot_coeffs_to_slots": [_vp, _vp, _vpp],
    "b200fhe_boot_slots_to_coeffs": [_vp, _vp, _vpp],
    "b200fhe_boot_precompute": [_vp, _vp],
    "b200fhe_boot_get_diagonal": [_vp, C.c_int, _sz, _sz, _vpp, C.POINTER(_sz), C.POINTER(_dbl)],
    "b200fhe_encoder_create": [_vp, _vpp],
    "b200fhe_encoder_destroy": [_vp],
    "b200fhe_encode": [_vp, C.POINTER(_dbl), _sz, _sz, _dbl, C.c_int, C.c_int, _vp, _vp],
    "b200fhe_decode": [_vp, _vp, _sz, C.c_int, _dbl, C.POINTER(_dbl), _vp],
    "b200fhe_encode_sparse_batch": [_vp, C.POINTER(_dbl), _sz, _sz, _sz, _dbl, _vp, _vp],
    "b200fhe_keygen_create": [_vp, _vp, _vp, _vpp],
    "b200fhe_keygen_create_reference_prng": [_vp, C.c_uint64, _vp, _vpp],
    "b200fhe_keygen_set_secret": [_vp, _vp],
    "b200fhe_keygen_get_secret": [_vp, _vp],
    "b200fhe_keygen_relin_key": [_vp, _vpp],
    "b200fhe_keygen_galois_key": [_vp, C.c_uint32, _vpp],
    "b200fhe_keygen_public_key": [_vp, _vp],
    "b200fhe_encrypt_symmetric": [_vp, _vp, _sz, _vp],
    "b200fhe_encrypt_asymmetric": [_vp, _vp, _vp, _sz, _vp],
    "b200fhe_decrypt": [_vp, _vp, _sz, C.c_uint32, _vp],
    "b200fhe_keygen_sample": [_vp, C.c_int, _vp, _sz, _sz],
    "b200fhe_chacha20_block": [_vp, _u32p, C.c_uint32, _u32p, _u32p],
    "b200fhe_ct_save": [_vp, _vp, _vp, _sz, C.POINTER(_sz)],
    "b200fhe_ct_load": [_vp, _sz, _vp, _vp, _sz, C.POINTER(_sz)],
    "b200fhe_plain_save": [C.c_uint64, C.c_uint64, C.c_uint64, _dbl, _vp, _vp, _sz, C.POINTER(_sz)],
    "b200fhe_plain_load": [_vp, _sz, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(_dbl),
                           _vp, _sz, C.POINTER(_sz)],
    "b200fhe_evk_save": [_vp, _vpp, _sz, _vp, _sz, C.POINTER(_sz)],
    "b200fhe_evk_load": [_vp, _vp, _sz, _vpp, _sz, C.POINTER(_sz), C.POINTER(_sz)],
    "b200fhe_secret_save": [_vp, _vp, _vp, _sz, C.POINTER(_sz)],
    "b200fhe_secret_load": [_vp, _vp, _sz, _vp, C.POINTER(_sz)],
})
class TensorDesc(C.Structure):
    _fields_ = [("width", C.c_size_t), ("slotstr", C.c_size_t), ("num_ch", C.c_size_t), ("num_cts", C.c_size_t)]


_i32p, _dp, _tdp = C.POINTER(C.c_int32), C.POINTER(C.c_double), C.POINTER(TensorDesc)
_SIG.update({
    "b200fhe_shard_ks_begin": [_vp, _sz, _sz, _sz, _vp, _vp, _vp],
    "b200fhe_shard_ks_middle": [_vp, _sz, _sz, _sz, _sz, _sz, _vp, _vp, _vpp, _vp, _vp],
    "b200fhe_shard_ks_end": [_vp, _sz, _sz, _sz, _vp, _vp, C.c_uint32, _vp, _vp],
    "b200fhe_shard_ks_begin_mirrored": [_vp, _sz, _sz, _sz, _vp, _vp, _vp, C.c_uint32, _vp],
    "b200fhe_shard_ks_middle_mirrored": [_vp, _sz, _sz, _sz, _sz, _sz, _vp, _vp, _vp, _vp, _vp, C.c_uint32, _vp],
    "b200fhe_peer_exchange_mc": [_vp, _vp, _vp, _vp, _vp, C.c_uint32, C.c_uint32, _vp, _vp, C.c_uint32, C.c_uint64, _vp],
    "b200fhe_peer_exchange": [_vp, _vp, _vp, _vp, C.c_uint32, C.c_uint32, _vp, _vp, C.c_uint32, C.c_uint64, _vp],
    "b200fhe_ct_pack": [_vp, _vpp, _sz, _vpp],
    "b200fhe_ct_unpack": [_vp, _vp, _vpp, _sz, C.POINTER(_sz)],
    "b200fhe_eval_set_seal_galois_key": [_vp, C.c_uint32, _vpp, _sz],
    "b200fhe_keygen_galois_key_seal": [_vp, C.c_uint32, _vpp],
    "b200fhe_eval_negate": [_vp, _vp],
    "b200fhe_eval_add_plain": [_vp, _vp, _vp, _dbl, C.c_int],
    "b200fhe_eval_multiply_plain": [_vp, _vp, _vp, _dbl],
    "b200fhe_eval_add_many": [_vp, _vpp, _sz, _vpp],
    "b200fhe_eval_multiply": [_vp, _vp, _vp, _vpp],
    "b200fhe_eval_relinearize": [_vp, _vp, _vpp],
    "b200fhe_eval_rescale_to_next": [_vp, _vp, _vpp],
    "b200fhe_eval_mod_switch_to_next": [_vp, _vp, _vpp],
    "b200fhe_eval_apply_galois": [_vp, _vp, C.c_uint32, _vpp],
    "b200fhe_eval_rotate": [_vp, _vp, C.c_int, _vpp],
    "b200fhe_eval_hoisting": [_vp, _vp, C.POINTER(C.c_int), _sz, _vpp],
    "b200fhe_dnn_create": [_vp, _vp, _dbl, _vpp],
    "b200fhe_dnn_set_relin_key": [_vp, _vpp, _sz],
    "b200fhe_dnn_set_lanes": [_vp, _vp, _sz],
    "b200fhe_dnn_set_batch": [_vp, _sz],
    "b200fhe_dnn_set_galois_key": [_vp, C.c_uint32, _vpp, _sz],
    "b200fhe_dnn_add_rotation_indices": [_i32p, _sz, C.POINTER(_sz), _sz, _sz, _sz],
    "b200fhe_dnn_add_avgpool_rotations": [_i32p, _sz, C.POINTER(_sz), _sz, _sz],
    "b200fhe_dnn_enc_tensor": [_vp, _vp, _vp, _dp, _sz, _sz, _vpp, _sz, _tdp],
    "b200fhe_dnn_dec_tensor": [_vp, _vp, _vpp, _tdp, _dp],
    "b200fhe_dnn_conv": [_vp, _vpp, _tdp, _dp, _sz, _sz, _sz, _vpp, _sz, _tdp],
    "b200fhe_dnn_batch_norm": [_vp, _vpp, _tdp, _dp, _dp, _dp, _dp, _vpp, _sz, _tdp],
    "b200fhe_dnn_add": [_vp, _vpp, _tdp, _vpp, _tdp, _vpp, _sz, _tdp],
    "b200fhe_dnn_relu": [_vp, _vpp, _tdp, C.c_int, C.c_int, C.c_int, _vpp, _sz, _tdp],
    "b200fhe_dnn_relu_deg_two": [_vp, _vpp, _tdp, _vpp, _sz, _tdp],
    "b200fhe_dnn_sign": [_vp, _vpp, _tdp, _sz, _vpp, _sz, _tdp],
    "b200fhe_dnn_relu_composite": [_vp, _vp, _vpp, _tdp, _vpp, _sz, _tdp],
    "b200fhe_dnn_bootstrap": [_vp, _vp, _vpp, _tdp, _vpp, _sz, _tdp],
    "b200fhe_dnn_avgpool_fullcon": [_vp, _vpp, _tdp, _dp, _sz, _sz, _dp, _vpp, _sz, _tdp],
    "b200fhe_dnn_softmax": [_vp, _vp, _vpp, _tdp, _sz, _sz, _sz, _sz, _vpp, _sz, _tdp],
    "b200fhe_ctx_create_from_tables": [C.c_int, _u64p, _sz, _sz, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vpp],
    "b200fhe_ctx_set_special": [_vp, _sz],
    "b200fhe_ctx_info": [_vp, C.POINTER(C.c_int), C.POINTER(_sz), C.POINTER(_sz), C.POINTER(C.c_int)],
    "b200fhe_ntt_forward_ex": [_vp, _vp, _vp, _vp, _vp],
    "b200fhe_ntt_inverse_ex": [_vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "b200fhe_inner_prod_devkeys": [_vp, _sz, _vp, _vp, _vp, _vp],
    "b200fhe_keyswitch_devkeys": [_vp, _sz, _vp, _vp, _vp, _vp],
    "b200fhe_resnet20_rotation_indices": [_sz, _i32p, _sz, C.POINTER(_sz)],
    "b200fhe_resnet20_infer": [_vp, _vp, _vpp, _tdp, _dp, _sz, _sz, _sz, _sz, _vpp, _sz, _tdp, _dp],
})
_OTHER = {  # non-int return types
    "b200fhe_last_error": C.c_char_p, "b200fhe_version": C.c_char_p, "b200fhe_launch_count": C.c_uint64,
    "b200fhe_auto_index": C.c_uint32, "b200fhe_boot_evaluator": C.c_void_p, "b200fhe_eval_pool_bytes": C.c_size_t,
    "b200fhe_boot_peer_arena_bytes": C.c_size_t,
    "b200fhe_encoder_slots": C.c_size_t, "b200fhe_keygen_secret": C.c_void_p, "b200fhe_evk_digits": C.c_size_t,
    "b200fhe_keygen_destroy": None, "b200fhe_evk_free": None, "b200fhe_dnn_destroy": None,
    "b200fhe_resnet20_param_count": C.c_size_t, "b200fhe_trace_reset": None, "b200fhe_trace_read": None,
}
EXPORTS = sorted(list(_SIG) + list(_OTHER))


def lib():
    """Load libb200fhe.so; fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise B200FheError("libb200fhe.so is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(_SO)
        for name, args in _SIG.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        for name, rt in _OTHER.items():
            getattr(L, name).restype = rt
        L.b200fhe_auto_index.argtypes = [C.c_int, C.c_int32]
        L.b200fhe_boot_evaluator.argtypes = [C.c_void_p]
        L.b200fhe_eval_pool_bytes.argtypes = [C.c_void_p]
        L.b200fhe_boot_peer_arena_bytes.argtypes = [C.c_void_p]
        L.b200fhe_keygen_secret.argtypes = [C.c_void_p]
        L.b200fhe_keygen_destroy.argtypes = [C.c_void_p]
        L.b200fhe_evk_digits.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_size_t]
        L.b200fhe_evk_free.argtypes = [C.c_void_p]
        L.b200fhe_dnn_destroy.argtypes = [C.c_void_p]
        L.b200fhe_resnet20_param_count.argtypes = [C.c_size_t] * 4
        L.b200fhe_trace_reset.argtypes = []
        L.b200fhe_trace_read.argtypes = [C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise B200FheError("b200fhe error %d: %s" % (rc, lib().b200fhe_last_error().decode()))


def trace_reset():
    lib().b200fhe_trace_reset()


def trace_read():
    """-> dict(hbm_bytes, butterflies, macs, calls) accumulated by the kernel launchers since trace_reset()"""
    out = (C.c_uint64 * 4)()
    lib().b200fhe_trace_read(out)
    return {"hbm_bytes": int(out[0]), "butterflies": int(out[1]), "macs": int(out[2]), "calls": int(out[3])}


def launch_count():
    return int(lib().b200fhe_launch_count())


def coeff_modulus_create(N, bits):
    out = np.zeros(len(bits), dtype=np.uint64)
    _check(lib().b200fhe_coeff_modulus_create(N, (C.c_int * len(bits))(*bits), len(bits), out.ctypes.data_as(_u64p)))
    return out


def galois_table(logN, elt):
    t = np.zeros(1 << logN, dtype=np.uint32)
    _check(lib().b200fhe_galois_table_host(logN, elt, t.ctypes.data_as(_u32p)))
    return t


def to_dev(a, device="cuda:0"):
    """numpy uint64 array -> torch int64 tensor on the GPU (same bits)."""
    import torch
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return torch.from_numpy(a.view(np.int64)).to(device)


def to_host(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint64)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    import torch
    if not torch.cuda.is_available():  # host-only contexts (setup logic, tests without a GPU): no stream to name
        return C.c_void_p(0)
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Context:
    """PhantomContext-like owner of the NTT tables and per-level RNS constants (reference src/context.cu:121)."""

    def __init__(self, logN, primes_QP, size_P, device=0, host_only=False):
        self.logN, self.N = logN, 1 << logN
        self.primes = np.ascontiguousarray(primes_QP, dtype=np.uint64)
        self.size_QP, self.size_P = len(self.primes), size_P
        self.size_Q = self.size_QP - size_P
        self.host_only = host_only
        h = C.c_void_p()
        if host_only:
            _check(lib().b200fhe_ctx_create_host(logN, self.primes.ctypes.data_as(_u64p), self.size_QP, size_P, C.byref(h)))
        else:
            _check(lib().b200fhe_ctx_create(logN, self.primes.ctypes.data_as(_u64p), self.size_QP, size_P, device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- tables / constants (host) ----
    def table(self, which):
        n = self.size_QP * (self.N if which < 4 else 1)
        out = np.zeros(n, dtype=np.uint64)
        _check(lib().b200fhe_ctx_export_table(self.h, which, out.ctypes.data_as(_u64p)))
        return out.reshape(self.size_QP, -1)

    def beta(self, size_Ql):
        b = C.c_size_t()
        _check(lib().b200fhe_level_beta(self.h, size_Ql, C.byref(b)))
        return b.value

    def level_const(self, size_Ql, which):
        n = C.c_size_t()
        _check(lib().b200fhe_level_export(self.h, size_Ql, which, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.uint64)
        _check(lib().b200fhe_level_export(self.h, size_Ql, which, out.ctypes.data_as(_u64p), n.value, C.byref(n)))
        return out[:n.value]

    # ---- device ops (torch int64 tensors carrying uint64 residues) ----
    def ntt_forward(self, x, limbs, start=0, size_Ql=0, npoly=1, poly_stride=0):
        _check(lib().b200fhe_ntt_forward(self.h, _ptr(x), limbs, start, size_Ql, npoly, poly_stride, _stream()))

    def ntt_inverse(self, x, limbs, start=0, size_Ql=0, npoly=1, poly_stride=0):
        _check(lib().b200fhe_ntt_inverse(self.h, _ptr(x), limbs, start, size_Ql, npoly, poly_stride, _stream()))

    def ntt_inverse_scale(self, out, x, limbs, start=0, size_Ql=0, scale=None, scale_shoup=None):
        _check(lib().b200fhe_ntt_inverse_scale(self.h, _ptr(out), _ptr(x), limbs, start, size_Ql, _ptr(scale),
                                               _ptr(scale_shoup), _stream()))

    def ntt_forward_fuse_moddown(self, ct, cx, pinv, pinv_shoup, delta, limbs, start=0):
        _check(lib().b200fhe_ntt_forward_fuse_moddown(self.h, _ptr(ct), _ptr(cx), _ptr(pinv), _ptr(pinv_shoup),
                                                      _ptr(delta), limbs, start, _stream()))

    @staticmethod
    def _evk(digits):
        return (C.c_void_p * len(digits))(*[d if isinstance(d, int) else d.data_ptr() for d in digits])

    def bconv_P_to_Ql(self, size_Ql, dst, src):
        _check(lib().b200fhe_bconv_P_to_Ql(self.h, size_Ql, _ptr(dst), _ptr(src), _stream()))

    def modup(self, size_Ql, dst, c2):
        _check(lib().b200fhe_modup(self.h, size_Ql, _ptr(dst), _ptr(c2), _stream()))

    def inner_prod(self, size_Ql, cx, up, evk):
        _check(lib().b200fhe_inner_prod(self.h, size_Ql, _ptr(cx), _ptr(up), self._evk(evk), _stream()))

    def moddown_from_ntt(self, size_Ql, ct, cx):
        _check(lib().b200fhe_moddown_from_ntt(self.h, size_Ql, _ptr(ct), _ptr(cx), _stream()))

    def keyswitch(self, size_Ql, ct, c2, evk):
        _check(lib().b200fhe_keyswitch(self.h, size_Ql, _ptr(ct), _ptr(c2), self._evk(evk), _stream()))

    def tensor(self, L, out, a, b):
        _check(lib().b200fhe_tensor(self.h, L, _ptr(out), _ptr(a), _ptr(b), _stream()))

    def hmult_relin(self, size_Ql, out, a, b, rlk):
        _check(lib().b200fhe_hmult_relin(self.h, size_Ql, _ptr(out), _ptr(a), _ptr(b), self._evk(rlk), _stream()))

    def set_lanes(self, on):
        _check(lib().b200fhe_ctx_set_lanes(self.h, int(on)))

    def rescale(self, size_Ql, dst, src, polys=2):
        _check(lib().b200fhe_rescale(self.h, size_Ql, _ptr(dst), _ptr(src), polys, _stream()))

    def rescale_noalloc(self, size_Ql, dst, src, polys=2):
        """b200fhe_rescale on a static device scratch array: allocates nothing (include/b200fhe.h)."""
        _check(lib().b200fhe_rescale_noalloc(self.h, size_Ql, _ptr(dst), _ptr(src), polys, _stream()))

    def automap_table(self, k, perm):
        _check(lib().b200fhe_automap_table(self.h, k, _ptr(perm), _stream()))

    def apply_galois_ntt(self, dst, src, perm, limbs, npoly=1, poly_stride=0):
        _check(lib().b200fhe_apply_galois_ntt(self.h, _ptr(dst), _ptr(src), _ptr(perm), limbs, npoly, poly_stride, _stream()))

    def rotate_fused(self, size_Ql, out, ct, k, glk):
        _check(lib().b200fhe_rotate_fused(self.h, size_Ql, _ptr(out), _ptr(ct), k, self._evk(glk), _stream()))

    def eltwise(self, op, out, a, b, limbs, size_Ql=0):
        _check(lib().b200fhe_eltwise(self.h, op, _ptr(out), _ptr(a), _ptr(b), limbs, size_Ql, _stream()))

    # ---- entries for the C++ shim over the reference's types (include/b200fhe.h, shim/) ----
    @classmethod
    def from_tables(cls, logN, primes_QP, size_P, tw, tw_shoup, itw, itw_shoup, ninv, ninv_shoup):
        """a context on caller-owned DEVICE tables in the reference's DNTTTable layout (torch tensors)"""
        self = cls.__new__(cls)
        self.logN, self.N = logN, 1 << logN
        self.primes = np.ascontiguousarray(primes_QP, dtype=np.uint64)
        self.size_QP, self.size_P = len(self.primes), size_P
        self.size_Q = self.size_QP - size_P
        self.host_only = False
        h = C.c_void_p()
        _check(lib().b200fhe_ctx_create_from_tables(logN, self.primes.ctypes.data_as(_u64p), self.size_QP, size_P, _ptr(tw),
                                                    _ptr(tw_shoup), _ptr(itw), _ptr(itw_shoup), _ptr(ninv), _ptr(ninv_shoup),
                                                    _stream(), C.byref(h)))
        self.h = h
        return self

    def set_special(self, size_P):
        _check(lib().b200fhe_ctx_set_special(self.h, size_P))
        self.size_P, self.size_Q = size_P, self.size_QP - size_P

    @staticmethod
    def _range(start, limbs, special_from=0, special_base=0, excl=(0, 0)):
        return (C.c_size_t * 6)(start, limbs, special_from, special_base, excl[0], excl[1])

    def ntt_forward_ex(self, out, x, start, limbs, special_from=0, special_base=0, excl=(0, 0)):
        r = self._range(start, limbs, special_from, special_base, excl)
        _check(lib().b200fhe_ntt_forward_ex(self.h, _ptr(out), _ptr(x), C.cast(r, C.c_void_p), _stream()))

    def ntt_inverse_ex(self, out, x, start, limbs, special_from=0, special_base=0, scale=None, scale_shoup=None):
        r = self._range(start, limbs, special_from, special_base)
        _check(lib().b200fhe_ntt_inverse_ex(self.h, _ptr(out), _ptr(x), C.cast(r, C.c_void_p), _ptr(scale), _ptr(scale_shoup),
                                            _stream()))

    def inner_prod_devkeys(self, size_Ql, cx, up, table):
        _check(lib().b200fhe_inner_prod_devkeys(self.h, size_Ql, _ptr(cx), _ptr(up), _ptr(table), _stream()))

    def keyswitch_devkeys(self, size_Ql, ct, c2, table):
        _check(lib().b200fhe_keyswitch_devkeys(self.h, size_Ql, _ptr(ct), _ptr(c2), _ptr(table), _stream()))

    # ---- host-buffer entry points (H2D + compute + D2H inside) ----
    def ntt_forward_host(self, a, limbs, start=0, size_Ql=0):
        _check(lib().b200fhe_ntt_forward_host(self.h, a.ctypes.data_as(_u64p), limbs, start, size_Ql))

    def ntt_inverse_host(self, a, limbs, start=0, size_Ql=0):
        _check(lib().b200fhe_ntt_inverse_host(self.h, a.ctypes.data_as(_u64p), limbs, start, size_Ql))

    def ntt_host_batch(self, host_ptr, limbs, npoly, ops=3, size_Ql=0):
        """host_ptr: address of [npoly][limbs][N] uint64 in (pinned) host memory"""
        _check(lib().b200fhe_ntt_host_batch(self.h, C.c_void_p(host_ptr), limbs, size_Ql, npoly, ops))


# ------------------------------------------------------------------------------------------ evaluator / bootstrap
class CtDesc(C.Structure):
    _fields_ = [("size_Ql", C.c_size_t), ("limbs", C.c_size_t), ("polys", C.c_uint32), ("noise_scale_deg", C.c_uint32),
                ("scale", C.c_double)]


def auto_index(logN, rotation):
    return int(lib().b200fhe_auto_index(logN, rotation))


class Ciphertext:
    """Opaque ciphertext handle of an Evaluator (reference PhantomCiphertext: limbs + chain index + scale + NoiseScaleDeg)."""

    def __init__(self, ev, h):
        self.ev, self.h = ev, C.c_void_p(h) if not isinstance(h, C.c_void_p) else h

    def desc(self):
        d = CtDesc()
        _check(lib().b200fhe_ct_describe(self.h, C.byref(d)))
        return d

    def tensor(self):
        """copy of the limbs as a torch int64 tensor [polys][limbs][N]"""
        import torch
        d = self.desc()
        n = d.polys * d.limbs * self.ev.ctx.N
        t = torch.empty(n, dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_ct_export(self.ev.h, self.h, _ptr(t), n))
        return t.view(d.polys, d.limbs, self.ev.ctx.N)

    def clone(self):
        h = C.c_void_p()
        _check(lib().b200fhe_ct_clone(self.ev.h, self.h, C.byref(h)))
        return Ciphertext(self.ev, h)

    def free(self):
        if self.h:
            lib().b200fhe_ct_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Evaluator:
    """FLEXIBLEAUTO evaluator (reference include/evaluate.cuh:262-470) on the current torch stream."""

    def __init__(self, ctx, handle=None, owner=None):
        self.ctx, self._owner = ctx, owner
        if handle is None:
            h = C.c_void_p()
            _check(lib().b200fhe_eval_create(ctx.h, _stream(), C.byref(h)))
            self.h, self._own = h, True
        else:
            self.h, self._own = C.c_void_p(handle), False
        self._keys = []  # keep key tensors alive

    def close(self):
        if getattr(self, "h", None) and self._own:
            lib().b200fhe_eval_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def scaling_factors(self, which=0):
        n = C.c_size_t()
        _check(lib().b200fhe_eval_scaling_factors(self.h, which, None, 0, C.byref(n)))
        out = (C.c_double * max(n.value, 1))()
        _check(lib().b200fhe_eval_scaling_factors(self.h, which, out, n.value, C.byref(n)))
        return list(out)[:n.value]

    @staticmethod
    def _digits(digits):
        ptrs = [d if isinstance(d, int) else d.data_ptr() for d in digits]
        return (C.c_void_p * len(ptrs))(*ptrs), len(ptrs)

    def set_relin_key(self, digits):
        a, n = self._digits(digits)
        self._keys.append(digits)
        _check(lib().b200fhe_eval_set_relin_key(self.h, a, n))

    def set_galois_key(self, k, digits):
        a, n = self._digits(digits)
        self._keys.append(digits)
        _check(lib().b200fhe_eval_set_galois_key(self.h, k, a, n))

    def import_ct(self, t, size_Ql, polys=2, extended=False, scale=1.0, noise=1):
        """t: torch int64 tensor or raw device pointer"""
        h = C.c_void_p()
        p = C.c_void_p(t) if isinstance(t, int) else _ptr(t)
        _check(lib().b200fhe_ct_import(self.h, p, size_Ql, polys, 1 if extended else 0, scale, noise, C.byref(h)))
        return Ciphertext(self, h)

    def _out(self, f, *args):
        h = C.c_void_p()
        _check(f(self.h, *args, C.byref(h)))
        return Ciphertext(self, h)

    def mod_reduce(self, ct, levels=1):
        _check(lib().b200fhe_eval_mod_reduce(self.h, ct.h, levels))

    def mod_switch_level(self, ct, levels):
        _check(lib().b200fhe_eval_mod_switch_level(self.h, ct.h, levels))

    def mult_const(self, ct, v):
        _check(lib().b200fhe_eval_mult_const(self.h, ct.h, v))

    def add_const(self, ct, v):
        _check(lib().b200fhe_eval_add_const(self.h, ct.h, v))

    def mult_by_integer(self, ct, v):
        _check(lib().b200fhe_eval_mult_by_integer(self.h, ct.h, v))

    def mult_by_monomial(self, ct, power):
        _check(lib().b200fhe_eval_mult_by_monomial(self.h, ct.h, power))

    def add_auto(self, a, b):
        _check(lib().b200fhe_eval_add_auto(self.h, a.h, b.h))

    def sub_auto(self, a, b):
        _check(lib().b200fhe_eval_sub_auto(self.h, a.h, b.h))

    def mult_auto(self, a, b):
        return self._out(lib().b200fhe_eval_mult_auto, a.h, b.h)

    def square(self, a):
        return self._out(lib().b200fhe_eval_square, a.h)

    def raise_mod(self, ct):
        return self._out(lib().b200fhe_eval_raise_mod, ct.h)

    def rotate_fused(self, ct, index):
        return self._out(lib().b200fhe_eval_rotate_fused, ct.h, index)

    def conj_fused(self, ct):
        return self._out(lib().b200fhe_eval_conj_fused, ct.h)

    def chebyshev(self, ct, coeffs, a=-1.0, b=1.0):
        arr = (C.c_double * len(coeffs))(*coeffs)
        return self._out(lib().b200fhe_eval_chebyshev, ct.h, arr, len(coeffs), a, b)

    def fast_rotations_ext(self, ct, indices, add_first=True):
        idx = (C.c_int32 * len(indices))(*indices)
        outs = (C.c_void_p * len(indices))()
        _check(lib().b200fhe_eval_fast_rotations_ext(self.h, ct.h, idx, len(indices), 1 if add_first else 0, outs))
        return [Ciphertext(self, o) for o in outs]

    def keyswitch_ext(self, ct):
        return self._out(lib().b200fhe_eval_keyswitch_ext, ct.h)

    def keyswitch_down(self, ct, first_only=False):
        return self._out(lib().b200fhe_eval_keyswitch_down, ct.h, 1 if first_only else 0)

    def mult_ext(self, ct, plain, plain_scale):
        p = C.c_void_p(plain) if isinstance(plain, int) else _ptr(plain)
        _check(lib().b200fhe_eval_mult_ext(self.h, ct.h, p, plain_scale))

    def add_ext(self, a, b):
        _check(lib().b200fhe_eval_add_ext(self.h, a.h, b.h))

    def pack(self, cts):
        """independent ciphertexts with identical bookkeeping -> one batched handle (one launch set per operation)"""
        arr = (C.c_void_p * len(cts))(*[c.h for c in cts])
        return self._out(lib().b200fhe_ct_pack, arr, len(cts))

    def unpack(self, ct):
        n = C.c_size_t()
        _check(lib().b200fhe_ct_unpack(self.h, ct.h, None, 0, C.byref(n)))
        outs = (C.c_void_p * n.value)()
        _check(lib().b200fhe_ct_unpack(self.h, ct.h, outs, n.value, C.byref(n)))
        return [Ciphertext(self, outs[i]) for i in range(n.value)]

    # ---- SEAL-style surface (reference src/evaluate.cu:95-2149, examples/3_ckks.cu)
    def set_seal_galois_key(self, elt, digits):
        a, n = self._digits(digits)
        self._keys.append(digits)
        _check(lib().b200fhe_eval_set_seal_galois_key(self.h, elt, a, n))

    def negate(self, ct):
        _check(lib().b200fhe_eval_negate(self.h, ct.h))

    def add_plain(self, ct, plain, plain_scale, sub=False):
        _check(lib().b200fhe_eval_add_plain(self.h, ct.h, _ptr(plain), float(plain_scale), 1 if sub else 0))

    def multiply_plain(self, ct, plain, plain_scale):
        _check(lib().b200fhe_eval_multiply_plain(self.h, ct.h, _ptr(plain), float(plain_scale)))

    def add_many(self, cts):
        arr = (C.c_void_p * len(cts))(*[c.h for c in cts])
        return self._out(lib().b200fhe_eval_add_many, arr, len(cts))

    def multiply(self, a, b):
        return self._out(lib().b200fhe_eval_multiply, a.h, b.h)

    def relinearize(self, ct):
        return self._out(lib().b200fhe_eval_relinearize, ct.h)

    def rescale_to_next(self, ct):
        return self._out(lib().b200fhe_eval_rescale_to_next, ct.h)

    def mod_switch_to_next(self, ct):
        return self._out(lib().b200fhe_eval_mod_switch_to_next, ct.h)

    def apply_galois(self, ct, elt):
        return self._out(lib().b200fhe_eval_apply_galois, ct.h, elt)

    def rotate(self, ct, step):
        return self._out(lib().b200fhe_eval_rotate, ct.h, step)

    def hoisting(self, ct, steps):
        arr = (C.c_int * len(steps))(*steps)
        return self._out(lib().b200fhe_eval_hoisting, ct.h, arr, len(steps))

    def pool_bytes(self):
        return int(lib().b200fhe_eval_pool_bytes(self.h))


class Bootstrapper:
    """FHECKKSRNS (reference include/bootstrap.cuh:98-160): setup parameters, diagonals and keys in, EvalBootstrap."""

    def __init__(self, ctx, level_budget=(2, 2), dim1=(0, 0), num_slots=0, correction_factor=0):
        self.ctx = ctx
        h = C.c_void_p()
        _check(lib().b200fhe_boot_create(ctx.h, _stream(), level_budget[0], level_budget[1], dim1[0], dim1[1], num_slots,
                                         correction_factor, C.byref(h)))
        self.h = h
        self.ev = Evaluator(ctx, handle=lib().b200fhe_boot_evaluator(h), owner=self)
        self._diag = []

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_boot_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def params(self, which):
        out = (C.c_int32 * 9)()
        _check(lib().b200fhe_boot_params(self.h, which, out))
        return list(out)

    def rotation_indices(self):
        n = C.c_size_t()
        _check(lib().b200fhe_boot_rotation_indices(self.h, None, 0, C.byref(n)))
        out = (C.c_int32 * max(n.value, 1))()
        _check(lib().b200fhe_boot_rotation_indices(self.h, out, n.value, C.byref(n)))
        return list(out)[:n.value]

    def stage_info(self, which, stage):
        a, b, c = C.c_size_t(), C.c_size_t(), C.c_size_t()
        _check(lib().b200fhe_boot_stage_info(self.h, which, stage, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def set_diagonal(self, which, stage, index, plain, size_Ql, scale):
        p = C.c_void_p(plain) if isinstance(plain, int) else _ptr(plain)
        if not isinstance(plain, int):
            self._diag.append(plain)
        _check(lib().b200fhe_boot_set_diagonal(self.h, which, stage, index, p, size_Ql, scale))

    def precompute(self, encoder):
        """EvalBootstrapSetup(precompute=True): compute + encode all diagonals on this GPU (device memory owned here)"""
        _check(lib().b200fhe_boot_precompute(self.h, encoder.h))

    def get_diagonal(self, which, stage, index):
        """-> (torch int64 view [Ql + P][N] on the installed diagonal, Ql, scale)"""
        import torch
        p, ql, sc = C.c_void_p(), C.c_size_t(), C.c_double()
        _check(lib().b200fhe_boot_get_diagonal(self.h, which, stage, index, C.byref(p), C.byref(ql), C.byref(sc)))
        if not p.value:
            return None, ql.value, sc.value
        n = (ql.value + self.ctx.size_P) * self.ctx.N

        class _E:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (int(p.value), False), "version": 2}
        return torch.as_tensor(_E(), device="cuda").view(ql.value + self.ctx.size_P, self.ctx.N), ql.value, sc.value

    def _out(self, f, ct):
        h = C.c_void_p()
        _check(f(self.h, ct.h, C.byref(h)))
        return Ciphertext(self.ev, h)

    def eval(self, ct, num_iterations=1, precision=0):
        if num_iterations <= 1:
            return self._out(lib().b200fhe_boot_eval, ct)
        h = C.c_void_p()
        _check(lib().b200fhe_boot_eval_iter(self.h, ct.h, num_iterations, precision, C.byref(h)))
        return Ciphertext(self.ev, h)

    def capture(self, ct):
        """EvalBootstrap of `ct`'s buffer as a CUDA graph (b200fhe_boot_graph_capture) -> BootGraph"""
        return BootGraph(self, ct)

    def peer_arena_bytes(self):
        return int(lib().b200fhe_boot_peer_arena_bytes(self.h))

    def set_peer_shard(self, world, rank=0, arena_bases=None, arena_bytes=0, counter=None, virtual_ranks=False):
        """the linear-transform stages limb-sharded over `world` ranks (b200fhe_boot_set_peer_shard); world <= 1: off.
        sharding.PeerShardedBootstrap sets up the arena over torch's symmetric memory."""
        bases = (C.c_void_p * max(world, 1))(*(arena_bases or [None] * max(world, 1)))
        _check(lib().b200fhe_boot_set_peer_shard(self.h, world, rank, bases, arena_bytes,
                                                 C.c_void_p(counter) if counter else None, 1 if virtual_ranks else 0))

    def coeffs_to_slots(self, ct):
        return self._out(lib().b200fhe_boot_coeffs_to_slots, ct)

    def slots_to_coeffs(self, ct):
        return self._out(lib().b200fhe_boot_slots_to_coeffs, ct)


class BootGraph:
    """A captured bootstrap: launch() replays it on the bootstrapper's stream; `output` is a Ciphertext on the graph's result
    buffer (rewritten by every launch), `input_ptr` the device address of the input limbs [2][Ql][N] the graph reads."""

    def __init__(self, boot, ct):
        self.boot, self._ct = boot, ct
        h = C.c_void_p()
        _check(lib().b200fhe_boot_graph_capture(boot.h, ct.h, C.byref(h)))
        self.h = h
        p, o = C.c_void_p(), C.c_void_p()
        _check(lib().b200fhe_boot_graph_io(h, C.byref(p), C.byref(o)))
        self.input_ptr = p.value
        self.output = Ciphertext(boot.ev, o)

    def launch(self):
        _check(lib().b200fhe_boot_graph_launch(self.h))
        return self.output

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_boot_graph_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Encoder:
    """CKKS encoder / decoder (PhantomCKKSEncoder: include/ckks.h:175-345).  Values are complex numpy arrays on the
    host; plaintexts are device tensors [limbs][N] in NTT form."""

    def __init__(self, ctx):
        self.ctx = ctx
        h = C.c_void_p()
        _check(lib().b200fhe_encoder_create(ctx.h, C.byref(h)))
        self.h = h
        self.slots = ctx.N // 2

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_encoder_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode(self, values, size_Ql, scale, ext=False, sparse=False):
        import numpy as np
        import torch
        v = np.ascontiguousarray(np.asarray(values, dtype=np.complex128))
        limbs = size_Ql + (self.ctx.size_P if ext else 0)
        out = torch.empty((limbs, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_encode(self.h, v.view(np.float64).ctypes.data_as(C.POINTER(C.c_double)), v.size, size_Ql,
                                    float(scale), int(ext), int(sparse), _ptr(out), _stream()))
        return out

    def encode_sparse_batch(self, values, size_Ql, scale):
        """values: [batch][n] complex -> [batch][size_Ql][N] (encode_sparse of every row in one pass)"""
        import numpy as np
        import torch
        v = np.ascontiguousarray(np.asarray(values, dtype=np.complex128))
        out = torch.empty((v.shape[0], size_Ql, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_encode_sparse_batch(self.h, v.view(np.float64).ctypes.data_as(C.POINTER(C.c_double)), v.shape[1],
                                                 v.shape[0], size_Ql, float(scale), _ptr(out), _stream()))
        return out

    def decode(self, plain, size_Ql, scale, ext=False):
        import numpy as np
        out = np.empty(self.slots, dtype=np.complex128)
        _check(lib().b200fhe_decode(self.h, _ptr(plain), size_Ql, int(ext), float(scale),
                                    out.view(np.float64).ctypes.data_as(C.POINTER(C.c_double)), _stream()))
        return out


class TensorCT:
    """One ciphertext per channel (reference TensorCT, include/dnn.cuh:33-39)."""

    def __init__(self, cts, width, slotstr, num_ch):
        self.cts, self.width, self.slotstr, self.num_ch = list(cts), width, slotstr, num_ch

    def _args(self):
        arr = (C.c_void_p * len(self.cts))(*[c.h for c in self.cts])
        return arr, TensorDesc(self.width, self.slotstr, self.num_ch, len(self.cts))


def _f64(a):
    import numpy as np
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


class DNN:
    """Encrypted CNN layers (reference phantom::DNN, include/dnn.cuh:41-80) on an Evaluator (the bootstrapper's when
    bootstraps are involved)."""

    def __init__(self, ev, encoder, scale):
        self.ev, self.encoder = ev, encoder
        h = C.c_void_p()
        _check(lib().b200fhe_dnn_create(ev.h, encoder.h, float(scale), C.byref(h)))
        self.h = h
        self._keys = []

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_dnn_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_relin_key(self, digits):
        a, n = Evaluator._digits(digits)
        self._keys.append(digits)
        _check(lib().b200fhe_dnn_set_relin_key(self.h, a, n))

    def set_galois_key(self, k, digits):
        a, n = Evaluator._digits(digits)
        self._keys.append(digits)
        _check(lib().b200fhe_dnn_set_galois_key(self.h, k, a, n))

    def set_lanes(self, boot, n):
        """run the independent channels of ReluComposite / BootStrap on n concurrent streams (same limbs)"""
        _check(lib().b200fhe_dnn_set_lanes(self.h, boot.h, n))
        self._boot = boot  # the lanes read the bootstrapper's diagonals: keep it alive

    def set_peer_ranks(self, world, rank=0, arena_bases=None, arena_bytes=0, counter=None):
        """ONE image on `world` GPUs: the channels of ReluComposite / BootStrap dealt to the ranks, results all-gathered over
        NVLink peer memory (b200fhe_dnn_set_peer_ranks); world <= 1: off.  sharding.PeerShardedDNN sets up the arena."""
        bases = (C.c_void_p * max(world, 1))(*(arena_bases or [None] * max(world, 1)))
        _check(lib().b200fhe_dnn_set_peer_ranks(self.h, world, rank, bases, arena_bytes, C.c_void_p(counter) if counter else None))

    def set_batch(self, n):
        """up to n channels of a layer go through ReluComposite / BootStrap as one batched ciphertext (same limbs)"""
        _check(lib().b200fhe_dnn_set_batch(self.h, n))

    @staticmethod
    def _indices(f, have, *args):
        cap = len(have) + 64
        buf = (C.c_int32 * cap)(*have)
        n = C.c_size_t(len(have))
        _check(f(buf, cap, C.byref(n), *args))
        return list(buf)[:n.value]

    @staticmethod
    def add_rotation_indices(have, width, kernel_h, slotstr):
        return DNN._indices(lib().b200fhe_dnn_add_rotation_indices, have, width, kernel_h, slotstr)

    @staticmethod
    def add_avgpool_rotations(have, width, slotstr):
        return DNN._indices(lib().b200fhe_dnn_add_avgpool_rotations, have, width, slotstr)

    def _call(self, f, pre, t, post, cap):
        arr, d = t._args()
        outs = (C.c_void_p * cap)()
        od = TensorDesc()
        _check(f(self.h, *pre, arr, C.byref(d), *post, outs, cap, C.byref(od)))
        return TensorCT([Ciphertext(self.ev, outs[i]) for i in range(od.num_cts)], od.width, od.slotstr, od.num_ch)

    def enc_tensor(self, kg, pk, image):
        a, p = _f64(image)
        w, ch = a.shape[0], a.shape[2]
        outs = (C.c_void_p * ch)()
        od = TensorDesc()
        _check(lib().b200fhe_dnn_enc_tensor(self.h, kg.h, _ptr(pk), p, w, ch, outs, ch, C.byref(od)))
        return TensorCT([Ciphertext(self.ev, outs[i]) for i in range(od.num_cts)], od.width, od.slotstr, od.num_ch)

    def dec_tensor(self, kg, t):
        import numpy as np
        out = np.zeros((t.width, t.width, t.num_ch))
        arr, d = t._args()
        _check(lib().b200fhe_dnn_dec_tensor(self.h, kg.h, arr, C.byref(d), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def conv(self, t, weight, stride=1):
        w, p = _f64(weight)  # [kh][kh][in][out]
        return self._call(lib().b200fhe_dnn_conv, (), t, (p, w.shape[0], w.shape[3], stride), w.shape[3])

    def batch_norm(self, t, weight, bias, mean, var):
        keep = [_f64(x) for x in (weight, bias, mean, var)]
        return self._call(lib().b200fhe_dnn_batch_norm, (), t, tuple(k[1] for k in keep), len(t.cts))

    def add(self, a, b):
        arr_b, db = b._args()
        return self._call(lib().b200fhe_dnn_add, (), a, (arr_b, C.byref(db)), len(a.cts))

    def relu(self, t, a=-1, b=1, deg=7):
        return self._call(lib().b200fhe_dnn_relu, (), t, (a, b, deg), len(t.cts))

    def relu_deg_two(self, t):
        return self._call(lib().b200fhe_dnn_relu_deg_two, (), t, (), len(t.cts))

    def sign(self, t, k):
        return self._call(lib().b200fhe_dnn_sign, (), t, (k,), len(t.cts))

    def relu_composite(self, t, boot):
        return self._call(lib().b200fhe_dnn_relu_composite, (boot.h,), t, (), len(t.cts))

    def bootstrap(self, t, boot):
        return self._call(lib().b200fhe_dnn_bootstrap, (boot.h,), t, (), len(t.cts))

    def avgpool_fullcon(self, t, weight, bias):
        w, p = _f64(weight)  # [T][t]
        b, pb = _f64(bias)
        return self._call(lib().b200fhe_dnn_avgpool_fullcon, (), t, (p, w.shape[0], w.shape[1], pb), max(w.shape[1], len(t.cts)))

    def softmax(self, t, boot, BoundB=64, BoundR=10000, GumbelLambda=4, GoldSchmidtD=16):
        return self._call(lib().b200fhe_dnn_softmax, (boot.h,), t, (BoundB, BoundR, GumbelLambda, GoldSchmidtD), len(t.cts))


def resnet20_rotation_indices(width):
    """PrePareResNet20's rotation set (Resnet/models/model_resnet20.cu:3-15) for a width x width image"""
    n = C.c_size_t()
    _check(lib().b200fhe_resnet20_rotation_indices(width, None, 0, C.byref(n)))
    buf = (C.c_int32 * n.value)()
    _check(lib().b200fhe_resnet20_rotation_indices(width, buf, n.value, C.byref(n)))
    return list(buf)


def resnet20_param_count(in_ch=3, base_ch=16, blocks=3, classes=10):
    return int(lib().b200fhe_resnet20_param_count(in_ch, base_ch, blocks, classes))


def resnet20_infer(dnn, boot, t, params, base_ch=16, blocks=3, classes=10, timings=False):
    """ResNet20_infer (Resnet/models/model_resnet20.cu:44-153); params: flat float64 array (see include/b200fhe.h)"""
    p, pp = _f64(params)
    arr, d = t._args()
    cap = max(classes, base_ch * 4)
    outs = (C.c_void_p * cap)()
    od = TensorDesc()
    ms = (C.c_double * 4)()
    _check(lib().b200fhe_resnet20_infer(dnn.h, boot.h, arr, C.byref(d), pp, p.size, base_ch, blocks, classes, outs, cap,
                                        C.byref(od), ms if timings else None))
    out = TensorCT([Ciphertext(dnn.ev, outs[i]) for i in range(od.num_cts)], od.width, od.slotstr, od.num_ch)
    return (out, dict(conv_ms=ms[0], bn_add_ms=ms[1], relu_ms=ms[2], pool_fc_ms=ms[3])) if timings else out


class EvalKey:
    """One evaluation key (dnum digits of [2][size_QP][N]) owned by the library."""

    def __init__(self, h):
        self.h = h
        n = lib().b200fhe_evk_digits(h, None, 0)
        a = (C.c_void_p * n)()
        lib().b200fhe_evk_digits(h, a, n)
        self.digits = [int(x) for x in a]

    def free(self):
        if getattr(self, "h", None):
            lib().b200fhe_evk_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class KeyGenerator:
    """Secret / public / evaluation keys, encryption and decryption (PhantomSecretKey, PhantomPublicKey:
    include/secretkey.h, src/secretkey.cu).  seed: 32 bytes (reproducible key set) or None (operating-system entropy)."""

    def __init__(self, ctx, seed=None, reference_prng_seed=None):
        self.ctx = ctx
        h = C.c_void_p()
        if reference_prng_seed is not None:   # the reference's own samplers and seed order (same-seed key parity)
            _check(lib().b200fhe_keygen_create_reference_prng(ctx.h, int(reference_prng_seed), _stream(), C.byref(h)))
            self.h = h
            return
        if seed is not None:
            seed = bytes(seed)
            if len(seed) != 32:
                raise ValueError("seed must be 32 bytes")
            sp = C.cast(C.c_char_p(seed), C.c_void_p)
        else:
            sp = None
        _check(lib().b200fhe_keygen_create(ctx.h, sp, _stream(), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_keygen_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def secret(self):
        """a COPY of the secret key, [size_QP][N] NTT form"""
        import torch
        out = torch.empty((self.ctx.size_QP, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_keygen_get_secret(self.h, _ptr(out)))
        return out

    def set_secret(self, t):
        _check(lib().b200fhe_keygen_set_secret(self.h, _ptr(t)))

    def relin_key(self):
        h = C.c_void_p()
        _check(lib().b200fhe_keygen_relin_key(self.h, C.byref(h)))
        return EvalKey(h)

    def galois_key(self, k):
        h = C.c_void_p()
        _check(lib().b200fhe_keygen_galois_key(self.h, k, C.byref(h)))
        return EvalKey(h)

    def galois_key_seal(self, elt):
        """PhantomGaloisKey-style key of Galois element elt (sigma_elt(s) -> s), for Evaluator.apply_galois / rotate / hoisting"""
        h = C.c_void_p()
        _check(lib().b200fhe_keygen_galois_key_seal(self.h, elt, C.byref(h)))
        return EvalKey(h)

    def rotation_keys(self, rotations, conj=True):
        """EvalAtIndexKeyGen: {automorphism index: EvalKey} for the slot rotations (+ the conjugation key)"""
        ks = {auto_index(self.ctx.logN, int(r)) for r in rotations}
        if conj:
            ks.add(2 * self.ctx.N - 1)
        return {k: self.galois_key(k) for k in sorted(ks)}

    def public_key(self):
        import torch
        pk = torch.empty((2, self.ctx.size_QP, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_keygen_public_key(self.h, _ptr(pk)))
        return pk

    def encrypt_symmetric(self, plain, size_Ql):
        import torch
        ct = torch.empty((2, size_Ql, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_encrypt_symmetric(self.h, _ptr(plain), size_Ql, _ptr(ct)))
        return ct

    def encrypt_asymmetric(self, pk, plain, size_Ql):
        import torch
        ct = torch.empty((2, size_Ql, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_encrypt_asymmetric(self.h, _ptr(pk), _ptr(plain), size_Ql, _ptr(ct)))
        return ct

    def decrypt(self, ct, size_Ql, polys=2):
        import torch
        out = torch.empty((size_Ql, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_decrypt(self.h, _ptr(ct), size_Ql, polys, _ptr(out)))
        return out

    def sample(self, kind, limbs, size_Ql=0):
        import torch
        out = torch.empty((limbs, self.ctx.N), dtype=torch.int64, device="cuda")
        _check(lib().b200fhe_keygen_sample(self.h, kind, _ptr(out), limbs, size_Ql))
        return out


def chacha20_block(ctx, key8, counter, nonce3):
    """one ChaCha20 block computed on the device (known-answer test of the samplers' generator)"""
    k = (C.c_uint32 * 8)(*key8)
    n = (C.c_uint32 * 3)(*nonce3)
    out = (C.c_uint32 * 16)()
    _check(lib().b200fhe_chacha20_block(ctx.h, k, counter, n, out))
    return list(out)


# ------------------------------------------------------------------------------------------- serialization
class CtHeader(C.Structure):
    """b200fhe_ct_header: the 58-byte header of PhantomCiphertext::save (include/ciphertext.h:184-193)"""
    _fields_ = [("chain_index", C.c_uint64), ("polys", C.c_uint64), ("N", C.c_uint64), ("limbs", C.c_uint64),
                ("scale", C.c_double), ("correction_factor", C.c_uint64), ("noise_scale_deg", C.c_uint64),
                ("is_ntt_form", C.c_uint8), ("is_asymmetric", C.c_uint8)]


def _save(call):
    """two-pass save: size query, then fill"""
    n = C.c_size_t()
    _check(call(None, 0, C.byref(n)))
    buf = (C.c_uint8 * n.value)()
    _check(call(buf, n.value, C.byref(n)))
    return bytes(buf)


def save_ciphertext(ctx, t, size_Ql, scale, noise=1, asymmetric=False):
    """t: [polys][size_Ql][N] device tensor -> bytes in the reference's PhantomCiphertext format"""
    h = CtHeader(ctx.size_Q - size_Ql + 1, t.shape[0], ctx.N, size_Ql, scale, 1, noise, 1, 1 if asymmetric else 0)
    return _save(lambda b, cap, n: lib().b200fhe_ct_save(C.byref(h), _ptr(t), b, cap, n))


def load_ciphertext(data):
    """bytes -> (header, device tensor [polys][limbs][N], bytes consumed)"""
    import torch
    h, used = CtHeader(), C.c_size_t()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    _check(lib().b200fhe_ct_load(buf, len(data), C.byref(h), None, 0, C.byref(used)))
    t = torch.empty((h.polys, h.limbs, h.N), dtype=torch.int64, device="cuda")
    _check(lib().b200fhe_ct_load(buf, len(data), C.byref(h), _ptr(t), t.numel(), C.byref(used)))
    return h, t, used.value


def save_plaintext(ctx, t, size_Ql, scale):
    return _save(lambda b, cap, n: lib().b200fhe_plain_save(ctx.size_Q - size_Ql + 1, ctx.N, t.shape[0], scale, _ptr(t), b, cap, n))


def load_plaintext(data):
    import torch
    ci, N, limbs, scale, used = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_double(), C.c_size_t()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    _check(lib().b200fhe_plain_load(buf, len(data), C.byref(ci), C.byref(N), C.byref(limbs), C.byref(scale), None, 0, C.byref(used)))
    t = torch.empty((limbs.value, N.value), dtype=torch.int64, device="cuda")
    _check(lib().b200fhe_plain_load(buf, len(data), None, None, None, None, _ptr(t), t.numel(), C.byref(used)))
    return ci.value, scale.value, t, used.value


def save_eval_key(ctx, digits):
    """digits: device pointers or tensors of one key -> bytes in the reference's PhantomRelinKey format"""
    a = (C.c_void_p * len(digits))(*[d if isinstance(d, int) else d.data_ptr() for d in digits])
    return _save(lambda b, cap, n: lib().b200fhe_evk_save(ctx.h, a, len(digits), b, cap, n))


def load_eval_key(ctx, data):
    """bytes -> (list of digit tensors [2][size_QP][N], bytes consumed)"""
    import torch
    dnum, used = C.c_size_t(), C.c_size_t()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    _check(lib().b200fhe_evk_load(ctx.h, buf, len(data), None, 0, C.byref(dnum), C.byref(used)))
    digits = [torch.empty((2, ctx.size_QP, ctx.N), dtype=torch.int64, device="cuda") for _ in range(dnum.value)]
    a = (C.c_void_p * dnum.value)(*[d.data_ptr() for d in digits])
    _check(lib().b200fhe_evk_load(ctx.h, buf, len(data), a, dnum.value, C.byref(dnum), C.byref(used)))
    return digits, used.value


def save_secret_key(ctx, secret):
    return _save(lambda b, cap, n: lib().b200fhe_secret_save(ctx.h, _ptr(secret), b, cap, n))


def load_secret_key(ctx, data):
    import torch
    t = torch.empty((ctx.size_QP, ctx.N), dtype=torch.int64, device="cuda")
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    _check(lib().b200fhe_secret_load(ctx.h, buf, len(data), _ptr(t), None))
    return t


_GK_MAGIC = b"B2GK"


def save_galois_keys(ctx, keys):
    """{automorphism index: digits} -> bytes.  The reference's PhantomGaloisKeyFused has no save(): this container is
    ours -- magic, key count (u64), then (index u32, PhantomRelinKey record) per key in ascending index order."""
    import struct
    out = [_GK_MAGIC, struct.pack("<Q", len(keys))]
    for k in sorted(keys):
        d = keys[k].digits if hasattr(keys[k], "digits") else keys[k]
        out += [struct.pack("<I", k), save_eval_key(ctx, d)]
    return b"".join(out)


def load_galois_keys(ctx, data):
    import struct
    if data[:4] != _GK_MAGIC or len(data) < 12:
        raise B200FheError("not a Galois-key container")
    (n,), off, keys = struct.unpack_from("<Q", data, 4), 12, {}
    for _ in range(n):
        if off + 4 > len(data):
            raise B200FheError("truncated Galois-key container")
        (k,) = struct.unpack_from("<I", data, off)
        digits, used = load_eval_key(ctx, data[off + 4:])
        keys[k] = digits
        off += 4 + used
    return keys
