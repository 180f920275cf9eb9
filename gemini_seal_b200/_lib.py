"""ctypes binding of the C ABI declared in include/gseal_b200.h.

The product only ever loads the nvcc-built gemini_seal_b200/_C/libgseal_b200.so.  There is no CPU
fallback: if the library is missing, `load()` raises.  `bind()` is shared with the test-only
simulator build (tests/cusim) so both expose identical prototypes.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_C", "libgseal_b200.so")

u64 = C.c_uint64
u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p
sz = C.c_size_t
i32 = C.c_int
u32 = C.c_uint32

# name -> (restype, argtypes).  Keep in the order of include/gseal_b200.h.
PROTOTYPES = {
    "gsb_last_error": (C.c_char_p, []),
    "gsb_version": (i32, []),
    "gsb_launch_count": (C.c_ulonglong, []),
    "gsb_prof_enable": (i32, [i32]),
    "gsb_prof_report": (C.c_long, [C.c_char_p, sz]),
    "gsb_ctx_create": (i32, [i32, i32, u64p, sz, sz, C.POINTER(vp)]),
    "gsb_ctx_destroy": (None, [vp]),
    "gsb_ctx_get_tables": (i32, [vp, sz, u64p, u64p, u64p, u64p, u64p]),
    "gsb_ctx_info": (i32, [vp, C.POINTER(i32), C.POINTER(sz), C.POINTER(sz)]),
    "gsb_ws_create": (i32, [vp, C.POINTER(vp)]),
    "gsb_ws_destroy": (None, [vp]),
    "gsb_malloc": (i32, [vp, sz, C.POINTER(vp)]),
    "gsb_free": (i32, [vp, vp]),
    "gsb_malloc_host": (i32, [sz, C.POINTER(vp)]),
    "gsb_free_host": (i32, [vp]),
    "gsb_memcpy_h2d": (i32, [vp, vp, vp, sz, vp]),
    "gsb_memcpy_d2h": (i32, [vp, vp, vp, sz, vp]),
    "gsb_stream_sync": (i32, [vp, vp]),
    "gsb_ntt_forward_dev": (i32, [vp, vp, sz, sz, sz, i32, vp]),
    "gsb_ntt_inverse_dev": (i32, [vp, vp, sz, sz, sz, i32, vp]),
    "gsb_ntt_forward_host": (i32, [vp, vp, sz, sz, sz, i32]),
    "gsb_ntt_inverse_host": (i32, [vp, vp, sz, sz, sz, i32]),
    "gsb_poly_op_dev": (i32, [vp, i32, vp, vp, u64p, vp, sz, sz, sz, vp]),
    "gsb_poly_op_host": (i32, [vp, i32, vp, vp, u64p, vp, sz, sz, sz]),
    "gsb_key_upload": (i32, [vp, i32, u32, vp, sz]),
    "gsb_key_upload_bundles": (i32, [vp, i32, u32, C.POINTER(vp), sz]),
    "gsb_key_drop": (i32, [vp, i32, u32]),
    "gsb_galois_elt_from_step": (i32, [vp, i32, C.POINTER(u32)]),
    "gsb_galois_table": (i32, [vp, u32, C.POINTER(u32)]),
    "gsb_ckks_multiply_dev": (i32, [vp, vp, vp, vp, sz, sz, vp]),
    "gsb_switch_key_dev": (i32, [vp, vp, vp, sz, vp, sz, i32, u32, sz, sz, vp]),
    "gsb_relinearize_dev": (i32, [vp, vp, vp, vp, sz, sz, vp]),
    "gsb_rescale_dev": (i32, [vp, vp, vp, vp, sz, sz, sz, vp]),
    "gsb_apply_galois_dev": (i32, [vp, vp, u32, vp, vp, sz, sz, vp]),
    "gsb_mul_relin_rescale_dev": (i32, [vp, vp, vp, vp, vp, sz, sz, vp]),
    "gsb_ckks_multiply_host": (i32, [vp, vp, vp, vp, sz, sz]),
    "gsb_relinearize_host": (i32, [vp, vp, vp, sz, sz]),
    "gsb_rescale_host": (i32, [vp, vp, vp, sz, sz, sz]),
    "gsb_apply_galois_host": (i32, [vp, u32, vp, vp, sz, sz]),
    "gsb_mul_relin_rescale_host": (i32, [vp, vp, vp, vp, sz, sz]),
    "gsb_bfv_multiply_dev": (i32, [vp, vp, u64, vp, vp, vp, sz, sz, vp]),
    "gsb_bfv_multiply_host": (i32, [vp, u64, vp, vp, vp, sz, sz]),
    "gsb_bfv_aux_base": (i32, [vp, sz, u64, u64p, sz, C.POINTER(sz)]),
    "gsb_bfv_mod_switch_dev": (i32, [vp, vp, vp, sz, sz, sz, vp]),
    "gsb_bfv_mod_switch_host": (i32, [vp, vp, vp, sz, sz, sz]),
    "gsb_bfv_relinearize_dev": (i32, [vp, vp, vp, vp, sz, sz, vp]),
    "gsb_bfv_apply_galois_dev": (i32, [vp, vp, u32, vp, vp, sz, sz, vp]),
    "gsb_bfv_relinearize_host": (i32, [vp, vp, vp, sz, sz]),
    "gsb_bfv_apply_galois_host": (i32, [vp, u32, vp, vp, sz, sz]),
    "gsb_ckks_multiply_sizes_dev": (i32, [vp, vp, sz, vp, sz, vp, sz, sz, vp]),
    "gsb_ckks_multiply_sizes_host": (i32, [vp, vp, sz, vp, sz, vp, sz, sz]),
    "gsb_bfv_multiply_sizes_dev": (i32, [vp, vp, u64, vp, sz, vp, sz, vp, sz, sz, vp]),
    "gsb_bfv_multiply_sizes_host": (i32, [vp, u64, vp, sz, vp, sz, vp, sz, sz]),
    "gsb_relinearize_sizes_dev": (i32, [vp, vp, vp, sz, vp, sz, sz, i32, vp]),
    "gsb_relinearize_sizes_host": (i32, [vp, vp, sz, vp, sz, sz, i32]),
    "gsb_bfv_add_plain_dev": (i32, [vp, vp, vp, sz, i32, u64p, u64, u64, u64, sz, vp]),
    "gsb_bfv_add_plain_host": (i32, [vp, vp, vp, sz, i32, u64p, u64, u64, u64, sz]),
    "gsb_bfv_plain_to_rns_dev": (i32, [vp, vp, sz, u64, u64p, vp, sz, i32, vp]),
    "gsb_bfv_plain_to_rns_host": (i32, [vp, vp, sz, u64, u64p, vp, sz, i32]),
    "gsb_multiply_plain_coeff_dev": (i32, [vp, vp, vp, sz, sz, sz, vp]),
    "gsb_multiply_plain_coeff_host": (i32, [vp, vp, sz, sz, vp, sz, u64, u64p, i32]),
    "gsb_kswitch_keygen_dev": (i32, [vp, vp, i32, C.POINTER(u32), sz, vp, vp, vp, i32, vp]),
    "gsb_kswitch_keygen_host": (i32, [vp, i32, C.POINTER(u32), sz, vp, vp, vp, i32]),
    "gsb_drop_last_limb_dev": (i32, [vp, vp, vp, sz, sz, vp]),
    "gsb_secret_key_set": (i32, [vp, vp]),
    "gsb_dot_product_ct_sk_dev": (i32, [vp, vp, vp, vp, sz, sz, sz, i32, vp]),
    "gsb_dot_product_ct_sk_host": (i32, [vp, vp, vp, sz, sz, sz, i32]),
    "gsb_bfv_decrypt_dev": (i32, [vp, vp, u64, vp, vp, sz, sz, sz, vp]),
    "gsb_bfv_decrypt_host": (i32, [vp, u64, vp, vp, sz, sz, sz]),
    "gsb_rescale_noalloc_dev": (i32, [vp, vp, vp, sz, sz, sz, vp]),
    "gsb_ctx_fast_ntt": (i32, [vp]),
    "gsb_ctx_shortcuts": (i32, [vp]),
}
# symbols that the test-only simulator build does not provide
CUDA_ONLY = {"gsb_rescale_noalloc_dev"}  # static __device__ scratch: CUDA build only


def bind(lib):
    """Attach prototypes; raises AttributeError if a declared symbol is not exported."""
    for name, (res, args) in PROTOTYPES.items():
        if name in CUDA_ONLY and not hasattr(lib, name):
            continue
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


_lib = None


def load(path=None):
    """Load the CUDA library (built by gemini_seal_b200.build.build()).  No fallback."""
    global _lib
    if path is None and _lib is not None:
        return _lib
    p = path or os.environ.get("GSB_B200_LIB") or LIB_PATH  # GSB_B200_LIB: an alternative nvcc build (tuning experiments)
    if not os.path.exists(p):
        raise RuntimeError(
            "gemini_seal_b200: CUDA library %s is missing; run `python -m gemini_seal_b200.build` "
            "(there is no CPU fallback)" % p)
    lib = bind(C.CDLL(p))
    if path is None:
        _lib = lib
    return lib
