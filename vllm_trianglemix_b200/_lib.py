"""ctypes binding of libtmx.so (the C ABI declared in include/tmx.h).

There is no CPU or eager fallback: if the shared library is missing the import of
any op raises, and every non-zero status becomes a RuntimeError carrying
tmx_strerror() (the reference raises plain torch errors).
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtmx.so")

_c = ctypes
_i32, _i64, _f32, _vp, _sz = _c.c_int, _c.c_int64, _c.c_float, _c.c_void_p, _c.c_size_t

# name -> (restype, argtypes); must list every symbol include/tmx.h declares
SIGNATURES = {
    "tmx_strerror": (_c.c_char_p, [_i32]),
    "tmx_last_cuda_error": (_c.c_char_p, []),
    "tmx_version": (_i32, []),
    "tmx_mask_bits": (_i32, [_vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "tmx_mask_dense01": (_i32, [_vp, _i32, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "tmx_mask_additive": (_i32, [_vp, _i32, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "tmx_mask_rewrite_neg_inf": (_i32, [_vp, _vp, _i32, _i64, _f32, _vp]),
    "tmx_mask_bits_from_additive": (_i32, [_vp, _i64, _vp, _i32, _i64, _i32, _vp, _vp]),
    "tmx_block_index": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "tmx_attn_workspace_bytes": (_sz, [_i32] * 11),
    "tmx_prefill_attn": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                _i64, _i64, _i64, _f32, _i32, _i32, _i32, _i32, _vp, _i64,
                                _vp, _vp, _sz, _i32, _vp]),
    "tmx_prefill_attn_static": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                       _i64, _i64, _i64, _f32, _i32, _i32, _i32, _i32, _vp, _i64,
                                       _vp, _i32, _vp]),
    "tmx_chunked_prefill_attn": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i64, _i64,
                                        _vp, _vp, _i64, _i64, _vp, _vp, _i64, _i32, _f32, _i32, _i32, _i32,
                                        _i32, _vp, _vp, _sz, _i32, _vp]),
    "tmx_qk_norm_rope": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32,
                                _i64, _i64, _i64, _i64, _f32, _vp, _i64, _vp, _vp, _vp,
                                _i64, _i64, _i64, _i32, _vp]),
    "tmx_reshape_and_cache_flash": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i64, _i64,
                                           _i64, _i64, _i64, _i32, _vp]),
    "tmx_linear_bf16": (_i32, [_vp, _i32, _i32, _i32, _i64, _vp, _vp, _i32, _i32, _i32, _i64, _i64, _i64, _vp]),
    "tmx_sum_partials": (_i32, [_vp, _vp, _i32, _i64, _i32, _i64, _vp]),
    "tmx_last_launch_count": (_i32, []),
}

_lib = None
_lock = threading.Lock()


class TmxError(RuntimeError):
    pass


def load():
    """Load libtmx.so once; raise (never fall back) if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise TmxError(
                    "libtmx.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; "
                    "g.build()'` or `make -C vllm_trianglemix_b200/csrc`. There is no CPU fallback." % LIB_PATH)
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def check(status: int, what: str):
    if status != 0:
        lib = load()
        msg = lib.tmx_strerror(status).decode()
        if status == 6:
            msg += " [" + lib.tmx_last_cuda_error().decode() + "]"
        raise TmxError("%s failed: %s (status %d)" % (what, msg, status))
