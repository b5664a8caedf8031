"""ctypes binding of libstransfer_b200.so (include/stransfer_b200.h).

There is deliberately NO fallback: if the library is missing or a call fails the
product path raises.  ``lib()`` loads (and, if the sources changed, rebuilds) the
in-tree shared object.
"""
import ctypes
import os
import sys
import time
from ctypes import c_char_p, c_double, c_float, c_int, c_size_t, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libstransfer_b200.so')

_P, _I, _F, _Z, _D = c_void_p, c_int, c_float, c_size_t, c_double

# name -> (restype, argtypes); mirrors include/stransfer_b200.h one to one
SIGNATURES = {
    'st_arch': (_I, []),
    'st_last_error': (c_char_p, []),
    'st_set_conv_path': (_I, [_I]),
    'st_last_conv_kernel': (_I, []),
    'st_launch_count': (ctypes.c_longlong, []),
    'st_nchw_to_nhwc': (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P]),
    'st_nhwc_to_nchw': (_I, [_P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _P]),
    'st_image_u8_to_nchw': (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    'st_image_nchw_to_u8': (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    'st_pack_conv_weight': (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P]),
    'st_conv2d': (_I, [_P, _P, _I, _I, _I, _I, _I, _I,      # x, planes, N H W Cin up
                       _P, _P, _I, _I, _I,                  # w, planes, rows_total, img_rows
                       _I, _I, _I, _I,                      # ks stride pad transposed
                       _I, _I, _I, _I,                      # Ho Wo Cout_pad Cbias
                       _P, _P, _P, _P, _I,                  # bias mask add_hi add_lo add_planes
                       _P, _P, _P, _P, _I,                  # out_hi out_lo relu_hi relu_lo out_planes
                       _P, _I, _F, _P, _P, _P]),            # out_nchw Cnchw acc_mul acc_mul_slot add_mul_slot stream
    'st_conv2d_routed': (_I, [_P, _P, _I, _I, _I, _I, _I,   # x, planes, N H W Cin
                              _P, _P, _I, _I, _I, _I, _I,   # w, planes, rows_total, img_rows, ks, Cout_pad
                              _P, _P, _P, _I, _P,           # mask gy_hi gy_lo gy_planes route
                              _P, _P, _I, _F, _P, _P, _P]), # out_hi out_lo out_planes acc_mul acc_mul_slot gy_mul_slot stream
    'st_conv2d_pool': (_I, [_P, _P, _I, _I, _I, _I, _I,     # x, planes, N H W Cin
                            _P, _P, _I, _I, _I, _I, _I, _P, # w, planes, rows_total, ks, Cout_pad, Cbias, bias
                            _P, _P, _I, _P, _P, _P,         # out_hi out_lo out_planes pool_hi pool_lo pool_idx
                            _F, _P, _P]),                   # acc_mul acc_mul_slot stream
    'st_conv3x3_image_fwd': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    'st_conv3x3_image_dgrad': (_I, [_P, _P, _I, _P, _P, _I, _I, _I, _I, _P]),
    'st_conv9_expand': (_I, [_P, _P, _I, _I, _P, _P, _I, _P, _P, _I, _I, _I, _I, _P]),
    'st_conv9_reduce_workspace_bytes': (_Z, [_I, _I, _I]),
    'st_conv9_reduce': (_I, [_P, _P, _I, _P, _P, _P, _P, _Z, _I, _I, _I, _P]),
    'st_conv9_wgrad_workspace_bytes': (_Z, [_I]),
    'st_conv9_wgrad': (_I, [_P, _P, _I, _P, _P, _P, _P, _I, _I, _I, _P, _P, _P, _Z, _I, _I, _I, _P]),
    'st_conv2d_wgrad_workspace_bytes': (_Z, [_I, _I, _I, _I, _I, _I]),
    'st_conv2d_wgrad': (_I, [_P, _P, _I, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I,
                             _P, _P, _P, _Z, _P]),
    'st_relu_maxpool2x2_fwd': (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    'st_relu_maxpool2x2_bwd': (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    'st_relu_maxpool2x2_fwd_idx': (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    'st_relu_maxpool2x2_bwd_idx': (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    'st_gram_workspace_bytes': (_Z, [_I, _I, _I]),
    'st_set_gram_max_splits': (_I, [_I]),
    'st_gram_fwd': (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _Z, _P]),
    'st_gram_style_workspace_bytes': (_Z, [_I, _I, _I]),
    'st_gram_style_fwd': (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _Z, _P]),
    'st_style_mse_bwd_seed_multi': (_I, [_I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    'st_style_mse_fwd': (_I, [_P, _P, _I, _I, _P, _P, _P, _P, _P]),
    'st_style_mse_bwd_seed': (_I, [_P, _I, _I, _I, _P, _P, _P, _I, _P, _P, _P]),
    'st_loss_total': (_I, [_P, _P, _I, _P, _F, _P, _F, _P, _P]),
    'st_absmax_slot': (_I, [_P, _Z, _F, _P, _P, _P]),
    'st_mse_fwd': (_I, [_P, _P, _I, _P, _P, _I, _Z, _I, _I, _P, _P, _P]),
    'st_mse_bwd': (_I, [_P, _P, _I, _P, _P, _I, _Z, _I, _F, _P, _P, _I, _P, _P, _P]),
    'st_tv_fwd': (_I, [_P, _I, _I, _I, _I, _F, _P, _P, _P]),
    'st_tv_bwd': (_I, [_P, _I, _I, _I, _I, _F, _P, _P, _P]),
    'st_temporal_fwd': (_I, [_P, _P, _P, _P, _Z, _F, _P, _P, _P]),
    'st_temporal_bwd': (_I, [_P, _P, _Z, _F, _P, _P, _P, _P]),
    'st_temporal_warp_fwd': (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _P, _P, _P]),
    'st_temporal_warp_bwd': (_I, [_P, _P, _P, _I, _I, _I, _I, _F, _P, _P, _P, _P]),
    'st_instnorm_workspace_bytes': (_Z, [_I, _I]),
    'st_instnorm_fwd': (_I, [_P, _P, _P, _P, _F, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    'st_instnorm_fwd_static': (_I, [_P, _P, _P, _P, _F, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _P]),
    'st_adam_chunk_count': (_Z, [_P, _I]),
    'st_adam_fill_chunks': (_I, [_P, _I, _P]),
    'st_adam_step': (_I, [_P, _P, _P, _P, _P, _Z, _P, _P, _I, _D, _D, _D, _D, _D, _D, _P, _P]),
    'st_lbfgs_state_bytes': (_Z, [_I, _Z]),
    'st_set_lbfgs_variant': (_I, [_I]),
    'st_lbfgs_init': (_I, [_P, _I, _Z, _D, _D, _D, _P]),
    'st_lbfgs_iteration': (_I, [_P, _P, _P, _P, _P, _P, _P, _Z, _Z, _I, _P, _I, _P]),
    'st_instnorm_bwd': (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
}

_lib = None


class StransferError(RuntimeError):
    pass


def lib():
    """Load the shared object (building it first when stale/missing)."""
    global _lib
    if _lib is not None:
        return _lib
    from . import build as _build
    if not _build.up_to_date():
        sys.stderr.write('styletransfer_b200: libstransfer_b200.so is stale or missing -- rebuilding with nvcc (about a minute)\n')
        _build.build()                      # serialised across processes by a file lock, published atomically
    if not os.path.exists(LIB_PATH):
        raise StransferError(f'{LIB_PATH} is missing: run `python -m styletransfer_b200.build`')
    handle = None
    for attempt in range(5):
        try:
            handle = ctypes.CDLL(LIB_PATH)
            break
        except OSError as e:                # e.g. a snapshot of the tree that is still being materialised
            if attempt == 4:
                raise StransferError(f'cannot load {LIB_PATH}: {e}') from e
            sys.stderr.write(f'styletransfer_b200: loading {LIB_PATH} failed ({e}); retrying\n')
            time.sleep(2.0 * (attempt + 1))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if handle.st_arch() != 100:
        raise StransferError('libstransfer_b200.so was not built for sm_100a')
    _lib = handle
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().st_last_error().decode(errors='replace')
        raise StransferError(f'{what} failed (code {rc}): {msg}')
