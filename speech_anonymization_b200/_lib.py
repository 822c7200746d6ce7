"""ctypes binding of libspanon.so (include/spanon.h).

There is NO fallback: if the shared library is missing or a call returns a non-zero
status the caller gets an exception.  PyTorch is only used for device memory and streams.
"""
import ctypes as C
import os

import torch  # noqa: F401  (loads the CUDA runtime the library links against)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libspanon.so")

c_f32p = C.POINTER(C.c_float)
c_i32p = C.POINTER(C.c_int32)


class SpanonError(RuntimeError):
    pass


class FbankCfg(C.Structure):
    _fields_ = [("n_fft", C.c_int32), ("win_length", C.c_int32), ("hop_length", C.c_int32),
                ("n_mels", C.c_int32), ("amin", C.c_float), ("multiplier", C.c_float),
                ("db_offset", C.c_float), ("top_db", C.c_float), ("window", C.c_void_p),
                ("mel_start", C.c_void_p), ("mel_len", C.c_void_p), ("mel_off", C.c_void_p),
                ("mel_w", C.c_void_p)]


class BatchNorm(C.Structure):
    _fields_ = [("weight", C.c_void_p), ("bias", C.c_void_p), ("running_mean", C.c_void_p),
                ("running_var", C.c_void_p)]


class FcaeWeights(C.Structure):
    _fields_ = [("in_dim", C.c_int32),
                ("enc_w", C.c_void_p * 3), ("enc_b", C.c_void_p * 3),
                ("dec_w", C.c_void_p * 3), ("dec_b", C.c_void_p * 3),
                ("clf_norm", BatchNorm),
                ("init_w", C.c_void_p * 2), ("init_b", C.c_void_p * 2),
                ("cls_w", C.c_void_p * 4), ("cls_b", C.c_void_p * 4),
                ("cls_bn1", BatchNorm), ("cls_bn6", BatchNorm)]


class ConvAEWeights(C.Structure):
    _fields_ = [("enc_w", C.c_void_p * 5), ("enc_b", C.c_void_p * 5),
                ("enc_in_w", C.c_void_p * 4), ("enc_in_b", C.c_void_p * 4),
                ("dec_w", C.c_void_p * 5), ("dec_b", C.c_void_p * 5),
                ("dec_in_w", C.c_void_p * 2), ("dec_in_b", C.c_void_p * 2),
                ("clf_norm", BatchNorm),
                ("tdnn_w", C.c_void_p * 3), ("tdnn_b", C.c_void_p * 3),
                ("tdnn_bn", BatchNorm * 3),
                ("cls_w", C.c_void_p * 3), ("cls_b", C.c_void_p * 3),
                ("cls_bn", BatchNorm * 2)]


class ConvReconWeights(C.Structure):
    _fields_ = [("conv_w", C.c_void_p * 5), ("conv_b", C.c_void_p * 5),
                ("in_w", C.c_void_p * 4), ("in_b", C.c_void_p * 4)]


class XvectorWeights(C.Structure):
    _fields_ = [("tdnn_w", C.c_void_p * 5), ("tdnn_b", C.c_void_p * 5), ("tdnn_bn", BatchNorm * 5),
                ("emb_w", C.c_void_p), ("emb_b", C.c_void_p), ("clf_norm", BatchNorm),
                ("dnn_w", C.c_void_p), ("dnn_b", C.c_void_p), ("dnn_bn", BatchNorm),
                ("out_w", C.c_void_p), ("out_b", C.c_void_p)]


class ConvAEOpts(C.Structure):
    _fields_ = [("precision", C.c_int32), ("chunk_utts", C.c_int32), ("skip_classifier", C.c_int32),
                ("weights_prepared", C.c_int32), ("recon_fp16", C.c_int32)]


class TapConv(C.Structure):
    _fields_ = [("kind", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32), ("k", C.c_int32), ("dilation", C.c_int32),
                ("padding", C.c_int32)]


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 56), ("launches", C.c_uint64), ("ms_total", C.c_double)]


PREC_FP32, PREC_FP16 = 0, 1

# name -> (restype, argtypes); must list EVERY function include/spanon.h declares
SIGNATURES = {
    "spanon_strerror": (C.c_char_p, [C.c_int]),
    "spanon_abi_version": (C.c_int, []),
    "spanon_launch_count": (C.c_uint64, []),
    "spanon_last_cuda_error": (C.c_char_p, []),
    "spanon_profile_begin": (C.c_int, [C.c_void_p]),
    "spanon_profile_end": (C.c_int, [C.POINTER(KernelTime), C.c_int]),
    "spanon_num_frames": (C.c_int64, [C.c_int64, C.c_int32]),
    "spanon_fbank_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                   C.POINTER(FbankCfg), C.c_void_p]),
    "spanon_fbank_i16": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                   C.POINTER(FbankCfg), C.c_void_p]),
    "spanon_fbank_clamp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_float,
                                     C.c_void_p]),
    "spanon_norm_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p,
                                    C.c_float, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]),
    "spanon_norm_apply": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                    C.c_float, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_fcae_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "spanon_fcae_fwd": (C.c_int, [C.c_void_p, C.POINTER(FcaeWeights), C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                  C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_fcae_classify": (C.c_int, [C.c_void_p, C.POINTER(FcaeWeights), C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                       C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_fcae_static_max_batch": (C.c_int64, []),
    "spanon_fcae_fwd_static": (C.c_int, [C.c_void_p, C.POINTER(FcaeWeights), C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                         C.c_void_p]),
    "spanon_convae_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.POINTER(ConvAEOpts)]),
    "spanon_convae_fwd": (C.c_int, [C.c_void_p, C.POINTER(ConvAEWeights), C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_float, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_size_t, C.c_int64, C.c_int64, C.c_int64, C.POINTER(ConvAEOpts),
                                    C.c_void_p]),
    "spanon_convrecon_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.POINTER(ConvAEOpts)]),
    "spanon_convrecon_fwd": (C.c_int, [C.c_void_p, C.POINTER(ConvReconWeights), C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_float, C.c_int64, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int64, C.c_int64,
                                       C.c_int64, C.POINTER(ConvAEOpts), C.c_void_p]),
    "spanon_xvector_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "spanon_xvector_classify": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(XvectorWeights), C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_size_t, C.c_int64, C.c_int64, C.c_int64, C.c_int32,
                                          C.c_void_p]),
    "spanon_tapconv_workspace_bytes": (C.c_size_t, [C.POINTER(TapConv)]),
    "spanon_tapconv_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.c_size_t, C.c_int64, C.c_int64, C.POINTER(TapConv), C.c_void_p]),
    "spanon_tapconv_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_size_t, C.c_int64, C.c_int64, C.POINTER(TapConv), C.c_void_p]),
    "spanon_conv1_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_conv1_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                   C.c_void_p]),
    "spanon_colstats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_norm_act_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                      C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_void_p]),
    "spanon_norm_act_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_int64, C.c_int64, C.c_int, C.c_double, C.c_void_p]),
    "spanon_pool_reinterp_fwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                           C.c_int64, C.c_int64, C.c_void_p]),
    "spanon_pool_reinterp_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                           C.c_int64, C.c_void_p]),
    "spanon_specaug_warp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "spanon_specaug_mask": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_int32, C.c_void_p, C.c_void_p]),
    "spanon_cosine_loss": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p]),
    "spanon_cosine_loss_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                         C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "spanon_knn_mi_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                      C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "spanon_anonymize_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                                      C.POINTER(ConvAEOpts)]),
    "spanon_anonymize_host": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(FbankCfg),
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(ConvAEOpts),
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "spanon_anonymize_host_i16": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(FbankCfg),
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(ConvAEOpts),
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "spanon_pipe_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int64, C.POINTER(FbankCfg),
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(ConvAEOpts), C.c_int32]),
    "spanon_pipe_submit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "spanon_pipe_submit_i16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "spanon_pipe_wait": (C.c_int, [C.c_void_p]),
    "spanon_pipe_drain": (C.c_int, [C.c_void_p]),
    "spanon_pipe_weights_changed": (C.c_int, [C.c_void_p, C.c_void_p]),
    "spanon_pipe_destroy": (None, [C.c_void_p]),
}

_lib = None


def lib():
    """Loads libspanon.so (once).  Raises if it has not been built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SpanonError(
                "libspanon.so not found at %s: build it with `python -m speech_anonymization_b200.build` "
                "(or __graft_entry__.build()); this package has no CPU / PyTorch fallback" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status, what=""):
    if status != 0:
        l = lib()
        msg = l.spanon_strerror(status).decode()
        extra = l.spanon_last_cuda_error().decode() if status == -6 else ""
        raise SpanonError("libspanon %s failed (%d): %s %s" % (what, status, msg, extra))


def profile_begin(device=None):
    check(lib().spanon_profile_begin(stream_ptr(device)), "profile_begin")


def profile_end(max_entries=64):
    """-> {kernel name: (launches, total ms)} since profile_begin (synchronises the stream).  Host latency ahead of an entry
    point's first launch is booked by the library under "(host)" and left out here."""
    arr = (KernelTime * max_entries)()
    n = lib().spanon_profile_end(arr, max_entries)
    if n < 0:
        check(n, "profile_end")
    return {arr[i].name.decode(): (int(arr[i].launches), float(arr[i].ms_total)) for i in range(n)
            if not arr[i].name.decode().startswith("(")}


def ptr(t):
    """Device (or host) pointer of a tensor, None -> NULL."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t, name, dtype=torch.float32):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise SpanonError("%s is on %s: speech_anonymization_b200 only runs on a CUDA (sm_100) device, "
                          "there is no CPU path" % (name, t.device))
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
