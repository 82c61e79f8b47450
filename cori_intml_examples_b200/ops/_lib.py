"""ctypes binding of libcib200.so (C ABI in include/cib200.h).

The library replaces what the reference reaches through Keras/TensorFlow (MKL-DNN / cuDNN / cuBLAS) and Horovod (MPI):
SURVEY 2.7 lists every implied native op (K1-K12) and collective (C1-C5) with its reference call site.

The library is built in-tree by ``__graft_entry__.build()`` / ``scripts/build.sh`` with
``nvcc -gencode arch=compute_100a,code=sm_100a``.  On a machine with a GPU a missing library is a hard
error: the device path never silently falls back to PyTorch ops.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libcib200.so")


class CibBlock(C.Structure):
    _fields_ = [("kind", C.c_int), ("path", C.c_int),
                ("Hi", C.c_int), ("Wi", C.c_int), ("Ci", C.c_int),
                ("Ho", C.c_int), ("Wo", C.c_int), ("Co", C.c_int),
                ("ksize", C.c_int), ("stride", C.c_int), ("pad", C.c_int), ("pool", C.c_int),
                ("K", C.c_int), ("N", C.c_int), ("head_kind", C.c_int), ("drop_rate", C.c_float),
                ("w_off", C.c_longlong), ("b_off", C.c_longlong)]


class CibHyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("rho", C.c_float), ("momentum", C.c_float), ("schedule_decay", C.c_float), ("decay", C.c_float),
                ("m_schedule", C.c_float), ("iterations", C.c_uint), ("seed", C.c_uint)]


class CibMetrics(C.Structure):
    _fields_ = [("loss_sum", C.c_float), ("correct_sum", C.c_float), ("sample_sum", C.c_float),
                ("step_loss", C.c_float), ("step_acc", C.c_float), ("iterations", C.c_uint),
                ("m_schedule", C.c_float), ("batch_cursor", C.c_uint)]


class CibError(RuntimeError):
    pass


_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CibError("libcib200.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "or scripts/build.sh (needs nvcc with sm_100a support)")
    L = C.CDLL(LIB_PATH)
    L.cib_last_error.restype = C.c_char_p
    vp, i, ll, f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    sig = {
        "cib_init": [i], "cib_ktrace_enable": [i], "cib_ktrace_read": [vp], "cib_set_conv_version": [i], "cib_version": [], "cib_exclusive_begin": [], "cib_exclusive_end": [],
        "cib_device_error": [C.POINTER(C.c_uint)],
        "cib_engine_create": [C.POINTER(vp), C.POINTER(CibBlock), i, i, i, i, i, ll, i, i],
        "cib_engine_destroy": [vp],
        "cib_engine_upload": [vp, i, vp, ll], "cib_engine_download": [vp, i, vp, ll],
        "cib_engine_set_hyper": [vp, C.POINTER(CibHyper)], "cib_engine_set_lr": [vp, f],
        "cib_engine_get_metrics": [vp, C.POINTER(CibMetrics)], "cib_engine_reset_metrics": [vp],
        "cib_engine_set_dataset": [vp, vp, vp, i, i], "cib_engine_set_perm": [vp, vp, i],
        "cib_engine_train_steps": [vp, i], "cib_engine_time_steps": [vp, i, C.POINTER(C.c_float)],
        "cib_engine_time_steps_aligned": [vp, i, i, C.POINTER(C.c_float)], "cib_engine_set_wrap": [vp, i], "cib_engine_eval_steps": [vp, i, vp],
        "cib_engine_train_batch_host": [vp, vp, vp, i, vp],
        "cib_engine_sync": [vp], "cib_engine_loss_ring": [vp, C.POINTER(C.POINTER(C.c_float)), C.POINTER(i)], "cib_engine_launch_count": [vp, C.POINTER(i), C.POINTER(i)], "cib_engine_check_canaries": [vp, C.POINTER(ll)],
        "cib_engine_buffer": [vp, i, C.POINTER(vp), C.POINTER(ll)],
        "cib_engine_forward_backward": [vp, i, i],
        "cib_engine_set_stream_priority": [vp, i], "cib_engine_get_stream_priority": [vp, C.POINTER(i), C.POINTER(i), C.POINTER(i)],
        "cib_comm_create_local": [C.POINTER(vp), C.POINTER(vp), i],
        "cib_engine_ipc_handle": [vp, vp], "cib_comm_create_ipc": [C.POINTER(vp), vp, i, i, vp],
        "cib_comm_destroy": [vp], "cib_engine_attach_comm": [vp, vp, i, i],
        "cib_engine_broadcast_state": [vp, i], "cib_engine_gather_slots": [vp], "cib_engine_ar_trace": [vp, i, vp],
        "cib_engine_nvls_step": [vp, i, i, i, C.POINTER(i)], "cib_engine_nvls_active": [vp],
        "cib_set_ar_pull": [i], "cib_set_pdl": [i], "cib_set_ar_ll": [i], "cib_set_par_ctas": [i], "cib_set_peer_timeout_s": [C.c_double],
        "cib_symbuf_create": [C.POINTER(vp), i, ll], "cib_symbuf_ipc_handle": [vp, vp],
        "cib_symbuf_attach_ipc": [vp, i, i, vp], "cib_symbuf_attach_local": [C.POINTER(vp), i],
        "cib_symbuf_ptr": [vp, C.POINTER(vp)], "cib_symbuf_allreduce": [vp, ll, i, i, vp],
        "cib_symbuf_broadcast": [vp, ll, i, vp], "cib_symbuf_destroy": [vp],
        "cib_nvls_supported": [i, C.POINTER(i)], "cib_nvls_create_local": [C.POINTER(vp), i, ll], "cib_nvls_ptr": [vp, i, C.POINTER(vp)],
        "cib_nvls_allreduce": [vp, i, ll, i, vp], "cib_nvls_destroy": [vp],
        "cib_gemm_tf32": [vp, vp, vp, i, i, i, i, i, i, i, vp],
        "cib_dense_relu_fwd": [vp, vp, vp, vp, vp, i, i, i, vp],
        "cib_dense_relu_fwd_noalloc": [vp, vp, vp, vp, i, i, i, vp],
        "cib_conv3x3_fwd_tf32": [vp, vp, vp, vp, i, i, i, i, i, i, i, vp],
        "cib_conv3x3_dgrad_tf32": [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, f, vp],
        "cib_conv3x3_wgrad_tf32": [vp, vp, vp, vp, i, i, i, i, i, i, vp],
        "cib_conv_fwd_direct": [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, vp],
        "cib_set_conv1_tc": [i], "cib_set_early_update": [i], "cib_set_head_version": [i],
        "cib_conv1_fwd_tc": [vp, vp, vp, vp, i, i, i, i, i, i, i, vp],
        "cib_conv1_wgrad_tc": [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, vp],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = L
    return L


def check(rc):
    """Raise CibError with the library's message when a call failed."""
    if rc != 0:
        raise CibError(lib().cib_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device/host pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def current_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class exclusive(object):
    """``with exclusive():`` -- the library's process-wide exclusive section (include/cib200.h): device-wide synchronisation
    and device allocation from Python must not overlap with another trial thread's CUDA-graph capture."""

    def __enter__(self):
        lib().cib_exclusive_begin()
        return self

    def __exit__(self, *a):
        lib().cib_exclusive_end()
        return False
