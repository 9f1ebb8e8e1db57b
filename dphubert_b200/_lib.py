"""ctypes binding of the C-ABI kernel library (`include/dphubert_b200.h`).

There is no CPU fallback: if the shared library is missing this module raises at first use, and every
wrapper refuses non-CUDA tensors.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdphubert_b200.so")

_lib = None


class DphError(RuntimeError):
    pass


class Operand(ctypes.Structure):
    _fields_ = [
        ("ptr", ctypes.c_void_p),
        ("major_mn", ctypes.c_int),
        ("dims", ctypes.c_uint64 * 5),
        ("strides", ctypes.c_uint64 * 4),
        ("b0_mul", ctypes.c_int),
        ("b1_mul", ctypes.c_int),
    ]


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("a", Operand),
        ("b", Operand),
        ("M", ctypes.c_int),
        ("N", ctypes.c_int),
        ("K", ctypes.c_int),
        ("nb0", ctypes.c_int),
        ("nb1", ctypes.c_int),
        ("block_n", ctypes.c_int),
        ("ksplit", ctypes.c_int),
        ("c", ctypes.c_void_p),
        ("c2", ctypes.c_void_p),
        ("ldc", ctypes.c_int64),
        ("c_sb0", ctypes.c_int64),
        ("c_sb1", ctypes.c_int64),
        ("c_dtype", ctypes.c_int),
        ("c_mode", ctypes.c_int),
        ("alpha", ctypes.c_float),
        ("bias", ctypes.c_void_p),
        ("act", ctypes.c_int),
        ("colmask", ctypes.c_void_p),
        ("scalar_mask", ctypes.c_void_p),
        ("residual", ctypes.c_void_p),
        ("ldr", ctypes.c_int64),
        ("r_sb0", ctypes.c_int64),
        ("r_sb1", ctypes.c_int64),
        ("r_dtype", ctypes.c_int),
        ("drop_p", ctypes.c_float),
        ("seed", ctypes.c_uint64),
        ("vec_sb0", ctypes.c_int64),
        ("scalar_sb0", ctypes.c_int64),
        ("cta_group", ctypes.c_int),
        ("act_bwd", ctypes.c_int),
        ("z", ctypes.c_void_p),
        ("ldz", ctypes.c_int64),
        ("z_sb0", ctypes.c_int64),
        ("z_sb1", ctypes.c_int64),
        ("dbias", ctypes.c_void_p),
        ("dmask", ctypes.c_void_p),
        ("b_nsplit", ctypes.c_int),
    ]


class OperandF32(ctypes.Structure):
    _fields_ = [
        ("ptr", ctypes.c_void_p),
        ("rs", ctypes.c_int64),
        ("cs", ctypes.c_int64),
        ("kxs", ctypes.c_int64),
        ("kin", ctypes.c_int),
        ("bs0", ctypes.c_int64),
        ("bs1", ctypes.c_int64),
    ]


class GemmF32Desc(ctypes.Structure):
    _fields_ = [
        ("a", OperandF32),
        ("b", OperandF32),
        ("M", ctypes.c_int),
        ("N", ctypes.c_int),
        ("K", ctypes.c_int),
        ("nb0", ctypes.c_int),
        ("nb1", ctypes.c_int),
        ("c", ctypes.c_void_p),
        ("c2", ctypes.c_void_p),
        ("ldc", ctypes.c_int64),
        ("c_sb0", ctypes.c_int64),
        ("c_sb1", ctypes.c_int64),
        ("accumulate", ctypes.c_int),
        ("alpha", ctypes.c_float),
        ("bias", ctypes.c_void_p),
        ("act", ctypes.c_int),
        ("colmask", ctypes.c_void_p),
        ("scalar_mask", ctypes.c_void_p),
        ("residual", ctypes.c_void_p),
        ("ldr", ctypes.c_int64),
        ("r_sb0", ctypes.c_int64),
        ("r_sb1", ctypes.c_int64),
        ("drop_p", ctypes.c_float),
        ("seed", ctypes.c_uint64),
        ("vec_sb0", ctypes.c_int64),
        ("scalar_sb0", ctypes.c_int64),
    ]


def lib():
    """Load (once) and return the ctypes handle. Raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DphError(
                "dphubert_b200: %s not found — run `python -m dphubert_b200._build` (there is no CPU fallback)" % LIB_PATH
            )
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.dph_last_error.restype = ctypes.c_char_p
        _lib.dph_launch_count.restype = ctypes.c_longlong
        from . import _sigs

        _sigs.declare(_lib)
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise DphError("%s failed (%d): %s" % (what, rc, lib().dph_last_error().decode()))


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr():
    """torch's current CUDA stream as a raw handle.  Called once per kernel launch: the raw getter avoids building a
    `torch.cuda.Stream` object each time (that alone was ~30 % of the host time of an eager training step)."""
    if _raw_stream is not None:
        return ctypes.c_void_p(_raw_stream(torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device pointer of a CUDA tensor (or NULL for None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise DphError("dphubert_b200 kernels need CUDA tensors (no CPU fallback); got a %s tensor" % t.device)
    return ctypes.c_void_p(t.data_ptr())


def launch_count():
    return int(lib().dph_launch_count())
