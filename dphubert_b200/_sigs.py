"""ctypes signatures of every exported C-ABI entry point (kept next to include/dphubert_b200.h)."""
import ctypes as C

vp = C.c_void_p
i32 = C.c_int
i64 = C.c_int64
u64 = C.c_uint64
f32 = C.c_float

# name -> argtypes (restype is int unless listed in RESTYPES)
SIGS = {
    "dph_launch_count": [],
    "dph_launch_count_reset": [],
    "dph_abi_version": [],
    "dph_last_error": [],
    "dph_gemm_bf16": [vp, vp],
    "dph_gemm_f32": [vp, vp],
}

RESTYPES = {
    "dph_launch_count": C.c_longlong,
    "dph_launch_count_reset": None,
    "dph_last_error": C.c_char_p,
}


def declare(lib):
    for name, args in SIGS.items():
        fn = getattr(lib, name)  # raises AttributeError if the symbol is missing
        fn.argtypes = args
        fn.restype = RESTYPES.get(name, C.c_int)
