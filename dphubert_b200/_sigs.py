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
    "dph_rng_set": [u64, vp],
    "dph_rng_advance": [vp],
    "dph_rng_get": [vp, vp],
    "dph_gemm_bf16": [vp, vp],
    "dph_gemm_f32": [vp, vp],
    # norm.cu
    "dph_layernorm_fwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, i32, f32, u64, i32, vp],
    "dph_layernorm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, u64, i32, i64, i32, i32,
                          vp],
    # elementwise.cu
    "dph_epilogue_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, u64, i32, i64, i32, i64, i64, i32, vp],
    "dph_pad_rows_zero": [vp, vp, i32, i32, i32, i32, vp],
    "dph_posconv_pad": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dph_posconv_unpad": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dph_posconv_tc": [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dph_conv_out_lengths": [vp, vp, i32, i32, vp, vp, vp],
    "dph_cast": [vp, i32, vp, i32, i64, vp],
    "dph_vec_mul": [vp, vp, vp, f32, f32, i32, vp],
    "dph_transpose": [vp, i32, vp, i32, i32, i32, i32, vp],
    "dph_memset": [vp, i32, i64, vp],
    "dph_memset2d": [vp, i64, i32, i64, i64, vp],
    "dph_axpy": [vp, vp, i64, i32, vp],
    "dph_group_mask_bwd": [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp],
    # conv0.cu
    "dph_conv0_stats": [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, vp],
    "dph_conv0_fwd": [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "dph_conv0_bwd": [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32,
                      i32, i32, vp],
    "dph_conv0_bwd_static": [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32,
                             i32, i32, i32, vp],
    "dph_wave_norm": [vp, i64, vp, i64, vp, i32, i32, f32, vp],
    # collate.cu
    "dph_collate_crop": [vp, i32, vp, vp, vp, vp, i32, i32, vp],
    # softmax.cu
    "dph_softmax_fwd": [vp, i64, vp, vp, i64, vp, i32, i32, i32, f32, u64, vp, vp, vp, vp, i32, i32, vp],
    "dph_softmax_bwd": [vp, vp, i64, vp, i64, i32, i32, i32, f32, u64, vp, vp, vp, vp, i32, vp, vp, i32, i32, vp],
    "dph_wavlm_gate_fwd": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "dph_wavlm_gate_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    # attention.cu
    "dph_attn_fwd": [vp, vp, vp, vp, i32, i32, i32, f32, vp, vp, vp, vp, vp, i32, f32, u64, vp, vp],
    "dph_attn_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, vp, vp, vp, vp, vp, i32, f32, u64, vp, vp, vp,
                     i32, vp],
    # hardconcrete.cu
    "dph_hardconcrete_fwd": [vp, vp, vp, vp, i64, f32, f32, f32, f32, u64, vp],
    "dph_hardconcrete_bwd": [vp, vp, vp, vp, i64, f32, f32, f32, vp],
    "dph_l0_norms": [vp, vp, vp, i32, f32, vp, vp],
    "dph_num_params": [vp, vp, i32, vp, vp, vp],
    "dph_lagrangian": [vp, vp, f32, vp, vp, vp],
    "dph_l0_norms_bwd": [vp, vp, vp, vp, i32, f32, vp, vp, vp, vp],
    # loss.cu
    "dph_distill_loss_fwd": [vp, vp, vp, vp, vp, i64, i32, f32, f32, f32, i32, i32, i32, i64, i32, vp],
    "dph_distill_loss_bwd": [vp, vp, vp, vp, vp, i64, i64, i32, f32, f32, f32, i32, i32, vp],
    # prep.cu
    "dph_conv_weight_prep": [vp, vp, vp, i32, i32, i32, vp, i32, vp],
    "dph_conv_wgrad_unpermute": [vp, vp, i32, i32, i32, vp],
    "dph_posconv_weight_prep": [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "dph_posconv_weight_bwd": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    # optim.cu
    "dph_sumsq": [vp, i64, vp, vp],
    "dph_counter_inc": [vp, vp],
    "dph_adamw": [vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, vp, vp, f32, vp, i32, vp],
    "dph_scale": [vp, i64, f32, vp],
    "dph_multimem_allreduce": [vp, i64, i32, i32, f32, i32, vp],
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
