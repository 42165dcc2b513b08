"""ctypes prototypes for every entry point in ``include/b200dp.h``."""
import ctypes as C

P = C.c_void_p
I = C.c_int
LL = C.c_longlong
SZ = C.c_size_t
F = C.c_float
D = C.c_double
U64 = C.c_uint64
PLL = C.POINTER(C.c_longlong)
PI = C.POINTER(C.c_int)

_SIGS = {
    "b200_last_error": (C.c_char_p, []),
    "b200_comm_pad_bytes": (SZ, []),
    "b200_comm_create": (I, [I, I, I, C.POINTER(P)]),
    "b200_comm_rank": (I, [P]),
    "b200_comm_nranks": (I, [P]),
    "b200_comm_is_emulated": (I, [P]),
    "b200_comm_set_timeout_ms": (I, [P, D]),
    "b200_comm_set_step_mirror": (I, [P, I, I, SZ, SZ, I, P, P]),
    "b200_comm_set_inbox": (I, [P, I, I, SZ]),
    "b200_comm_set_coef_cache": (I, [P, I, SZ]),
    "b200_comm_set_debug_skew_ns": (I, [P, C.c_uint]),
    "b200_comm_set_debug_poison": (I, [P, I]),
    "b200_comm_seg_alloc": (I, [P, SZ, PI, P]),
    "b200_comm_seg_open": (I, [P, I, P]),
    "b200_comm_seg_ptr": (P, [P, I, I]),
    "b200_comm_destroy": (I, [P]),
    "b200_fused_allreduce_adam": (I, [P, I, I, I, I, I, SZ, SZ, I, SZ, SZ, SZ, I, SZ, I, SZ, SZ, F, I, PLL, PLL, PI, PI, P]),
    "b200_allreduce": (I, [P, I, I, I, I, I, SZ, I, SZ, LL, F, P]),
    "b200_broadcast": (I, [P, I, I, I, I, I, SZ, I, SZ, LL, P]),
    "b200_allgather_slices": (I, [P, I, I, I, I, SZ, LL, P]),
    "b200_comm_pull_from_peer": (I, [P, I, SZ, SZ, I, P]),
    "b200_allgather_acts": (I, [P, I, I, I, SZ, I, SZ, I, SZ, SZ, I, I, I, P]),
    "b200_gemm_tc_wgrad_adam": (I, [P, P, P, I, I, I, I, I, I, SZ, SZ, SZ, I, SZ, I, SZ, SZ, LL, I, I, F, I, P]),
    "b200_nvls_supported": (I, []),
    "b200_comm_mcseg_alloc": (I, [P, SZ, PI, PI, PI, C.POINTER(SZ)]),
    "b200_comm_mcseg_open": (I, [P, I, PI, I]),
    "b200_comm_mcseg_bind": (I, [P, I]),
    "b200_comm_seg_mc_ptr": (P, [P, I]),
    "b200_allreduce_nvls": (I, [P, I, I, I, I, SZ, I, SZ, LL, F, P]),
    "b200_broadcast_nvls": (I, [P, I, I, I, I, I, SZ, I, SZ, LL, P]),
    "b200_fused_allreduce_adam_nvls": (I, [P, I, I, I, I, SZ, SZ, I, SZ, SZ, SZ, I, SZ, I, SZ, SZ, F, LL, LL, I, P]),
    "b200_adasum_scratch_bytes": (SZ, [I, I]),
    "b200_adasum_level": (I, [P, I, I, I, I, SZ, I, SZ, I, SZ, I, PLL, PLL, P]),
    "b200_adasum_pair_static": (I, [P, P, P, I, I, PLL, PLL, P]),
    "b200_conv1_pool_fwd": (I, [P, P, P, P, P, I, I, P, I, P, P, I, P]),
    "b200_conv1_wgrad": (I, [P, P, P, P, P, P, I, I, P]),
    "b200_head_fwd_bwd": (I, [P, P, P, P, P, P, P, P, P, P, P, I, I, I, F, I, U64, P, I, I, I, P]),
    "b200_head_wgrad": (I, [P, P, P, P, P, P, I, I, I, P]),
    "b200_unpool_relu_bwd": (I, [P, P, P, P, P, I, I, I, I, P]),
    "b200_maxpool_fwd": (I, [P, P, P, I, I, I, I, P]),
    "b200_unpool_relu_bwd_scaled": (I, [P, P, P, P, P, I, I, I, I, F, P]),
    "b200_conv1_valid_fwd": (I, [P, P, P, P, I, I, I, P, I, P, P, I, P]),
    "b200_conv1_valid_wgrad": (I, [P, P, P, P, P, I, I, I, P]),
    "b200_dropout_bf16": (I, [P, P, LL, F, U64, P, I, P]),
    "b200_conv_direct_fwd": (I, [P, P, P, P, I, I, I, I, I, I, I, I, P]),
    "b200_conv_direct_dgrad": (I, [P, P, P, I, I, I, I, I, I, I, P]),
    "b200_conv_direct_wgrad": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "b200_gemm_simple": (I, [P, P, P, I, I, I, I, I, I, I, P]),
    "b200_cast_f32_to_bf16": (I, [P, P, LL, P]),
    "b200_adam_local": (I, [P, P, P, P, P, LL, I, D, D, D, D, I, F, P]),
    "b200_bias_relu_bf16": (I, [P, P, P, I, I, I, P]),
    "b200_gemm_tc": (I, [P, P, P, I, I, I, I, I, I, I, I, I, I, I, P]),
    "b200_gemm_tc_reduce_scatter": (I, [P, P, P, I, SZ, I, I, I, I, I, I, I, I, I, I, P]),
    "b200_conv_tc_fwd": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, P]),
    "b200_conv_tc_dgrad": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "b200_conv_tc_wgrad": (I, [P, P, P, P, I, I, I, I, I, I, P]),
    "b200_conv_tc_fwd_pad": (I, [P, P, P, P, P, I, I, I, I, I, I, I, I, I, P]),
    "b200_conv_tc_dgrad_pad": (I, [P, P, P, I, I, I, I, I, I, I, P]),
    "b200_conv_tc_wgrad_pad": (I, [P, P, P, P, I, I, I, I, I, I, I, P]),
    "b200_loader_create": (I, [P, P, LL, SZ, I, I, I, U64, C.POINTER(P), C.POINTER(P), C.POINTER(P)]),
    "b200_loader_next": (I, [P, PI, PLL]),
    "b200_loader_release": (I, [P, I]),
    "b200_loader_batches_per_epoch": (LL, [P]),
    "b200_loader_destroy": (I, [P]),
    "b200_topology_json": (I, [C.c_char_p, I]),
    "b200_l2_flush": (I, [P, SZ, P]),
    "b200_stream_l2_persist": (I, [P, SZ, F, P]),
    "b200_set_pdl": (I, [I]),
    "b200_set_stable_weights": (I, [I]),
}


def declare(lib) -> None:
    for name, (res, args) in _SIGS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.restype = res
        fn.argtypes = args
