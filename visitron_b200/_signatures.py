"""argtypes / restype for every symbol declared in include/visitron_b200.h."""
import ctypes as C

P, I, I64, F, U64, U32, D = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64, C.c_uint32, C.c_double

SIGNATURES = {
    "vt_init": (I, [I]),
    "vt_num_sms": (I, []),
    "vt_launch_count": (I64, []),
    "vt_reset_launch_count": (None, []),
    "vt_debug_force_block_n": (None, [I]),
    "vt_reserve_sms": (None, [I]),
    "vt_set_seed_source": (None, [P]),
    "vt_seed_advance": (I, [P, P]),
    "vt_set_row_limit": (None, [P]),
    "vt_compact_labels": (I, [P, I64, P, I64, I64, I64, I64, P, P, P, P, P, P]),
    "vt_pretrain_finalize_counted": (I, [P, P, P, P, P, P, P, P, P, I64, P, P, P]),
    "vt_gemm_bf16": (I, [P, I64, I, P, I64, I, P, I64, I64, I64, I64, P, P]),
    "vt_layernorm_fwd": (I, [P, I64, P, P, F, I64, I64, P, I64, F, U64, U32, I64, I64, I64, P]),
    "vt_row_partials_rows": (I64, [I64]),
    "vt_layernorm_bwd": (I, [P, I64, P, I64, P, F, I64, I64, P, I64, P, I64, P, P, P, P, F, U32, F, U32, U64, I64, I64, I64, P]),
    "vt_embed_ln_fwd": (I, [P, P, P, P, P, P, P, P, F, I64, I64, I64, I64, P, F, U64, U32, P]),
    "vt_embed_ln_bwd": (I, [P, P, P, P, P, P, P, P, F, I64, I64, I64, I64, I64, P, P, P, P, P, P, F, U64, U32, P]),
    "vt_colsum_partials_rows": (I64, [I64]),
    "vt_colsum_bf16": (I, [P, I64, I64, I64, P, P, I, P]),
    "vt_cast_f32_bf16": (I, [P, I64, I64, I64, P, I64, P, I64, P]),
    "vt_pad_cast_concat": (I, [P, I64, P, I64, I64, P, I64, P]),
    "vt_pad_cast_concat_mixed": (I, [P, I, I64, P, I, I64, I64, P, I64, P]),
    "vt_dropout_rows": (I, [P, I64, I64, I64, P, I64, F, U64, U32, I64, I64, I64, P]),
    "vt_gather_rows": (I, [P, I64, P, I64, I64, P, I64, P]),
    "vt_scatter_add_rows": (I, [P, I64, P, I64, I64, P, I64, P]),
    "vt_mask_bias": (I, [P, I, I64, P, P]),
    "vt_attention_key_tiles": (I, [P, I64, I64, P, P]),
    "vt_attention_fwd": (I, [P, I64, P, P, I64, I64, I64, I64, P, P, F, U64, U32, P]),
    "vt_attention_bwd_workspace": (I64, [I64, I64, I64]),
    "vt_attention_bwd": (I, [P, I64, P, P, P, P, P, P, I64, I64, I64, I64, P, F, U64, U32, P]),
    "vt_softmax_ce_fwd": (I, [P, I64, I64, I64, P, I, P, P, P, P]),
    "vt_softmax_ce_bwd": (I, [P, I64, I64, I64, P, I, P, F, P, P, I64, P]),
    "vt_action_head_fwd": (I, [P, P, P, I64, I64, I64, P, P, P, P, P]),
    "vt_action_head_bwd": (I, [P, P, I64, I64, I64, P, P, P, P, P, P, P, P]),
    "vt_pretrain_finalize": (I, [P, P, P, I64, P, P, P, I64, P, P, P, I64, P, P]),
    "vt_tanh_bwd": (I, [P, P, I64, P, P]),
    "vt_gelu_bwd": (I, [P, P, I64, P, P]),
    "vt_cast_bf16_f32": (I, [P, I64, P, P]),
    "vt_cast_f32_bf16_add": (I, [P, P, I64, P, P]),
    "vt_small_linear_fwd": (I, [P, I64, I64, P, I64, I64, P, I64, P, I, P, I64, I64, I64, P]),
    "vt_small_linear_bwd": (I, [P, I64, P, I64, I, P, I64, I64, P, I64, I64, P, I64, P, I64, P, I64, P, I64, P, I64, I64, P]),
    "vt_softdot_splits": (I64, [I64, I64]),
    "vt_softdot_fwd": (I, [P, P, P, I64, I64, I64, P, P, P, P, P]),
    "vt_softdot_bwd": (I, [P, P, P, P, P, P, P, I64, I64, I64, P, P, P, P]),
    "vt_head_ce_fwd": (I, [P, I64, I64, I64, P, P, P, I64, P, I64, P, P, P, P]),
    "vt_head_ce_bwd": (I, [P, I64, I64, I64, P, I64, P, P, P, P, I64, P]),
    "vt_bce_logits_fwd": (I, [P, P, P, I64, F, P, P, P]),
    "vt_bce_logits_bwd": (I, [P, P, P, I64, F, P, P, P, P]),
    "vt_lstm_seq_fwd": (I, [P, P, P, I64, I64, I64, P, P, P, P, P, P]),
    "vt_lstm_seq_bwd": (I, [P, P, P, P, P, P, P, I64, I64, I64, P, P]),
    "vt_lstm_cell_fwd": (I, [P, P, P, I64, I64, P, P, P, P]),
    "vt_lstm_cell_bwd": (I, [P, P, P, P, P, I64, I64, P, P, P]),
    "vt_dropout_f32": (I, [P, I64, F, U64, U32, P, P]),
    "vt_opt_chunk": (I64, []),
    "vt_adamw_step": (I, [P, I64, I64, I64, P, P, D, D, D, P, P]),
    "vt_clip_grad_norm": (I, [P, I64, I64, F, P, P, I, P]),
    "vt_mlm_mask": (I, [P, P, P, I64, I64, I64, I64, F, U64, P, P, P, P]),
    "vt_mlm_draws": (I, [I64, I64, U64, P, P, P]),
    "vt_pack_regions": (I, [P, I64, P, I64, P, I64, I64, P, P, P, I64, P]),
}


def declare(lib):
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
