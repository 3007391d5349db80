"""ctypes binding of libnes_b200.so (include/nes_b200.h).  There is no CPU fallback: if the library
is missing, or no B200 is visible when a context is created, the call raises."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libnes_b200.so')

c_i32, c_i64, c_u64, c_f32, c_f64, c_vp = (ctypes.c_int32, ctypes.c_int64, ctypes.c_uint64, ctypes.c_float,
                                           ctypes.c_double, ctypes.c_void_p)


class FcDims(ctypes.Structure):
    _fields_ = [('vocab_size', c_i32), ('input_encoding_size', c_i32), ('rnn_size', c_i32), ('fc_feat_size', c_i32),
                ('seq_length', c_i32), ('layer_n', c_i32), ('layer_n_affine', c_i32), ('vbn_e', c_i32), ('vbn_e_affine', c_i32)]


# name -> (restype, argtypes); every symbol include/nes_b200.h declares
SIGNATURES = {
    'nes_last_error': (ctypes.c_char_p, []),
    'nes_version': (c_i32, []),
    'nes_ctx_create_fc': (c_i32, [ctypes.POINTER(c_vp), c_i32, ctypes.POINTER(FcDims), c_u64, c_i64, c_i32, c_i32, c_i32]),
    'nes_ctx_create_flat': (c_i32, [ctypes.POINTER(c_vp), c_i32, ctypes.POINTER(c_i64), c_i32, c_i64, c_u64, c_i64, c_i32]),
    'nes_ctx_destroy': (c_i32, [c_vp]),
    'nes_cider_set_tables': (c_i32, [c_vp, c_vp, c_vp, c_i64, ctypes.c_double, c_vp, c_i32, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    'nes_shaped_fitness': (c_i32, [c_vp, c_i32, c_i64, c_i32, c_vp, c_vp]),
    'nes_rollout_sample': (c_i32, [c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_f32, c_vp, c_vp, c_vp]),
    'nes_set_sensitivity': (c_i32, [c_vp, c_i32, c_vp]),
    'nes_mnist_rollout': (c_i32, [c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_f32, c_vp, c_vp, c_vp]),
    'nes_evolve_batch': (c_i32, [c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    'nes_set_mutation': (c_i32, [c_vp, c_i32]),
    'nes_cider_scores': (c_i32, [c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp]),
    'nes_set_stream': (c_i32, [c_vp, c_vp]),
    'nes_sync': (c_i32, [c_vp]),
    'nes_num_params': (c_i64, [c_vp]),
    'nes_padded_params': (c_i64, [c_vp]),
    'nes_granule': (c_i64, [c_vp]),
    'nes_table_len': (c_i64, [c_vp]),
    'nes_gpu_launches': (c_i64, [c_vp]),
    'nes_fill_table': (c_i32, [c_vp]),
    'nes_table_slice': (c_i32, [c_vp, c_i64, c_i64, c_vp]),
    'nes_sample_offsets': (c_i32, [c_vp, c_u64, c_i64, c_vp]),
    'nes_sample_parents': (c_i32, [c_vp, c_u64, c_i64, c_i32, c_vp]),
    'nes_sample_parents_tournament': (c_i32, [c_vp, c_u64, c_i64, c_i32, c_i32, c_vp]),
    'nes_es_select_evolve': (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_f32, c_i32, c_vp]),
    'nes_copy_parent': (c_i32, [c_vp, c_i32, c_i32]),
    'nes_set_theta': (c_i32, [c_vp, c_i32, c_vp, c_i32]),
    'nes_get_theta': (c_i32, [c_vp, c_i32, c_vp, c_i32]),
    'nes_evolve': (c_i32, [c_vp, c_i32, c_i64, c_f32, c_f32, c_vp, c_vp, c_i32]),
    'nes_rollout': (c_i32, [c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_f32, c_i32, c_vp, c_vp]),
    'nes_rollout_trace': (c_i32, [c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_f32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    'nes_logprob_fitness': (c_i32, [c_vp, c_i64, c_i32, c_vp]),
    'nes_rank_weights': (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    'nes_rank_weights_static': (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    'nes_compute_ranks': (c_i32, [c_vp, c_vp, c_i64, c_vp]),
    'nes_gradient': (c_i32, [c_vp, c_vp, c_vp, c_i64, c_f32, c_vp]),
    'nes_gradient_shard': (c_i32, [c_vp, c_vp, c_vp, c_i64, c_f32, c_i32, c_i32, c_vp]),
    'nes_opt_init': (c_i32, [c_vp, c_i32, c_f64, c_f64, c_f64, c_f64]),
    'nes_opt_set_stepsize': (c_i32, [c_vp, c_f64]),
    'nes_opt_update': (c_i32, [c_vp, c_vp, c_f64, ctypes.POINTER(c_f64)]),
    'nes_opt_update_direct': (c_i32, [c_vp, c_vp, c_i32, ctypes.POINTER(c_f64)]),
    'nes_opt_get_state': (c_i32, [c_vp, c_vp, c_vp, ctypes.POINTER(c_i64)]),
    'nes_opt_set_state': (c_i32, [c_vp, c_vp, c_vp, c_i64]),
    'nes_master_step': (c_i32, [c_vp, c_vp, c_vp, c_i64, c_f32, c_f64, ctypes.POINTER(c_f64)]),
    'nes_last_gradient_ms': (c_i32, [c_vp, ctypes.POINTER(c_f32)]),
    'nes_last_rollout_ms': (c_i32, [c_vp, ctypes.POINTER(c_f32)]),
    'nes_profile_enable': (c_i32, [c_vp, c_i32]),
    'nes_profile_read': (c_i32, [c_vp, ctypes.POINTER(c_f32), ctypes.POINTER(c_i32)]),
    'nes_hybrid_stats': (c_i32, [c_vp, ctypes.POINTER(c_i64), c_i32]),
    'nes_gemm_selftest': (c_i32, [c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_f32, c_i32, c_vp]),
}

_lib = None


class NesError(RuntimeError):
    pass


def load():
    """dlopen the in-tree library and bind every prototype.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NesError('libnes_b200.so is not built: run `python -c "import __graft_entry__ as g; g.build()"` '
                           '(there is no CPU fallback)')
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise NesError(load().nes_last_error().decode('utf-8', 'replace'))


def ptr(x):
    """raw address of a torch tensor / numpy array / None (host or device, the C side sorts it out)"""
    if x is None:
        return None
    if hasattr(x, 'data_ptr'):
        assert x.is_contiguous(), 'tensor must be contiguous'
        return c_vp(x.data_ptr())
    assert x.flags['C_CONTIGUOUS'], 'array must be C-contiguous'
    return c_vp(x.ctypes.data)
