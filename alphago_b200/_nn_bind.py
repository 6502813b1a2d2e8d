"""ctypes signatures of the network entry points (stage 2 of include/alphago_b200.h)."""
import ctypes as C


def bind(L):
    vp, i = C.c_void_p, C.c_int
    L.ag_net_create.restype = vp
    L.ag_net_create.argtypes = [i, i, i, vp, vp, i, i]
    L.ag_net_destroy.argtypes = [vp]; L.ag_net_destroy.restype = None
    L.ag_net_layer_info.argtypes = [vp, i, i]
    L.ag_net_set_conv.argtypes = [vp, i, vp, vp]
    L.ag_net_get_conv.argtypes = [vp, i, vp, vp]
    L.ag_net_set_head.argtypes = [vp, i, vp, vp]
    L.ag_net_get_head.argtypes = [vp, i, vp, vp]
    L.ag_net_forward.argtypes = [vp, vp, i, vp, vp]
    L.ag_net_forward_loss.argtypes = [vp, vp, vp, i, vp, vp, vp]
    L.ag_net_backward_layer.argtypes = [vp, i, i, vp]
    L.ag_net_backward_range.argtypes = [vp, i, i, i, vp]
    L.ag_net_set_sm_budget.argtypes = [vp, i]
    L.ag_net_set_loss.argtypes = [vp, i]
    L.ag_net_wgrad_done.argtypes = [vp, i, C.POINTER(vp), C.POINTER(C.c_uint)]
    L.ag_stream_wait_value.argtypes = [vp, vp, C.c_uint]
    L.ag_net_wait_first_layer.argtypes = [vp, vp]
    L.ag_net_grads.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_longlong), vp]
    L.ag_net_sgd.argtypes = [vp, C.c_float, C.c_float, C.c_float, vp]
    L.ag_net_train_step.argtypes = [vp, vp, vp, i, C.c_float, C.c_float, vp, vp]
    L.ag_net_debug_read.argtypes = [vp, i, i, i, vp]
    L.ag_net_run_layer.argtypes = [vp, i, i, i, vp]
    L.ag_debug_hold_sms.argtypes = [i, C.c_longlong, vp]
    L.ag_wgrad_scratch_bytes.restype = C.c_size_t
    L.ag_wgrad_scratch_bytes.argtypes = [i, i, i, C.c_longlong]
    L.ag_wgrad_partial_bytes.restype = C.c_size_t
    L.ag_wgrad_partial_bytes.argtypes = [i, i, i]
    L.ag_wgrad.argtypes = [vp, vp, C.c_longlong, i, i, i, i, vp, C.c_size_t, vp, vp]
    L.ag_wgrad_static_max_bytes.restype = C.c_size_t
    L.ag_wgrad_static.argtypes = [vp, vp, C.c_longlong, i, i, i, i, vp, vp]
    L.ag_xchg_create.restype = vp
    L.ag_xchg_create.argtypes = [i, i, vp, C.c_longlong, i, vp]
    L.ag_xchg_connect.argtypes = [vp, vp]
    L.ag_xchg_allreduce.argtypes = [vp, C.c_longlong, C.c_longlong, vp]
    L.ag_xchg_wait_readers.argtypes = [vp, vp]
    L.ag_xchg_destroy.argtypes = [vp]; L.ag_xchg_destroy.restype = None
    L.ag_selfplay_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, C.c_longlong, vp, C.c_float, i, i, vp, vp]
    L.ag_selfplay_winner.argtypes = [vp, vp, i, i, C.c_float, vp, vp]
    L.ag_augment.argtypes = [vp, vp, vp, vp, i, i, i, vp, vp, vp]
    return L
