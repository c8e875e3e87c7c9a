"""ctypes binding of libkge_b200.so (include/kge_abi.h).  No CPU fallback: a missing
library or a failing call raises."""
import ctypes as C
import os

KGE_ABI_VERSION = 12
KGE_MAX_TABLES = 4
(UM, SE, TRANSE, TRANSH, TRANSR, TRANSD, ROTATE, DISTMULT, RESCAL) = range(9)
SCORE_LP, SCORE_LP_POW, SCORE_DOT = 0, 1, 2
LOSS_HINGE, LOSS_LOGISTIC, LOSS_BCE, LOSS_SELF_ADV, LOSS_SQUARE = range(5)
SIDE_H, SIDE_T, SIDE_HT = 0, 1, 2
NS_UNIFORM, NS_TYPED = 0, 1
T_ENT, T_ENT_PROJ, T_REL_A, T_REL_B = range(4)
CON_NORMALIZE, CON_CLIP = 0, 1
OPTIM_ADAM, OPTIM_SGD, OPTIM_ADAGRAD = 0, 1, 2
OPT_RADIX_SORT = 1
OPT_NO_SHADOW = 2
OPT_NO_FUSED_PUSH = 3
OPT_RING_SLOTS = 4
OPT_GRAPH_PROFILE = 5
OPT_RANK_TC_PASSES = 6
OPT_CLUSTER_GROUP = 7
OPT_PRE_EARLY = 8
N_PHASES = 10
PHASES = ["constraint", "step", "reduce", "sort", "update", "rank", "rank_gemm", "rank_fix", "apply", "barrier"]
PEER_DEAD = 0xdead


class Table(C.Structure):
    _fields_ = [("w", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("wn", C.c_void_p),
                ("last", C.c_void_p), ("rows", C.c_int64), ("cols", C.c_int32), ("ld", C.c_int32)]


class Tables(C.Structure):
    _fields_ = [("t", Table * KGE_MAX_TABLES), ("wn_valid", C.c_int32 * KGE_MAX_TABLES),
                ("lazy_t", C.c_int64)]


class Cfg(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("model", C.c_int32), ("score", C.c_int32),
                ("loss", C.c_int32), ("sampler", C.c_int32), ("corrupt_side", C.c_int32),
                ("negative_ratio", C.c_int32), ("constraint", C.c_int32), ("id_dtype", C.c_int32),
                ("d_ent", C.c_int32), ("d_rel", C.c_int32), ("optimizer", C.c_int32),
                ("score_p", C.c_float), ("margin", C.c_float), ("temperature", C.c_float),
                ("constraint_weight", C.c_float), ("lr", C.c_float), ("beta1", C.c_float),
                ("beta2", C.c_float), ("eps", C.c_float), ("clip_norm", C.c_float),
                ("rotate_scale", C.c_float), ("n_ent", C.c_int64), ("n_rel", C.c_int64)]


class Sampler(C.Structure):
    _fields_ = [("pool", C.c_void_p), ("pool_len", C.c_int64), ("ent_type", C.c_void_p),
                ("type_off", C.c_void_p), ("type_members", C.c_void_p), ("ent_pos", C.c_void_p),
                ("n_types", C.c_int64)]


class Filter(C.Structure):
    _fields_ = [("off", C.c_void_p), ("ids", C.c_void_p), ("end", C.c_void_p)]


class Dp(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("B_global", C.c_int64), ("i_base", C.c_int64)]


KGE_MAX_PEERS = 8


class Grads(C.Structure):
    _fields_ = [("dense", C.c_void_p * KGE_MAX_TABLES), ("n_peers", C.c_int32),
                ("peer", (C.c_void_p * KGE_MAX_TABLES) * KGE_MAX_PEERS), ("peer_scal", C.c_void_p * KGE_MAX_PEERS),
                ("owner_rank", C.c_int32), ("owner_rows", C.c_int32 * KGE_MAX_TABLES),
                ("w_peer", (C.c_void_p * KGE_MAX_TABLES) * KGE_MAX_PEERS), ("wn_peer", C.c_void_p * KGE_MAX_PEERS),
                ("abort_word", C.c_void_p), ("w_mc", C.c_void_p * KGE_MAX_TABLES), ("wn_mc", C.c_void_p),
                ("w_private", C.c_int32)]


class PeerFlags(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("flags", C.c_void_p * KGE_MAX_PEERS)]


class Shard(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("shard_rows", C.c_int64), ("cap", C.c_int64),
                ("w_peer", (C.c_void_p * 2) * KGE_MAX_PEERS), ("ids_peer", C.c_void_p * KGE_MAX_PEERS),
                ("cnt_peer", C.c_void_p * KGE_MAX_PEERS), ("val_peer", (C.c_void_p * 2) * KGE_MAX_PEERS),
                ("stamp", C.c_void_p), ("idx_of", C.c_void_p), ("work", C.c_void_p), ("contrib", C.c_void_p)]


SYMBOLS = ["kge_abi_version", "kge_last_error", "kge_ctx_create", "kge_ctx_destroy",
           "kge_workspace_bytes", "kge_sample_negatives", "kge_score", "kge_constrain_table",
           "kge_train_step", "kge_train_step_host", "kge_train_grad", "kge_train_apply", "kge_rank", "kge_profile_begin", "kge_profile_end",
           "kge_set_option", "kge_flush_rows",
           "kge_shard_lookup", "kge_shard_pre", "kge_shard_grad", "kge_shard_apply",
           "kge_peer_barrier", "kge_train_step_dp", "kge_rank_lp_static",
           "kge_filter_index_bytes", "kge_filter_index_build", "kge_filter_lookup",
           "kge_train_step_next", "kge_train_step_host_next",
           "kge_step_graph_create", "kge_step_graph_launch", "kge_step_graph_destroy",
           "kge_train_step_dp_next", "kge_step_graph_create_dp", "kge_step_graph_launch_dp", "kge_step_graph_phases", "kge_step_graph_wait_loss",
           "kge_epoch_perm_bytes", "kge_epoch_perm", "kge_batch_gather"]
RANK_STATIC_BYTES = 4 << 20

_LIB = None


def lib_path():
    here = os.path.dirname(os.path.abspath(__file__))
    return os.environ.get("KGE_B200_LIB",
                          os.path.normpath(os.path.join(here, "..", "..", "lib", "libkge_b200.so")))


def load():
    """dlopen the library and declare prototypes. Raises if it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not found: build it with __graft_entry__.build() "
                           "(KGE has no CPU fallback)")
    L = C.CDLL(path)
    vp, i64, u64, i32 = C.c_void_p, C.c_int64, C.c_uint64, C.c_int
    L.kge_abi_version.restype = i32
    L.kge_last_error.restype = C.c_char_p
    L.kge_ctx_create.argtypes = [i32, C.POINTER(vp)]
    L.kge_ctx_destroy.argtypes = [vp]
    L.kge_workspace_bytes.argtypes = [C.POINTER(Cfg), i64, i64, C.POINTER(C.c_size_t)]
    L.kge_sample_negatives.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Sampler), vp, i64, u64, u64, vp, vp]
    L.kge_score.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), vp, vp, vp, i64, vp, vp]
    L.kge_constrain_table.argtypes = [vp, C.POINTER(Tables), i32, i32, vp]
    L.kge_train_step.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), vp, i64,
                                 u64, u64, u64, vp, i32, vp, vp, vp, C.c_size_t, vp]
    L.kge_train_step_host.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), vp, vp, i64,
                                      u64, u64, u64, i32, vp, vp, vp, C.c_size_t, vp]
    L.kge_train_step_next.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), vp, i64, u64, u64, u64, vp,
                                      vp, u64, vp, C.c_size_t, vp]
    L.kge_train_step_host_next.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), vp, vp, i64, u64, u64,
                                           u64, vp, vp, vp, vp, u64, vp, C.c_size_t, vp]
    L.kge_step_graph_create.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), vp, vp, i64, u64, vp, vp,
                                        C.c_size_t, C.POINTER(vp)]
    L.kge_step_graph_launch.argtypes = [vp, C.POINTER(Tables), vp, u64, u64, vp, u64, vp, vp]
    L.kge_step_graph_destroy.argtypes = [vp]
    L.kge_step_graph_phases.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.kge_step_graph_wait_loss.argtypes = [vp]
    L.kge_train_grad.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Dp), vp, i64,
                                 u64, u64, vp, C.POINTER(Grads), vp, vp, vp, C.c_size_t, vp]
    L.kge_train_apply.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Grads), vp, i64, u64, vp, vp,
                                  C.c_size_t, vp]
    L.kge_rank.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), vp, i64, i32, C.POINTER(Filter),
                           i64, i64, vp, vp, C.c_size_t, vp]
    L.kge_rank_lp_static.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), vp, i64, i32, C.POINTER(Filter),
                                     i64, i64, vp, vp]
    L.kge_filter_index_bytes.argtypes = [C.POINTER(Cfg), i64, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    L.kge_filter_index_build.argtypes = [vp, C.POINTER(Cfg), vp, i64, i32, vp, C.c_size_t, vp, C.c_size_t, vp]
    L.kge_filter_lookup.argtypes = [vp, C.POINTER(Cfg), vp, i64, vp, i64, i32, vp, vp, C.POINTER(Filter), vp]
    L.kge_epoch_perm_bytes.argtypes = [i64, C.POINTER(C.c_size_t)]
    L.kge_epoch_perm.argtypes = [vp, i64, u64, u64, vp, vp, C.c_size_t, vp]
    L.kge_batch_gather.argtypes = [vp, i32, vp, i64, vp, i64, i64, i32, vp, vp]
    L.kge_shard_lookup.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Shard),
                                   C.POINTER(Dp), vp, i64, u64, u64, vp, vp, C.c_size_t, vp]
    L.kge_shard_pre.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Shard), u64, vp]
    L.kge_shard_grad.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Shard),
                                 C.POINTER(Dp), vp, i64, u64, u64, vp, C.POINTER(Grads), vp, vp, vp, C.c_size_t, vp]
    L.kge_shard_apply.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Shard), C.POINTER(Grads), vp, i64,
                                  u64, vp, vp, C.c_size_t, vp]
    L.kge_peer_barrier.argtypes = [vp, C.POINTER(PeerFlags), C.c_uint32, vp]
    L.kge_train_step_dp.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Dp), vp, i64,
                                    u64, u64, u64, C.POINTER(Grads), C.POINTER(Grads), C.POINTER(PeerFlags), C.c_uint32,
                                    vp, vp, C.c_size_t, vp]
    L.kge_train_step_dp_next.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Dp), vp, i64,
                                         u64, u64, u64, C.POINTER(Grads), C.POINTER(Grads), C.POINTER(PeerFlags), C.c_uint32,
                                         vp, vp, u64, vp, C.c_size_t, vp]
    L.kge_step_graph_create_dp.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), C.POINTER(Sampler), C.POINTER(Dp),
                                           C.POINTER(Grads), C.POINTER(Grads), C.POINTER(PeerFlags), vp, vp, i64, u64, vp, vp,
                                           C.c_size_t, C.POINTER(vp)]
    L.kge_step_graph_launch_dp.argtypes = [vp, C.POINTER(Tables), vp, u64, u64, C.c_uint32, i32, vp, u64, vp, vp]
    L.kge_flush_rows.argtypes = [vp, C.POINTER(Cfg), C.POINTER(Tables), vp]
    L.kge_set_option.argtypes = [vp, i32, i32]
    L.kge_profile_begin.argtypes = [vp, i32]
    L.kge_profile_end.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]
    for s in SYMBOLS:
        getattr(L, s)
        if s not in ("kge_last_error",):
            getattr(L, s).restype = i32
    if L.kge_abi_version() != KGE_ABI_VERSION:
        raise RuntimeError("libkge_b200 ABI version mismatch")
    _LIB = L
    return L


def check(rc):
    if rc != 0:
        raise RuntimeError("libkge_b200: " + load().kge_last_error().decode() + f" (code {rc})")
