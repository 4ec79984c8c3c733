"""ctypes binding of libb200gat.so - the C-ABI declared in include/b200gat.h.

This is the binding a Wheatley maintainer would add.  There is no
fallback: if the library is missing the import fails loudly."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200GAT_LIB") or os.path.join(_HERE, "libb200gat.so")   # env: development builds

EDGE_NONE, EDGE_TABLE, EDGE_DENSE, EDGE_AFFINE = 0, 1, 2, 3
MAX_ATTR = 8
HUB_DEGREE = 256
HUB_PARTS = 16        # CTAs a hub destination's in-edges are split over
GRAPH_LOW_DEGREE, GRAPH_HUB_COUNTERS = 1, 2  # b200gat_graph.flags
LOW_DEGREE_MAX = 32   # longest non-hub adjacency list of a graph flagged GRAPH_LOW_DEGREE
HUB_REC = 256 + 64    # floats per partial record (include/b200gat.h)
FLAVOUR_ADDITIVE, FLAVOUR_DOT = 0, 1

# every symbol include/b200gat.h declares (tests check that the library exports all of them)
SYMBOLS = (
    "b200gat_last_error", "b200gat_version", "b200gat_launch_count", "b200gat_timing_enable",
    "b200gat_timing_read", "b200gat_plan_scratch_elems", "b200gat_plan_build",
    "b200gat_node_order_scratch_elems", "b200gat_node_order", "b200gat_hub_list",
    "b200gat_gather_i32", "b200gat_gather_i32_range", "b200gat_gather_rows_f32", "b200gat_narrow_i64", "b200gat_conv_fwd", "b200gat_conv_bwd",
    "b200gat_attention", "b200gat_edge_out", "b200gat_mask_head_fwd", "b200gat_mask_head_bwd",
    "b200gat_clique_scratch_elems", "b200gat_clique_build", "b200gat_conv_uses_cliques",
    "b200gat_timing_collect_graph", "b200gat_timing_clear_graph",
    "b200gat_gemm_supported", "b200gat_gemm_workspace_elems", "b200gat_gemm_tf32",
    "b200gat_gemm_precision", "b200gat_gemm_precision_get",
    "b200gat_colsum_scratch_elems", "b200gat_colsum", "b200gat_colsum_noalloc",
    "b200gat_mlp_fwd", "b200gat_mlp_bwd_workspace_elems", "b200gat_mlp_bwd",
    "b200gat_jssp_batch_edges", "b200gat_jssp_batch_build",
    "b200gat_psp_slot_words", "b200gat_psp_batch_sizes", "b200gat_psp_batch_build",
    "b200gat_readout_fwd", "b200gat_readout_bwd",
    "b200gat_gae", "b200gat_ppo_adv_moments", "b200gat_ppo_loss",
    "b200gat_trunk_supported", "b200gat_trunk_wcat_elems", "b200gat_trunk_bwd_workspace_elems",
    "b200gat_trunk_fwd", "b200gat_trunk_bwd",
    "b200gat_agent_bwd_workspace_elems", "b200gat_agent_fwd", "b200gat_agent_bwd",
)
MLP_MAX_LAYERS = 8
EPI_NONE, EPI_GELU, EPI_DGELU = 0, 1, 2
MAX_CLIQUE = 128
MIN_CLIQUE = 8        # smaller cliques stay on the sparse kernels


class Cliques(C.Structure):
    _fields_ = [("n_blocks", C.c_int32), ("max_block", C.c_int32), ("n_members", C.c_int32),
                ("s_elems", C.c_int32),
                ("blk_ptr", C.c_void_p), ("blk_nodes", C.c_void_p), ("blk_type", C.c_void_p),
                ("blk_soff", C.c_void_p), ("node_blk", C.c_void_p), ("node_run", C.c_void_p),
                ("node_rank", C.c_void_p)]


class Graph(C.Structure):
    pass


Graph._fields_ = [("n_nodes", C.c_int32), ("n_edges", C.c_int32),
                ("in_ptr", C.c_void_p), ("in_src", C.c_void_p), ("in_eid", C.c_void_p),
                ("out_ptr", C.c_void_p), ("out_dst", C.c_void_p), ("out_pos", C.c_void_p),
                ("dst_order", C.c_void_p), ("src_order", C.c_void_p),
                ("hub_degree", C.c_int32), ("hub_dst", C.c_void_p), ("n_hub_dst", C.c_void_p),
                ("hub_src", C.c_void_p), ("n_hub_src", C.c_void_p),
                ("hub_scratch", C.c_void_p), ("hub_parts", C.c_int32), ("hub_cap", C.c_int32),
                ("in_pos", C.c_void_p), ("cliques", C.POINTER(Cliques)), ("residual", C.POINTER(Graph)),
                ("flags", C.c_int32)]


class Attn(C.Structure):
    _fields_ = [("flavour", C.c_int32), ("heads", C.c_int32), ("dim", C.c_int32),
                ("slope", C.c_float), ("softmax_eps", C.c_float),
                ("attn", C.c_void_p), ("attn_bias", C.c_void_p), ("out_bias", C.c_void_p)]


class EdgeFeats(C.Structure):
    _fields_ = [("mode", C.c_int32), ("n_types", C.c_int32),
                ("table", C.c_void_p), ("etype", C.c_void_p), ("dense", C.c_void_p),
                ("n_attr", C.c_int32), ("attr", C.c_void_p), ("affine", C.c_void_p)]


class Gemm(C.Structure):
    """``b200gat_gemm`` (include/b200gat.h)."""
    _fields_ = [("a_major", C.c_int32), ("b_major", C.c_int32),
                ("M", C.c_int64), ("N", C.c_int64), ("K", C.c_int64),
                ("A", C.c_void_p), ("lda", C.c_int64), ("a_parts", C.c_int32), ("a_part_stride", C.c_int64),
                ("B", C.c_void_p), ("ldb", C.c_int64),
                ("D", C.c_void_p), ("ldd", C.c_int64), ("d_parts", C.c_int32), ("d_part_stride", C.c_int64),
                ("bias", C.c_void_p), ("residual", C.c_void_p), ("ldr", C.c_int64),
                ("epilogue", C.c_int32), ("D2", C.c_void_p), ("Zin", C.c_void_p), ("ld2", C.c_int64),
                ("workspace", C.c_void_p), ("workspace_elems", C.c_int64), ("a_rowsum", C.c_void_p)]


class Mlp(C.Structure):
    """``b200gat_mlp`` (include/b200gat.h)."""
    _fields_ = [("n_layers", C.c_int32), ("M", C.c_int64), ("dims", C.c_int32 * 9),
                ("W", C.c_void_p * 8), ("b", C.c_void_p * 8), ("x", C.c_void_p), ("residual", C.c_void_p),
                ("z", C.c_void_p * 8), ("h", C.c_void_p * 8), ("y", C.c_void_p)]


class JsspObs(C.Structure):
    """``b200gat_jssp_obs`` (include/b200gat.h)."""
    _fields_ = [("n_graphs", C.c_int32), ("max_nodes", C.c_int32), ("n_feat", C.c_int32), ("max_edges", C.c_int32),
                ("features", C.c_void_p), ("edge_index", C.c_void_p), ("meta", C.c_void_p),
                ("clique_conflicts", C.c_int32), ("mid_in_edges", C.c_int32), ("bidir", C.c_int32),
                ("self_loops", C.c_int32), ("no_tct", C.c_int32), ("pooling", C.c_int32),
                ("n_machine_ids", C.c_int32), ("n_nodes_total", C.c_int64), ("n_edges_graphs", C.c_int64),
                ("node_conflicts", C.c_int32), ("max_n_machines", C.c_int32), ("n_affected", C.c_int64)]


POOLING = {None: 0, "learn": 1, "learninv": 2}
CONFLICTS = {"att": 0, "clique": 1, "node": 2}
PSP_MAGIC, PSP_HEADER_WORDS, PSP_N_FEAT, PSP_MAX_RES = 0x31505350, 16, 16, 32
PSP_BASE_CODES = {"prec": 0, "rprec": 1, "rp": 2, "rrp": 3, "rc": 4}


class PspLayout(C.Structure):
    """``b200gat_psp_layout`` (include/b200gat.h)."""
    _fields_ = [("max_nodes", C.c_int32), ("max_res", C.c_int32), ("max_prec", C.c_int32), ("max_rp", C.c_int32),
                ("slot_words", C.c_int64)]


class PspObs(C.Structure):
    """``b200gat_psp_obs`` (include/b200gat.h)."""
    _fields_ = [("layout", PspLayout), ("n_graphs", C.c_int32), ("arena", C.c_void_p), ("meta", C.c_void_p),
                ("conflicts", C.c_int32), ("pooling", C.c_int32), ("self_loops", C.c_int32),
                ("base_order", C.c_int32 * 5), ("n_res", C.c_int32), ("unused_res_mask", C.c_uint32),
                ("seg_start", C.c_int64 * 16), ("n_edges", C.c_int64), ("n_task_nodes", C.c_int64),
                ("n_total_nodes", C.c_int64)]


class MlpGrads(C.Structure):
    """``b200gat_mlp_grads`` (include/b200gat.h)."""
    _fields_ = [("gy", C.c_void_p), ("dx", C.c_void_p), ("dW", C.c_void_p * 8), ("db", C.c_void_p * 8),
                ("dz", C.c_void_p * 8), ("workspace", C.c_void_p), ("workspace_elems", C.c_int64)]


READOUT_MAX_SRC = 16
READOUT_NONE, READOUT_MAX, READOUT_AVG, READOUT_ROWS = 0, 1, 2, 3


class Readout(C.Structure):
    """``b200gat_readout`` (include/b200gat.h)."""
    _fields_ = [("n_src", C.c_int32), ("src", C.c_void_p * 16), ("width", C.c_int32 * 16), ("ld", C.c_int64 * 16),
                ("n_graphs", C.c_int32), ("max_n_nodes", C.c_int32), ("row_start", C.c_void_p),
                ("n_task_rows", C.c_int64), ("pooling", C.c_int32), ("pool_row0", C.c_int64), ("out_ld", C.c_int64)]


TRUNK_MAX_LAYERS = 16


class TrunkLayer(C.Structure):
    """``b200gat_trunk_layer`` (include/b200gat.h)."""
    _fields_ = [("n_proj", C.c_int32), ("Wp", C.c_void_p * 3), ("bp", C.c_void_p * 3), ("attn", Attn),
                ("ef", EdgeFeats), ("n_mlp", C.c_int32), ("dims", C.c_int32 * 9), ("W", C.c_void_p * 8),
                ("b", C.c_void_p * 8), ("P", C.c_void_p), ("conv_out", C.c_void_p), ("s_save", C.c_void_p),
                ("m_save", C.c_void_p), ("l_save", C.c_void_p), ("agg_save", C.c_void_p), ("s_dense", C.c_void_p),
                ("z", C.c_void_p * 8), ("h", C.c_void_p * 8), ("y", C.c_void_p)]


class Trunk(C.Structure):
    """``b200gat_trunk`` (include/b200gat.h)."""
    _fields_ = [("n_layers", C.c_int32), ("in_dim", C.c_int32), ("residual", C.c_int32), ("n_nodes", C.c_int64),
                ("graph", C.POINTER(Graph)), ("x0", C.c_void_p), ("wcat", C.c_void_p),
                ("layer", TrunkLayer * TRUNK_MAX_LAYERS)]


class TrunkLayerGrads(C.Structure):
    """``b200gat_trunk_layer_grads`` (include/b200gat.h)."""
    _fields_ = [("g_y", C.c_void_p), ("dWp", C.c_void_p), ("dbp", C.c_void_p), ("g_attn", C.c_void_p),
                ("g_table", C.c_void_p), ("g_affine", C.c_void_p), ("g_out_bias", C.c_void_p),
                ("dW", C.c_void_p * 8), ("db", C.c_void_p * 8)]


class TrunkGrads(C.Structure):
    """``b200gat_trunk_grads`` (include/b200gat.h)."""
    _fields_ = [("g_x0", C.c_void_p), ("g_x0_in", C.c_void_p), ("workspace", C.c_void_p),
                ("workspace_elems", C.c_int64), ("layer", TrunkLayerGrads * TRUNK_MAX_LAYERS)]


CRITIC = {"l2": 0, "l1": 1, "sl1": 2}


class PpoLossArgs(C.Structure):
    """``b200gat_ppo_loss_args`` (include/b200gat.h)."""
    _fields_ = [("n", C.c_int32), ("logprob", C.c_void_p), ("old_logprob", C.c_void_p), ("advantages", C.c_void_p),
                ("entropy", C.c_void_p), ("value", C.c_void_p), ("returns", C.c_void_p), ("old_values", C.c_void_p),
                ("clip_coef", C.c_float), ("norm_adv", C.c_int32), ("clip_vloss", C.c_int32),
                ("critic_loss", C.c_int32), ("ent_coef", C.c_float), ("vf_coef", C.c_float), ("moments", C.c_void_p)]


AGENT_MAX_PAD = 8
SRC_EXTERNAL, SRC_X0, SRC_LAYER0 = 0, 1, 2


class PadJob(C.Structure):
    """``b200gat_pad_job`` (include/b200gat.h)."""
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("ld_src", C.c_int64), ("ld_dst", C.c_int64),
                ("rows", C.c_int32), ("cols", C.c_int32), ("rows_dst", C.c_int32), ("reserved", C.c_int32)]


class AgentStep(C.Structure):
    """``b200gat_agent_step`` (include/b200gat.h)."""
    _fields_ = [("n_pad", C.c_int32), ("pad", PadJob * AGENT_MAX_PAD), ("has_embedder", C.c_int32),
                ("embedder", Mlp), ("trunk", Trunk), ("readout", Readout), ("src_kind", C.c_int32 * 16),
                ("node_features", C.c_void_p), ("graph_embedding", C.c_void_p), ("argmax", C.c_void_p),
                ("action_net", Mlp), ("value_net", Mlp), ("mask", C.c_void_p), ("ld_mask", C.c_int64),
                ("action", C.c_void_p), ("fill", C.c_float), ("logits", C.c_void_p), ("logp", C.c_void_p),
                ("head", C.c_void_p), ("loss", PpoLossArgs), ("loss_out", C.c_void_p), ("g_head", C.c_void_p)]


class AgentStepGrads(C.Structure):
    """``b200gat_agent_step_grads`` (include/b200gat.h)."""
    _fields_ = [("g_loss", C.c_void_p), ("embedder", MlpGrads), ("action_net", MlpGrads), ("value_net", MlpGrads),
                ("trunk", TrunkGrads), ("g_src_ext", C.c_void_p * 16), ("g_x0", C.c_void_p),
                ("workspace", C.c_void_p), ("workspace_elems", C.c_int64)]


class B200GatError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m wheatley_b200.build` "
            "(or __graft_entry__.build()).  wheatley_b200 has no CPU / PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int32, C.c_float
    lib.b200gat_last_error.restype = C.c_char_p
    lib.b200gat_last_error.argtypes = []
    lib.b200gat_version.restype = C.c_int
    lib.b200gat_launch_count.restype = i64
    lib.b200gat_launch_count.argtypes = []
    lib.b200gat_timing_enable.restype = None
    lib.b200gat_timing_enable.argtypes = [C.c_int]
    lib.b200gat_timing_read.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(i64)]
    lib.b200gat_timing_collect_graph.argtypes = []
    lib.b200gat_timing_clear_graph.argtypes = []
    lib.b200gat_timing_clear_graph.restype = None
    lib.b200gat_plan_scratch_elems.restype = i64
    lib.b200gat_plan_scratch_elems.argtypes = [i64, i64]
    lib.b200gat_plan_build.argtypes = [vp, vp, i64, i64] + [vp] * 9
    lib.b200gat_node_order_scratch_elems.restype = i64
    lib.b200gat_node_order_scratch_elems.argtypes = [i64]
    lib.b200gat_node_order.argtypes = [vp, vp, i64, vp, vp, vp]
    lib.b200gat_hub_list.argtypes = [vp, i64, i32, vp, vp, vp]
    lib.b200gat_gather_i32.argtypes = [vp, vp, i64, vp, vp]
    lib.b200gat_gather_i32_range.argtypes = [vp, vp, i64, vp, vp, vp]
    lib.b200gat_gather_rows_f32.argtypes = [vp, vp, i64, i32, vp, vp]
    lib.b200gat_narrow_i64.argtypes = [vp, i64, vp, vp]
    gp, ap, ep = C.POINTER(Graph), C.POINTER(Attn), C.POINTER(EdgeFeats)
    lib.b200gat_conv_fwd.argtypes = [gp, ap, ep] + [vp] * 10
    lib.b200gat_conv_bwd.argtypes = [gp, ap, ep] + [vp] * 19
    lib.b200gat_attention.argtypes = [gp, i32, f32, vp, vp, vp, vp, vp, vp]
    lib.b200gat_clique_scratch_elems.restype = i64
    lib.b200gat_clique_scratch_elems.argtypes = [i64, i64]
    lib.b200gat_clique_build.argtypes = [gp, vp, i32] + [vp] * 16
    lib.b200gat_conv_uses_cliques.argtypes = [gp, ap, ep]
    lib.b200gat_edge_out.argtypes = [gp, ap, ep, vp, vp, vp, vp]
    lib.b200gat_mask_head_fwd.argtypes = [vp, i64, vp, i64, i32, i32, f32, vp, vp, vp, vp]
    lib.b200gat_mask_head_bwd.argtypes = [vp, vp, i64, vp, vp, vp, vp, i32, i32, vp, vp]
    mp = C.POINTER(Gemm)
    lib.b200gat_gemm_supported.argtypes = [mp]
    lib.b200gat_gemm_workspace_elems.argtypes = [mp]
    lib.b200gat_gemm_workspace_elems.restype = i64
    lib.b200gat_gemm_tf32.argtypes = [mp, vp]
    lib.b200gat_gemm_precision.argtypes = [C.c_int]
    lib.b200gat_gemm_precision.restype = None
    lib.b200gat_gemm_precision_get.argtypes = []
    lib.b200gat_colsum_scratch_elems.argtypes = [i64, i64]
    lib.b200gat_colsum_scratch_elems.restype = i64
    lib.b200gat_colsum.argtypes = [vp, i64, i64, i64, vp, vp, vp]
    lib.b200gat_colsum_noalloc.argtypes = [vp, i64, i64, i64, vp, vp]
    lp, lg = C.POINTER(Mlp), C.POINTER(MlpGrads)
    lib.b200gat_mlp_fwd.argtypes = [lp, vp]
    lib.b200gat_mlp_bwd_workspace_elems.argtypes = [lp, lg]
    lib.b200gat_mlp_bwd_workspace_elems.restype = i64
    lib.b200gat_mlp_bwd.argtypes = [lp, lg, vp]
    jp = C.POINTER(JsspObs)
    lib.b200gat_jssp_batch_edges.argtypes = [jp]
    lib.b200gat_jssp_batch_edges.restype = i64
    lib.b200gat_jssp_batch_build.argtypes = [jp, vp, vp, vp, vp, vp]
    pl, po = C.POINTER(PspLayout), C.POINTER(PspObs)
    lib.b200gat_psp_slot_words.argtypes = [pl]
    lib.b200gat_psp_slot_words.restype = i64
    lib.b200gat_psp_batch_sizes.argtypes = [po, vp]
    lib.b200gat_psp_batch_build.argtypes = [po, vp, vp, vp, vp, vp, vp, vp]
    lib.b200gat_gae.argtypes = [vp] * 5 + [i32, i32, i32, f32, f32, f32, vp, vp, vp]
    lib.b200gat_ppo_adv_moments.argtypes = [vp, i32, vp, vp]
    lib.b200gat_ppo_loss.argtypes = [C.POINTER(PpoLossArgs), vp, vp, vp, vp, vp, vp]
    tp, tg = C.POINTER(Trunk), C.POINTER(TrunkGrads)
    lib.b200gat_trunk_supported.argtypes = [tp]
    lib.b200gat_trunk_wcat_elems.argtypes = [tp]
    lib.b200gat_trunk_wcat_elems.restype = i64
    lib.b200gat_trunk_bwd_workspace_elems.argtypes = [tp]
    lib.b200gat_trunk_bwd_workspace_elems.restype = i64
    lib.b200gat_trunk_fwd.argtypes = [tp, vp]
    lib.b200gat_trunk_bwd.argtypes = [tp, tg, vp]
    sp, sg = C.POINTER(AgentStep), C.POINTER(AgentStepGrads)
    lib.b200gat_agent_bwd_workspace_elems.argtypes = [sp]
    lib.b200gat_agent_bwd_workspace_elems.restype = i64
    lib.b200gat_agent_fwd.argtypes = [sp, vp]
    lib.b200gat_agent_bwd.argtypes = [sp, sg, vp]
    rp = C.POINTER(Readout)
    lib.b200gat_readout_fwd.argtypes = [rp, vp, vp, vp, vp]
    lib.b200gat_readout_bwd.argtypes = [rp, vp, i64, vp, i64, vp, C.POINTER(C.c_void_p), i64, vp]
    for name in SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("b200gat_last_error", "b200gat_plan_scratch_elems", "b200gat_launch_count",
                        "b200gat_timing_enable", "b200gat_node_order_scratch_elems",
                        "b200gat_clique_scratch_elems", "b200gat_timing_clear_graph",
                        "b200gat_gemm_workspace_elems", "b200gat_colsum_scratch_elems",
                        "b200gat_mlp_bwd_workspace_elems", "b200gat_jssp_batch_edges", "b200gat_psp_slot_words",
                        "b200gat_trunk_wcat_elems", "b200gat_trunk_bwd_workspace_elems",
                        "b200gat_agent_bwd_workspace_elems", "b200gat_gemm_precision"):
            fn.restype = C.c_int
    return lib


lib = _load()


def check(rc, what):
    if rc != 0:
        raise B200GatError(f"{what} failed (code {rc}): {lib.b200gat_last_error().decode()}")


TIMING_CLASSES = {"conv_fwd": 0, "conv_bwd_dst": 1, "conv_bwd_src": 2, "plan": 3, "head": 4, "gemm": 5,
                  "batch": 6, "gemm_dx": 7, "clq_fwd": 8, "clq_bwd_mma": 9, "clq_bwd_elem": 10, "gemm_dw": 11}


def launch_count() -> int:
    return int(lib.b200gat_launch_count())


def timing_enable(on: bool):
    lib.b200gat_timing_enable(1 if on else 0)


def timing_read():
    """{class: (total_ms, launches)} since the last ``timing_enable``."""
    out = {}
    for name, cls in TIMING_CLASSES.items():
        ms, n = C.c_double(0), C.c_int64(0)
        check(lib.b200gat_timing_read(cls, C.byref(ms), C.byref(n)), "b200gat_timing_read")
        out[name] = (ms.value, n.value)
    return out


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


GEMM_TF32, GEMM_FP32X3 = 0, 1
_gemm_mode = [GEMM_TF32]


def sync_gemm_precision():
    """Copy ``torch.backends.cuda.matmul.allow_tf32`` into the library's process-wide GEMM arithmetic switch:
    TF32 operands when matmuls may use TF32 (the reference's live trainer, psp/train_psp.py:47), 3xTF32 split
    arithmetic (fp32 matmul accuracy, three tcgen05.mma per k step) when they may not."""
    import torch
    mode = GEMM_TF32 if torch.backends.cuda.matmul.allow_tf32 else GEMM_FP32X3
    if mode != _gemm_mode[0]:
        lib.b200gat_gemm_precision(mode)
        _gemm_mode[0] = mode


def stream_ptr():
    """Raw ``cudaStream_t`` of torch's current stream on the current device (the cheap accessor:
    ``torch.cuda.current_stream()`` builds a Python object, ~10 us per call on the eager path).  Every C-ABI call
    fetches its stream here right before the call, so this is also where the GEMM arithmetic switch is kept in
    step with torch's matmul-precision flag."""
    import torch
    sync_gemm_precision()
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


class _NoGuard:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NO_GUARD = _NoGuard()


def device_guard(device):
    """``torch.cuda.device(device)`` only when ``device`` is not already current (the common case costs
    two attribute reads instead of a context-manager round trip through the CUDA runtime)."""
    import torch
    idx = device.index
    if idx is None or idx == torch._C._cuda_getDevice():
        return _NO_GUARD
    return torch.cuda.device(idx)


def eager_only(fn):
    """The fused ops call the C-ABI through ctypes with raw device pointers: nothing a tracing compiler can look
    into.  The reference wraps its trunks in ``torch.compile(dynamic=True)`` (psp/models/gnn_mp.py:139-300); marked
    ``torch.compiler.disable`` these functions run eagerly inside such a wrapper (a clean graph break around one
    call) instead of failing in the tracer."""
    import torch
    dis = getattr(getattr(torch, "compiler", None), "disable", None)
    return dis(fn) if dis is not None else fn
