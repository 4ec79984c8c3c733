"""GraphPlan: the int32 CSR-by-destination / CSR-by-source view of one batched graph,
built on the GPU by ``b200gat_plan_build`` once per batch and shared by all layers and
by forward and backward.

Replaces what DGL builds behind ``apply_edges / edge_softmax / update_all``
(generic/eegatconv.py:269,280,292) and PyG behind ``propagate``
(generic/gatv3_conv.py:155-160)."""
import torch

from . import _lib


class GraphPlan:
    def __init__(self, src: torch.Tensor, dst: torch.Tensor, n_nodes: int):
        if not (src.is_cuda and dst.is_cuda):
            raise _lib.B200GatError("GraphPlan needs CUDA tensors: wheatley_b200 has no CPU path")
        if src.dtype != torch.int64:
            src = src.long()
        if dst.dtype != torch.int64:
            dst = dst.long()
        src, dst = src.contiguous(), dst.contiguous()
        E, N = int(src.numel()), int(n_nodes)
        dev = src.device
        i32 = dict(dtype=torch.int32, device=dev)
        self.n_nodes, self.n_edges, self.device = N, E, dev
        self.in_ptr = torch.empty(N + 1, **i32)
        self.out_ptr = torch.empty(N + 1, **i32)
        self.in_src = torch.empty(E, **i32)
        self.in_eid = torch.empty(E, **i32)
        self.out_dst = torch.empty(E, **i32)
        self.out_pos = torch.empty(E, **i32)
        self.stats = torch.empty(4, **i32)
        scratch = torch.empty(_lib.lib.b200gat_plan_scratch_elems(N, E), **i32)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib.b200gat_plan_build(
                _lib.ptr(src), _lib.ptr(dst), E, N, _lib.ptr(self.in_ptr), _lib.ptr(self.in_src),
                _lib.ptr(self.in_eid), _lib.ptr(self.out_ptr), _lib.ptr(self.out_dst),
                _lib.ptr(self.out_pos), _lib.ptr(scratch), _lib.ptr(self.stats),
                _lib.stream_ptr()), "b200gat_plan_build")
            # locality order of destinations / sources (a permutation: affects scheduling only)
            self.dst_order = torch.empty(N, **i32)
            self.src_order = torch.empty(N, **i32)
            oscr = torch.empty(_lib.lib.b200gat_node_order_scratch_elems(N), **i32)
            for ptr_, nbr_, out_ in ((self.in_ptr, self.in_src, self.dst_order),
                                     (self.out_ptr, self.out_dst, self.src_order)):
                _lib.check(_lib.lib.b200gat_node_order(_lib.ptr(ptr_), _lib.ptr(nbr_), N, _lib.ptr(out_),
                                                       _lib.ptr(oscr), _lib.stream_ptr()),
                           "b200gat_node_order")
            # hub nodes (segments longer than HUB_DEGREE are split over a CTA's warps)
            self.hub_dst = torch.empty(N, **i32)
            self.hub_src = torch.empty(N, **i32)
            self.n_hub = torch.empty(2, **i32)
            for k, (ptr_, lst_) in enumerate(((self.in_ptr, self.hub_dst), (self.out_ptr, self.hub_src))):
                _lib.check(_lib.lib.b200gat_hub_list(_lib.ptr(ptr_), N, _lib.HUB_DEGREE, _lib.ptr(lst_),
                                                     self.n_hub.data_ptr() + 4 * k, _lib.stream_ptr()),
                           "b200gat_hub_list")
        self._stats_host = None
        self._c = _lib.Graph(N, E, _lib.ptr(self.in_ptr), _lib.ptr(self.in_src), _lib.ptr(self.in_eid),
                             _lib.ptr(self.out_ptr), _lib.ptr(self.out_dst), _lib.ptr(self.out_pos),
                             _lib.ptr(self.dst_order), _lib.ptr(self.src_order), _lib.HUB_DEGREE,
                             _lib.ptr(self.hub_dst), self.n_hub.data_ptr(), _lib.ptr(self.hub_src),
                             self.n_hub.data_ptr() + 4)
        self._sorted_cache = {}

    @classmethod
    def from_edge_index(cls, edge_index: torch.Tensor, n_nodes: int):
        """PyG convention: row 0 = source, row 1 = target (psp/graph/graph_conv.py:53)."""
        return cls(edge_index[0], edge_index[1], n_nodes)

    @property
    def c(self):
        return self._c

    def stats_host(self):
        """(n_zero_in_degree, max_in_degree, n_bad_edges, max_out_degree); one device sync,
        cached."""
        if self._stats_host is None:
            self._stats_host = tuple(int(x) for x in self.stats.tolist())
        return self._stats_host

    def validate(self):
        bad = self.stats_host()[2]
        if bad:
            raise _lib.B200GatError(f"{bad} edges have an endpoint outside [0, {self.n_nodes})")

    def sort_edge_ints(self, per_edge: torch.Tensor, key=None) -> torch.Tensor:
        """Permute an int per-edge attribute (e.g. edge type) from the original edge order
        into the plan's destination-sorted order (int32)."""
        key = key if key is not None else (per_edge.data_ptr(), per_edge._version)
        hit = self._sorted_cache.get(key)
        if hit is not None:
            return hit
        t = per_edge
        if t.dtype != torch.int32:
            t32 = torch.empty(t.numel(), dtype=torch.int32, device=t.device)
            if t.dtype == torch.int64 and t.is_contiguous() and t.data_ptr() % 16 == 0:
                _lib.check(_lib.lib.b200gat_narrow_i64(_lib.ptr(t), t.numel(), _lib.ptr(t32),
                                                       _lib.stream_ptr()), "b200gat_narrow_i64")
            else:
                t32.copy_(t)
            t = t32
        out = torch.empty(self.n_edges, dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib.b200gat_gather_i32(_lib.ptr(t.contiguous()), _lib.ptr(self.in_eid),
                                               self.n_edges, _lib.ptr(out), _lib.stream_ptr()),
                   "b200gat_gather_i32")
        self._sorted_cache[key] = out
        return out


    def sort_edge_rows(self, per_edge: torch.Tensor, key=None) -> torch.Tensor:
        """Permute fp32 per-edge attribute rows ``[E, K]`` into destination-sorted order."""
        key = key if key is not None else ("rows", per_edge.data_ptr(), per_edge._version)
        hit = self._sorted_cache.get(key)
        if hit is not None:
            return hit
        t = per_edge.detach()
        if t.dtype != torch.float32:
            t = t.float()
        t = t.contiguous().view(self.n_edges, -1)
        out = torch.empty_like(t)
        _lib.check(_lib.lib.b200gat_gather_rows_f32(_lib.ptr(t), _lib.ptr(self.in_eid), self.n_edges,
                                                    t.shape[1], _lib.ptr(out), _lib.stream_ptr()),
                   "b200gat_gather_rows_f32")
        self._sorted_cache[key] = out
        return out


def plan_for_edge_index(edge_index: torch.Tensor, n_nodes: int, holder=None) -> GraphPlan:
    """Plan cache: the plan is stored on ``holder`` (the graph object the reference passes
    around, or the edge_index tensor itself) and reused while the tensor is unchanged."""
    holder = edge_index if holder is None else holder
    tag = (edge_index.data_ptr(), edge_index._version, tuple(edge_index.shape), int(n_nodes))
    cached = getattr(holder, "_b200gat_plan", None)
    if cached is not None and cached[0] == tag:
        return cached[1]
    plan = GraphPlan.from_edge_index(edge_index, n_nodes)
    try:
        holder._b200gat_plan = (tag, plan)
    except AttributeError:
        pass
    return plan
