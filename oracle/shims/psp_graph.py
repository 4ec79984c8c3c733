"""Stand-in for ``psp/graph/pyg_graph.py`` PYGGraph / PYGBatchGraph WITHOUT torch_geometric.
TEST INFRASTRUCTURE - see oracle/__init__.py.

The reference's PyG-flavour batch construction (``psp/models/agent.py:298-406``,
``psp/models/rewiring.py``, ``psp/models/agent_graph_observation.py``) is pure torch code written against
the abstract graph interface of ``psp/graph/graph.py:4-88``; its only implementation at this commit wraps
``torch_geometric.data.HeteroData`` (un-vendored, un-pinned third party, not installable here).  This file
restates, method by method, what ``pyg_graph.py`` does on a HeteroData with ONE node type ``"n"``, keeping the
HeteroData behaviours the batching result depends on (published PyG 2.5 / 2.6 ``torch_geometric/data``
semantics, cited per method):

* stores are created on first access and keep INSERTION ORDER (``HeteroData.__getitem__`` ->
  ``get_edge_store``): the order of the edge types in the final homogeneous graph;
* ``EdgeStorage.num_edges`` is ``edge_index.size(1)`` when the store has an ``edge_index`` and 0 otherwise
  (a store holding only ``type`` / ``rid`` / ``att_rc`` / ``att_rp`` has no key containing "edge");
* ``Batch.from_data_list`` (``data/collate.py``): the batch gets the stores of the FIRST graph, in its order,
  with that graph's attribute keys; tensors are concatenated along dim 0, ``edge_index`` along dim 1 with the
  cumulative node count added; the node store gets ``batch`` (graph id per node) and ``ptr``;
* ``HeteroData.to_homogeneous(node_attrs, edge_attrs, add_node_type=False, add_edge_type=False,
  dummy_values=False)``: ``edge_index`` = concatenation over the edge stores THAT HAVE an ``edge_index``
  (``collect('edge_index', allow_missing=True)``), in store order; every requested edge attribute =
  concatenation along dim 0 over ALL edge stores in store order (every store must hold it).

Everything else (``add_edges``, ``add_nodes``, ``set_edata`` with ``index`` ...) is the body of the
corresponding ``pyg_graph.py`` method on those containers.  ``file:line`` = psp/graph/pyg_graph.py."""
import torch


class _Store(dict):
    """One HeteroData storage: attribute dict; ``num_edges`` as EdgeStorage defines it."""

    @property
    def num_edges(self):
        ei = self.get("edge_index")
        return 0 if ei is None else int(ei.shape[1])

    @property
    def edge_index(self):
        return self["edge_index"]


class _Hetero:
    """HeteroData with one node type: ``node`` store, insertion-ordered ``edges`` stores, ``globals``."""

    def __init__(self):
        self.node = _Store()
        self.edges = {}
        self.globals = {}

    def edge_store(self, etype) -> _Store:
        st = self.edges.get(etype)
        if st is None:                      # HeteroData creates the store on access
            st = self.edges[etype] = _Store()
        return st

    def clone(self):
        h = _Hetero()
        cp = lambda v: v.clone() if isinstance(v, torch.Tensor) else v   # noqa: E731
        h.node = _Store({k: cp(v) for k, v in self.node.items()})
        h.edges = {t: _Store({k: cp(v) for k, v in st.items()}) for t, st in self.edges.items()}
        h.globals = {k: cp(v) for k, v in self.globals.items()}
        return h


class StandInGraph:
    """pyg_graph.py PYGGraph (single graph)."""

    def __init__(self, problem_edges, num_nodes, factored_rp=False, observe_conflicts_as_cliques=False):
        self._graph = _Hetero()                                           # :20-25
        self._graph.node["num_nodes"] = num_nodes
        self._graph.edge_store("prec")["edge_index"] = torch.tensor(problem_edges, dtype=torch.int64).t().reshape(2, -1)
        self.factored_rp = factored_rp
        self.observe_conflicts_as_cliques = observe_conflicts_as_cliques

    # ---- nodes
    def ndata(self, featname=None):                                        # :63-70
        if isinstance(self._graph, _Homo):
            return self._graph if featname is None else self._graph[featname]
        return self._graph.node if featname is None else self._graph.node[featname]

    def set_ndata(self, featname, t):                                      # :72-73
        self._graph.node[featname] = t

    def num_nodes(self):                                                   # :81-84
        n = self._graph.node.get("num_nodes")
        return 0 if n is None else n

    def add_nodes(self, num_nodes, featname, data):                        # :161-162 (num_nodes NOT updated)
        self._graph.node[featname] = torch.cat([self._graph.node[featname], data])

    def add_batchinfo(self, batchinfo):                                    # :164-165
        self._graph.node["batch"] = torch.cat([self._graph.node["batch"], batchinfo])

    # ---- globals
    def set_global_data(self, featname, data):                             # :118-119
        self._graph.globals[featname] = data

    def global_data(self, featname=None):                                  # :121-124
        return self._graph.globals if featname is None else self._graph.globals[featname]

    # ---- edges
    def edges(self, etype):                                                # :126-145
        st = self._graph.edge_store(etype)
        if st.num_edges == 0:
            e = torch.tensor([], dtype=torch.int64)
            return e, e.clone(), e.clone()
        ei = st["edge_index"]
        return ei[0], ei[1], torch.arange(ei.shape[1], dtype=torch.int64)

    def add_edges(self, sources, destinations, etype, data=None):          # :167-185
        new_edges = torch.stack([sources, destinations], dim=0)
        st = self._graph.edge_store(etype)
        if st.num_edges == 0:
            st["edge_index"] = new_edges
            if data is not None:
                for key in data:
                    st[key] = data[key]
        else:
            st["edge_index"] = torch.cat([st["edge_index"], new_edges], dim=1)
            if data is not None:
                for key in data:
                    st[key] = torch.cat([st[key], data[key]], dim=0)

    def num_edges(self, etype):                                            # :187-188
        return self._graph.edge_store(etype).num_edges

    def edata(self, etype, dataid):                                        # :190-197
        if isinstance(self._graph, _Homo):
            return self._graph[dataid] if etype is None else self._graph[etype][dataid]
        if etype is None:
            return self._graph.node[dataid]
        return self._graph.edge_store(etype)[dataid]

    def set_edata(self, etype, featname, data, index=None):                # :75-79
        st = self._graph.edge_store(etype)
        if index is None:
            st[featname] = data
        else:
            st[featname][:, index] = data

    def fill_void(self):                                                   # :283-301
        for t in ("prec", "rprec", "rp", "rrp"):
            st = self._graph.edge_store(t)
            if "edge_index" not in st:
                st["edge_index"] = torch.tensor([[], []], dtype=torch.int64)
                if t in ("rp", "rrp"):
                    st["r"] = torch.empty(0, 4)

    def clone(self):                                                       # :217-227
        g = StandInGraph.__new__(StandInGraph)
        g._graph = self._graph.clone()
        g.factored_rp = self.factored_rp
        g.observe_conflicts_as_cliques = self.observe_conflicts_as_cliques
        return g

    # ---- batching
    @classmethod
    def batch(cls, graphlist, num_nodes, num_edges):                       # :199-201
        return StandInBatchGraph(graphlist)

    @classmethod
    def pad_resources(cls, graphlist):                                     # :203-215
        num_res = 0
        for g in graphlist:
            num_res = max(num_res, g.ndata("resources").shape[1])
        for g in graphlist:
            res = g.ndata("resources")
            g.set_ndata("resources", torch.nn.functional.pad(res, (0, num_res - res.shape[1])))
        return num_res, graphlist


class _Homo(dict):
    """``torch_geometric.data.Data`` after to_homogeneous: flat attribute dict + ``edge_index``."""

    @property
    def edge_index(self):
        return self["edge_index"]


class StandInBatchGraph(StandInGraph):
    """pyg_graph.py PYGBatchGraph (:319-346) on ``Batch.from_data_list``."""

    def __init__(self, graphs):
        first = graphs[0]._graph
        b = _Hetero()
        counts = [g.num_nodes() for g in graphs]
        offs = [0]
        for c in counts:
            offs.append(offs[-1] + c)
        for k in first.node:                                     # node store: attributes of the first graph
            if k == "num_nodes":
                b.node[k] = offs[-1]
            else:
                b.node[k] = torch.cat([g._graph.node[k] for g in graphs], dim=0)
        b.node["batch"] = torch.repeat_interleave(torch.arange(len(graphs)), torch.tensor(counts))
        b.node["ptr"] = torch.tensor(offs, dtype=torch.int64)
        for t, st0 in first.edges.items():                       # edge stores of the first graph, in its order
            st = b.edge_store(t)
            for k in st0:
                vals = [g._graph.edges[t][k] for g in graphs]
                if k == "edge_index":
                    st[k] = torch.cat([v + o for v, o in zip(vals, offs)], dim=1)
                else:
                    st[k] = torch.cat(vals, dim=0)
        for k in first.globals:
            b.globals[k] = torch.cat([g._graph.globals[k] for g in graphs], dim=0)
        self._graph = b

    def batch_num_nodes(self):                                             # :323-326
        ptr = self._graph.node["ptr"]
        return ptr[1:] - ptr[:-1]

    def batch_id(self):                                                    # :328-329
        return self._graph.node["batch"]

    def num_nodes(self):
        return self._graph.node["num_nodes"]

    def add_self_loops(self):                                              # :331-335 (one loop per node of the batch)
        loop = torch.arange(self.num_nodes(), dtype=torch.int64)
        self._graph.edge_store("self")["edge_index"] = torch.stack([loop, loop])

    def to_homogeneous(self, ndata, edata):                                # :337-340
        h = self._graph
        out = _Homo()
        for k in ndata:
            out[k] = h.node[k]
        out["edge_index"] = torch.cat([st["edge_index"] for st in h.edges.values() if "edge_index" in st], dim=1)
        for k in edata:
            out[k] = torch.cat([st[k] for st in h.edges.values()], dim=0)
        out.store_order = list(h.edges.keys())
        self._graph = out

    def to(self, device):                                                  # :342-344
        return self
