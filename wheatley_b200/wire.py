"""Observation wire format env worker -> learner (layout in include/b200gat.h "PSP observation
wire format").

The reference serialises every observation with ``torch.save`` of the HeteroData dict into a shared byte array
(psp/graph/pyg_graph.py:240-255 ``serialize``, psp/env/graphgym/async_vector_env.py:59-103 ``write_to_shared_memory``)
and unpickles it in the learner (``deserialize`` :271-283, async_vector_env.py ``read_from_shared_memory``): a
pickle round trip plus ~25 small tensor allocations per env step.  Here an observation is ONE fixed-layout slot
of 4-byte words inside a flat arena: the worker writes numpy views in place (no pickling, no allocation), the
learner copies the slots it needs to the device as they are (one pinned ``memcpy`` per batch) and
``b200gat_psp_batch_build`` reads them there.  A slot carries exactly what ``AgentGraphObservation.rewire_internal``
and ``Agent.preprocess`` consume (psp/models/agent_graph_observation.py:66-118, psp/models/agent.py:298-406): the
16 feature columns, ``resources``, ``res_cal``, the ``prec`` / ``rp`` edge lists and the ``rp`` attributes;
``rprec`` / ``rrp`` / ``rc`` are functions of those and are generated on the device.

Pure numpy + torch (pinned host memory); no kernel is involved in this file."""
import numpy as np
import torch

from . import _lib

FEAT_COLUMNS = ("affected", "selectable", "job", "type", "normalized_durations[3]", "normalized_tct[3]",
                "normalized_tardiness[3]", "has_due_date", "normalized_due_dates", "past")   # 16 columns


class PspObsLayout:
    """Word offsets of the fields of one slot for given capacities."""

    def __init__(self, max_nodes, max_res, max_prec, max_rp):
        if max_res > _lib.PSP_MAX_RES:
            raise ValueError(f"at most {_lib.PSP_MAX_RES} resources per graph")
        self.max_nodes, self.max_res, self.max_prec, self.max_rp = (int(max_nodes), int(max_res), int(max_prec),
                                                                     int(max_rp))
        self.c = _lib.PspLayout(self.max_nodes, self.max_res, self.max_prec, self.max_rp, 0)
        self.c.slot_words = int(_lib.lib.b200gat_psp_slot_words(self.c))
        self.slot_words = int(self.c.slot_words)
        o = _lib.PSP_HEADER_WORDS
        self.off_feat = o
        o += self.max_nodes * _lib.PSP_N_FEAT
        self.off_res = o
        o += self.max_nodes * self.max_res
        self.off_cal = o
        o += self.max_res * 3
        self.off_prec = o
        o += 2 * self.max_prec
        self.off_rp = o
        o += 2 * self.max_rp
        self.off_rpr = o
        o += 4 * self.max_rp
        assert o <= self.slot_words
        # [max_nodes * max_res, max_res] one-hot: column of each entry of the row-major ``resources`` block
        self.res_select = np.tile(np.eye(self.max_res, dtype=np.float32), (self.max_nodes, 1))

    @classmethod
    def for_graphs(cls, graphs):
        """Smallest layout that holds every graph of ``graphs`` (dicts as ``encode`` takes them)."""
        return cls(max(g["feat"].shape[0] for g in graphs), max(g["resources"].shape[1] for g in graphs),
                   max(g["prec"].shape[1] for g in graphs), max(g["rp"].shape[1] for g in graphs))

    # ---- worker side
    def encode(self, slot: np.ndarray, feat, resources, res_cal, prec, rp=None, rp_r=None):
        """Write one observation into ``slot`` (int32 view of ``slot_words`` words), in place."""
        n, R = int(feat.shape[0]), int(resources.shape[1])
        n_prec = int(prec.shape[1])
        n_rp = 0 if rp is None else int(rp.shape[1])
        if n > self.max_nodes or R > self.max_res or n_prec > self.max_prec or n_rp > self.max_rp:
            raise ValueError("observation exceeds the slot capacities "
                             f"(nodes {n}/{self.max_nodes}, resources {R}/{self.max_res}, prec {n_prec}/"
                             f"{self.max_prec}, rp {n_rp}/{self.max_rp})")        # async_vector_env.py:98-101
        if feat.shape[1] != _lib.PSP_N_FEAT:
            raise ValueError(f"feat must have {_lib.PSP_N_FEAT} columns")
        slot[: _lib.PSP_HEADER_WORDS] = 0
        slot[0:6] = (_lib.PSP_MAGIC, n, R, n_prec, n_rp, _lib.PSP_N_FEAT)
        f = slot.view(np.float32)
        v = f[self.off_feat: self.off_res].reshape(self.max_nodes, _lib.PSP_N_FEAT)
        v[:n] = feat
        v[n:] = 0
        v = f[self.off_res: self.off_cal].reshape(self.max_nodes, self.max_res)
        v[:] = 0
        v[:n, :R] = resources
        v = f[self.off_cal: self.off_prec].reshape(self.max_res, 3)
        v[:] = 0
        v[:R] = res_cal
        v = slot[self.off_prec: self.off_rp].reshape(2, self.max_prec)
        v[:, :n_prec] = prec
        if n_rp:
            slot[self.off_rp: self.off_rpr].reshape(2, self.max_rp)[:, :n_rp] = rp
            f[self.off_rpr: self.off_rpr + 4 * self.max_rp].reshape(self.max_rp, 4)[:n_rp] = rp_r

    # ---- learner side (host views; the device never needs this)
    def decode(self, slot: np.ndarray) -> dict:
        if int(slot[0]) != _lib.PSP_MAGIC:
            raise ValueError("not a PSP observation slot (bad magic)")
        n, R, n_prec, n_rp = (int(x) for x in slot[1:5])
        f = slot.view(np.float32)
        return dict(
            feat=f[self.off_feat: self.off_res].reshape(self.max_nodes, _lib.PSP_N_FEAT)[:n],
            resources=f[self.off_res: self.off_cal].reshape(self.max_nodes, self.max_res)[:n, :R],
            res_cal=f[self.off_cal: self.off_prec].reshape(self.max_res, 3)[:R],
            prec=slot[self.off_prec: self.off_rp].reshape(2, self.max_prec)[:, :n_prec].astype(np.int64),
            rp=slot[self.off_rp: self.off_rpr].reshape(2, self.max_rp)[:, :n_rp].astype(np.int64),
            rp_r=f[self.off_rpr: self.off_rpr + 4 * self.max_rp].reshape(self.max_rp, 4)[:n_rp])


class PspArena:
    """``n_slots`` observation slots in one pinned host buffer (``pin=False``: plain memory, e.g. a view onto a
    ``multiprocessing`` shared array).  ``slot(i)`` is the int32 view a worker writes; ``gather(idx)`` packs a
    minibatch's slots into another arena with one fancy-index copy (the rollout buffer -> minibatch step of
    alg/ppo.py:556-575, psp/models/agent.py:271-283 ``_get_obs_graph``)."""

    def __init__(self, layout: PspObsLayout, n_slots: int, pin=True, buffer=None):
        self.layout, self.n_slots = layout, int(n_slots)
        if buffer is not None:
            self.words = np.frombuffer(buffer, dtype=np.int32, count=self.n_slots * layout.slot_words)
            self.words = self.words.reshape(self.n_slots, layout.slot_words)
            self.tensor = torch.from_numpy(self.words)
        else:
            self.tensor = torch.zeros((self.n_slots, layout.slot_words), dtype=torch.int32,
                                      pin_memory=bool(pin and torch.cuda.is_available()))
            self.words = self.tensor.numpy()

    def slot(self, i):
        return self.words[i]

    def write(self, i, graph: dict):
        self.layout.encode(self.words[i], graph["feat"], graph["resources"], graph["res_cal"], graph["prec"],
                           graph.get("rp"), graph.get("rp_r"))

    def read(self, i) -> dict:
        return self.layout.decode(self.words[i])

    def gather(self, idx, out: "PspArena" = None) -> "PspArena":
        idx = np.asarray(idx, dtype=np.int64)
        if out is None:
            out = PspArena(self.layout, len(idx))
        np.take(self.words, idx, axis=0, out=out.words[: len(idx)])
        return out

    @classmethod
    def from_graphs(cls, graphs, layout=None, pin=True):
        layout = layout or PspObsLayout.for_graphs(graphs)
        a = cls(layout, len(graphs), pin=pin)
        for i, g in enumerate(graphs):
            a.write(i, g)
        return a
