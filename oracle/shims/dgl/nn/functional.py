import math
import torch


def edge_softmax(graph, logits, eids=None, norm_by="dst"):
    """DGL edge_softmax: a_e = exp(s_e - max_{e'->dst} s_e') / sum_{e'->dst} exp(...)."""
    assert norm_by == "dst" and eids is None
    dst = graph._dst
    n = graph.num_nodes()
    shp = logits.shape
    flat = logits.reshape(shp[0], -1)
    idx = dst.view(-1, 1).expand_as(flat)
    smax = torch.full((n, flat.shape[1]), -math.inf, dtype=flat.dtype)
    smax = smax.scatter_reduce(0, idx, flat.detach(), reduce="amax", include_self=True)
    ex = torch.exp(flat - smax[dst])
    ssum = torch.zeros((n, flat.shape[1]), dtype=flat.dtype).index_add(0, dst, ex)
    return (ex / ssum[dst]).reshape(shp)
