import math
import torch


def remove_self_loops(edge_index, edge_attr=None):
    mask = edge_index[0] != edge_index[1]
    edge_index = edge_index[:, mask]
    return edge_index, (None if edge_attr is None else edge_attr[mask])


def add_self_loops(edge_index, edge_attr=None, fill_value=None, num_nodes=None):
    assert edge_attr is None
    loop = torch.arange(0, num_nodes, dtype=edge_index.dtype, device=edge_index.device)
    loop = loop.unsqueeze(0).repeat(2, 1)
    return torch.cat([edge_index, loop], dim=1), None


def softmax(src, index=None, ptr=None, num_nodes=None, dim=0):
    """torch_geometric.utils.softmax: max-subtracted (detached max), +1e-16 in the
    denominator."""
    assert ptr is None and dim == 0
    n = int(index.max()) + 1 if num_nodes is None else num_nodes
    flat = src.reshape(src.shape[0], -1)
    idx = index.view(-1, 1).expand_as(flat)
    smax = torch.full((n, flat.shape[1]), -math.inf, dtype=flat.dtype)
    smax = smax.scatter_reduce(0, idx, flat.detach(), reduce="amax", include_self=True)
    out = (flat - smax[index]).exp()
    ssum = torch.zeros((n, flat.shape[1]), dtype=flat.dtype).index_add(0, index, out) + 1e-16
    return (out / ssum[index]).reshape(src.shape)
