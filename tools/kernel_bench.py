#!/usr/bin/env python
"""Times the fused conv kernels alone (CUDA events, L2 flushed between launches) on a batched
Taillard graph; prints one JSON line per variant.  Development tool, not the bench contract."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nj", type=int, default=100)
    ap.add_argument("--nm", type=int, default=20)
    ap.add_argument("--graphs", type=int, default=8)
    ap.add_argument("--heads", type=int, default=4)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--flavour", default="egat", choices=["egat", "gatv2"])
    ap.add_argument("--no-order", action="store_true")
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    import wheatley_b200 as wb
    from wheatley_b200 import _lib
    from wheatley_b200.workload import jssp_batch
    dev = torch.device("cuda", 0)
    hb = jssp_batch(args.nj, args.nm, args.graphs)
    N, E = hb["n_nodes"], hb["n_edges"]
    plan = wb.GraphPlan(hb["src"].to(dev), hb["dst"].to(dev), N)
    if args.no_order:
        plan._c.dst_order = None
        plan._c.src_order = None
    H, D = args.heads, args.dim
    HD = H * D
    torch.manual_seed(0)
    Pi = torch.randn(N, HD, device=dev, requires_grad=True)
    Pj = torch.randn(N, HD, device=dev, requires_grad=True)
    V = torch.randn(N, HD, device=dev, requires_grad=True) if args.flavour == "egat" else None
    attn = (0.1 * torch.randn(1, H, D, device=dev)).requires_grad_(True)
    table = torch.randn(6, HD, device=dev, requires_grad=True)
    ef = wb.EdgeFeatures(_lib.EDGE_TABLE, table=table, etype_sorted=plan.sort_edge_ints(hb["etype"].to(dev)))
    R = torch.randn(N, HD, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def run():
        out, _ = wb.fused_gat_conv(plan, Pi, Pj, V, attn, heads=H, dim=D, slope=0.01 if V is not None else 0.2,
                                   edge_feats=ef)
        (out * R).sum().backward()

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    _lib.timing_enable(True)
    for _ in range(args.iters):
        flush.zero_()
        run()
    torch.cuda.synchronize()
    tm = _lib.timing_read()
    _lib.timing_enable(False)
    P = 3 if V is not None else 2
    row = N * HD * 4
    byt = {"conv_fwd": row * (P + 1) + E * (8 + 4 + 4 * H),
           "conv_bwd": row * (1 + 2 * P) + E * (8 + 4 + 4 * H)}
    f = tm["conv_fwd"][0] / max(tm["conv_fwd"][1], 1)
    bd = tm["conv_bwd_dst"][0] / max(tm["conv_bwd_dst"][1], 1)
    bs = tm["conv_bwd_src"][0] / max(tm["conv_bwd_src"][1], 1)
    print(json.dumps({"tag": args.tag, "flavour": args.flavour, "graphs": args.graphs, "N": N, "E": E,
                      "order": not args.no_order, "generic": os.environ.get("B200GAT_FORCE_GENERIC", "0"),
                      "fwd_ms": f, "bwd_dst_ms": bd, "bwd_src_ms": bs,
                      "fwd_GBps": byt["conv_fwd"] / f / 1e6, "bwd_GBps": byt["conv_bwd"] / (bd + bs) / 1e6,
                      "fwd_us_per_graph": 1e3 * f / args.graphs,
                      "fwdbwd_us_per_graph": 1e3 * (f + bd + bs) / args.graphs,
                      "edges_per_s_fwdbwd": E / ((f + bd + bs) * 1e-3)}), flush=True)


if __name__ == "__main__":
    main()
