"""Development probe: where the host time of one eager end-to-end step goes (cProfile) and how long the
step takes with the dense path on tcgen05 vs the library GEMM."""
import cProfile, io, os, pstats, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import wheatley_b200 as wb
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.ppo import FlatGradAllReduce, ppo_like_loss
    from wheatley_b200.workload import jssp_batch
    torch.backends.cuda.matmul.allow_tf32 = True
    dev = torch.device("cuda", 0)
    n_j, n_m, B = 100, 20, 8
    torch.manual_seed(0)
    agent = JsspAgent(6 + n_m + 4, n_j * n_m, n_m, 64, 6, 4).to(dev)
    bucket = FlatGradAllReduce(list(agent.parameters()))
    opt = torch.optim.Adam(agent.parameters(), lr=2e-4, fused=True, capturable=True)
    hb = jssp_batch(n_j, n_m, B, seed0=0, progress=0.0)
    g = torch.Generator().manual_seed(1)
    masks = torch.rand(B, n_j * n_m, generator=g) < 0.3
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    host = dict(src=hb["src"], dst=hb["dst"], etype=hb["etype"], feat=hb["feat"], masks=masks, action=action,
                adv=torch.randn(B, generator=g), ret=torch.randn(B, generator=g))
    host = {k: v.contiguous().pin_memory() for k, v in host.items()}
    N, NT = hb["n_nodes"], hb["n_task_nodes"]

    def step():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        plan = wb.GraphPlan(d["src"], d["dst"], N)
        jb = JsspBatch(plan, d["etype"], d["feat"], B, n_j * n_m, NT, poolnodes=torch.arange(NT, NT + B, device=dev))
        bucket.zero()
        _, lp, ent, val, _ = agent.get_action_and_value(jb, d["action"], d["masks"])
        loss = ppo_like_loss(lp, ent, val, d["adv"], d["ret"])
        loss.backward()
        bucket.reduce()
        opt.step()
        return loss.item()

    for backend in ("tcgen05",):
        os.environ["WHEATLEY_B200_GEMM"] = backend
        for _ in range(4):
            step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            step()
        torch.cuda.synchronize()
        print(f"{backend}: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms per eager e2e step")
        # host-only time: launch everything, measure before the sync
        t0 = time.perf_counter()
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        plan = wb.GraphPlan(d["src"], d["dst"], N)
        jb = JsspBatch(plan, d["etype"], d["feat"], B, n_j * n_m, NT, poolnodes=torch.arange(NT, NT + B, device=dev))
        t1 = time.perf_counter()
        bucket.zero()
        _, lp, ent, val, _ = agent.get_action_and_value(jb, d["action"], d["masks"])
        loss = ppo_like_loss(lp, ent, val, d["adv"], d["ret"])
        t2 = time.perf_counter()
        loss.backward()
        t3 = time.perf_counter()
        torch.cuda.synchronize()
        t4 = time.perf_counter()
        print(f"  host: batch {1e3 * (t1 - t0):.2f} fwd {1e3 * (t2 - t1):.2f} bwd {1e3 * (t3 - t2):.2f} drain {1e3 * (t4 - t3):.2f} ms")
        pr = cProfile.Profile()
        pr.enable()
        step()
        pr.disable()
        sio = io.StringIO()
        pstats.Stats(pr, stream=sio).sort_stats("tottime").print_stats(28)
        print("\n".join(sio.getvalue().splitlines()[:50]))


if __name__ == "__main__":
    main()
