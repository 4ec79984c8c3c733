"""Development probe: host time (launch side only, no sync inside the step) of each section of one eager
end-to-end PPO-update step on the bench workload.  Sections are timed with perf_counter wrappers around the
modules' forward and the autograd Functions' forward / backward."""
import collections, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

T = collections.defaultdict(float)
C = collections.defaultdict(int)


def wrap(owner, name, label, static=False):
    fn = getattr(owner, name)

    def w(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            T[label] += time.perf_counter() - t0
            C[label] += 1
    setattr(owner, name, staticmethod(w) if static else w)


def main():
    import wheatley_b200 as wb
    from wheatley_b200 import dense, functional, eegatconv, mlp, plan as plan_mod, _lib
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.gnn_dgl import JsspBatch, GnnDGL
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

    for cls, lab in ((functional._FusedGATConv, "F.conv"), (dense._GeluMLPFused, "F.mlp"), (dense._Proj3, "F.proj3"),
                     (functional._MaskedHeadEval, "F.head"), (dense._GeluMLP, "F.mlp_percall")):
        wrap(cls, "forward", lab + ".fwd", static=True)
        wrap(cls, "backward", lab + ".bwd", static=True)
    wrap(eegatconv.EEGATConv, "forward", "M.conv_module")
    wrap(mlp.MLP, "forward", "M.mlp_module")
    wrap(GnnDGL, "node_embeddings", "M.trunk")
    wrap(_lib.lib, "b200gat_mlp_fwd", "C.mlp_fwd") if False else None

    from wheatley_b200.prefetch import BatchPrefetcher
    from wheatley_b200.workload import jssp_observation_batch
    obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B).items()}
    extras = {k: host[k] for k in ("masks", "action", "adv", "ret")}
    pf = BatchPrefetcher(dev, side_stream=os.environ.get("HB_SIDE", "0") == "1")
    mode = os.environ.get("HB_MODE", "serial").replace("bg", "")
    if mode == "obs":
        pf.submit_observation(obs, n_j * n_m, extras, conflicts="clique", graph_pooling="learn")
    elif mode == "coo":
        pf.submit(host, N)

    def step_pf(rec):
        t0 = time.perf_counter()
        d, plan = pf.get()
        t1 = time.perf_counter()
        if mode == "obs":
            pf.submit_observation(obs, n_j * n_m, extras, conflicts="clique", graph_pooling="learn")
            jb = d["batch"]
        else:
            pf.submit(host, N)
            jb = JsspBatch(plan, d["etype"], d["feat"], B, n_j * n_m, NT, poolnodes=torch.arange(NT, NT + B, device=dev))
        t2 = time.perf_counter()
        bucket.zero()
        _, lp, ent, val, _ = agent.get_action_and_value(jb, d["action"], d["masks"])
        t3 = time.perf_counter()
        loss = ppo_like_loss(lp, ent, val, d["adv"], d["ret"])
        t4 = time.perf_counter()
        loss.backward()
        t5 = time.perf_counter()
        bucket.reduce()
        opt.step()
        t6 = time.perf_counter()
        v = loss.item()
        t7 = time.perf_counter()
        if rec:
            for k, a, b in (("pf.get", t0, t1), ("pf.submit", t1, t2), ("forward", t2, t3), ("loss", t3, t4),
                            ("backward", t4, t5), ("reduce+adam", t5, t6), ("drain(item)", t6, t7),
                            ("TOTAL", t0, t7)):
                T["S." + k] += b - a
                C["S." + k] += 1
        return v

    def step(rec):
        if mode != "serial":
            return step_pf(rec)
        t0 = time.perf_counter()
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        t1 = time.perf_counter()
        plan = wb.GraphPlan(d["src"], d["dst"], N)
        jb = JsspBatch(plan, d["etype"], d["feat"], B, n_j * n_m, NT, poolnodes=torch.arange(NT, NT + B, device=dev))
        t2 = time.perf_counter()
        bucket.zero()
        _, lp, ent, val, _ = agent.get_action_and_value(jb, d["action"], d["masks"])
        t3 = time.perf_counter()
        loss = ppo_like_loss(lp, ent, val, d["adv"], d["ret"])
        t4 = time.perf_counter()
        loss.backward()
        t5 = time.perf_counter()
        bucket.reduce()
        opt.step()
        t6 = time.perf_counter()
        v = loss.item()
        t7 = time.perf_counter()
        if rec:
            for k, a, b in (("h2d", t0, t1), ("plan", t1, t2), ("forward", t2, t3), ("loss", t3, t4),
                            ("backward", t4, t5), ("reduce+adam", t5, t6), ("drain(item)", t6, t7),
                            ("TOTAL", t0, t7)):
                T["S." + k] += b - a
                C["S." + k] += 1
        return v

    import gc
    gc.collect(); gc.freeze()
    for _ in range(6):
        step(False)
    torch.cuda.synchronize()
    T.clear(); C.clear()
    n = 30
    for _ in range(n):
        step(True)
    for k in sorted(T):
        print(f"{k:22s} {1e6 * T[k] / n:9.1f} us/step  ({C[k] / n:.0f} calls/step, {1e6 * T[k] / max(C[k], 1):7.1f} us/call)")


if __name__ == "__main__":
    main()
