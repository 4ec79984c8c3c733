"""Development probe: launch-side (host) time of every section of the eager e2e step, no profiler in the way:
perf_counter wrappers around the autograd Functions' forward / backward and the batch / plan builders."""
import collections, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
T, C = collections.defaultdict(float), collections.defaultdict(int)


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
    from wheatley_b200 import _lib, batch, dense, edge_feats, functional, plan as plan_mod, ppo, readout, trunk
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.prefetch import BatchPrefetcher, HostReadback
    from wheatley_b200.workload import jssp_observation_batch
    torch.backends.cuda.matmul.allow_tf32 = True
    dev = torch.device("cuda", 0)
    n_j, n_m, B = 100, 20, 8
    torch.manual_seed(0)
    agent = JsspAgent(6 + n_m + 4, n_j * n_m, n_m, 64, 6, 4).to(dev)
    bucket = ppo.FlatGradAllReduce(list(agent.parameters()))
    opt = ppo.FlatAdam(agent.parameters(), lr=2e-4, bucket=bucket)
    obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B).items()}
    g = torch.Generator().manual_seed(1)
    masks = torch.rand(B, n_j * n_m, generator=g) < 0.3
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    ex = dict(masks=masks, action=action, adv=torch.randn(B, generator=g), ret=torch.randn(B, generator=g),
              old_lp=-torch.rand(B, generator=g) - 4, old_v=torch.randn(B, generator=g))
    ex = {k: v.contiguous().pin_memory() for k, v in ex.items()}
    kw = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.05, vf_coef=0.5)
    from wheatley_b200 import agent_step
    wrap(type(agent), "_step_inputs", "M.step_inputs")
    wrap(type(agent.gnn), "input_features", "M.input_features")
    wrap(type(agent.gnn), "edge_features", "M.edge_features")
    for cls, lab in ((agent_step._AgentStepFn, "F.step"), (trunk._TrunkFn, "F.trunk"), (dense._GeluMLPFused, "F.mlp"), (readout._Readout, "F.readout"),
                     (functional._MaskedHeadEval, "F.head"), (ppo._PpoLoss, "F.loss"),
                     (edge_feats._EdgeTables, "F.tables")):
        wrap(cls, "forward", lab + ".fwd", static=True)
        wrap(cls, "backward", lab + ".bwd", static=True)
    wrap(batch, "jssp_batch_on_device", "P.batch_build")
    wrap(plan_mod.GraphPlan, "__init__", "P.plan_init")
    wrap(plan_mod.GraphPlan, "prepare", "P.plan_prepare")
    wrap(plan_mod.GraphPlan, "finish", "P.plan_finish")
    wrap(plan_mod.CliqueView, "__init__", "P.clique_init")
    wrap(plan_mod.CliqueView, "finish", "P.clique_finish")
    wrap(plan_mod.CliqueView, "_finish", "P.clique__finish")
    wrap(plan_mod.GraphPlan, "sort_edge_ints", "P.sort_edge_ints")
    wrap(plan_mod.GraphPlan, "stats_host", "P.stats_host")
    wrap(plan_mod, "_orders_and_hubs", "P.orders_hubs")
    wrap(plan_mod, "_attach_hub_scratch", "P.hub_scratch")
    wrap(batch, "jssp_sizes", "P.jssp_sizes")
    wrap(torch.cuda.Event, "synchronize", "X.event_sync")
    wrap(type(agent.gnn), "_trunk_specs", "M.trunk_specs")
    wrap(edge_feats, "edge_tables", "M.edge_tables")
    wrap(ppo.FlatGradAllReduce, "flat_grads", "O.flat_grads")
    wrap(ppo.FlatAdam, "step", "O.adam_step")
    wrap(ppo.FlatGradAllReduce, "zero", "O.zero")
    wrap(type(agent.gnn), "layer_embeddings", "M.layer_embeddings")
    wrap(type(agent.gnn), "forward_split", "M.gnn")
    wrap(type(agent), "get_action_and_value", "M.agent_fwd")
    pf = BatchPrefetcher(dev)
    rb = HostReadback()
    pf.submit_observation(obs, n_j * n_m, ex, conflicts="clique", graph_pooling="learn")

    def step():
        t0 = time.perf_counter()
        d, _ = pf.get()
        pf.submit_observation(obs, n_j * n_m, ex, conflicts="clique", graph_pooling="learn")
        t1 = time.perf_counter()
        bucket.zero()
        loss, _ = agent.ppo_minibatch(d["batch"], d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"], **kw)
        t2 = time.perf_counter()
        loss.backward()
        t3 = time.perf_counter()
        bucket.reduce()
        opt.step()
        rb.push(loss)
        if len(rb) > 1:
            rb.pop()
        t4 = time.perf_counter()
        for k, a, b in (("S.prefetch", t0, t1), ("S.fwd", t1, t2), ("S.bwd", t2, t3), ("S.opt+read", t3, t4)):
            T[k] += b - a
            C[k] += 1

    import gc
    for _ in range(60):
        step()
    torch.cuda.synchronize()
    gc.collect(); gc.freeze()
    T.clear(); C.clear()
    n = 100
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    print(f"{(time.perf_counter() - t0) / n * 1e3:.2f} ms per step (prefetch + async read)")
    for k in sorted(T):
        print(f"  {k:22s} {T[k] / n * 1e3:7.3f} ms/step  ({C[k] / n:.1f} calls/step)")


if __name__ == "__main__":
    main()
