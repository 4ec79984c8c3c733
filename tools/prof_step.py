"""Development probe: host-side cost of one eager end-to-end PPO-update step of the bench workload
(observation in pinned memory -> device batch build -> plan -> fused trunk -> heads -> PPO loss -> backward -> Adam).
Prints launch-side section times (no sync inside) and a cProfile of one step."""
import cProfile, io, os, pstats, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import wheatley_b200 as wb
    from wheatley_b200 import _lib
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.ppo import FlatAdam, FlatGradAllReduce, ppo_loss
    from wheatley_b200.workload import jssp_observation_batch
    torch.backends.cuda.matmul.allow_tf32 = True
    dev = torch.device("cuda", 0)
    n_j, n_m, B = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (100, 20, 8)))
    torch.manual_seed(0)
    agent = JsspAgent(6 + n_m + 4, n_j * n_m, n_m, 64, 6, 4).to(dev)
    bucket = FlatGradAllReduce(list(agent.parameters()))
    opt = FlatAdam(agent.parameters(), lr=2e-4, bucket=bucket)
    obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B).items()}
    g = torch.Generator().manual_seed(1)
    masks = torch.rand(B, n_j * n_m, generator=g) < 0.3
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    ex = dict(masks=masks, action=action, adv=torch.randn(B, generator=g), ret=torch.randn(B, generator=g),
              old_lp=-torch.rand(B, generator=g) - 4, old_v=torch.randn(B, generator=g))
    ex = {k: v.contiguous().pin_memory() for k, v in ex.items()}
    kw = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.05, vf_coef=0.5)
    T = {}

    def step(timing=False):
        t = [time.perf_counter()]
        d = {k: v.to(dev, non_blocking=True) for k, v in ex.items()}
        jb, _ = jssp_batch_from_observation(obs, dev, n_j * n_m, conflicts="clique", graph_pooling="learn")
        t.append(time.perf_counter())
        bucket.zero()
        loss, _ = agent.ppo_minibatch(jb, d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"], **kw)
        t.append(time.perf_counter())
        loss.backward()
        t.append(time.perf_counter())
        bucket.reduce()
        opt.step()
        t.append(time.perf_counter())
        v = loss.item()
        t.append(time.perf_counter())
        if timing:
            for k, a, b in zip(("batch+plan", "fwd", "bwd", "adam", "drain"), t[:-1], t[1:]):
                T[k] = T.get(k, 0.0) + (b - a) * 1e3
        return v

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    import gc
    gc.collect(); gc.freeze()
    n = 10
    t0 = time.perf_counter()
    for _ in range(n):
        step(True)
    torch.cuda.synchronize()
    print(f"{(time.perf_counter() - t0) / n * 1e3:.2f} ms per eager e2e step;  host sections (ms/step): "
          + ", ".join(f"{k} {v / n:.2f}" for k, v in T.items()))
    if os.environ.get("PROF", "cprofile") == "torch":
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            for _ in range(3):
                step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=70, max_name_column_width=70))
        return
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        step()
    pr.disable()
    sio = io.StringIO()
    pstats.Stats(pr, stream=sio).sort_stats("cumulative").print_stats(60)
    print("\n".join(l[:170] for l in sio.getvalue().splitlines()[:90]))
    sio = io.StringIO()
    pstats.Stats(pr, stream=sio).sort_stats("tottime").print_stats(30)
    print("\n".join(l[:170] for l in sio.getvalue().splitlines()[:50]))


if __name__ == "__main__":
    main()
