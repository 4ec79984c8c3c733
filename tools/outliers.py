"""Development probe: per-step wall time of the prefetched e2e loop over many steps, with the caching allocator's
device-allocation counter, to find what the occasional 20-200 ms steps are."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from wheatley_b200 import ppo
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
    pf, rb = BatchPrefetcher(dev), HostReadback()
    pf.submit_observation(obs, n_j * n_m, ex, conflicts="clique", graph_pooling="learn")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    import gc
    gc.collect(); gc.freeze()
    rows = []
    for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 300):
        flush.zero_()
        st0 = torch.cuda.memory_stats()
        t0 = time.perf_counter()
        d, _ = pf.get()
        t1 = time.perf_counter()
        pf.submit_observation(obs, n_j * n_m, ex, conflicts="clique", graph_pooling="learn")
        t2 = time.perf_counter()
        bucket.zero()
        loss, _ = agent.ppo_minibatch(d["batch"], d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"], **kw)
        t3 = time.perf_counter()
        loss.backward()
        t4 = time.perf_counter()
        opt.step()
        rb.push(loss)
        if len(rb) > 1:
            rb.pop()
        t5 = time.perf_counter()
        st1 = torch.cuda.memory_stats()
        rows.append((1e3 * (t5 - t0), 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), 1e3 * (t5 - t4),
                     st1["num_device_alloc"] - st0["num_device_alloc"], st1["num_alloc_retries"] - st0["num_alloc_retries"],
                     st1["reserved_bytes.all.current"] >> 20))
    torch.cuda.synchronize()
    tot = sorted(r[0] for r in rows[20:])
    print(f"steps {len(rows)}  median {tot[len(tot)//2]:.2f}  mean {sum(tot)/len(tot):.2f}  p99 {tot[int(len(tot)*0.99)]:.2f}  max {tot[-1]:.2f} ms; reserved {rows[-1][8]} MiB")
    print("outliers (> 2x median): step total get submit fwd bwd opt+read dev_allocs retries reserved")
    med = tot[len(tot) // 2]
    for i, r in enumerate(rows):
        if i >= 20 and r[0] > 2 * med:
            print(i, " ".join(f"{x:.2f}" if isinstance(x, float) else str(x) for x in r))
    print("first 20 steps:", " ".join(f"{r[0]:.1f}/{r[6]}" for r in rows[:20]))


if __name__ == "__main__":
    main()
