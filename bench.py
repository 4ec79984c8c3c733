#!/usr/bin/env python
"""bench.py - GNN fwd+bwd graph-edges/sec of the B200-native Wheatley hot path.

    python bench.py --gpus N --steps K --warmup W            # this implementation
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

A *step* is one PPO-minibatch pass of the hot path over one batch of synthetic schedule graphs:
batched-graph plan -> L x (projections, fused edge-attention conv, MLP) -> masked action head ->
loss -> fused backward -> gradient all-reduce (N > 1) -> Adam step.  Metric: graph edges of the
batch (all ranks) / step time.  Default workload = BASELINE.json north_star's target config:
Taillard 100x20 schedule graphs, DGL-flavour clique conflicts + learned pooling node
(2,001 nodes / 205,801 edges per graph), 8 graphs per GPU, 6 layers x 4 heads x 64.

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {  # name: (n_jobs, n_machines, default graphs per GPU)
    "taillard100x20": (100, 20, 8),
    "taillard15x15": (15, 15, 128),
    "taillard30x20": (30, 20, 64),
    "jssp10x10": (10, 10, 128),
    # PSP / PyG flavour = the reference's LIVE trainer (psp.train_psp): GATv2 layers, PspEdgeEmbedder
    # edge features, resource-node conflicts, learned pool node
    "psp272": (4, 4, 1536),                # 272.sh:1: --batch_size 1536 --hidden_dim 32 --n_layers 4 --conflicts node
    "psp_taillard100x20": (100, 20, 8),
}
PSP_DEFAULTS = {"psp272": dict(hidden=32, layers=4, conflicts="node"),
                "psp_taillard100x20": dict(hidden=64, layers=6, conflicts="node")}
METRIC = "gnn_fwd_bwd_graph_edges_per_sec"
UNIT = "edges/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="taillard100x20", choices=sorted(WORKLOADS))
    ap.add_argument("--graphs-per-gpu", type=int, default=0)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --graphs-per-gpu graphs on every rank; strong: --graphs graphs in total, split over "
                         "the ranks in contiguous shards balanced by edge count (ppo.shard_by_edges)")
    ap.add_argument("--graphs", type=int, default=0, help="strong scaling: graphs of the global minibatch")
    ap.add_argument("--layers", type=int, default=0, help="0 = the workload's default (6; 4 for psp272)")
    ap.add_argument("--hidden", type=int, default=0, help="0 = the workload's default (64; 32 for psp272)")
    ap.add_argument("--conflicts", default=None, help="PSP workloads: node (default) | clique")
    ap.add_argument("--heads", type=int, default=4)
    ap.add_argument("--progress", type=float, default=0.0, help="episode progress of the graph snapshot")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--gemm", default=None, help="override WHEATLEY_B200_GEMM (cublas|tcgen05)")
    ap.add_argument("--max-seconds", type=float, default=900.0, help="watchdog: hard exit after this long")
    ap.add_argument("--no-graph", action="store_true",
                    help="launch the resident-input step eagerly instead of replaying one captured CUDA graph")
    return ap.parse_args()


def model_config(args):
    n_j, n_m, b = WORKLOADS[args.workload]
    psp = args.workload.startswith("psp")
    dflt = PSP_DEFAULTS.get(args.workload, dict(hidden=64, layers=6, conflicts="clique"))
    return dict(n_j=n_j, n_m=n_m, graphs=args.graphs_per_gpu or b, input_dim=16 if psp else 6 + n_m + 4,
                max_n_nodes=n_j * n_m + (2 if psp else 0), layers=args.layers or dflt["layers"],
                hidden=args.hidden or dflt["hidden"], heads=args.heads, psp=psp,
                conflicts=(args.conflicts or dflt["conflicts"]) if psp else "clique")


PPO_KW = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.05, vf_coef=0.5)   # alg/ppo.py:85-102


def synth_targets(batch, n_nodes, seed):
    """Rollout-buffer side of one minibatch (alg/ppo.py:556-564): masks, actions, advantages, returns, old
    log-probabilities and old values."""
    g = torch.Generator().manual_seed(seed)
    masks = torch.rand(batch, n_nodes, generator=g) < 0.3
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    adv = torch.randn(batch, generator=g)
    ret = torch.randn(batch, generator=g)
    old_lp = -torch.rand(batch, generator=g) * 2.0 - 4.0
    old_v = torch.randn(batch, generator=g)
    return masks, action, adv, ret, old_lp, old_v


# ------------------------------------------------------------------------- reference / CPU arm
def build_oracle(cfg, seed=0):
    from oracle import trunk_oracle
    torch.manual_seed(seed)
    gnn = trunk_oracle.OracleGnnDGL(cfg["input_dim"], "learn", cfg["max_n_nodes"], cfg["n_m"], 3,
                                    cfg["layers"], cfg["hidden"], "gelu", cfg["heads"])
    return trunk_oracle.OracleAgent(gnn, cfg["input_dim"] + cfg["hidden"] * (cfg["layers"] + 1))


def cpu_step_fn_psp(cfg, n_graphs, progress, seed0=0):
    """PSP flavour on the host cores: the numpy batch oracle's graph (built once, outside the timed step, like the
    GPU arm's resident leg) through the oracle GnnMP trunk, heads and the PPO loss, fwd + bwd."""
    import numpy as np
    from oracle import head_oracle, ppo_oracle, psp_batch_oracle as pbo, trunk_oracle
    from wheatley_b200_workload import psp_taillard_graph
    graphs = [psp_taillard_graph(cfg["n_j"], cfg["n_m"], seed0 + b, progress=progress) for b in range(n_graphs)]
    b = pbo.psp_batch(graphs, cfg["conflicts"], "learn", True)
    torch.manual_seed(0)
    gnn = trunk_oracle.OracleGnnMP(16, "learn", cfg["max_n_nodes"], 3, cfg["layers"], cfg["hidden"], "gelu",
                                   cfg["heads"], conflicts=cfg["conflicts"])
    F = 16 + cfg["hidden"] * (cfg["layers"] + 1)
    vn, an = trunk_oracle.OracleMLP(3, F, 64, 1, False, "gelu"), trunk_oracle.OracleMLP(3, F, 64, 1, False, "gelu")
    params = list(gnn.parameters()) + list(vn.parameters()) + list(an.parameters())
    masks, action, adv, ret, old_lp, old_v = synth_targets(n_graphs, cfg["max_n_nodes"], 1)
    T = lambda a: torch.from_numpy(np.asarray(a))                                          # noqa: E731
    rn = b["resource_nodes"]
    args_ = (T(b["edge_index"]), T(b["type"]), T(b["att_rc"]).float(), T(b["att_rp"]).float(), T(b["feat"]).float(),
             n_graphs, b["num_nodes"], b["n_nodes"].tolist() if n_graphs > 1 else int(b["n_nodes"][0]),
             b["poolnodes"].tolist(), None if rn is None else rn.tolist(), T(b["res_cal_id"]).float())

    def step():
        for p in params:
            p.grad = None
        nf, ge = gnn(*args_)
        _, logprob, entropy, _ = head_oracle.action_and_logprob(an(nf).squeeze(-1), action, masks)
        loss, _ = ppo_oracle.ppo_loss(logprob, entropy, vn(ge), old_lp, adv, ret, **PPO_KW)
        loss.backward()
        return loss.item()
    return step, int(b["edge_index"].shape[1])


def cpu_step_fn(cfg, n_graphs, progress, seed0=0):
    """The oracle's fwd+bwd over ``n_graphs`` graphs of the workload on the host cores."""
    if cfg["psp"]:
        return cpu_step_fn_psp(cfg, n_graphs, progress, seed0)
    from wheatley_b200_workload import jssp_batch
    agent = build_oracle(cfg)
    b = jssp_batch(cfg["n_j"], cfg["n_m"], n_graphs, seed0, progress)
    from oracle import ppo_oracle
    masks, action, adv, ret, old_lp, old_v = synth_targets(n_graphs, cfg["max_n_nodes"], 1)
    args_ = (b["src"], b["dst"], b["etype"], b["feat"], n_graphs, cfg["max_n_nodes"], b["n_task_nodes"])

    def step():
        agent.zero_grad(set_to_none=True)
        logprob, entropy, value = agent.evaluate(args_, action, masks)
        loss, _ = ppo_oracle.ppo_loss(logprob, entropy, value, old_lp, adv, ret, **PPO_KW)
        loss.backward()
        return loss.item()
    return step, b["n_edges"]


def time_cpu(cfg, args, steps, warmup, budget_s=None):
    """Bounded CPU sample: graphs per step chosen so a step is ~1-3 s of host work."""
    torch.set_num_threads(os.cpu_count() or 1)
    n_graphs = 1 if cfg["n_j"] * cfg["n_m"] >= 1000 else min(cfg["graphs"], 256 if cfg["psp"] else 16)
    step, edges = cpu_step_fn(cfg, n_graphs, args.progress)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = (time.perf_counter() - t0) / done
    return dict(value=edges / dt, unit=UNIT, cores=torch.get_num_threads(), kind="port",
                sample=f"{done} steps x {n_graphs} graph(s) of {args.workload} ({edges} edges/step), "
                       f"oracle trunk fwd+bwd fp32 on CPU, {dt * 1e3:.1f} ms/step"), dt, n_graphs


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = model_config(args)
    cb, dt, n_graphs = time_cpu(cfg, args, args.steps, args.warmup, budget_s=150.0)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, cfg, n_graphs, note="CPU oracle port of the reference path "
                                      "(DGL/PyG not installable here); bounded sample"),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(json.dumps(line))


def workload_config(args, cfg, graphs, note=None):
    what = (f"PSP / PyG flavour (GATv2 layers, PspEdgeEmbedder), conflicts={cfg['conflicts']} + learned pool node"
            if cfg["psp"] else "DGL-flavour clique conflicts + learned pool node")
    c = {"workload": f"{args.workload} {what}, progress {args.progress}",
         "graphs_per_gpu": graphs, "nodes_per_graph": cfg["max_n_nodes"] + 1 + (cfg["n_m"] if cfg["psp"] and
                                                                               cfg["conflicts"] == "node" else 0),
         "layers": cfg["layers"],
         "hidden": cfg["hidden"], "heads": cfg["heads"], "mlp_layers": 3, "activation": "gelu"}
    if note:
        c["note"] = note
    return c


# ------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.f.read().splitlines():
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------- B200 arm
def algorithmic_bytes(kind, N, E, HD, H, P=3, T=4):
    """DESIGN.md 'Algorithmic bytes' (SURVEY 8d): fp32 node rows, int32 indices, type-table edge
    features, [E,H,D] edge output not materialised."""
    row = N * HD * 4
    if kind == "conv_fwd":            # read Pi,Pj,V ; write out ; per edge: src+dst ids, type, saved logits
        return row * (P + 1) + E * (8 + T + 4 * H)
    if kind == "conv_bwd":            # read dOut + P saved ; write P grads ; per edge ids, type, alpha
        return row * (1 + 2 * P) + E * (8 + T + 4 * H)
    raise KeyError(kind)


def run_b200(args):
    import torch.distributed as dist
    if args.gemm:
        os.environ["WHEATLEY_B200_GEMM"] = args.gemm
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: wheatley_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL writes its version / debug lines to stdout by default: keep stdout for the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    import wheatley_b200 as wb
    from wheatley_b200 import _lib
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.ppo import FlatAdam, FlatGradAllReduce, ppo_loss
    from wheatley_b200.workload import jssp_batch

    torch.backends.cuda.matmul.allow_tf32 = True       # the reference's matmul precision (psp/train_psp.py:47)
    from wheatley_b200 import dense as _dense
    dense_backend = ("tcgen05 TF32 (b200gat_gemm_tf32; library GEMM for uncovered shapes)"
                     if _dense.tf32_path_enabled() else "cublas TF32")
    cfg = model_config(args)
    B = cfg["graphs"]
    seed_base = rank * B                               # weak scaling: every rank its own B graphs
    if args.scaling == "strong":
        # one global minibatch of G graphs (BASELINE config 5: "minibatches sharded across 8 B200"); ranks take
        # contiguous shards of whole graphs balanced by edge count; the loss is evaluated as this rank's share of
        # the GLOBAL mean (ppo_minibatch(sharded=True): advantage moments all-reduced), the gradients are SUMMED
        from wheatley_b200.ppo import shard_by_edges
        from wheatley_b200.workload import jssp_observation as _one_obs
        G = args.graphs or cfg["graphs"] * 1
        if cfg["psp"]:
            per_graph = [1] * G
        else:
            per_graph = [int(_one_obs(cfg["n_j"], cfg["n_m"], g_, None, args.progress)["n_edges"]) for g_ in range(G)]
        lo, hi = shard_by_edges(per_graph, world)[rank]
        B, seed_base = hi - lo, lo
        if B <= 0:
            raise SystemExit(f"rank {rank}: empty shard ({G} graphs over {world} ranks)")
        cfg["graphs"] = B
    sharded = args.scaling == "strong" and world > 1
    torch.manual_seed(0)                               # identical replicas on every rank
    psp = cfg["psp"]
    if psp:
        from wheatley_b200.agent import PspAgent
        agent = PspAgent(cfg["max_n_nodes"], cfg["n_m"], hidden_dim=cfg["hidden"], n_layers=cfg["layers"],
                         n_heads=cfg["heads"], conflicts=cfg["conflicts"]).to(dev)
    else:
        agent = JsspAgent(cfg["input_dim"], cfg["max_n_nodes"], cfg["n_m"], cfg["hidden"], cfg["layers"],
                          cfg["heads"]).to(dev)
    bucket = FlatGradAllReduce(list(agent.parameters()))
    # Adam of alg/ppo.py on ONE flat parameter / the all-reduced flat gradient (ppo.FlatAdam)
    opt = FlatAdam(agent.parameters(), lr=2e-4, bucket=bucket)

    masks, action, adv, ret, old_lp, old_v = synth_targets(B, cfg["max_n_nodes"], 1 + seed_base)
    extras = dict(masks=masks, action=action, adv=adv, ret=ret, old_lp=old_lp, old_v=old_v)
    extras = {k: v.contiguous().pin_memory() for k, v in extras.items()}
    from wheatley_b200.prefetch import BatchPrefetcher
    if psp:
        # the e2e / rollout legs start from the wire-format observation slots the env workers write
        # (wheatley_b200.wire); the batched graph is built on the device (wheatley_b200.psp_batch)
        from wheatley_b200.wire import PspArena
        from wheatley_b200.workload import psp_taillard_graph
        arena = PspArena.from_graphs([psp_taillard_graph(cfg["n_j"], cfg["n_m"], seed_base + b, progress=args.progress)
                                      for b in range(B)])
        h2d_bytes = (arena.tensor[:B].numel() * arena.tensor.element_size() + 5 * (B + 1) * 8
                     + sum(v.numel() * v.element_size() for v in extras.values()))
        input_note = ("arena of wire-format observation slots (wheatley_b200.wire: features, resources, precedence / "
                      "resource-priority edges per graph) + masks / actions / advantages / returns / old log-probs / "
                      "old values in pinned host memory; self loops, reverse edges, resource / pool nodes and their "
                      "wiring are built on the device by b200gat_psp_batch_build")

        def build_batch():
            return agent.preprocess(arena)

        def submit(pf):
            pf.submit_psp(agent, arena, extras)
        probe = build_batch()
        N, E, NT = int(probe[0].ndata("feat").shape[0]), int(probe[0].edge_index.shape[1]), int(probe[2])
        del probe
    else:
        # the e2e / rollout legs start from the padded gym observation (what the reference's batching receives,
        # jssp/models/agent_observation.py:127-131): the batched graph is built on the device (wheatley_b200.batch)
        from wheatley_b200.workload import jssp_observation_batch
        from wheatley_b200.batch import jssp_batch_from_observation
        obs = {k: v.contiguous().pin_memory() for k, v in
               jssp_observation_batch(cfg["n_j"], cfg["n_m"], B, seed0=seed_base, progress=args.progress).items()}
        builder_kw = dict(conflicts="clique", graph_pooling="learn")
        h2d_bytes = (sum(v.numel() * v.element_size() for v in obs.values())
                     + sum(v.numel() * v.element_size() for v in extras.values()) + 4 * (B + 1) * 8)
        input_note = ("padded gym observation (features, precedence edge_index, n_nodes, n_edges) + masks / actions / "
                      "advantages / returns / old log-probs / old values in pinned host memory; the batched graph "
                      "(clique conflict edges, pool wiring) is built on the device by b200gat_jssp_batch_build")

        def build_batch():
            return jssp_batch_from_observation(obs, dev, cfg["max_n_nodes"], **builder_kw)[0]

        def submit(pf):
            pf.submit_observation(obs, cfg["max_n_nodes"], extras, **builder_kw)
        probe = build_batch()
        N, E, NT = int(probe.plan.n_nodes), int(probe.plan.n_edges), int(probe.num_task_nodes)
        del probe

    def to_dev():
        return {k: v.to(dev, non_blocking=True) for k, v in extras.items()}

    def train_step(jb, d):
        bucket.zero()
        # alg/ppo.py:575-711 as one call: get_action_and_value + the clipped-PPO minibatch loss (statistics stay on
        # the device); loss.backward() is the whole backward pass
        loss, _stats = agent.ppo_minibatch(jb, d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"],
                                           sharded=sharded, **PPO_KW)
        loss.backward()
        bucket.reduce(average=not sharded)     # sharded: the loss is already this rank's share of the global mean
        opt.step()
        return loss

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, steps, warmup, sampler=False, graphed=False, kernel_timing=False):
        # nvidia-smi needs ~100 ms to start: launch it before the warm-up so that it samples the timed region
        smp = ClockSampler(local) if sampler and rank == 0 else None
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if kernel_timing:
            _lib.timing_enable(True)                       # per-kernel events over the timed steps only
        l0 = _lib.launch_count()
        total = 0.0
        evs, host_ms = [], []
        ms0 = torch.cuda.memory_stats(dev)
        allocs0, frees0 = ms0.get("num_device_alloc", 0), ms0.get("num_device_free", 0)
        reserved0 = torch.cuda.memory_reserved(dev)
        for _ in range(steps):
            flush.zero_()                                  # L2 flush, outside the event pair
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            th = time.perf_counter()
            e0.record()
            fn()
            e1.record()
            host_ms.append((time.perf_counter() - th) * 1e3)
            evs.append((e0, e1))
            if graphed:                                    # read the graph's per-kernel event pairs of this replay
                torch.cuda.synchronize()
                _lib.lib.b200gat_timing_collect_graph()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        clocks = smp.stop() if smp else None
        raw = [a.elapsed_time(b) for a, b in evs]
        per_step = sorted(raw)
        imax = raw.index(per_step[-1])
        timed.last_spread = {"min": per_step[0], "median": per_step[len(per_step) // 2], "max": per_step[-1],
                             "max_at_step": imax, "host_ms_of_that_step": host_ms[imax],
                             "host_ms_median": sorted(host_ms)[len(host_ms) // 2],
                             "cudaMalloc_calls_in_timed_region":
                                 torch.cuda.memory_stats(dev).get("num_device_alloc", 0) - allocs0,
                             "cudaFree_calls_in_timed_region":
                                 torch.cuda.memory_stats(dev).get("num_device_free", 0) - frees0,
                             "reserved_mib_start_end": [reserved0 >> 20, torch.cuda.memory_reserved(dev) >> 20]}
        total = sum(per_step)
        launches = _lib.launch_count() - l0
        t = torch.tensor([total], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) / steps, launches // steps, clocks

    # ---- value: inputs resident in HBM, plan prebuilt (topology is layer- and epoch-invariant)
    d_res = to_dev()
    jb_res = build_batch()
    torch.cuda.synchronize()
    step_fn, graph, graph_note = (lambda: train_step(jb_res, d_res)), None, "eager launches"
    tgraph, captured = None, 0
    if not args.no_graph:
        # the whole step (fwd + head + loss + bwd + all-reduce + Adam) as ONE CUDA graph: shapes and the plan
        # are fixed for resident inputs, so replaying it removes the per-launch host cost of ~350 kernels.
        # Two captures: a plain one, replayed for `value`, and an instrumented one (every library launch
        # bracketed by external event nodes) replayed afterwards for the per-kernel times of the roofline.
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    train_step(jb_res, d_res)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            l0 = _lib.launch_count()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, capture_error_mode="relaxed"):
                train_step(jb_res, d_res)
            captured = _lib.launch_count() - l0
            _lib.timing_enable(True)                       # captured launches carry external event pairs
            tgraph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(tgraph, capture_error_mode="relaxed"):
                train_step(jb_res, d_res)
            _lib.timing_enable(False)
            step_fn, graph_note = graph.replay, "one CUDA graph replay per step"
        except Exception as exc:                           # keep the eager path measurable
            graph, tgraph = None, None
            graph_note = f"eager launches (graph capture failed: {type(exc).__name__}: {exc})"[:200]
            _lib.lib.b200gat_timing_clear_graph()
            torch.cuda.synchronize()
    ms_dev, launches, clocks = timed(step_fn, args.steps, args.warmup, sampler=True, graphed=False,
                                     kernel_timing=graph is None)
    if graph is not None:
        launches = captured
        # per-kernel times: replay the instrumented graph (not part of `value`)
        _lib.timing_enable(True)
        ms_instr, _, _ = timed(tgraph.replay, args.steps, 1, sampler=False, graphed=True, kernel_timing=False)
    tm = _lib.timing_read()
    _lib.timing_enable(False)

    # ---- the same step with fp32-accurate matmuls (the JSSP entry point of the reference never sets
    # torch.set_float32_matmul_precision("high"), only psp/train_psp.py:47 does): allow_tf32 = False switches
    # b200gat_gemm_tf32 to 3xTF32 split arithmetic (three tcgen05.mma per k step, fp32 matmul accuracy: the
    # configuration the 1e-4 parity tests run).  Same fused step, eager launches.  Reported beside `value`.
    ms_fp32 = None
    if os.environ.get("BENCH_SKIP_FP32_DENSE", "0") != "1":
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            ms_fp32, _, _ = timed(lambda: train_step(jb_res, d_res), max(3, args.steps // 2), 3)
        finally:
            torch.backends.cuda.matmul.allow_tf32 = True

    # The eager legs below are host-bound (~350 launches per step): a full cyclic-GC pass over the ~10^6 long-lived
    # objects of torch + this process costs 10-40 ms and showed up as 2-4 outlier steps in 20.  Freeze what exists
    # (standard practice for latency-sensitive loops); objects created from here on are still collected.
    import gc
    gc.collect()
    gc.freeze()
    # ... and hand the caching allocator one large free block to carve the eager legs' tensors from: its pools keep
    # growing for the first dozens of steps (a cudaMalloc costs 0.2 ms - tens of ms of host time inside a step)
    reserve = torch.empty(6 << 30, dtype=torch.uint8, device=dev)
    del reserve

    # ---- e2e: host buffers in, loss out, through the public module API, every step
    def e2e_step_serial():
        d = {k: v.to(dev, non_blocking=True) for k, v in extras.items()}
        return train_step(build_batch(), d).item()         # .item(): device->host read of the loss
    E2E_WARMUP = max(args.warmup, 20)     # the caching allocator still grows (cudaMalloc: 5-200 ms) for ~15 steps
    timed(e2e_step_serial, max(3, args.steps), E2E_WARMUP)
    ms_e2e_serial = timed.last_spread["median"]

    # The update loop as a user writes it with wheatley_b200.prefetch: every step still copies ITS inputs from
    # pinned host memory and reads ITS loss back, but the copy + plan build of step i+1 are queued on a side
    # stream while step i trains (the copy engine and the plan kernels run under the step's kernels, and the
    # plan's device->host read of its counts no longer waits on an idle GPU).  K timed steps = K copies.
    from wheatley_b200.prefetch import HostReadback
    DEPTH = 2                      # batches / loss reads in flight: the launch thread may run two steps ahead of the
                                   # GPU, so a short stall of the (shared) host does not idle the device

    def run_e2e(side):
        pf = BatchPrefetcher(dev, side_stream=side)
        for _ in range(DEPTH):
            submit(pf)
        rb = HostReadback(depth=DEPTH + 1)
        losses = []

        def e2e_step():
            d, _ = pf.get()
            submit(pf)                                         # next step's inputs
            rb.push(train_step(d["batch"], d))                 # device->host read of THIS step's loss, asynchronous
            if len(rb) > DEPTH:
                losses.append(float(rb.pop()))                 # ... consumed DEPTH steps later (HostReadback)
        mean, _, _ = timed(e2e_step, max(3, args.steps), E2E_WARMUP)
        spread = dict(timed.last_spread)
        while len(rb):
            losses.append(float(rb.pop()))
        assert len(losses) == max(3, args.steps) + E2E_WARMUP and all(x == x for x in losses)   # every loss read
        for _ in range(DEPTH):
            pf.get()                                           # drain the unused prefetches
        torch.cuda.synchronize()
        return mean, spread

    # Placement of the prefetched copies / builds is the user's choice (BatchPrefetcher(side_stream=...)): on a side
    # stream they run under the previous step's kernels (wins when the step is GPU-bound: 4.6 vs 5.2 ms at 8 x Taillard
    # 100x20), at the price of record_stream bookkeeping per tensor on the host (loses when the loop is host-bound:
    # 6.8 vs 5.6 ms at psp272).  Both are measured; the better one is the headline, both are reported.
    forced = os.environ.get("BENCH_PREFETCH_SIDE_STREAM")
    e2e_runs = {}
    for side in ([forced == "1"] if forced in ("0", "1") else [False, True]):
        e2e_runs[side] = run_e2e(side)
    choice = torch.tensor([min(e2e_runs, key=lambda k: e2e_runs[k][0])], device=dev, dtype=torch.int32)
    if world > 1:
        dist.broadcast(choice, 0)                              # every rank reports the placement rank 0 picked
    SIDE = bool(int(choice))
    ms_e2e_mean, e2e_spread = e2e_runs[SIDE]
    e2e_by_placement = {("side_stream" if k else "same_stream"): v[0] for k, v in e2e_runs.items()}
    e2e_spread["other_placement"] = {("side_stream" if k else "same_stream"): v[1]
                                     for k, v in e2e_runs.items() if k != SIDE}
    # every step is bracketed by its own event pair: the headline is the MEAN over the timed steps; the median
    # (slowest rank's), min and max are reported beside it in ms_per_step_median / ms_per_step_spread
    t_med = torch.tensor([e2e_spread["median"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_med, op=dist.ReduceOp.MAX)
    ms_e2e_median = float(t_med)
    ms_e2e = ms_e2e_mean                                   # headline = MEAN over the timed steps (max over ranks)

    # ---- PPO env-steps/sec (BASELINE.json metric, second half): a TIMED PPO iteration of alg/ppo.py on this workload,
    # per rank, with an in-process dummy environment (the CPU simulators jssp/env, psp/env are out of scope, SURVEY
    # 8f N2: the dummy env returns the same observation arrays and a seeded reward every step):
    #   rollout  (alg/ppo.py:157-206): T steps x num_envs graphs: host observation -> device batch build -> trunk +
    #            heads under no_grad -> sampled action read back to the host (env.step needs it) -> reward / done in
    #   GAE      (:276-299) on the device (b200gat_gae)
    #   update   (:539-728): n_epochs passes over the T minibatches (one rollout step's graphs each, shuffled): batch
    #            build, agent.ppo_minibatch, backward, gradient all-reduce, Adam; the statistics of an epoch are read
    #            back ONCE for the KL early stop (:741-749), all-reduced so that every rank stops together
    from wheatley_b200.ppo import gae as device_gae
    T_roll, n_epochs = 8, 3
    rng_env = torch.Generator().manual_seed(7 + rank)

    def rollout_step():
        masks_d = extras["masks"].to(dev, non_blocking=True)
        jb = build_batch()
        with torch.no_grad():
            action, logprob, _, value, _ = agent.get_action_and_value(jb, None, masks_d)
        return action.cpu(), logprob, value.reshape(-1)

    pf_upd = BatchPrefetcher(dev, side_stream=SIDE)      # update phase: the placement the e2e leg found better

    def ppo_iteration():
        lps, vals, acts, rews = [], [], [], []
        for _t in range(T_roll):
            a_host, lp, v = rollout_step()                               # device->host sync, like alg/ppo.py:206
            rews.append(torch.rand(B, generator=rng_env).pin_memory().to(dev, non_blocking=True))   # env.step
            lps.append(lp); vals.append(v); acts.append(a_host.pin_memory().to(dev, non_blocking=True))
        rewards = torch.stack(rews).unsqueeze(-1)
        values = torch.stack(vals).unsqueeze(-1)
        dones = torch.zeros((T_roll, B), device=dev)
        _, _, next_v = rollout_step()
        adv, ret = device_gae(rewards, values, dones, next_v.unsqueeze(-1), torch.ones(B, device=dev), 1.0, 1.0)
        masks_d = extras["masks"].to(dev, non_blocking=True)
        stop = False
        for _ep in range(n_epochs):
            kl = torch.zeros(1, device=dev)
            order = torch.randperm(T_roll, generator=rng_env).tolist()
            submit(pf_upd)                                               # the epoch's first minibatch
            for i, mb in enumerate(order):
                d_mb, _ = pf_upd.get()
                if i + 1 < len(order):
                    submit(pf_upd)                                       # next minibatch's copies / batch / plan build
                bucket.zero()
                loss, stats = agent.ppo_minibatch(d_mb["batch"], acts[mb], masks_d, lps[mb], adv[mb, :, 0],
                                                  ret[mb, :, 0], values[mb, :, 0], sharded=sharded, **PPO_KW)
                loss.backward()
                bucket.reduce(average=not sharded)
                opt.step()
                kl += stats[3]
            if world > 1:
                dist.all_reduce(kl)
            stop = float(kl) / (T_roll * world) > 1e9                    # one read-back per epoch (target_kl check)
            if stop:
                break
        return T_roll * B

    ppo_iteration()                                                      # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_ppo = time.perf_counter()
    n_it = 2
    env_steps = sum(ppo_iteration() for _ in range(n_it))
    torch.cuda.synchronize()
    t_ppo = time.perf_counter() - t_ppo
    timed(lambda: rollout_step()[0], max(3, args.steps), 5)
    ms_roll = timed.last_spread["median"]
    ppo_all = torch.tensor([env_steps / t_ppo], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ppo_all)

    edges_all = torch.tensor([E], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(edges_all)
    edges_total = float(edges_all)
    # tear the graphs down before anything else: a live CUDA graph holding captured NCCL kernels makes the
    # communicator teardown block (seen at N=2: every rank printed / finished, then hung in shutdown)
    step_fn = None
    if graph is not None:
        graph.reset()
        tgraph.reset()
        _lib.lib.b200gat_timing_clear_graph()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    if rank != 0:
        finish(world)
        return

    import json as _json
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = _json.load(open(pk))
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    HD, H = cfg["hidden"] * cfg["heads"], cfg["heads"]
    # one "launch" of the conv pass = the sparse kernel(s) on the residual CSR + the dense clique kernel(s)
    fwd_n, bd_n = tm["conv_fwd"][1], tm["conv_bwd_dst"][1]
    fwd_ms = tm["conv_fwd"][0] + tm.get("clq_fwd", (0.0, 0))[0]
    bwd_ms = (tm["conv_bwd_dst"][0] + tm["conv_bwd_src"][0] + tm.get("clq_bwd_mma", (0.0, 0))[0]
              + tm.get("clq_bwd_elem", (0.0, 0))[0])
    per = {}
    # P projected node tensors (3 EGAT, 2 GATv2), T bytes of edge features per edge (type id; + 5 fp32
    # attributes for the PSP edge embedder)
    PT = dict(P=2, T=24) if psp else dict(P=3, T=4)
    if fwd_n:
        per["conv_fwd"] = (fwd_ms / fwd_n, algorithmic_bytes("conv_fwd", N, E, HD, H, **PT))
    if bd_n:
        per["conv_bwd"] = (bwd_ms / bd_n, algorithmic_bytes("conv_bwd", N, E, HD, H, **PT))
    dom = max(per, key=lambda k: per[k][0] * (fwd_n if k == "conv_fwd" else bd_n)) if per else None
    ms_share = ms_instr if graph is not None else ms_dev
    share_note = (f"per-kernel CUDA events inside an instrumented replay of the same graph ({ms_share:.3f} ms/step; "
                  "the event nodes serialise launches, `value` is timed on the plain graph)"
                  if graph is not None else "per-kernel CUDA events in the timed eager steps")
    roofline = None
    # DRAM traffic of the kernel class per launch, from the committed ncu --set full capture of this workload -
    # accepted only when it was made with the attention kernels this run uses (SHA-1 of their sources)
    traffic, traffic_src = None, None
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from make_traffic import kernel_source_sha
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            tr = json.load(f)
        same = (tr["workload"], tr["graphs_per_gpu"], tr["hidden"], tr["heads"]) == (args.workload, B, cfg["hidden"],
                                                                                 cfg["heads"])
        if same and dom in tr and tr.get("kernel_source_sha") == kernel_source_sha():
            traffic, traffic_src = tr[dom], tr["source"]
        elif same:
            traffic_src = "stale: profiles/r2_traffic.json was captured with other kernel sources"
    except (OSError, KeyError, ValueError, ImportError):
        pass
    if dom:
        ms, byt = per[dom]
        ach = byt / (ms * 1e-3) / 1e9
        roofline = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                    "frac": ach / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": byt, "avg_launch_ms": ms,
                    "kernels": {k: {"avg_ms": v[0], "algorithmic_bytes": v[1],
                                    "gbps": v[1] / (v[0] * 1e-3) / 1e9} for k, v in per.items()},
                    "share_of_step": {k: (t_ms / args.steps) / ms_share for k, (t_ms, _) in tm.items() if t_ms},
                    "share_basis": share_note}
        # The second roofline DESIGN.md section 5 states beside the HBM one: the logit is non-linear inside the sum, so
        # every feature-edge (E x H x D) costs >= 3 fp32 issue slots forward (add, max, fma; the aggregation is on the
        # tensor cores) and >= 6 backward; 148 SMs x 4 schedulers x 32 lanes issue per clock.
        clk = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
        fe = float(E) * cfg["heads"] * cfg["hidden"]
        lanes = 148 * 4 * 32
        floors = {"conv_fwd": 3 * fe / lanes / clk * 1e3, "conv_bwd": 6 * fe / lanes / clk * 1e3}
        roofline["fp32_issue_bound"] = {
            "model": "E*H*D feature-edges x (3 forward | 6 backward) fp32 issue slots / (148 SMs x 4 schedulers x 32 "
                     "lanes x measured SM clock), per conv launch group (one layer, this rank's batch)",
            "floor_ms": floors,
            "frac": {k: floors[k] / per[k][0] for k in floors if k in per}}
    line = {"metric": METRIC, "value": edges_total / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32 (attention kernels fp32; dense projections TF32 = the matmul precision of the reference's "
                     "live PSP trainer, psp/train_psp.py:47; its JSSP entry point runs fp32 matmuls: see "
                     "fp32_dense for that configuration)",
            "fp32_dense": None if ms_fp32 is None else {
                "ms_per_step": ms_fp32, "value": edges_total / (ms_fp32 * 1e-3), "unit": UNIT,
                "note": "same step, torch.backends.cuda.matmul.allow_tf32 = False: the tcgen05 GEMM runs 3xTF32 split "
                        "arithmetic (B200GAT_GEMM_FP32X3, fp32 matmul accuracy; eager launches): the configuration "
                        "the 1e-4 parity tests run"},
            "data": "synthetic",
            "config": dict(workload_config(args, cfg, B), edges_per_gpu=E, nodes_per_gpu=N,
                           l2="flushed between timed steps (256 MiB write)",
                           parallelism=f"dp{world}", gemm=dense_backend,
                           step="plan reuse + trunk fwd + masked head + clipped-PPO loss (alg/ppo.py:582-705) + bwd + grad allreduce "
                                "+ Adam",
                           launch=graph_note),
            "e2e": {"value": edges_total / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "ms_per_step_median": ms_e2e_median, "ms_per_step_spread": e2e_spread,
                    "statistic": "mean of the per-step CUDA-event times over the timed steps (max over ranks; rank 0 "
                                 "spread shown)",
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
                    "input": input_note,
                    "pipeline": "wheatley_b200.prefetch.BatchPrefetcher(side_stream=%s): step i+1's host->device "
                                "copy, batch and plan build are queued %s; every timed step copies its inputs "
                                % (SIDE, "on a side stream and run under step i's kernels" if SIDE
                                   else "ahead of step i's kernels on the same stream") +
                                "from pinned host memory and reads its loss back (wheatley_b200.prefetch.HostReadback: "
                                "asynchronous copy into pinned memory, consumed one step later, so the launch thread "
                                "never waits on the GPU; ms_per_step_no_prefetch is the serial loop with loss.item())",
                    "ms_per_step_no_prefetch": ms_e2e_serial, "ms_per_step_by_prefetch_placement": e2e_by_placement,
                    "warmup_steps": E2E_WARMUP},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "ppo": {"env_steps_per_sec": float(ppo_all), "unit": "env-steps/s", "num_envs_per_gpu": B,
                    "minibatch_graphs": B, "n_epochs": n_epochs, "rollout_steps": T_roll,
                    "rollout_ms_per_step": ms_roll, "iteration_ms": t_ppo / n_it * 1e3,
                    "definition": "TIMED loop, wall clock: (rollout of rollout_steps x num_envs env steps with the "
                                  "sampled action read back every step, GAE on the device, n_epochs x rollout_steps "
                                  "minibatch updates, the next minibatch's batch build prefetched, with one "
                                  "statistics read-back per epoch) / elapsed; dummy "
                                  "in-process environment, CPU simulators excluded; all ranks summed",
                    "rollout_graph_edges_per_sec": edges_per_rank_note(E, ms_roll, world)}}
    if world == 1 and not args.no_cpu_baseline:
        cb, _, _ = time_cpu(cfg, args, 1000, 1, budget_s=args.cpu_seconds)
        line["cpu_baseline"] = cb
    emit(json.dumps(line))
    finish(world)


def edges_per_rank_note(edges, ms, world):
    """Forward-only (inference) throughput of the rollout step, all ranks (same shard size on every rank)."""
    return world * edges / (ms * 1e-3)


_JSON_FD = None


def guard_stdout():
    """stdout carries exactly one JSON line.  Libraries write there behind Python's back (NCCL prints its version
    line to fd 1 on some boxes whatever NCCL_DEBUG_FILE says): point fd 1 at stderr for the run and keep the
    real stdout for emit()."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(text):
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (text + "\n").encode())


def finish(world):
    """End of a rank's run.  Multi-rank runs leave through os._exit after flushing: the NCCL communicator
    teardown at interpreter shutdown has nothing left to do for the bench and is the one place a finished run
    could still block."""
    sys.stdout.flush()
    sys.stderr.flush()
    if world > 1:
        os._exit(0)


def start_watchdog(seconds):
    """A bench run that has not finished after ``seconds`` is stuck: exit instead of holding the GPUs."""
    import threading

    def fire():
        sys.stderr.write(f"bench.py: watchdog fired after {seconds} s, exiting\n")
        sys.stderr.flush()
        os._exit(3)
    t = threading.Timer(seconds, fire)
    t.daemon = True
    t.start()


def main():
    args = parse()
    guard_stdout()
    start_watchdog(args.max_seconds)
    # the workload generator is plain numpy: load it by path so the reference arm does not need
    # the CUDA library (wheatley_b200/__init__ refuses to import without it)
    import importlib.util
    spec = importlib.util.spec_from_file_location("wheatley_b200_workload",
                                                  os.path.join(ROOT, "wheatley_b200", "workload.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["wheatley_b200_workload"] = mod
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
