#!/usr/bin/env python
"""profiles/<tag>_traffic.json from a conv-kernel ncu summary (tools/ncu_summary.py full ...): DRAM bytes per conv
launch of the whole batch, keyed to the SHA-1 of the attention kernels' sources so that bench.py refuses a capture
made with other kernels (a committed constant goes stale silently).
    python tools/make_traffic.py profiles/r2s_conv_kernels_full.txt profiles/r2_traffic.json taillard100x20 8 64 4"""
import hashlib, json, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KERNEL_SOURCES = ("gat_fast.cu", "gat_win.cuh", "gat_rows.cuh", "gat_clique.cu", "gat_generic.cu", "gat_api.cu",
                  "gat_kernels.cuh", "common.cuh")


def kernel_source_sha():
    h = hashlib.sha1()
    for f in KERNEL_SOURCES:
        h.update(open(os.path.join(ROOT, "wheatley_b200", "csrc", f), "rb").read())
    return h.hexdigest()


def main():
    summary, out, workload, graphs, hidden, heads = sys.argv[1:7]
    rows = {}
    names = None
    for ln in open(summary):
        parts = [p.strip() for p in ln.split("|")]
        if parts[0] == "metric":
            names = parts[2:]
        elif parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[parts[1]]
            rows[parts[0]] = [float(x) * scale for x in parts[2:]]
    tot = [r + w for r, w in zip(rows["dram__bytes_read.sum"], rows["dram__bytes_write.sum"])]
    fwd = sum(t for n, t in zip(names, tot) if "fwd" in n)
    bwd = sum(t for n, t in zip(names, tot) if "bwd" in n)
    json.dump({"workload": workload, "graphs_per_gpu": int(graphs), "hidden": int(hidden), "heads": int(heads),
               "conv_fwd": int(fwd), "conv_bwd": int(bwd), "kernel_source_sha": kernel_source_sha(),
               "per_kernel": {n: int(t) for n, t in zip(names, tot)},
               "source": f"{summary} (ncu --set full --clock-control none, tools/gpu_ncu_r2.sh: the conv launches of "
                         "the last layer of a warm eager step)"}, open(out, "w"), indent=1)
    print(open(out).read())


if __name__ == "__main__":
    main()
