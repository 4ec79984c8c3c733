/* b200gat - C ABI of the B200-native graph-attention hot path for Wheatley.
 *
 * Wheatley (jolibrain/wheatley @ 7143392) is pure Python: the "FFI" of this path in
 * the reference is the set of third-party ops its layers dispatch to (SURVEY.md 2.2,
 * K1-K9).  Each entry point below replaces one of those dispatch sites; the reference
 * file:line is cited per function.  INTEGRATION.md shows the ctypes binding a Wheatley
 * maintainer adds (wheatley_b200/_lib.py is that binding).
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless stated otherwise; tensors are dense,
 *    row-major, fp32 / int32 (int64 only where the reference hands over int64 COO);
 *  - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, no call
 *    synchronises the device;
 *  - return value: 0 on success, a B200GAT_ERR_* code otherwise; b200gat_last_error()
 *    returns a host string describing the last failure of the calling thread;
 *  - an edge e goes src[e] -> dst[e]; messages are reduced at dst.
 */
#ifndef B200GAT_H
#define B200GAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200GAT_OK 0
#define B200GAT_ERR_INVALID 1     /* bad argument (null pointer, negative size, ...)      */
#define B200GAT_ERR_CUDA 2        /* a CUDA runtime call / launch failed                   */
#define B200GAT_ERR_UNSUPPORTED 3 /* shape not supported by any kernel variant             */

#define B200GAT_FLAVOUR_ADDITIVE 0 /* EGAT / GATv2: s = attn . leaky_relu(Pi[u]+Pj[v]+Pe) */
#define B200GAT_FLAVOUR_DOT 1      /* GATv3: s = (w_c . (Pj[v]*Pi[u]) + b_c) / sqrt(D)    */

#define B200GAT_EDGE_NONE 0
#define B200GAT_EDGE_TABLE 1 /* Pe[e] = table[etype[e]]   (edge features are an Embedding(type)) */
#define B200GAT_EDGE_DENSE 2 /* Pe[e] = dense[e]          (caller projected [E,Fe] itself)       */
#define B200GAT_EDGE_AFFINE 3 /* Pe[e] = table[etype[e]] + sum_k attr[e,k] * affine[k]          */
                              /* (psp/models/edge_embedder.py:130-143 folded through lin_edge)   */
#define B200GAT_MAX_ATTR 8

const char* b200gat_last_error(void);
int b200gat_version(void);

/* Number of CUDA kernels this library has launched in the calling process (monotonic):
 * bench.py reports the difference over its timed region as "gpu_launches". */
int64_t b200gat_launch_count(void);

/* Optional device timing per kernel class: when enabled, every launch of that class is
 * bracketed by CUDA events recorded on the launch stream.  b200gat_timing_enable(on) clears
 * the accumulators; b200gat_timing_read synchronises the recorded events and returns the
 * summed kernel time and the number of launches since the last enable.  Used by bench.py for
 * the live roofline figure; off by default (no events, no overhead). */
#define B200GAT_TIMING_CONV_FWD 0
#define B200GAT_TIMING_CONV_BWD_DST 1
#define B200GAT_TIMING_CONV_BWD_SRC 2
#define B200GAT_TIMING_PLAN 3
#define B200GAT_TIMING_HEAD 4
#define B200GAT_TIMING_GEMM 5
#define B200GAT_TIMING_BATCH 6
#define B200GAT_TIMING_GEMM_DX 7
#define B200GAT_TIMING_CLQ_FWD 8
#define B200GAT_TIMING_CLQ_BWD_MMA 9
#define B200GAT_TIMING_CLQ_BWD_ELEM 10
#define B200GAT_TIMING_GEMM_DW 11
#define B200GAT_TIMING_CLASSES 12
void b200gat_timing_enable(int on);
int b200gat_timing_read(int cls, double* total_ms, int64_t* count);
/* CUDA graphs: launches captured while timing is enabled record their event pairs as external event nodes of
 * the graph.  After a replay has completed (synchronise first), b200gat_timing_collect_graph() adds the pairs'
 * elapsed times to the accumulators (returns the number of pairs read); b200gat_timing_clear_graph() destroys
 * the pairs (call it when the graph is destroyed). */
int b200gat_timing_collect_graph(void);
void b200gat_timing_clear_graph(void);

/* ------------------------------------------------------------------ graph plan
 * Topology of one batched graph in both orders, built once per batch and reused by
 * every layer and by forward and backward (the topology is layer-invariant,
 * psp/models/gnn_flat.py:204-208, jssp/models/gnn_dgl.py:366-377).
 * Replaces: DGL's CSR/CSC materialisation behind apply_edges / edge_softmax /
 * update_all (generic/eegatconv.py:269,280,292) and PyG's scatter index handling
 * (generic/gatv3_conv.py:155-160,193).                                              */
struct b200gat_cliques_s;
typedef struct b200gat_graph_s {
  int32_t n_nodes;
  int32_t n_edges;
  const int32_t* in_ptr;  /* [N+1] CSR by destination                                    */
  const int32_t* in_src;  /* [E]   source node of each dst-sorted edge                   */
  const int32_t* in_eid;  /* [E]   original edge id of each dst-sorted edge (stable:     */
                          /*       ascending inside a destination segment)               */
  const int32_t* out_ptr; /* [N+1] CSR by source                                         */
  const int32_t* out_dst; /* [E]   destination node of each src-sorted edge              */
  const int32_t* out_pos; /* [E]   position of that edge in the dst-sorted order         */
  const int32_t* dst_order; /* [N] processing order of destinations (b200gat_node_order on   */
                            /*     in_ptr/in_src), or NULL = identity.  A permutation: it   */
                            /*     changes which warps run side by side, never a result.    */
  const int32_t* src_order; /* [N] same for the source-side backward pass, or NULL          */
  /* Hub nodes: adjacency segments longer than hub_degree are split over the warps of a CTA    */
  /* (the learned pooling node of a 100x20 schedule graph has 2,000 in-edges,                   */
  /* jssp/models/gnn_dgl.py:263-286).  Lists built by b200gat_hub_lists; NULL = no splitting.  */
  int32_t hub_degree;
  const int32_t* hub_dst;   /* [<=N] destinations with in-degree  > hub_degree               */
  const int32_t* n_hub_dst; /* [1]   device-side count                                       */
  const int32_t* hub_src;   /* [<=N] sources with out-degree > hub_degree                    */
  const int32_t* n_hub_src; /* [1]                                                           */
  /* Optional scratch for splitting hub destinations over several CTAs: hub_cap * hub_parts records   */
  /* of (256 + 64) floats; used when the device-side hub count fits hub_cap, ignored (NULL/0) else.   */
  /* One conv call at a time may use a graph's scratch (single-stream use, like the reference).       */
  /* With flag B200GAT_GRAPH_HUB_COUNTERS the records are followed by hub_cap int32 arrival counters, ZERO when a conv */
  /* call starts (the kernels leave them at zero): the slice that arrives last merges its hub inside the kernel       */
  /* instead of a merge launch of its own.                                                                             */
  float* hub_scratch;
  int32_t hub_parts;
  int32_t hub_cap;
  /* Residual graphs only (b200gat_clique_build): position of each in-edge in the destination-   */
  /* sorted order of the FULL graph (index into etype / attr / s_save); NULL = identity.          */
  const int32_t* in_pos;
  /* Dense clique blocks of this graph and the CSR of the edges outside them, or NULL/NULL.      */
  const struct b200gat_cliques_s* cliques;
  const struct b200gat_graph_s* residual;
  /* B200GAT_GRAPH_LOW_DEGREE: every node that is not a hub has at most ~32 in- and out-edges (host knowledge: the  */
  /* plan's degree statistics, or the residual of a clique view).  The attention kernels then take runs of 16        */
  /* consecutive nodes per CTA with the runs' rows and adjacency staged in shared memory (csrc/gat_win.cu) instead   */
  /* of one warp per node in the locality order.  A scheduling hint: it never changes a result beyond rounding.      */
  int32_t flags;
} b200gat_graph;
#define B200GAT_GRAPH_LOW_DEGREE 1
#define B200GAT_GRAPH_HUB_COUNTERS 2
#define B200GAT_HUB_RECORD (256 + 64)   /* floats per hub_scratch record */

/* ------------------------------------------------------------------ clique blocks
 * Machine-conflict cliques (jssp/utils/utils.py:460-478; psp/utils/utils.py:62-92) are complete
 * digraphs: every member reads the same source rows.  A block is a node set S,
 * min_size <= |S| <= B200GAT_MAX_CLIQUE, such that every v in S has the sources S \ {v}, in ascending
 * node id, as ONE contiguous same-typed run of its destination-sorted in-edges.  Blocks are computed
 * as dense tiles (shared-memory staged rows, register-tiled logits, tensor-core aggregation); all other
 * edges stay on the residual CSR.  Blocks are numbered by ascending smallest member.                */
#define B200GAT_MAX_CLIQUE 128
typedef struct b200gat_cliques_s {
  int32_t n_blocks;          /* host copies of counts[0], counts[1], counts[3]                      */
  int32_t max_block;
  int32_t n_members;
  int32_t s_elems;           /* per-head elements of the dense logit buffer = blk_soff[n_blocks]    */
  const int32_t* blk_ptr;    /* [n_blocks+1] offsets into blk_nodes                                  */
  const int32_t* blk_nodes;  /* [n_members]  members of each block, ascending                        */
  const int32_t* blk_type;   /* [n_blocks]   edge type of the block's edges (0 when untyped)         */
  const int32_t* blk_soff;   /* [n_blocks+1] prefix of n_b * roundup4(n_b): block b / head h owns     */
                             /* s_dense[blk_soff[b]*H + h*n_b*roundup4(n_b) ...][dst rank][src rank]   */
  const int32_t* node_blk;   /* [N] block of each node or -1                                         */
  const int32_t* node_run;   /* [N] position (full dst-sorted order) of the first block in-edge      */
  const int32_t* node_rank;  /* [N] rank of the node inside its block                                */
} b200gat_cliques;

int64_t b200gat_clique_scratch_elems(int64_t n_nodes, int64_t n_edges);

/* Detect the blocks of `g` (exact, deterministic) and build the residual CSR.
 * etype_sorted: [E] edge types in destination-sorted order, or NULL (untyped).
 * Outputs (caller-allocated, device): node_blk/node_run/node_rank [N]; blk_ptr/blk_soff [N+1];
 * blk_nodes/blk_type [N]; residual CSR r_in_ptr/r_out_ptr [N+1], r_in_src/r_in_pos/r_out_dst/
 * r_out_pos [E] (upper bounds); counts int32[8] = {n_blocks, max_block, n_residual_edges, n_members,
 * s_elems, 0, 0, 0};
 * scratch: b200gat_clique_scratch_elems int32, 8-byte aligned.                                        */
int b200gat_clique_build(const b200gat_graph* g, const int32_t* etype_sorted, int32_t min_size,
                         int32_t* node_blk, int32_t* node_run, int32_t* node_rank, int32_t* blk_ptr,
                         int32_t* blk_nodes, int32_t* blk_type, int32_t* blk_soff, int32_t* r_in_ptr,
                         int32_t* r_in_src, int32_t* r_in_pos, int32_t* r_out_ptr, int32_t* r_out_dst,
                         int32_t* r_out_pos, int32_t* counts, int32_t* scratch, void* stream);

/* Number of int32 scratch elements b200gat_plan_build needs. */
int64_t b200gat_plan_scratch_elems(int64_t n_nodes, int64_t n_edges);

/* Build the plan from int64 COO (what dgl.batch / Batch.from_data_list hand over,
 * jssp/models/agent_observation.py:223, psp/graph/pyg_graph.py:321-323).
 * edge_src/edge_dst: [E] int64.  Outputs as in b200gat_graph (caller-allocated).
 * stats (int32[4], device): [0] #nodes with zero in-degree, [1] max in-degree,
 * [2] #edges with an endpoint outside [0,N), [3] max out-degree.                     */
int b200gat_plan_build(const int64_t* edge_src, const int64_t* edge_dst, int64_t n_edges,
                       int64_t n_nodes, int32_t* in_ptr, int32_t* in_src, int32_t* in_eid,
                       int32_t* out_ptr, int32_t* out_dst, int32_t* out_pos, int32_t* scratch,
                       int32_t* stats, void* stream);

/* Locality order of the nodes of one CSR side: nodes sorted (stably) by the neighbour in the
 * middle of their adjacency segment, so that nodes with overlapping neighbourhoods (the tasks of
 * one machine clique, jssp/utils/utils.py:460-478) are processed by neighbouring warps.
 * ptr [N+1], nbr [E] (in_ptr/in_src or out_ptr/out_dst) -> order [N]. */
int64_t b200gat_node_order_scratch_elems(int64_t n_nodes);
int b200gat_node_order(const int32_t* ptr, const int32_t* nbr, int64_t n_nodes, int32_t* order,
                       int32_t* scratch, void* stream);

/* Nodes whose adjacency segment is longer than `hub_degree`: list [<=N] + device count [1]. */
int b200gat_hub_list(const int32_t* ptr, int64_t n_nodes, int32_t hub_degree, int32_t* list,
                     int32_t* count, void* stream);

/* out[i] = table[idx[i]] for int32 (permute per-edge attributes into dst-sorted order). */
int b200gat_gather_i32(const int32_t* table, const int32_t* idx, int64_t n, int32_t* out,
                       void* stream);
/* Same, and range[0] = min, range[1] = max of the gathered values (device int32[2]; n == 0 leaves
 * INT_MAX / INT_MIN).  The host mirror checks edge types against the rows of the edge-type table before the
 * first layer uses them: the reference's nn.Embedding raises IndexError on an out-of-range type
 * (jssp/models/gnn_dgl.py:364, psp/models/edge_embedder.py:131), the kernels would read / accumulate out of
 * bounds. */
int b200gat_gather_i32_range(const int32_t* table, const int32_t* idx, int64_t n, int32_t* out,
                             int32_t* range, void* stream);
/* out[i, :] = table[idx[i], :] for fp32 rows of `width` elements (per-edge attributes). */
int b200gat_gather_rows_f32(const float* table, const int32_t* idx, int64_t n, int32_t width,
                            float* out, void* stream);
/* out[i] = (int32) in[i]  (int64 -> int32 narrowing with 16-byte loads). */
int b200gat_narrow_i64(const int64_t* in, int64_t n, int32_t* out, void* stream);

/* ------------------------------------------------------------------ batch builder
 * Device-side replacement of the host batching of JSSP observations:
 *   jssp/models/agent_observation.py:110-229 (per-graph DGL graphs: precedence edges forward (type 1) and
 *   reversed (type 2) :53-72, machine-clique conflict edges :81-88,160-176 = jssp/utils/utils.py:460-478
 *   (type 5, or machine id + 5 with mid_in_edges), self loops (type 0) :210-211, `dgl.batch` :223, the
 *   feature fix-up feat[:,1] = (feat[:,1] == -1 ? -1 : feat[:,4]) :133-135 and no_tct :138-139)
 *   + jssp/models/gnn_dgl.py:263-286 (one pool node per graph appended after all graph nodes: pool self
 *   loops (type 0), [pool -> node (type 3) for "learninv"], node -> pool (type 4)).
 * The reference does this with Python loops over the graphs on the CPU and copies the batched graph; here
 * the padded observation arrays are copied as they are and the COO (same edge ORDER, bit-exact) is written
 * by two kernels.  Per-graph sizes are host knowledge (the observation carries n_nodes / n_edges, and the
 * clique edge counts follow from a histogram of the machine ids), so no device->host read is needed.
 *   features   [B, max_nodes, n_feat] fp32 (column 6 = machine id, -1 = unaffected)
 *   edge_index [B, 2, max_edges] int64 (precedence edges, node ids local to the graph)
 *   meta       [4, B+1] int64, device: row 0 node offsets (exclusive prefix of n_nodes, [B] = total),
 *              row 1 edge offsets of the per-graph segments (prefix of per-graph edge counts),
 *              row 2 n_edges per graph (precedence edges), row 3 clique edges per graph
 * Outputs: src / dst / etype [E_total] int64, feat [N_total (+ B pool rows) (+ B*M machine rows), n_feat] fp32. */
#define B200GAT_POOL_NONE 0
#define B200GAT_POOL_LEARN 1
#define B200GAT_POOL_LEARNINV 2
typedef struct {
  int32_t n_graphs, max_nodes, n_feat, max_edges;
  const float* features;
  const int64_t* edge_index;
  const int64_t* meta;
  int32_t clique_conflicts; /* 1: conflicts == "clique"; 0: no conflict edges ("att")      */
  int32_t mid_in_edges, bidir, self_loops, no_tct;
  int32_t pooling;          /* B200GAT_POOL_*                                               */
  int32_t n_machine_ids;    /* machine ids are in [-1, n_machine_ids)                       */
  int64_t n_nodes_total;    /* meta[0][B]                                                   */
  int64_t n_edges_graphs;   /* meta[1][B]                                                   */
  /* conflicts == "node" (jssp/models/gnn_dgl.py:288-335): max_n_machines machine nodes per graph appended after
   * the pool nodes, wired in this order after the pool edges: machine self loops (type 0), task -> machine
   * (type mid+5) and machine -> task (type mid+5+M) for every affected task in ascending node order,
   * machine -> pool (type 5+2M, "learn" pooling only).  meta row 3 then holds the exclusive prefix of the
   * per-graph affected-task counts ([B] = n_affected) instead of clique edge counts. */
  int32_t node_conflicts;
  int32_t max_n_machines;
  int64_t n_affected;
} b200gat_jssp_obs;
/* Total edge count of the batch (graph segments + pool wiring). */
int64_t b200gat_jssp_batch_edges(const b200gat_jssp_obs* o);
int b200gat_jssp_batch_build(const b200gat_jssp_obs* o, int64_t* src, int64_t* dst, int64_t* etype,
                             float* feat, void* stream);

/* ------------------------------------------------------------------ PSP observation wire format + batch builder
 * (SURVEY rows a5, 8f N1 and N2).
 *
 * Wire format.  The reference ships every observation from an env worker to the learner as a torch.save pickle of
 * a HeteroData dict through a shared byte array (psp/graph/pyg_graph.py:240-283,
 * psp/env/graphgym/async_vector_env.py:59-103) and rebuilds tensors from it per step.  Here an observation is ONE
 * fixed-layout slot of 4-byte little-endian words, written in place by the worker and copied to the device as it is:
 *     int32 header[16]: [0] magic 'PSP1', [1] n_nodes, [2] n_resources, [3] n_prec, [4] n_rp, [5] 16, rest 0
 *     fp32  feat      [max_nodes][16]   the 16 columns AgentGraphObservation.rewire_internal concatenates
 *                                       (psp/models/agent_graph_observation.py:66-86), rows >= n_nodes zero
 *     fp32  resources [max_nodes][max_res]   normalised consumptions (psp/env/gstate.py:260-277), zero padded
 *     fp32  res_cal   [max_res][3]           calendar ids (gstate.py:332-341)
 *     int32 prec      [2][max_prec]          precedence edges, node ids local to the graph (row 0 = source)
 *     int32 rp        [2][max_rp]            resource-priority edges (gstate.py:891-934)
 *     fp32  rp_r      [max_rp][4]            their attributes "r"
 * rprec / rrp (reversed copies) and, in clique mode, the rc conflict edges are functions of these fields
 * (agent_graph_observation.py:88-118) and are generated on the device.  slot_words is rounded up to 16 bytes.
 *
 * Builder.  b200gat_psp_batch_build replaces psp/models/agent.py:298-406 Agent.preprocess (AgentGraphObservation
 * batching :300-309 = Batch.from_data_list per edge type, add_self_loops :339-344, rewire :353-368 =
 * psp/models/rewiring.py:27-363, homogeneous_edges :371-378 = rewiring.py:366-440, to_homogeneous) and the
 * rewire_internal edge additions before it: same node order (task nodes of all graphs, resource nodes, pool
 * nodes), same EDGE ORDER, same per-edge type / rid / att_rc / att_rp, bit-exact.  The reference runs this in
 * Python on the CPU per minibatch (6 DataLoader workers, alg/ppo.py:547-555) and per rollout step.
 *   meta [5][B+1] int64, device: exclusive prefixes over the graphs of n_nodes, n_prec, n_rp, rc edges
 *        (sum_r c_r (c_r - 1), clique mode), resource users (non-zeros of resources, node mode); host knowledge.
 *   base_order: the batch's store order of the base edge types, codes 0 prec, 1 rprec, 2 rp, 3 rrp, 4 rc
 *        (-1 = skip): the order in which the FIRST graph got its edge stores (Batch.from_data_list inherits it).
 * Outputs: edge_index [2][E] int64, type [E] int64, rid [E] int32, attr [E][5] fp32 (att_rc | att_rp),
 *          feat [n_total][16] fp32 (resource-node rows zero: uninitialised in the reference, rewiring.py:132-141),
 *          res_cal_id [B * n_res][3] fp32 (agent_graph_observation.py:264-269). */
#define B200GAT_PSP_MAGIC 0x31505350
#define B200GAT_PSP_HEADER_WORDS 16
#define B200GAT_PSP_N_FEAT 16
#define B200GAT_PSP_MAX_RES 32
#define B200GAT_CONFLICTS_NONE 0
#define B200GAT_CONFLICTS_CLIQUE 1
#define B200GAT_CONFLICTS_NODE 2
typedef struct {
  int32_t max_nodes, max_res, max_prec, max_rp;
  int64_t slot_words;
} b200gat_psp_layout;
typedef struct {
  b200gat_psp_layout layout;
  int32_t n_graphs;
  const int32_t* arena;       /* device: n_graphs consecutive slots */
  const int64_t* meta;        /* device [5][B+1] */
  int32_t conflicts;          /* B200GAT_CONFLICTS_* */
  int32_t pooling;            /* B200GAT_POOL_* */
  int32_t self_loops;
  int32_t base_order[5];
  int32_t n_res;              /* resources per graph after padding = max over the batch (pyg_graph.py:203-215) */
  uint32_t unused_res_mask;   /* bit r: no graph of the batch uses resource r (rewiring.py:218-228) */
  /* filled by b200gat_psp_batch_sizes: first edge of each segment by edge-type id (agent_graph_observation.py:36-53;
   * -1 = absent), totals */
  int64_t seg_start[16];
  int64_t n_edges, n_task_nodes, n_total_nodes;
} b200gat_psp_obs;
int64_t b200gat_psp_slot_words(const b200gat_psp_layout* l);
int b200gat_psp_batch_sizes(b200gat_psp_obs* o, const int64_t* meta_host);
int b200gat_psp_batch_build(const b200gat_psp_obs* o, int64_t* edge_index, int64_t* type, int32_t* rid,
                            float* attr, float* feat, float* res_cal_id, void* stream);

/* ------------------------------------------------------------------ attention layer */
typedef struct {
  int32_t flavour; /* B200GAT_FLAVOUR_*                                                  */
  int32_t heads;   /* H                                                                  */
  int32_t dim;     /* D  (per-head feature size; De == Do is required)                   */
  float slope;     /* leaky_relu slope: 0.01 EGAT (generic/eegatconv.py:276), 0.2 GATv2  */
  float softmax_eps;      /* 0 for DGL edge_softmax, 1e-16 for PyG softmax               */
  const float* attn;      /* additive: [H*D];  dot: w_c [D] (shared by heads)            */
  const float* attn_bias; /* dot: b_c [1]; else NULL                                     */
  const float* out_bias;  /* [H*D] added after aggregation (GATv2 `bias`), or NULL       */
} b200gat_attn;

typedef struct {
  int32_t mode;         /* B200GAT_EDGE_*                                                */
  int32_t n_types;      /* rows of `table`                                               */
  const float* table;   /* [T, H*D]; fold the EGAT pre-activation bias into it           */
  const int32_t* etype; /* [E] edge type in DST-SORTED order (see b200gat_gather_i32)    */
  const float* dense;   /* [E, H*D] in ORIGINAL edge order                               */
  int32_t n_attr;       /* K <= B200GAT_MAX_ATTR (EDGE_AFFINE)                           */
  const float* attr;    /* [E, K] per-edge scalars in DST-SORTED order (b200gat_gather_rows_f32) */
  const float* affine;  /* [K, H*D]                                                      */
} b200gat_edge_feats;

/* Fused per-edge logit + segmented edge-softmax + weighted aggregation.
 * Replaces generic/eegatconv.py:266-292 (apply_edges u_add_v, leaky_relu, attn dot,
 * edge_softmax, update_all u_mul_e/sum) and PyG GATv2Conv/GATv3Conv propagate
 * (psp/graph/graph_conv.py:51-53, generic/gatv3_conv.py:155-198).
 *   Pi, Pj, V : [N, H*D] projected node tensors (V may alias Pi: GATv2/GATv3 x_l)
 *   out       : [N, H*D]
 *   s_save    : [E, H] raw logits in dst-sorted order   (saved for backward)
 *   m_save,l_save : [N, H] per-destination max and sum of exp(s-m)
 *   agg_save  : [N, H*D] the aggregate WITHOUT out_bias, a private copy for the backward
 *               (callers add residuals into `out` in place, psp/models/gnn_flat.py:230), or NULL
 *   s_dense   : [H * g->cliques->s_elems] raw logits of the clique-block edges (those are NOT written
 *               to s_save); required when g->cliques has blocks and the dense path applies
 *               (additive flavour, D == 64, edge mode NONE/TABLE), else may be NULL */
int b200gat_conv_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                     const float* Pi, const float* Pj, const float* V, float* out,
                     float* s_save, float* m_save, float* l_save, float* agg_save, float* s_dense,
                     void* stream);

/* Matching fused backward (replaces autograd through the ops above).
 *   g_out            : [N, H*D]
 *   gPi, gPj, gV     : [N, H*D] written (not accumulated).  If gV == gPi the sum of both
 *                      contributions is written there (GATv2: x_l is logit term and value).
 *   g_attn           : [H*D] (additive) / [D] (dot), ACCUMULATED with atomics: zero it first
 *   g_attn_bias      : [1] (dot), accumulated; else NULL
 *   g_table          : [T, H*D] accumulated (EDGE_TABLE / EDGE_AFFINE) ; g_dense: [E, H*D]
 *                      written (EDGE_DENSE) ; g_affine: [K, H*D] accumulated (EDGE_AFFINE)
 *   gs_scratch       : [2, E, H] scratch (dst-sorted order): d loss / d logit, attention weight;
 *                      when the dense clique path applies (b200gat_conv_uses_cliques) it must hold
 *                      2*E*H + roundup4(H * s_elems) + 2*N*H floats, 16-byte aligned
 *   agg_save         : [N, H*D] from the forward, or NULL (then recomputed with one more pass)  */
int b200gat_conv_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                     const float* Pi, const float* Pj, const float* V, const float* s_save,
                     const float* m_save, const float* l_save, const float* agg_save,
                     const float* s_dense, const float* g_out, float* gPi,
                     float* gPj, float* gV, float* g_attn, float* g_attn_bias, float* g_table,
                     float* g_dense, float* g_affine, float* gs_scratch, void* stream);

/* alpha[e,h] in ORIGINAL edge order from the saved logits (get_attention=True,
 * generic/eegatconv.py:304-308; return_attention_weights, generic/gatv3_conv.py:174-179). */
int b200gat_attention(const b200gat_graph* g, int32_t heads, float softmax_eps,
                      const float* s_save, const float* m_save, const float* l_save,
                      const float* s_dense, float* alpha, void* stream);
/* 1 when b200gat_conv_fwd/bwd would take the dense clique path for this configuration. */
int b200gat_conv_uses_cliques(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef);

/* f_out[e] = leaky_relu(Pi[src]+Pj[dst]+Pe[e]) in ORIGINAL edge order: the edge output of
 * EGATConv (generic/eegatconv.py:276-277,313), produced on request only.                  */
int b200gat_edge_out(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                     const float* Pi, const float* Pj, float* f_out, void* stream);

/* ------------------------------------------------------------------ masked action head
 * Replaces the torch.where / logsumexp / Categorical chain of
 * generic/agent.py:197-231 (fill -1e20), :238-254 (fill -1e10) and :256-277.
 *   logits [B, n] (row stride ld_logits), mask [B, >=n] bytes (row stride ld_mask) or NULL
 *   logp   [B, n] normalised log-probabilities of the masked distribution
 *   entropy[B]    -sum_mask p log p   (all entries when mask == NULL)
 *   argmax [B]    first index of the maximum masked logit (int64, like torch.argmax)     */
int b200gat_mask_head_fwd(const float* logits, int64_t ld_logits, const uint8_t* mask,
                          int64_t ld_mask, int32_t batch, int32_t n, float fill, float* logp,
                          float* entropy, int64_t* argmax, void* stream);
/* g_logits[b,j] = mask ? (1[j==action_b] - p_bj) * g_logprob_b - p_bj (logp_bj + H_b) * g_entropy_b : 0 */
int b200gat_mask_head_bwd(const float* logp, const uint8_t* mask, int64_t ld_mask,
                          const int64_t* action, const float* entropy, const float* g_logprob,
                          const float* g_entropy, int32_t batch, int32_t n, float* g_logits,
                          void* stream);

/* ------------------------------------------------------------------ dense projections (tcgen05, TF32)
 * Replaces the torch.nn.Linear GEMMs around every conv (generic/mlp.py:70-77; fc_ni / fc_nj /
 * fc_node_src of generic/eegatconv.py:260-262,288; lin_l / lin_r of PyG GATv2Conv) at the matmul
 * precision the reference runs them in (TF32, psp/train_psp.py:47):
 *     D[m, n] = sum_k A(m, k) * B(n, k)  (+ bias[n])  -> epilogue
 * a_major / b_major: 0 = the operand is stored [mn rows, k cols] (reduction index contiguous),
 *                    1 = stored [k rows, mn cols].  So with row-major x [M,K], W [N,K], dy [M,N]:
 *     y  = x W^T   : A = x  (0), B = W (0)
 *     dx = dy W    : A = dy (0), B = W (1)            (reduction over N)
 *     dW = dy^T x  : A = dy (1), B = x (1)            (reduction over the M rows, split-K + workspace)
 * a_parts > 1: A is a_parts equally shaped matrices a_part_stride floats apart, concatenated along k
 * (a_major 0) or along m (a_major 1); d_parts > 1: the output columns are split over d_parts matrices
 * [M, N / d_parts] d_part_stride floats apart (the Pi | Pj | V projections in one launch).
 * Epilogues: NONE: D = acc + bias (+ residual); GELU: D = acc + bias (pre-activation, may be NULL),
 * D2 = gelu(D) (exact erf form, torch.nn.GELU default); DGELU: D = (acc + bias) * gelu'(Zin), no residual.
 * Requirements: N a multiple of 4 (64 per part when d_parts > 1), leading dimensions multiples of 4 floats, 16-byte aligned
 * pointers; b200gat_gemm_supported() tells whether a descriptor is covered (callers use the library
 * GEMM otherwise, e.g. for the width-1 head outputs).                                               */
#define B200GAT_EPI_NONE 0
#define B200GAT_EPI_GELU 1
#define B200GAT_EPI_DGELU 2
typedef struct b200gat_gemm {
  int32_t a_major, b_major;
  int64_t M, N, K;
  const float* A;
  int64_t lda;
  int32_t a_parts;
  int64_t a_part_stride;
  const float* B;
  int64_t ldb;
  float* D;
  int64_t ldd;
  int32_t d_parts;
  int64_t d_part_stride;
  const float* bias;
  const float* residual;
  int64_t ldr;
  int32_t epilogue;
  float* D2;
  const float* Zin;
  int64_t ld2;
  float* workspace;         /* split-K partials (b200gat_gemm_workspace_elems floats) or NULL */
  int64_t workspace_elems;
  float* a_rowsum;          /* optional [M], a_major 1 only: sum_k A(m, k) - the bias gradient when A = dy  */
                            /* (dW GEMM); exact fp32 sums of the staged tiles by the idle epilogue warps */
} b200gat_gemm;
/* Arithmetic of b200gat_gemm_tf32 and of everything built on it (MLP / trunk / agent calls), process-wide like
 * torch.backends.cuda.matmul.allow_tf32, which the host mirror copies into it before every call:
 *   B200GAT_GEMM_TF32   operands truncated to TF32 by the tensor core (allow_tf32 = True; psp/train_psp.py:47)
 *   B200GAT_GEMM_FP32X3 3xTF32 split arithmetic: a_hi*b_hi + a_hi*b_lo + a_lo*b_hi with lo = x - trunc(x), three
 *                       tcgen05.mma per k step, ~2^-20 relative per product = fp32 matmul accuracy (allow_tf32 =
 *                       False: the reference's JSSP entry point, and the 1e-4 parity tests)                        */
#define B200GAT_GEMM_TF32 0
#define B200GAT_GEMM_FP32X3 1
void b200gat_gemm_precision(int mode);
int b200gat_gemm_precision_get(void);
int b200gat_gemm_supported(const b200gat_gemm* g);
int64_t b200gat_gemm_workspace_elems(const b200gat_gemm* g);
int b200gat_gemm_tf32(const b200gat_gemm* g, void* stream);
/* out[n] = sum_m X[m, n] (bias gradients), deterministic; scratch: b200gat_colsum_scratch_elems floats */
int64_t b200gat_colsum_scratch_elems(int64_t M, int64_t N);
int b200gat_colsum(const float* X, int64_t M, int64_t N, int64_t ld, float* out, float* scratch,
                   void* stream);
/* b200gat_colsum without a scratch argument and without any allocation, for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * The slab partials live in a file-scope device array of
 * 2 * 148 * B200GAT_COLSUM_NOALLOC_MAX_N floats (1.16 MiB) that every GPU loading the library holds until exit.
 * N > B200GAT_COLSUM_NOALLOC_MAX_N -> B200GAT_ERR_UNSUPPORTED (b200gat_last_error says so); any M.  Same summation
 * order as b200gat_colsum, so the results are bitwise equal.  Every call shares the array (it keeps the last call's
 * partials; nothing needs to start at zero): calls that overlap on the GPU - two streams, two host threads - race.
 * Calls no cudaMalloc / cudaFree / allocating library routine; the array's device address is looked up once per
 * device (cudaGetSymbolAddress) with that device current.                                                        */
#define B200GAT_COLSUM_NOALLOC_MAX_N 1024
int b200gat_colsum_noalloc(const float* X, int64_t M, int64_t N, int64_t ld, float* out, void* stream);

/* ------------------------------------------------------------------ fused GELU MLP (one call per pass)
 * generic/mlp.py:70-77 with activation GELU and no LayerNorm (the default configuration, args.py:335-348):
 *     h_0 = x ; z_i = h_i W_i^T + b_i ; h_{i+1} = gelu(z_i) (i < n-1) ; y = z_{n-1} (+ residual)
 * b200gat_mlp_fwd issues the n GEMMs of the forward pass, b200gat_mlp_bwd the 2n - 1 (or 2n) GEMMs of the
 * backward pass (dW_i = dz_i^T h_i with the bias row sums, dz_{i-1} = (dz_i W_i) * gelu'(z_{i-1}), dx = dz_0 W_0)
 * on b200gat_gemm_tf32, from ONE host call each: the eager PPO step is host-bound, and one binding call per
 * pass instead of one per GEMM is what the host mirror needs.  Every width must be a multiple of 4 floats
 * (the mirror zero-pads the 30-wide JSSP input and the width-1 head outputs).  z / h may be NULL in b200gat_mlp
 * for inference (no pre-activations kept); h is always needed as the next layer's input.                  */
#define B200GAT_MLP_MAX_LAYERS 8
typedef struct b200gat_mlp {
  int32_t n_layers;
  int64_t M;
  int32_t dims[B200GAT_MLP_MAX_LAYERS + 1]; /* dims[0] input width ... dims[n] output width               */
  const float* W[B200GAT_MLP_MAX_LAYERS];   /* [dims[i+1], dims[i]] row-major                             */
  const float* b[B200GAT_MLP_MAX_LAYERS];   /* [dims[i+1]] or NULL                                        */
  const float* x;                           /* [M, dims[0]]                                               */
  const float* residual;                    /* [M, dims[n]] or NULL                                       */
  float* z[B200GAT_MLP_MAX_LAYERS];         /* i < n-1: pre-activations [M, dims[i+1]], NULL = not kept   */
  float* h[B200GAT_MLP_MAX_LAYERS];         /* i < n-1: gelu(z_i)       [M, dims[i+1]]                    */
  float* y;                                 /* [M, dims[n]]                                               */
} b200gat_mlp;
typedef struct b200gat_mlp_grads {
  const float* gy;                          /* [M, dims[n]]                                               */
  float* dx;                                /* [M, dims[0]] or NULL                                       */
  float* dW[B200GAT_MLP_MAX_LAYERS];        /* [dims[i+1], dims[i]] or NULL                               */
  float* db[B200GAT_MLP_MAX_LAYERS];        /* [dims[i+1]] or NULL (needs dW[i])                          */
  float* dz[B200GAT_MLP_MAX_LAYERS];        /* i < n-1: scratch [M, dims[i+1]]                            */
  float* workspace;                         /* b200gat_mlp_bwd_workspace_elems floats                     */
  int64_t workspace_elems;
} b200gat_mlp_grads;
int b200gat_mlp_fwd(const b200gat_mlp* m, void* stream);
int64_t b200gat_mlp_bwd_workspace_elems(const b200gat_mlp* m, const b200gat_mlp_grads* g);
int b200gat_mlp_bwd(const b200gat_mlp* m, const b200gat_mlp_grads* g, void* stream);

/* ------------------------------------------------------------------ trunk readout
 * Replaces the tail of the reference's trunks: torch.cat over the per-layer node tensors
 * (psp/models/gnn_mp.py:419-425, jssp/models/gnn_dgl.py:401), the per-graph Python loop that slices, pools
 * (torch.max(dim=0) / matmul with 1/n / the learned pool node's row) and F.pad's every graph to max_n_nodes
 * followed by torch.stack (gnn_mp.py:429-470,472-500; gnn_dgl.py:402-425).  The sources are read where they lie:
 *   node_features  [B, max_n_nodes, F] = rows row_start[b] .. row_start[b+1]-1 of concat_i src_i, zero rows after
 *   graph_embedding[B, F]              = max / mean over the graph's rows, or row pool_row0 + b of the concat
 *   argmax         [B, F] int32        = first maximal row (local index) per column, kept for the backward
 * (the reference's gradient follows torch.max's index, the first maximum on the CPU).
 * Backward writes every source's gradient [n_rows_total, width_i] completely (rows nothing reads get zeros);
 * g_src[i] == NULL skips a source (the raw input features need no gradient). */
#define B200GAT_READOUT_MAX_SRC 16
#define B200GAT_READOUT_NONE 0
#define B200GAT_READOUT_MAX 1
#define B200GAT_READOUT_AVG 2
#define B200GAT_READOUT_ROWS 3 /* graph_embedding[b] = row pool_row0 + b ("learn" / "learninv" pool nodes) */
typedef struct b200gat_readout {
  int32_t n_src;
  const float* src[B200GAT_READOUT_MAX_SRC]; /* [rows, width[i]], row stride ld[i]                          */
  int32_t width[B200GAT_READOUT_MAX_SRC];
  int64_t ld[B200GAT_READOUT_MAX_SRC];
  int32_t n_graphs, max_n_nodes;
  const int32_t* row_start;                  /* device [B+1]: exclusive prefix of the graphs' task-node counts */
  int64_t n_task_rows;                       /* host copy of row_start[B]                                     */
  int32_t pooling;                           /* B200GAT_READOUT_*                                             */
  int64_t pool_row0;                         /* ROWS: pool node of graph b is row pool_row0 + b               */
  int64_t out_ld;                            /* row stride of node_features / graph_embedding (F <= out_ld <  */
                                             /* F + 128; the pad columns are written as zeros); 0 = F.  The   */
                                             /* heads' first GEMM needs 16-byte aligned rows: F = 478 -> 480  */
} b200gat_readout;
int b200gat_readout_fwd(const b200gat_readout* r, float* node_features, float* graph_embedding, int32_t* argmax,
                        void* stream);
/* ld_g_nf / ld_g_ge: row strides of the incoming gradients (0 = F) */
int b200gat_readout_bwd(const b200gat_readout* r, const float* g_node_features, int64_t ld_g_nf,
                        const float* g_graph_embedding, int64_t ld_g_ge, const int32_t* argmax, float* const* g_src,
                        int64_t n_rows_total, void* stream);

/* ------------------------------------------------------------------ PPO update tail
 * b200gat_gae replaces the Python loop over the rollout steps of alg/ppo.py:276-299:
 *   delta_t = (r_t + gamma * V_{t+1} * nonterminal_{t+1} - V_t) / sigma
 *   A_t     = delta_t + gamma * lambda * nonterminal_{t+1} * A_{t+1} ;  returns = A + V
 * rewards / values / advantages / returns [T, E, R] (steps, envs, reward components), dones [T, E],
 * next_value [E, R], next_done [E]; same operation order as the reference, no fused multiply-adds: bit-exact
 * against the fp32 torch loop. */
int b200gat_gae(const float* rewards, const float* values, const float* dones, const float* next_value,
                const float* next_done, int32_t n_steps, int32_t n_envs, int32_t reward_dim, float gamma,
                float gae_lambda, float sigma, float* advantages, float* returns, void* stream);

/* Clipped-PPO minibatch loss of alg/ppo.py:582-705 and its gradient in ONE launch:
 *   ratio = exp(logprob - old_logprob) ; a = norm_adv ? (adv - mean) / (std + 1e-8) : adv   (:602-605, unbiased std)
 *   pg_loss = mean(max(-a ratio, -a clamp(ratio, 1 - clip, 1 + clip)))   (:608-620; clip_coef < 0: mean(-a ratio))
 *   v_loss  = mse / l1 of (value, returns)  (:632-641);  clip_vloss: mean(max(v_loss_unclipped (the scalar mean),
 *             per-sample loss of old_value + clamp(value - old_value, -clip, clip)))  (:665-690; the reference
 *             hard-codes clip_vloss = False, alg/ppo.py:89)
 *   loss    = pg_loss - ent_coef * mean(entropy) + vf_coef * v_loss      (:701-705, update > 0)
 * stats [8] = {pg_loss, v_loss, entropy_loss, approx_kl (:588-590), clipfrac (:591-597), mean |ratio - 1| (:616),
 *              advantage mean, advantage std}: the values the reference reads back with .item() one by one.
 * g_logprob / g_entropy / g_value [n] = d loss / d input (the loss is a scalar: the autograd backward is a scale).
 * Sharded minibatch: `moments` = device double[3] {count, sum, sum of squares} of the advantages of the
 * GLOBAL minibatch (b200gat_ppo_adv_moments on every rank, all-reduced): normalisation then uses the global mean /
 * std, and loss, stats and gradients are this rank's share sum_local / count_global, so a SUM all-reduce of the
 * gradients (and of loss / stats) gives exactly the single-process values. */
#define B200GAT_CRITIC_L2 0
#define B200GAT_CRITIC_L1 1
#define B200GAT_CRITIC_SL1 2 /* clipped value loss only (:679-684) */
typedef struct b200gat_ppo_loss_args {
  int32_t n;
  const float* logprob;     /* [n] new log-probabilities of the taken actions          */
  const float* old_logprob; /* [n]                                                     */
  const float* advantages;  /* [n] aggregated over reward components                   */
  const float* entropy;     /* [n] or NULL                                             */
  const float* value;       /* [n] new values                                          */
  const float* returns;     /* [n]                                                     */
  const float* old_values;  /* [n], clip_vloss only                                    */
  float clip_coef;          /* < 0: None                                               */
  int32_t norm_adv, clip_vloss, critic_loss;
  float ent_coef, vf_coef;
  const double* moments;    /* device double[3] or NULL (local statistics)             */
} b200gat_ppo_loss_args;
int b200gat_ppo_adv_moments(const float* advantages, int32_t n, double* moments, void* stream);
int b200gat_ppo_loss(const b200gat_ppo_loss_args* p, float* loss, float* stats, float* g_logprob, float* g_entropy,
                     float* g_value, void* stream);

/* ------------------------------------------------------------------ whole-trunk passes (one call per pass)
 * The L message-passing layers of the reference's trunks,
 *     jssp/models/gnn_dgl.py:366-399 :  h = EGATConv(g, x, emb) ; x = mlps[l](h.flatten) (+ x, residual :395-396)
 *     psp/models/gnn_flat.py:170-253  :  h = GraphConv(g, x, edge_embedders[l](g)) ; x = mlps[l](h) (+ x, :228-230)
 * issued from ONE host call per pass.  Per layer: the projections fc_ni | fc_nj | fc_node_src (or lin_l | lin_r with
 * their biases, V = x_l) as one tcgen05 GEMM writing P [n_proj, N, H*D]; the fused attention (b200gat_conv_fwd); the
 * GELU MLP with the residual in its last epilogue (b200gat_mlp_fwd).  Backward: MLP backward, fused attention
 * backward writing gPi | gPj | gV into one buffer, the projection input- and weight-gradient GEMMs reading it as one
 * operand, with the gradient that reaches a layer's output from outside (the readout) and the residual path added in
 * the input-gradient GEMM's epilogue.  The eager PPO step of the host mirror was host-bound (~300 launches from
 * Python, 7.4 ms per step against 4.6 ms of device work); the same launches issued from C cost ~1 ms of host time.
 * Covered configuration: additive flavour, edge mode NONE / TABLE / AFFINE, GELU MLPs without LayerNorm, widths that
 * b200gat_gemm_tf32 covers (b200gat_trunk_supported says so; the host mirror runs layer by layer otherwise).
 * All buffers are caller-allocated; nothing is allocated or synchronised inside. */
#define B200GAT_TRUNK_MAX_LAYERS 16
typedef struct b200gat_trunk_layer {
  int32_t n_proj;            /* 3: Pi | Pj | V (EGAT) ; 2: x_l | x_r, V aliases Pi (GATv2)                        */
  const float* Wp[3];        /* [H*D, in_dim] each                                                               */
  const float* bp[3];        /* [H*D] each, all or none (GATv2 lin_l / lin_r biases)                             */
  b200gat_attn attn;
  b200gat_edge_feats ef;     /* table / affine already projected through fc_fij / lin_edge; not EDGE_DENSE       */
  int32_t n_mlp;
  int32_t dims[B200GAT_MLP_MAX_LAYERS + 1]; /* dims[0] = H*D ... dims[n_mlp] = in_dim (the next layer's input)   */
  const float* W[B200GAT_MLP_MAX_LAYERS];
  const float* b[B200GAT_MLP_MAX_LAYERS];
  /* written by the forward, read by the backward */
  float* P;                  /* [n_proj, N, H*D]                                                                 */
  float* conv_out;           /* [N, H*D]                                                                         */
  float *s_save, *m_save, *l_save, *agg_save, *s_dense;  /* as b200gat_conv_fwd                                  */
  float* z[B200GAT_MLP_MAX_LAYERS];         /* hidden pre-activations (i < n_mlp - 1); NULL = inference          */
  float* h[B200GAT_MLP_MAX_LAYERS];         /* hidden activations                                                */
  float* y;                  /* [N, in_dim] the layer's output                                                   */
} b200gat_trunk_layer;
typedef struct b200gat_trunk {
  int32_t n_layers, in_dim, residual;
  int64_t n_nodes;
  const b200gat_graph* graph;
  const float* x0;           /* [N, in_dim]                                                                      */
  float* wcat;               /* b200gat_trunk_wcat_elems floats: the layers' concatenated projection weights     */
  b200gat_trunk_layer layer[B200GAT_TRUNK_MAX_LAYERS];
} b200gat_trunk;
typedef struct b200gat_trunk_layer_grads {
  const float* g_y;          /* [N, in_dim] gradient reaching this layer's output from outside the trunk, or NULL */
  float* dWp;                /* [n_proj * H*D, in_dim] (the n_proj weight gradients, concatenated)               */
  float* dbp;                /* [n_proj * H*D] or NULL                                                            */
  float *g_attn, *g_table, *g_affine; /* ACCUMULATED (atomics): zero them first; g_table / g_affine by edge mode */
  float* g_out_bias;         /* [H*D] written, or NULL                                                            */
  float* dW[B200GAT_MLP_MAX_LAYERS];
  float* db[B200GAT_MLP_MAX_LAYERS];
} b200gat_trunk_layer_grads;
typedef struct b200gat_trunk_grads {
  float* g_x0;               /* [N, in_dim] or NULL                                                               */
  const float* g_x0_in;      /* [N, in_dim] gradient reaching x0 from outside the trunk (added into g_x0) or NULL */
  float* workspace;          /* b200gat_trunk_bwd_workspace_elems floats, 16-byte aligned                         */
  int64_t workspace_elems;
  b200gat_trunk_layer_grads layer[B200GAT_TRUNK_MAX_LAYERS];
} b200gat_trunk_grads;
int b200gat_trunk_supported(const b200gat_trunk* t);
int64_t b200gat_trunk_wcat_elems(const b200gat_trunk* t);
int64_t b200gat_trunk_bwd_workspace_elems(const b200gat_trunk* t);
int b200gat_trunk_fwd(const b200gat_trunk* t, void* stream);
int b200gat_trunk_bwd(const b200gat_trunk* t, const b200gat_trunk_grads* g, void* stream);

/* ------------------------------------------------------------------ whole PPO-minibatch evaluation (one call per pass)
 * Everything alg/ppo.py:575-711 runs on the device for one minibatch - agent.get_action_and_value
 * (generic/agent.py:190-231: trunk, value_net, action_net, masked categorical) and the clipped-PPO loss (:582-705) -
 * from ONE host call forward and ONE backward:
 *   pad jobs (16-byte aligned copies of ragged inputs / weights) -> [embedder MLP] -> b200gat_trunk_fwd ->
 *   b200gat_readout_fwd -> action_net / value_net MLPs -> masked head -> b200gat_ppo_loss
 * and the matching chain backward.  The host mirror's step was host-bound (5.6 ms of launch-side time per step
 * through ~10 autograd nodes against 4.5 ms of device work); issued from C the same launches cost ~1 ms. */
#define B200GAT_AGENT_MAX_PAD 8
#define B200GAT_SRC_EXTERNAL 0 /* readout source the trunk did not produce (raw features) */
#define B200GAT_SRC_X0 1       /* the trunk's input (embedder output)                      */
#define B200GAT_SRC_LAYER0 2   /* + l: output of trunk layer l                             */
typedef struct b200gat_pad_job {
  const float* src;          /* [rows, cols], row stride ld_src                                                  */
  float* dst;                /* [rows, ld_dst]: the copy, columns cols .. ld_dst-1 zero                          */
  int64_t ld_src, ld_dst;
  int32_t rows, cols;
  int32_t rows_dst;          /* >= rows: rows .. rows_dst-1 of dst are zero too (0 = rows)                        */
  int32_t reserved;
} b200gat_pad_job;
typedef struct b200gat_agent_step {
  int32_t n_pad;
  b200gat_pad_job pad[B200GAT_AGENT_MAX_PAD];
  int32_t has_embedder;      /* 1: trunk.x0 = embedder(x) (jssp/models/gnn_dgl.py:355); 0: the caller built x0   */
  b200gat_mlp embedder;      /* y must be trunk.x0                                                                */
  b200gat_trunk trunk;
  b200gat_readout readout;   /* out_ld must be a multiple of 4 (it is the action_net's input width)              */
  int32_t src_kind[B200GAT_READOUT_MAX_SRC]; /* B200GAT_SRC_* of every readout source                            */
  float* node_features;      /* [B, max_n_nodes, out_ld]                                                          */
  float* graph_embedding;    /* [B, out_ld]                                                                       */
  int32_t* argmax;           /* [B, F] (max pooling) or NULL                                                      */
  b200gat_mlp action_net;    /* x = node_features, M = B * max_n_nodes; output width padded to 4, column 0 used  */
  b200gat_mlp value_net;     /* x = graph_embedding, M = B                                                        */
  const uint8_t* mask;       /* [B, >= max_n_nodes] or NULL                                                       */
  int64_t ld_mask;
  const int64_t* action;     /* [B]                                                                               */
  float fill;                /* -1e20 (generic/agent.py:206)                                                      */
  float* logits;             /* [B, max_n_nodes]                                                                  */
  float* logp;               /* [B, max_n_nodes]                                                                  */
  float* head;               /* [3, B]: logprob | entropy | value                                                 */
  b200gat_ppo_loss_args loss;/* logprob / entropy / value are set by the call (rows of `head`)                    */
  float* loss_out;           /* [9]: loss | the 8 statistics of b200gat_ppo_loss                                  */
  float* g_head;             /* [3, B]: d loss / d (logprob, entropy, value)                                      */
} b200gat_agent_step;
typedef struct b200gat_agent_step_grads {
  const float* g_loss;       /* device [1]: upstream gradient of the loss, or NULL (= 1)                          */
  b200gat_mlp_grads embedder, action_net, value_net; /* dW / db (and embedder.dx) from the caller; the rest is   */
                                                     /* carved out of `workspace` by the call                    */
  b200gat_trunk_grads trunk; /* the layers' parameter gradients from the caller; g_y / g_x0 / workspace: the call */
  float* g_src_ext[B200GAT_READOUT_MAX_SRC]; /* gradient of each EXTERNAL readout source [rows, width], or NULL  */
  float* g_x0;               /* [N, in_dim] total gradient of trunk.x0 (has_embedder == 0), or NULL               */
  float* workspace;
  int64_t workspace_elems;
} b200gat_agent_step_grads;
int64_t b200gat_agent_bwd_workspace_elems(const b200gat_agent_step* a);
int b200gat_agent_fwd(const b200gat_agent_step* a, void* stream);
int b200gat_agent_bwd(const b200gat_agent_step* a, const b200gat_agent_step_grads* g, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GAT_H */
