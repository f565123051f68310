/* lvb_gpu.h — C-ABI of the B200 (sm_100a) Fitch tree-length engine for LVB.
 *
 * This is the whole drop-in boundary: plain pointers and sizes, no C++ or torch
 * types.  LVB's host code (C compiled as C++, see /root/reference/Makefile:74-81)
 * keeps its structs (TREESTACK_TREE_NODES, DataStructure) and calls these entry
 * points where it used to touch state rows on the CPU.  Reference locations each
 * entry point replaces are given as file:line relative to the reference tree.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; the message is in
 *     lvb_gpu_last_error() (per thread).  LVB host code turns non-zero into
 *     crash() (src/Error.c:49-62); see INTEGRATION.md.
 *   - nodes are numbered as in LVB: 0..n-1 leaves (taxon i), n..2n-4 internal; the
 *     root is a leaf index with two children (src/TreeOperations.c:897-900, :588).
 *   - topologies are passed as two strided int64 views so that LVB's own array of
 *     40-byte TREESTACK_TREE_NODES structs (src/LVB.h:121-128) can be handed over
 *     without conversion: left = &tree[0].left, right = &tree[0].right,
 *     stride_bytes = sizeof(TREESTACK_TREE_NODES).  Plain int64 arrays use stride 8.
 *   - one host thread per context; all device work of a context runs on one stream.
 *   - there is no CPU fallback anywhere behind this interface.
 */
#ifndef LVB_GPU_H
#define LVB_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lvb_gpu_ctx lvb_gpu_ctx;

/* A tree (or candidate tree) as getplen() sees it (src/TreeEvaluation.c:49-62):
 * child links, the root, and the set of internal nodes whose state rows must be
 * recomputed — LVB's "dirty" nodes, i.e. those whose sitestate[0] == 0
 * (src/TreeOperations.c:49-59, 88-103, 631-635).  Only left/right of the dirty nodes
 * and of the root are read. */
typedef struct lvb_gpu_topology
{
    const int64_t *left;   /* left[i]  at byte offset i*stride_bytes */
    const int64_t *right;  /* right[i] at byte offset i*stride_bytes */
    int64_t stride_bytes;  /* 40 for TREESTACK_TREE_NODES[], 8 for int64_t[] */
    int64_t root;          /* leaf index */
    const int64_t *dirty;  /* internal node numbers, any order; ignored if n_dirty < 0 */
    int64_t n_dirty;       /* < 0: every internal node (what ss_init / lvb_reroot leave) */
    /* Optional overlay: the tree is left/right with the child links of n_mod nodes replaced.
     * A proposal then is "the current tree + the handful of links mutate_* rewired" and the
     * host need not materialise (treecopy, TreeOperations.c:737-773) a tree per proposal. */
    const int64_t *mod_node;
    const int64_t *mod_left;
    const int64_t *mod_right;
    int64_t n_mod;         /* 0: no overlay */
} lvb_gpu_topology;

/* ---- device / context:  Main.c:82 lvb_initialize() gains device init ---- */
int lvb_gpu_device_count(int *count);
int lvb_gpu_init(int device, lvb_gpu_ctx **ctx);
int lvb_gpu_shutdown(lvb_gpu_ctx *ctx);
const char *lvb_gpu_last_error(void);
/* the CUDA stream all work of this context is enqueued on (a cudaStream_t) */
void *lvb_gpu_stream(lvb_gpu_ctx *ctx);

/* ---- packed state matrix:  DataOperations.c:164-249 DNAToBinary output, and
 * TreeOperations.c:1500-1513 ss_init (leaf rows installed once, on the device) ----
 * enc_rows: n pointers to nwords 64-bit words each (Solve.c:516-519 layout).
 * [word_begin, word_begin+word_count) is the block of words this context owns
 * (site-block sharding, same decomposition as TreeEvaluation.c:95-97); pass 0,nwords
 * for the whole alignment.  max_trees bounds lvb_gpu_tree_create(). */
int lvb_gpu_matrix_upload(lvb_gpu_ctx *ctx, int64_t n, int64_t nwords, const uint64_t *const *enc_rows,
                          int64_t word_begin, int64_t word_count, int64_t max_trees);
/* same, from one contiguous n x nwords row-major array */
int lvb_gpu_matrix_upload_flat(lvb_gpu_ctx *ctx, int64_t n, int64_t nwords, const uint64_t *enc,
                               int64_t word_begin, int64_t word_count, int64_t max_trees);
/* DNAToBinary on the device: rows = n pointers to m characters (after matchange);
 * returns non-zero on a symbol DNAToBinary would crash on (DataOperations.c:235-241). */
int lvb_gpu_matrix_pack(lvb_gpu_ctx *ctx, int64_t n, int64_t m, const char *const *rows,
                        int64_t word_begin, int64_t word_count, int64_t max_trees);

/* the alignment of another context (same or another GPU of this process), copied device to
 * device: how the additional driver contexts of a many-chain run get their leaf rows without a
 * second DNAToBinary + upload.  Same word block as `src`. */
int lvb_gpu_matrix_clone(lvb_gpu_ctx *ctx, lvb_gpu_ctx *src, int64_t max_trees);

/* ---- column scans of the text alignment, before it is packed: constchar
 * (DataOperations.c:272-307; matchange, :309-349, cuts the columns where keep[k] == 0) and
 * MinimumTreeLength (:53-105, literal-character semantics: ambiguity letters count as states of
 * their own, '-' '?' 'N' 'X' do not count, more than MAXSTATES = 5 states add 5, a column with
 * no countable character adds -1).  rows = n pointers to m characters as read_file() left them.
 * keep (m bytes, may be NULL) <- 1 where the column is NOT constant; *n_variable = their number;
 * *min_len_all = MinimumTreeLength of the alignment as given, *min_len_variable = of the
 * alignment after cutcols (what matchange stores in min_len_tree, :404).  Needs no uploaded
 * matrix. ---- */
int lvb_gpu_column_stats(lvb_gpu_ctx *ctx, int64_t n, int64_t m, const char *const *rows, unsigned char *keep,
                         int64_t *n_variable, int64_t *min_len_all, int64_t *min_len_variable);

/* ---- the same statistics for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc.  This call allocates nothing (no cudaMalloc / cudaFree / allocating library routine):
 * d_text (n x m characters, row-major), d_keep (m bytes, may be NULL) and d_out3 (3 x int64: variable columns,
 * MinimumTreeLength over all columns, over the variable columns) are the CALLER's device memory on the
 * context's device; the scratch of the reduction is a fixed array inside the library (393,472 bytes, held by
 * every GPU that loads the library until the process exits).  Limit: m <= LVB_GPU_NOALLOC_MAX_COLUMNS, more is
 * refused with an error.  The kernel is enqueued on the context's stream (lvb_gpu_stream) and the call returns
 * without waiting: read d_out3 / d_keep after lvb_gpu_synchronize or in stream order.  All calls on one GPU
 * share the scratch array: calls that overlap on the GPU (contexts on different streams) race — serialise them.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress); make it before a capture. */
#define LVB_GPU_NOALLOC_MAX_COLUMNS 4194304
int lvb_gpu_column_stats_noalloc(lvb_gpu_ctx *ctx, int64_t n, int64_t m, const unsigned char *d_text,
                                 unsigned char *d_keep, int64_t *d_out3);

/* ---- trees: device-resident state rows + cached per-node change counts of one
 * search chain's current tree (what treealloc gives Anneal, TreeOperations.c:914-955;
 * Solve.c:194-197).  A new tree has every internal node dirty. ---- */
int lvb_gpu_tree_create(lvb_gpu_ctx *ctx, int *tree);
int lvb_gpu_tree_destroy(lvb_gpu_ctx *ctx, int tree);
/* treecopy with state sets (TreeOperations.c:737-773), device to device */
int lvb_gpu_tree_copy(lvb_gpu_ctx *ctx, int dst_tree, int src_tree);

/* ---- getplen that stores (TreeEvaluation.c:49-270): recompute the dirty nodes of
 * `topo` in post-order, write their state rows and change counts into `tree`, add the
 * root step (:238-264), return the tree length.  Used for the first evaluation
 * (n_dirty < 0), for an accepted proposal (replaces treecopy + SwapTrees,
 * TreeOperations.c:737-773, 1062-1080; Solve.c:323, 370) and for re-rooting
 * (TreeOperations.c:576-656: only nodes on the old-root..new-root path change).
 * length may be NULL, in which case the call does not wait for the device. ---- */
int lvb_gpu_tree_commit(lvb_gpu_ctx *ctx, int tree, const lvb_gpu_topology *topo, int64_t *length);

/* the same for several trees at once (independent SA replicates): trees[] must be distinct */
int lvb_gpu_commit_batch(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                         int64_t *lengths);

/* ---- getplen on proposals, score only (Solve.c:300, StartingTemperature.c:128):
 * candidate k is `topos[k]`, a rearrangement of the current tree `trees[k]` whose
 * dirty list names every node whose subtree changed (what mutate_nni/spr/tbr mark,
 * TreeOperations.c:207, 330-334, 520-537).  Nothing is written to the trees.
 * All candidates are evaluated in one launch; lengths[k] is what getplen would
 * return for candidate k. ---- */
int lvb_gpu_score_batch(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                        int64_t *lengths);

/* The same in two halves, so that a host thread that drives several SA chains can prepare the
 * proposals of one set of chains while the device scores another set's: submit enqueues the
 * batch and returns a ticket, wait blocks until that batch is done and delivers its n_cand
 * lengths.  Batches run in submission order on the context's stream; commits enqueued after a
 * submit do not disturb it (a chain's commit always follows the wait of its own proposals).
 * At most 7 other commit/score calls may lie between a submit and its wait. */
int lvb_gpu_score_batch_submit(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                               int *ticket);
int lvb_gpu_batch_wait(lvb_gpu_ctx *ctx, int ticket, int64_t *lengths);

/* lvb_gpu_commit_batch in the same two halves: a caller that evaluates many independent trees (start
 * trees of replicates, the trees of a stack) compiles and enqueues tree k+1 while the device still
 * works on tree k, and collects the lengths with lvb_gpu_batch_wait.  Commits run in submission
 * order on the context's stream. */
int lvb_gpu_commit_batch_submit(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                int *ticket);

/* The same two calls for a caller that compiles the programs itself, on ITS threads: a driver of
 * many SA chains generates a chain's proposals and compiles them (lvb_gpu_compile_schedule, which
 * is thread-safe and keeps one compiler per calling thread) in the same parallel loop, so that the
 * batch call only lays the batch out and launches it.  pre[k] is the program of topos[k] exactly as
 * lvb_gpu_compile_schedule(n, &topos[k], store, ...) returned it (store = 1 for commit, 0 for
 * score); pre[k].instr == NULL: compile topos[k] here, as the plain calls do (a first evaluation
 * with n_dirty < 0 must take this route: it is cut into pieces). */
typedef struct lvb_gpu_program_view
{
    const int32_t *instr; /* n_instr x {a, b, node, ctl} */
    int64_t n_instr;
    int64_t n_slots;
} lvb_gpu_program_view;
int lvb_gpu_commit_batch_pre(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                             const lvb_gpu_program_view *pre, int64_t *lengths);
int lvb_gpu_score_batch_submit_pre(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                   const lvb_gpu_program_view *pre, int *ticket);

/* ---- score AND keep: Solve.c:300-323, 366-370 without the second evaluation.  The reference scores a
 * proposal with getplen on a copy of the tree (treecopy + mutate, :288-300) and, if it is accepted,
 * SwapTrees makes the copy the tree (:323, :370) — the rows getplen wrote are kept.  The plain calls above
 * score read-only and pay a commit (the same rows again, now stored) for an accepted proposal.  Here a
 * candidate with keep[k] != 0 is scored with its rows STORED — into spare rows of its tree, which nothing
 * else reads — and lvb_gpu_accept_kept() afterwards makes it the tree by switching the tree's nodes over to
 * those rows and installing the candidate's per-node counts: no second pass over the rows.  A candidate that
 * is not accepted costs nothing more; the tree is as it was.  Worth it when a proposal is more likely
 * accepted than not (the early, hot phase of the annealing): a kept candidate writes a row per node it
 * reads.  At most ONE kept candidate per tree and batch; candidates with keep[k] == 0 are scored read-only
 * as by lvb_gpu_score_batch.  The lengths are the ones lvb_gpu_score_batch returns.
 *   lvb_gpu_score_keep_submit_pre  asynchronous, programs from the caller (pre[k] compiled with store = 0;
 *                                  pre may be NULL or pre[k].instr NULL: compiled here); lengths through
 *                                  lvb_gpu_batch_wait(ticket)
 *   lvb_gpu_score_keep             the same, synchronous, compiled here
 *   lvb_gpu_accept_kept            cands[i] = index of a kept candidate in the batch named by `ticket`; the
 *                                  batch must still be held: at most 7 device calls after its submit, and
 *                                  after its lengths were fetched.  Returns without waiting for the device;
 *                                  the trees count as changed for every later call. */
/* measuring aid (LVB_GPU_TRACE=1 when the context is made): six device time stamps in ns per work item — (piece,
 * column block), in hand-out order — of the last whole-tree store launch: ticket drawn, instructions decoded,
 * published, first operand taken, last step done, books kept.  out == NULL: only the counts. */
int lvb_gpu_trace_download(lvb_gpu_ctx *ctx, uint64_t *out, int64_t cap_items, int64_t *n_items, int64_t *n_colblocks);

int lvb_gpu_score_keep_submit_pre(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                  const lvb_gpu_program_view *pre, const uint8_t *keep, int *ticket);
int lvb_gpu_score_keep(lvb_gpu_ctx *ctx, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                       const uint8_t *keep, int64_t *lengths, int *ticket);
int lvb_gpu_accept_kept(lvb_gpu_ctx *ctx, int ticket, int64_t n_accept, const int64_t *cands);

/* ---- the exhaustive NNI neighbourhood of one evaluated tree, in one call: what the loop of
 * deterministic_hillclimb (Solve.c:100-111) computes with 2(n-3) x (mutate_deterministic,
 * TreeOperations.c:105-158, + getplen).  lengths[2*(u-n) + j] is the length getplen would return
 * for mutate_deterministic(p = u, left = j) of `topo`, for every internal node u in n..2n-4 and
 * j = 0 (left == LVB_FALSE: u keeps its left child and takes its sister) then 1 — the order of
 * the reference's two loops.
 * `tree` must hold the evaluated current tree (`topo` as last committed; its dirty list is
 * ignored); nothing in it is written.  `scratch_tree` is another device tree of this context whose
 * rows are overwritten with the "up" state sets (the Fitch set of the rest of the tree seen from
 * each internal branch); afterwards it counts as never evaluated.
 * Method: tree lengths do not depend on the root, so each candidate is evaluated on the branch it
 * rearranges from four state rows (two children, sister, up set): 2 launches of row-streaming
 * kernels for all 2(n-3) candidates instead of 2(n-3) dirty paths. ---- */
int lvb_gpu_nni_neighbourhood(lvb_gpu_ctx *ctx, int tree, int scratch_tree, const lvb_gpu_topology *topo,
                              int64_t *lengths);

/* ---- parity / debugging (treedump, TreeOperations.c:1082-1112) ---- */
/* state row of `node` (leaf or internal) of `tree`: word_count words of this context's block */
int lvb_gpu_download_row(lvb_gpu_ctx *ctx, int tree, int64_t node, uint64_t *out);
/* cached change counts of all 2n-3 nodes (leaves give 0), summed over this context's block */
int lvb_gpu_tree_changes(lvb_gpu_ctx *ctx, int tree, int64_t *changes);

/* ---- multi-GPU site-block sharding: one context per rank, each owning a block of
 * words; per-tree partial lengths are summed with one NCCL all-reduce over NVLink
 * inside commit/score_batch once a communicator is attached.  The 128-byte id comes
 * from rank 0 and is distributed by the host program (MPI, torch.distributed, a file). ---- */
int lvb_gpu_comm_unique_id(void *id128);
int lvb_gpu_comm_init(lvb_gpu_ctx *ctx, int nranks, int rank, const void *id128);

/* ---- introspection for tests and bench ---- */
/* record CUDA events around the Fitch kernels (needed by lvb_gpu_last_kernel_ms / lvb_gpu_kernel_time;
 * off by default: two event records per call are not free on the latency-bound SA path) */
int lvb_gpu_set_timing(lvb_gpu_ctx *ctx, int on);
/* number of kernel launches issued by this context so far */
int64_t lvb_gpu_launch_count(lvb_gpu_ctx *ctx);
/* device time (ms) of the Fitch kernel of the most recent commit/score_batch call, measured
 * with CUDA events on the context's stream; waits for the kernel to finish */
int lvb_gpu_last_kernel_ms(lvb_gpu_ctx *ctx, float *ms);
/* sum of the device times (ms) of all Fitch kernel launches of this context since the last reset,
 * from CUDA events recorded on the context's stream around each launch; waits for pending ones */
int lvb_gpu_kernel_time(lvb_gpu_ctx *ctx, double *total_ms, int64_t *launches, int reset);
/* launch again the batch left on the device by the last commit/score call (same programs, no
 * host compile, no host-to-device copy): device-resident timing, and re-scoring after the trees
 * changed.  lengths may be NULL (no wait: consecutive re-runs queue up on the stream).  With a
 * communicator attached, every re-run ends with the all-reduce of the partial lengths. */
int lvb_gpu_rerun_last_batch(lvb_gpu_ctx *ctx, int64_t *lengths);
/* wait for all work enqueued on the context's stream */
int lvb_gpu_synchronize(lvb_gpu_ctx *ctx);
/* Host-only: compile `topo` into the instruction stream the kernels execute, without a device.
 * instr receives up to cap 16-byte instructions {a, b, node, ctl}; returns the count in *n_instr
 * and the number of on-chip stack slots in *n_slots (at most n_dirty + 2 instructions).
 * Thread-safe; used by callers that compile on their own threads (the *_pre calls) and by the CPU
 * tests of the schedule compiler. */
int lvb_gpu_compile_schedule(int64_t n, const lvb_gpu_topology *topo, int store, int32_t *instr, int64_t cap,
                             int64_t *n_instr, int64_t *n_slots);

/* Host-only: the same for a store program cut into independent subtrees of about piece_size
 * nodes (how large commits are spread over the GPU).  pieces receives {instr_begin, n_instr,
 * level} triples; pieces of one level may run concurrently, levels run in order. */
int lvb_gpu_compile_pieces(int64_t n, const lvb_gpu_topology *topo, int piece_size, int32_t *instr, int64_t cap,
                           int64_t *n_instr, int64_t *n_slots, int32_t *pieces, int64_t pieces_cap,
                           int64_t *n_pieces);

/* Host-only: the up-pass program lvb_gpu_nni_neighbourhood runs — one store instruction per
 * internal node v, in pre-order, whose result is the Fitch set of the tree WITHOUT the subtree
 * below v.  Operand rows >= n are rows of the evaluated tree; results belong under the same node
 * numbers in the scratch tree.
 * piece_size > 0 cuts it into a top piece (level 0) and independent subtrees of at most piece_size
 * nodes (level 1, run after level 0), reported as {instr_begin, n_instr, level} triples; an
 * operand (2n-3) + (v - n) then names the row of v in the scratch tree. */
int lvb_gpu_compile_up_pass(int64_t n, const lvb_gpu_topology *topo, int piece_size, int32_t *instr, int64_t cap,
                            int64_t *n_instr, int64_t *n_slots, int32_t *pieces, int64_t pieces_cap, int64_t *n_pieces);

#ifdef __cplusplus
}
#endif
#endif /* LVB_GPU_H */
