/* fitch_kernels.h — launch wrappers of the sm_100a kernels (fitch_kernels.cu). */
#ifndef LVB_B200_FITCH_KERNELS_H
#define LVB_B200_FITCH_KERNELS_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "fitch_isa.h"

namespace lvbgpu
{

/* Everything a batch launch needs; lives in kernel parameter space. */
/* Sum of K int64 over the ranks of a site-block run through PEER MEMORY (NVLink / NVSwitch loads and
 * stores between the GPUs of one box) instead of a collective library call: every rank owns a slab
 * that all ranks have mapped (CUDA IPC); rank r writes its K partial values into its own row of slot
 * (seq mod slots) of EVERY rank's slab, then sums the nranks rows that arrived in its own slab.
 * An entry is two 8-byte words {low half | seq << 32}, {high half | seq << 32}: an 8-byte store is
 * atomic, so the sequence number doubles as the arrival flag (no separate flag, no fence between
 * data and flag).  Every rank adds the rows in the same order.  A rank can run at most one call
 * ahead of the slowest one, so two slots would do; there are four. */
#define LVB_P2P_MAX_RANKS 16
#define LVB_P2P_SLOTS 4
struct P2PParams
{
    unsigned long long *peer[LVB_P2P_MAX_RANKS]; /* peer[r]: rank r's slab as mapped on this rank (peer[rank] = own) */
    int32_t nranks, rank;
    int32_t slot;               /* seq % LVB_P2P_SLOTS */
    uint32_t seq;               /* call number, never 0 */
    int64_t kmax;               /* entries per row */
    int64_t K;
    const long long *in;        /* [K] this rank's partial values */
    long long *out;             /* [K] the sums (may alias in) */
    long long *out2;            /* optional second destination (mapped host memory), or NULL */
    unsigned *err;              /* mapped host word: set to 0x80000000 | rank that did not arrive */
    unsigned spin_limit;
};

#define LVB_DEP_LEVELS 8 /* piece levels of one launch (schedule.cpp cuts at most 7) */

struct ExecParams
{
    uint4 *pool;                 /* row pool: rows [0,n) leaves, then I rows per tree */
    int64_t stride_u4;           /* row stride in 16-byte columns */
    int32_t ncols;               /* columns to process (= stride_u4) */
    int32_t n;                   /* leaves */
    int32_t n_internal;          /* I = n - 3 */
    int32_t tree_stride;         /* pool rows per device tree: 2 I (two banks of I rows; a node's current row is in one
                                  * of them, the other takes the rows of a KEPT candidate, see peek) */
    int32_t peek;                /* store-mode launch whose finalize step only LOOKS: out[k] = the length the candidate
                                  * would have, kept[] <- its per-node counts, newsum[k] <- its sum of counts; chg and
                                  * sum_chg stay as they are (lvb_gpu_score_keep_*: candidates scored with their rows
                                  * stored in the spare bank; fitch_apply_kept_kernel makes one of them the tree) */
    int32_t *kept;               /* [n_instr of batch] */
    long long *newsum;           /* [n_cand] */
    int32_t n_prog;
    int32_t n_colblocks;
    int32_t n_slots;             /* stack slots of the deepest program in the batch */
    int32_t score_v1;            /* score mode: use the one-instruction-look-ahead kernel (A/B switch) */
    const FitchInstr *instr;
    const FitchProgram *prog;
    unsigned long long *totals;  /* [n_prog] score mode: partial sums of the column blocks; left zero by the kernel */
    unsigned *done;              /* [n_prog] score mode: column blocks finished; left zero by the kernel */
    long long *out;              /* [n_prog] score mode: the length, written by the last column block
                                  * (pinned host memory mapped into the device, or device memory) */
    int32_t *cnts;               /* [n_instr of batch] per-instruction counts (store mode); zeroed by finalize */
    unsigned *ticket;            /* store mode, dependent pieces: ticket[0] CTA start-order counter (left zero by the
                                  * kernel) */
    unsigned *dep;               /* store mode: [LVB_DEP_LEVELS][n_colblocks] pieces of a level finished per column
                                  * block, counted over all launches (never reset while the geometry stays); NULL:
                                  * the pieces of the launch are independent */
    unsigned dep_target[LVB_DEP_LEVELS]; /* value dep[L][cb] has when every piece of level L of THIS launch has
                                  * stored its rows for column block cb: a piece of level L + 1 waits for it */
    int32_t fused_finalize;      /* store mode: the finalize step (chg_rw / sum_chg_rw / out) runs inside the Fitch
                                  * kernel instead of a second launch.  Single-level batch: by the last column block
                                  * of a program.  Dependent pieces: the last column block of a PIECE folds that
                                  * piece's counts, the last piece of a candidate publishes its length */
    const FitchProgram *cand;    /* store mode, dependent pieces, fused finalize: [n_cand] one record per candidate;
                                  * cand[k].n_wait = number of pieces of candidate k; a piece names its candidate
                                  * in FitchProgram::pad1 */
    int32_t n_cand;
    int32_t pipe_pf;             /* store mode, dependent pieces: 0 = register-pipelined kernel (fitch_exec_kernel),
                                  * 2/3/5/7 = cp.async ring with that many instructions of look-ahead */
    int32_t level_off[LVB_DEP_LEVELS + 1]; /* streaming kernel, dependent pieces: pieces of level L are prog[level_off[L]
                                  * .. level_off[L + 1]) (the pieces are sorted by level) */
    int32_t n_levels;
    int32_t lag;                 /* streaming kernel: work order.  < 0 (default): level by level, piece by piece, all column
                                  * blocks of a piece together.  >= 0 (LVB_GPU_STREAM_LAG, measured: no gain): item (piece of level L, column block cb) has the key
                                  * cb + L * lag; items are handed out by ascending key, so the upper levels of a column
                                  * block follow `lag` column blocks behind its level below instead of after ALL column
                                  * blocks of that level: by the time they are drawn their operands are normally there,
                                  * and the end of the launch has only the last column blocks' small upper pieces left */
    unsigned long long *trace;   /* streaming kernel, LVB_GPU_TRACE=1 (a measuring aid, NULL otherwise): per work item
                                  * (ticket) six %globaltimer stamps: ticket drawn, instructions decoded, published (after
                                  * the wait for the level below), first operand in registers, last step done, books kept */
    int32_t fuse_p2p;            /* streaming kernel, fused finalize, site-block mode: the thread that completes a
                                  * candidate's partial length exchanges it with the other ranks through peer memory
                                  * right there (x: in/out/K unused; the candidate number is the entry) — the launch
                                  * IS the all-reduce, no second kernel.  out[k] and x.out2[k] receive the sum */
    P2PParams x;
    int32_t stream_nt;           /* store mode: > 0 = the streaming kernel (TMA bulk loads, persistent CTAs) with this many
                                  * consumer threads per CTA; `block` of launch_fitch_exec is then ignored and
                                  * n_colblocks must have been computed with stream_nt columns per block; ticket must
                                  * be set (dep may be NULL: independent programs) */
    unsigned *err;               /* mapped host word: 1 + piece whose dependency wait was abandoned (0 = none) */
    unsigned spin_limit;         /* polls of a dependency counter after which the wait is abandoned */
    int32_t *chg_rw;             /* = chg, writable (fused finalize) */
    long long *sum_chg_rw;       /* = sum_chg, writable */
    const int32_t *chg;          /* [trees][I] cached per-node change counts */
    const long long *sum_chg;    /* [trees] sum of chg over internal nodes */
};

/* block: threads per CTA (multiple of 32).  per_node: store-mode batch (per-instruction
 * counts are produced, base length is not added). */
cudaError_t launch_fitch_exec(const ExecParams &p, bool per_node, int block, cudaStream_t stream);

/* The same launch(es) with the batch INSIDE the kernel parameters (no staged copy on the device):
 * for the small batches of a single SA chain, where a host-to-device copy per call is a large
 * part of a round.  prog/cand/instr are HOST pointers.  per_node: store batch, followed by the
 * finalize launch (lengths as for launch_commit_finalize); else score batch.
 * inline_batch_fits tells whether a batch is small enough. */
bool inline_batch_fits(int n_prog, int n_cand, size_t n_instr);
cudaError_t launch_fitch_inline(const ExecParams &p, bool per_node, int block, cudaStream_t stream,
                                const FitchProgram *prog, int n_prog, const FitchProgram *cand, int n_cand,
                                const FitchInstr *instr, size_t n_instr, int32_t *chg, long long *sum_chg,
                                long long *lengths, int max_instr);

/* A kept candidate becomes its tree: chg[tree][node] <- kept count of every node instruction, sum_chg[tree] <-
 * newsum[k]; which[] lists the candidates (indices into cand), one CTA each */
cudaError_t launch_apply_kept(const FitchInstr *instr, const FitchProgram *cand, const int32_t *which, int n_which,
                              const int32_t *kept, const long long *newsum, int32_t *chg, long long *sum_chg, int n,
                              int n_internal, cudaStream_t stream);

/* the same for a batch that went out inside kernel parameters (inline_batch_fits): cand / instr are HOST pointers, the
 * candidate list travels in the parameters too — one launch, nothing staged */
cudaError_t launch_apply_kept_inline(const FitchProgram *cand, int n_cand, const FitchInstr *instr, size_t n_instr,
                                     const int64_t *which, int n_which, const int32_t *kept, const long long *newsum,
                                     int32_t *chg, long long *sum_chg, int n, int n_internal, cudaStream_t stream);

/* After a store-mode batch: chg[tree][node] <- new count, sum_chg[tree] += delta,
 * lengths[p] <- sum_chg[tree] + root step counts.  Trees of the batch must be distinct. */
cudaError_t launch_commit_finalize(const FitchInstr *instr, const FitchProgram *prog, int n_prog, int32_t *cnts,
                                   int32_t *chg, long long *sum_chg, int n, int n_internal, long long *lengths,
                                   int max_instr, cudaStream_t stream, int32_t *kept = nullptr,
                                   long long *newsum = nullptr);

/* Exhaustive NNI neighbourhood (Solve.c:64-142, TreeOperations.c:105-158): branch e joins the
 * pool rows rows[e] = {A, B, C, D}; delta[2e] += length(((A,C),(B,D))) - length(((A,B),(C,D))),
 * delta[2e+1] += length(((B,C),(A,D))) - the same, over this context's columns.  The current
 * arrangement is taken from what the two trees already hold for u (rows2, chg_low, chg_up).
 * delta must be zero before the launch. */
struct NniParams
{
    const uint4 *pool;
    int64_t stride_u4;
    int32_t ncols;
    int32_t n_edges;
    int32_t n_colblocks;
    const uint4 *rows;      /* [n_edges] A, B, C, D */
    const uint2 *rows2;     /* [n_edges] low(u), up(u): the rows of u in the evaluated and in the scratch tree */
    const int32_t *chg_low; /* [n_edges] cached counts of u in the evaluated tree = changes(A,B) over all columns */
    const int32_t *chg_up;  /* [n_edges] the same in the scratch tree (left by the up-pass) = changes(C,D) */
    long long *delta;       /* [2 * n_edges] */
};
cudaError_t launch_nni_neighbourhood(const NniParams &p, int block, cudaStream_t stream);

/* DNAToBinary on the device (reference: src/DataOperations.c:164-249).  text: n x m bytes
 * row-major; writes words [word_begin, word_begin + word_count) of each leaf row into the pool;
 * *bad is set to 1 + the offending byte if a symbol is not allowed. */
cudaError_t launch_pack_dna(const unsigned char *text, int64_t n, int64_t m, int64_t word_begin, int64_t word_count,
                            uint4 *pool, int64_t stride_u4, int *bad, cudaStream_t stream);

/* constchar + MinimumTreeLength on the device (src/DataOperations.c:272-307, 53-105).  text: n x m
 * bytes row-major.  keep[k] (may be NULL) <- 1 where column k is not constant; sums[0] += kept
 * columns, sums[1] += MinimumTreeLength over all columns, sums[2] += over the kept columns. */
cudaError_t launch_column_stats(const unsigned char *text, int64_t n, int64_t m, unsigned char *keep, long long *sums,
                                cudaStream_t stream);

/* The same statistics with NO allocation: text, keep (may be NULL) and out[3] (= what launch_column_stats adds into
 * sums, here written) are the caller's device memory; the scratch of the reduction is a file-scope device array of
 * fitch_kernels.cu, shared by all calls on a device.  m <= LVB_NOALLOC_MAX_COLUMNS, else cudaErrorInvalidValue. */
#define LVB_NOALLOC_MAX_COLUMNS (1 << 22)
cudaError_t launch_column_stats_noalloc(const unsigned char *text, int64_t n, int64_t m, unsigned char *keep, long long *out,
                                        cudaStream_t stream);

cudaError_t launch_p2p_allreduce(const P2PParams &p, cudaStream_t stream);

} // namespace lvbgpu
#endif
