/* tests/mock_gpu.c — TEST INFRASTRUCTURE ONLY.
 *
 * A stand-in for liblvbgpu.so on machines without a GPU: the entry points of include/lvb_gpu.h
 * that the host layer (lvb_b200/host/) calls, answered by the CPU oracle (oracle/fitch_oracle.c).
 * tests/test_host_search.py loads it with RTLD_GLOBAL BEFORE liblvbhost.so, so that the host
 * layer's unresolved lvb_gpu_* symbols bind here; the product libraries contain no seam, switch
 * or fallback for this, and nothing under lvb_b200/ refers to this file.  It lets the `-m "not
 * gpu"` suite check the host logic (speculative batching, acceptance replay, cooling, tree
 * stack) against trajectories recorded from the reference; the `-m gpu` suite runs the same
 * checks on the real library.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/lvb_gpu.h"
#include "../oracle/fitch_oracle.h"

struct lvb_gpu_ctx
{
    long n, nwords, B, I, max_trees;
    uint64_t *leaves;   /* n x nwords */
    uint64_t **rows;    /* per tree: B x nwords (leaf rows included) */
    long **changes;     /* per tree: B */
    char *used, *ready;
    long *left, *right; /* scratch topology */
    uint64_t *scratch;
    long *scratch_chg;
    int64_t launches;
    int64_t *ticket_len[8]; /* submitted batches waiting for lvb_gpu_batch_wait */
    int64_t ticket_n[8];
    int ticket_next;
    /* lvb_gpu_score_keep_*: the evaluated state of every kept candidate of a batch, until the ticket is reused */
    int64_t keep_n[8];
    int *keep_tree[8];
    uint64_t **keep_rows[8];
    long **keep_chg[8];
};
#define MOCK_TICKETS 8

static char g_err[512];
static int fail(const char *msg)
{
    snprintf(g_err, sizeof g_err, "mock: %s", msg);
    return 1;
}

static void drop_kept(lvb_gpu_ctx *c, int t);
const char *lvb_gpu_last_error(void) { return g_err; }
int lvb_gpu_device_count(int *count)
{
    *count = 1;
    return 0;
}
int lvb_gpu_init(int device, lvb_gpu_ctx **ctx)
{
    (void)device;
    *ctx = (lvb_gpu_ctx *)calloc(1, sizeof(lvb_gpu_ctx));
    return 0;
}
int lvb_gpu_shutdown(lvb_gpu_ctx *c)
{
    if (!c)
        return 0;
    for (long t = 0; t < c->max_trees; t++)
    {
        free(c->rows[t]);
        free(c->changes[t]);
    }
    free(c->rows);
    free(c->changes);
    free(c->leaves);
    free(c->used);
    free(c->ready);
    free(c->left);
    free(c->right);
    free(c->scratch);
    free(c->scratch_chg);
    for (int t = 0; t < MOCK_TICKETS; t++)
        drop_kept(c, t);
    free(c);
    return 0;
}

int lvb_gpu_matrix_upload(lvb_gpu_ctx *c, int64_t n, int64_t nwords, const uint64_t *const *enc_rows, int64_t word_begin,
                          int64_t word_count, int64_t max_trees)
{
    if (word_begin < 0 || word_count < 1 || word_begin + word_count > nwords)
        return fail("bad word block");
    /* a site block is simply a narrower alignment: words [word_begin, word_begin + word_count) */
    c->n = n;
    c->nwords = word_count;
    c->B = 2 * n - 3;
    c->I = n - 3;
    c->max_trees = max_trees;
    c->leaves = (uint64_t *)malloc((size_t)n * word_count * 8);
    for (long i = 0; i < n; i++)
        memcpy(c->leaves + i * word_count, enc_rows[i] + word_begin, (size_t)word_count * 8);
    c->rows = (uint64_t **)calloc((size_t)max_trees, sizeof(uint64_t *));
    c->changes = (long **)calloc((size_t)max_trees, sizeof(long *));
    c->used = (char *)calloc((size_t)max_trees, 1);
    c->ready = (char *)calloc((size_t)max_trees, 1);
    c->left = (long *)malloc((size_t)c->B * sizeof(long));
    c->right = (long *)malloc((size_t)c->B * sizeof(long));
    c->scratch = (uint64_t *)malloc((size_t)c->B * word_count * 8);
    c->scratch_chg = (long *)malloc((size_t)c->B * sizeof(long));
    return 0;
}

int lvb_gpu_matrix_upload_flat(lvb_gpu_ctx *c, int64_t n, int64_t nwords, const uint64_t *enc, int64_t word_begin,
                               int64_t word_count, int64_t max_trees)
{
    const uint64_t **rows = (const uint64_t **)malloc((size_t)n * sizeof(*rows));
    for (long i = 0; i < n; i++)
        rows[i] = enc + i * nwords;
    const int rc = lvb_gpu_matrix_upload(c, n, nwords, rows, word_begin, word_count, max_trees);
    free(rows);
    return rc;
}

int lvb_gpu_matrix_clone(lvb_gpu_ctx *c, lvb_gpu_ctx *src, int64_t max_trees)
{
    /* src->nwords is the block width; upload it as a whole alignment of that width */
    return lvb_gpu_matrix_upload_flat(c, src->n, src->nwords, src->leaves, 0, src->nwords, max_trees);
}

/* cached change counts of all 2n-3 nodes of this context's block */
int lvb_gpu_tree_changes(lvb_gpu_ctx *c, int tree, int64_t *changes)
{
    for (long v = 0; v < c->B; v++)
        changes[v] = v < c->n ? 0 : c->changes[tree][v];
    return 0;
}

int lvb_gpu_tree_create(lvb_gpu_ctx *c, int *tree)
{
    for (long t = 0; t < c->max_trees; t++)
        if (!c->used[t])
        {
            c->used[t] = 1;
            c->ready[t] = 0;
            if (!c->rows[t])
            {
                c->rows[t] = (uint64_t *)calloc((size_t)c->B * c->nwords, 8);
                c->changes[t] = (long *)calloc((size_t)c->B, sizeof(long));
            }
            memcpy(c->rows[t], c->leaves, (size_t)c->n * c->nwords * 8);
            *tree = (int)t;
            return 0;
        }
    return fail("all device trees are in use");
}

int lvb_gpu_tree_destroy(lvb_gpu_ctx *c, int tree)
{
    c->used[tree] = 0;
    return 0;
}

int lvb_gpu_tree_copy(lvb_gpu_ctx *c, int dst, int src)
{
    memcpy(c->rows[dst], c->rows[src], (size_t)c->B * c->nwords * 8);
    memcpy(c->changes[dst], c->changes[src], (size_t)c->B * sizeof(long));
    c->ready[dst] = c->ready[src];
    return 0;
}

static long strided(const int64_t *base, int64_t stride, long i)
{
    return (long)*(const int64_t *)((const char *)base + i * stride);
}

/* evaluate topo against `rows`/`chg` (modified in place) */
static long evaluate(lvb_gpu_ctx *c, const lvb_gpu_topology *tp, uint64_t *rows, long *chg)
{
    for (long v = 0; v < c->B; v++)
    {
        c->left[v] = strided(tp->left, tp->stride_bytes, v);
        c->right[v] = strided(tp->right, tp->stride_bytes, v);
    }
    for (long k = 0; k < tp->n_mod; k++)
    {
        c->left[tp->mod_node[k]] = (long)tp->mod_left[k];
        c->right[tp->mod_node[k]] = (long)tp->mod_right[k];
    }
    if (tp->n_dirty < 0)
        for (long v = c->n; v < c->B; v++)
            rows[v * c->nwords] = 0;
    else
        for (long k = 0; k < tp->n_dirty; k++)
            rows[tp->dirty[k] * c->nwords] = 0;
    c->launches++;
    return oracle_getplen(c->n, c->nwords, c->left, c->right, (long)tp->root, rows, chg);
}

int lvb_gpu_commit_batch(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                         int64_t *lengths)
{
    for (int64_t k = 0; k < n_cand; k++)
    {
        const int t = trees[k];
        if (!c->ready[t] && topos[k].n_dirty >= 0)
            return fail("first commit must recompute every internal node");
        const long len = evaluate(c, &topos[k], c->rows[t], c->changes[t]);
        c->ready[t] = 1;
        if (lengths)
            lengths[k] = len;
    }
    return 0;
}

int lvb_gpu_tree_commit(lvb_gpu_ctx *c, int tree, const lvb_gpu_topology *topo, int64_t *length)
{
    return lvb_gpu_commit_batch(c, 1, &tree, topo, length);
}

int lvb_gpu_score_batch(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                        int64_t *lengths)
{
    for (int64_t k = 0; k < n_cand; k++)
    {
        const int t = trees[k];
        if (!c->ready[t])
            return fail("tree has never been evaluated");
        memcpy(c->scratch, c->rows[t], (size_t)c->B * c->nwords * 8);
        memcpy(c->scratch_chg, c->changes[t], (size_t)c->B * sizeof(long));
        lengths[k] = evaluate(c, &topos[k], c->scratch, c->scratch_chg);
    }
    return 0;
}

int lvb_gpu_column_stats(lvb_gpu_ctx *c, int64_t n, int64_t m, const char *const *rows, unsigned char *keep,
                         int64_t *n_variable, int64_t *min_len_all, int64_t *min_len_variable)
{
    (void)c;
    unsigned char *k = keep ? keep : (unsigned char *)malloc((size_t)m);
    const long kept = oracle_variable_columns((long)n, (long)m, rows, k);
    char **cut = (char **)malloc((size_t)n * sizeof(char *));
    for (long i = 0; i < n; i++)
    {
        cut[i] = (char *)malloc((size_t)kept + 1);
        long w = 0;
        for (long col = 0; col < m; col++)
            if (k[col])
                cut[i][w++] = rows[i][col];
    }
    if (n_variable)
        *n_variable = kept;
    if (min_len_all)
        *min_len_all = oracle_min_tree_length((long)n, (long)m, rows);
    if (min_len_variable)
        *min_len_variable = kept ? oracle_min_tree_length((long)n, kept, (const char *const *)cut) : 0;
    for (long i = 0; i < n; i++)
        free(cut[i]);
    free(cut);
    if (!keep)
        free(k);
    return 0;
}

/* asynchronous halves: the mock computes at submit time and hands the lengths over at wait time */
int lvb_gpu_score_batch_submit(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos, int *ticket)
{
    const int t = c->ticket_next;
    c->ticket_next = (c->ticket_next + 1) % MOCK_TICKETS;
    if (c->ticket_n[t])
        return fail("too many submitted batches");
    c->ticket_len[t] = (int64_t *)realloc(c->ticket_len[t], (size_t)n_cand * sizeof(int64_t));
    if (lvb_gpu_score_batch(c, n_cand, trees, topos, c->ticket_len[t]))
        return 1;
    c->ticket_n[t] = n_cand;
    *ticket = t;
    return 0;
}
int lvb_gpu_batch_wait(lvb_gpu_ctx *c, int ticket, int64_t *lengths)
{
    if (ticket < 0 || ticket >= MOCK_TICKETS || !c->ticket_n[ticket])
        return fail("no such ticket");
    memcpy(lengths, c->ticket_len[ticket], (size_t)c->ticket_n[ticket] * sizeof(int64_t));
    c->ticket_n[ticket] = 0;
    return 0;
}

/* score and keep: a kept candidate's rows and counts are put aside; accepting it installs them */
static void drop_kept(lvb_gpu_ctx *c, int t)
{
    for (int64_t k = 0; k < c->keep_n[t]; k++)
    {
        free(c->keep_rows[t][k]);
        free(c->keep_chg[t][k]);
    }
    free(c->keep_rows[t]);
    free(c->keep_chg[t]);
    free(c->keep_tree[t]);
    c->keep_rows[t] = NULL;
    c->keep_chg[t] = NULL;
    c->keep_tree[t] = NULL;
    c->keep_n[t] = 0;
}
int lvb_gpu_score_keep_submit_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                  const lvb_gpu_program_view *pre, const uint8_t *keep, int *ticket)
{
    (void)pre;
    if (lvb_gpu_score_batch_submit(c, n_cand, trees, topos, ticket))
        return 1;
    const int t = *ticket;
    drop_kept(c, t);
    c->keep_n[t] = n_cand;
    c->keep_tree[t] = (int *)calloc((size_t)n_cand, sizeof(int));
    c->keep_rows[t] = (uint64_t **)calloc((size_t)n_cand, sizeof(uint64_t *));
    c->keep_chg[t] = (long **)calloc((size_t)n_cand, sizeof(long *));
    for (int64_t k = 0; k < n_cand; k++)
    {
        c->keep_tree[t][k] = trees[k];
        if (!keep[k])
            continue;
        c->keep_rows[t][k] = (uint64_t *)malloc((size_t)c->B * c->nwords * 8);
        c->keep_chg[t][k] = (long *)malloc((size_t)c->B * sizeof(long));
        memcpy(c->keep_rows[t][k], c->rows[trees[k]], (size_t)c->B * c->nwords * 8);
        memcpy(c->keep_chg[t][k], c->changes[trees[k]], (size_t)c->B * sizeof(long));
        if (evaluate(c, &topos[k], c->keep_rows[t][k], c->keep_chg[t][k]) != c->ticket_len[t][k])
            return fail("kept candidate: lengths differ");
    }
    return 0;
}
int lvb_gpu_score_keep(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                       const uint8_t *keep, int64_t *lengths, int *ticket)
{
    if (lvb_gpu_score_keep_submit_pre(c, n_cand, trees, topos, NULL, keep, ticket))
        return 1;
    return lvb_gpu_batch_wait(c, *ticket, lengths);
}
int lvb_gpu_accept_kept(lvb_gpu_ctx *c, int ticket, int64_t n_accept, const int64_t *cands)
{
    if (ticket < 0 || ticket >= MOCK_TICKETS)
        return fail("accept: no such ticket");
    for (int64_t i = 0; i < n_accept; i++)
    {
        const int64_t k = cands[i];
        if (k < 0 || k >= c->keep_n[ticket] || !c->keep_rows[ticket][k])
            return fail("accept: candidate was not kept");
        const int t = c->keep_tree[ticket][k];
        memcpy(c->rows[t], c->keep_rows[ticket][k], (size_t)c->B * c->nwords * 8);
        memcpy(c->changes[t], c->keep_chg[ticket][k], (size_t)c->B * sizeof(long));
        free(c->keep_rows[ticket][k]);
        free(c->keep_chg[ticket][k]);
        c->keep_rows[ticket][k] = NULL;
        c->keep_chg[ticket][k] = NULL;
        c->launches++;
    }
    return 0;
}

/* precompiled programs: the mock evaluates the topologies and ignores the programs */
int lvb_gpu_commit_batch_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                             const lvb_gpu_program_view *pre, int64_t *lengths)
{
    (void)pre;
    return lvb_gpu_commit_batch(c, n_cand, trees, topos, lengths);
}
int lvb_gpu_score_batch_submit_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                   const lvb_gpu_program_view *pre, int *ticket)
{
    (void)pre;
    return lvb_gpu_score_batch_submit(c, n_cand, trees, topos, ticket);
}

/* every mutate_deterministic candidate (TreeOperations.c:105-158), each by a FULL oracle pass:
 * the slowest, most literal reading of deterministic_hillclimb's loop (Solve.c:100-111) */
int lvb_gpu_nni_neighbourhood(lvb_gpu_ctx *c, int tree, int scratch_tree, const lvb_gpu_topology *tp, int64_t *lengths)
{
    if (!c->ready[tree] || scratch_tree == tree || !c->used[scratch_tree])
        return fail("neighbourhood: bad trees");
    long *l0 = (long *)malloc((size_t)c->B * sizeof(long)), *r0 = (long *)malloc((size_t)c->B * sizeof(long));
    long *par = (long *)malloc((size_t)c->B * sizeof(long));
    for (long v = 0; v < c->B; v++)
    {
        l0[v] = strided(tp->left, tp->stride_bytes, v);
        r0[v] = strided(tp->right, tp->stride_bytes, v);
    }
    for (long k = 0; k < tp->n_mod; k++)
    {
        l0[tp->mod_node[k]] = (long)tp->mod_left[k];
        r0[tp->mod_node[k]] = (long)tp->mod_right[k];
    }
    for (long v = 0; v < c->B; v++)
        if (v >= c->n || v == tp->root)
            par[l0[v]] = par[r0[v]] = v;
    for (long u = c->n; u < c->B; u++)
        for (int left = 0; left < 2; left++)
        {
            memcpy(c->left, l0, (size_t)c->B * sizeof(long));
            memcpy(c->right, r0, (size_t)c->B * sizeof(long));
            const long v = par[u], a = l0[u], b = r0[u];
            const long sis = l0[v] == u ? r0[v] : l0[v];
            const long moved = left ? a : b;
            if (l0[v] == u)
                c->right[v] = moved;
            else
                c->left[v] = moved;
            c->left[u] = left ? b : a;
            c->right[u] = sis;
            memcpy(c->scratch, c->rows[tree], (size_t)c->B * c->nwords * 8);
            memcpy(c->scratch_chg, c->changes[tree], (size_t)c->B * sizeof(long));
            oracle_dirty_all(c->n, c->nwords, c->scratch);
            lengths[2 * (u - c->n) + left] =
                oracle_getplen(c->n, c->nwords, c->left, c->right, (long)tp->root, c->scratch, c->scratch_chg);
        }
    c->ready[scratch_tree] = 0;
    c->launches += 3;
    free(l0);
    free(r0);
    free(par);
    return 0;
}

int64_t lvb_gpu_launch_count(lvb_gpu_ctx *c) { return c ? c->launches : 0; }
int lvb_gpu_synchronize(lvb_gpu_ctx *c)
{
    (void)c;
    return 0;
}

/* parity helpers the Python binding (lvb_b200/gpu.py: Engine) calls; they let the logic of the
 * `-m gpu` tests be tried on a CPU box by pointing the binding at this library */
int lvb_gpu_set_timing(lvb_gpu_ctx *c, int on)
{
    (void)c;
    (void)on;
    return 0;
}
int lvb_gpu_download_row(lvb_gpu_ctx *c, int tree, int64_t node, uint64_t *out)
{
    if (node < 0 || node >= c->B)
        return fail("node out of range");
    memcpy(out, (node < c->n ? c->leaves : c->rows[tree]) + (size_t)node * c->nwords, (size_t)c->nwords * 8);
    return 0;
}
/* entry points the mock does not model: present so that the binding loads, failing loudly */
#define ABSENT return fail("not modelled by the mock device")
int lvb_gpu_comm_init(lvb_gpu_ctx *c, int nranks, int rank, const void *id) { (void)c; (void)nranks; (void)rank; (void)id; ABSENT; }
int lvb_gpu_comm_unique_id(void *id) { (void)id; ABSENT; }
/* the host-only compilers (lvb_gpu_compile_schedule/_pieces/_up_pass) are NOT overridden: the host
 * layer uses the product's own, which need no device */
int lvb_gpu_kernel_time(lvb_gpu_ctx *c, double *ms, int64_t *k, int reset) { (void)c; (void)ms; (void)k; (void)reset; ABSENT; }
int lvb_gpu_last_kernel_ms(lvb_gpu_ctx *c, float *ms) { (void)c; (void)ms; ABSENT; }
int lvb_gpu_matrix_pack(lvb_gpu_ctx *c, int64_t n, int64_t m, const char *const *rows, int64_t wb, int64_t wc, int64_t mt)
{ (void)c; (void)n; (void)m; (void)rows; (void)wb; (void)wc; (void)mt; ABSENT; }
int lvb_gpu_rerun_last_batch(lvb_gpu_ctx *c, int64_t *lengths) { (void)c; (void)lengths; ABSENT; }
void *lvb_gpu_stream(lvb_gpu_ctx *c)
{
    (void)c;
    return NULL;
}
