/* TreeEvaluation_gpu.c — drop-in replacement for the reference's src/TreeEvaluation.c.
 *
 * Compiled like every reference source (as C++, /root/reference/Makefile:74-81) against the
 * reference's OWN headers (-I <reference>/src: LVB.h, DataStructure.h), it defines getplen()
 * with the exact contract of src/LVB.h:187 / src/TreeEvaluation.c:49-270 and takes the place of
 * TreeEvaluation.o in libLVB.a (Makefile:113-130).  The callers stay unmodified:
 * Solve.c:95,110,200,300,559 and StartingTemperature.c:98,128 call it as before, TreeOperations.c
 * keeps marking dirty nodes through sitestate[0] == 0 (:49-59, 88-103, 631-635, 946-950).
 * Underneath, every length comes from the B200 engine through include/lvb_gpu.h; there is no
 * CPU path here (alloc/free_memory_to_getplen live in MemoryOperations.c and are untouched:
 * their scratch arrays are accepted and ignored).
 *
 * Protocol (the same as lvb_b200/host/device.cpp:getplen, INTEGRATION.md route A):
 *   - leaf rows are uploaded once, lazily, from BranchArray[0..n-1].sitestate (ss_init,
 *     TreeOperations.c:1500-1513, put the DNAToBinary rows there);
 *   - each host tree array is bound to one device tree on first use;
 *   - a CLEAN internal node carries in sitestate[0] a non-zero tag (version << 8 | 1 + device tree)
 *     naming the device tree, and the evaluation of it, whose row is valid for that node.  The
 *     reference's treecopy (TreeOperations.c:737-773) copies the tag with the row, so a tree that
 *     was copied from another one says so itself and the copy of the rows is repeated device to
 *     device (lvb_gpu_tree_copy).  A tag whose version is not the current one of the tree it names
 *     (that tree was re-evaluated since) makes every node dirty: correct, merely slower;
 *   - dirty nodes (tag 0) become the explicit dirty list of lvb_gpu_tree_commit, which
 *     recomputes and stores exactly those rows and returns the length with the root step
 *     (TreeEvaluation.c:238-264); afterwards every internal node is tagged clean, as the
 *     reference leaves them (:125-130, :228-229).
 * The host copies of the internal state rows are never read by anyone but this file
 * (SURVEY.md §8(b) "Outputs / side effects"), so only word 0 of them is maintained.
 *
 * Errors of the device layer surface as crash() (src/Error.c:49-62).
 */
#include "LVB.h"

#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "lvb_gpu.h"

#define LVB_DROPIN_MAX_TREES 16

static lvb_gpu_ctx *g_ctx = NULL;
static long g_n = 0, g_nwords = 0;
static const void *g_leaf0 = NULL;     /* identity of the alignment that is on the device */
static Lvb_bit_length g_leaf0_word = 0;
static struct
{
    const TREESTACK_TREE_NODES *host;  /* host array bound to device tree i, or NULL */
    uint64_t version;                  /* evaluation counter value of its last commit */
    int used;
} g_tree[LVB_DROPIN_MAX_TREES];
static uint64_t g_version = 0;
static int64_t *g_dirty = NULL;
static long g_dirty_cap = 0;

static void gpu_check(int rc)
{
    if (rc != 0)
        crash("GPU error: %s", lvb_gpu_last_error());
}

static void report_at_exit(void)
{
    if (g_ctx && getenv("LVB_GPU_REPORT"))
        fprintf(stderr, "lvb_gpu kernel launches: %lld\n", (long long)lvb_gpu_launch_count(g_ctx));
    if (g_ctx)
        lvb_gpu_shutdown(g_ctx);
    g_ctx = NULL;
}

static void open_device(Dataptr MSA, const TREESTACK_TREE_NODES *BranchArray)
{
    if (!g_ctx)
    {
        const char *dev = getenv("LVB_GPU_DEVICE");
        gpu_check(lvb_gpu_init(dev ? atoi(dev) : 0, &g_ctx));
        atexit(report_at_exit);
    }
    /* the packed alignment: DataOperations.c:164-249 output as ss_init left it in the leaves */
    const uint64_t **rows = (const uint64_t **)malloc((size_t)MSA->n * sizeof *rows);
    if (!rows)
        crash("out of memory: leaf row pointers");
    for (long i = 0; i < MSA->n; i++)
        rows[i] = (const uint64_t *)BranchArray[i].sitestate;
    gpu_check(lvb_gpu_matrix_upload(g_ctx, MSA->n, MSA->nwords, rows, 0, MSA->nwords, LVB_DROPIN_MAX_TREES));
    free(rows);
    g_n = MSA->n;
    g_nwords = MSA->nwords;
    memset(g_tree, 0, sizeof g_tree);
}

static int bind_tree(const TREESTACK_TREE_NODES *BranchArray)
{
    for (int i = 0; i < LVB_DROPIN_MAX_TREES; i++)
        if (g_tree[i].used && g_tree[i].host == BranchArray)
            return i;
    int dt = -1;
    if (lvb_gpu_tree_create(g_ctx, &dt) != 0)
    {
        /* every device tree is bound to some host array, most of them freed long ago (Anneal and
         * StartingTemperature free theirs, Solve.c:475-476): forget all bindings.  Tags that still
         * name them fail the version check below and cost one full pass. */
        for (int i = 0; i < LVB_DROPIN_MAX_TREES; i++)
            if (g_tree[i].used)
            {
                gpu_check(lvb_gpu_tree_destroy(g_ctx, i));
                g_tree[i].used = 0;
                g_tree[i].host = NULL;
            }
        gpu_check(lvb_gpu_tree_create(g_ctx, &dt));
    }
    if (dt < 0 || dt >= LVB_DROPIN_MAX_TREES)
        crash("GPU error: device tree %d out of range", dt);
    g_tree[dt].used = 1;
    g_tree[dt].host = BranchArray;
    g_tree[dt].version = 0; /* never evaluated */
    return dt;
}

long getplen(Dataptr restrict MSA, TREESTACK_TREE_NODES *BranchArray, Parameters rcstruct, const long root,
             long *restrict p_todo_arr, long *p_todo_arr_sum_changes, int *p_runs)
{
    (void)rcstruct; /* unused by the reference's body too */
    (void)p_todo_arr;
    (void)p_todo_arr_sum_changes;
    (void)p_runs;
    const long n = MSA->n, B = MSA->numberofpossiblebranches;
    /* a new alignment (or the first call): upload the leaf rows */
    if (!g_ctx || g_n != n || g_nwords != MSA->nwords || g_leaf0 != (const void *)MSA ||
        g_leaf0_word != BranchArray[0].sitestate[0])
    {
        open_device(MSA, BranchArray);
        g_leaf0 = (const void *)MSA;
        g_leaf0_word = BranchArray[0].sitestate[0];
    }
    const int dt = bind_tree(BranchArray);
    if (B - n > g_dirty_cap)
    {
        free(g_dirty);
        g_dirty = (int64_t *)malloc((size_t)(B - n) * sizeof *g_dirty);
        if (!g_dirty)
            crash("out of memory: dirty list");
        g_dirty_cap = B - n;
    }
    const Lvb_bit_length mine = g_tree[dt].version ? ((Lvb_bit_length)g_tree[dt].version << 8 | (Lvb_bit_length)(dt + 1)) : 0;
    long n_dirty = 0;
    Lvb_bit_length foreign = 0;
    int mixed = 0;
    for (long v = n; v < B; v++) /* TreeEvaluation.c:191-202: which nodes are dirty */
    {
        const Lvb_bit_length f = BranchArray[v].sitestate[0];
        if (f == 0)
            g_dirty[n_dirty++] = v;
        else if (f != mine || mine == 0)
        {
            if (foreign != 0 && foreign != f)
                mixed = 1;
            foreign = f;
        }
    }
    int all = n_dirty == B - n;
    if (!all && foreign != 0)
    {
        /* clean nodes that name another evaluation: the tree was made by treecopy from it */
        const int src = (int)(foreign & 0xff) - 1;
        const uint64_t ver = (uint64_t)(foreign >> 8);
        if (!mixed && src >= 0 && src < LVB_DROPIN_MAX_TREES && src != dt && g_tree[src].used &&
            g_tree[src].version == ver && ver != 0)
            gpu_check(lvb_gpu_tree_copy(g_ctx, dt, src)); /* the rows treecopy copied on the host, device to device */
        else
            all = 1; /* unknown or outdated provenance: recompute everything */
    }
    if (!all && foreign == 0 && mine == 0)
        all = 1;
    lvb_gpu_topology tp;
    memset(&tp, 0, sizeof tp);
    tp.left = (const int64_t *)&BranchArray[0].left;
    tp.right = (const int64_t *)&BranchArray[0].right;
    tp.stride_bytes = (int64_t)sizeof(TREESTACK_TREE_NODES);
    tp.root = root;
    tp.dirty = g_dirty;
    tp.n_dirty = all ? -1 : (int64_t)n_dirty;
    int64_t length = 0;
    gpu_check(lvb_gpu_tree_commit(g_ctx, dt, &tp, &length));
    g_tree[dt].version = ++g_version;
    const Lvb_bit_length tag = (Lvb_bit_length)g_tree[dt].version << 8 | (Lvb_bit_length)(dt + 1);
    for (long v = n; v < B; v++)
        BranchArray[v].sitestate[0] = tag; /* clean, TreeEvaluation.c:125-130 */
    lvb_assert(length > 0); /* :267 */
    return (long)length;
}
