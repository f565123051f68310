/* device.cpp — binding of an alignment to the GPU engine, and the reference-compatible,
 * one-tree-per-call getplen() (src/TreeEvaluation.c:49-270) on top of it. */
#include <cstdlib>
#include <map>

#include "lvb_host.h"

namespace lvbhost
{

static void gpu_check(int rc)
{
    if (rc != 0)
        crash("GPU error: %s", lvb_gpu_last_error()); /* errors of the device layer surface as crash(), src/Error.c:49 */
}

int64_t Device::launches() const { return (ctx ? lvb_gpu_launch_count(ctx) : 0) + extra_launches; }

/* Main.c:82 lvb_initialize() + Solve.c:516-519 (enc_mat, DNAToBinary): the packed leaf rows
 * are built once with the reference's layout and live on the device from here on */
void device_open(Device &dev, int device_index, Dataptr MSA, int max_trees, int64_t word_begin, int64_t word_count)
{
    /* large alignments: DNAToBinary on the device (pack_dna_kernel) — the characters go up once and are
     * packed there; 1000 x 200k: 0.03 s against 0.2 s for the host loop + upload of the words.  A symbol
     * the device rejects sends the call down the host route, which crashes with the reference's message */
    static const long long pack_min = [] {
        const char *e = getenv("LVB_GPU_DEVICE_PACK_MIN");
        return e ? atoll(e) : 8LL * 1000 * 1000;
    }();
    if ((long long)MSA->n * (long long)MSA->m >= pack_min)
    {
        dev.MSA = MSA;
        dev.index = device_index;
        gpu_check(lvb_gpu_init(device_index, &dev.ctx));
        if (word_count < 0)
            word_count = MSA->nwords - word_begin;
        dev.word_count = word_count;
        std::vector<const char *> text((size_t)MSA->n);
        for (long i = 0; i < MSA->n; i++)
            text[(size_t)i] = MSA->row[(size_t)i].data();
        if (lvb_gpu_matrix_pack(dev.ctx, MSA->n, MSA->m, text.data(), word_begin, word_count, max_trees) == 0)
            return;
        lvb_gpu_shutdown(dev.ctx);
        dev.ctx = nullptr;
    }
    std::vector<Lvb_bit_length> flat((size_t)MSA->n * (size_t)MSA->nwords);
    std::vector<Lvb_bit_length *> rows((size_t)MSA->n);
    for (long i = 0; i < MSA->n; i++)
        rows[(size_t)i] = flat.data() + (size_t)i * (size_t)MSA->nwords;
    DNAToBinary(MSA, rows.data());
    device_open_encoded(dev, device_index, MSA, rows.data(), max_trees, word_begin, word_count);
}

/* the same with the DNAToBinary rows already made (several contexts on one alignment) */
void device_open_encoded(Device &dev, int device_index, Dataptr MSA, Lvb_bit_length *const *rows, int max_trees,
                         int64_t word_begin, int64_t word_count)
{
    dev.MSA = MSA;
    dev.index = device_index;
    gpu_check(lvb_gpu_init(device_index, &dev.ctx));
    if (word_count < 0)
        word_count = MSA->nwords - word_begin;
    dev.word_count = word_count;
    gpu_check(lvb_gpu_matrix_upload(dev.ctx, MSA->n, MSA->nwords, (const uint64_t *const *)rows, word_begin,
                                    word_count, max_trees));
}

/* matchange (src/DataOperations.c:309-349) with the two O(n*m) text scans — constchar (:272-307)
 * and the MinimumTreeLength of the cut alignment (:53-105, :404) — done by the device in one pass
 * over the characters (lvb_gpu_column_stats); cutting the columns stays on the host. */
void matchange_device(Dataptr MSA, const Parameters &rc, int device_index)
{
    lvb_gpu_ctx *ctx = nullptr;
    gpu_check(lvb_gpu_init(device_index, &ctx));
    std::vector<const char *> rows((size_t)MSA->n);
    for (long i = 0; i < MSA->n; i++)
        rows[(size_t)i] = MSA->row[(size_t)i].data();
    std::vector<unsigned char> keep((size_t)MSA->m, 0);
    int64_t kept = 0, min_all = 0, min_kept = 0;
    const int rc_ = lvb_gpu_column_stats(ctx, MSA->n, MSA->m, rows.data(), keep.data(), &kept, &min_all, &min_kept);
    if (rc_ != 0)
    {
        const std::string msg = lvb_gpu_last_error();
        lvb_gpu_shutdown(ctx);
        crash("GPU error: %s", msg.c_str());
    }
    gpu_check(lvb_gpu_shutdown(ctx));
    matchange_apply(MSA, rc, keep.data(), (long)kept, (long)min_kept);
}

/* a further context on device_index with the alignment copied from `src` device to device */
void device_open_clone(Device &dev, int device_index, const Device &src, int max_trees)
{
    dev.MSA = src.MSA;
    dev.index = device_index;
    dev.word_count = src.word_count;
    gpu_check(lvb_gpu_init(device_index, &dev.ctx));
    gpu_check(lvb_gpu_matrix_clone(dev.ctx, src.ctx, max_trees));
}

void device_close(Device &dev)
{
    if (dev.ctx)
        lvb_gpu_shutdown(dev.ctx);
    dev.ctx = nullptr;
}

/* ---- getplen, the compatible route ----
 * The reference marks what must be recomputed by zeroing sitestate[0] of a node and makes a
 * proposal by copying the whole current tree, state rows included (treecopy).  The host copy
 * now moves one flag word per node; a clean flag holds 1 + the number of the device tree
 * whose row for that node is valid.  getplen(tree):
 *   - finds the device tree bound to this host array (created on first use),
 *   - if the clean flags name ANOTHER device tree, the host tree was produced by treecopy from
 *     that tree: the copy of the rows the reference did on the host is done now, device to
 *     device (lvb_gpu_tree_copy),
 *   - recomputes and stores the dirty nodes (lvb_gpu_tree_commit) and returns the length,
 *   - leaves every flag clean, as the reference does (TreeEvaluation.c:125-130).
 */
static Device *g_dev = nullptr;
static std::map<const TREESTACK_TREE_NODES *, int> g_bound;

void getplen_bind_device(Device *dev)
{
    g_dev = dev;
    g_bound.clear();
}

long getplen(Dataptr MSA, TREESTACK_TREE_NODES *tree, const Parameters &rc, long root, long *p_todo_arr,
             long *p_todo_arr_sum_changes, int *p_runs)
{
    (void)rc;
    (void)p_todo_arr;
    (void)p_todo_arr_sum_changes;
    (void)p_runs;
    if (!g_dev || !g_dev->ctx)
        crash("getplen: no GPU device is bound (there is no CPU path)");
    const long n = MSA->n, B = MSA->numberofpossiblebranches;
    int dt;
    bool fresh = false;
    auto it = g_bound.find(tree);
    if (it == g_bound.end())
    {
        gpu_check(lvb_gpu_tree_create(g_dev->ctx, &dt));
        g_bound[tree] = dt;
        fresh = true;
    }
    else
        dt = it->second;
    const Lvb_bit_length mine = (Lvb_bit_length)dt + 1;
    std::vector<int64_t> dirty;
    Lvb_bit_length foreign = 0;
    bool mixed = false;
    for (long v = n; v < B; v++)
    {
        const Lvb_bit_length f = tree[v].sitestate[0];
        if (f == 0)
            dirty.push_back(v);
        else if (f != mine)
        {
            if (foreign != 0 && foreign != f)
                mixed = true;
            foreign = f;
        }
    }
    bool all = (long)dirty.size() == B - n;
    if (!all && foreign != 0)
    {
        const bool known = !mixed && foreign - 1 < (Lvb_bit_length)INT32_MAX;
        if (known && lvb_gpu_tree_copy(g_dev->ctx, dt, (int)(foreign - 1)) == 0)
            fresh = false; /* the rows now are those of the source tree */
        else
            all = true; /* unknown provenance: recompute everything */
    }
    if (fresh && !all)
        all = true;
    lvb_gpu_topology tp = lvb_gpu_topology();
    tp.left = &tree[0].left;
    tp.right = &tree[0].right;
    tp.stride_bytes = (int64_t)sizeof(TREESTACK_TREE_NODES);
    tp.root = root;
    tp.dirty = dirty.data();
    tp.n_dirty = all ? -1 : (int64_t)dirty.size();
    int64_t length = 0;
    gpu_check(lvb_gpu_tree_commit(g_dev->ctx, dt, &tp, &length));
    for (long v = n; v < B; v++)
        tree[v].sitestate[0] = mine;
    lvb_assert(length > 0); /* TreeEvaluation.c:267 */
    return (long)length;
}

/* src/MemoryOperations.c:120-132: the scratch arrays belong to the CPU algorithm; kept so that
 * callers compile unchanged */
void alloc_memory_to_getplen(Dataptr MSA, long **p_todo_arr, long **p_todo_arr_sum_changes, int **p_runs)
{
    (void)MSA;
    *p_todo_arr = nullptr;
    *p_todo_arr_sum_changes = nullptr;
    *p_runs = nullptr;
}
void free_memory_to_getplen(long **p_todo_arr, long **p_todo_arr_sum_changes, int **p_runs)
{
    free(*p_todo_arr);
    free(*p_todo_arr_sum_changes);
    free(*p_runs);
}

} // namespace lvbhost
