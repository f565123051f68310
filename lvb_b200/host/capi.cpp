/* capi.cpp — C linkage over the host layer so that tests (ctypes) and bench.py can drive the
 * same functions the `lvb` binary uses.  Thin call-throughs only. */
#include <cstring>

#include <chrono>
#include <cstdlib>

#include "lvb_host.h"

using namespace lvbhost;

namespace
{
struct Handle
{
    DataStructure msa;
    Parameters rc;
    Uni rng;
};
thread_local std::string g_msg;

template <class F>
int guarded(F f)
{
    set_crash_throws(true);
    try
    {
        f();
        return 0;
    }
    catch (const LvbError &e)
    {
        g_msg = e.message;
        return 1;
    }
}
} // namespace

extern "C"
{

const char *lvbh_last_error(void) { return g_msg.c_str(); }

/* ---- random numbers ---- */
int lvbh_rng_stream(int seed, long count, double *uni_out, long upper, long *randpint_out)
{
    return guarded([&] {
        Uni r;
        r.rinit(seed);
        for (long i = 0; i < count; i++)
            uni_out[i] = r.uni();
        for (long i = 0; i < count; i++)
            randpint_out[i] = r.randpint(upper);
    });
}

/* ---- alignment ---- */
void *lvbh_msa_from_rows(long n, long m, const char *const *rows)
{
    Handle *h = new Handle();
    if (guarded([&] {
            matrix_from_rows(&h->msa, n, m, rows, nullptr);
            matchange(&h->msa, h->rc);
        }))
    {
        delete h;
        return nullptr;
    }
    return h;
}

void *lvbh_msa_load(const char *path, int format)
{
    Handle *h = new Handle();
    if (guarded([&] {
            phylip_dna_matrin(path, format, &h->msa);
            matchange(&h->msa, h->rc);
        }))
    {
        delete h;
        return nullptr;
    }
    return h;
}

void lvbh_msa_free(void *hv) { delete (Handle *)hv; }

/* out[0..6] = n, m, original_m, nwords, bytes, numberofpossiblebranches, min_len_tree */
void lvbh_msa_dims(void *hv, long *out)
{
    const DataStructure &d = ((Handle *)hv)->msa;
    out[0] = d.n;
    out[1] = d.m;
    out[2] = d.original_m;
    out[3] = d.nwords;
    out[4] = d.bytes;
    out[5] = d.numberofpossiblebranches;
    out[6] = d.min_len_tree;
}

void lvbh_msa_row(void *hv, long i, char *out)
{
    const DataStructure &d = ((Handle *)hv)->msa;
    memcpy(out, d.row[(size_t)i].data(), (size_t)d.m);
}

void lvbh_msa_title(void *hv, long i, char *out, long cap)
{
    const DataStructure &d = ((Handle *)hv)->msa;
    strncpy(out, d.rowtitle[(size_t)i].c_str(), (size_t)cap - 1);
    out[cap - 1] = '\0';
}

int lvbh_dna_to_binary(void *hv, uint64_t *out)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        std::vector<Lvb_bit_length *> rows((size_t)h->msa.n);
        for (long i = 0; i < h->msa.n; i++)
            rows[(size_t)i] = out + i * h->msa.nwords;
        DNAToBinary(&h->msa, rows.data());
    });
}

/* ---- trees (the handle's own generator supplies the draws) ---- */
void lvbh_rinit(void *hv, int seed) { ((Handle *)hv)->rng.rinit(seed); }
double lvbh_uni(void *hv) { return ((Handle *)hv)->rng.uni(); }

void *lvbh_tree_alloc(void *hv) { return treealloc(&((Handle *)hv)->msa, true); }
void lvbh_tree_free(void *t) { free(t); }

int lvbh_pull_random_tree(void *hv, void *t)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        PullRandomTree(&h->msa, h->rng, (TREESTACK_TREE_NODES *)t);
        ss_init(&h->msa, (TREESTACK_TREE_NODES *)t);
    });
}

void lvbh_tree_mark_clean(void *hv, void *t) { mark_all_clean(&((Handle *)hv)->msa, (TREESTACK_TREE_NODES *)t); }

/* kind: 0 nni, 1 spr, 2 tbr */
int lvbh_mutate(void *hv, int kind, void *dst, const void *src, long root)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        TREESTACK_TREE_NODES *d = (TREESTACK_TREE_NODES *)dst;
        const TREESTACK_TREE_NODES *s = (const TREESTACK_TREE_NODES *)src;
        if (kind == 0)
            mutate_nni(&h->msa, h->rng, d, s, root);
        else if (kind == 1)
            mutate_spr(&h->msa, h->rng, d, s, root);
        else
            mutate_tbr(&h->msa, h->rng, d, s, root);
    });
}

int lvbh_mutate_deterministic(void *hv, void *dst, const void *src, long root, long p, int left)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        mutate_deterministic(&h->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, root, p, left != 0);
    });
}

/* returns the new root; path_out receives the internal nodes whose rows change */
long lvbh_arbreroot(void *hv, void *t, long root, long *path_out, long *n_path)
{
    Handle *h = (Handle *)hv;
    long newroot = -1;
    std::vector<long> path;
    if (guarded([&] { newroot = arbreroot(&h->msa, h->rng, (TREESTACK_TREE_NODES *)t, root, &path); }))
        return -1;
    if (n_path)
        *n_path = (long)path.size();
    if (path_out)
        for (size_t i = 0; i < path.size(); i++)
            path_out[i] = path[i];
    return newroot;
}

int lvbh_reroot(void *hv, void *t, long oldroot, long newroot)
{
    Handle *h = (Handle *)hv;
    return guarded([&] { lvb_reroot(&h->msa, (TREESTACK_TREE_NODES *)t, oldroot, newroot, true); });
}

void lvbh_treecopy(void *hv, void *dst, const void *src)
{
    treecopy(&((Handle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, true);
}

void lvbh_tree_get(void *hv, const void *tv, long *parent, long *left, long *right, int *dirty)
{
    const DataStructure &d = ((Handle *)hv)->msa;
    const TREESTACK_TREE_NODES *t = (const TREESTACK_TREE_NODES *)tv;
    for (long i = 0; i < d.numberofpossiblebranches; i++)
    {
        parent[i] = t[i].parent;
        left[i] = t[i].left;
        right[i] = t[i].right;
        dirty[i] = t[i].sitestate[0] == 0;
    }
}

void lvbh_tree_set(void *hv, void *tv, const long *parent, const long *left, const long *right)
{
    const DataStructure &d = ((Handle *)hv)->msa;
    TREESTACK_TREE_NODES *t = (TREESTACK_TREE_NODES *)tv;
    for (long i = 0; i < d.numberofpossiblebranches; i++)
    {
        t[i].parent = parent[i];
        t[i].left = left[i];
        t[i].right = right[i];
    }
    ss_init(&((Handle *)hv)->msa, t);
}

/* a sequence of pushes into one fresh tree stack: results[i] = what the push of trees[i] returned (1 = new, 0 = a
 * duplicate); returns the final size.  The tests run the same sequence through the reference. */
long lvbh_treestack_sequence(void *hv, long count, void *const *trees, const long *roots, int hash, long *results)
{
    Handle *h = (Handle *)hv;
    TREESTACK stack = CreateNewTreestack();
    for (long i = 0; i < count; i++)
        results[i] = hash ? CompareHashTreeToHashstack(&h->msa, &stack, (const TREESTACK_TREE_NODES *)trees[i], roots[i])
                          : CompareTreeToTreestack(&h->msa, &stack, (const TREESTACK_TREE_NODES *)trees[i], roots[i]);
    return stack.next;
}

int lvbh_same_topology(void *hv, const void *t1, long r1, const void *t2, long r2)
{
    Handle *h = (Handle *)hv;
    return topology_key(&h->msa, (const TREESTACK_TREE_NODES *)t1, r1) ==
           topology_key(&h->msa, (const TREESTACK_TREE_NODES *)t2, r2);
}

int lvbh_treeprint(void *hv, const void *t, long root, const char *path)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        FILE *fp = fopen(path, "a");
        if (!fp)
            crash("cannot open %s", path);
        lvb_treeprint(&h->msa, fp, (const TREESTACK_TREE_NODES *)t, root);
        fclose(fp);
    });
}

/* ---- getplen, compatible route (needs a GPU) ---- */
static Device g_capi_dev;
int lvbh_device_open(void *hv, int device, int max_trees)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        device_open(g_capi_dev, device, &h->msa, max_trees);
        getplen_bind_device(&g_capi_dev);
    });
}
void lvbh_device_close(void)
{
    getplen_bind_device(nullptr);
    device_close(g_capi_dev);
}
long lvbh_getplen(void *hv, void *t, long root)
{
    Handle *h = (Handle *)hv;
    long len = -1;
    if (guarded([&] { len = getplen(&h->msa, (TREESTACK_TREE_NODES *)t, h->rc, root, nullptr, nullptr, nullptr); }))
        return -1;
    return len;
}

/* ---- the search: GetSoln for `replicates` chains on one device.
 * out (per chain, 7 longs): final length, rearrangements, topologies, initial length, trees scored, score launches, groups
 * start_t (per chain): starting temperature.  tsv_path / outtree_path may be NULL (chain 0 only;
 * further chains write outtree_path.r<k>). ---- */
int lvbh_search(void *hv, int device, int seed, int algorithm, int cooling, int max_trees_kept, int replicates,
                int speculation, long max_rearrangements, int host_threads, int groups, const char *tsv_path, const char *outtree_path, long *out,
                double *start_t, double *seconds)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        Parameters rc = h->rc;
        rc.seed = seed;
        rc.algorithm_selection = algorithm;
        rc.cooling_schedule = cooling;
        rc.n_number_max_trees = max_trees_kept;
        rc.replicates = replicates;
        rc.speculation = speculation;
        rc.max_rearrangements = max_rearrangements;
        rc.n_processors_available = host_threads < 1 ? 1 : host_threads;
        rc.groups = groups < 0 ? 0 : groups;
        rc.device = device;
        Device dev;
        const auto t_open = std::chrono::steady_clock::now();
        device_open(dev, device, &h->msa, replicates);
        if (getenv("LVB_GPU_PROFILE"))
            fprintf(stderr, "[lvbh_search] device_open (alignment packed and on the device, %d trees allocated): %.3f s\n", replicates,
                    std::chrono::duration<double>(std::chrono::steady_clock::now() - t_open).count());
        FILE *tsv = tsv_path ? fopen(tsv_path, "w") : nullptr;
        std::vector<TREESTACK> stacks;
        std::vector<SearchResult> results;
        GetSolnBatched(&h->msa, rc, dev, replicates, stacks, results, nullptr, tsv);
        if (tsv)
            fclose(tsv);
        for (int r = 0; r < replicates; r++)
        {
            out[7 * r + 0] = results[(size_t)r].final_length;
            out[7 * r + 1] = results[(size_t)r].rearrangements;
            out[7 * r + 2] = stacks[(size_t)r].next;
            out[7 * r + 3] = results[(size_t)r].initial_length;
            out[7 * r + 4] = results[(size_t)r].getplen_calls;
            out[7 * r + 5] = results[(size_t)r].launches_scored;
            out[7 * r + 6] = results[(size_t)r].groups;
            start_t[r] = results[(size_t)r].start_temperature;
            if (outtree_path)
            {
                const std::string nm = r == 0 ? std::string(outtree_path) : std::string(outtree_path) + ".r" + std::to_string(r);
                FILE *fp = fopen(nm.c_str(), "w");
                if (!fp)
                    crash("cannot open %s", nm.c_str());
                PrintTreestack(&h->msa, &stacks[(size_t)r], fp);
                fclose(fp);
            }
        }
        if (seconds)
            *seconds = results[0].seconds;
        device_close(dev);
    });
}

/* ---- deterministic_hillclimb (Solve.c:64-142) from `tree`/`root`.  out: final length, iterations
 * (current_iter), trees on the stack, neighbourhood calls; outtree_path (may be NULL) receives
 * the stack as PrintTreestack writes it. ---- */
int lvbh_hillclimb(void *hv, int device, const void *tree, long root, const char *outtree_path, long *out)
{
    Handle *h = (Handle *)hv;
    return guarded([&] {
        Device dev;
        device_open(dev, device, &h->msa, 2);
        TREESTACK stack = CreateNewTreestack();
        long iter = 0, calls = 0;
        const long len = deterministic_hillclimb(&h->msa, dev, &stack, (const TREESTACK_TREE_NODES *)tree, h->rc, root,
                                                 nullptr, &iter, false, &calls);
        out[0] = len;
        out[1] = iter;
        out[2] = stack.next;
        out[3] = calls;
        if (outtree_path)
        {
            FILE *fp = fopen(outtree_path, "w");
            if (!fp)
                crash("cannot open %s", outtree_path);
            PrintTreestack(&h->msa, &stack, fp);
            fclose(fp);
        }
        device_close(dev);
    });
}

} /* extern "C" */
