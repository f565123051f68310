/* tree.cpp — LVB's tree model and proposal operators on the host (src/TreeOperations.c of the
 * reference), kept in C-style so that Solve-level code reads like the reference's.
 *
 * A host tree is B = 2n-3 TREESTACK_TREE_NODES followed, in the same allocation, by B flag
 * words (sitestate[0] of the reference: 0 = dirty).  The real state rows are on the device.
 * Every operator leaves parent/left/right and the dirty flags exactly as the reference's
 * operator does for the same random draws; tests/test_host_parity.py checks that against the
 * reference itself.
 */
#include <cstdlib>
#include <cstring>

#include "lvb_host.h"

namespace lvbhost
{

static const Lvb_bit_length CLEAN = ~(Lvb_bit_length)0; /* any non-zero word means "clean" */

/* While a ProposalScratch is installed, every node record is saved the first time an operator
 * is about to change it, so that the operator can run directly on the current tree and be
 * undone afterwards (propose()).  No journal: the operators behave exactly as before. */
static thread_local ProposalScratch *tl_journal = nullptr;

static inline void jt(TREESTACK_TREE_NODES *tree, long v)
{
    ProposalScratch *j = tl_journal;
    if (!j || j->stamp[(size_t)v] == j->cur)
        return;
    j->stamp[(size_t)v] = j->cur;
    j->node.push_back(v);
    j->before.push_back(tree[v]);
    j->flag_before.push_back(tree[v].sitestate[0]);
}

/* src/TreeOperations.c:914-955: one block, freed with a single free() */
TREESTACK_TREE_NODES *treealloc(Dataptr MSA, bool b_with_sitestate)
{
    const long B = MSA->numberofpossiblebranches;
    lvb_assert(B >= 2 * MIN_N - 3);
    lvb_assert(B <= 2 * MAX_N - 3);
    const size_t bytes = (size_t)B * sizeof(TREESTACK_TREE_NODES) + (b_with_sitestate ? (size_t)B * sizeof(Lvb_bit_length) : 0);
    TREESTACK_TREE_NODES *tree = (TREESTACK_TREE_NODES *)malloc(bytes);
    if (!tree)
        crash("out of memory: cannot allocate for %s", b_with_sitestate ? "tree with statesets" : "tree without statesets");
    Lvb_bit_length *flags = b_with_sitestate ? (Lvb_bit_length *)(tree + B) : nullptr;
    for (long i = 0; i < B; i++)
    {
        tree[i].parent = tree[i].left = tree[i].right = tree[i].changes = UNSET;
        tree[i].sitestate = flags ? flags + i : nullptr;
        if (flags)
            flags[i] = 0; /* dirty */
    }
    return tree;
}

/* src/TreeOperations.c:737-773: scalars always, flag words (the reference's state rows) on request;
 * dest keeps its own sitestate pointers */
void treecopy(Dataptr MSA, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, bool b_with_sitestate)
{
    const long B = MSA->numberofpossiblebranches;
    for (long i = 0; i < B; i++)
    {
        Lvb_bit_length *keep = dest[i].sitestate;
        dest[i] = src[i];
        dest[i].sitestate = keep;
    }
    if (b_with_sitestate)
        memcpy((void *)(dest + B), (const void *)(src + B), (size_t)B * sizeof(Lvb_bit_length));
}

/* src/TreeOperations.c:49-59 */
void nodeclear(TREESTACK_TREE_NODES *tree, long brnch)
{
    jt(tree, brnch);
    tree[brnch].left = tree[brnch].right = tree[brnch].parent = tree[brnch].changes = UNSET;
    tree[brnch].sitestate[0] = 0;
}

/* src/TreeOperations.c:78-86 */
void treeclear(Dataptr MSA, TREESTACK_TREE_NODES *tree)
{
    for (long i = 0; i < MSA->numberofpossiblebranches; i++)
        nodeclear(tree, i);
}

[[noreturn]] static void cr_bpnc(const TREESTACK_TREE_NODES *tree, long branch)
{
    const long p = tree[branch].parent;
    crash("internal error in tree array %p: branch record %ld says\n"
          "it has parent %ld, but branch record %ld has left child %ld\n"
          "and right child %ld",
          (const void *)tree, branch, p, p, tree[p].left, tree[p].right);
}

/* src/TreeOperations.c:657-676 */
long getsister(const TREESTACK_TREE_NODES *tree, long branch)
{
    const long p = tree[branch].parent;
    if (p == UNSET)
        return UNSET;
    if (tree[p].left == branch)
        return tree[p].right;
    if (tree[p].right == branch)
        return tree[p].left;
    cr_bpnc(tree, branch);
}

/* src/TreeOperations.c:88-103: dirty from `node` up to, not including, the root */
static void make_dirty_below(Dataptr MSA, TREESTACK_TREE_NODES *tree, long node)
{
    lvb_assert(node >= MSA->n);
    lvb_assert(tree[node].parent != UNSET);
    do
    {
        jt(tree, node);
        tree[node].sitestate[0] = 0;
        node = tree[node].parent;
    } while (tree[node].parent != UNSET);
}

/* the interchange shared by mutate_nni and mutate_deterministic (src/TreeOperations.c:124-155,
 * 176-205): u keeps `kept` and takes its former sibling c; `moved` goes up to u's parent v */
static void nni_at(Dataptr MSA, TREESTACK_TREE_NODES *tree, long u, bool move_right_child)
{
    const long v = tree[u].parent;
    const long a = tree[u].left, b = tree[u].right;
    const bool u_is_left = tree[v].left == u;
    const long c = u_is_left ? tree[v].right : tree[v].left;
    const long moved = move_right_child ? b : a;
    const long kept = move_right_child ? a : b;
    jt(tree, v);
    jt(tree, u);
    jt(tree, kept);
    jt(tree, c);
    jt(tree, moved);
    if (u_is_left)
        tree[v].right = moved;
    else
        tree[v].left = moved;
    tree[u].left = kept;
    tree[u].right = c;
    tree[kept].parent = tree[c].parent = u;
    tree[moved].parent = v;
    make_dirty_below(MSA, tree, u);
}

/* src/TreeOperations.c:105-158 */
void mutate_deterministic(Dataptr MSA, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root, long p,
                          bool left)
{
    lvb_assert(p != root);
    lvb_assert(p >= MSA->n);
    treecopy(MSA, dest, src, true);
    nni_at(MSA, dest, p, !left);
}

static void nni_inplace(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree)
{
    const long u = rng.randpint(MSA->numberofpossiblebranches - MSA->n - 1) + MSA->n;
    nni_at(MSA, tree, u, rng.uni() < 0.5);
}

/* src/TreeOperations.c:160-209: two draws, the branch then the side */
void mutate_nni(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root)
{
    (void)root;
    treecopy(MSA, dest, src, true);
    nni_inplace(MSA, rng, dest);
}

/* src/TreeOperations.c:211-234 */
static bool is_descendant(const TREESTACK_TREE_NODES *tree, long ancestor, long candidate)
{
    for (long p = tree[candidate].parent; p != UNSET; p = tree[p].parent)
        if (p == ancestor)
            return true;
    return false;
}

/* What mutate_spr and mutate_tbr share (src/TreeOperations.c:254-327 == :369-434): draw the
 * subtree `src` and the edge above `dest`, cut src out together with its parent node, and
 * put that node (excess) on the edge above dest with dest as its left child.  src is left
 * detached (parent UNSET); excess.right is still UNSET. */
struct Prune
{
    long src, dest, excess, parents_par;
};
static Prune prune_and_open(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree, long root)
{
    const long top = MSA->numberofpossiblebranches - 1;
    Prune p;
    do
        p.src = rng.randpint(top);
    while (p.src == root || p.src == tree[root].left || p.src == tree[root].right);
    const long src_parent = tree[p.src].parent;
    lvb_assert(src_parent != UNSET);
    const long src_sister = getsister(tree, p.src);
    lvb_assert(src_sister != UNSET);
    do
        p.dest = rng.randpint(top);
    while (p.dest == p.src || p.dest == src_parent || p.dest == src_sister || p.dest == root ||
           is_descendant(tree, p.src, p.dest));

    /* excise */
    jt(tree, src_parent);
    jt(tree, p.src);
    jt(tree, tree[src_parent].parent);
    jt(tree, src_sister);
    jt(tree, tree[p.dest].parent);
    jt(tree, p.dest);
    if (tree[src_parent].left == p.src)
        tree[src_parent].left = UNSET;
    else if (tree[src_parent].right == p.src)
        tree[src_parent].right = UNSET;
    else
        cr_bpnc(tree, p.src);
    tree[p.src].parent = UNSET;
    p.parents_par = tree[src_parent].parent;
    lvb_assert(p.parents_par != UNSET);
    if (tree[p.parents_par].left == src_parent)
        tree[p.parents_par].left = src_sister;
    else
        tree[p.parents_par].right = src_sister;
    tree[src_sister].parent = p.parents_par;
    p.excess = src_parent;
    nodeclear(tree, p.excess);

    /* open the destination edge */
    const long dest_parent = tree[p.dest].parent;
    if (tree[dest_parent].left == p.dest)
        tree[dest_parent].left = p.excess;
    else if (tree[dest_parent].right == p.dest)
        tree[dest_parent].right = p.excess;
    else
        crash("destination %ld is not a child of it's parent %ld\n", p.dest, dest_parent);
    tree[p.excess].parent = dest_parent;
    tree[p.excess].left = p.dest;
    tree[p.dest].parent = p.excess;
    return p;
}

/* src/TreeOperations.c:236-335 */
static void spr_inplace(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree, long root)
{
    const Prune p = prune_and_open(MSA, rng, tree, root);
    tree[p.excess].right = p.src;
    tree[p.src].parent = p.excess;
    make_dirty_below(MSA, tree, p.excess);
    if (p.parents_par != root)
        make_dirty_below(MSA, tree, p.parents_par);
}

void mutate_spr(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root)
{
    treecopy(MSA, dest, src, true);
    spr_inplace(MSA, rng, dest, root);
}

/* leaves under `top` in left-first depth-first order (count + addtoarray,
 * src/TreeOperations.c:544-574), following child links only */
static void subtree_leaves(const TREESTACK_TREE_NODES *tree, long top, std::vector<long> &leaves, std::vector<long> &work)
{
    leaves.clear();
    work.clear();
    work.push_back(top);
    while (!work.empty())
    {
        const long v = work.back();
        work.pop_back();
        const long l = tree[v].left, r = tree[v].right;
        if (l == UNSET && r == UNSET)
        {
            leaves.push_back(v);
            continue;
        }
        if (r != UNSET)
            work.push_back(r);
        if (l != UNSET)
            work.push_back(l);
    }
}

/* src/TreeOperations.c:337-541: SPR whose pruned subtree is first re-rooted on the edge above
 * a random one of its leaves (when it has more than two) */
static void tbr_inplace(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree, long root)
{
    const Prune p = prune_and_open(MSA, rng, tree, root);
    long src = p.src;

    static thread_local std::vector<long> leaves, work, up;
    subtree_leaves(tree, src, leaves, work);
    const long size = (long)leaves.size();
    const long oldroot = src;
    bool rerooted = false;
    size_t n_mid = 0;
    if (size > 2)
    {
        rerooted = true;
        long newroot;
        do
            newroot = leaves[(size_t)rng.randpint(size - 1)];
        while (newroot == tree[oldroot].left || newroot == tree[oldroot].right);

        /* the nodes from newroot's parent up to oldroot, as linked before anything is changed */
        up.clear();
        for (long v = tree[newroot].parent; v != oldroot; v = tree[v].parent)
            up.push_back(v);
        up.push_back(oldroot);

        for (long v : up)
            jt(tree, v);
        /* newroot's parent becomes the top of the subtree: children (its old parent, newroot) */
        const long first = up[0];
        long carried = tree[first].left == newroot ? tree[first].right : tree[first].left; /* newroot's sister */
        tree[first].parent = UNSET;
        tree[first].left = up[1];
        tree[first].right = newroot;
        long previous = first;
        /* every node strictly between: children become (its old parent, the sister carried
         * down from the node below), and it hands its own other child upwards */
        for (size_t k = 1; k + 1 < up.size(); k++)
        {
            const long cur = up[k], above = up[k + 1];
            long other = UNSET;
            if (tree[cur].left == previous)
                other = tree[cur].right;
            else if (tree[cur].right == previous)
                other = tree[cur].left;
            lvb_assert(other != UNSET);
            jt(tree, carried);
            tree[cur].right = carried;
            tree[cur].parent = previous;
            tree[cur].left = above;
            tree[above].parent = cur;
            tree[carried].parent = cur;
            carried = other;
            previous = cur;
        }
        n_mid = up.size() - 2;
        /* the old top keeps its off-path child and takes the last carried sister */
        tree[oldroot].parent = previous;
        if (tree[oldroot].left == previous)
            tree[oldroot].left = carried;
        else if (tree[oldroot].right == previous)
            tree[oldroot].right = carried;
        jt(tree, carried);
        tree[carried].parent = oldroot;
        src = tree[newroot].parent;
    }

    jt(tree, src);
    tree[p.excess].right = src;
    tree[src].parent = p.excess;

    if (rerooted)
    {
        make_dirty_below(MSA, tree, oldroot);
        for (size_t j = n_mid; j-- > 0;)
            make_dirty_below(MSA, tree, up[1 + j]);
        make_dirty_below(MSA, tree, src);
    }
    make_dirty_below(MSA, tree, p.excess);
    if (p.parents_par != root)
        make_dirty_below(MSA, tree, p.parents_par);
}

void mutate_tbr(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src_tree, long root)
{
    treecopy(MSA, dest, src_tree, true);
    tbr_inplace(MSA, rng, dest, root);
}

/* A proposal without the whole-tree copy: run the operator on the current tree itself with
 * the journal on, keep what it changed (records, child links that differ, nodes to recompute),
 * put the tree back.  Cost is O(nodes touched), not O(n).  `current` must be all clean. */
void propose(Dataptr MSA, Uni &rng, int op, TREESTACK_TREE_NODES *current, long root, ProposalScratch &scratch,
             Proposal &out)
{
    if (scratch.stamp.size() != (size_t)MSA->numberofpossiblebranches)
    {
        scratch.stamp.assign((size_t)MSA->numberofpossiblebranches, 0);
        scratch.cur = 0;
    }
    if (++scratch.cur == 0)
    {
        std::fill(scratch.stamp.begin(), scratch.stamp.end(), 0u);
        scratch.cur = 1;
    }
    scratch.node.clear();
    scratch.before.clear();
    scratch.flag_before.clear();
    tl_journal = &scratch;
    try
    {
        if (op == 0)
            nni_inplace(MSA, rng, current);
        else if (op == 1)
            spr_inplace(MSA, rng, current, root);
        else
            tbr_inplace(MSA, rng, current, root);
    }
    catch (...)
    {
        tl_journal = nullptr;
        throw;
    }
    tl_journal = nullptr;
    out.op = op;
    out.node = scratch.node;
    out.after.clear();
    out.dirty.clear();
    out.mod_node.clear();
    out.mod_left.clear();
    out.mod_right.clear();
    for (size_t i = 0; i < scratch.node.size(); i++)
    {
        const long v = scratch.node[i];
        const TREESTACK_TREE_NODES &was = scratch.before[i];
        out.after.push_back(current[v]);
        if (current[v].left != was.left || current[v].right != was.right)
        {
            out.mod_node.push_back(v);
            out.mod_left.push_back(current[v].left);
            out.mod_right.push_back(current[v].right);
        }
        if (current[v].sitestate[0] == 0 && v >= MSA->n)
            out.dirty.push_back(v);
        current[v] = was;
        current[v].sitestate[0] = scratch.flag_before[i];
    }
}

/* the accepted proposal becomes the current tree (SwapTrees, src/TreeOperations.c:1062-1080); flags stay clean */
void apply_proposal(TREESTACK_TREE_NODES *current, const Proposal &p)
{
    for (size_t i = 0; i < p.node.size(); i++)
    {
        Lvb_bit_length *keep = current[p.node[i]].sitestate;
        current[p.node[i]] = p.after[i];
        current[p.node[i]].sitestate = keep;
    }
}

/* src/TreeOperations.c:576-637.  Walking from the new root leaf up to the old one, every
 * node takes (its old parent, its old sister) as children; node numbers thereby shift one
 * vertex towards the old root along the path, and only they get different subtrees. */
long lvb_reroot(Dataptr MSA, TREESTACK_TREE_NODES *tree, long oldroot, long newroot, bool b_with_sitestate,
                std::vector<long> *path)
{
    lvb_assert(newroot < MSA->n);
    lvb_assert(newroot != oldroot);
    static thread_local std::vector<long> chain;
    chain.clear();
    for (long v = newroot; v != oldroot; v = tree[v].parent)
    {
        lvb_assert(v != UNSET);
        chain.push_back(v);
    }
    chain.push_back(oldroot);
    if (path)
        path->assign(chain.begin() + 1, chain.end() - 1);
    long previous = UNSET;
    for (size_t k = 0; k + 1 < chain.size(); k++)
    {
        const long cur = chain[k], above = chain[k + 1];
        long sister;
        if (tree[above].left == cur)
            sister = tree[above].right;
        else if (tree[above].right == cur)
            sister = tree[above].left;
        else
            crash("internal error in function lvb_reroot(): current\n"
                  "branch %ld has old parent %ld, but old parent does not\n"
                  "have it as a child",
                  cur, above);
        tree[cur].parent = previous;
        tree[cur].left = above;
        tree[cur].right = sister;
        tree[above].parent = cur;
        tree[sister].parent = cur;
        previous = cur;
    }
    tree[oldroot].left = tree[oldroot].right = UNSET;
    if (b_with_sitestate)
        for (long v = MSA->n; v < MSA->numberofpossiblebranches; v++)
            tree[v].sitestate[0] = 0;
    return oldroot;
}

/* src/TreeOperations.c:639-655 */
long arbreroot(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree, long oldroot, std::vector<long> *path)
{
    long newroot;
    do
        newroot = rng.randpint(MSA->n - 1);
    while (newroot == oldroot);
    lvb_reroot(MSA, tree, oldroot, newroot, true, path);
    return newroot;
}

/* src/TreeOperations.c:799-1060.  Three stages, each drawing from the chain's generator in
 * the reference's order: a random shape grown by splitting random leaves
 * (GenerateRandomTopology), a random assignment of taxa to its leaves (randleaf), and a
 * renumbering that puts taxon i in node i (tree_make_canonical) followed by rooting at taxon 0. */
void PullRandomTree(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree)
{
    const long B = MSA->numberofpossiblebranches, n = MSA->n;
    treeclear(MSA, tree);

    /* shape: nodes 0,1,2 are the root and its two children; split a random non-root leaf */
    std::vector<char> isleaf((size_t)B, 0);
    long nextfree = 0, leaves = 3;
    tree[0].parent = UNSET;
    isleaf[(size_t)nextfree++] = 1;
    tree[0].left = nextfree;
    tree[nextfree].parent = 0;
    isleaf[(size_t)nextfree++] = 1;
    tree[0].right = nextfree;
    tree[nextfree].parent = 0;
    isleaf[(size_t)nextfree++] = 1;
    while (leaves < n)
    {
        long grow;
        do
            grow = 1 + rng.randpint(nextfree - 2);
        while (!isleaf[(size_t)grow]);
        tree[grow].left = nextfree;
        tree[nextfree].parent = grow;
        isleaf[(size_t)nextfree++] = 1;
        tree[grow].right = nextfree;
        tree[nextfree].parent = grow;
        isleaf[(size_t)nextfree++] = 1;
        isleaf[(size_t)grow] = 0;
        leaves++;
    }

    /* taxa to leaves, in node order, by rejection */
    std::vector<char> used((size_t)n, 0);
    std::vector<long> object((size_t)B, UNSET);
    long assigned = 0;
    for (long i = 0; i < B; i++)
        if (isleaf[(size_t)i])
        {
            long cand;
            do
                cand = rng.randpint(n - 1);
            while (used[(size_t)cand]);
            object[(size_t)i] = cand;
            used[(size_t)cand] = 1;
            assigned++;
        }
    lvb_assert(assigned == n);

    /* renumber: sweep the nodes in order; a leaf holding taxon t != its own number changes
     * places with node t (every reference to either number follows); repeat until a sweep
     * makes no exchange.  Only the two records and their neighbours mention the two numbers. */
    auto exchange = [&](long a, long b) {
        long touched[8];
        int nt = 0;
        auto touch = [&](long x) {
            if (x == UNSET)
                return;
            for (int k = 0; k < nt; k++)
                if (touched[k] == x)
                    return;
            touched[nt++] = x;
        };
        const long ends[2] = {a, b};
        for (long e : ends)
        {
            touch(e);
            touch(tree[e].parent);
            touch(tree[e].left);
            touch(tree[e].right);
        }
        auto swapped = [&](long x) { return x == a ? b : (x == b ? a : x); };
        for (int k = 0; k < nt; k++)
        {
            TREESTACK_TREE_NODES &rec = tree[touched[k]];
            rec.parent = swapped(rec.parent);
            rec.left = swapped(rec.left);
            rec.right = swapped(rec.right);
        }
        TREESTACK_TREE_NODES ra = tree[a], rb = tree[b];
        Lvb_bit_length *sa = ra.sitestate, *sb = rb.sitestate;
        tree[a] = rb;
        tree[b] = ra;
        tree[a].sitestate = sa; /* flag words stay with the node number */
        tree[b].sitestate = sb;
    };
    bool swap_made;
    do
    {
        swap_made = false;
        for (long i = 0; i < B; i++)
        {
            const long t = object[(size_t)i];
            if (t != UNSET && t != i)
            {
                exchange(i, t);
                object[(size_t)i] = object[(size_t)t];
                object[(size_t)t] = t;
                swap_made = true;
            }
        }
    } while (swap_made);

    long root = UNSET;
    for (long i = 0; i < n; i++)
        if (tree[i].parent == UNSET)
        {
            lvb_assert(root == UNSET);
            root = i;
        }
    if (root != 0)
        lvb_reroot(MSA, tree, root, 0, true);
    for (long i = 0; i < n; i++)
        lvb_assert(object[(size_t)i] == i);
    for (long i = n; i < B; i++)
        lvb_assert(object[(size_t)i] == UNSET);
}

/* src/TreeOperations.c:1062-1080 */
void SwapTrees(TREESTACK_TREE_NODES **t1, long *r1, TREESTACK_TREE_NODES **t2, long *r2)
{
    TREESTACK_TREE_NODES *t = *t1;
    *t1 = *t2;
    *t2 = t;
    const long r = *r1;
    *r1 = *r2;
    *r2 = r;
}

/* src/TreeOperations.c:1500-1513.  The leaf rows were installed on the device once
 * (device_open); on the host only the flags remain: leaves clean, internal nodes dirty. */
void ss_init(Dataptr MSA, TREESTACK_TREE_NODES *tree)
{
    for (long i = 0; i < MSA->n; i++)
        tree[i].sitestate[0] = CLEAN;
    for (long i = MSA->n; i < MSA->numberofpossiblebranches; i++)
        tree[i].sitestate[0] = 0;
}

void collect_dirty(const Dataptr MSA, const TREESTACK_TREE_NODES *tree, std::vector<int64_t> &out)
{
    out.clear();
    for (long i = MSA->n; i < MSA->numberofpossiblebranches; i++)
        if (tree[i].sitestate[0] == 0)
            out.push_back(i);
}

void mark_all_clean(Dataptr MSA, TREESTACK_TREE_NODES *tree)
{
    for (long i = 0; i < MSA->numberofpossiblebranches; i++)
        tree[i].sitestate[0] = CLEAN;
}

/* src/TreeOperations.c:1144-1220 (ur_print): "(rootname,<left>,<right>);" with every other
 * internal node as "(<left>,<right>)"; trailing blanks of names dropped */
void lvb_treeprint(Dataptr MSA, FILE *stream, const TREESTACK_TREE_NODES *tree, long root)
{
    auto name = [&](long obj) {
        std::string s = MSA->rowtitle[(size_t)obj];
        while (!s.empty() && s.back() == ' ')
            s.pop_back();
        return s;
    };
    fprintf(stream, "(%s", name(root).c_str());
    /* explicit stack: positive = node to open, negative marker = close bracket */
    struct Item
    {
        long node;
        bool close;
    };
    std::vector<Item> st;
    st.push_back({tree[root].right, false});
    st.push_back({tree[root].left, false});
    bool comma = true;
    while (!st.empty())
    {
        const Item it = st.back();
        st.pop_back();
        if (it.close)
        {
            fputc(')', stream);
            comma = true;
            continue;
        }
        if (comma)
            fputc(',', stream);
        if (it.node < MSA->n)
        {
            fputs(name(it.node).c_str(), stream);
            comma = true;
        }
        else
        {
            fputc('(', stream);
            comma = false;
            st.push_back({0, true});
            st.push_back({tree[it.node].right, false});
            st.push_back({tree[it.node].left, false});
        }
    }
    fprintf(stream, ");\n");
    if (ferror(stream))
        crash("file error when writing unrooted tree");
}

} // namespace lvbhost
