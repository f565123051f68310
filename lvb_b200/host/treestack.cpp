/* treestack.cpp — the stack of best trees (src/Treestack.c, src/Hash.cpp of the reference).
 *
 * Host bookkeeping, touched only when a proposal is accepted and is at least as good as the
 * best so far (src/Solve.c:309-320).  It is kept because the SA trajectory depends on it:
 * `accepted` counts pushes of NEW topologies and drives the cooling schedule.
 *
 * Duplicate rule of the reference: two trees are the same when, both re-rooted at taxon 0,
 * their sorted lists of sorted clade sets are equal (makesets/fillsets/Sort/setstcmp,
 * src/TreeOperations.c:1222-1265, 1344-1498) — i.e. equal unrooted topologies.  Here the same
 * equivalence is decided on a canonical text of the topology: children ordered by the smallest
 * taxon beneath them, O(n) per tree instead of O(n^2) set building.
 */
#include <algorithm>
#include <cstring>

#include "lvb_host.h"

namespace lvbhost
{

std::string topology_key(Dataptr MSA, const TREESTACK_TREE_NODES *tree_in, long root)
{
    const long B = MSA->numberofpossiblebranches, n = MSA->n;
    static thread_local std::vector<TREESTACK_TREE_NODES> copy;
    copy.assign(tree_in, tree_in + B);
    TREESTACK_TREE_NODES *tree = copy.data();
    if (root != 0)
        lvb_reroot(MSA, tree, root, 0, false);
    /* smallest taxon under every node, children before parents */
    static thread_local std::vector<int32_t> least, order;
    least.assign((size_t)B, 0);
    order.clear();
    order.push_back((int32_t)tree[0].left);
    order.push_back((int32_t)tree[0].right);
    for (size_t h = 0; h < order.size(); h++)
    {
        const long v = order[h];
        if (v >= n)
        {
            order.push_back((int32_t)tree[v].left);
            order.push_back((int32_t)tree[v].right);
        }
    }
    lvb_assert((long)order.size() == B - 1);
    for (size_t h = order.size(); h-- > 0;)
    {
        const long v = order[h];
        least[(size_t)v] = v < n ? (int32_t)v : std::min(least[(size_t)tree[v].left], least[(size_t)tree[v].right]);
    }
    /* pre-order text: taxon numbers, -1 opens a clade, -2 closes it */
    std::vector<int32_t> text;
    text.reserve((size_t)(3 * n));
    std::vector<int32_t> st;
    auto push_children = [&](long v) {
        long a = tree[v].left, b = tree[v].right;
        if (least[(size_t)a] > least[(size_t)b])
            std::swap(a, b);
        st.push_back((int32_t)b);
        st.push_back((int32_t)a);
    };
    push_children(0);
    while (!st.empty())
    {
        const int32_t v = st.back();
        st.pop_back();
        if (v == -2)
            text.push_back(-2);
        else if (v < n)
            text.push_back(v);
        else
        {
            text.push_back(-1);
            st.push_back(-2);
            push_children(v);
        }
    }
    return std::string((const char *)text.data(), text.size() * sizeof(int32_t));
}

TREESTACK CreateNewTreestack(void) { return TREESTACK(); }

long CountTreestack(const TREESTACK &s) { return s.next; }

void ClearTreestack(TREESTACK *sp) { sp->next = 0; }

/* src/Treestack.c:78-95 (PushCurrentTreeToStack): copy without state flags */
static void push(Dataptr MSA, TREESTACK *sp, const TREESTACK_TREE_NODES *tree, long root, std::string &&key)
{
    lvb_assert(root < MSA->n);
    if ((size_t)sp->next == sp->stack.size())
        sp->stack.emplace_back();
    TREESTACK_TREES &slot = sp->stack[(size_t)sp->next];
    slot.tree.assign(tree, tree + MSA->numberofpossiblebranches);
    for (auto &nd : slot.tree)
        nd.sitestate = nullptr;
    slot.root = root;
    slot.key = std::move(key);
    sp->next++;
}

/* src/Treestack.c:231-306: 1 if pushed, 0 if the topology is already on the stack */
long CompareTreeToTreestack(Dataptr MSA, TREESTACK *sp, const TREESTACK_TREE_NODES *tree, long root)
{
    std::string key = topology_key(MSA, tree, root);
    for (long i = sp->next - 1; i >= 0; i--)
        if (sp->stack[(size_t)i].key == key)
            return 0;
    push(MSA, sp, tree, root, std::move(key));
    return 1;
}

/* src/Hash.cpp:49-90.  Decides on a separate set of keys (the reference's static hashSet),
 * which is emptied only when the stack is empty at call time and is not told about trees
 * pushed through CompareTreeToTreestack — kept that way, see SURVEY.md Appendix C. */
long CompareHashTreeToHashstack(Dataptr MSA, TREESTACK *sp, const TREESTACK_TREE_NODES *tree, long root)
{
    std::string key = topology_key(MSA, tree, root);
    if (sp->next == 0)
        sp->hash_keys.clear();
    else if (std::find(sp->hash_keys.begin(), sp->hash_keys.end(), key) != sp->hash_keys.end())
        return 0;
    sp->hash_keys.push_back(key);
    push(MSA, sp, tree, root, std::move(key));
    return 1;
}

/* src/Treestack.c:352-374 */
long PullTreefromTreestack(Dataptr MSA, TREESTACK_TREE_NODES *tree, long *root, TREESTACK *sp)
{
    if (sp->next < 1)
        return 0;
    sp->next--;
    treecopy(MSA, tree, sp->stack[(size_t)sp->next].tree.data(), false);
    *root = sp->stack[(size_t)sp->next].root;
    return 1;
}

/* src/Treestack.c:376-414: every tree re-rooted at taxon 0, in push order */
int PrintTreestack(Dataptr MSA, TREESTACK *sp, FILE *outfp)
{
    std::vector<TREESTACK_TREE_NODES> work((size_t)MSA->numberofpossiblebranches);
    for (long i = 0; i < sp->next; i++)
    {
        const TREESTACK_TREES &e = sp->stack[(size_t)i];
        std::copy(e.tree.begin(), e.tree.end(), work.begin());
        if (e.root != 0)
            lvb_reroot(MSA, work.data(), e.root, 0, false);
        lvb_treeprint(MSA, outfp, work.data(), 0);
    }
    if (fflush(outfp) != 0)
        crash("file write error when writing best trees");
    return (int)sp->next;
}

} // namespace lvbhost
