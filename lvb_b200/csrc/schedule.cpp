/* schedule.cpp — see schedule.h. */
#include "schedule.h"

#include <algorithm>

namespace lvbgpu
{

static inline int64_t strided(const int64_t *base, int64_t stride_bytes, int64_t i)
{
    return *reinterpret_cast<const int64_t *>(reinterpret_cast<const char *>(base) + i * stride_bytes);
}

ScheduleCompiler::ScheduleCompiler(int64_t n)
    : n_(n), B_(2 * n - 3), cur_(0), stamp_(B_ > 0 ? B_ : 0, 0), child_(stamp_.size(), 0), seen_(stamp_.size(), 0),
      mod_(stamp_.size(), 0), mod_at_(stamp_.size(), 0), need_(stamp_.size(), 0), size_(stamp_.size(), 0),
      cl_(stamp_.size(), 0), cr_(stamp_.size(), 0), pos_(stamp_.size(), 0), korder_(stamp_.size(), 0),
      k0_(stamp_.size(), 0), k1_(stamp_.size(), 0), ksize_(stamp_.size(), 0), kneed_(stamp_.size(), 0),
      klevel_(stamp_.size(), 0), kstart_(stamp_.size(), 0), kdepth_(stamp_.size(), 0), kpush_(stamp_.size(), 0),
      kcut_(stamp_.size(), 0)
{
}

bool ScheduleCompiler::compile(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                               std::string &err, int piece_size, std::vector<Piece> *pieces)
{
    return compile_impl(topo, store, out, n_slots, err, piece_size, pieces, topo.n_dirty < 0);
}

/* whole: try the whole-tree route (every internal node dirty, one sweep from the root's children);
 * anything unusual there re-enters with whole = false, where the general route names the fault */
bool ScheduleCompiler::compile_impl(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                                    std::string &err, int piece_size, std::vector<Piece> *pieces, bool whole)
{
    if (piece_size > 0 && (!store || !pieces))
    {
        err = "schedule: only store programs can be cut into pieces";
        return false;
    }
    if (n_ < 3)
    {
        err = "schedule: need at least 3 leaves";
        return false;
    }
    if (++cur_ == 0)
    { /* stamp wrapped: start over */
        std::fill(stamp_.begin(), stamp_.end(), 0u);
        std::fill(child_.begin(), child_.end(), 0u);
        std::fill(seen_.begin(), seen_.end(), 0u);
        std::fill(mod_.begin(), mod_.end(), 0u);
        cur_ = 1;
    }
    const int64_t root = topo.root;
    if (root < 0 || root >= n_)
    {
        err = "schedule: root " + std::to_string(root) + " is not a leaf";
        return false;
    }
    /* child links: the overlay wins over the base arrays */
    for (int64_t i = 0; i < topo.n_mod; i++)
    {
        const int64_t v = topo.mod_node[i];
        if (v < 0 || v >= B_)
        {
            err = "schedule: overlay names node " + std::to_string(v);
            return false;
        }
        mod_[v] = cur_;
        mod_at_[v] = (int32_t)i;
    }
    auto L = [&](int64_t v) { return mod_[v] == cur_ ? topo.mod_left[mod_at_[v]] : strided(topo.left, topo.stride_bytes, v); };
    auto R = [&](int64_t v) { return mod_[v] == cur_ ? topo.mod_right[mod_at_[v]] : strided(topo.right, topo.stride_bytes, v); };

    const int64_t rl = L(root), rr = R(root);
    const bool root_ok = !(rl < 0 || rl >= B_ || rr < 0 || rr >= B_ || rl == rr || rl == root || rr == root);
    if (whole && !root_ok)
        return compile_impl(topo, store, out, n_slots, err, piece_size, pieces, false);
    /* whole-tree route: "dirty" is "internal"; no dirty list, no marks */
    auto dirtyp = [&](int64_t v) { return whole ? v >= n_ : is_dirty(v); };
    if (!whole)
    {
        /* the dirty set, de-duplicated */
        dirty_.clear();
        if (topo.n_dirty < 0)
        {
            dirty_.resize((size_t)(B_ - n_));
            for (int64_t v = n_; v < B_; v++)
            {
                stamp_[v] = cur_;
                dirty_[(size_t)(v - n_)] = (int32_t)v;
            }
        }
        else
        {
            for (int64_t k = 0; k < topo.n_dirty; k++)
            {
                const int64_t v = topo.dirty[k];
                if (v < n_ || v >= B_)
                {
                    err = "schedule: dirty node " + std::to_string(v) + " is not an internal node";
                    return false;
                }
                if (stamp_[v] != cur_)
                {
                    stamp_[v] = cur_;
                    dirty_.push_back((int32_t)v);
                }
            }
        }
        /* which dirty nodes hang under another dirty node; the child links are resolved (overlay,
         * stride) once, here */
        for (int32_t v : dirty_)
        {
            const int64_t l = L(v), r = R(v);
            if (l < 0 || l >= B_ || r < 0 || r >= B_ || l == r || l == v || r == v)
            {
                err = "schedule: node " + std::to_string(v) + " has bad children " + std::to_string(l) + "," +
                      std::to_string(r);
                return false;
            }
            cl_[v] = (int32_t)l;
            cr_[v] = (int32_t)r;
            if (is_dirty(l))
                child_[l] = cur_;
            if (is_dirty(r))
                child_[r] = cur_;
        }
        if (!root_ok)
        {
            err = "schedule: root " + std::to_string(root) + " has bad children";
            return false;
        }
    }

    /* discovery order of every dirty node, from the tops down (parents before children,
     * breadth first).  From here on a dirty node is named by its POSITION in that order: the two
     * children of a node sit next to each other, so the sweeps below walk their arrays almost
     * sequentially.  A child is kept as its position (dirty) or as ~row (clean operand row). */
    size_t no = 0;
    auto discover = [&](int64_t top) -> bool {
        if (!dirtyp(top))
            return true;
        if (seen_[top] == cur_)
            return false;
        seen_[top] = cur_;
        size_t head = no;
        pos_[top] = (int32_t)no;
        korder_[no++] = (int32_t)top;
        while (head < no)
        {
            const int32_t v = korder_[head];
            int32_t kid[2];
            if (whole && head + 6 < no)
            { /* the caller's arrays are cold for a topology that was not seen before (2 x 2000 scattered reads):
               * ask for the links of a node a few places down the queue now */
                const int64_t w = korder_[head + 6];
                __builtin_prefetch(reinterpret_cast<const char *>(topo.left) + w * topo.stride_bytes);
                __builtin_prefetch(reinterpret_cast<const char *>(topo.right) + w * topo.stride_bytes);
                __builtin_prefetch(&seen_[w]);
            }
            if (whole)
            { /* links are resolved and checked on the way */
                const int64_t l = L(v), r = R(v);
                if (l < 0 || l >= B_ || r < 0 || r >= B_ || l == r || l == v || r == v)
                    return false;
                kid[0] = (int32_t)l;
                kid[1] = (int32_t)r;
            }
            else
            {
                kid[0] = cl_[v];
                kid[1] = cr_[v];
            }
            for (int k = 0; k < 2; k++)
            {
                const int32_t c = kid[k];
                if (dirtyp(c))
                {
                    if (seen_[c] == cur_)
                        return false;
                    seen_[c] = cur_;
                    pos_[c] = (int32_t)no;
                    kid[k] = (int32_t)no;
                    korder_[no++] = c;
                }
                else
                    kid[k] = ~c;
            }
            k0_[head] = kid[0];
            k1_[head] = kid[1];
            head++;
        }
        return true;
    };
    if (whole)
    { /* tops in ascending node order, as the general route meets them */
        const int64_t t0 = std::min(rl, rr), t1 = std::max(rl, rr);
        if (!discover(t0) || !discover(t1) || (int64_t)no != B_ - n_)
            return compile_impl(topo, store, out, n_slots, err, piece_size, pieces, false);
    }
    else
    {
        bool ok = true;
        for (int32_t v : dirty_)
            if (child_[v] != cur_)
                ok = ok && discover(v);
        if (!ok || no != dirty_.size())
        {
            err = "schedule: dirty nodes do not form a forest (a node is reachable twice or lies on a cycle)";
            return false;
        }
    }
    /* bottom-up over the dirty forest: subtree sizes and stack need of the uncut program.
     * size: nodes of the node's own piece at and below it; need: stack slots in use at the peak while
     * that part is evaluated. */
    auto size_and_need = [&](size_t k) {
        const int32_t c[2] = {k0_[k], k1_[k]};
        int32_t sz = 1, nd[2] = {-1, -1};
        for (int j = 0; j < 2; j++)
            if (c[j] >= 0 && !kcut_[c[j]])
            {
                sz += ksize_[c[j]];
                nd[j] = kneed_[c[j]];
            }
        if (nd[0] >= 0 && nd[1] >= 0)
            kneed_[k] = std::max(std::max(nd[0], nd[1]), std::min(nd[0], nd[1]) + 1);
        else
            kneed_[k] = std::max(0, std::max(nd[0], nd[1]));
        ksize_[k] = sz;
    };
    for (size_t k = no; k-- > 0;)
    {
        kcut_[k] = 0;
        klevel_[k] = 0;
        size_and_need(k);
    }

    /* Pieces.  Level 0: every MAXIMAL subtree of at most piece_size dirty nodes, packed into bins of
     * about piece_size instructions — a piece is a forest, so the bins are balanced whatever the
     * shape of the tree.  What is left (the "crown": nodes with more than piece_size nodes below
     * them, about 2 d / piece_size of them in a random tree) is cut the same way with a limit of
     * sqrt(2 W) while more than kCrownTop nodes are left, and the rest runs as the top piece with
     * the root step.  Pieces of level L read, as ordinary HBM rows, rows stored by levels < L.
     * The dependent chain of a launch is therefore one bin per level plus a small top piece,
     * instead of one piece_size..2*piece_size subtree per level of a greedy bottom-up cut. */
    const uint32_t node_ctl = store ? LVB_CTL_STORE : 0u;
    int peak = 0;
    size_t cursor = out.size();
    int top_level = 0;
    tops_.clear();
    if (piece_size > 0 && no > (size_t)piece_size) /* what fits one piece stays one piece */
    {
        constexpr size_t kCrownTop = 32; /* a crown this small is not worth another dependency hop */
        constexpr int kMaxLevels = 6;
        /* positions that have no dirty parent: the root's children and, in the general route,
         * detached dirty subtrees */
        if (dirtyp(rl))
            tops_.push_back(pos_[rl]);
        if (dirtyp(rr))
            tops_.push_back(pos_[rr]);
        for (size_t i = 0; !whole && i < dirty_.size(); i++)
            if (const int32_t v = dirty_[i]; child_[v] != cur_ && v != rl && v != rr)
                tops_.push_back(pos_[v]);
        crown_.clear();
        int64_t S = piece_size;
        bool first = true;
        while (top_level < kMaxLevels)
        {
            items_.clear();
            next_.clear();
            auto visit = [&](size_t k) {
                if (ksize_[k] <= S)
                    return;
                next_.push_back((int32_t)k);
                const int32_t c[2] = {k0_[k], k1_[k]};
                for (int j = 0; j < 2; j++)
                    if (c[j] >= 0 && !kcut_[c[j]] && ksize_[c[j]] <= S)
                    {
                        kcut_[c[j]] = 1;
                        items_.push_back(c[j]);
                    }
            };
            if (first)
                for (size_t k = 0; k < no; k++)
                    visit(k);
            else
                for (int32_t k : crown_)
                    visit((size_t)k);
            for (int32_t t : tops_)
                if (!kcut_[t] && ksize_[t] <= S)
                {
                    kcut_[t] = 1;
                    items_.push_back(t);
                }
            first = false;
            crown_.swap(next_);
            if (!items_.empty())
            {
                /* pack the subtrees of this level into bins, largest first.  Level 0 is handed out piece by piece
                 * (all column blocks of a bin, then the next bin), so its bins get SMALLER towards the end —
                 * full size for the first 3/4 of the work, then half, a quarter, an eighth of it (guided
                 * self-scheduling): the launch runs dry on small items instead of on whole bins.  Upper levels:
                 * equal bins, each subtree into the emptiest one. */
                int64_t total = 0;
                for (int32_t k : items_)
                    total += ksize_[k];
                std::sort(items_.begin(), items_.end(), [&](int32_t a, int32_t b) {
                    return ksize_[a] != ksize_[b] ? ksize_[a] > ksize_[b] : a < b;
                });
                size_t n_bins = 0;
                if (top_level == 0 && taper_ && S >= 48)
                {
                    int64_t filled = 0, load = 0, cap = S;
                    for (int32_t k : items_)
                    {
                        if (load > 0 && load + ksize_[k] > cap + cap / 4)
                        { /* next bin */
                            n_bins++;
                            load = 0;
                            const int64_t left = total - filled;
                            cap = left * 4 > total ? S : left * 8 > total ? S / 2 : left * 16 > total ? S / 4 : (taper_ < 2 || left * 32 > total) ? S / 8 : S / 16;
                            cap = std::max<int64_t>(cap, 4);
                        }
                        klevel_[k] = (int32_t)n_bins; /* for now: the bin */
                        load += ksize_[k];
                        filled += ksize_[k];
                    }
                    n_bins++;
                }
                else
                {
                    n_bins = (size_t)std::max<int64_t>(1, (total + S - 1) / S);
                    bins_.assign(n_bins, std::pair<int64_t, int32_t>(0, 0));
                    for (size_t b = 0; b < n_bins; b++)
                        bins_[b].second = (int32_t)b;
                    auto heavier = [](const std::pair<int64_t, int32_t> &a, const std::pair<int64_t, int32_t> &b) {
                        return a.first != b.first ? a.first > b.first : a.second > b.second;
                    };
                    for (int32_t k : items_)
                    { /* bins_ is a min-heap on the load */
                        std::pop_heap(bins_.begin(), bins_.end(), heavier);
                        klevel_[k] = bins_.back().second; /* for now: the bin */
                        bins_.back().first += ksize_[k];
                        std::push_heap(bins_.begin(), bins_.end(), heavier);
                    }
                }
                /* lay the bins out one after the other */
                bin_at_.assign(n_bins + 1, 0);
                for (int32_t k : items_)
                    bin_at_[(size_t)klevel_[k] + 1] += ksize_[k];
                for (size_t b = 0; b < n_bins; b++)
                {
                    if (pieces && bin_at_[b + 1] > 0)
                        pieces->push_back(Piece{(int32_t)(cursor + (size_t)bin_at_[b]), bin_at_[b + 1], top_level});
                    bin_at_[b + 1] += bin_at_[b];
                }
                for (int32_t k : items_)
                {
                    const size_t b = (size_t)klevel_[k];
                    kstart_[k] = (int32_t)(cursor + (size_t)bin_at_[b]);
                    bin_at_[b] += ksize_[k];
                    kdepth_[k] = 0;
                    kpush_[k] = 0;
                    klevel_[k] = top_level;
                }
                cursor += (size_t)total;
                top_level++;
                /* sizes and stack need of what is left, without the subtrees just cut */
                for (size_t i = crown_.size(); i-- > 0;)
                    size_and_need((size_t)crown_[i]);
            }
            if (crown_.size() <= kCrownTop || items_.empty())
                break;
            int64_t s2 = 6;
            while (s2 * s2 < 2 * (int64_t)crown_.size())
                s2++;
            S = s2;
        }
    }
    const size_t last_begin = cursor;
    /* dirty subtrees that do not hang under the root's children (only for dirty sets that are
     * not closed towards the root; LVB's operators never produce them, getplen would still
     * recompute them) */
    for (size_t i = 0; !whole && i < dirty_.size(); i++)
        if (const int32_t v = dirty_[i]; child_[v] != cur_ && v != rl && v != rr)
        {
            const int32_t k = pos_[v];
            if (!kcut_[k])
            {
                kstart_[k] = (int32_t)cursor;
                kdepth_[k] = 0;
                kpush_[k] = 0;
                cursor += (size_t)ksize_[k];
            }
        }
    /* root step, TreeEvaluation.c:238-264: fitch(root.left, root.right), never stored, then the
     * count against the root's own row */
    const int32_t krl = dirtyp(rl) ? pos_[rl] : ~(int32_t)rl, krr = dirtyp(rr) ? pos_[rr] : ~(int32_t)rr;
    auto inside = [&](int32_t c) { return c >= 0 && !kcut_[c]; }; /* evaluated inside the piece being emitted */
    auto row_of = [&](int32_t c) { return c >= 0 ? korder_[c] : ~c; };
    const size_t root_start = cursor;
    cursor += 1 + (size_t)(inside(krl) ? ksize_[krl] : 0) + (size_t)(inside(krr) ? ksize_[krr] : 0);
    const size_t root_pos = cursor - 1;
    cursor++; /* the count-only instruction */
    out.resize(cursor);
    FitchInstr *const o = out.data();

    /* one node: place the dirty children, write the node's own instruction at `at` */
    auto place = [&](int32_t l, int32_t r, int32_t node, uint32_t ctl, size_t start, int depth, size_t at, uint32_t push) {
        const bool dl = inside(l), dr = inside(r);
        const uint32_t pushbits = push << LVB_CTL_PUSH_SHIFT;
        if (dl && dr)
        {
            const bool left_first = kneed_[l] >= kneed_[r];
            const int32_t first = left_first ? l : r, second = left_first ? r : l;
            kstart_[first] = (int32_t)start;
            kdepth_[first] = depth;
            kpush_[first] = depth + 1; /* its result waits in slot `depth` */
            kstart_[second] = (int32_t)(start + (size_t)ksize_[first]);
            kdepth_[second] = depth + 1;
            kpush_[second] = 0;
            peak = std::max(peak, depth + 1);
            o[at] = FitchInstr{LVB_OPER_SLOT(depth), LVB_OPER_PREV, node, ctl | pushbits};
        }
        else if (dl || dr)
        {
            const int32_t c = dl ? l : r, other = dl ? r : l;
            kstart_[c] = (int32_t)start;
            kdepth_[c] = depth;
            kpush_[c] = 0;
            o[at] = FitchInstr{LVB_OPER_PREV, row_of(other), node, ctl | pushbits};
        }
        else
            o[at] = FitchInstr{row_of(l), row_of(r), node, ctl | pushbits};
    };
    place(krl, krr, -1, 0u, root_start, 0, root_pos, 0u);
    o[root_pos + 1] = FitchInstr{LVB_OPER_PREV, (int32_t)root, -1, LVB_CTL_COUNT_ONLY};
    for (size_t k = 0; k < no; k++) /* parents before children: start/depth/push of k are set by now */
        place(k0_[k], k1_[k], korder_[k], node_ctl, (size_t)kstart_[k], kdepth_[k], (size_t)kstart_[k] + (size_t)ksize_[k] - 1,
              (uint32_t)kpush_[k]);
    if (pieces && cursor > last_begin)
        pieces->push_back(Piece{(int32_t)last_begin, (int32_t)(cursor - last_begin), top_level});

    if (peak > LVB_MAX_SLOTS)
    {
        err = "schedule: program needs " + std::to_string(peak) + " stack slots (max " +
              std::to_string(LVB_MAX_SLOTS) + ")";
        return false;
    }
    n_slots = peak;
    return true;
}

bool ScheduleCompiler::compile_up_pass(const lvb_gpu_topology &topo, std::vector<FitchInstr> &out, int &n_slots,
                                       std::vector<int32_t> &parent, std::vector<int32_t> &left,
                                       std::vector<int32_t> &right, std::string &err, int piece_size,
                                       std::vector<Piece> *pieces)
{
    if (piece_size > 0 && !pieces)
    {
        err = "up-pass: pieces wanted but no place to put them";
        return false;
    }
    const int64_t root = topo.root;
    if (n_ < 4 || root < 0 || root >= n_)
    {
        err = "up-pass: bad root or fewer than 4 leaves";
        return false;
    }
    /* materialise the child links (overlay applied) and the parents */
    left.assign((size_t)B_, -1);
    right.assign((size_t)B_, -1);
    parent.assign((size_t)B_, -1);
    for (int64_t v = 0; v < B_; v++)
    {
        if (v < n_ && v != root)
            continue;
        left[v] = (int32_t)strided(topo.left, topo.stride_bytes, v);
        right[v] = (int32_t)strided(topo.right, topo.stride_bytes, v);
    }
    for (int64_t i = 0; i < topo.n_mod; i++)
    {
        const int64_t v = topo.mod_node[i];
        if (v < 0 || v >= B_ || (v < n_ && v != root))
        {
            err = "up-pass: overlay names node " + std::to_string(v);
            return false;
        }
        left[v] = (int32_t)topo.mod_left[i];
        right[v] = (int32_t)topo.mod_right[i];
    }
    for (int64_t v = 0; v < B_; v++)
    {
        if (v < n_ && v != root)
            continue;
        const int64_t c[2] = {left[v], right[v]};
        for (int k = 0; k < 2; k++)
        {
            if (c[k] < 0 || c[k] >= B_ || c[k] == root || c[k] == v || parent[c[k]] != -1)
            {
                err = "up-pass: node " + std::to_string(v) + " has bad children";
                return false;
            }
            parent[c[k]] = (int32_t)v;
        }
    }
    /* pre-order of the internal nodes from the root (also proves that the links form one tree) */
    order_.clear();
    for (int k = 0; k < 2; k++)
    {
        const int32_t c = k ? right[root] : left[root];
        if (c >= n_)
            order_.push_back(c);
    }
    for (size_t head = 0; head < order_.size(); head++)
    {
        const int32_t v = order_[head];
        if (left[v] >= n_)
            order_.push_back(left[v]);
        if (right[v] >= n_)
            order_.push_back(right[v]);
    }
    if ((int64_t)order_.size() != B_ - n_)
    {
        err = "up-pass: the links do not form one tree";
        return false;
    }
    /* stack slots the pre-order walk below v needs: a node with two internal children parks
     * up(v) while the first child's subtree is walked; the child that needs fewer goes first */
    for (size_t k = order_.size(); k-- > 0;)
    {
        const int32_t v = order_[k];
        const int32_t a = left[v], b = right[v];
        const int na = a >= n_ ? need_[a] : -1, nb = b >= n_ ? need_[b] : -1;
        if (na >= 0 && nb >= 0)
            need_[v] = std::max(std::max(na, nb), std::min(na, nb) + 1);
        else
            need_[v] = std::max(0, std::max(na, nb));
        size_[v] = 1 + (a >= n_ ? size_[a] : 0) + (b >= n_ ? size_[b] : 0); /* internal nodes at and below v */
    }
    /* pieces: the walk stops in front of every maximal subtree of at most piece_size internal
     * nodes (tiny ones stay inline); up(c) of such a subtree's top c is still computed and stored
     * by the piece above, and the subtree becomes a program of its own (level 1) that starts by
     * reading that row back.  All level-1 programs are independent of each other. */
    std::vector<uint8_t> is_cut;
    std::vector<int32_t> cut_roots;
    if (piece_size > 0)
    {
        is_cut.assign((size_t)B_, 0);
        const int min_size = std::max(2, piece_size / 4);
        for (int32_t v : order_)
        {
            const int32_t pv = parent[v];
            const bool parent_big = pv == root || size_[pv] > piece_size;
            if (parent_big && size_[v] <= piece_size && size_[v] >= min_size)
            {
                is_cut[(size_t)v] = 1;
                cut_roots.push_back(v);
            }
        }
    }
    auto enter = [&](int32_t c, bool top_piece) { return !(top_piece && piece_size > 0 && is_cut[(size_t)c]); };
    int depth = 0, peak = 0;
    auto walk = [&](int32_t top, bool top_piece) { /* the instruction that produced up(top) was emitted last */
        frames_.clear();
        frames_.push_back(Frame{top, -1, -1, 0});
        while (!frames_.empty())
        {
            Frame &f = frames_.back();
            const int32_t v = f.v, a = left[v], b = right[v];
            if (f.stage == 0)
            {
                const bool ia = a >= n_, ib = b >= n_;
                if (ia && ib)
                {
                    const bool a_first = need_[a] <= need_[b];
                    f.first = a_first ? a : b;
                    f.second = a_first ? b : a;
                    f.stage = 1;
                    out.back().ctl |= (uint32_t)(depth + 1) << LVB_CTL_PUSH_SHIFT;
                    depth++;
                    peak = std::max(peak, depth);
                    const int32_t c = f.first, o = f.second;
                    out.push_back(FitchInstr{LVB_OPER_PREV, o, c, LVB_CTL_STORE});
                    if (enter(c, top_piece))
                        frames_.push_back(Frame{c, -1, -1, 0});
                }
                else if (ia || ib)
                {
                    const int32_t c = ia ? a : b, o = ia ? b : a;
                    out.push_back(FitchInstr{LVB_OPER_PREV, o, c, LVB_CTL_STORE});
                    if (enter(c, top_piece))
                        f = Frame{c, -1, -1, 0}; /* tail call */
                    else
                        frames_.pop_back();
                }
                else
                    frames_.pop_back();
            }
            else
            { /* back from the first child's subtree: up(v) is in the top slot */
                depth--;
                const int32_t c = f.second, o = f.first;
                out.push_back(FitchInstr{LVB_OPER_SLOT(depth), o, c, LVB_CTL_STORE});
                if (enter(c, top_piece))
                    f = Frame{c, -1, -1, 0};
                else
                    frames_.pop_back();
            }
        }
    };
    const size_t begin0 = out.size();
    const int32_t rc[2] = {left[root], right[root]};
    for (int k = 0; k < 2; k++)
        if (rc[k] >= n_)
        {
            out.push_back(FitchInstr{(int32_t)root, rc[1 - k], rc[k], LVB_CTL_STORE});
            if (enter(rc[k], true))
                walk(rc[k], true);
        }
    if (pieces)
        pieces->push_back(Piece{(int32_t)begin0, (int32_t)(out.size() - begin0), 0});
    /* level 1: operand B + (c - n) names the row of c in the tree the program STORES to */
    const int32_t own = (int32_t)(B_ - n_);
    for (int32_t c : cut_roots)
    {
        const size_t begin = out.size();
        out.push_back(FitchInstr{c + own, c + own, -1, 0u}); /* up(c) into the previous-result register */
        walk(c, false);
        pieces->push_back(Piece{(int32_t)begin, (int32_t)(out.size() - begin), 1});
    }
    if (peak > LVB_MAX_SLOTS)
    {
        err = "up-pass: program needs " + std::to_string(peak) + " stack slots";
        return false;
    }
    n_slots = peak;
    return true;
}

} // namespace lvbgpu
