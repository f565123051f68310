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
      need_(stamp_.size(), 0)
{
}

bool ScheduleCompiler::compile(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                               std::string &err)
{
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
        cur_ = 1;
    }
    const int64_t root = topo.root;
    if (root < 0 || root >= n_)
    {
        err = "schedule: root " + std::to_string(root) + " is not a leaf";
        return false;
    }
    auto L = [&](int64_t v) { return strided(topo.left, topo.stride_bytes, v); };
    auto R = [&](int64_t v) { return strided(topo.right, topo.stride_bytes, v); };

    /* the dirty set, de-duplicated */
    dirty_.clear();
    if (topo.n_dirty < 0)
    {
        for (int64_t v = n_; v < B_; v++)
        {
            stamp_[v] = cur_;
            dirty_.push_back((int32_t)v);
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
    /* which dirty nodes hang under another dirty node */
    for (int32_t v : dirty_)
    {
        const int64_t l = L(v), r = R(v);
        if (l < 0 || l >= B_ || r < 0 || r >= B_ || l == r || l == v || r == v)
        {
            err = "schedule: node " + std::to_string(v) + " has bad children " + std::to_string(l) + "," +
                  std::to_string(r);
            return false;
        }
        if (is_dirty(l))
            child_[l] = cur_;
        if (is_dirty(r))
            child_[r] = cur_;
    }
    const int64_t rl = L(root), rr = R(root);
    if (rl < 0 || rl >= B_ || rr < 0 || rr >= B_ || rl == rr || rl == root || rr == root)
    {
        err = "schedule: root " + std::to_string(root) + " has bad children";
        return false;
    }

    /* discovery order of every dirty node, from the tops down (parents before children) */
    order_.clear();
    auto discover = [&](int64_t top) -> bool {
        if (!is_dirty(top))
            return true;
        if (seen_[top] == cur_)
            return false;
        seen_[top] = cur_;
        size_t head = order_.size();
        order_.push_back((int32_t)top);
        while (head < order_.size())
        {
            const int32_t v = order_[head++];
            const int64_t c[2] = {L(v), R(v)};
            for (int k = 0; k < 2; k++)
            {
                if (is_dirty(c[k]))
                {
                    if (seen_[c[k]] == cur_)
                        return false;
                    seen_[c[k]] = cur_;
                    order_.push_back((int32_t)c[k]);
                }
            }
        }
        return true;
    };
    bool ok = true;
    for (int32_t v : dirty_)
        if (child_[v] != cur_)
            ok = ok && discover(v);
    if (!ok || order_.size() != dirty_.size())
    {
        err = "schedule: dirty nodes do not form a forest (a node is reachable twice or lies on a cycle)";
        return false;
    }
    /* stack need: slots in use at the peak while a node's dirty subtree is evaluated */
    for (size_t k = order_.size(); k-- > 0;)
    {
        const int32_t v = order_[k];
        const int64_t l = L(v), r = R(v);
        const bool dl = is_dirty(l), dr = is_dirty(r);
        if (dl && dr)
        {
            const int32_t big = std::max(need_[l], need_[r]), small = std::min(need_[l], need_[r]);
            need_[v] = std::max(big, small + 1);
        }
        else if (dl)
            need_[v] = need_[l];
        else if (dr)
            need_[v] = need_[r];
        else
            need_[v] = 0;
    }

    /* emission */
    const uint32_t node_ctl = store ? LVB_CTL_STORE : 0u;
    int depth = 0, peak = 0;
    auto emit_from = [&](int32_t top) {
        /* v == -1 is the root step's first application: fitch(root.left, root.right), never stored */
        frames_.clear();
        frames_.push_back(Frame{top, -1, -1, 0});
        while (!frames_.empty())
        {
            Frame &f = frames_.back();
            const int32_t v = f.v;
            const int64_t l = (v < 0) ? rl : L(v), r = (v < 0) ? rr : R(v);
            const uint32_t ctl = (v < 0) ? 0u : node_ctl;
            switch (f.stage)
            {
            case 0: {
                const bool dl = is_dirty(l), dr = is_dirty(r);
                if (dl && dr)
                {
                    const bool left_first = need_[l] >= need_[r];
                    f.first = (int32_t)(left_first ? l : r);
                    f.second = (int32_t)(left_first ? r : l);
                    f.stage = 1;
                    const int32_t c = f.first;
                    frames_.push_back(Frame{c, -1, -1, 0});
                }
                else if (dl || dr)
                {
                    f.first = (int32_t)(dl ? l : r);
                    f.stage = 3;
                    const int32_t c = f.first;
                    frames_.push_back(Frame{c, -1, -1, 0});
                }
                else
                {
                    out.push_back(FitchInstr{(int32_t)l, (int32_t)r, v, ctl});
                    frames_.pop_back();
                }
                break;
            }
            case 1: { /* first child is in PREV: park it in a slot, then do the second */
                out.back().ctl |= (uint32_t)(depth + 1) << LVB_CTL_PUSH_SHIFT;
                depth++;
                peak = std::max(peak, depth);
                f.stage = 2;
                const int32_t c = f.second;
                frames_.push_back(Frame{c, -1, -1, 0});
                break;
            }
            case 2: /* second child is in PREV, first in the top slot */
                depth--;
                out.push_back(FitchInstr{LVB_OPER_SLOT(depth), LVB_OPER_PREV, v, ctl});
                frames_.pop_back();
                break;
            default: { /* 3: the one dirty child is in PREV, the other child is a clean row */
                const int64_t other = (f.first == l) ? r : l;
                out.push_back(FitchInstr{LVB_OPER_PREV, (int32_t)other, v, ctl});
                frames_.pop_back();
                break;
            }
            }
        }
    };
    /* dirty subtrees that do not hang under the root's children (only for dirty sets that are
     * not closed towards the root; LVB's operators never produce them, getplen would still
     * recompute them) */
    for (int32_t v : dirty_)
        if (child_[v] != cur_ && v != rl && v != rr)
            emit_from(v);
    /* root step, TreeEvaluation.c:238-264 */
    emit_from(-1);
    out.push_back(FitchInstr{LVB_OPER_PREV, (int32_t)root, -1, LVB_CTL_COUNT_ONLY});

    if (peak > LVB_MAX_SLOTS)
    {
        err = "schedule: program needs " + std::to_string(peak) + " stack slots (max " +
              std::to_string(LVB_MAX_SLOTS) + ")";
        return false;
    }
    n_slots = peak;
    return true;
}

} // namespace lvbgpu
