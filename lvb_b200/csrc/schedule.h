/* schedule.h — host-side compiler from (topology, dirty set) to a Fitch program.
 *
 * Replaces, on the host, the dependency sweeps of getplen
 * (/root/reference/src/TreeEvaluation.c:191-236): the dirty set that LVB's proposal
 * operators leave behind (make_dirty_below, src/TreeOperations.c:88-103) is turned into one
 * post-order instruction list; see fitch_isa.h.
 */
#ifndef LVB_B200_SCHEDULE_H
#define LVB_B200_SCHEDULE_H

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/lvb_gpu.h"
#include "fitch_isa.h"

namespace lvbgpu
{

class ScheduleCompiler
{
  public:
    /* n = number of leaves; reusable across calls (O(d) work per compile, no O(n) clearing) */
    explicit ScheduleCompiler(int64_t n);

    /* Appends the program for `topo` to `out`.  store: set LVB_CTL_STORE on every node
     * instruction.  Returns false and sets err on a malformed topology.  n_slots receives the
     * number of stack slots the program uses. */
    bool compile(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                 std::string &err);

    int64_t n() const { return n_; }

  private:
    struct Frame
    {
        int32_t v;
        int32_t first, second; /* dirty children in evaluation order, -1 if none */
        int32_t stage;
    };
    bool is_dirty(int64_t v) const { return v >= n_ && v < B_ && stamp_[v] == cur_; }
    int64_t n_, B_;
    uint32_t cur_;
    std::vector<uint32_t> stamp_;   /* dirty marks */
    std::vector<uint32_t> child_;   /* "is a child of a dirty node" marks */
    std::vector<uint32_t> seen_;    /* visited marks (cycle / duplicate guard) */
    std::vector<int32_t> need_;     /* stack need per dirty node */
    std::vector<int32_t> order_;    /* scratch: discovery order */
    std::vector<int32_t> dirty_;    /* scratch: de-duplicated dirty list */
    std::vector<Frame> frames_;
};

} // namespace lvbgpu
#endif
