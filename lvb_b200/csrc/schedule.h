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
#include <utility>
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

    /* One independently schedulable part of a program.  Pieces of level L read, as ordinary
     * HBM rows, the rows stored by pieces of levels < L, so all pieces of one level can run
     * concurrently and levels run as consecutive launches. */
    struct Piece
    {
        int32_t instr_begin; /* index into `out` */
        int32_t n_instr;
        int32_t level;
    };

    /* Appends the program for `topo` to `out`.  store: set LVB_CTL_STORE on every node
     * instruction.  Returns false and sets err on a malformed topology.  n_slots receives the
     * number of stack slots the program uses.
     * piece_size > 0 (store programs only): cut the dirty forest into pieces; `pieces` receives
     * them in level order (the last one holds the root step).  Level 0 holds every maximal subtree of
     * at most piece_size nodes, packed into bins (a piece is a FOREST of such subtrees) of about
     * piece_size instructions; the nodes above them are cut again with a smaller limit while there
     * are many, and the rest is the top piece.  piece_size == 0: one piece. */
    bool compile(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                 std::string &err, int piece_size = 0, std::vector<Piece> *pieces = nullptr);

    int64_t n() const { return n_; }

    /* level-0 bins: equal (0) or smaller towards the end, down to 1/8 (1) or 1/16 (2) of the full size (A/B switch) */
    void set_taper(int depth) { taper_ = depth; }

    /* The up-pass of a whole tree: one store instruction per internal node v computing
     * up(v) = the Fitch set of everything that is NOT below v (the tree seen from v's parent
     * branch): up(v) = fitch(up(parent), low(sister)), and fitch(row(root), low(sister)) for
     * the two children of the root.  Operand rows >= n are rows of the evaluated tree (low
     * sets), results are stored under the same node numbers in another device tree.  The
     * traversal is pre-order with up(parent) in the previous-result register or in a stack
     * slot, so no instruction reads a row the program itself stores.
     * parent receives the parent of every node (root: -1).  Used by the exhaustive NNI
     * neighbourhood (src/Solve.c:64-142).
     * piece_size > 0: the program is cut into a top piece (level 0) and independent subtrees of at
     * most piece_size internal nodes (level 1, run after level 0); an operand B + (v - n) then names
     * the row of v in the tree the program stores to (up(v), stored by the top piece). */
    bool compile_up_pass(const lvb_gpu_topology &topo, std::vector<FitchInstr> &out, int &n_slots,
                         std::vector<int32_t> &parent, std::vector<int32_t> &left, std::vector<int32_t> &right,
                         std::string &err, int piece_size = 0, std::vector<Piece> *pieces = nullptr);

  private:
    bool compile_impl(const lvb_gpu_topology &topo, bool store, std::vector<FitchInstr> &out, int &n_slots,
                      std::string &err, int piece_size, std::vector<Piece> *pieces, bool whole);
    struct Frame
    {
        int32_t v;
        int32_t first, second; /* dirty children in evaluation order, -1 if none */
        int32_t stage;
    };
    bool is_dirty(int64_t v) const { return v >= n_ && v < B_ && stamp_[v] == cur_; }
    int64_t n_, B_;
    int taper_ = 1; /* 0: equal bins; 1: bins down to an eighth; 2: down to a sixteenth (for the last 1/32 of the work) */
    uint32_t cur_;
    /* indexed by node */
    std::vector<uint32_t> stamp_;   /* dirty marks */
    std::vector<uint32_t> child_;   /* "is a child of a dirty node" marks */
    std::vector<uint32_t> seen_;    /* visited marks (cycle / duplicate guard) */
    std::vector<uint32_t> mod_;     /* "child links come from the overlay" marks */
    std::vector<int32_t> mod_at_;   /* index into the overlay arrays */
    std::vector<int32_t> need_;     /* up-pass: stack need per node */
    std::vector<int32_t> size_;     /* up-pass: internal nodes at and below a node */
    std::vector<int32_t> cl_, cr_;  /* resolved child links of the dirty nodes */
    std::vector<int32_t> pos_;      /* position of a dirty node in the discovery order */
    /* indexed by position in the discovery order (compile) */
    std::vector<int32_t> korder_;   /* the node at a position */
    std::vector<int32_t> k0_, k1_;  /* children: position of a dirty child, or ~row of a clean one */
    std::vector<int32_t> ksize_;    /* nodes of the piece hanging under the node */
    std::vector<int32_t> kneed_;    /* stack need */
    std::vector<int32_t> klevel_;   /* level of the piece the node belongs to */
    std::vector<int32_t> kstart_;   /* emission: index of the first instruction of the node's subtree */
    std::vector<int32_t> kdepth_;   /* emission: stack slots in use while the subtree is evaluated */
    std::vector<int32_t> kpush_;    /* emission: 1 + slot the node's result is parked in, 0 = none */
    std::vector<uint8_t> kcut_;     /* the node is the root of a piece of its own */
    std::vector<int32_t> order_;    /* scratch (up-pass): discovery order */
    std::vector<int32_t> dirty_;    /* scratch: de-duplicated dirty list */
    /* scratch of the piece cut: positions without a dirty parent, the crown (what is not cut yet),
     * the subtrees cut at the current level and their bins (load, bin) */
    std::vector<int32_t> tops_, crown_, next_, items_, bin_at_;
    std::vector<std::pair<int64_t, int32_t>> bins_;
    std::vector<Frame> frames_;
};

} // namespace lvbgpu
#endif
