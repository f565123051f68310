/* fitch_isa.h — the instruction stream the Fitch kernels execute.
 *
 * A tree evaluation (LVB's getplen, /root/reference/src/TreeEvaluation.c:182-265) is
 * compiled on the host into a straight-line post-order "program": one instruction per
 * dirty internal node plus two for the root step (:238-264).  Every thread of the grid
 * owns one 16-byte column (32 sites) of the state matrix and runs the whole program on
 * it, so the reference's repeated sweeps of the to-do list (:204-236) become a single
 * pass with no inter-thread dependency.
 *
 * Shared between host (schedule.cpp) and device (fitch_kernels.cu).
 */
#ifndef LVB_B200_FITCH_ISA_H
#define LVB_B200_FITCH_ISA_H

#include <stdint.h>

/* Operand encoding of FitchInstr::a / ::b
 *   v >= 0        state row of LVB node v (leaf: the alignment row; internal: the row the
 *                 program's tree currently holds for v) — read from HBM.  Store programs that
 *                 read one tree and write another (FitchProgram::src_tree1) name a row of the
 *                 tree they write as (2n-3) + (v - n).  This is what the COMPILER emits; before a
 *                 batch goes to the device the device layer replaces every operand >= n by the
 *                 absolute row of the pool (tree, bank of the node: see lvb_gpu.cu)
 *   OPER_PREV     the result of the previous instruction (still in registers)
 *   OPER_SLOT(s)  on-chip stack slot s (shared memory), written by an earlier push
 */
#define LVB_OPER_PREV (-1)
#define LVB_OPER_SLOT(s) (-2 - (s))
#define LVB_OPER_IS_SLOT(x) ((x) <= -2)
#define LVB_OPER_SLOT_INDEX(x) (-2 - (x))

/* FitchInstr::ctl bits */
#define LVB_CTL_STORE 0x1u      /* write the result to the row of `node` in the program's tree */
#define LVB_CTL_COUNT_ONLY 0x2u /* only the change count is needed (2nd root-step application) */
#define LVB_CTL_BANK1 0x4u      /* (set by the device layer, not by the compiler) the result row is the node's row in
                                 * the SECOND bank of the program's tree */
#define LVB_CTL_PUSH_SHIFT 8    /* bits 8..15: 1 + stack slot to push the result to; 0 = no push */
#define LVB_CTL_PUSH_MASK 0xff00u

#define LVB_MAX_SLOTS 24 /* enough for > 16M leaves with deeper-child-first ordering */

typedef struct FitchInstr
{
    int32_t a;    /* operand */
    int32_t b;    /* operand */
    int32_t node; /* LVB node number this instruction computes (>= n), or -1 for the root step */
    uint32_t ctl;
} FitchInstr;

/* One program = one (candidate) tree evaluated against one device tree. */
typedef struct FitchProgram
{
    int32_t instr_begin; /* index of the first instruction in the batch's instruction buffer */
    int32_t n_instr;
    int32_t tree;        /* device tree the internal operands/stores refer to */
    int32_t flags;       /* bit 0: per-node counts wanted (store programs); pieces: their level */
    int32_t parent;      /* piece (index into the launch's program array) that reads the row this piece's
                          * top node stores, -1 if none */
    int32_t n_wait;      /* number of pieces of the same launch whose stored rows this piece reads */
    int32_t src_tree1;   /* store programs: 1 + device tree the internal-row OPERANDS are read from, when
                          * that is not `tree` (up-pass: reads the evaluated tree, stores into a scratch
                          * tree); 0 = same tree */
    int32_t pad1;
} FitchProgram;

#endif
