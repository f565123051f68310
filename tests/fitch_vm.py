"""Test infrastructure: a numpy interpreter of the Fitch instruction stream
(lvb_b200/csrc/fitch_isa.h), used on the CPU to check the host schedule compiler
against the oracle without a GPU.  Never imported by the product."""
from __future__ import annotations

import numpy as np

LOW3 = np.uint64(0x7777777777777777)
TOP1 = np.uint64(0x8888888888888888)
PREV = -1
CTL_STORE, CTL_COUNT_ONLY = 1, 2


def fitch_rows(x: np.ndarray, y: np.ndarray):
    """Fitch step on whole rows of packed words (TreeEvaluation.c:219-230)."""
    t = x & y
    hit = (((t & LOW3) + LOW3) | t) & TOP1
    ch = 16 * x.size - int(sum(bin(int(w)).count("1") for w in hit))
    nib = hit >> np.uint64(3)
    out = t | ((x | y) & ((nib + LOW3) ^ TOP1))
    return out, ch


def run_program(instr: np.ndarray, rows: np.ndarray, n: int, store: bool):
    """rows: B x nwords current state (modified in place when store).  Returns
    (sum of counts, {node: count}, root-step count)."""
    acc = None
    slots = {}
    total, per_node, rootc = 0, {}, 0

    def operand(v):
        if v >= 0:
            return rows[v]
        if v == PREV:
            return acc
        return slots[-2 - v]

    for a, b, node, ctl in instr.tolist():
        out, ch = fitch_rows(operand(a), operand(b))
        total += ch
        if node >= 0:
            assert node >= n and node not in per_node
            per_node[node] = ch
        else:
            rootc += ch
        if ctl & CTL_STORE:
            assert store and node >= n
            rows[node] = out
        push = (ctl >> 8) & 0xFF
        if push:
            slots[push - 1] = out
        acc = out
    return total, per_node, rootc
