/* fitch_kernels.cu — hand-written sm_100a kernels of the Fitch tree-length path.
 *
 * Arithmetic restated from the reference's 64-bit SWAR word step
 * (/root/reference/src/TreeEvaluation.c:219-230, masks src/LVB.h:88-89) for 32-bit lanes
 * (a 4-bit state set never straddles a lane):
 *     t   = x & y                                 intersection
 *     e8  = ~(((t & 0x7777..) + 0x7777..) | t) & 0x8888..   bit 3 of a nibble set <=> intersection EMPTY
 *     ch  = popc(e8)                              one change per empty intersection
 *     out = t | ((x | y) & 0xF-where-empty)       union where empty, else intersection
 * The reference counts 16 - popc(non-empty flags); counting the empty flags directly is the
 * same number.  Work is bitwise (LOP3/IADD/SHF/IMAD/POPC): tensor cores do not apply.
 *
 * Parallelisation: thread = one 16-byte column (32 sites) of every row; CTA = one program
 * (candidate tree) x one block of columns.  The program (fitch_isa.h) is staged in shared
 * memory in chunks; results needed later sit in registers (previous result) or in a
 * per-thread shared-memory stack, so every HBM load is of a row that is read-only during
 * the launch and every produced row is written at most once.
 */
#include "fitch_kernels.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>

namespace lvbgpu
{

#define LVB_CHUNK 128 /* instructions staged in shared memory at a time */
#define LVB_MAX_DEVICES 64

/* An instruction as the inner loop wants it (decoded once per CTA while the chunk is staged):
 *   x = operand a: pool ROW number (tree offset applied); ignored if D_APREV (a = previous result)
 *   y = operand b: pool row, or if D_BSLOT the byte offset of the stack slot in the per-thread stack
 *   z = pool row the result is stored to (D_STORE)
 *   w = D_* flags in bits 0..3 | byte offset of the stack slot to push the result to (D_PUSH)
 * Fitch's step is symmetric in its operands, so the host's (slot, prev) becomes (prev, slot). */
#define D_STORE 0x1u
#define D_APREV 0x2u
#define D_BSLOT 0x4u
#define D_PUSH 0x8u

/* The inner loop is written with explicit PTX for its memory operations: nvcc otherwise
 * re-derives the shared-memory window base and the operand initialisers on every trip. */
__device__ __forceinline__ unsigned long long row_addr(unsigned row, unsigned stride_b, unsigned long long colbase)
{
    unsigned long long d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(row), "r"(stride_b), "l"(colbase));
    return d;
}
/* predicated 16-byte load through the non-coherent path; v keeps its old value when !pred.
 * Rows read as operands are never written in the same launch. */
template <bool NC>
__device__ __forceinline__ void ldg_if(uint4 &v, unsigned long long ptr, unsigned pred)
{
    if (NC)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                     "@p ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                     : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)
                     : "l"(ptr), "r"(pred));
    else /* store mode: a row may have been written by another CTA of this launch (dependent
          * pieces), so it is read at L2, never through the non-coherent L1 path */
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                     "@p ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                     : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)
                     : "l"(ptr), "r"(pred)
                     : "memory");
}
__device__ __forceinline__ void stg_if(unsigned long long ptr, const uint4 &v, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p st.global.v4.u32 [%4], {%0,%1,%2,%3};\n\t}"
                 :
                 : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(ptr), "r"(pred)
                 : "memory");
}
__device__ __forceinline__ uint4 lds128(unsigned addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds128_if(uint4 &v, unsigned addr, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)
                 : "r"(addr), "r"(pred));
}
__device__ __forceinline__ void sts128_if(unsigned addr, const uint4 &v, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p st.shared.v4.u32 [%4], {%0,%1,%2,%3};\n\t}"
                 :
                 : "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(addr), "r"(pred)
                 : "memory");
}
__device__ __forceinline__ void sts16_if(unsigned addr, unsigned v, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b16 h;\n\tsetp.ne.b32 p, %2, 0;\n\tcvt.u16.u32 h, %0;\n\t"
                 "@p st.shared.u16 [%1], h;\n\t}"
                 :
                 : "r"(v), "r"(addr), "r"(pred)
                 : "memory");
}
__device__ __forceinline__ void sts32_if(unsigned addr, unsigned v, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t"
                 "@p st.shared.u32 [%1], %0;\n\t}"
                 :
                 : "r"(v), "r"(addr), "r"(pred)
                 : "memory");
}

/* one 32-bit lane (8 sites): e8 has bit 3 of a nibble set <=> the intersection is EMPTY */
__device__ __forceinline__ unsigned fitch_lane(unsigned x, unsigned y, unsigned &out)
{
    const unsigned t = x & y;
    const unsigned a = (t & 0x77777777u) + 0x77777777u;
    const unsigned e8 = ~(a | t) & 0x88888888u;
    const unsigned s = e8 >> 3; /* 1 per site whose intersection is empty */
    out = t | ((x | y) & (s * 15u));
    return s;
}

/* returns the flag word of this 32-site column: one bit per site whose intersection is empty */
__device__ __forceinline__ unsigned fitch_step(const uint4 &x, const uint4 &y, uint4 &out)
{
    const unsigned s0 = fitch_lane(x.x, y.x, out.x);
    const unsigned s1 = fitch_lane(x.y, y.y, out.y);
    const unsigned s2 = fitch_lane(x.z, y.z, out.z);
    const unsigned s3 = fitch_lane(x.w, y.w, out.w);
    /* the four flag words occupy bit 0 of each nibble: interleave them, count once */
    return s0 + 2u * s1 + 4u * s2 + 8u * s3;
}

/* store mode, after the counts of a program are complete: one CTA folds them into chg / sum_chg
 * and publishes the length (its own launch, or the last column block of a single-level batch) */
__device__ __forceinline__ void fitch_commit_finalize_body(const FitchInstr *instr, const FitchProgram *progs,
                                                           int32_t *cnts, int32_t *chg, long long *sum_chg, int n,
                                                           int n_internal, long long *lengths, int pi,
                                                           int32_t *kept = nullptr, long long *newsum = nullptr)
{
    __shared__ long long s_delta, s_root;
    const FitchProgram prog = progs[pi];
    int32_t *tchg = chg + (int64_t)prog.tree * n_internal;
    if (threadIdx.x == 0)
    {
        s_delta = 0;
        s_root = 0;
    }
    __syncthreads();
    long long delta = 0, rootc = 0;
    for (int i = threadIdx.x; i < prog.n_instr; i += blockDim.x)
    {
        const int node = instr[prog.instr_begin + i].node;
        const int c = __ldcg(&cnts[prog.instr_begin + i]);
        cnts[prog.instr_begin + i] = 0; /* ready for the next store batch */
        if (node >= 0)
        {
            delta += (long long)c - tchg[node - n];
            if (kept) /* peek: the tree keeps its counts, the candidate's are put aside */
                kept[prog.instr_begin + i] = c;
            else
                tchg[node - n] = c;
        }
        else
            rootc += c; /* root step: never cached (TreeEvaluation.c:238-264) */
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        delta += __shfl_xor_sync(0xffffffffu, delta, o);
        rootc += __shfl_xor_sync(0xffffffffu, rootc, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd((unsigned long long *)&s_delta, (unsigned long long)delta);
        atomicAdd((unsigned long long *)&s_root, (unsigned long long)rootc);
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const long long s = sum_chg[prog.tree] + s_delta;
        if (kept)
            newsum[pi] = s;
        else
            sum_chg[prog.tree] = s;
        lengths[pi] = s + s_root;
    }
}

/* one value of the peer-memory all-reduce (see P2PParams): this rank's partial sum `v` of entry k goes into
 * its row of every rank's slab, then the rows of all ranks that arrived in the own slab are added, in rank
 * order.  Returns the sum; on a wait that was abandoned sets *q.err and returns the partial value. */
__device__ __forceinline__ long long p2p_exchange_one(const P2PParams &q, int64_t k, long long pv)
{
    const unsigned long long v = (unsigned long long)pv;
    const unsigned long long tag = (unsigned long long)q.seq << 32;
    const unsigned long long w0 = tag | (v & 0xffffffffull), w1 = tag | (v >> 32);
    const size_t mine = 2 * (((size_t)q.slot * q.nranks + q.rank) * (size_t)q.kmax + (size_t)k);
    for (int r = 0; r < q.nranks; r++)
    {
        /* start with the next rank so that the ranks do not all write to rank 0 first */
        const int d = (q.rank + 1 + r) % q.nranks;
        volatile unsigned long long *dst = q.peer[d] + mine;
        dst[0] = w0;
        dst[1] = w1;
    }
    long long sum = 0;
    for (int r = 0; r < q.nranks; r++)
    {
        const volatile unsigned long long *e =
            q.peer[q.rank] + 2 * (((size_t)q.slot * q.nranks + r) * (size_t)q.kmax + (size_t)k);
        unsigned long long a = e[0], b = e[1];
        unsigned spins = 0;
        while ((a >> 32) != q.seq || (b >> 32) != q.seq)
        {
            __nanosleep(50);
            if ((++spins & 1023u) == 0u && (spins >= q.spin_limit || *(volatile unsigned *)q.err != 0u))
            {
                atomicCAS(q.err, 0u, 0x80000000u | (unsigned)r);
                return pv;
            }
            a = e[0];
            b = e[1];
        }
        sum += (long long)(((b & 0xffffffffull) << 32) | (a & 0xffffffffull));
    }
    return sum;
}

/* ---- dependent pieces in one launch (store mode) -------------------------------------------
 * A piece waits, per column block, for the pieces whose stored rows it reads.  CTAs are numbered
 * in the order they START (pieces are sorted by level, children before parents), so a waiting CTA
 * never occupies a slot that a CTA it waits for still needs. */
__device__ __forceinline__ unsigned piece_start_ticket(const ExecParams &p)
{
    __shared__ unsigned s_vb;
    if (threadIdx.x == 0)
    {
        const unsigned t = atomicAdd(p.ticket, 1u);
        if (t == gridDim.x - 1)
            *p.ticket = 0u; /* every ticket is taken: ready for the next launch */
        s_vb = t;
    }
    __syncthreads();
    return s_vb;
}
/* A piece of level L > 0 waits, per column block, until every piece of level L - 1 of this launch
 * has stored its rows (the levels below that were complete before the first piece of level L - 1
 * went on).  dep[L][cb] only ever counts up: the host hands every launch the value it will have
 * reached (p.dep_target), so nothing has to be reset between launches.
 * The wait is bounded: after spin_limit polls (or as soon as another CTA gave up) the CTA records
 * 1 + its piece in the mapped host word p.err and leaves; the host turns that into an error
 * instead of a hung launch.  Returns false when the wait was abandoned. */
__device__ __forceinline__ bool piece_wait_children(const ExecParams &p, int prog_id, int cb, int level)
{
    __shared__ unsigned s_abort;
    if (threadIdx.x == 0)
    {
        volatile unsigned *f = p.dep + (size_t)(level - 1) * p.n_colblocks + cb;
        const unsigned target = p.dep_target[level - 1];
        unsigned spins = 0, ab = 0;
        while ((int)(*f - target) < 0)
        {
            __nanosleep(64);
            if ((++spins & 1023u) == 0u && (spins >= p.spin_limit || *(volatile unsigned *)p.err != 0u))
            {
                ab = 1u;
                break;
            }
        }
        if (ab)
            atomicCAS(p.err, 0u, 1u + (unsigned)prog_id);
        else
            __threadfence();
        s_abort = ab;
    }
    __syncthreads();
    return s_abort == 0u;
}
/* after the CTA's last row store: one more piece of this level is complete for the column block */
__device__ __forceinline__ void piece_release_parent(const ExecParams &p, const FitchProgram &prog, int cb)
{
    if (prog.parent < 0)
        return;
    __syncthreads(); /* every thread's row stores are ordered before the release below */
    if (threadIdx.x == 0)
    {
        __threadfence();
        atomicAdd(p.dep + (size_t)prog.parent * p.n_colblocks + cb, 1u);
    }
}
/* fused finalize of a launch of dependent pieces: the column block of a piece that finishes last
 * folds that piece's counts into chg / sum_chg (what fitch_commit_finalize_body does for a whole
 * program), and the piece of a candidate that is folded last publishes the candidate's length —
 * no second launch, and the folding of the lower pieces overlaps the upper ones.
 * p.done[q] counts the column blocks of piece q, p.done[n_prog + k] the pieces of candidate k,
 * p.totals[k] collects the root-step counts of candidate k; all are left zero. */
__device__ __forceinline__ void piece_finalize(const ExecParams &p, const FitchInstr *instrs, const FitchProgram &prog,
                                               int prog_id)
{
    __shared__ unsigned s_last;
    __shared__ long long s_delta, s_root;
    const int tid = threadIdx.x;
    __syncthreads(); /* this CTA's count atomics are issued */
    if (tid == 0)
    {
        __threadfence();
        const unsigned last = atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u ? 1u : 0u;
        if (last)
            p.done[prog_id] = 0u;
        s_last = last;
        s_delta = 0;
        s_root = 0;
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    int32_t *tchg = p.chg_rw + (int64_t)prog.tree * p.n_internal;
    long long delta = 0, rootc = 0;
    for (int i = tid; i < prog.n_instr; i += blockDim.x)
    {
        const int node = instrs[prog.instr_begin + i].node;
        const int c = __ldcg(&p.cnts[prog.instr_begin + i]);
        p.cnts[prog.instr_begin + i] = 0; /* ready for the next store batch */
        if (node >= 0)
        {
            delta += (long long)c - tchg[node - p.n];
            tchg[node - p.n] = c;
        }
        else
            rootc += c; /* root step: never cached (TreeEvaluation.c:238-264) */
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        delta += __shfl_xor_sync(0xffffffffu, delta, o);
        rootc += __shfl_xor_sync(0xffffffffu, rootc, o);
    }
    if ((tid & 31) == 0)
    {
        if (delta)
            atomicAdd((unsigned long long *)&s_delta, (unsigned long long)delta);
        if (rootc)
            atomicAdd((unsigned long long *)&s_root, (unsigned long long)rootc);
    }
    __syncthreads();
    if (tid == 0)
    {
        const int k = prog.pad1;
        const int ctree = p.cand[k].tree;
        const unsigned n_pieces = (unsigned)p.cand[k].n_wait;
        if (s_delta)
            atomicAdd((unsigned long long *)&p.sum_chg_rw[ctree], (unsigned long long)s_delta);
        if (s_root)
            atomicAdd(&p.totals[k], (unsigned long long)s_root);
        __threadfence();
        if (atomicAdd(&p.done[p.n_prog + k], 1u) == n_pieces - 1u)
        {
            __threadfence();
            p.done[p.n_prog + k] = 0u;
            const long long sum = (long long)atomicAdd((unsigned long long *)&p.sum_chg_rw[ctree], 0ull);
            long long len = sum + (long long)atomicExch(&p.totals[k], 0ull);
            if (p.fuse_p2p)
            { /* site blocks: this rank's partial length -> every rank, their partial lengths -> the sum */
                len = p2p_exchange_one(p.x, k, len);
                if (p.x.out2)
                    p.x.out2[k] = len;
            }
            p.out[k] = len;
        }
    }
}

template <bool PER_NODE>
__device__ __forceinline__ void fitch_exec_body(const ExecParams &p, const FitchProgram *progs, const FitchInstr *instrs)
{
    extern __shared__ uint4 smem_u4[];
    uint4 *s_instr = smem_u4;                                          /* LVB_CHUNK + 2 decoded instructions */
    int *s_cnt = reinterpret_cast<int *>(smem_u4 + LVB_CHUNK + 2);     /* PER_NODE: [LVB_CHUNK][warps] */
    uint4 *s_stack = smem_u4 + LVB_CHUNK + 2 + (PER_NODE ? LVB_CHUNK * 8 / 4 : 0); /* n_slots x blockDim */
    __shared__ unsigned long long s_total;

    const int tid = threadIdx.x;
    const int nthr = blockDim.x;
    const int nwarps = nthr >> 5;
    unsigned vb = blockIdx.x;
    if (PER_NODE && p.dep)
        vb = piece_start_ticket(p);
    /* independent programs: column-block-major, CTAs of different programs over the same columns
     * run together, so rows shared between candidates of one tree are served from L2.
     * dependent pieces: piece-major (pieces are sorted by level), so that by the time a piece
     * starts the pieces below it have normally finished — a soft barrier between levels */
    const bool piece_major = PER_NODE && p.dep;
    const int prog_id = piece_major ? vb / p.n_colblocks : vb % p.n_prog;
    const int cb = piece_major ? vb % p.n_colblocks : vb / p.n_prog;
    const FitchProgram prog = progs[prog_id];
    if (PER_NODE && p.dep && prog.flags > 0 && !piece_wait_children(p, prog_id, cb, prog.flags))
        return;
    const int col = cb * nthr + tid;
    const unsigned active = col < p.ncols ? 1u : 0u;
    const unsigned amask = active ? ~0u : 0u;
    /* operand rows are absolute pool rows (the host resolved tree and bank); node v of the program's tree is stored
     * to row v + tree_off (+ I for the tree's second bank, LVB_CTL_BANK1) */
    const unsigned tree_off = (unsigned)prog.tree * (unsigned)p.tree_stride;
    const FitchInstr *gi = instrs + prog.instr_begin;
    /* threads past the end of the row (last column block only) redo the last column; their
     * counts are masked and their stores suppressed */
    const unsigned long long colbase = reinterpret_cast<unsigned long long>(p.pool + min(col, p.ncols - 1));
    const unsigned stride_b = (unsigned)p.stride_u4 * 16u;
    const unsigned slot_bytes = (unsigned)nthr * 16u;
    const unsigned sa_instr = (unsigned)__cvta_generic_to_shared(s_instr);
    const unsigned sa_stack = (unsigned)__cvta_generic_to_shared(s_stack + tid);
    const unsigned sa_cnt = (unsigned)__cvta_generic_to_shared(s_cnt + (tid >> 5));
    const unsigned cnt_step = (unsigned)nwarps * 4u;
    const unsigned lane0 = (tid & 31) == 0 ? 1u : 0u;

    if (tid == 0)
        s_total = 0ull;

    uint4 acc = make_uint4(~0u, ~0u, ~0u, ~0u);
    unsigned total = 0;

    for (int base = 0; base < prog.n_instr; base += LVB_CHUNK)
    {
        const int m = min(LVB_CHUNK, prog.n_instr - base);
        __syncthreads();
        for (int i = tid; i <= m + 1; i += nthr)
        {
            uint4 d = make_uint4(0u, 0u, 0u, D_APREV | D_BSLOT); /* sentinels after the last one: no loads */
            if (i < m)
            {
                const FitchInstr in = gi[base + i];
                int a = in.a, b = in.b;
                if (LVB_OPER_IS_SLOT(a) || b == LVB_OPER_PREV)
                {
                    const int t = a;
                    a = b;
                    b = t;
                }
                unsigned w = in.ctl & LVB_CTL_STORE ? D_STORE : 0u;
                const unsigned push = (in.ctl & LVB_CTL_PUSH_MASK) >> LVB_CTL_PUSH_SHIFT;
                if (push)
                    w |= D_PUSH | ((push - 1u) * slot_bytes);
                if (a < 0)
                {
                    w |= D_APREV;
                    a = 0;
                }
                if (b < 0)
                {
                    w |= D_BSLOT;
                    b = LVB_OPER_SLOT_INDEX(b) * (int)slot_bytes;
                }
                d = make_uint4((unsigned)a, (unsigned)b, in.node >= 0 ? (unsigned)in.node + tree_off + ((in.ctl & LVB_CTL_BANK1) ? (unsigned)p.n_internal : 0u) : 0u, w);
            }
            s_instr[i] = d;
        }
        __syncthreads();

        /* software pipeline, depth 1, two instructions per trip (registers ping-pong, no moves):
         * the HBM operands of instruction i+1 are requested before instruction i is computed */
        uint4 c0 = lds128(sa_instr), c1;
        uint4 a0 = acc, b0 = acc, a1 = acc, b1 = acc;
        ldg_if<!PER_NODE>(a0, row_addr(c0.x, stride_b, colbase), !(c0.w & D_APREV));
        ldg_if<!PER_NODE>(b0, row_addr(c0.y, stride_b, colbase), !(c0.w & D_BSLOT));

#define LVB_STEP(IADDR, CADDR, CI, GA, GB, NI, NA, NB)                                                                \
    {                                                                                                                 \
        NI = lds128(IADDR);                                                                                           \
        ldg_if<!PER_NODE>(NA, row_addr(NI.x, stride_b, colbase), !(NI.w & D_APREV));                                             \
        ldg_if<!PER_NODE>(NB, row_addr(NI.y, stride_b, colbase), !(NI.w & D_BSLOT));                                             \
        const bool ap = CI.w & D_APREV;                                                                               \
        uint4 x;                                                                                                      \
        x.x = ap ? acc.x : GA.x;                                                                                      \
        x.y = ap ? acc.y : GA.y;                                                                                      \
        x.z = ap ? acc.z : GA.z;                                                                                      \
        x.w = ap ? acc.w : GA.w;                                                                                      \
        lds128_if(GB, sa_stack + CI.y, CI.w & D_BSLOT);                                                               \
        const unsigned f = fitch_step(x, GB, acc) & amask;                                                            \
        if (PER_NODE)                                                                                                 \
            sts32_if(CADDR, __reduce_add_sync(0xffffffffu, __popc(f)), lane0);                                        \
        else                                                                                                          \
            total += __popc(f);                                                                                       \
        stg_if(row_addr(CI.z, stride_b, colbase), acc, (CI.w & D_STORE) & active);                                    \
        sts128_if(sa_stack + (CI.w & ~15u), acc, CI.w & D_PUSH);                                                      \
    }

        unsigned ia = sa_instr + 16u, ca = sa_cnt;
        int i = 0;
        for (; i + 1 < m; i += 2)
        {
            LVB_STEP(ia, ca, c0, a0, b0, c1, a1, b1)
            LVB_STEP(ia + 16u, ca + cnt_step, c1, a1, b1, c0, a0, b0)
            ia += 32u;
            ca += 2u * cnt_step;
        }
        if (i < m)
            LVB_STEP(ia, ca, c0, a0, b0, c1, a1, b1)
#undef LVB_STEP

        if (PER_NODE)
        {
            __syncthreads();
            for (int k = tid; k < m; k += nthr)
            {
                int sum = 0;
                for (int w = 0; w < nwarps; w++)
                    sum += s_cnt[k * nwarps + w];
                if (sum)
                    atomicAdd(&p.cnts[prog.instr_begin + base + k], sum);
            }
        }
    }

    if (PER_NODE)
    { /* store mode: lengths come from the per-instruction counts (commit_finalize) */
        if (p.dep)
        {
            piece_release_parent(p, prog, cb);
            if (p.fused_finalize)
                piece_finalize(p, instrs, prog, prog_id);
            return;
        }
        if (p.fused_finalize)
        { /* single-level batch (program == candidate): the column block that finishes last folds the
           * counts and publishes the length — no second launch on the latency path of a round */
            __shared__ unsigned s_last;
            __syncthreads(); /* this CTA's count atomics are issued */
            if (tid == 0)
            {
                __threadfence();
                const unsigned last = atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u ? 1u : 0u;
                if (last)
                    p.done[prog_id] = 0u;
                s_last = last;
            }
            __syncthreads();
            if (s_last)
            {
                __threadfence();
                fitch_commit_finalize_body(instrs, progs, p.cnts, p.chg_rw, p.sum_chg_rw, p.n, p.n_internal, p.out, prog_id,
                                           p.peek ? p.kept : nullptr, p.newsum);
            }
        }
        return;
    }
    /* per-program length: block reduce, one 64-bit atomic per CTA */
    const unsigned w = __reduce_add_sync(0xffffffffu, total);
    if ((tid & 31) == 0 && w)
        atomicAdd(&s_total, (unsigned long long)w);
    __syncthreads();
    if (cb == 0)
    {
        /* score mode: add the part of the length that does not change, once per program:
         * sum of cached counts of clean nodes = sum_chg - sum over recomputed nodes of chg
         * (reference: TreeEvaluation.c:191-202 adds .changes of every clean node) */
        long long basev = 0;
        const int32_t *chg = p.chg + (int64_t)prog.tree * p.n_internal;
        for (int i = tid; i < prog.n_instr; i += nthr)
        {
            const int node = gi[i].node;
            if (node >= 0)
                basev -= chg[node - p.n];
        }
        for (int o = 16; o > 0; o >>= 1)
            basev += __shfl_xor_sync(0xffffffffu, basev, o);
        if ((tid & 31) == 0 && basev)
            atomicAdd(&s_total, (unsigned long long)basev);
        __syncthreads();
        if (tid == 0)
            s_total += (unsigned long long)p.sum_chg[prog.tree];
    }
    if (p.n_colblocks == 1)
    { /* one CTA holds the whole program (narrow alignments): nothing to combine across blocks */
        if (tid == 0)
            p.out[prog_id] = (long long)s_total;
        return;
    }
    if (tid == 0)
    {
        /* the column block that finishes last publishes the length and leaves totals/done zero
         * for the next batch: no memset before and no copy after the launch */
        if (s_total)
            atomicAdd(&p.totals[prog_id], s_total);
        __threadfence();
        if (atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u)
        {
            __threadfence();
            p.out[prog_id] = (long long)atomicExch(&p.totals[prog_id], 0ull);
            p.done[prog_id] = 0u;
        }
    }
}

/* Programs and instructions either in device memory (staged by the host with one copy per
 * batch) or, for the small batches of a single SA chain, INSIDE the kernel parameters: the
 * launch carries them, no host-to-device copy sits on the latency path of a round. */
template <int NP, int NI>
struct InlineBatch
{
    FitchProgram prog[NP]; /* pieces / programs of the launch */
    FitchProgram cand[NP]; /* one record per candidate (finalize) */
    FitchInstr instr[NI];
};
#define LVB_INL_SMALL_P 4
#define LVB_INL_SMALL_I 112
#define LVB_INL_LARGE_P 16
#define LVB_INL_LARGE_I 480

template <bool PER_NODE>
__global__ void __launch_bounds__(256, 6) fitch_exec_kernel(const ExecParams p)
{
    fitch_exec_body<PER_NODE>(p, p.prog, p.instr);
}
template <bool PER_NODE, int NP, int NI>
__global__ void __launch_bounds__(256, 6)
    fitch_exec_kernel_inl(const __grid_constant__ ExecParams p, const __grid_constant__ InlineBatch<NP, NI> b)
{
    fitch_exec_body<PER_NODE>(p, b.prog, b.instr);
}

/* ---- score mode, deep pipeline -------------------------------------------------------------
 * A proposal's program is a short chain (its dirty path): every instruction needs the result of
 * the one before, but NONE of its HBM operands depends on a result — all addresses are known
 * when the CTA starts.  With one instruction of look-ahead a CTA pays one memory latency per
 * node, which is what a small batch (an SA chain scoring a few proposals per launch) is made
 * of.  Here every thread streams its own 16-byte column of the operand rows through a private
 * ring in shared memory with cp.async (LDGSTS), LVB_PF instructions ahead of the arithmetic;
 * no cross-thread synchronisation is needed because a thread only ever reads what it fetched. */
#define LVB_PF 3          /* instructions in flight ahead of the one being computed: large batches */
#define LVB_RING 4        /* ring slots (LVB_PF + 1: a slot is refilled one step after it was read) */
#define LVB_PF_DEEP 7     /* the same for small batches (latency-bound) */
#define D_NOP 0x10000000u /* padding up to a multiple of LVB_RING (bit outside the slot-offset range) */

__device__ __forceinline__ void cp_async16_if(unsigned dst, unsigned long long src, unsigned pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t"
                 "@p cp.async.cg.shared.global [%0], [%1], 16;\n\t}"
                 :
                 : "r"(dst), "l"(src), "r"(pred)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

/* STORE: the same pipeline for single-level store batches (an accepted move, a re-root path):
 * results marked LVB_CTL_STORE are written to their rows, per-instruction counts are collected
 * as in fitch_exec_kernel<true>, and the last column block of a program finalizes it.
 * PF: instructions in flight ahead of the arithmetic; 3 keeps 7 CTAs per SM for the large,
 * bandwidth-bound batches, 7 is for small batches, whose duration is (instructions of the longest
 * program) x (memory latency / PF). */
template <bool STORE, int PF>
__device__ __forceinline__ void fitch_score_body(const ExecParams &p, const FitchProgram *progs, const FitchInstr *instrs)
{
    constexpr int RING = PF + 1; /* a slot is refilled one step after it was read */
    extern __shared__ uint4 smem_u4[];
    uint4 *s_instr = smem_u4; /* LVB_CHUNK + 2 * RING decoded instructions */
    const int tid = threadIdx.x;
    const int nthr = blockDim.x;
    int *s_cnt = reinterpret_cast<int *>(smem_u4 + LVB_CHUNK + 2 * RING); /* STORE: [LVB_CHUNK][warps] */
    uint4 *s_ring = smem_u4 + LVB_CHUNK + 2 * RING + (STORE ? LVB_CHUNK * 8 / 4 : 0); /* [RING][2][nthr] */
    uint4 *s_stack = s_ring + RING * 2 * nthr;                                       /* [n_slots][nthr] */
    __shared__ unsigned long long s_total;

    /* column-block-major: CTAs of different programs over the same columns run together, so
     * rows shared between candidates of one tree are served from L2.  Dependent pieces of a store
     * launch: piece-major in start order, see fitch_exec_body */
    const bool piece_major = STORE && p.dep;
    unsigned vb = blockIdx.x;
    if (piece_major)
        vb = piece_start_ticket(p);
    const int prog_id = piece_major ? vb / p.n_colblocks : vb % p.n_prog;
    const int cb = piece_major ? vb % p.n_colblocks : vb / p.n_prog;
    const FitchProgram prog = progs[prog_id];
    const int col = cb * nthr + tid;
    const unsigned active = col < p.ncols ? 1u : 0u;
    const unsigned amask = active ? ~0u : 0u;
    const unsigned tree_off = (unsigned)prog.tree * (unsigned)p.tree_stride; /* see fitch_exec_body */
    const FitchInstr *gi = instrs + prog.instr_begin;
    const unsigned long long colbase = reinterpret_cast<unsigned long long>(p.pool + min(col, p.ncols - 1));
    const unsigned stride_b = (unsigned)p.stride_u4 * 16u;
    const unsigned slot_bytes = (unsigned)nthr * 16u;
    const unsigned sa_instr = (unsigned)__cvta_generic_to_shared(s_instr);
    const unsigned sa_ring = (unsigned)__cvta_generic_to_shared(s_ring + tid);
    const unsigned sa_stack = (unsigned)__cvta_generic_to_shared(s_stack + tid);
    const int nwarps = nthr >> 5;
    const unsigned sa_cnt = (unsigned)__cvta_generic_to_shared(s_cnt + (tid >> 5));
    const unsigned cnt_step = (unsigned)nwarps * 4u;
    const unsigned lane0 = (tid & 31) == 0 ? 1u : 0u;

    if (tid == 0)
        s_total = 0ull;
    uint4 acc = make_uint4(~0u, ~0u, ~0u, ~0u);
    unsigned total = 0;

    for (int base = 0; base < prog.n_instr; base += LVB_CHUNK)
    {
        const int m = min(LVB_CHUNK, prog.n_instr - base);
        const int m_pad = (m + RING - 1) / RING * RING;
        __syncthreads();
        for (int i = tid; i < m_pad + RING; i += nthr)
        {
            uint4 d = make_uint4(0u, 0u, 0u, D_APREV | D_NOP); /* padding: no loads, no arithmetic */
            if (i < m)
            {
                const FitchInstr in = gi[base + i];
                int a = in.a, b = in.b;
                if (LVB_OPER_IS_SLOT(a) || b == LVB_OPER_PREV)
                {
                    const int t = a;
                    a = b;
                    b = t;
                }
                unsigned w = (STORE && (in.ctl & LVB_CTL_STORE)) ? D_STORE : 0u;
                const unsigned push = (in.ctl & LVB_CTL_PUSH_MASK) >> LVB_CTL_PUSH_SHIFT;
                if (push)
                    w |= D_PUSH | ((push - 1u) * slot_bytes);
                if (a < 0)
                {
                    w |= D_APREV;
                    a = 0;
                }
                if (b < 0)
                {
                    w |= D_BSLOT;
                    b = LVB_OPER_SLOT_INDEX(b) * (int)slot_bytes;
                }
                d = make_uint4((unsigned)a, (unsigned)b, in.node >= 0 ? (unsigned)in.node + tree_off + ((in.ctl & LVB_CTL_BANK1) ? (unsigned)p.n_internal : 0u) : 0u, w);
            }
            s_instr[i] = d;
        }
        __syncthreads();
        /* dependent piece: its instructions are staged, now wait for the rows of the level below */
        if (piece_major && base == 0 && prog.flags > 0 && !piece_wait_children(p, prog_id, cb, prog.flags))
            return;

#define LVB_ISSUE(J, S)                                                                                               \
    {                                                                                                                 \
        const uint4 ni = lds128(sa_instr + (unsigned)(J) * 16u);                                                      \
        cp_async16_if(sa_ring + (2u * (S)) * slot_bytes, row_addr(ni.x, stride_b, colbase), !(ni.w & D_APREV));       \
        cp_async16_if(sa_ring + (2u * (S) + 1u) * slot_bytes, row_addr(ni.y, stride_b, colbase),            \
                      !(ni.w & (D_BSLOT | D_NOP)));                                                                   \
        cp_async_commit();                                                                                            \
    }
#define LVB_SCORE_STEP(J, S)                                                                                          \
    {                                                                                                                 \
        cp_async_wait<PF - 1>(); /* this thread's operands of instruction J have landed */                        \
        const uint4 ci = lds128(sa_instr + (unsigned)(J) * 16u);                                                      \
        uint4 y;                                                                                                      \
        /* operand a: the previous result stays in acc, or acc is replaced by the fetched row */                      \
        lds128_if(acc, sa_ring + (2u * (S)) * slot_bytes, !(ci.w & D_APREV));                                         \
        y = lds128((ci.w & D_BSLOT) ? sa_stack + ci.y : sa_ring + (2u * (S) + 1u) * slot_bytes);                      \
        LVB_ISSUE((J) + PF, ((S) + PF) % RING) /* refills the slot read one step ago */                   \
        if (!(ci.w & D_NOP))                                                                                          \
        {                                                                                                             \
            const unsigned f = fitch_step(acc, y, acc) & amask;                                                       \
            if (STORE)                                                                                                \
            {                                                                                                         \
                sts32_if(sa_cnt + (unsigned)(J) * cnt_step, __reduce_add_sync(0xffffffffu, __popc(f)), lane0);        \
                stg_if(row_addr(ci.z, stride_b, colbase), acc, (ci.w & D_STORE) & active);                            \
            }                                                                                                         \
            else                                                                                                      \
                total += __popc(f);                                                                                   \
            sts128_if(sa_stack + (ci.w & 0x0ffffff0u), acc, ci.w & D_PUSH);                                           \
        }                                                                                                             \
    }

#pragma unroll
        for (int s = 0; s < PF; s++)
            LVB_ISSUE(s, s)
        for (int i = 0; i < m_pad; i += RING)
        {
#pragma unroll
            for (int s = 0; s < RING; s++)
                LVB_SCORE_STEP(i + s, s)
        }
        cp_async_wait<0>();
        if (STORE)
        {
            __syncthreads();
            for (int k = tid; k < m; k += nthr)
            {
                int sum = 0;
                for (int w = 0; w < nwarps; w++)
                    sum += s_cnt[k * nwarps + w];
                if (sum)
                    atomicAdd(&p.cnts[prog.instr_begin + base + k], sum);
            }
        }
#undef LVB_SCORE_STEP
#undef LVB_ISSUE
    }

    if (STORE && p.dep)
    {
        piece_release_parent(p, prog, cb);
        if (p.fused_finalize)
            piece_finalize(p, instrs, prog, prog_id);
        return;
    }
    if (STORE && !p.fused_finalize)
        return; /* the finalize launch follows */
    if (STORE)
    { /* the column block that finishes last folds the counts and publishes the length */
        __shared__ unsigned s_last;
        __syncthreads();
        if (tid == 0)
        {
            __threadfence();
            const unsigned last = atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u ? 1u : 0u;
            if (last)
                p.done[prog_id] = 0u;
            s_last = last;
        }
        __syncthreads();
        if (s_last)
        {
            __threadfence();
            fitch_commit_finalize_body(instrs, progs, p.cnts, p.chg_rw, p.sum_chg_rw, p.n, p.n_internal, p.out, prog_id,
                                           p.peek ? p.kept : nullptr, p.newsum);
        }
        return;
    }
    /* per-program length: block reduce, one 64-bit atomic per CTA */
    const unsigned w = __reduce_add_sync(0xffffffffu, total);
    if ((tid & 31) == 0 && w)
        atomicAdd(&s_total, (unsigned long long)w);
    __syncthreads();
    if (cb == 0)
    {
        /* add the part of the length that does not change, once per program: sum of cached
         * counts of clean nodes = sum_chg - sum over recomputed nodes of chg
         * (reference: TreeEvaluation.c:191-202 adds .changes of every clean node) */
        long long basev = 0;
        const int32_t *chg = p.chg + (int64_t)prog.tree * p.n_internal;
        for (int i = tid; i < prog.n_instr; i += nthr)
        {
            const int node = gi[i].node;
            if (node >= 0)
                basev -= chg[node - p.n];
        }
        for (int o = 16; o > 0; o >>= 1)
            basev += __shfl_xor_sync(0xffffffffu, basev, o);
        if ((tid & 31) == 0 && basev)
            atomicAdd(&s_total, (unsigned long long)basev);
        __syncthreads();
        if (tid == 0)
            s_total += (unsigned long long)p.sum_chg[prog.tree];
    }
    if (p.n_colblocks == 1)
    { /* one CTA holds the whole program (narrow alignments): nothing to combine across blocks */
        if (tid == 0)
            p.out[prog_id] = (long long)s_total;
        return;
    }
    if (tid == 0)
    {
        /* the column block that finishes last publishes the length and leaves totals/done zero
         * for the next batch: no memset before and no copy after the launch */
        if (s_total)
            atomicAdd(&p.totals[prog_id], s_total);
        __threadfence();
        if (atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u)
        {
            __threadfence();
            p.out[prog_id] = (long long)atomicExch(&p.totals[prog_id], 0ull);
            p.done[prog_id] = 0u;
        }
    }
}

/* ---- score mode, large batches, TWO columns per thread ---------------------------------------
 * A launch of hundreds of candidates of one tree is bound by instruction issue, not by HBM (ncu, 512 NNI
 * candidates of a 2000 x 1M tree: issue slots 75 %, ALU pipe 73 %, DRAM 15 % — four fifths of the operand
 * sectors hit L2): of the 70 SASS instructions of a node step only 33 are the Fitch arithmetic of a 16-byte
 * column, the rest fetches and decodes the instruction, forms addresses, issues and awaits the copies.  Here a
 * thread owns two columns (tid and tid + blockDim of its column block of 2 x blockDim columns), so that the
 * per-step overhead is paid once per 32 bytes.  Same ring, same look-ahead as fitch_score_body<false, LVB_PF>. */
__device__ __forceinline__ void fitch_score2_body(const ExecParams &p, const FitchProgram *progs, const FitchInstr *instrs)
{
    constexpr int PF = LVB_PF, RING = LVB_RING;
    extern __shared__ uint4 smem_u4[];
    uint4 *s_instr = smem_u4; /* LVB_CHUNK + 2 * RING decoded instructions */
    const int tid = threadIdx.x;
    const int nthr = blockDim.x;
    uint4 *s_ring = smem_u4 + LVB_CHUNK + 2 * RING; /* [RING][2 operands][2 columns][nthr] */
    uint4 *s_stack = s_ring + RING * 4 * nthr;      /* [n_slots][2 columns][nthr] */
    __shared__ unsigned long long s_total;

    const int prog_id = blockIdx.x % p.n_prog; /* column-block-major, see fitch_score_body */
    const int cb = blockIdx.x / p.n_prog;
    const FitchProgram prog = progs[prog_id];
    const int col0 = cb * 2 * nthr + tid, col1 = col0 + nthr;
    const unsigned amask0 = col0 < p.ncols ? ~0u : 0u, amask1 = col1 < p.ncols ? ~0u : 0u;
    const FitchInstr *gi = instrs + prog.instr_begin;
    const unsigned long long colbase0 = reinterpret_cast<unsigned long long>(p.pool + min(col0, p.ncols - 1));
    const unsigned long long colbase1 = reinterpret_cast<unsigned long long>(p.pool + min(col1, p.ncols - 1));
    const unsigned stride_b = (unsigned)p.stride_u4 * 16u;
    const unsigned half = (unsigned)nthr * 16u; /* one column set of one operand */
    const unsigned slot_bytes = 2u * half;      /* one operand / one stack slot: both column sets */
    const unsigned sa_instr = (unsigned)__cvta_generic_to_shared(s_instr);
    const unsigned sa_ring = (unsigned)__cvta_generic_to_shared(s_ring + tid);
    const unsigned sa_stack = (unsigned)__cvta_generic_to_shared(s_stack + tid);

    if (tid == 0)
        s_total = 0ull;
    uint4 acc0 = make_uint4(~0u, ~0u, ~0u, ~0u), acc1 = acc0;
    unsigned total = 0;

    for (int base = 0; base < prog.n_instr; base += LVB_CHUNK)
    {
        const int m = min(LVB_CHUNK, prog.n_instr - base);
        const int m_pad = (m + RING - 1) / RING * RING;
        __syncthreads();
        for (int i = tid; i < m_pad + RING; i += nthr)
        {
            uint4 d = make_uint4(0u, 0u, 0u, D_APREV | D_NOP); /* padding: no loads, no arithmetic */
            if (i < m)
            {
                const FitchInstr in = gi[base + i];
                int a = in.a, b = in.b;
                if (LVB_OPER_IS_SLOT(a) || b == LVB_OPER_PREV)
                {
                    const int t = a;
                    a = b;
                    b = t;
                }
                unsigned w = 0u;
                const unsigned push = (in.ctl & LVB_CTL_PUSH_MASK) >> LVB_CTL_PUSH_SHIFT;
                if (push)
                    w |= D_PUSH | ((push - 1u) * slot_bytes);
                if (a < 0)
                {
                    w |= D_APREV;
                    a = 0;
                }
                if (b < 0)
                {
                    w |= D_BSLOT;
                    b = LVB_OPER_SLOT_INDEX(b) * (int)slot_bytes;
                }
                d = make_uint4((unsigned)a, (unsigned)b, 0u, w);
            }
            s_instr[i] = d;
        }
        __syncthreads();

#define LVB_ISSUE2(J, S)                                                                                              \
    {                                                                                                                 \
        const uint4 ni = lds128(sa_instr + (unsigned)(J) * 16u);                                                      \
        const unsigned pa = !(ni.w & D_APREV), pb = !(ni.w & (D_BSLOT | D_NOP));                                      \
        cp_async16_if(sa_ring + (2u * (S)) * slot_bytes, row_addr(ni.x, stride_b, colbase0), pa);                     \
        cp_async16_if(sa_ring + (2u * (S)) * slot_bytes + half, row_addr(ni.x, stride_b, colbase1), pa);              \
        cp_async16_if(sa_ring + (2u * (S) + 1u) * slot_bytes, row_addr(ni.y, stride_b, colbase0), pb);                \
        cp_async16_if(sa_ring + (2u * (S) + 1u) * slot_bytes + half, row_addr(ni.y, stride_b, colbase1), pb);         \
        cp_async_commit();                                                                                            \
    }
#define LVB_SCORE2_STEP(J, S)                                                                                         \
    {                                                                                                                 \
        cp_async_wait<PF - 1>(); /* this thread's operands of instruction J have landed */                        \
        const uint4 ci = lds128(sa_instr + (unsigned)(J) * 16u);                                                      \
        lds128_if(acc0, sa_ring + (2u * (S)) * slot_bytes, !(ci.w & D_APREV));                                        \
        lds128_if(acc1, sa_ring + (2u * (S)) * slot_bytes + half, !(ci.w & D_APREV));                                 \
        const unsigned ya = (ci.w & D_BSLOT) ? sa_stack + ci.y : sa_ring + (2u * (S) + 1u) * slot_bytes;              \
        const uint4 y0 = lds128(ya), y1 = lds128(ya + half);                                                          \
        LVB_ISSUE2((J) + PF, ((S) + PF) % RING) /* refills the slot read one step ago */                              \
        if (!(ci.w & D_NOP))                                                                                          \
        {                                                                                                             \
            const unsigned f0 = fitch_step(acc0, y0, acc0) & amask0;                                                  \
            const unsigned f1 = fitch_step(acc1, y1, acc1) & amask1;                                                  \
            total += __popc(f0) + __popc(f1);                                                                         \
            const unsigned pa = sa_stack + (ci.w & 0x0ffffff0u);                                                      \
            sts128_if(pa, acc0, ci.w & D_PUSH);                                                                       \
            sts128_if(pa + half, acc1, ci.w & D_PUSH);                                                                \
        }                                                                                                             \
    }

#pragma unroll
        for (int s = 0; s < PF; s++)
            LVB_ISSUE2(s, s)
        for (int i = 0; i < m_pad; i += RING)
        {
#pragma unroll
            for (int s = 0; s < RING; s++)
                LVB_SCORE2_STEP(i + s, s)
        }
        cp_async_wait<0>();
#undef LVB_SCORE2_STEP
#undef LVB_ISSUE2
    }

    /* per-program length: block reduce, one 64-bit atomic per CTA (as fitch_score_body) */
    const unsigned w = __reduce_add_sync(0xffffffffu, total);
    if ((tid & 31) == 0 && w)
        atomicAdd(&s_total, (unsigned long long)w);
    __syncthreads();
    if (cb == 0)
    {
        long long basev = 0;
        const int32_t *chg = p.chg + (int64_t)prog.tree * p.n_internal;
        for (int i = tid; i < prog.n_instr; i += nthr)
        {
            const int node = gi[i].node;
            if (node >= 0)
                basev -= chg[node - p.n];
        }
        for (int o = 16; o > 0; o >>= 1)
            basev += __shfl_xor_sync(0xffffffffu, basev, o);
        if ((tid & 31) == 0 && basev)
            atomicAdd(&s_total, (unsigned long long)basev);
        __syncthreads();
        if (tid == 0)
            s_total += (unsigned long long)p.sum_chg[prog.tree];
    }
    if (p.n_colblocks == 1)
    {
        if (tid == 0)
            p.out[prog_id] = (long long)s_total;
        return;
    }
    if (tid == 0)
    {
        if (s_total)
            atomicAdd(&p.totals[prog_id], s_total);
        __threadfence();
        if (atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u)
        {
            __threadfence();
            p.out[prog_id] = (long long)atomicExch(&p.totals[prog_id], 0ull);
            p.done[prog_id] = 0u;
        }
    }
}

__global__ void __launch_bounds__(256) fitch_score2_kernel(const ExecParams p)
{
    fitch_score2_body(p, p.prog, p.instr);
}

__global__ void __launch_bounds__(256) fitch_score_kernel(const ExecParams p)
{
    fitch_score_body<false, LVB_PF>(p, p.prog, p.instr);
}
template <int NP, int NI>
__global__ void __launch_bounds__(256)
    fitch_score_kernel_inl(const __grid_constant__ ExecParams p, const __grid_constant__ InlineBatch<NP, NI> b)
{
    fitch_score_body<false, LVB_PF>(p, b.prog, b.instr);
}
/* small batches: deep pipeline; STORE: single-level store batches */
template <bool STORE>
__global__ void __launch_bounds__(256) fitch_pipe_kernel(const ExecParams p)
{
    fitch_score_body<STORE, LVB_PF_DEEP>(p, p.prog, p.instr);
}
template <bool STORE, int NP, int NI>
__global__ void __launch_bounds__(256)
    fitch_pipe_kernel_inl(const __grid_constant__ ExecParams p, const __grid_constant__ InlineBatch<NP, NI> b)
{
    fitch_score_body<STORE, LVB_PF_DEEP>(p, b.prog, b.instr);
}

/* store launches of dependent pieces (a first pass, a re-root of a big tree) through the cp.async
 * ring with PF instructions of look-ahead: every HBM operand of a piece is a row that was complete
 * before the piece started, so its loads can run any distance ahead of the arithmetic */
template <int PF>
__global__ void __launch_bounds__(256) fitch_store_pipe_kernel(const ExecParams p)
{
    fitch_score_body<true, PF>(p, p.prog, p.instr);
}

static size_t exec_smem_bytes(bool per_node, int n_slots, int block)
{
    return (size_t)(LVB_CHUNK + 2 + (per_node ? LVB_CHUNK * 8 / 4 : 0) + (size_t)n_slots * block) * sizeof(uint4);
}

static size_t pipe_smem_bytes(bool store, int pf, int n_slots, int block)
{
    return (size_t)(LVB_CHUNK + 2 * (pf + 1) + (store ? LVB_CHUNK * 8 / 4 : 0) + (size_t)(2 * (pf + 1) + n_slots) * block) *
           sizeof(uint4);
}
/* small launches are latency-bound: the deep pipeline (and, for single-level store batches, the
 * pipelined store kernel); LVB_GPU_DEEP=0 keeps the round-1 kernels for A/B runs */
static bool deep_enabled()
{
    static const int on = [] {
        const char *e = getenv("LVB_GPU_DEEP");
        return (e && atoi(e) == 0) ? 0 : 1;
    }();
    return on != 0;
}
/* large score launches: two columns per thread (LVB_GPU_SCORE2=0: the one-column kernel, for A/B runs) */
static bool score2_enabled()
{
    static const int on = [] {
        const char *e = getenv("LVB_GPU_SCORE2");
        return (e && atoi(e) == 0) ? 0 : 1;
    }();
    return on != 0;
}
static bool small_grid(unsigned grid)
{
    static const unsigned limit = [] {
        const char *e = getenv("LVB_GPU_SMALL_GRID"); /* CTAs up to which a launch counts as latency-bound */
        return (e && atoi(e) > 0) ? (unsigned)atoi(e) : 148u * 4u;
    }();
    return grid <= limit;
}

/* Dynamic shared memory above 48 KB must be allowed per kernel (and device) first.  The limit only
 * ever goes UP: several host threads (driver groups of one GPU) launch the same kernels with
 * different sizes, and a thread that lowered the limit between another thread's call here and its
 * launch would make that launch fail. */
template <class K>
static cudaError_t want_smem(K kernel, size_t smem)
{
    if (smem <= 48 * 1024)
        return cudaSuccess;
    struct Key
    {
        const void *f;
        int dev;
        bool operator<(const Key &o) const { return f != o.f ? f < o.f : dev < o.dev; }
    };
    static std::mutex mu;
    static std::map<Key, size_t> allowed;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return e;
    std::lock_guard<std::mutex> lock(mu);
    size_t &cur = allowed[Key{reinterpret_cast<const void *>(kernel), dev}];
    if (cur >= smem)
        return cudaSuccess;
    /* in steps of 16 KB so that a slowly growing stack does not come back here every time */
    const size_t want = std::min<size_t>((smem + 16383) / 16384 * 16384, 227 * 1024);
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(want, smem))) != cudaSuccess)
        return e;
    cur = std::max(want, smem);
    return cudaSuccess;
}

/* ---- store mode, streaming kernel: TMA bulk loads + mbarrier ring, warp-specialised, persistent ----
 * The store launches that matter for bandwidth (a first pass, a re-root of a big tree, many commits at
 * once) are streams of rows: every HBM operand is a whole row segment of the CTA's column block,
 * contiguous in memory (a CTA of NT consumer threads owns NT x 16 bytes of every row), and none of
 * the addresses depends on a result.  So one PRODUCER warp per CTA walks the program ahead of the
 * arithmetic and moves each operand segment with ONE bulk asynchronous copy (cp.async.bulk, the TMA
 * engine) into a ring of LVB_SR segments in shared memory, completion counted in bytes on an mbarrier
 * per ring slot; the NT CONSUMER threads wait for a slot, take their 16 bytes, hand the slot back
 * (one mbarrier arrive per warp) and do the Fitch step.  Against the per-thread cp.async ring of
 * fitch_score_body this removes the per-thread address arithmetic, LDGSTS issue and group
 * bookkeeping from the inner loop, and the ring only holds REAL loads (half of the operands of a
 * post-order program are the previous result or a stack slot), so the same look-ahead costs half the
 * shared memory and twice as many CTAs fit an SM.  A third role, the BOOKKEEPER warp, does what follows the
 * last step of a piece (counts to global memory, the piece counted in for the level above, the finalize).
 * The CTAs are persistent: the producer takes the next (piece, column block) ticket, stages and
 * decodes its instructions into the other instruction buffer and starts its loads while the
 * consumers still work on the current piece — no start-up or drain bubble between pieces.
 * Dependent pieces: the PRODUCER waits for the level below (piece_wait_children) before it
 * publishes a piece, the consumers never wait for anything but their operands. */
#define LVB_SR 8       /* ring slots (power of two) */
#define LVB_SCHUNK 128 /* instructions per staged chunk (96, which lets a CTA with 4 stack slots fit 4 to an SM, measured no better) */

struct StreamWork
{
    int32_t n;     /* instructions in this chunk; 0 = no more work */
    int32_t base;  /* index of the chunk's first instruction in its program */
    int32_t last;  /* last chunk of its program */
    int32_t prog_id;
    int32_t cb;
    int32_t pad[3];
    FitchProgram prog;
};

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    /* labels inside a { } block are local to it */
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "LVB_WAIT:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra LVB_DONE;\n\t"
                 "bra LVB_WAIT;\n\t"
                 "LVB_DONE:\n\t}"
                 :
                 : "r"(bar), "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void bulk_load(unsigned dst, unsigned long long src, unsigned bytes, unsigned bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :
                 : "r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
/* piece_finalize done by ONE warp (warp 0 of the consumers) while the other warps go on with the next
 * piece: the column block of a piece that finishes last folds that piece's counts into chg / sum_chg,
 * the piece of a candidate that is folded last publishes the candidate's length. */
__device__ __forceinline__ void stream_finalize_warp(const ExecParams &p, const FitchInstr *instrs, const FitchProgram &prog,
                                                     int prog_id, int lane)
{
    unsigned last = 0u;
    if (lane == 0)
    {
        __threadfence(); /* the count atomics of this CTA (issued by this warp, ordered by __syncwarp) first */
        last = atomicAdd(&p.done[prog_id], 1u) == (unsigned)p.n_colblocks - 1u ? 1u : 0u;
        if (last)
            p.done[prog_id] = 0u;
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last)
        return;
    __threadfence();
    int32_t *tchg = p.chg_rw + (int64_t)prog.tree * p.n_internal;
    long long delta = 0, rootc = 0;
    for (int i = lane; i < prog.n_instr; i += 32)
    {
        const int node = instrs[prog.instr_begin + i].node;
        const int c = __ldcg(&p.cnts[prog.instr_begin + i]);
        p.cnts[prog.instr_begin + i] = 0; /* ready for the next store batch */
        if (node >= 0)
        {
            delta += (long long)c - tchg[node - p.n];
            tchg[node - p.n] = c;
        }
        else
            rootc += c; /* root step: never cached (TreeEvaluation.c:238-264) */
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        delta += __shfl_xor_sync(0xffffffffu, delta, o);
        rootc += __shfl_xor_sync(0xffffffffu, rootc, o);
    }
    if (lane == 0)
    {
        const int k = prog.pad1;
        const int ctree = p.cand[k].tree;
        const unsigned n_pieces = (unsigned)p.cand[k].n_wait;
        if (delta)
            atomicAdd((unsigned long long *)&p.sum_chg_rw[ctree], (unsigned long long)delta);
        if (rootc)
            atomicAdd(&p.totals[k], (unsigned long long)rootc);
        __threadfence();
        if (atomicAdd(&p.done[p.n_prog + k], 1u) == n_pieces - 1u)
        {
            __threadfence();
            p.done[p.n_prog + k] = 0u;
            const long long sum = (long long)atomicAdd((unsigned long long *)&p.sum_chg_rw[ctree], 0ull);
            long long len = sum + (long long)atomicExch(&p.totals[k], 0ull);
            if (p.fuse_p2p)
            { /* site blocks: this rank's partial length -> every rank, their partial lengths -> the sum */
                len = p2p_exchange_one(p.x, k, len);
                if (p.x.out2)
                    p.x.out2[k] = len;
            }
            p.out[k] = len;
        }
    }
}

__global__ void __launch_bounds__(320) fitch_stream_kernel(const ExecParams p)
{
    extern __shared__ uint4 smem_u4[];
    /* full[SR], empty[SR], ifull[2], iempty[2], cdone[2] */
    __shared__ __align__(8) unsigned long long s_bar[2 * LVB_SR + 6];
    __shared__ StreamWork s_work[2];

    const int nt = (int)blockDim.x - 64; /* consumer threads; then the producer warp, then the bookkeeper warp */
    const int nwarps = nt >> 5;
    const int tid = threadIdx.x;
    uint4 *s_instr = smem_u4;                                              /* [2][LVB_SCHUNK] decoded instructions */
    /* [2][LVB_SCHUNK][nwarps] 16-bit: a warp's count of one step is at most 32 x 32 sites */
    unsigned short *s_cnt = reinterpret_cast<unsigned short *>(smem_u4 + 2 * LVB_SCHUNK);
    uint4 *s_ring = smem_u4 + 2 * LVB_SCHUNK + (2 * LVB_SCHUNK * nwarps + 7) / 8; /* [LVB_SR][nt] */
    uint4 *s_stack = s_ring + LVB_SR * nt;                                  /* [n_slots][nt] */

    /* made opaque to the compiler, which otherwise re-derives the shared window base at every use */
    unsigned bar0;
    asm volatile("mov.u32 %0, %1;" : "=r"(bar0) : "r"((unsigned)__cvta_generic_to_shared(s_bar)));
    auto full_bar = [&](unsigned s) { return bar0 + 8u * s; };
    auto empty_bar = [&](unsigned s) { return bar0 + 8u * (LVB_SR + s); };
    auto ifull_bar = [&](unsigned b) { return bar0 + 8u * (2 * LVB_SR + b); };
    auto iempty_bar = [&](unsigned b) { return bar0 + 8u * (2 * LVB_SR + 2 + b); };
    auto cdone_bar = [&](unsigned b) { return bar0 + 8u * (2 * LVB_SR + 4 + b); };
    if (tid == 0)
    {
        for (unsigned s = 0; s < LVB_SR; s++)
        {
            mbar_init(full_bar(s), 1u);
            mbar_init(empty_bar(s), (unsigned)nwarps);
        }
        for (unsigned b = 0; b < 2; b++)
        {
            mbar_init(ifull_bar(b), 1u);
            mbar_init(iempty_bar(b), 1u);               /* the bookkeeper hands the buffer back */
            mbar_init(cdone_bar(b), (unsigned)nwarps);  /* every consumer warp is through with the chunk */
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const unsigned stride_b = (unsigned)p.stride_u4 * 16u;
    const unsigned slot_bytes = (unsigned)nt * 16u;
    const unsigned total = (unsigned)p.n_prog * (unsigned)p.n_colblocks;
    const bool piece_major = p.dep != nullptr;

    if (tid >= nt + 32)
    { /* ---------------- bookkeeper warp ----------------
       * What has to happen after the last step of a chunk — the per-instruction counts of the CTA into global
       * memory, the piece counted in for the level above (after a fence behind every row store of the CTA), the
       * fused finalize — is a chain of fences and atomic round trips (~10 us under load, measured with
       * LVB_GPU_TRACE).  Done by a consumer warp it stalls the whole CTA through the operand ring; here the
       * consumers only arrive on an mbarrier and go on with the next piece. */
        const int lane = tid - nt - 32;
        for (unsigned q = 0;; q++)
        {
            const unsigned buf = q & 1u;
            mbar_wait(ifull_bar(buf), (q >> 1) & 1u);
            const int m = s_work[buf].n;
            if (m == 0)
                break;
            const FitchProgram prog = s_work[buf].prog;
            const int base = s_work[buf].base, last = s_work[buf].last, prog_id = s_work[buf].prog_id, cb = s_work[buf].cb;
            const unsigned item = (unsigned)s_work[buf].pad[0];
            mbar_wait(cdone_bar(buf), (q >> 1) & 1u); /* counts and row stores of every consumer warp are issued */
            /* first of all the piece is counted in: pieces of the level above wait for nothing but this */
            if (last && piece_major && prog.parent >= 0 && lane == 0)
            {
                __threadfence();
                atomicAdd(p.dep + (size_t)prog.parent * p.n_colblocks + cb, 1u);
            }
            const unsigned short *cnt = s_cnt + buf * (LVB_SCHUNK * nwarps);
            for (int k = lane; k < m; k += 32)
            {
                int sum = 0;
                for (int w = 0; w < nwarps; w++)
                    sum += cnt[k * nwarps + w];
                if (sum)
                    atomicAdd(&p.cnts[prog.instr_begin + base + k], sum);
            }
            __syncwarp();
            if (lane == 0)
                mbar_arrive(iempty_bar(buf)); /* the producer may stage the chunk after next here */
            if (last)
            {
                if (p.fused_finalize)
                    stream_finalize_warp(p, p.instr, prog, prog_id, lane);
                if (p.trace && lane == 0)
                    p.trace[6 * (size_t)item + 5] = global_ns();
            }
        }
        return;
    }
    if (tid >= nt)
    { /* ---------------- producer warp ---------------- */
        const int lane = tid - nt;
        const unsigned sa_ring = (unsigned)__cvta_generic_to_shared(s_ring);
        unsigned jl = 0; /* loads issued */
        unsigned q = 0;  /* chunks published */
        bool first_item = true;
        for (;;)
        {
            /* the first item of a CTA is its own number (no round trip to the counter before the first load is
             * issued); the counter hands out the items from gridDim.x on */
            unsigned vb = blockIdx.x;
            if (!first_item)
            {
                if (lane == 0)
                    vb = atomicAdd(p.ticket, 1u) + gridDim.x;
                vb = __shfl_sync(0xffffffffu, vb, 0);
            }
            first_item = false;
            bool stop = vb >= total;
            if (p.trace && lane == 0 && !stop)
                p.trace[6 * (size_t)vb + 0] = global_ns();
            if (stop && lane == 0 && vb == total + gridDim.x - 1u)
                *p.ticket = 0u; /* every CTA has drawn its last ticket: ready for the next launch */
            int prog_id = 0, cb = 0;
            FitchProgram prog = {};
            if (!stop)
            {
                if (piece_major && p.lag < 0)
                { /* level by level, piece by piece (the level-0 bins get smaller towards the end), all column blocks */
                    prog_id = (int)(vb / (unsigned)p.n_colblocks);
                    cb = (int)(vb % (unsigned)p.n_colblocks);
                }
                else if (piece_major)
                { /* ticket -> (piece, column block) in key order, key = cb + level * lag (ExecParams::lag) */
                    const int ncb = p.n_colblocks, nl = p.n_levels, lag = p.lag;
                    auto upto = [&](int k) { /* items with key <= k */
                        unsigned c = 0;
                        for (int L = 0; L < nl; L++)
                            c += (unsigned)(p.level_off[L + 1] - p.level_off[L]) * (unsigned)min(max(k - L * lag + 1, 0), ncb);
                        return c;
                    };
                    int lo = 0, hi = ncb - 1 + (nl - 1) * lag; /* smallest k with upto(k) > vb */
                    while (lo < hi)
                    {
                        const int mid = (lo + hi) >> 1;
                        if (upto(mid) > vb)
                            hi = mid;
                        else
                            lo = mid + 1;
                    }
                    unsigned r = vb - (lo > 0 ? upto(lo - 1) : 0u);
                    for (int L = 0; L < nl; L++)
                    {
                        const int c = lo - L * lag;
                        if (c < 0 || c >= ncb)
                            continue;
                        const unsigned pl = (unsigned)(p.level_off[L + 1] - p.level_off[L]);
                        if (r < pl)
                        {
                            prog_id = p.level_off[L] + (int)r;
                            cb = c;
                            break;
                        }
                        r -= pl;
                    }
                }
                else
                {
                    prog_id = (int)(vb % (unsigned)p.n_prog);
                    cb = (int)(vb / (unsigned)p.n_prog);
                }
                prog = p.prog[prog_id];
            }
            const unsigned tree_off = (unsigned)prog.tree * (unsigned)p.tree_stride;
            const FitchInstr *gi = p.instr + prog.instr_begin;
            const unsigned bytes = (unsigned)min(nt, p.ncols - cb * nt) * 16u;
            const unsigned long long colbase = reinterpret_cast<unsigned long long>(p.pool + (size_t)cb * nt);
            bool waited = false;
            for (int base = 0; base < (stop ? 1 : prog.n_instr); base += LVB_SCHUNK)
            {
                const unsigned buf = q & 1u;
                mbar_wait(iempty_bar(buf), ((q >> 1) & 1u) ^ 1u); /* the consumers are done with this buffer */
                const int m = stop ? 0 : min(LVB_SCHUNK, prog.n_instr - base);
                uint4 *dst = s_instr + buf * LVB_SCHUNK;
                for (int i = lane; i < m; i += 32)
                {
                    const FitchInstr in = gi[base + i];
                    int a = in.a, b = in.b;
                    if (LVB_OPER_IS_SLOT(a) || b == LVB_OPER_PREV)
                    {
                        const int t = a;
                        a = b;
                        b = t;
                    }
                    unsigned w = (in.ctl & LVB_CTL_STORE) ? D_STORE : 0u;
                    const unsigned push = (in.ctl & LVB_CTL_PUSH_MASK) >> LVB_CTL_PUSH_SHIFT;
                    if (push)
                        w |= D_PUSH | ((push - 1u) * slot_bytes);
                    if (a < 0)
                    {
                        w |= D_APREV;
                        a = 0;
                    }
                    if (b < 0)
                    {
                        w |= D_BSLOT;
                        b = LVB_OPER_SLOT_INDEX(b) * (int)slot_bytes;
                    }
                    dst[i] = make_uint4((unsigned)a, (unsigned)b, in.node >= 0 ? (unsigned)in.node + tree_off + ((in.ctl & LVB_CTL_BANK1) ? (unsigned)p.n_internal : 0u) : 0u, w);
                }
                if (p.trace && lane == 0 && !stop && base == 0)
                    p.trace[6 * (size_t)vb + 1] = global_ns();
                /* a dependent piece is published only when the rows of the level below are there */
                unsigned ok = 1u;
                if (!stop && !waited && piece_major && prog.flags > 0)
                {
                    if (lane == 0)
                    {
                        volatile unsigned *f = p.dep + (size_t)(prog.flags - 1) * p.n_colblocks + cb;
                        const unsigned target = p.dep_target[prog.flags - 1];
                        unsigned spins = 0;
                        while ((int)(*f - target) < 0)
                        {
                            __nanosleep(64);
                            if ((++spins & 1023u) == 0u && (spins >= p.spin_limit || *(volatile unsigned *)p.err != 0u))
                            {
                                ok = 0u;
                                break;
                            }
                        }
                        if (!ok)
                            atomicCAS(p.err, 0u, 1u + (unsigned)prog_id);
                        __threadfence();
                        /* rows written by other CTAs (generic proxy) are read by the bulk copies (async proxy) */
                        asm volatile("fence.proxy.async;" ::: "memory");
                    }
                    ok = __shfl_sync(0xffffffffu, ok, 0);
                    waited = true;
                }
                if (p.trace && lane == 0 && !stop && base == 0)
                    p.trace[6 * (size_t)vb + 2] = global_ns();
                if (lane == 0)
                {
                    StreamWork &wk = s_work[buf];
                    wk.pad[0] = (int32_t)vb;
                    wk.n = ok ? m : 0;
                    wk.base = base;
                    wk.last = base + LVB_SCHUNK >= prog.n_instr ? 1 : 0;
                    wk.prog_id = prog_id;
                    wk.cb = cb;
                    wk.prog = prog;
                }
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(ifull_bar(buf));
                q++;
                if (!ok)
                    stop = true;
                if (stop)
                    break;
                if (lane == 0)
                { /* the loads of this chunk, in the order the consumers take them */
                    for (int i = 0; i < m; i++)
                    {
                        const uint4 ci = dst[i];
                        if (!(ci.w & D_APREV))
                        {
                            const unsigned s = jl & (LVB_SR - 1u);
                            mbar_wait(empty_bar(s), ((jl / LVB_SR) & 1u) ^ 1u);
                            mbar_expect_tx(full_bar(s), bytes);
                            bulk_load(sa_ring + s * slot_bytes, row_addr(ci.x, stride_b, colbase), bytes, full_bar(s));
                            jl++;
                        }
                        if (!(ci.w & D_BSLOT))
                        {
                            const unsigned s = jl & (LVB_SR - 1u);
                            mbar_wait(empty_bar(s), ((jl / LVB_SR) & 1u) ^ 1u);
                            mbar_expect_tx(full_bar(s), bytes);
                            bulk_load(sa_ring + s * slot_bytes, row_addr(ci.y, stride_b, colbase), bytes, full_bar(s));
                            jl++;
                        }
                    }
                }
                __syncwarp();
            }
            if (stop)
                break;
        }
        return;
    }

    /* ---------------- consumer threads ---------------- */
    const unsigned sa_instr = (unsigned)__cvta_generic_to_shared(s_instr);
    const unsigned sa_ring = (unsigned)__cvta_generic_to_shared(s_ring + tid);
    const unsigned sa_stack = (unsigned)__cvta_generic_to_shared(s_stack + tid);
    const unsigned lane0 = (tid & 31) == 0 ? 1u : 0u;
    const unsigned cnt_step = (unsigned)nwarps * 2u;
    uint4 acc = make_uint4(~0u, ~0u, ~0u, ~0u);
    unsigned jc = 0; /* loads taken */
    for (unsigned q = 0;; q++)
    {
        const unsigned buf = q & 1u;
        mbar_wait(ifull_bar(buf), (q >> 1) & 1u);
        const int m = s_work[buf].n;
        if (m == 0)
            break;
        const int cb = s_work[buf].cb;
        const int col = cb * nt + tid;
        const unsigned active = col < p.ncols ? 1u : 0u;
        const unsigned amask = active ? ~0u : 0u;
        const unsigned long long colbase = reinterpret_cast<unsigned long long>(p.pool + min(col, p.ncols - 1));
        const unsigned ia0 = sa_instr + buf * (LVB_SCHUNK * 16u);
        const unsigned sa_cnt = (unsigned)__cvta_generic_to_shared(s_cnt + buf * (LVB_SCHUNK * nwarps) + (tid >> 5));
        const unsigned item = (unsigned)s_work[buf].pad[0];
        for (int i = 0; i < m; i++)
        {
            const uint4 ci = lds128(ia0 + (unsigned)i * 16u);
            uint4 y;
            if (!(ci.w & D_APREV))
            {
                const unsigned s = jc & (LVB_SR - 1u);
                mbar_wait(bar0 + 8u * s, (jc / LVB_SR) & 1u);
                acc = lds128(sa_ring + s * slot_bytes);
                __syncwarp();
                if (lane0)
                    mbar_arrive(bar0 + 8u * (LVB_SR + s));
                jc++;
            }
            if (ci.w & D_BSLOT)
                y = lds128(sa_stack + ci.y);
            else
            {
                const unsigned s = jc & (LVB_SR - 1u);
                mbar_wait(bar0 + 8u * s, (jc / LVB_SR) & 1u);
                y = lds128(sa_ring + s * slot_bytes);
                __syncwarp();
                if (lane0)
                    mbar_arrive(bar0 + 8u * (LVB_SR + s));
                jc++;
            }
            if (p.trace && tid == 0 && i == 0 && s_work[buf].base == 0)
                p.trace[6 * (size_t)item + 3] = global_ns();
            const unsigned f = fitch_step(acc, y, acc) & amask;
            sts16_if(sa_cnt + (unsigned)i * cnt_step, __reduce_add_sync(0xffffffffu, __popc(f)), lane0);
            stg_if(row_addr(ci.z, stride_b, colbase), acc, (ci.w & D_STORE) & active);
            sts128_if(sa_stack + (ci.w & 0x0ffffff0u), acc, ci.w & D_PUSH);
        }
        if (p.trace && tid == 0 && s_work[buf].last)
            p.trace[6 * (size_t)item + 4] = global_ns();
        /* chunk end: this warp's counts and row stores are issued (the arrive releases them to the bookkeeper) */
        __syncwarp();
        if (lane0)
            mbar_arrive(cdone_bar(buf));
    }
}


static size_t stream_smem_bytes(int n_slots, int nt)
{
    const int nwarps = nt / 32;
    return (size_t)(2 * LVB_SCHUNK + (2 * LVB_SCHUNK * nwarps + 7) / 8 + (size_t)(LVB_SR + n_slots) * nt) * sizeof(uint4);
}

cudaError_t launch_fitch_stream(const ExecParams &p, int nt, cudaStream_t stream)
{
    static int sms[LVB_MAX_DEVICES] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return e;
    if (dev < 0 || dev >= LVB_MAX_DEVICES)
        return cudaErrorInvalidDevice;
    if (!sms[dev] && (e = cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess)
        return e;
    const size_t smem = stream_smem_bytes(p.n_slots, nt);
    if ((e = want_smem(fitch_stream_kernel, smem)) != cudaSuccess)
        return e;
    int per_sm = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fitch_stream_kernel, nt + 64, smem)) != cudaSuccess)
        return e;
    if (per_sm < 1)
        return cudaErrorLaunchOutOfResources;
    const long long total = (long long)p.n_prog * p.n_colblocks;
    const unsigned grid = (unsigned)std::min<long long>(total, (long long)sms[dev] * per_sm);
    fitch_stream_kernel<<<grid, nt + 64, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_fitch_exec(const ExecParams &p, bool per_node, int block, cudaStream_t stream)
{
    const size_t smem = exec_smem_bytes(per_node, p.n_slots, block);
    const unsigned grid = (unsigned)p.n_prog * (unsigned)p.n_colblocks;
    const bool deep = deep_enabled() && small_grid(grid);
    cudaError_t e;
    if (per_node && p.stream_nt > 0)
        return launch_fitch_stream(p, p.stream_nt, stream);
    if (per_node && p.dep && p.pipe_pf > 0)
    {
        const int pf = p.pipe_pf >= 7 ? 7 : p.pipe_pf >= 5 ? 5 : 3;
        const size_t sm = pipe_smem_bytes(true, pf, p.n_slots, block);
        void (*k)(const ExecParams) = pf == 7   ? fitch_store_pipe_kernel<7>
                                      : pf == 5 ? fitch_store_pipe_kernel<5>
                                                : fitch_store_pipe_kernel<3>;
        if ((e = want_smem(k, sm)) != cudaSuccess)
            return e;
        k<<<grid, block, sm, stream>>>(p);
    }
    else if (per_node && !p.dep && deep)
    {
        const size_t sm = pipe_smem_bytes(true, LVB_PF_DEEP, p.n_slots, block);
        if ((e = want_smem(fitch_pipe_kernel<true>, sm)) != cudaSuccess)
            return e;
        fitch_pipe_kernel<true><<<grid, block, sm, stream>>>(p);
    }
    else if (per_node)
    {
        if ((e = want_smem(fitch_exec_kernel<true>, smem)) != cudaSuccess)
            return e;
        fitch_exec_kernel<true><<<grid, block, smem, stream>>>(p);
    }
    else if (p.score_v1)
    {
        if ((e = want_smem(fitch_exec_kernel<false>, smem)) != cudaSuccess)
            return e;
        fitch_exec_kernel<false><<<grid, block, smem, stream>>>(p);
    }
    else if (deep)
    {
        const size_t sm = pipe_smem_bytes(false, LVB_PF_DEEP, p.n_slots, block);
        if ((e = want_smem(fitch_pipe_kernel<false>, sm)) != cudaSuccess)
            return e;
        fitch_pipe_kernel<false><<<grid, block, sm, stream>>>(p);
    }
    else if (score2_enabled() && (long long)p.n_prog * ((p.ncols + 511) / 512) >= 4ll * 148 * 6)
    { /* many candidates: two columns per thread (issue-bound regime); the column blocks are twice as wide */
        static const int b2 = [] {
            const char *e = getenv("LVB_GPU_SCORE2_BLOCK"); /* threads per CTA of this kernel (A/B runs) */
            return (e && atoi(e) >= 32) ? atoi(e) / 32 * 32 : 0;
        }();
        /* 128 threads: 5 CTAs of 4 warps per SM instead of 2 of 8 (the ring of a CTA holds 4 operands x 2 columns per
         * thread): 0.683 -> 0.645 ms per 512 candidates of a 2000 x 1M tree (profiles/r02am_score2_block.log) */
        const int blk = b2 ? std::min(b2, 256) : std::min(block, 128);
        ExecParams q = p;
        q.n_colblocks = (p.ncols + 2 * blk - 1) / (2 * blk);
        const size_t sm = (size_t)(LVB_CHUNK + 2 * LVB_RING + (size_t)(4 * LVB_RING + 2 * p.n_slots) * blk) * sizeof(uint4);
        if ((e = want_smem(fitch_score2_kernel, sm)) != cudaSuccess)
            return e;
        fitch_score2_kernel<<<(unsigned)q.n_prog * (unsigned)q.n_colblocks, blk, sm, stream>>>(q);
    }
    else
    {
        const size_t sm = pipe_smem_bytes(false, LVB_PF, p.n_slots, block);
        if ((e = want_smem(fitch_score_kernel, sm)) != cudaSuccess)
            return e;
        fitch_score_kernel<<<grid, block, sm, stream>>>(p);
    }
    return cudaGetLastError();
}

__global__ void __launch_bounds__(1024) fitch_commit_finalize_kernel(const FitchInstr *instr, const FitchProgram *progs,
                                                                    int32_t *cnts, int32_t *chg,
                                                                    long long *sum_chg, int n, int n_internal,
                                                                    long long *lengths, int32_t *kept, long long *newsum)
{
    fitch_commit_finalize_body(instr, progs, cnts, chg, sum_chg, n, n_internal, lengths, blockIdx.x, kept, newsum);
}

/* a kept candidate becomes its tree: its counts and its sum replace the tree's */
__global__ void __launch_bounds__(256) fitch_apply_kept_kernel(const FitchInstr *instr, const FitchProgram *cand,
                                                               const int32_t *which, const int32_t *kept,
                                                               const long long *newsum, int32_t *chg, long long *sum_chg,
                                                               int n, int n_internal)
{
    const int k = which[blockIdx.x];
    const FitchProgram prog = cand[k];
    int32_t *tchg = chg + (int64_t)prog.tree * n_internal;
    for (int i = threadIdx.x; i < prog.n_instr; i += blockDim.x)
    {
        const int node = instr[prog.instr_begin + i].node;
        if (node >= 0)
            tchg[node - n] = kept[prog.instr_begin + i];
    }
    if (threadIdx.x == 0)
        sum_chg[prog.tree] = newsum[k];
}
template <int NP, int NI>
__global__ void __launch_bounds__(1024)
    fitch_commit_finalize_kernel_inl(const __grid_constant__ InlineBatch<NP, NI> b, int32_t *cnts, int32_t *chg,
                                     long long *sum_chg, int n, int n_internal, long long *lengths, int32_t *kept,
                                     long long *newsum)
{
    fitch_commit_finalize_body(b.instr, b.cand, cnts, chg, sum_chg, n, n_internal, lengths, blockIdx.x, kept, newsum);
}

/* the same with the batch and the list of candidates inside the kernel parameters */
struct WhichList
{
    int32_t k[LVB_INL_LARGE_P];
};
template <int NP, int NI>
__global__ void __launch_bounds__(128)
    fitch_apply_kept_kernel_inl(const __grid_constant__ InlineBatch<NP, NI> b, const __grid_constant__ WhichList w,
                                const int32_t *kept, const long long *newsum, int32_t *chg, long long *sum_chg, int n,
                                int n_internal)
{
    const int k = w.k[blockIdx.x];
    const FitchProgram prog = b.cand[k];
    int32_t *tchg = chg + (int64_t)prog.tree * n_internal;
    for (int i = threadIdx.x; i < prog.n_instr; i += blockDim.x)
    {
        const int node = b.instr[prog.instr_begin + i].node;
        if (node >= 0)
            tchg[node - n] = kept[prog.instr_begin + i];
    }
    if (threadIdx.x == 0)
        sum_chg[prog.tree] = newsum[k];
}

template <int NP, int NI>
static cudaError_t launch_apply_kept_inline_t(const FitchProgram *cand, int n_cand, const FitchInstr *instr, size_t n_instr,
                                              const int64_t *which, int n_which, const int32_t *kept, const long long *newsum,
                                              int32_t *chg, long long *sum_chg, int n, int n_internal, cudaStream_t stream)
{
    static thread_local InlineBatch<NP, NI> tb;
    WhichList w;
    memcpy(tb.cand, cand, (size_t)n_cand * sizeof(FitchProgram));
    memcpy(tb.instr, instr, n_instr * sizeof(FitchInstr));
    for (int i = 0; i < n_which; i++)
        w.k[i] = (int32_t)which[i];
    fitch_apply_kept_kernel_inl<NP, NI><<<n_which, 128, 0, stream>>>(tb, w, kept, newsum, chg, sum_chg, n, n_internal);
    return cudaGetLastError();
}

cudaError_t launch_apply_kept_inline(const FitchProgram *cand, int n_cand, const FitchInstr *instr, size_t n_instr,
                                     const int64_t *which, int n_which, const int32_t *kept, const long long *newsum,
                                     int32_t *chg, long long *sum_chg, int n, int n_internal, cudaStream_t stream)
{
    if (n_cand <= LVB_INL_SMALL_P && n_instr <= (size_t)LVB_INL_SMALL_I)
        return launch_apply_kept_inline_t<LVB_INL_SMALL_P, LVB_INL_SMALL_I>(cand, n_cand, instr, n_instr, which, n_which, kept,
                                                                            newsum, chg, sum_chg, n, n_internal, stream);
    return launch_apply_kept_inline_t<LVB_INL_LARGE_P, LVB_INL_LARGE_I>(cand, n_cand, instr, n_instr, which, n_which, kept, newsum,
                                                                        chg, sum_chg, n, n_internal, stream);
}

/* ---- launches with the batch inside the kernel parameters ---- */
bool inline_batch_fits(int n_prog, int n_cand, size_t n_instr)
{
    return n_prog <= LVB_INL_LARGE_P && n_cand <= LVB_INL_LARGE_P && n_instr <= (size_t)LVB_INL_LARGE_I;
}

template <int NP, int NI>
static cudaError_t launch_inline_t(const ExecParams &p, bool per_node, int block, cudaStream_t stream,
                                   const FitchProgram *prog, int n_prog, const FitchProgram *cand, int n_cand,
                                   const FitchInstr *instr, size_t n_instr, int32_t *chg, long long *sum_chg,
                                   long long *lengths, int max_instr)
{
    static thread_local InlineBatch<NP, NI> tb; /* host copy; the launch copies it into the command stream */
    memcpy(tb.prog, prog, (size_t)n_prog * sizeof(FitchProgram));
    memcpy(tb.cand, cand, (size_t)n_cand * sizeof(FitchProgram));
    memcpy(tb.instr, instr, n_instr * sizeof(FitchInstr));
    const unsigned grid = (unsigned)p.n_prog * (unsigned)p.n_colblocks;
    const bool deep = deep_enabled();
    cudaError_t e;
    if (per_node)
    {
        if (deep)
        {
            const size_t sm = pipe_smem_bytes(true, LVB_PF_DEEP, p.n_slots, block);
            if ((e = want_smem(fitch_pipe_kernel_inl<true, NP, NI>, sm)) != cudaSuccess)
                return e;
            fitch_pipe_kernel_inl<true, NP, NI><<<grid, block, sm, stream>>>(p, tb);
        }
        else
        {
            const size_t smem = exec_smem_bytes(true, p.n_slots, block);
            if ((e = want_smem(fitch_exec_kernel_inl<true, NP, NI>, smem)) != cudaSuccess)
                return e;
            fitch_exec_kernel_inl<true, NP, NI><<<grid, block, smem, stream>>>(p, tb);
        }
        if ((e = cudaGetLastError()) != cudaSuccess)
            return e;
        if (!p.fused_finalize)
        {
            int threads = 32;
            while (threads < max_instr && threads < 1024)
                threads *= 2;
            fitch_commit_finalize_kernel_inl<NP, NI><<<n_cand, threads, 0, stream>>>(tb, p.cnts, chg, sum_chg, p.n,
                                                                                    p.n_internal, lengths,
                                                                                    p.peek ? p.kept : nullptr, p.newsum);
        }
    }
    else if (deep)
    {
        const size_t sm = pipe_smem_bytes(false, LVB_PF_DEEP, p.n_slots, block);
        if ((e = want_smem(fitch_pipe_kernel_inl<false, NP, NI>, sm)) != cudaSuccess)
            return e;
        fitch_pipe_kernel_inl<false, NP, NI><<<grid, block, sm, stream>>>(p, tb);
    }
    else
    {
        const size_t sm = pipe_smem_bytes(false, LVB_PF, p.n_slots, block);
        if ((e = want_smem(fitch_score_kernel_inl<NP, NI>, sm)) != cudaSuccess)
            return e;
        fitch_score_kernel_inl<NP, NI><<<grid, block, sm, stream>>>(p, tb);
    }
    return cudaGetLastError();
}

cudaError_t launch_fitch_inline(const ExecParams &p, bool per_node, int block, cudaStream_t stream,
                                const FitchProgram *prog, int n_prog, const FitchProgram *cand, int n_cand,
                                const FitchInstr *instr, size_t n_instr, int32_t *chg, long long *sum_chg,
                                long long *lengths, int max_instr)
{
    if (n_prog <= LVB_INL_SMALL_P && n_cand <= LVB_INL_SMALL_P && n_instr <= (size_t)LVB_INL_SMALL_I)
        return launch_inline_t<LVB_INL_SMALL_P, LVB_INL_SMALL_I>(p, per_node, block, stream, prog, n_prog, cand, n_cand,
                                                                 instr, n_instr, chg, sum_chg, lengths, max_instr);
    return launch_inline_t<LVB_INL_LARGE_P, LVB_INL_LARGE_I>(p, per_node, block, stream, prog, n_prog, cand, n_cand, instr,
                                                             n_instr, chg, sum_chg, lengths, max_instr);
}

cudaError_t launch_commit_finalize(const FitchInstr *instr, const FitchProgram *prog, int n_prog, int32_t *cnts,
                                   int32_t *chg, long long *sum_chg, int n, int n_internal, long long *lengths,
                                   int max_instr, cudaStream_t stream, int32_t *kept, long long *newsum)
{
    /* one thread per instruction of the longest program where possible: the loads are dependent
     * (instruction -> count, cached count), so width, not work, sets the duration */
    const int threads = max_instr >= 1024 ? 1024 : max_instr <= 128 ? 128 : (max_instr + 31) / 32 * 32;
    fitch_commit_finalize_kernel<<<n_prog, threads, 0, stream>>>(instr, prog, cnts, chg, sum_chg, n, n_internal, lengths,
                                                                 kept, newsum);
    return cudaGetLastError();
}

cudaError_t launch_apply_kept(const FitchInstr *instr, const FitchProgram *cand, const int32_t *which, int n_which,
                              const int32_t *kept, const long long *newsum, int32_t *chg, long long *sum_chg, int n,
                              int n_internal, cudaStream_t stream)
{
    fitch_apply_kept_kernel<<<n_which, 128, 0, stream>>>(instr, cand, which, kept, newsum, chg, sum_chg, n, n_internal);
    return cudaGetLastError();
}

/* ---- exhaustive NNI neighbourhood --------------------------------------------------------
 * deterministic_hillclimb (src/Solve.c:64-142) scores the 2(n-3) trees mutate_deterministic
 * (src/TreeOperations.c:105-158) makes from one current tree.  A Fitch length does not depend
 * on where the tree is rooted, so a candidate can be evaluated on the branch that is being
 * rearranged: with A, B the subtrees below u, C the sister of u and D everything above the
 * parent of u (its "up" set, produced by the up-pass program), the current tree is
 * ((A,B),(C,D)), the candidate for left == FALSE is ((A,C),(B,D)) and the one for left == TRUE
 * is ((B,C),(A,D)); the lengths inside A, B, C and D are common to all three.  One thread
 * reads its 16-byte column of the four rows (and of the two rows the current tree has for this
 * branch) and returns both length differences: 6 rows of traffic per PAIR of candidates, whatever
 * the depth of u.  The kernel is bound by integer issue, not by HBM (ncu: ALU pipe 86 %), so
 * reading low(u)/up(u) instead of recomputing fitch(A,B)/fitch(C,D) is a net gain. */
/* (measured: sending the additions of these steps to the FMA pipe as t * one + 0x7777... with an opaque multiplier
 * takes the kernel from 0.776 to 0.823 ms — the ALU pipe is 82 % busy, but the longer IMAD latency costs more) */
__device__ __forceinline__ unsigned fitch_lane_flags(unsigned x, unsigned y)
{
    const unsigned t = x & y;
    const unsigned a = (t & 0x77777777u) + 0x77777777u;
    return ~(a | t) & 0x88888888u;
}
/* number of sites of the column whose intersection is empty (no result row wanted) */
__device__ __forceinline__ int fitch_count(const uint4 &x, const uint4 &y)
{
    const unsigned f = (fitch_lane_flags(x.x, y.x) >> 3) | (fitch_lane_flags(x.y, y.y) >> 2) |
                       (fitch_lane_flags(x.z, y.z) >> 1) | fitch_lane_flags(x.w, y.w);
    return __popc(f);
}
__device__ __forceinline__ int join_cost(const uint4 &p, const uint4 &q, const uint4 &r, const uint4 &s)
{
    /* ((p,q),(r,s)) evaluated on its middle branch */
    uint4 pq, rs;
    int c = __popc(fitch_step(p, q, pq));
    c += __popc(fitch_step(r, s, rs));
    return c + fitch_count(pq, rs);
}

__global__ void __launch_bounds__(256) nni_neighbourhood_kernel(const NniParams p)
{
    __shared__ int s_d[2];
    const int tid = threadIdx.x;
    /* column-block-major: the CTAs of neighbouring branches over the same columns run together,
     * and a row is an operand of up to four branches (its own, its sister's, its children's) */
    const int edge = blockIdx.x % p.n_edges;
    const int cb = blockIdx.x / p.n_edges;
    const int col = cb * blockDim.x + tid;
    if (tid < 2)
        s_d[tid] = 0;
    __syncthreads();
    int d0 = 0, d1 = 0;
    if (col < p.ncols)
    {
        const uint4 r = p.rows[edge];  /* pool rows of A, B, C, D */
        const uint2 r2 = p.rows2[edge]; /* pool rows of low(u) = fitch(A,B) and up(u) = fitch(C,D) */
        const uint4 *base = p.pool + col;
        const uint4 A = __ldg(base + (size_t)r.x * p.stride_u4);
        const uint4 B = __ldg(base + (size_t)r.y * p.stride_u4);
        const uint4 C = __ldg(base + (size_t)r.z * p.stride_u4);
        const uint4 D = __ldg(base + (size_t)r.w * p.stride_u4);
        const uint4 LU = __ldg(base + (size_t)r2.x * p.stride_u4);
        const uint4 UU = __ldg(base + (size_t)r2.y * p.stride_u4);
        /* the current tree on this branch costs changes(A,B) + changes(C,D) + changes(low(u), up(u)):
         * the first two are the cached counts of u in the two trees (subtracted once per branch
         * below), both sets are rows already — only the third term is computed here */
        const int cur = fitch_count(LU, UU);
        d0 = join_cost(A, C, B, D) - cur;
        d1 = join_cost(B, C, A, D) - cur;
    }
    d0 = __reduce_add_sync(0xffffffffu, d0);
    d1 = __reduce_add_sync(0xffffffffu, d1);
    if ((tid & 31) == 0)
    {
        if (d0)
            atomicAdd(&s_d[0], d0);
        if (d1)
            atomicAdd(&s_d[1], d1);
    }
    __syncthreads();
    if (tid < 2)
    {
        long long v = s_d[tid];
        if (cb == 0)
            v -= (long long)p.chg_low[edge] + (long long)p.chg_up[edge];
        if (v)
            atomicAdd(reinterpret_cast<unsigned long long *>(p.delta + 2 * (size_t)edge + tid), (unsigned long long)v);
    }
}

cudaError_t launch_nni_neighbourhood(const NniParams &p, int block, cudaStream_t stream)
{
    const unsigned grid = (unsigned)p.n_edges * (unsigned)p.n_colblocks;
    nni_neighbourhood_kernel<<<grid, block, 0, stream>>>(p);
    return cudaGetLastError();
}

/* ---- DNAToBinary on the device ---- */
__constant__ unsigned char c_state_of[256];

static void build_state_table(unsigned char *t)
{
    /* src/DataOperations.c:192-232; A=1 C=2 G=4 T=8 (src/LVB.h:73-76); 0 = not allowed */
    for (int i = 0; i < 256; i++)
        t[i] = 0;
    t['A'] = 1; t['C'] = 2; t['G'] = 4; t['T'] = 8; t['U'] = 8;
    t['Y'] = 2 | 8; t['R'] = 1 | 4; t['W'] = 1 | 8; t['S'] = 2 | 4; t['K'] = 8 | 4; t['M'] = 2 | 1;
    t['B'] = 2 | 4 | 8; t['D'] = 1 | 4 | 8; t['H'] = 1 | 2 | 8; t['V'] = 1 | 2 | 4;
    t['N'] = 15; t['X'] = 15; t['?'] = 15; t['-'] = 15;
}

__global__ void __launch_bounds__(256) pack_dna_kernel(const unsigned char *text, int64_t n, int64_t m,
                                                       int64_t word_begin, int64_t word_count, uint4 *pool,
                                                       int64_t stride_u4, int *bad)
{
    /* one thread = one 64-bit word = 16 sites of one row */
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= word_count)
        return;
    const int64_t site0 = (word_begin + w) * 16;
    for (int64_t row = blockIdx.y; row < n; row += gridDim.y)
    {
        const unsigned char *src = text + row * m;
        unsigned long long word = 0;
#pragma unroll
        for (int k = 0; k < 16; k++)
        {
            const int64_t site = site0 + k;
            unsigned s = 15u; /* padding is 'N' */
            if (site < m)
            {
                const unsigned char c = src[site];
                s = c_state_of[c];
                if (s == 0)
                    atomicMax(bad, 1 + (int)c);
            }
            word |= (unsigned long long)s << (4 * k);
        }
        reinterpret_cast<unsigned long long *>(pool + row * stride_u4)[w] = word;
    }
}

cudaError_t launch_pack_dna(const unsigned char *text, int64_t n, int64_t m, int64_t word_begin, int64_t word_count,
                            uint4 *pool, int64_t stride_u4, int *bad, cudaStream_t stream)
{
    /* __constant__ memory is per device: the table is installed once on every GPU this process
     * packs an alignment on (lvb_gpu -D 2 opens its contexts from several host threads) */
    static std::mutex table_mutex;
    static bool table_done[LVB_MAX_DEVICES] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return e;
    {
        std::lock_guard<std::mutex> lock(table_mutex);
        if (dev < 0 || dev >= LVB_MAX_DEVICES || !table_done[dev])
        {
            unsigned char t[256];
            build_state_table(t);
            if ((e = cudaMemcpyToSymbolAsync(c_state_of, t, 256, 0, cudaMemcpyHostToDevice, stream)) != cudaSuccess)
                return e;
            if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) /* t lives on this stack frame */
                return e;
            if (dev >= 0 && dev < LVB_MAX_DEVICES)
                table_done[dev] = true;
        }
    }
    dim3 grid((unsigned)((word_count + 255) / 256), (unsigned)(n < 65535 ? n : 65535));
    pack_dna_kernel<<<grid, 256, 0, stream>>>(text, n, m, word_begin, word_count, pool, stride_u4, bad);
    return cudaGetLastError();
}

/* ---- column statistics of the text alignment --------------------------------------------
 * constchar (src/DataOperations.c:272-307): a column is kept when some row differs from row 0
 * in the literal character.  MinimumTreeLength / getstatev (:53-105): per column the number of
 * distinct literal characters other than '-', '?', 'N', 'X', minus one — or MAXSTATES = 5 when
 * more than five are met (the order of the rows does not matter: only the set of characters
 * does), and -1 for a column with no countable character (strlen("") - 1, kept).
 * One thread = one column; a warp reads 32 consecutive characters of each row. */
__global__ void __launch_bounds__(256) column_stats_kernel(const unsigned char *text, int64_t n, int64_t m,
                                                           unsigned char *keep, long long *sums)
{
    __shared__ long long s_sum[3];
    if (threadIdx.x < 3)
        s_sum[threadIdx.x] = 0;
    __syncthreads();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    long long kept = 0, all = 0, of_kept = 0;
    if (k < m)
    {
        unsigned mask[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u}; /* one bit per byte value */
        const unsigned char *p = text + k;
        for (int64_t i = 0; i < n; i++, p += m)
        {
            const unsigned c = *p;
#pragma unroll
            for (int w = 0; w < 8; w++)
                mask[w] |= (c >> 5) == (unsigned)w ? 1u << (c & 31u) : 0u;
        }
        int present = 0;
#pragma unroll
        for (int w = 0; w < 8; w++)
            present += __popc(mask[w]);
        mask['-' >> 5] &= ~(1u << ('-' & 31));
        mask['?' >> 5] &= ~(1u << ('?' & 31));
        mask['N' >> 5] &= ~(1u << ('N' & 31));
        mask['X' >> 5] &= ~(1u << ('X' & 31));
        int states = 0;
#pragma unroll
        for (int w = 0; w < 8; w++)
            states += __popc(mask[w]);
        const int contrib = states > 5 ? 5 : states - 1;
        const int variable = present > 1 ? 1 : 0;
        if (keep)
            keep[k] = (unsigned char)variable;
        kept = variable;
        all = contrib;
        of_kept = variable ? contrib : 0;
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        kept += __shfl_xor_sync(0xffffffffu, kept, o);
        all += __shfl_xor_sync(0xffffffffu, all, o);
        of_kept += __shfl_xor_sync(0xffffffffu, of_kept, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd((unsigned long long *)&s_sum[0], (unsigned long long)kept);
        atomicAdd((unsigned long long *)&s_sum[1], (unsigned long long)all);
        atomicAdd((unsigned long long *)&s_sum[2], (unsigned long long)of_kept);
    }
    __syncthreads();
    if (threadIdx.x < 3 && s_sum[threadIdx.x])
        atomicAdd((unsigned long long *)&sums[threadIdx.x], (unsigned long long)s_sum[threadIdx.x]);
}

cudaError_t launch_column_stats(const unsigned char *text, int64_t n, int64_t m, unsigned char *keep, long long *sums,
                                cudaStream_t stream)
{
    const unsigned grid = (unsigned)((m + 255) / 256);
    column_stats_kernel<<<grid, 256, 0, stream>>>(text, n, m, keep, sums);
    return cudaGetLastError();
}

/* ---- column statistics without any allocation (lvb_gpu_column_stats_noalloc) -------------------------
 * The same statistics as column_stats_kernel for callers that must not allocate (code under CUDA graph
 * capture, serving loops, harnesses that forbid cudaMalloc): text, keep and the three results are the
 * CALLER's device memory, and the scratch the reduction needs — three partial sums per CTA and the ticket that
 * tells the last CTA to add them up — lives in this file-scope array instead of a cudaMalloc'd buffer.
 * Every GPU that loads the library holds the array (LVB_NOALLOC_SCRATCH_BYTES = 393,472 bytes) until exit.
 * All calls on a device share it: calls that overlap on the GPU (different streams) race; it keeps the last
 * call's contents, so the ticket is cleared on the stream before every launch (and left zero by the kernel).
 * In host code g_noalloc_scratch is only a stand-in for the symbol, not a device address: the launcher takes
 * the address with cudaGetSymbolAddress once per device, with that device current. */
#define LVB_NOALLOC_BLOCK 256
#define LVB_NOALLOC_MAX_BLOCKS (LVB_NOALLOC_MAX_COLUMNS / LVB_NOALLOC_BLOCK)
#define LVB_NOALLOC_HEAD 32 /* long longs in front of the partial sums: [0] the ticket */
#define LVB_NOALLOC_SCRATCH_WORDS (LVB_NOALLOC_HEAD + 3 * LVB_NOALLOC_MAX_BLOCKS)
static __device__ __align__(256) long long g_noalloc_scratch[LVB_NOALLOC_SCRATCH_WORDS];

/* one column: what column_stats_kernel computes per thread (DataOperations.c:272-307, 53-105) */
__device__ __forceinline__ void column_stats_of(const unsigned char *text, int64_t n, int64_t m, int64_t k,
                                                unsigned char *keep, long long &kept, long long &all, long long &of_kept)
{
    unsigned mask[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u}; /* one bit per byte value */
    const unsigned char *p = text + k;
    for (int64_t i = 0; i < n; i++, p += m)
    {
        const unsigned c = *p;
#pragma unroll
        for (int w = 0; w < 8; w++)
            mask[w] |= (c >> 5) == (unsigned)w ? 1u << (c & 31u) : 0u;
    }
    int present = 0;
#pragma unroll
    for (int w = 0; w < 8; w++)
        present += __popc(mask[w]);
    mask['-' >> 5] &= ~(1u << ('-' & 31));
    mask['?' >> 5] &= ~(1u << ('?' & 31));
    mask['N' >> 5] &= ~(1u << ('N' & 31));
    mask['X' >> 5] &= ~(1u << ('X' & 31));
    int states = 0;
#pragma unroll
    for (int w = 0; w < 8; w++)
        states += __popc(mask[w]);
    const int contrib = states > 5 ? 5 : states - 1;
    const int variable = present > 1 ? 1 : 0;
    if (keep)
        keep[k] = (unsigned char)variable;
    kept = variable;
    all = contrib;
    of_kept = variable ? contrib : 0;
}

__global__ void __launch_bounds__(LVB_NOALLOC_BLOCK)
    column_stats_noalloc_kernel(const unsigned char *text, int64_t n, int64_t m, unsigned char *keep, long long *out,
                                long long *scratch)
{
    __shared__ long long s_sum[3];
    __shared__ unsigned s_last;
    if (threadIdx.x < 3)
        s_sum[threadIdx.x] = 0;
    __syncthreads();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    long long v[3] = {0, 0, 0};
    if (k < m)
        column_stats_of(text, n, m, k, keep, v[0], v[1], v[2]);
    for (int j = 0; j < 3; j++)
    {
        for (int o = 16; o > 0; o >>= 1)
            v[j] += __shfl_xor_sync(0xffffffffu, v[j], o);
        if ((threadIdx.x & 31) == 0 && v[j])
            atomicAdd((unsigned long long *)&s_sum[j], (unsigned long long)v[j]);
    }
    __syncthreads();
    /* this CTA's partial sums into the scratch; the CTA that draws the last ticket adds all of them up */
    long long *part = scratch + LVB_NOALLOC_HEAD;
    if (threadIdx.x < 3)
        part[3 * (size_t)blockIdx.x + threadIdx.x] = s_sum[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0)
    {
        __threadfence();
        s_last = atomicAdd((unsigned *)scratch, 1u) == gridDim.x - 1u ? 1u : 0u;
        s_sum[0] = s_sum[1] = s_sum[2] = 0;
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    long long t[3] = {0, 0, 0};
    for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x)
        for (int j = 0; j < 3; j++)
            t[j] += __ldcg(&part[3 * (size_t)b + j]);
    for (int j = 0; j < 3; j++)
    {
        for (int o = 16; o > 0; o >>= 1)
            t[j] += __shfl_xor_sync(0xffffffffu, t[j], o);
        if ((threadIdx.x & 31) == 0 && t[j])
            atomicAdd((unsigned long long *)&s_sum[j], (unsigned long long)t[j]);
    }
    __syncthreads();
    if (threadIdx.x < 3)
        out[threadIdx.x] = s_sum[threadIdx.x];
    if (threadIdx.x == 0)
        *(unsigned *)scratch = 0u; /* ready for the next call */
}

cudaError_t launch_column_stats_noalloc(const unsigned char *text, int64_t n, int64_t m, unsigned char *keep, long long *out,
                                        cudaStream_t stream)
{
    if (m < 1 || m > (int64_t)LVB_NOALLOC_MAX_COLUMNS)
        return cudaErrorInvalidValue;
    /* the device address of the scratch array, taken once per device with that device current (the caller has
     * made the context's device current); no allocation anywhere on this path */
    static std::mutex mu;
    static long long *addr[LVB_MAX_DEVICES] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return e;
    if (dev < 0 || dev >= LVB_MAX_DEVICES)
        return cudaErrorInvalidDevice;
    long long *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!addr[dev])
        {
            void *ptr = nullptr;
            if ((e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch)) != cudaSuccess)
                return e;
            addr[dev] = static_cast<long long *>(ptr);
        }
        scratch = addr[dev];
    }
    /* the array keeps the last call's contents: the ticket must start at zero */
    if ((e = cudaMemsetAsync(scratch, 0, LVB_NOALLOC_HEAD * sizeof(long long), stream)) != cudaSuccess)
        return e;
    const unsigned grid = (unsigned)((m + LVB_NOALLOC_BLOCK - 1) / LVB_NOALLOC_BLOCK);
    column_stats_noalloc_kernel<<<grid, LVB_NOALLOC_BLOCK, 0, stream>>>(text, n, m, keep, out, scratch);
    return cudaGetLastError();
}

/* ---- all-reduce of partial lengths over peer memory (site-block mode) ---- */
__global__ void __launch_bounds__(256) p2p_allreduce_kernel(const __grid_constant__ P2PParams q)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= q.K)
        return;
    const long long sum = p2p_exchange_one(q, k, q.in[k]);
    if (*(volatile unsigned *)q.err != 0u)
        return;
    q.out[k] = sum;
    if (q.out2)
        q.out2[k] = sum;
}

cudaError_t launch_p2p_allreduce(const P2PParams &p, cudaStream_t stream)
{
    const unsigned grid = (unsigned)((p.K + 255) / 256);
    p2p_allreduce_kernel<<<grid, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

} // namespace lvbgpu
