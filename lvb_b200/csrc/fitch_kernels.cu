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

namespace lvbgpu
{

#define LVB_CHUNK 256 /* instructions staged in shared memory at a time */

__device__ __forceinline__ uint4 ld_row(const uint4 *p)
{
    /* rows read as operands are never written in the same launch: non-coherent path */
    return __ldg(p);
}

__device__ __forceinline__ unsigned fitch_lane(unsigned x, unsigned y, unsigned &out)
{
    const unsigned t = x & y;
    const unsigned a = (t & 0x77777777u) + 0x77777777u;
    const unsigned e8 = ~(a | t) & 0x88888888u;
    const unsigned s = e8 >> 3; /* 1 per site whose intersection is empty */
    out = t | ((x | y) & (s * 15u));
    return s;
}

/* returns the number of changes (0..32) of this 32-site column */
__device__ __forceinline__ unsigned fitch_step(const uint4 &x, const uint4 &y, uint4 &out)
{
    const unsigned s0 = fitch_lane(x.x, y.x, out.x);
    const unsigned s1 = fitch_lane(x.y, y.y, out.y);
    const unsigned s2 = fitch_lane(x.z, y.z, out.z);
    const unsigned s3 = fitch_lane(x.w, y.w, out.w);
    /* the four flag words occupy bit 0 of each nibble: interleave them and count once */
    return __popc(s0 + 2u * s1 + 4u * s2 + 8u * s3);
}

template <bool PER_NODE>
__global__ void __launch_bounds__(256) fitch_exec_kernel(const ExecParams p)
{
    extern __shared__ uint4 smem_u4[];
    FitchInstr *s_instr = reinterpret_cast<FitchInstr *>(smem_u4);             /* LVB_CHUNK x 16 B */
    int *s_cnt = reinterpret_cast<int *>(smem_u4 + LVB_CHUNK);                 /* LVB_CHUNK ints   */
    uint4 *s_stack = smem_u4 + LVB_CHUNK + LVB_CHUNK / 4;                      /* n_slots x blockDim */
    __shared__ unsigned long long s_total;

    const int tid = threadIdx.x;
    const int nthr = blockDim.x;
    /* column-block-major: CTAs of different programs over the same columns run together,
     * so rows shared between candidates of one tree are served from L2 */
    const int prog_id = blockIdx.x % p.n_prog;
    const int cb = blockIdx.x / p.n_prog;
    const FitchProgram prog = p.prog[prog_id];
    const int col = cb * nthr + tid;
    const bool active = col < p.ncols;
    const int64_t tree_off = (int64_t)prog.tree * p.n_internal; /* internal node v lives in row v + tree_off */
    const uint4 *pool = p.pool;
    const FitchInstr *gi = p.instr + prog.instr_begin;

    if (tid == 0)
        s_total = 0ull;

    const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4 acc = ones;
    unsigned total = 0;

    for (int base = 0; base < prog.n_instr; base += LVB_CHUNK)
    {
        const int m = min(LVB_CHUNK, prog.n_instr - base);
        __syncthreads();
        for (int i = tid; i < m; i += nthr)
        {
            s_instr[i] = gi[base + i];
            if (PER_NODE)
                s_cnt[i] = 0;
        }
        __syncthreads();

        /* software pipeline, depth 1: the HBM operands of instruction i+1 are requested
         * before instruction i is computed */
        FitchInstr cur = s_instr[0];
        uint4 ga = ones, gb = ones;
        if (active)
        {
            if (cur.a >= 0)
                ga = ld_row(pool + (cur.a + (cur.a >= p.n ? tree_off : 0)) * p.stride_u4 + col);
            if (cur.b >= 0)
                gb = ld_row(pool + (cur.b + (cur.b >= p.n ? tree_off : 0)) * p.stride_u4 + col);
        }
        for (int i = 0; i < m; i++)
        {
            FitchInstr nxt = cur;
            uint4 na = ones, nb = ones;
            if (i + 1 < m)
            {
                nxt = s_instr[i + 1];
                if (active)
                {
                    if (nxt.a >= 0)
                        na = ld_row(pool + (nxt.a + (nxt.a >= p.n ? tree_off : 0)) * p.stride_u4 + col);
                    if (nxt.b >= 0)
                        nb = ld_row(pool + (nxt.b + (nxt.b >= p.n ? tree_off : 0)) * p.stride_u4 + col);
                }
            }
            uint4 x, y;
            if (cur.a >= 0)
                x = ga;
            else if (cur.a == LVB_OPER_PREV)
                x = acc;
            else
                x = s_stack[LVB_OPER_SLOT_INDEX(cur.a) * nthr + tid];
            if (cur.b >= 0)
                y = gb;
            else if (cur.b == LVB_OPER_PREV)
                y = acc;
            else
                y = s_stack[LVB_OPER_SLOT_INDEX(cur.b) * nthr + tid];

            uint4 out;
            const unsigned ch = fitch_step(x, y, out);
            if (!PER_NODE)
                total += ch;
            if (PER_NODE)
            {
                const unsigned w = __reduce_add_sync(0xffffffffu, ch);
                if ((tid & 31) == 0 && w)
                    atomicAdd(&s_cnt[i], (int)w);
            }
            if ((cur.ctl & LVB_CTL_STORE) && active)
                p.pool[(cur.node + tree_off) * p.stride_u4 + col] = out;
            const unsigned push = (cur.ctl & LVB_CTL_PUSH_MASK) >> LVB_CTL_PUSH_SHIFT;
            if (push)
                s_stack[(push - 1) * nthr + tid] = out;
            acc = out;
            cur = nxt;
            ga = na;
            gb = nb;
        }
        if (PER_NODE)
        {
            __syncthreads();
            for (int i = tid; i < m; i += nthr)
                if (s_cnt[i])
                    atomicAdd(&p.cnts[prog.instr_begin + base + i], s_cnt[i]);
        }
    }

    if (PER_NODE)
        return; /* store mode: lengths come from the per-instruction counts (commit_finalize) */
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

cudaError_t launch_fitch_exec(const ExecParams &p, bool per_node, int block, cudaStream_t stream)
{
    const size_t smem = (size_t)(LVB_CHUNK + LVB_CHUNK / 4 + (size_t)p.n_slots * block) * sizeof(uint4);
    const unsigned grid = (unsigned)p.n_prog * (unsigned)p.n_colblocks;
    cudaError_t e;
    if (per_node)
    {
        if (smem > 48 * 1024 &&
            (e = cudaFuncSetAttribute(fitch_exec_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem)) != cudaSuccess)
            return e;
        fitch_exec_kernel<true><<<grid, block, smem, stream>>>(p);
    }
    else
    {
        if (smem > 48 * 1024 &&
            (e = cudaFuncSetAttribute(fitch_exec_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem)) != cudaSuccess)
            return e;
        fitch_exec_kernel<false><<<grid, block, smem, stream>>>(p);
    }
    return cudaGetLastError();
}

/* one CTA per program of a store-mode batch */
__global__ void __launch_bounds__(128) fitch_commit_finalize_kernel(const FitchInstr *instr, const FitchProgram *progs,
                                                                    int32_t *cnts, int32_t *chg,
                                                                    long long *sum_chg, int n, int n_internal,
                                                                    long long *lengths)
{
    __shared__ long long s_delta, s_root;
    const FitchProgram prog = progs[blockIdx.x];
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
        const int c = cnts[prog.instr_begin + i];
        cnts[prog.instr_begin + i] = 0; /* ready for the next store batch */
        if (node >= 0)
        {
            delta += (long long)c - tchg[node - n];
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
        sum_chg[prog.tree] = s;
        lengths[blockIdx.x] = s + s_root;
    }
}

cudaError_t launch_commit_finalize(const FitchInstr *instr, const FitchProgram *prog, int n_prog, int32_t *cnts,
                                   int32_t *chg, long long *sum_chg, int n, int n_internal, long long *lengths,
                                   cudaStream_t stream)
{
    fitch_commit_finalize_kernel<<<n_prog, 128, 0, stream>>>(instr, prog, cnts, chg, sum_chg, n, n_internal, lengths);
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
    static bool table_done = false;
    if (!table_done)
    {
        unsigned char t[256];
        build_state_table(t);
        cudaError_t e = cudaMemcpyToSymbol(c_state_of, t, 256);
        if (e != cudaSuccess)
            return e;
        table_done = true;
    }
    dim3 grid((unsigned)((word_count + 255) / 256), (unsigned)(n < 65535 ? n : 65535));
    pack_dna_kernel<<<grid, 256, 0, stream>>>(text, n, m, word_begin, word_count, pool, stride_u4, bad);
    return cudaGetLastError();
}

} // namespace lvbgpu
