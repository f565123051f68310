/* lvb_gpu.cu — context, device memory and the C-ABI of include/lvb_gpu.h.
 *
 * Device-resident layout (HBM), per context:
 *   pool      one allocation of (n + max_trees*(n-3)) rows x stride_u4 16-byte columns.
 *             rows [0,n)            : the packed alignment (DNAToBinary words of this
 *                                     context's word block), padded to a 128-byte multiple
 *                                     with all-ones ('N', DataOperations.c:187-188)
 *             rows n + t*(n-3) + k  : state row of internal node n+k of device tree t
 *   chg       int32 [max_trees][n-3] cached per-node change counts (.changes, LVB.h:125)
 *   sum_chg   int64 [max_trees]      their sum per tree
 * One state row per node: an accepted proposal is committed in place (only its dirty
 * rows are rewritten), rejected proposals are scored without writing, so the reference's
 * whole-tree copy per proposal (treecopy, TreeOperations.c:737-773) has no counterpart.
 */
#include <cuda_runtime.h>
#include <dlfcn.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <climits>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/lvb_gpu.h"
#include "fitch_isa.h"
#include "fitch_kernels.h"
#include "schedule.h"

using namespace lvbgpu;

static thread_local std::string g_err;

static int fail(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CU(call)                                                                                                      \
    do                                                                                                                \
    {                                                                                                                 \
        cudaError_t e_ = (call);                                                                                      \
        if (e_ != cudaSuccess)                                                                                        \
            return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);                  \
    } while (0)

/* ---- NCCL, bound at run time so that the library loads where NCCL is absent ---- */
struct NcclId
{
    char internal[128];
};
struct NcclApi
{
    void *lib = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(void **, int, NcclId, int) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
enum
{
    kNcclInt8 = 0,
    kNcclInt32 = 2,
    kNcclInt64 = 4,
    kNcclSum = 0
};

static int nccl_load()
{
    if (g_nccl.lib)
        return 0;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names)
    {
        g_nccl.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib)
            break;
    }
    if (!g_nccl.lib)
        return fail("NCCL: cannot dlopen libnccl.so.2: %s", dlerror());
    *(void **)&g_nccl.GetUniqueId = dlsym(g_nccl.lib, "ncclGetUniqueId");
    *(void **)&g_nccl.CommInitRank = dlsym(g_nccl.lib, "ncclCommInitRank");
    *(void **)&g_nccl.AllReduce = dlsym(g_nccl.lib, "ncclAllReduce");
    *(void **)&g_nccl.AllGather = dlsym(g_nccl.lib, "ncclAllGather");
    *(void **)&g_nccl.CommDestroy = dlsym(g_nccl.lib, "ncclCommDestroy");
    *(void **)&g_nccl.GetErrorString = dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy)
        return fail("NCCL: missing symbols in libnccl");
    return 0;
}

/* One in-flight batch: pinned host staging + its device copy + where the kernels leave the
 * lengths (pinned, device-mapped: no device-to-host copy is enqueued) */
struct Staging
{
    unsigned char *h = nullptr, *d = nullptr;
    size_t cap = 0;
    long long *h_len = nullptr; /* pinned, mapped */
    long long *d_len = nullptr; /* device alias of h_len */
    size_t cap_len = 0;
    cudaEvent_t done = nullptr; /* recorded after the last kernel that reads this slot */
    bool pending = false;
    int64_t unclaimed = 0;      /* > 0: a submitted batch whose lengths lvb_gpu_batch_wait has not fetched yet */
    /* a batch of lvb_gpu_score_keep_*: what lvb_gpu_accept_kept needs to make a kept candidate its tree */
    bool keep_batch = false, staged = false;
    size_t k_prog = 0, k_instr = 0, k_cand = 0;
    std::vector<uint8_t> kept_flag; /* per candidate: 1 = kept, 2 = accepted */
    int32_t *d_kept = nullptr;      /* [cap_kept] per-instruction counts of the batch, put aside by the peek finalize */
    size_t cap_kept = 0;
    long long *d_newsum = nullptr;  /* [cap_newsum] per candidate: the sum of counts its tree would have */
    size_t cap_newsum = 0;
    int32_t *h_which = nullptr, *d_which = nullptr; /* pinned, mapped: the candidates of one accept call */
    size_t cap_which = 0;
};

struct lvb_gpu_ctx
{
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    bool timing = false; /* record CUDA events around the Fitch kernels (lvb_gpu_set_timing) */

    int64_t n = 0, nwords = 0, word_begin = 0, word_count = 0, B = 0, I = 0;
    int64_t stride_u4 = 0, max_trees = 0;
    uint4 *pool = nullptr;
    int32_t *chg = nullptr;
    long long *sum_chg = nullptr;
    std::vector<uint8_t> tree_used, tree_ready;
    /* Every device tree owns TWO banks of I rows; bank[t * I + (v - n)] says which one holds the row of node v.
     * A commit rewrites rows in place.  A KEPT candidate (lvb_gpu_score_keep_*) is scored with its rows stored
     * into the OTHER bank of its dirty nodes; accepting it flips those bits — the tree changes without a second
     * pass over the rows.  Programs name nodes; the operands become absolute pool rows when a batch is laid out. */
    std::vector<uint8_t> bank;
    std::vector<ScheduleCompiler *> sc; /* one per host thread that compiles */

    /* batch staging */
    static const int kRing = 8;
    Staging ring[kRing];
    int ring_next = 0;
    std::vector<std::vector<FitchInstr>> c_instr;              /* per candidate */
    std::vector<std::vector<ScheduleCompiler::Piece>> c_pieces; /* per candidate */
    std::vector<int> c_slots;
    std::vector<std::string> c_err;
    std::vector<FitchProgram> prog;                    /* pieces, sorted by level */
    std::vector<std::pair<int32_t, int32_t>> levels;   /* (first piece, count) per launch */
    int piece_min = 64;                                /* smallest piece size asked of the compiler (LVB_GPU_PIECE_MIN) */
    int64_t target_warps = 148 * 96;                   /* resident warps wanted per launch */
    unsigned long long *d_totals = nullptr;            /* [cap_tot] per-program sums; zero between batches */
    unsigned *d_done = nullptr;                        /* [cap_tot] CTAs finished per program; zero between batches */
    long long *d_lengths = nullptr;                    /* [cap_tot] lengths on the device (site-block mode: all-reduced here) */
    size_t cap_tot = 0;
    unsigned *d_dep = nullptr;                         /* [cap_dep] + 1 ticket: dependency counters of one store launch; zero between */
    size_t cap_dep = 0;
    size_t dep_ncb = 0;                                /* column blocks the counters in d_dep are laid out for */
    unsigned dep_total[LVB_DEP_LEVELS] = {};           /* value of d_dep[L][*] once every launch so far is complete */
    unsigned last_level_count[LVB_DEP_LEVELS] = {};    /* pieces per level of the resident batch that count themselves in */
    unsigned last_level_pieces[LVB_DEP_LEVELS] = {};   /* pieces per level of the resident batch */
    int sm_count = 148;
    int taper = 1;                                     /* level-0 bins get smaller towards the end (LVB_GPU_TAPER=0: equal bins) */
    int stream_lag = -1;                               /* LVB_GPU_STREAM_LAG: column-block lag between levels (-1 = automatic) */
    long long *d_nni = nullptr, *h_nni = nullptr;      /* [cap_nni] length differences of an NNI neighbourhood */
    size_t cap_nni = 0;
    int64_t compile_par_work = 32768; /* ... and with at least this many dirty nodes in all */
    int compile_par_min = 8;  /* host compile: batches of >= min candidates are spread over
                                                        * OpenMP threads, >= grain candidates each */
    bool inline_ok = true;   /* small batches travel inside the kernel parameters (LVB_GPU_INLINE=0: always staged) */
    bool last_inline = false;
    bool up_pieces = true;   /* the up-pass of lvb_gpu_nni_neighbourhood is cut into pieces (LVB_GPU_UP_PIECES=0: one program) */
    bool fuse_ok = true;     /* single-level store batches finalize inside the Fitch kernel (LVB_GPU_FUSE=0: own launch) */
    bool score_v1 = false;
    bool level_launches = false;                       /* one launch per level instead of dependent pieces in one launch */
    int stream_nt = -1;                                /* dependent-piece store launches: consumer threads per CTA of the
                                                        * streaming kernel (TMA bulk loads); 0 = the cp.async / register
                                                        * pipelined kernels below (LVB_GPU_STORE_STREAM); -1 = by the row
                                                        * width, see stream_threads() */
    int store_pf = 7;                                  /* dependent-piece store launches: 0 = register-pipelined kernel,
                                                        * 2/3/5/7 = cp.async ring with that look-ahead (LVB_GPU_STORE_PF) */
    unsigned spin_limit = 1u << 26;                    /* polls (>= 100 ns each) before a dependency wait is abandoned */
    unsigned *h_err = nullptr, *d_err = nullptr;       /* mapped host word the kernels report an abandoned wait in */
    int fault_dep = 0;                                 /* LVB_GPU_FAULT=dep: the next dependent launch gets a wait that
                                                        * can never be satisfied (test of the bounded wait) */
    bool last_dep = false;
    int32_t *d_cnts = nullptr;                         /* [cap_cnts] per-instruction counts; zero between batches */
    size_t cap_cnts = 0;
    int block_override = 0;

    int64_t launches = 0;

    /* the batch that is resident on the device (for lvb_gpu_rerun_last_batch) */
    int64_t last_K = 0;
    bool last_store = false;
    int last_slots = 0;
    int last_slot = -1;
    size_t last_n_instr = 0, last_n_prog = 0;
    int last_max_instr = 0;
    /* device time of the Fitch kernel, accumulated over launches (events on the stream) */
    static const int kEvRing = 32;
    cudaEvent_t ring0[kEvRing] = {}, ring1[kEvRing] = {};
    int ring_used = 0;
    uint8_t ring_kind[kEvRing] = {};       /* 1: store launch, 0: score launch */
    int64_t ring_cand[kEvRing] = {}, ring_instr[kEvRing] = {};
    double kind_ms[2] = {0, 0};            /* profile: device time, launches, candidates, instructions per kind */
    int64_t kind_launches[2] = {0, 0}, kind_cand[2] = {0, 0}, kind_instr[2] = {0, 0};
    double kernel_ms_total = 0;
    int64_t kernel_launches_timed = 0;

    void *comm = nullptr;
    int nranks = 1, rank = 0;
    /* all-reduce of partial lengths through peer memory (see fitch_kernels.h: P2PParams); falls back to
     * ncclAllReduce when the slabs cannot be mapped (LVB_GPU_P2P=0 forces that) */
    bool p2p = false;
    unsigned long long *p2p_slab = nullptr;
    unsigned long long *p2p_peer[LVB_P2P_MAX_RANKS] = {};
    int64_t p2p_kmax = 8192;
    uint32_t p2p_seq = 0;
    /* LVB_GPU_TRACE=1: time stamps of every work item of the last streaming store launch (lvb_gpu_trace_download) */
    bool trace = false;
    unsigned long long *d_trace = nullptr;
    size_t cap_trace = 0, n_trace = 0;
    bool p2p_fuse = true;                              /* LVB_GPU_P2P_FUSE=0: always the separate all-reduce kernel */

    /* LVB_GPU_PROFILE=1: host seconds per stage of commit/score calls, printed at shutdown */
    bool profile = false;
    double t_compile = 0, t_layout = 0, t_submit = 0, t_wait = 0;
    int64_t n_calls = 0, n_cand_total = 0;
};

static inline double wall_now()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

/* the pool row that holds node v (>= n) of device tree t */
static inline int64_t pool_row(const lvb_gpu_ctx *c, int64_t t, int64_t v)
{
    return c->n + t * 2 * c->I + (int64_t)c->bank[(size_t)(t * c->I + (v - c->n))] * c->I + (v - c->n);
}

static size_t grow(size_t have, size_t need, size_t start)
{
    size_t cap = have ? have : start;
    while (cap < need)
        cap *= 2;
    return cap;
}

/* device arrays indexed by program / instruction: grown with the stream idle, zero-filled */
static int ensure_device_arrays(lvb_gpu_ctx *c, size_t n_instr, size_t n_cand)
{
    if (n_cand > c->cap_tot)
    {
        const size_t cap = grow(c->cap_tot, n_cand, 256);
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_totals);
        cudaFree(c->d_done);
        cudaFree(c->d_lengths);
        CU(cudaMalloc(&c->d_totals, cap * sizeof(unsigned long long)));
        CU(cudaMalloc(&c->d_done, cap * sizeof(unsigned)));
        CU(cudaMalloc(&c->d_lengths, cap * sizeof(long long)));
        CU(cudaMemsetAsync(c->d_totals, 0, cap * sizeof(unsigned long long), c->stream));
        CU(cudaMemsetAsync(c->d_done, 0, cap * sizeof(unsigned), c->stream));
        c->cap_tot = cap;
    }
    if (n_instr > c->cap_cnts)
    {
        const size_t cap = grow(c->cap_cnts, n_instr, 4096);
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_cnts);
        CU(cudaMalloc(&c->d_cnts, cap * sizeof(int32_t)));
        CU(cudaMemsetAsync(c->d_cnts, 0, cap * sizeof(int32_t), c->stream));
        c->cap_cnts = cap;
    }
    return 0;
}

/* next staging slot, large enough; waits only if the batch that used it 4 calls ago is still running */
static int acquire_slot(lvb_gpu_ctx *c, size_t bytes, size_t n_cand, int *slot_out)
{
    const int slot = c->ring_next;
    c->ring_next = (c->ring_next + 1) % lvb_gpu_ctx::kRing;
    Staging &st = c->ring[slot];
    if (st.unclaimed)
        return fail("too many submitted batches are waiting for lvb_gpu_batch_wait (at most %d calls may lie between "
                    "a submit and its wait)", lvb_gpu_ctx::kRing - 1);
    if (st.pending)
    {
        CU(cudaEventSynchronize(st.done));
        st.pending = false;
    }
    if (bytes > st.cap)
    {
        const size_t cap = grow(st.cap, bytes, 64 << 10);
        if (st.h)
            CU(cudaFreeHost(st.h));
        if (st.d)
            CU(cudaFree(st.d));
        CU(cudaMallocHost(&st.h, cap));
        CU(cudaMalloc(&st.d, cap));
        st.cap = cap;
    }
    if (n_cand > st.cap_len)
    {
        const size_t cap = grow(st.cap_len, n_cand, 256);
        if (st.h_len)
            CU(cudaFreeHost(st.h_len));
        CU(cudaHostAlloc(&st.h_len, cap * sizeof(long long), cudaHostAllocMapped));
        CU(cudaHostGetDevicePointer(&st.d_len, st.h_len, 0));
        st.cap_len = cap;
    }
    if (!st.done)
        CU(cudaEventCreateWithFlags(&st.done, cudaEventDisableTiming));
    *slot_out = slot;
    return 0;
}

static int pick_block(const lvb_gpu_ctx *c)
{
    if (c->block_override)
        return c->block_override;
    if (c->stride_u4 >= 4096)
        return 256;
    if (c->stride_u4 >= 1024)
        return 128;
    return 64;
}

/* consumer threads per CTA of the streaming kernel: 256 (4 KB operand segments) for wide rows, 128 when that would
 * leave under 16 column blocks — the word blocks of an 8-GPU run of 2000 x 1M: 0.0705 ms against 0.0733 per pass; at
 * a quarter of that width (31 column blocks of 256) 256 is still the better one: 0.1185 against 0.128 ms */
static int stream_threads(const lvb_gpu_ctx *c)
{
    if (c->stream_nt >= 0)
        return c->stream_nt;
    return c->stride_u4 >= 4096 ? 256 : 128;
}

/* column-block width of a store launch of dependent pieces: the streaming kernel has its own */
static int dep_block(const lvb_gpu_ctx *c)
{
    return stream_threads(c) > 0 ? stream_threads(c) : pick_block(c);
}

/* fold finished event pairs into the running kernel-time total */
static int drain_kernel_events(lvb_gpu_ctx *c)
{
    for (int i = 0; i < c->ring_used; i++)
    {
        float ms = 0;
        CU(cudaEventSynchronize(c->ring1[i]));
        CU(cudaEventElapsedTime(&ms, c->ring0[i], c->ring1[i]));
        c->kernel_ms_total += ms;
        c->kernel_launches_timed++;
        const int kd = c->ring_kind[i];
        c->kind_ms[kd] += ms;
        c->kind_launches[kd]++;
        c->kind_cand[kd] += c->ring_cand[i];
        c->kind_instr[kd] += c->ring_instr[i];
    }
    c->ring_used = 0;
    return 0;
}

/* after a host-side wait: did a kernel abandon a dependency wait?  If so the scratch counters of
 * that launch are in an unknown state: clear them, forget the resident batch, report. */
static int check_device_error(lvb_gpu_ctx *c)
{
    if (!c->h_err || *c->h_err == 0u)
        return 0;
    const unsigned code = *c->h_err;
    const unsigned piece = code - 1u;
    cudaStreamSynchronize(c->stream);
    *c->h_err = 0u;
    if (c->d_dep)
        cudaMemsetAsync(c->d_dep, 0, (c->cap_dep + 1) * sizeof(unsigned), c->stream);
    for (int L = 0; L < LVB_DEP_LEVELS; L++)
        c->dep_total[L] = 0u;
    if (c->d_totals)
        cudaMemsetAsync(c->d_totals, 0, c->cap_tot * sizeof(unsigned long long), c->stream);
    if (c->d_done)
        cudaMemsetAsync(c->d_done, 0, c->cap_tot * sizeof(unsigned), c->stream);
    if (c->d_cnts)
        cudaMemsetAsync(c->d_cnts, 0, c->cap_cnts * sizeof(int32_t), c->stream);
    cudaStreamSynchronize(c->stream);
    c->last_K = 0;
    /* the trees of the batch hold partly written rows: they must be evaluated from scratch */
    std::fill(c->tree_ready.begin(), c->tree_ready.end(), 0);
    if (code & 0x80000000u)
        return fail("device error: the partial lengths of rank %u never arrived in this rank's slab (peer-memory "
                    "all-reduce): the ranks did not make the same calls, or a rank died; every device tree must be "
                    "committed again with n_dirty < 0", code & 0x7fffffffu);
    return fail("device error: piece %u gave up waiting for the rows of the pieces below it (a CTA of the launch "
                "faulted or never ran); every device tree must be committed again with n_dirty < 0", piece);
}

/* sum buf[0..K) over the ranks, in place, on the context's stream; host_out (optional, mapped host
 * memory) also receives the sums */
static int allreduce_i64(lvb_gpu_ctx *c, long long *buf, int64_t K, long long *host_out, bool *wrote_host)
{
    if (wrote_host)
        *wrote_host = false;
    if (c->p2p && K <= c->p2p_kmax)
    {
        P2PParams q;
        for (int r = 0; r < LVB_P2P_MAX_RANKS; r++)
            q.peer[r] = c->p2p_peer[r];
        q.nranks = c->nranks;
        q.rank = c->rank;
        if (++c->p2p_seq == 0u)
            c->p2p_seq = 1u;
        q.seq = c->p2p_seq;
        q.slot = (int32_t)(c->p2p_seq % LVB_P2P_SLOTS);
        q.kmax = c->p2p_kmax;
        q.K = K;
        q.in = buf;
        q.out = buf;
        q.out2 = host_out;
        q.err = c->d_err;
        q.spin_limit = c->spin_limit;
        CU(launch_p2p_allreduce(q, c->stream));
        c->launches++;
        if (wrote_host && host_out)
            *wrote_host = true;
        return 0;
    }
    int r = g_nccl.AllReduce(buf, buf, (size_t)K, kNcclInt64, kNcclSum, c->comm, c->stream);
    if (r != 0)
        return fail("ncclAllReduce failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 0;
}

/* launch the batch that is resident in staging slot c->last_slot */
static int launch_resident(lvb_gpu_ctx *c, int64_t *lengths, bool reduce_always = false, bool defer = false)
{
    const int64_t K = c->last_K;
    const bool store = c->last_store;
    Staging &st = c->ring[c->last_slot];
    const FitchProgram *d_prog = reinterpret_cast<const FitchProgram *>(st.d);
    const FitchProgram *d_cand = d_prog + c->last_n_prog;
    const FitchInstr *d_instr = reinterpret_cast<const FitchInstr *>(d_cand + K);
    /* where the lengths go: straight to pinned host memory, or to the device when they still
     * have to be summed over the ranks */
    long long *out = c->comm ? c->d_lengths : st.d_len;

    ExecParams p;
    p.pool = c->pool;
    p.stride_u4 = c->stride_u4;
    p.ncols = (int32_t)c->stride_u4;
    p.n = (int32_t)c->n;
    p.n_internal = (int32_t)c->I;
    p.tree_stride = (int32_t)(2 * c->I);
    p.peek = st.keep_batch ? 1 : 0;
    p.kept = st.d_kept;
    p.newsum = st.d_newsum;
    const bool stream = store && c->last_dep && stream_threads(c) > 0;
    const int block = stream ? stream_threads(c) : pick_block(c);
    p.stream_nt = stream ? stream_threads(c) : 0;
    p.n_colblocks = (int32_t)((c->stride_u4 + block - 1) / block);
    p.n_slots = c->last_slots;
    p.score_v1 = c->score_v1 ? 1 : 0;
    p.instr = d_instr;
    p.totals = c->d_totals;
    p.done = c->d_done;
    p.out = out;
    p.cnts = c->d_cnts;
    p.ticket = c->last_dep ? c->d_dep + c->cap_dep : nullptr;
    p.dep = c->last_dep ? c->d_dep : nullptr;
    p.chg = c->chg;
    p.sum_chg = c->sum_chg;
    p.chg_rw = c->chg;
    p.sum_chg_rw = c->sum_chg;
    /* a store batch without pieces (program == candidate) or one launch of dependent pieces: the
     * kernel finalizes by itself */
    const bool fused = store && c->fuse_ok && c->levels.size() == 1 && (c->last_dep || c->last_n_prog == (size_t)K);
    p.fused_finalize = fused ? 1 : 0;
    p.cand = d_cand;
    p.n_cand = (int32_t)K;
    p.pipe_pf = c->store_pf;
    p.err = c->d_err;
    p.spin_limit = c->spin_limit;
    for (int L = 0; L < LVB_DEP_LEVELS; L++)
    {
        if (c->last_dep)
            c->dep_total[L] += c->last_level_count[L];
        p.dep_target[L] = c->dep_total[L];
    }
    /* site blocks, dependent pieces with the finalize inside the kernel: the kernel also exchanges the partial
     * lengths over peer memory (no all-reduce launch behind it) */
    const bool want_reduce = c->comm && (lengths || !store || reduce_always || defer);
    const bool fuse_p2p = want_reduce && fused && c->last_dep && c->p2p && c->p2p_fuse && K <= c->p2p_kmax;
    p.fuse_p2p = fuse_p2p ? 1 : 0;
    if (fuse_p2p)
    {
        for (int r = 0; r < LVB_P2P_MAX_RANKS; r++)
            p.x.peer[r] = c->p2p_peer[r];
        p.x.nranks = c->nranks;
        p.x.rank = c->rank;
        if (++c->p2p_seq == 0u)
            c->p2p_seq = 1u;
        p.x.seq = c->p2p_seq;
        p.x.slot = (int32_t)(c->p2p_seq % LVB_P2P_SLOTS);
        p.x.kmax = c->p2p_kmax;
        p.x.K = K;
        p.x.in = nullptr;
        p.x.out = nullptr;
        p.x.out2 = (lengths || defer) ? st.d_len : nullptr;
        p.x.err = c->d_err;
        p.x.spin_limit = c->spin_limit;
    }
    p.trace = nullptr;
    if (c->trace && stream)
    {
        const size_t items = (size_t)c->last_n_prog * (size_t)p.n_colblocks;
        if (items > c->cap_trace)
        {
            CU(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_trace);
            c->d_trace = nullptr;
            CU(cudaMalloc(&c->d_trace, items * 6 * sizeof(unsigned long long)));
            c->cap_trace = items;
        }
        CU(cudaMemsetAsync(c->d_trace, 0, items * 6 * sizeof(unsigned long long), c->stream));
        c->n_trace = items;
        p.trace = c->d_trace;
    }
    p.n_levels = 1;
    p.lag = -1;
    p.level_off[0] = 0;
    p.level_off[1] = (int32_t)c->last_n_prog;
    if (c->last_dep)
    { /* the pieces are sorted by level; last_level_pieces[] counts them (the top level included) */
        int nl = 0;
        for (int L = 0; L < LVB_DEP_LEVELS; L++)
        {
            p.level_off[L + 1] = p.level_off[L] + (int32_t)c->last_level_pieces[L];
            if (c->last_level_pieces[L])
                nl = L + 1;
        }
        p.n_levels = nl;
        p.lag = c->stream_lag < 0 ? -1 : (int32_t)std::min<int64_t>(c->stream_lag, p.n_colblocks);
    }
    if (c->last_dep && c->fault_dep)
    { /* test hook: the pieces above level 0 wait for a piece that does not exist */
        p.dep_target[0]++;
        c->fault_dep = 0;
    }

    const double ts0 = c->profile ? wall_now() : 0;
    int slot = -1;
    if (c->timing)
    {
        if (c->ring_used == lvb_gpu_ctx::kEvRing && drain_kernel_events(c))
            return 1;
        slot = c->ring_used++;
        c->ring_kind[slot] = store ? 1 : 0;
        c->ring_cand[slot] = K;
        c->ring_instr[slot] = (int64_t)c->last_n_instr;
        CU(cudaEventRecord(c->ring0[slot], c->stream));
    }
    if (c->last_inline)
    {
        const FitchProgram *h_prog = reinterpret_cast<const FitchProgram *>(st.h);
        const FitchProgram *h_cand = h_prog + c->last_n_prog;
        const FitchInstr *h_instr = reinterpret_cast<const FitchInstr *>(h_cand + K);
        p.prog = nullptr;
        p.instr = nullptr;
        p.n_prog = (int32_t)c->last_n_prog;
        CU(launch_fitch_inline(p, store, block, c->stream, h_prog, (int)c->last_n_prog, h_cand, (int)K, h_instr,
                               c->last_n_instr, c->chg, c->sum_chg, out, c->last_max_instr));
        c->launches += store && !fused ? 2 : 1;
    }
    /* one launch per dependency level: pieces of a level only read rows stored by earlier levels */
    for (const auto &lv : c->levels)
    {
        if (c->last_inline)
            break;
        p.prog = d_prog + lv.first;
        p.n_prog = lv.second;
        if ((int64_t)p.n_prog * p.n_colblocks > (int64_t)INT32_MAX)
            return fail("batch too large: %lld CTAs", (long long)p.n_prog * p.n_colblocks);
        CU(launch_fitch_exec(p, store, block, c->stream));
        c->launches++;
    }
    if (c->timing)
    {
        CU(cudaEventRecord(c->ring1[slot], c->stream));
        c->ev0 = c->ring0[slot];
        c->ev1 = c->ring1[slot];
        c->timed = true;
    }
    if (store && !c->last_inline && !fused)
    {
        CU(launch_commit_finalize(d_instr, d_cand, (int)K, c->d_cnts, c->chg, c->sum_chg, (int)c->n, (int)c->I, out,
                                  c->last_max_instr, c->stream, st.keep_batch ? st.d_kept : nullptr, st.d_newsum));
        c->launches++;
    }
    if (want_reduce && !fuse_p2p)
    {
        bool wrote = false;
        if (allreduce_i64(c, c->d_lengths, K, (lengths || defer) ? st.d_len : nullptr, &wrote))
            return 1;
        if ((lengths || defer) && !wrote)
            CU(cudaMemcpyAsync(st.h_len, c->d_lengths, K * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
    }
    CU(cudaEventRecord(st.done, c->stream));
    st.pending = true;
    const double ts1 = c->profile ? wall_now() : 0;
    if (c->profile)
        c->t_submit += ts1 - ts0;
    if (defer)
        st.unclaimed = K;
    if (lengths)
    {
        CU(cudaEventSynchronize(st.done));
        if (c->profile)
            c->t_wait += wall_now() - ts1;
        st.pending = false;
        if (check_device_error(c))
            return 1;
        for (int64_t k = 0; k < K; k++)
            lengths[k] = (int64_t)st.h_len[k];
    }
    return 0;
}

/* compile + launch one batch; store: commit semantics, else score semantics */
static int run_batch(lvb_gpu_ctx *c, int64_t K, const int *trees, const lvb_gpu_topology *topos, bool store,
                     int64_t *lengths, int *ticket = nullptr, const lvb_gpu_program_view *pre = nullptr,
                     const uint8_t *keep = nullptr)
{
    /* keep: a SCORE batch (nothing about the trees changes) that runs as a store launch: candidates with keep[k] != 0
     * leave their rows in the spare bank of their tree and their counts aside (peek finalize) */
    const bool keeping = keep != nullptr;
    if (keeping)
        store = true;
    if (!c->pool)
        return fail("no alignment on the device: call lvb_gpu_matrix_upload first");
    if (K <= 0)
        return 0;
    CU(cudaSetDevice(c->device));
    c->last_K = 0;
    const double tp0 = c->profile ? wall_now() : 0;
    /* Large store programs are cut into independent subtrees so that enough warps are resident:
     * a program owns ceil(cols/32) warps per piece. */
    const int64_t warps_per_piece = (c->stride_u4 + 31) / 32;
    const int64_t want_pieces = store && !keeping ? std::max<int64_t>(1, c->target_warps / (warps_per_piece * K)) : 1;
    for (int64_t k = 0; k < K; k++)
    {
        const int t = trees[k];
        if (t < 0 || t >= c->max_trees || !c->tree_used[t])
            return fail("candidate %lld: tree %d does not exist", (long long)k, t);
        if ((!store || keeping) && !c->tree_ready[t])
            return fail("candidate %lld: tree %d has never been evaluated (commit it first)", (long long)k, t);
        if (store && !keeping && !c->tree_ready[t] && topos[k].n_dirty >= 0)
            return fail("tree %d: first commit must recompute every internal node (n_dirty < 0)", t);
    }
    if (store)
    { /* in-place commits of one tree must not share a launch; nor two kept candidates of one tree the spare bank */
        std::vector<uint8_t> seen(c->max_trees, 0);
        for (int64_t k = 0; k < K; k++)
        {
            if (keeping && !keep[k])
                continue;
            if (seen[trees[k]])
                return fail(keeping ? "two kept candidates name tree %d" : "commit batch names tree %d twice", trees[k]);
            seen[trees[k]] = 1;
        }
    }
    if ((size_t)K > c->c_instr.size())
    {
        c->c_instr.resize((size_t)K);
        c->c_pieces.resize((size_t)K);
        c->c_slots.resize((size_t)K);
        c->c_err.resize((size_t)K);
    }
    /* host compile, one candidate at a time; spread over host threads when the batch is large */
    int nthreads = 1;
#ifdef _OPENMP
    /* always the caller's whole team: libgomp re-creates threads when consecutive parallel regions
     * of one thread differ in size (measured: 3x slower rounds with teams alternating 8 / 16) */
    /* ... and only when there is enough to compile: a dirty node costs ~40 ns, waking a sleeping team 0.1-0.6 ms
     * (512 NNI candidates of a 2000-taxon tree: 0.26 ms on one thread) */
    int64_t to_compile = 0, work = 0;
    for (int64_t k = 0; k < K; k++)
        if (!(pre && pre[k].instr))
        {
            to_compile++;
            work += topos[k].n_dirty < 0 ? c->I : topos[k].n_dirty;
        }
    if (to_compile >= c->compile_par_min && work >= c->compile_par_work)
        nthreads = std::min<int>((int)c->sc.size(), omp_get_max_threads());
#endif
    bool bad = false;
    auto compile_one = [&](int64_t k, ScheduleCompiler *sc) {
        if (pre && pre[k].instr)
        { /* compiled by the caller (lvb_gpu_compile_schedule on one of its threads): one piece */
            const FitchInstr *src = reinterpret_cast<const FitchInstr *>(pre[k].instr);
            c->c_instr[k].assign(src, src + pre[k].n_instr);
            c->c_pieces[k].assign(1, ScheduleCompiler::Piece{0, (int32_t)pre[k].n_instr, 0});
            c->c_err[k].clear();
            c->c_slots[k] = (int)pre[k].n_slots;
            if (pre[k].n_instr < 2 || pre[k].n_slots < 0 || pre[k].n_slots > LVB_MAX_SLOTS)
            {
                c->c_err[k] = "malformed precompiled program";
                bad = true;
            }
            return;
        }
        int piece_size = 0;
        if (want_pieces > 1)
        {
            const int64_t d = topos[k].n_dirty < 0 ? c->I : topos[k].n_dirty;
            /* not worth dependent pieces for short programs; and a piece costs its CTA a start-up and a drain
             * (ticket, decode, refill of the operand ring, bookkeeping), which pieces under piece_min
             * instructions do not earn back even when that leaves fewer CTAs than the device can hold
             * (measured on the word blocks of an 8-GPU run, profiles/r02*_sweep*) */
            if (d >= 64)
                piece_size = (int)std::max<int64_t>(c->piece_min, (d + want_pieces - 1) / want_pieces);
        }
        c->c_instr[k].clear();
        c->c_pieces[k].clear();
        c->c_err[k].clear();
        int slots = 0;
        if (!sc->compile(topos[k], store && !keeping, c->c_instr[k], slots, c->c_err[k], piece_size, &c->c_pieces[k]))
            bad = true;
        c->c_slots[k] = slots;
    };
    if (nthreads > 1)
    {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 2) num_threads(nthreads)
        for (int64_t k = 0; k < K; k++)
            compile_one(k, c->sc[omp_get_thread_num()]);
#endif
    }
    else
        for (int64_t k = 0; k < K; k++)
            compile_one(k, c->sc[0]);
    if (bad)
        for (int64_t k = 0; k < K; k++)
            if (!c->c_err[k].empty())
                return fail("candidate %lld: %s", (long long)k, c->c_err[k].c_str());
    const double tp1 = c->profile ? wall_now() : 0;

    /* lay the batch out: [pieces sorted by level][one record per candidate][instructions] */
    size_t n_instr = 0, n_prog = 0;
    int n_slots = 0;
    bool multilevel = false;
    for (int64_t k = 0; k < K; k++)
    {
        n_instr += c->c_instr[k].size();
        n_prog += c->c_pieces[k].size();
        n_slots = std::max(n_slots, c->c_slots[k]);
    }
    if (n_instr > (size_t)INT32_MAX)
        return fail("batch too large: %zu instructions", n_instr);
    if (ensure_device_arrays(c, n_instr, n_prog + (size_t)K))
        return 1;
    const size_t bytes = (n_prog + (size_t)K) * sizeof(FitchProgram) + n_instr * sizeof(FitchInstr);
    int slot;
    if (acquire_slot(c, bytes, (size_t)K, &slot))
        return 1;
    Staging &st = c->ring[slot];
    FitchProgram *h_prog = reinterpret_cast<FitchProgram *>(st.h);
    FitchProgram *h_cand = h_prog + n_prog;
    FitchInstr *h_instr = reinterpret_cast<FitchInstr *>(h_cand + K);
    c->prog.clear();
    size_t off = 0;
    for (int64_t k = 0; k < K; k++)
    {
        const auto &ins = c->c_instr[k];
        memcpy(h_instr + off, ins.data(), ins.size() * sizeof(FitchInstr));
        { /* nodes -> pool rows: operands read the bank that holds the node's row; a commit stores in place, a kept
           * candidate into the other bank (and every node of it is stored) */
            const int64_t t = trees[k];
            const uint8_t *bk = c->bank.data() + (size_t)(t * c->I);
            const int32_t n32 = (int32_t)c->n, I32 = (int32_t)c->I;
            const int32_t base = (int32_t)(c->n + t * 2 * c->I);
            const bool kp = keeping && keep[k];
            FitchInstr *ip = h_instr + off;
            for (size_t i = 0; i < ins.size(); i++)
            {
                FitchInstr &in = ip[i];
                if (in.a >= n32)
                    in.a = base + (int32_t)bk[in.a - n32] * I32 + (in.a - n32);
                if (in.b >= n32)
                    in.b = base + (int32_t)bk[in.b - n32] * I32 + (in.b - n32);
                if (in.node >= 0)
                {
                    const uint8_t b = bk[in.node - n32];
                    if (kp)
                        in.ctl |= LVB_CTL_STORE | (b ? 0u : LVB_CTL_BANK1);
                    else if (b)
                        in.ctl |= LVB_CTL_BANK1;
                }
            }
        }
        for (const auto &pc : c->c_pieces[k])
        {
            c->prog.push_back(FitchProgram{(int32_t)(off + pc.instr_begin), pc.n_instr, trees[k], pc.level, -1, 0, 0, (int32_t)k});
            multilevel = multilevel || pc.level != 0;
        }
        /* n_wait of a candidate record: the number of its pieces (fused finalize of dependent pieces) */
        h_cand[k] = FitchProgram{(int32_t)off, (int32_t)ins.size(), trees[k], store ? 1 : 0, -1,
                                 (int32_t)c->c_pieces[k].size(), 0, 0};
        off += ins.size();
    }
    /* pieces sorted by level.  Either one launch per level, or (default) ONE launch in which a
     * piece waits, per column block, for the pieces whose stored rows it reads. */
    c->levels.clear();
    c->last_dep = false;
    if (multilevel)
    {
        std::stable_sort(c->prog.begin(), c->prog.end(),
                         [](const FitchProgram &a, const FitchProgram &b) { return a.flags < b.flags; });
        if (c->level_launches)
        {
            for (size_t i = 0; i < c->prog.size();)
            {
                size_t j = i;
                while (j < c->prog.size() && c->prog[j].flags == c->prog[i].flags)
                    j++;
                c->levels.push_back({(int32_t)i, (int32_t)(j - i)});
                i = j;
            }
        }
        else
        {
            /* every piece of level L > 0 waits, per column block, for all pieces of level L - 1
             * (schedule.cpp: a level only reads leaf rows and rows stored by the levels below, and the
             * levels are contiguous); a piece below the top level counts itself in when its rows are
             * stored.  The counters run on from launch to launch (launch_resident hands out targets). */
            int top = 0;
            for (const FitchProgram &pg : c->prog)
                top = std::max(top, pg.flags);
            if (top >= LVB_DEP_LEVELS)
                return fail("internal error: %d piece levels", top + 1);
            for (int L = 0; L < LVB_DEP_LEVELS; L++)
                c->last_level_count[L] = 0;
            for (int L = 0; L < LVB_DEP_LEVELS; L++)
                c->last_level_pieces[L] = 0;
            for (FitchProgram &pg : c->prog)
            {
                c->last_level_count[pg.flags]++;
                c->last_level_pieces[pg.flags]++;
                pg.parent = pg.flags < top ? pg.flags : -1;
                pg.n_wait = 0;
            }
            c->last_level_count[top] = 0; /* nobody waits for the top level: its pieces do not count themselves in */
            const size_t n_cb = (size_t)((c->stride_u4 + dep_block(c) - 1) / dep_block(c));
            const size_t need = (size_t)LVB_DEP_LEVELS * n_cb;
            if (need > c->cap_dep || n_cb != c->dep_ncb)
            {
                CU(cudaStreamSynchronize(c->stream));
                if (need > c->cap_dep)
                {
                    const size_t cap = grow(c->cap_dep, need, 1 << 12);
                    cudaFree(c->d_dep);
                    c->d_dep = nullptr;
                    CU(cudaMalloc(&c->d_dep, (cap + 1) * sizeof(unsigned)));
                    c->cap_dep = cap;
                }
                CU(cudaMemsetAsync(c->d_dep, 0, (c->cap_dep + 1) * sizeof(unsigned), c->stream));
                for (int L = 0; L < LVB_DEP_LEVELS; L++)
                    c->dep_total[L] = 0u;
                c->dep_ncb = n_cb;
            }
            c->last_dep = true;
            c->levels.push_back({0, (int32_t)c->prog.size()});
        }
    }
    else
        c->levels.push_back({0, (int32_t)c->prog.size()});
    memcpy(h_prog, c->prog.data(), c->prog.size() * sizeof(FitchProgram));
    /* small single-level batches (one SA chain's proposals, an accepted move) ride in the kernel
     * parameters; everything else is staged on the device with one copy */
    c->last_inline = c->inline_ok && !multilevel && c->levels.size() == 1 && inline_batch_fits((int)n_prog, (int)K, n_instr);
    if (!c->last_inline)
        CU(cudaMemcpyAsync(st.d, st.h, bytes, cudaMemcpyHostToDevice, c->stream));
    if (store && !keeping)
        for (int64_t k = 0; k < K; k++)
            c->tree_ready[trees[k]] = 1;
    st.keep_batch = keeping;
    st.staged = !c->last_inline;
    if (keeping)
    {
        st.k_prog = n_prog;
        st.k_instr = n_instr;
        st.k_cand = (size_t)K;
        st.kept_flag.assign(keep, keep + K);
        if (n_instr > st.cap_kept)
        {
            const size_t cap = grow(st.cap_kept, n_instr, 4096);
            cudaFree(st.d_kept);
            st.d_kept = nullptr;
            CU(cudaMalloc(&st.d_kept, cap * sizeof(int32_t)));
            st.cap_kept = cap;
        }
        if ((size_t)K > st.cap_newsum)
        {
            const size_t cap = grow(st.cap_newsum, (size_t)K, 256);
            cudaFree(st.d_newsum);
            st.d_newsum = nullptr;
            CU(cudaMalloc(&st.d_newsum, cap * sizeof(long long)));
            st.cap_newsum = cap;
        }
    }
    c->last_K = K;
    c->last_store = store;
    c->last_slots = n_slots;
    c->last_slot = slot;
    c->last_n_instr = n_instr;
    c->last_max_instr = 0;
    for (int64_t k = 0; k < K; k++)
        c->last_max_instr = std::max(c->last_max_instr, (int)c->c_instr[k].size());
    c->last_n_prog = n_prog;
    if (c->profile)
    {
        c->t_compile += tp1 - tp0;
        c->t_layout += wall_now() - tp1;
        c->n_calls++;
        c->n_cand_total += K;
    }
    if (ticket)
    { /* asynchronous: the lengths are fetched later through the staging slot */
        *ticket = slot;
        return launch_resident(c, nullptr, false, true);
    }
    return launch_resident(c, lengths);
}

static void free_matrix(lvb_gpu_ctx *c)
{
    cudaFree(c->pool);
    cudaFree(c->chg);
    cudaFree(c->sum_chg);
    c->pool = nullptr;
    c->chg = nullptr;
    c->sum_chg = nullptr;
    for (auto *s : c->sc)
        delete s;
    c->sc.clear();
}

static int alloc_matrix(lvb_gpu_ctx *c, int64_t n, int64_t nwords, int64_t word_begin, int64_t word_count,
                        int64_t max_trees)
{
    if (n < 5)
        return fail("alignment needs at least 5 rows (MIN_N, LVB.h:107), got %lld", (long long)n);
    if (nwords < 1 || word_begin < 0 || word_count < 1 || word_begin + word_count > nwords)
        return fail("bad word block [%lld,+%lld) of %lld", (long long)word_begin, (long long)word_count,
                    (long long)nwords);
    if (max_trees < 1)
        return fail("max_trees must be >= 1");
    if (2 * n - 3 > (int64_t)INT32_MAX / 2)
        return fail("too many rows");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    free_matrix(c);
    c->last_K = 0;
    c->n = n;
    c->nwords = nwords;
    c->word_begin = word_begin;
    c->word_count = word_count;
    c->B = 2 * n - 3;
    c->I = n - 3;
    c->max_trees = max_trees;
    /* two words per 16-byte column; rows padded to 128 bytes so that every warp access is line-aligned */
    const int64_t cols = (word_count + 1) / 2;
    c->stride_u4 = (cols + 7) / 8 * 8;
    const int64_t rows = n + max_trees * 2 * c->I;
    if (rows > (int64_t)INT32_MAX)
        return fail("row pool too large");
    CU(cudaMalloc(&c->pool, (size_t)rows * c->stride_u4 * sizeof(uint4)));
    CU(cudaMalloc(&c->chg, (size_t)max_trees * c->I * sizeof(int32_t)));
    CU(cudaMalloc(&c->sum_chg, (size_t)max_trees * sizeof(long long)));
    CU(cudaMemsetAsync(c->pool, 0xFF, (size_t)n * c->stride_u4 * sizeof(uint4), c->stream));
    CU(cudaMemsetAsync(c->chg, 0, (size_t)max_trees * c->I * sizeof(int32_t), c->stream));
    CU(cudaMemsetAsync(c->sum_chg, 0, (size_t)max_trees * sizeof(long long), c->stream));
    c->tree_used.assign(max_trees, 0);
    c->tree_ready.assign(max_trees, 0);
    c->bank.assign((size_t)max_trees * (size_t)c->I, 0);
    int nsc = 1;
#ifdef _OPENMP
    nsc = std::max(1, std::min(16, omp_get_max_threads()));
#endif
    for (int i = 0; i < nsc; i++)
    {
        c->sc.push_back(new ScheduleCompiler(n));
        c->sc.back()->set_taper(c->taper);
    }
    return 0;
}

/* rows(i) gives a pointer to the nwords words of row i */
template <class RowFn>
static int upload_rows(lvb_gpu_ctx *c, RowFn rows)
{
    const size_t row_bytes = (size_t)c->word_count * 8;
    size_t chunk_rows = (64u << 20) / row_bytes;
    if (chunk_rows < 1)
        chunk_rows = 1;
    if (chunk_rows > (size_t)c->n)
        chunk_rows = (size_t)c->n;
    unsigned char *stage = nullptr;
    CU(cudaMallocHost(&stage, chunk_rows * row_bytes));
    for (int64_t r0 = 0; r0 < c->n; r0 += (int64_t)chunk_rows)
    {
        const int64_t nr = std::min<int64_t>((int64_t)chunk_rows, c->n - r0);
        for (int64_t r = 0; r < nr; r++)
            memcpy(stage + r * row_bytes, rows(r0 + r) + c->word_begin, row_bytes);
        cudaError_t e = cudaMemcpy2DAsync(c->pool + r0 * c->stride_u4, (size_t)c->stride_u4 * sizeof(uint4), stage,
                                          row_bytes, row_bytes, (size_t)nr, cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess)
        {
            cudaFreeHost(stage);
            return fail("leaf row upload failed: %s", cudaGetErrorString(e));
        }
    }
    cudaFreeHost(stage);
    return 0;
}

extern "C"
{

    const char *lvb_gpu_last_error(void) { return g_err.c_str(); }

    int lvb_gpu_device_count(int *count)
    {
        CU(cudaGetDeviceCount(count));
        return 0;
    }

    int lvb_gpu_init(int device, lvb_gpu_ctx **out)
    {
        if (!out)
            return fail("lvb_gpu_init: null out pointer");
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0)
            return fail("no CUDA device: %s (this library has no CPU path)", cudaGetErrorString(e));
        if (device < 0 || device >= count)
            return fail("device %d out of range (have %d)", device, count);
        CU(cudaSetDevice(device));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10)
            return fail("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major,
                        prop.minor);
        lvb_gpu_ctx *c = new lvb_gpu_ctx();
        c->device = device;
        CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        for (int i = 0; i < lvb_gpu_ctx::kEvRing; i++)
        {
            CU(cudaEventCreate(&c->ring0[i]));
            CU(cudaEventCreate(&c->ring1[i]));
        }
        const char *tw = getenv("LVB_GPU_TARGET_WARPS");
        if (tw && atoll(tw) > 0)
            c->target_warps = atoll(tw);
        const char *cp = getenv("LVB_GPU_COMPILE_PAR");
        if (cp && atoi(cp) > 0)
            c->compile_par_min = atoi(cp);
        const char *il = getenv("LVB_GPU_INLINE");
        c->inline_ok = !(il && atoi(il) == 0);
        const char *up = getenv("LVB_GPU_UP_PIECES");
        c->up_pieces = !(up && atoi(up) == 0);
        const char *fz = getenv("LVB_GPU_FUSE");
        c->fuse_ok = !(fz && atoi(fz) == 0);
        const char *v1 = getenv("LVB_GPU_SCORE_V1");
        c->score_v1 = v1 && atoi(v1) != 0;
        const char *pf = getenv("LVB_GPU_PROFILE");
        c->profile = pf && atoi(pf) != 0;
        if (pf && atoi(pf) >= 2) /* also device time of the Fitch kernels (events around every launch) */
            c->timing = true;
        const char *ll = getenv("LVB_GPU_LEVEL_LAUNCHES");
        c->level_launches = ll && atoi(ll) != 0;
        const char *sst = getenv("LVB_GPU_STORE_STREAM");
        if (sst)
        {
            const int v = atoi(sst);
            c->stream_nt = v <= 0 ? 0 : std::min(256, std::max(32, v / 32 * 32));
        }
        const char *pm = getenv("LVB_GPU_PIECE_MIN");
        if (pm && atoi(pm) >= 1)
            c->piece_min = atoi(pm);
        const char *tr = getenv("LVB_GPU_TRACE");
        c->trace = tr && atoi(tr) != 0;
        const char *tp = getenv("LVB_GPU_TAPER");
        c->taper = tp ? std::max(0, std::min(2, atoi(tp))) : 1;
        const char *slg = getenv("LVB_GPU_STREAM_LAG");
        if (slg)
            c->stream_lag = atoi(slg);
        cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
        const char *spf = getenv("LVB_GPU_STORE_PF");
        if (spf)
            c->store_pf = atoi(spf);
        const char *sl = getenv("LVB_GPU_SPIN_LIMIT");
        if (sl && atoll(sl) >= 1024)
            c->spin_limit = (unsigned)std::min<long long>(atoll(sl), 0xffffffffll);
        const char *ft = getenv("LVB_GPU_FAULT");
        c->fault_dep = ft && strcmp(ft, "dep") == 0;
        CU(cudaHostAlloc(&c->h_err, sizeof(unsigned), cudaHostAllocMapped));
        *c->h_err = 0u;
        CU(cudaHostGetDevicePointer(&c->d_err, c->h_err, 0));
        const char *b = getenv("LVB_GPU_BLOCK");
        if (b)
        {
            const int v = atoi(b);
            if (v >= 32 && v <= 256 && v % 32 == 0)
                c->block_override = v;
        }
        *out = c;
        return 0;
    }

    int lvb_gpu_shutdown(lvb_gpu_ctx *c)
    {
        if (!c)
            return 0;
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
        if (c->profile)
            fprintf(stderr, "[lvb_gpu profile] %lld calls, %lld candidates: compile %.3f s, layout+H2D enqueue %.3f s, "
                            "launch enqueue %.3f s, wait %.3f s\n", (long long)c->n_calls, (long long)c->n_cand_total,
                    c->t_compile, c->t_layout, c->t_submit, c->t_wait);
        if (c->profile && c->timing && drain_kernel_events(c) == 0)
            for (int kd = 0; kd < 2; kd++)
                fprintf(stderr, "[lvb_gpu profile] %s kernel: %.3f s on the device, %lld launches, %lld programs, %lld instructions\n",
                        kd ? "store" : "score", c->kind_ms[kd] * 1e-3, (long long)c->kind_launches[kd],
                        (long long)c->kind_cand[kd], (long long)c->kind_instr[kd]);
        for (int r = 0; r < LVB_P2P_MAX_RANKS; r++)
            if (r != c->rank && c->p2p_peer[r])
                cudaIpcCloseMemHandle(c->p2p_peer[r]);
        cudaFree(c->p2p_slab);
        if (c->comm && g_nccl.CommDestroy)
            g_nccl.CommDestroy(c->comm);
        free_matrix(c);
        cudaFree(c->d_cnts);
        cudaFree(c->d_dep);
        cudaFree(c->d_totals);
        cudaFree(c->d_done);
        cudaFree(c->d_lengths);
        cudaFree(c->d_nni);
        cudaFree(c->d_trace);
        if (c->h_nni)
            cudaFreeHost(c->h_nni);
        if (c->h_err)
            cudaFreeHost(c->h_err);
        for (auto &st : c->ring)
        {
            cudaFreeHost(st.h);
            cudaFree(st.d);
            cudaFreeHost(st.h_len);
            cudaFree(st.d_kept);
            cudaFree(st.d_newsum);
            if (st.h_which)
                cudaFreeHost(st.h_which);
            if (st.done)
                cudaEventDestroy(st.done);
        }
        for (int i = 0; i < lvb_gpu_ctx::kEvRing; i++)
        {
            cudaEventDestroy(c->ring0[i]);
            cudaEventDestroy(c->ring1[i]);
        }
        cudaStreamDestroy(c->stream);
        delete c;
        return 0;
    }

    void *lvb_gpu_stream(lvb_gpu_ctx *c) { return c ? (void *)c->stream : nullptr; }

    int lvb_gpu_matrix_upload(lvb_gpu_ctx *c, int64_t n, int64_t nwords, const uint64_t *const *enc_rows,
                              int64_t word_begin, int64_t word_count, int64_t max_trees)
    {
        if (!c || !enc_rows)
            return fail("lvb_gpu_matrix_upload: null argument");
        if (alloc_matrix(c, n, nwords, word_begin, word_count, max_trees))
            return 1;
        return upload_rows(c, [&](int64_t i) { return enc_rows[i]; });
    }

    int lvb_gpu_matrix_upload_flat(lvb_gpu_ctx *c, int64_t n, int64_t nwords, const uint64_t *enc, int64_t word_begin,
                                   int64_t word_count, int64_t max_trees)
    {
        if (!c || !enc)
            return fail("lvb_gpu_matrix_upload_flat: null argument");
        if (alloc_matrix(c, n, nwords, word_begin, word_count, max_trees))
            return 1;
        return upload_rows(c, [&](int64_t i) { return enc + i * nwords; });
    }

    int lvb_gpu_matrix_pack(lvb_gpu_ctx *c, int64_t n, int64_t m, const char *const *rows, int64_t word_begin,
                            int64_t word_count, int64_t max_trees)
    {
        if (!c || !rows)
            return fail("lvb_gpu_matrix_pack: null argument");
        if (m < 1)
            return fail("alignment needs at least 1 column (MIN_M, LVB.h:103)");
        const int64_t nwords = (m + 15) / 16;
        if (alloc_matrix(c, n, nwords, word_begin, word_count, max_trees))
            return 1;
        /* only the characters of this context's word block travel */
        const int64_t site0 = word_begin * 16;
        const int64_t site1 = std::min<int64_t>(m, (word_begin + word_count) * 16);
        const int64_t mloc = site1 - site0;
        unsigned char *d_text = nullptr;
        int *d_bad = nullptr;
        CU(cudaMalloc(&d_text, (size_t)n * mloc));
        CU(cudaMalloc(&d_bad, sizeof(int)));
        CU(cudaMemsetAsync(d_bad, 0, sizeof(int), c->stream));
        /* row by row straight from the caller's (pageable) memory: the driver stages each copy itself;
         * a pinned staging buffer of our own costs more to allocate than it saves (measured) */
        int rc = 0;
        for (int64_t r = 0; r < n && !rc; r++)
        {
            const cudaError_t e = cudaMemcpyAsync(d_text + r * mloc, rows[r] + site0, (size_t)mloc, cudaMemcpyHostToDevice,
                                                  c->stream);
            if (e != cudaSuccess)
                rc = fail("text upload failed: %s", cudaGetErrorString(e));
        }
        int bad = 0;
        if (!rc)
        {
            /* the kernel sees a local alignment of mloc columns starting at word 0 */
            cudaError_t e = launch_pack_dna(d_text, n, mloc, 0, word_count, c->pool, c->stride_u4, d_bad, c->stream);
            c->launches++;
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess)
                e = cudaStreamSynchronize(c->stream);
            if (e != cudaSuccess)
                rc = fail("pack kernel failed: %s", cudaGetErrorString(e));
        }
        cudaFree(d_text);
        cudaFree(d_bad);
        if (!rc && bad)
            rc = fail("bad base symbol in data MSA: %c (cannot read data MSA)", (char)(bad - 1));
        return rc;
    }

    int lvb_gpu_matrix_clone(lvb_gpu_ctx *c, lvb_gpu_ctx *src, int64_t max_trees)
    {
        if (!c || !src || c == src)
            return fail("lvb_gpu_matrix_clone: bad argument");
        if (!src->pool)
            return fail("lvb_gpu_matrix_clone: the source context holds no alignment");
        if (alloc_matrix(c, src->n, src->nwords, src->word_begin, src->word_count, max_trees))
            return 1;
        /* leaf rows device to device (peer copy when the contexts sit on different GPUs) */
        CU(cudaSetDevice(src->device));
        CU(cudaStreamSynchronize(src->stream));
        CU(cudaSetDevice(c->device));
        const size_t bytes = (size_t)src->n * (size_t)src->stride_u4 * sizeof(uint4);
        if (c->device == src->device)
            CU(cudaMemcpyAsync(c->pool, src->pool, bytes, cudaMemcpyDeviceToDevice, c->stream));
        else
            CU(cudaMemcpyPeerAsync(c->pool, c->device, src->pool, src->device, bytes, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        return 0;
    }

    int lvb_gpu_column_stats(lvb_gpu_ctx *c, int64_t n, int64_t m, const char *const *rows, unsigned char *keep,
                             int64_t *n_variable, int64_t *min_len_all, int64_t *min_len_variable)
    {
        if (!c || !rows)
            return fail("lvb_gpu_column_stats: null argument");
        if (n < 1 || m < 1)
            return fail("lvb_gpu_column_stats: empty alignment");
        CU(cudaSetDevice(c->device));
        unsigned char *d_text = nullptr, *h_text = nullptr, *d_keep = nullptr;
        long long *d_sums = nullptr;
        long long sums[3] = {0, 0, 0};
        int rc = 0;
        auto check = [&](cudaError_t e, const char *what) {
            if (e != cudaSuccess && !rc)
                rc = fail("%s: %s", what, cudaGetErrorString(e));
            return e == cudaSuccess;
        };
        const size_t chunk_rows = std::max<size_t>(1, std::min<size_t>((size_t)n, (64u << 20) / (size_t)m));
        if (check(cudaMalloc(&d_text, (size_t)n * m), "text allocation") &&
            check(cudaMalloc(&d_keep, (size_t)m), "keep allocation") &&
            check(cudaMalloc(&d_sums, sizeof sums), "sums allocation") &&
            check(cudaMallocHost(&h_text, chunk_rows * (size_t)m), "staging allocation") &&
            check(cudaMemsetAsync(d_sums, 0, sizeof sums, c->stream), "memset"))
        {
            for (int64_t r0 = 0; r0 < n && !rc; r0 += (int64_t)chunk_rows)
            {
                const int64_t nr = std::min<int64_t>((int64_t)chunk_rows, n - r0);
                for (int64_t r = 0; r < nr; r++)
                    memcpy(h_text + r * m, rows[r0 + r], (size_t)m);
                if (check(cudaMemcpyAsync(d_text + r0 * m, h_text, (size_t)nr * m, cudaMemcpyHostToDevice, c->stream),
                          "text upload"))
                    check(cudaStreamSynchronize(c->stream), "text upload");
            }
            if (!rc && check(launch_column_stats(d_text, n, m, d_keep, d_sums, c->stream), "column_stats kernel"))
            {
                c->launches++;
                if (keep)
                    check(cudaMemcpyAsync(keep, d_keep, (size_t)m, cudaMemcpyDeviceToHost, c->stream), "keep download");
                check(cudaMemcpyAsync(sums, d_sums, sizeof sums, cudaMemcpyDeviceToHost, c->stream), "sums download");
                check(cudaStreamSynchronize(c->stream), "column_stats kernel");
            }
        }
        cudaFree(d_text);
        cudaFree(d_keep);
        cudaFree(d_sums);
        if (h_text)
            cudaFreeHost(h_text);
        if (rc)
            return rc;
        if (n_variable)
            *n_variable = (int64_t)sums[0];
        if (min_len_all)
            *min_len_all = (int64_t)sums[1];
        if (min_len_variable)
            *min_len_variable = (int64_t)sums[2];
        return 0;
    }

    /* no counterpart in the reference: added (callers that cannot allocate) */
    int lvb_gpu_column_stats_noalloc(lvb_gpu_ctx *c, int64_t n, int64_t m, const unsigned char *d_text,
                                     unsigned char *d_keep, int64_t *d_out3)
    {
        if (!c || !d_text || !d_out3)
            return fail("lvb_gpu_column_stats_noalloc: null argument");
        if (n < 1 || m < 1)
            return fail("lvb_gpu_column_stats_noalloc: empty alignment");
        if (m > (int64_t)LVB_NOALLOC_MAX_COLUMNS)
            return fail("lvb_gpu_column_stats_noalloc: %lld columns, the fixed scratch array holds the partial sums of at "
                        "most %lld (LVB_GPU_NOALLOC_MAX_COLUMNS); use lvb_gpu_column_stats", (long long)m,
                        (long long)LVB_NOALLOC_MAX_COLUMNS);
        CU(cudaSetDevice(c->device));
        static_assert(sizeof(long long) == sizeof(int64_t), "out3 is written as long long");
        CU(launch_column_stats_noalloc(d_text, n, m, d_keep, reinterpret_cast<long long *>(d_out3), c->stream));
        c->launches++;
        return 0;
    }

    int lvb_gpu_tree_create(lvb_gpu_ctx *c, int *tree)
    {
        if (!c || !tree)
            return fail("lvb_gpu_tree_create: null argument");
        if (!c->pool)
            return fail("no alignment on the device: call lvb_gpu_matrix_upload first");
        for (int64_t t = 0; t < c->max_trees; t++)
        {
            if (!c->tree_used[t])
            {
                c->tree_used[t] = 1;
                c->tree_ready[t] = 0;
                std::fill(c->bank.begin() + t * c->I, c->bank.begin() + (t + 1) * c->I, 0);
                CU(cudaSetDevice(c->device));
                CU(cudaMemsetAsync(c->chg + t * c->I, 0, (size_t)c->I * sizeof(int32_t), c->stream));
                CU(cudaMemsetAsync(c->sum_chg + t, 0, sizeof(long long), c->stream));
                *tree = (int)t;
                return 0;
            }
        }
        return fail("all %lld device trees are in use (max_trees of lvb_gpu_matrix_upload)", (long long)c->max_trees);
    }

    int lvb_gpu_tree_destroy(lvb_gpu_ctx *c, int tree)
    {
        if (!c || tree < 0 || tree >= c->max_trees || !c->tree_used[tree])
            return fail("lvb_gpu_tree_destroy: tree %d does not exist", tree);
        c->tree_used[tree] = 0;
        c->tree_ready[tree] = 0;
        return 0;
    }

    int lvb_gpu_tree_copy(lvb_gpu_ctx *c, int dst, int src)
    {
        if (!c || dst < 0 || src < 0 || dst >= c->max_trees || src >= c->max_trees || !c->tree_used[dst] ||
            !c->tree_used[src])
            return fail("lvb_gpu_tree_copy: bad tree %d <- %d", dst, src);
        if (dst == src)
            return 0;
        CU(cudaSetDevice(c->device));
        /* both banks as they are, and which bank holds what */
        const size_t bytes = (size_t)2 * c->I * c->stride_u4 * sizeof(uint4);
        CU(cudaMemcpyAsync(c->pool + (c->n + (int64_t)dst * 2 * c->I) * c->stride_u4,
                           c->pool + (c->n + (int64_t)src * 2 * c->I) * c->stride_u4, bytes, cudaMemcpyDeviceToDevice,
                           c->stream));
        std::copy(c->bank.begin() + (int64_t)src * c->I, c->bank.begin() + (int64_t)(src + 1) * c->I,
                  c->bank.begin() + (int64_t)dst * c->I);
        CU(cudaMemcpyAsync(c->chg + dst * c->I, c->chg + src * c->I, (size_t)c->I * sizeof(int32_t),
                           cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaMemcpyAsync(c->sum_chg + dst, c->sum_chg + src, sizeof(long long), cudaMemcpyDeviceToDevice,
                           c->stream));
        c->tree_ready[dst] = c->tree_ready[src];
        return 0;
    }

    int lvb_gpu_tree_commit(lvb_gpu_ctx *c, int tree, const lvb_gpu_topology *topo, int64_t *length)
    {
        if (!c || !topo)
            return fail("lvb_gpu_tree_commit: null argument");
        return run_batch(c, 1, &tree, topo, true, length);
    }

    int lvb_gpu_commit_batch(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                             int64_t *lengths)
    {
        if (!c || !trees || !topos)
            return fail("lvb_gpu_commit_batch: null argument");
        return run_batch(c, n_cand, trees, topos, true, lengths);
    }

    int lvb_gpu_score_batch(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                            int64_t *lengths)
    {
        if (!c || !trees || !topos || !lengths)
            return fail("lvb_gpu_score_batch: null argument");
        return run_batch(c, n_cand, trees, topos, false, lengths);
    }

    int lvb_gpu_score_batch_submit(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                   int *ticket)
    {
        if (!c || !trees || !topos || !ticket)
            return fail("lvb_gpu_score_batch_submit: null argument");
        if (n_cand <= 0)
            return fail("lvb_gpu_score_batch_submit: empty batch");
        return run_batch(c, n_cand, trees, topos, false, nullptr, ticket);
    }

    int lvb_gpu_commit_batch_submit(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                    int *ticket)
    {
        if (!c || !trees || !topos || !ticket)
            return fail("lvb_gpu_commit_batch_submit: null argument");
        if (n_cand <= 0)
            return fail("lvb_gpu_commit_batch_submit: empty batch");
        return run_batch(c, n_cand, trees, topos, true, nullptr, ticket);
    }

    int lvb_gpu_batch_wait(lvb_gpu_ctx *c, int ticket, int64_t *lengths)
    {
        if (!c || !lengths || ticket < 0 || ticket >= lvb_gpu_ctx::kRing)
            return fail("lvb_gpu_batch_wait: bad argument");
        Staging &st = c->ring[ticket];
        if (!st.unclaimed)
            return fail("lvb_gpu_batch_wait: ticket %d names no submitted batch", ticket);
        CU(cudaSetDevice(c->device));
        const double t0 = c->profile ? wall_now() : 0;
        CU(cudaEventSynchronize(st.done));
        if (c->profile)
            c->t_wait += wall_now() - t0;
        st.pending = false;
        if (check_device_error(c))
        {
            st.unclaimed = 0;
            return 1;
        }
        for (int64_t k = 0; k < st.unclaimed; k++)
            lengths[k] = (int64_t)st.h_len[k];
        st.unclaimed = 0;
        return 0;
    }

    int lvb_gpu_nni_neighbourhood(lvb_gpu_ctx *c, int tree, int scratch_tree, const lvb_gpu_topology *topo,
                                  int64_t *lengths)
    {
        if (!c || !topo || !lengths)
            return fail("lvb_gpu_nni_neighbourhood: null argument");
        if (!c->pool)
            return fail("no alignment on the device: call lvb_gpu_matrix_upload first");
        if (tree < 0 || tree >= c->max_trees || !c->tree_used[tree] || !c->tree_ready[tree])
            return fail("lvb_gpu_nni_neighbourhood: tree %d has not been evaluated", tree);
        if (scratch_tree < 0 || scratch_tree >= c->max_trees || !c->tree_used[scratch_tree] || scratch_tree == tree)
            return fail("lvb_gpu_nni_neighbourhood: scratch tree %d does not exist or is the tree itself", scratch_tree);
        if (c->I < 1)
            return fail("lvb_gpu_nni_neighbourhood: the tree has no internal branch");
        /* length of the current tree as getplen returns it: cached counts + root step */
        lvb_gpu_topology clean = *topo;
        clean.n_dirty = 0;
        clean.dirty = nullptr;
        int64_t base_len = 0;
        if (run_batch(c, 1, &tree, &clean, false, &base_len))
            return 1;
        c->last_K = 0;

        std::vector<FitchInstr> instr;
        std::vector<int32_t> parent, left, right;
        std::string err;
        int n_slots = 0;
        /* like a large commit, the up-pass is cut so that enough warps are resident: a top piece, then
         * every cut-off subtree as a program of its own in a second launch */
        std::vector<ScheduleCompiler::Piece> pcs;
        int piece_size = 0;
        if (c->I >= 64 && c->up_pieces)
        {
            const int64_t warps_per_piece = (c->stride_u4 + 31) / 32;
            const int64_t want = std::max<int64_t>(1, c->target_warps / warps_per_piece);
            if (want > 1)
                piece_size = (int)std::max<int64_t>(16, (c->I + want - 1) / want);
        }
        if (!c->sc[0]->compile_up_pass(*topo, instr, n_slots, parent, left, right, err, piece_size, &pcs))
            return fail("%s", err.c_str());
        const size_t n_pcs = pcs.size();
        const int64_t E = c->I;
        const size_t n_instr = instr.size();
        if (ensure_device_arrays(c, n_instr, 1))
            return 1;
        if ((size_t)(2 * E) > c->cap_nni)
        {
            CU(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_nni);
            if (c->h_nni)
                CU(cudaFreeHost(c->h_nni));
            c->cap_nni = grow(c->cap_nni, (size_t)(2 * E), 1024);
            CU(cudaMalloc(&c->d_nni, c->cap_nni * sizeof(long long)));
            CU(cudaMallocHost(&c->h_nni, c->cap_nni * sizeof(long long)));
        }
        const size_t bytes = (n_pcs + 1) * sizeof(FitchProgram) + n_instr * sizeof(FitchInstr) +
                             (size_t)E * (sizeof(uint4) + sizeof(uint2));
        int slot;
        if (acquire_slot(c, bytes, 1, &slot))
            return 1;
        Staging &st = c->ring[slot];
        FitchProgram *h_prog = reinterpret_cast<FitchProgram *>(st.h);
        FitchInstr *h_instr = reinterpret_cast<FitchInstr *>(h_prog + n_pcs + 1);
        uint4 *h_rows = reinterpret_cast<uint4 *>(h_instr + n_instr);
        uint2 *h_rows2 = reinterpret_cast<uint2 *>(h_rows + E);
        /* up-pass: operands from `tree`, stores (and the per-node counts nobody reads) into scratch_tree;
         * pieces are in level order (the top piece first), the last record covers the whole program
         * for the finalize step */
        size_t n_level0 = 0;
        for (size_t i = 0; i < n_pcs; i++)
        {
            h_prog[i] = FitchProgram{pcs[i].instr_begin, pcs[i].n_instr, scratch_tree, 0, -1, 0, tree + 1, 0};
            n_level0 += pcs[i].level == 0 ? 1 : 0;
        }
        h_prog[n_pcs] = FitchProgram{0, (int32_t)n_instr, scratch_tree, 1, -1, 0, tree + 1, 0};
        if (n_instr)
            memcpy(h_instr, instr.data(), n_instr * sizeof(FitchInstr));
        /* the scratch tree takes the up sets in its first bank; operands: rows of `tree` where they are now, and
         * (numbered from B up by the compiler) rows of the scratch tree */
        std::fill(c->bank.begin() + (int64_t)scratch_tree * c->I, c->bank.begin() + (int64_t)(scratch_tree + 1) * c->I, 0);
        auto low_row = [&](int32_t v) { return (uint32_t)(v < c->n ? v : pool_row(c, tree, v)); };
        auto up_row = [&](int32_t v) { return (uint32_t)(c->n + (int64_t)scratch_tree * 2 * c->I + (v - c->n)); };
        for (size_t i = 0; i < n_instr; i++)
        {
            FitchInstr &in = h_instr[i];
            int32_t *ops[2] = {&in.a, &in.b};
            for (int32_t *op : ops)
                if (*op >= c->B)
                    *op = (int32_t)up_row((int32_t)(*op - c->B + c->n));
                else if (*op >= c->n)
                    *op = (int32_t)low_row(*op);
        }
        st.keep_batch = false;
        const int64_t root = topo->root;
        for (int64_t e = 0; e < E; e++)
        {
            const int32_t u = (int32_t)(c->n + e), v = parent[u];
            const int32_t sis = left[v] == u ? right[v] : left[v];
            const uint32_t up = v == root ? (uint32_t)root : up_row(v);
            h_rows[e] = make_uint4(low_row(left[u]), low_row(right[u]), low_row(sis), up);
            h_rows2[e] = make_uint2(low_row(u), up_row(u));
        }
        CU(cudaMemcpyAsync(st.d, st.h, bytes, cudaMemcpyHostToDevice, c->stream));
        c->tree_ready[scratch_tree] = 0; /* its rows now hold up sets, not an evaluated tree */
        const FitchProgram *d_prog = reinterpret_cast<const FitchProgram *>(st.d);
        const FitchInstr *d_instr = reinterpret_cast<const FitchInstr *>(d_prog + n_pcs + 1);
        const uint4 *d_rows = reinterpret_cast<const uint4 *>(d_instr + n_instr);
        const int block = pick_block(c);
        const int n_cb = (int)((c->stride_u4 + block - 1) / block);
        if (n_instr)
        {
            ExecParams p;
            p.pool = c->pool;
            p.stride_u4 = c->stride_u4;
            p.ncols = (int32_t)c->stride_u4;
            p.n = (int32_t)c->n;
            p.n_internal = (int32_t)c->I;
            p.tree_stride = (int32_t)(2 * c->I);
            p.peek = 0;
            p.kept = nullptr;
            p.newsum = nullptr;
            p.n_colblocks = n_cb;
            p.n_slots = n_slots;
            p.score_v1 = 0;
            p.instr = d_instr;
            p.totals = c->d_totals;
            p.done = c->d_done;
            p.out = c->d_lengths;
            p.cnts = c->d_cnts;
            p.ticket = nullptr;
            p.dep = nullptr;
            p.chg = c->chg;
            p.sum_chg = c->sum_chg;
            p.chg_rw = c->chg;
            p.sum_chg_rw = c->sum_chg;
            p.fused_finalize = 0;
            p.cand = nullptr;
            p.n_cand = 0;
            p.pipe_pf = 0;
            p.stream_nt = 0;
            p.fuse_p2p = 0;
            p.trace = nullptr;
            p.err = c->d_err;
            p.spin_limit = c->spin_limit;
            const size_t level_first[2] = {0, n_level0}, level_count[2] = {n_level0, n_pcs - n_level0};
            for (int lv = 0; lv < 2; lv++)
            {
                if (!level_count[lv])
                    continue;
                p.prog = d_prog + level_first[lv];
                p.n_prog = (int32_t)level_count[lv];
                if ((int64_t)p.n_prog * n_cb > (int64_t)INT32_MAX)
                    return fail("lvb_gpu_nni_neighbourhood: %lld CTAs", (long long)p.n_prog * n_cb);
                CU(launch_fitch_exec(p, true, block, c->stream));
                c->launches++;
            }
            CU(launch_commit_finalize(d_instr, d_prog + n_pcs, 1, c->d_cnts, c->chg, c->sum_chg, (int)c->n, (int)c->I,
                                      c->d_lengths, (int)n_instr, c->stream));
            c->launches++;
        }
        CU(cudaMemsetAsync(c->d_nni, 0, (size_t)(2 * E) * sizeof(long long), c->stream));
        NniParams q;
        q.pool = c->pool;
        q.stride_u4 = c->stride_u4;
        q.ncols = (int32_t)c->stride_u4;
        q.n_edges = (int32_t)E;
        q.n_colblocks = n_cb;
        q.rows = d_rows;
        q.rows2 = reinterpret_cast<const uint2 *>(d_rows + E);
        q.chg_low = c->chg + (int64_t)tree * c->I;
        q.chg_up = c->chg + (int64_t)scratch_tree * c->I;
        q.delta = c->d_nni;
        if ((int64_t)E * n_cb > (int64_t)INT32_MAX)
            return fail("lvb_gpu_nni_neighbourhood: %lld CTAs", (long long)E * n_cb);
        int ev = -1;
        if (c->timing)
        {
            if (c->ring_used == lvb_gpu_ctx::kEvRing && drain_kernel_events(c))
                return 1;
            ev = c->ring_used++;
            c->ring_kind[ev] = 0;
            c->ring_cand[ev] = 2 * E;
            c->ring_instr[ev] = 0;
            CU(cudaEventRecord(c->ring0[ev], c->stream));
        }
        CU(launch_nni_neighbourhood(q, block, c->stream));
        c->launches++;
        if (c->timing)
        {
            CU(cudaEventRecord(c->ring1[ev], c->stream));
            c->ev0 = c->ring0[ev];
            c->ev1 = c->ring1[ev];
            c->timed = true;
        }
        if (c->comm && allreduce_i64(c, c->d_nni, 2 * E, nullptr, nullptr))
            return 1;
        CU(cudaMemcpyAsync(c->h_nni, c->d_nni, (size_t)(2 * E) * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(st.done, c->stream));
        CU(cudaEventSynchronize(st.done));
        if (check_device_error(c))
            return 1;
        for (int64_t k = 0; k < 2 * E; k++)
            lengths[k] = base_len + (int64_t)c->h_nni[k];
        return 0;
    }

    int lvb_gpu_download_row(lvb_gpu_ctx *c, int tree, int64_t node, uint64_t *out)
    {
        if (!c || !out || !c->pool)
            return fail("lvb_gpu_download_row: null argument or no alignment");
        if (node < 0 || node >= c->B)
            return fail("lvb_gpu_download_row: node %lld out of range", (long long)node);
        int64_t row = node;
        if (node >= c->n)
        {
            if (tree < 0 || tree >= c->max_trees || !c->tree_used[tree])
                return fail("lvb_gpu_download_row: tree %d does not exist", tree);
            row = pool_row(c, tree, node);
        }
        CU(cudaSetDevice(c->device));
        CU(cudaMemcpyAsync(out, c->pool + row * c->stride_u4, (size_t)c->word_count * 8, cudaMemcpyDeviceToHost,
                           c->stream));
        CU(cudaStreamSynchronize(c->stream));
        return 0;
    }

    int lvb_gpu_tree_changes(lvb_gpu_ctx *c, int tree, int64_t *changes)
    {
        if (!c || !changes || tree < 0 || tree >= c->max_trees || !c->tree_used[tree])
            return fail("lvb_gpu_tree_changes: bad argument");
        CU(cudaSetDevice(c->device));
        std::vector<int32_t> h((size_t)c->I);
        const int32_t *src = c->chg + (int64_t)tree * c->I;
        int32_t *tmp = nullptr;
        if (c->comm)
        {
            CU(cudaMalloc(&tmp, (size_t)c->I * sizeof(int32_t)));
            int r = g_nccl.AllReduce(src, tmp, (size_t)c->I, kNcclInt32, kNcclSum, c->comm, c->stream);
            if (r != 0)
            {
                cudaFree(tmp);
                return fail("ncclAllReduce failed");
            }
            src = tmp;
        }
        CU(cudaMemcpyAsync(h.data(), src, (size_t)c->I * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        if (tmp)
            cudaFree(tmp);
        for (int64_t i = 0; i < c->n; i++)
            changes[i] = 0;
        for (int64_t i = 0; i < c->I; i++)
            changes[c->n + i] = h[i];
        return 0;
    }

    int lvb_gpu_comm_unique_id(void *id128)
    {
        if (!id128)
            return fail("lvb_gpu_comm_unique_id: null argument");
        if (nccl_load())
            return 1;
        int r = g_nccl.GetUniqueId(id128);
        if (r != 0)
            return fail("ncclGetUniqueId failed: %d", r);
        return 0;
    }

    int lvb_gpu_comm_init(lvb_gpu_ctx *c, int nranks, int rank, const void *id128)
    {
        if (!c || !id128)
            return fail("lvb_gpu_comm_init: null argument");
        if (nccl_load())
            return 1;
        CU(cudaSetDevice(c->device));
        NcclId id;
        memcpy(&id, id128, sizeof id);
        int r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
        if (r != 0)
        {
            c->comm = nullptr;
            return fail("ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
        }
        c->nranks = nranks;
        c->rank = rank;
        /* peer-memory all-reduce: one slab per rank, mapped by every rank through CUDA IPC; the 64-byte
         * handles travel through the communicator that was just made */
        const char *pp = getenv("LVB_GPU_P2P");
        const char *pfz = getenv("LVB_GPU_P2P_FUSE");
        c->p2p_fuse = !(pfz && atoi(pfz) == 0);
        c->p2p = false;
        if (!(pp && atoi(pp) == 0) && nranks <= LVB_P2P_MAX_RANKS && g_nccl.AllGather)
        {
            const size_t slab_bytes = (size_t)LVB_P2P_SLOTS * (size_t)nranks * (size_t)c->p2p_kmax * 16;
            unsigned char *d_h = nullptr;
            std::vector<cudaIpcMemHandle_t> hs((size_t)nranks);
            bool ok = cudaMalloc(&c->p2p_slab, slab_bytes) == cudaSuccess &&
                      cudaMemset(c->p2p_slab, 0, slab_bytes) == cudaSuccess &&
                      cudaMalloc(&d_h, (size_t)nranks * sizeof(cudaIpcMemHandle_t)) == cudaSuccess;
            if (ok && nranks > 1)
            {
                ok = cudaIpcGetMemHandle(&hs[(size_t)rank], c->p2p_slab) == cudaSuccess &&
                     cudaMemcpy(d_h + (size_t)rank * sizeof(cudaIpcMemHandle_t), &hs[(size_t)rank], sizeof(cudaIpcMemHandle_t),
                                cudaMemcpyHostToDevice) == cudaSuccess;
                /* every rank takes part in the gather whether or not its own set-up worked: a handle of
                 * zeros tells the others */
                if (!ok)
                    cudaMemset(d_h + (size_t)rank * sizeof(cudaIpcMemHandle_t), 0, sizeof(cudaIpcMemHandle_t));
                int gr = g_nccl.AllGather(d_h + (size_t)rank * sizeof(cudaIpcMemHandle_t), d_h, sizeof(cudaIpcMemHandle_t),
                                          kNcclInt8, c->comm, c->stream);
                ok = ok && gr == 0 && cudaStreamSynchronize(c->stream) == cudaSuccess &&
                     cudaMemcpy(hs.data(), d_h, (size_t)nranks * sizeof(cudaIpcMemHandle_t), cudaMemcpyDeviceToHost) == cudaSuccess;
                for (int r = 0; ok && r < nranks; r++)
                {
                    if (r == rank)
                        continue;
                    static const cudaIpcMemHandle_t zero = {};
                    void *ptr = nullptr;
                    if (memcmp(&hs[(size_t)r], &zero, sizeof zero) == 0 ||
                        cudaIpcOpenMemHandle(&ptr, hs[(size_t)r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
                        ok = false;
                    c->p2p_peer[r] = (unsigned long long *)ptr;
                }
                /* all ranks must agree on the route: one more tiny gather of the outcome */
                unsigned char mine = ok ? 1 : 0;
                std::vector<unsigned char> all((size_t)nranks, 0);
                if (cudaMemcpy(d_h + rank, &mine, 1, cudaMemcpyHostToDevice) == cudaSuccess &&
                    g_nccl.AllGather(d_h + rank, d_h, 1, kNcclInt8, c->comm, c->stream) == 0 &&
                    cudaStreamSynchronize(c->stream) == cudaSuccess &&
                    cudaMemcpy(all.data(), d_h, (size_t)nranks, cudaMemcpyDeviceToHost) == cudaSuccess)
                {
                    for (int r = 0; r < nranks; r++)
                        ok = ok && all[(size_t)r] == 1;
                }
                else
                    ok = false;
            }
            (void)cudaGetLastError(); /* a failed IPC attempt must not poison later calls */
            cudaFree(d_h);
            if (ok)
            {
                c->p2p_peer[rank] = c->p2p_slab;
                c->p2p = true;
            }
            else
            {
                for (int r = 0; r < nranks; r++)
                    if (r != rank && c->p2p_peer[r])
                    {
                        cudaIpcCloseMemHandle(c->p2p_peer[r]);
                        c->p2p_peer[r] = nullptr;
                    }
                cudaFree(c->p2p_slab);
                c->p2p_slab = nullptr;
                (void)cudaGetLastError();
            }
        }
        if (getenv("LVB_GPU_PROFILE"))
            fprintf(stderr, "[lvb_gpu] rank %d/%d: all-reduce of partial lengths through %s\n", rank, nranks,
                    c->p2p ? "peer memory (CUDA IPC slabs over NVLink)" : "ncclAllReduce");
        return 0;
    }

    int64_t lvb_gpu_launch_count(lvb_gpu_ctx *c) { return c ? c->launches : 0; }

    int lvb_gpu_trace_download(lvb_gpu_ctx *c, uint64_t *out, int64_t cap_items, int64_t *n_items, int64_t *n_colblocks)
    {
        if (!c || !n_items)
            return fail("lvb_gpu_trace_download: null argument");
        *n_items = (int64_t)c->n_trace;
        if (n_colblocks)
            *n_colblocks = c->n_trace && c->last_n_prog ? (int64_t)(c->n_trace / c->last_n_prog) : 0;
        if (!out || !c->n_trace)
            return 0;
        if (cap_items < (int64_t)c->n_trace)
            return fail("lvb_gpu_trace_download: room for %lld items, have %zu", (long long)cap_items, c->n_trace);
        CU(cudaSetDevice(c->device));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaMemcpy(out, c->d_trace, c->n_trace * 6 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        return 0;
    }

    int lvb_gpu_set_timing(lvb_gpu_ctx *c, int on)
    {
        if (!c)
            return fail("lvb_gpu_set_timing: null argument");
        c->timing = on != 0;
        return 0;
    }

    int lvb_gpu_last_kernel_ms(lvb_gpu_ctx *c, float *ms)
    {
        if (!c || !ms)
            return fail("lvb_gpu_last_kernel_ms: null argument");
        if (!c->timed)
            return fail("no Fitch kernel has been timed yet (lvb_gpu_set_timing)");
        CU(cudaEventSynchronize(c->ev1));
        CU(cudaEventElapsedTime(ms, c->ev0, c->ev1));
        return 0;
    }

    int lvb_gpu_compile_pieces(int64_t n, const lvb_gpu_topology *topo, int piece_size, int32_t *instr, int64_t cap,
                               int64_t *n_instr, int64_t *n_slots, int32_t *pieces, int64_t pieces_cap,
                               int64_t *n_pieces)
    {
        if (!topo || !n_instr || !n_slots || !n_pieces)
            return fail("lvb_gpu_compile_pieces: null argument");
        if (n < 3)
            return fail("lvb_gpu_compile_pieces: n too small");
        ScheduleCompiler sc(n);
        std::vector<FitchInstr> out;
        std::vector<ScheduleCompiler::Piece> pcs;
        int slots = 0;
        std::string err;
        if (!sc.compile(*topo, true, out, slots, err, piece_size, &pcs))
            return fail("%s", err.c_str());
        *n_instr = (int64_t)out.size();
        *n_slots = slots;
        *n_pieces = (int64_t)pcs.size();
        if (instr)
        {
            if ((int64_t)out.size() > cap || (int64_t)pcs.size() > pieces_cap || !pieces)
                return fail("buffers too small: need %zu instructions, %zu pieces", out.size(), pcs.size());
            memcpy(instr, out.data(), out.size() * sizeof(FitchInstr));
            memcpy(pieces, pcs.data(), pcs.size() * sizeof(ScheduleCompiler::Piece));
        }
        return 0;
    }

    int lvb_gpu_rerun_last_batch(lvb_gpu_ctx *c, int64_t *lengths)
    {
        if (!c)
            return fail("lvb_gpu_rerun_last_batch: null argument");
        if (c->last_K <= 0)
            return fail("no batch is resident on the device");
        CU(cudaSetDevice(c->device));
        /* nothing is re-staged, so the launches of consecutive re-runs queue up behind each other on
         * the stream without a host wait.  With a communicator attached the partial lengths are
         * all-reduced on every re-run (the site-block step of BASELINE.json configs[3]). */
        if (c->last_inline)
        { /* the batch went out inside kernel parameters: stage it now, re-runs read device memory */
            Staging &st = c->ring[c->last_slot];
            const size_t bytes = (c->last_n_prog + (size_t)c->last_K) * sizeof(FitchProgram) + c->last_n_instr * sizeof(FitchInstr);
            CU(cudaMemcpyAsync(st.d, st.h, bytes, cudaMemcpyHostToDevice, c->stream));
            c->last_inline = false;
        }
        return launch_resident(c, lengths, true);
    }

    int lvb_gpu_kernel_time(lvb_gpu_ctx *c, double *total_ms, int64_t *launches, int reset)
    {
        if (!c || !total_ms || !launches)
            return fail("lvb_gpu_kernel_time: null argument");
        CU(cudaSetDevice(c->device));
        if (drain_kernel_events(c))
            return 1;
        *total_ms = c->kernel_ms_total;
        *launches = c->kernel_launches_timed;
        if (reset)
        {
            c->kernel_ms_total = 0;
            c->kernel_launches_timed = 0;
        }
        return 0;
    }

    int lvb_gpu_synchronize(lvb_gpu_ctx *c)
    {
        if (!c)
            return fail("lvb_gpu_synchronize: null argument");
        CU(cudaSetDevice(c->device));
        CU(cudaStreamSynchronize(c->stream));
        return check_device_error(c);
    }

    int lvb_gpu_compile_schedule(int64_t n, const lvb_gpu_topology *topo, int store, int32_t *instr, int64_t cap,
                                 int64_t *n_instr, int64_t *n_slots)
    {
        if (!topo || !n_instr || !n_slots)
            return fail("lvb_gpu_compile_schedule: null argument");
        if (n < 3)
            return fail("lvb_gpu_compile_schedule: n too small");
        /* one compiler per calling thread, kept between calls: O(dirty nodes) per call, no allocation */
        static thread_local ScheduleCompiler *sc = nullptr;
        static thread_local std::vector<FitchInstr> out;
        if (!sc || sc->n() != n)
        {
            delete sc;
            sc = new ScheduleCompiler(n);
        }
        out.clear();
        int slots = 0;
        std::string err;
        if (!sc->compile(*topo, store != 0, out, slots, err))
            return fail("%s", err.c_str());
        *n_instr = (int64_t)out.size();
        *n_slots = slots;
        if (instr)
        {
            if ((int64_t)out.size() > cap)
                return fail("instruction buffer too small: need %zu", out.size());
            memcpy(instr, out.data(), out.size() * sizeof(FitchInstr));
        }
        return 0;
    }

    int lvb_gpu_commit_batch_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                 const lvb_gpu_program_view *pre, int64_t *lengths)
    {
        if (!c || !trees || !topos || !pre)
            return fail("lvb_gpu_commit_batch_pre: null argument");
        return run_batch(c, n_cand, trees, topos, true, lengths, nullptr, pre);
    }

    int lvb_gpu_score_batch_submit_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                       const lvb_gpu_program_view *pre, int *ticket)
    {
        if (!c || !trees || !topos || !pre || !ticket)
            return fail("lvb_gpu_score_batch_submit_pre: null argument");
        if (n_cand <= 0)
            return fail("lvb_gpu_score_batch_submit_pre: empty batch");
        return run_batch(c, n_cand, trees, topos, false, nullptr, ticket, pre);
    }

    int lvb_gpu_score_keep_submit_pre(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                                      const lvb_gpu_program_view *pre, const uint8_t *keep, int *ticket)
    {
        if (!c || !trees || !topos || !keep || !ticket)
            return fail("lvb_gpu_score_keep_submit_pre: null argument");
        if (n_cand <= 0)
            return fail("lvb_gpu_score_keep_submit_pre: empty batch");
        return run_batch(c, n_cand, trees, topos, false, nullptr, ticket, pre, keep);
    }

    int lvb_gpu_score_keep(lvb_gpu_ctx *c, int64_t n_cand, const int *trees, const lvb_gpu_topology *topos,
                           const uint8_t *keep, int64_t *lengths, int *ticket)
    {
        if (!c || !trees || !topos || !keep || !lengths || !ticket)
            return fail("lvb_gpu_score_keep: null argument");
        if (n_cand <= 0)
            return fail("lvb_gpu_score_keep: empty batch");
        if (run_batch(c, n_cand, trees, topos, false, nullptr, ticket, nullptr, keep))
            return 1;
        return lvb_gpu_batch_wait(c, *ticket, lengths);
    }

    int lvb_gpu_accept_kept(lvb_gpu_ctx *c, int ticket, int64_t n_accept, const int64_t *cands)
    {
        if (!c || ticket < 0 || ticket >= lvb_gpu_ctx::kRing || (n_accept > 0 && !cands))
            return fail("lvb_gpu_accept_kept: bad argument");
        if (n_accept <= 0)
            return 0;
        Staging &st = c->ring[ticket];
        if (!st.keep_batch)
            return fail("lvb_gpu_accept_kept: ticket %d does not name a batch of lvb_gpu_score_keep_* that is still held "
                        "(at most %d device calls may lie between the submit and the accept)", ticket, lvb_gpu_ctx::kRing - 1);
        if (st.unclaimed)
            return fail("lvb_gpu_accept_kept: fetch the lengths of the batch first (lvb_gpu_batch_wait)");
        const FitchProgram *h_cand = reinterpret_cast<const FitchProgram *>(st.h) + st.k_prog;
        const FitchInstr *h_instr = reinterpret_cast<const FitchInstr *>(h_cand + st.k_cand);
        std::vector<uint8_t> seen(c->max_trees, 0);
        for (int64_t i = 0; i < n_accept; i++)
        {
            const int64_t k = cands[i];
            if (k < 0 || (size_t)k >= st.k_cand || st.kept_flag[(size_t)k] != 1)
                return fail("lvb_gpu_accept_kept: candidate %lld of the batch was not kept (or was accepted already)", (long long)k);
            if (seen[h_cand[k].tree])
                return fail("lvb_gpu_accept_kept: two candidates of tree %d", h_cand[k].tree);
            seen[h_cand[k].tree] = 1;
        }
        CU(cudaSetDevice(c->device));
        /* a small batch went out inside kernel parameters: so does its accept — one launch, nothing staged */
        const bool inl = !st.staged && inline_batch_fits((int)st.k_prog, (int)st.k_cand, st.k_instr) && st.k_prog == st.k_cand;
        if (!inl)
        {
            if ((size_t)n_accept > st.cap_which)
            {
                const size_t cap = grow(st.cap_which, (size_t)n_accept, 64);
                if (st.h_which)
                    CU(cudaFreeHost(st.h_which));
                st.h_which = nullptr;
                CU(cudaHostAlloc(&st.h_which, cap * sizeof(int32_t), cudaHostAllocMapped));
                CU(cudaHostGetDevicePointer(&st.d_which, st.h_which, 0));
                st.cap_which = cap;
            }
            else if (st.pending)
            { /* an earlier accept of this batch may still read the list */
                CU(cudaEventSynchronize(st.done));
                st.pending = false;
            }
        }
        for (int64_t i = 0; i < n_accept; i++)
        {
            const int64_t k = cands[i];
            if (!inl)
                st.h_which[i] = (int32_t)k;
            st.kept_flag[(size_t)k] = 2;
            /* the rows of the candidate's nodes are in the other bank: from now on they are the tree's */
            uint8_t *bk = c->bank.data() + (size_t)((int64_t)h_cand[k].tree * c->I);
            for (int32_t j = 0; j < h_cand[k].n_instr; j++)
            {
                const int32_t node = h_instr[h_cand[k].instr_begin + j].node;
                if (node >= 0)
                    bk[node - c->n] ^= 1;
            }
        }
        if (inl)
        {
            CU(launch_apply_kept_inline(h_cand, (int)st.k_cand, h_instr, st.k_instr, cands, (int)n_accept, st.d_kept, st.d_newsum,
                                        c->chg, c->sum_chg, (int)c->n, (int)c->I, c->stream));
            c->launches++;
            c->last_K = 0;
            return 0;
        }
        if (!st.staged)
        { /* the batch went out inside kernel parameters: the apply kernel reads it from device memory */
            const size_t bytes = (st.k_prog + st.k_cand) * sizeof(FitchProgram) + st.k_instr * sizeof(FitchInstr);
            CU(cudaMemcpyAsync(st.d, st.h, bytes, cudaMemcpyHostToDevice, c->stream));
            st.staged = true;
        }
        const FitchProgram *d_cand = reinterpret_cast<const FitchProgram *>(st.d) + st.k_prog;
        const FitchInstr *d_instr = reinterpret_cast<const FitchInstr *>(d_cand + st.k_cand);
        CU(launch_apply_kept(d_instr, d_cand, st.d_which, (int)n_accept, st.d_kept, st.d_newsum, c->chg, c->sum_chg, (int)c->n,
                             (int)c->I, c->stream));
        c->launches++;
        /* the slot's host and device staging are read by what was just enqueued */
        CU(cudaEventRecord(st.done, c->stream));
        st.pending = true;
        c->last_K = 0; /* the resident batch no longer describes the trees */
        return 0;
    }

    int lvb_gpu_compile_up_pass(int64_t n, const lvb_gpu_topology *topo, int piece_size, int32_t *instr, int64_t cap,
                                int64_t *n_instr, int64_t *n_slots, int32_t *pieces, int64_t pieces_cap, int64_t *n_pieces)
    {
        if (!topo || !n_instr || !n_slots)
            return fail("lvb_gpu_compile_up_pass: null argument");
        if (n < 4)
            return fail("lvb_gpu_compile_up_pass: n too small");
        ScheduleCompiler sc(n);
        std::vector<FitchInstr> out;
        std::vector<int32_t> parent, left, right;
        int slots = 0;
        std::string err;
        std::vector<ScheduleCompiler::Piece> pcs;
        if (!sc.compile_up_pass(*topo, out, slots, parent, left, right, err, piece_size, &pcs))
            return fail("%s", err.c_str());
        *n_instr = (int64_t)out.size();
        *n_slots = slots;
        if (n_pieces)
            *n_pieces = (int64_t)pcs.size();
        if (pieces)
        {
            if ((int64_t)pcs.size() > pieces_cap)
                return fail("piece buffer too small: need %zu", pcs.size());
            for (size_t i = 0; i < pcs.size(); i++)
            {
                pieces[3 * i] = pcs[i].instr_begin;
                pieces[3 * i + 1] = pcs[i].n_instr;
                pieces[3 * i + 2] = pcs[i].level;
            }
        }
        if (instr)
        {
            if ((int64_t)out.size() > cap)
                return fail("instruction buffer too small: need %zu", out.size());
            memcpy(instr, out.data(), out.size() * sizeof(FitchInstr));
        }
        return 0;
    }

} /* extern "C" */
