/* lvb_host.h — LVB's host side of the tree-length path, re-driven for the B200 engine.
 *
 * The reference keeps all of this in C compiled as C++ (/root/reference/Makefile:74-81); so
 * does this layer.  Names, argument meaning and error behaviour mirror the reference so that
 * its callers read the same (file:line given per item); what differs is where the Fitch state
 * rows live: on the device, behind include/lvb_gpu.h.  A host tree therefore carries ONE
 * 64-bit word per node where the reference carries a whole state row: the word that doubles
 * as LVB's "dirty" flag (sitestate[0] == 0  <=>  dirty, src/TreeOperations.c:49-59).  With
 * that, nodeclear / make_dirty_below / treecopy / lvb_reroot / ss_init keep their protocol
 * unchanged and getplen() turns the flags into an explicit dirty list for the device.
 *
 * Nothing here computes Fitch arithmetic: tree lengths come from liblvbgpu.so only.
 */
#ifndef LVB_B200_HOST_H
#define LVB_B200_HOST_H

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/lvb_gpu.h"

namespace lvbhost
{

/* ---- constants: src/LVB.h:72-150 ---- */
typedef uint64_t Lvb_bit_length;
enum : long
{
    UNSET = -1,
    MIN_N = 5,
    MAX_N = 1000000,
    MIN_M = 1,
    MAX_M = 5000000,
    MAXSTATES = 5,
    REROOT_INTERVAL = 1000,
    STAT_LOG_INTERVAL = 50000,
    MAXACCEPT_SLOW = 5,
    MAXPROPOSE_SLOW = 2000,
    MAXFAIL_SLOW = 40,
};
static const double LVB_EPS = 1E-11;
static const double FROZEN_T = 0.0001;
enum
{
    FORMAT_PHYLIP = 0,
    FORMAT_FASTA = 1,
    FORMAT_NEXUS = 2,
    FORMAT_CLUSTAL = 3
};

/* ---- errors: src/Error.c:49-67.  crash() prints "\nFATAL ERROR: <msg>\n" to stdout, flushes
 * and exits with EXIT_FAILURE; when a handler is installed (library use, tests) it throws
 * LvbError instead so that the caller survives. ---- */
struct LvbError
{
    std::string message;
};
[[noreturn]] void crash(const char *fmt, ...);
void set_crash_throws(bool on);
#define lvb_assert(test) ((void)((test) || (::lvbhost::lvb_assertion_fail(#test, __FILE__, __LINE__), 0)))
[[noreturn]] void lvb_assertion_fail(const char *test, const char *file, int line);

/* ---- tree node: src/LVB.h:121-128, same 40-byte layout so that arrays of it can be handed
 * to lvb_gpu_topology with stride_bytes = sizeof(TREESTACK_TREE_NODES) ---- */
struct TREESTACK_TREE_NODES
{
    long parent;
    long left;
    long right;
    long changes;
    Lvb_bit_length *sitestate; /* one flag word per node on the host; 0 = dirty */
};
static_assert(sizeof(TREESTACK_TREE_NODES) == 40, "layout of LVB.h:121-128");

/* ---- alignment: src/DataStructure.h:63-81 (fields this path uses) ---- */
struct DataStructure
{
    long m = 0;          /* columns after matchange */
    long original_m = 0; /* columns read */
    long n = 0;          /* rows */
    long max_length_seq_name = 0;
    long numberofpossiblebranches = 0; /* 2n-3 */
    long bytes = 0;                    /* bytes of one DEVICE state row: 8 * nwords */
    long nwords = 0;
    long min_len_tree = 0;
    long nsets = 0; /* n-3 */
    long mssz = 0;  /* n-2 */
    std::vector<std::string> row;
    std::vector<std::string> rowtitle;
};
typedef DataStructure *Dataptr;

/* ---- parameters: src/DataStructure.h:86-97 ---- */
struct Parameters
{
    int seed = 0;
    int cooling_schedule = 0;    /* 0 geometric, 1 linear */
    int algorithm_selection = 1; /* 0 SN, 1 SEQ-TNS, 2 PBS */
    int n_file_format = FORMAT_PHYLIP;
    int n_processors_available = 1;
    long verbose = 0;
    std::string file_name_in = "infile";
    std::string file_name_out = "outtree";
    int n_number_max_trees = 0;
    /* additions of this build */
    int device = 0;            /* CUDA device of a single-chain run */
    int speculation = 0;       /* max proposals scored per launch and chain; 0 = adaptive */
    int replicates = 1;        /* independent SA chains (seed, seed+1, ...) */
    int groups = 0;            /* device contexts / host driver threads the chains are spread over; 0 = auto */
    int devices = 1;           /* GPUs (device, device+1, ...) the replicates are spread over */
    long max_rearrangements = 0; /* > 0: stop every chain after this many proposals (benchmarks) */
};

/* ---- random numbers: src/RandomNumberGenerator.c:87-256 (Marsaglia UNI).  The reference keeps
 * the state in file statics; here it is an object so that every SA chain has its own stream,
 * and can be snapshotted for speculative proposal batches. ---- */
struct Uni
{
    double u[98];
    double c;
    int ui, uj;
    bool ready = false;
    void rinit(int ijkl);
    double uni();
    long randpint(long upper);
};

/* ---- alignment operations: src/DataOperations.c ---- */
long words_per_row(long m);                                     /* :446-458 */
long bytes_per_row(long m);                                     /* :460-467 */
long MinimumTreeLength(const Dataptr MSA);                      /* :53-105 */
void matchange(Dataptr MSA, const Parameters &rc);              /* :309-349 (+ constchar, cutcols) */
void DNAToBinary(const Dataptr MSA, Lvb_bit_length **enc_mat);  /* :164-249 */
/* matchange after the column scan: keep[k] != 0 where column k stays, kept = their number,
 * min_len_after = MinimumTreeLength of the cut alignment if known, else UNSET */
void matchange_apply(Dataptr MSA, const Parameters &rc, const unsigned char *keep, long kept, long min_len_after);
/* matchange with both text scans on the device (lvb_gpu_column_stats); same result */
void matchange_device(Dataptr MSA, const Parameters &rc, int device_index);
/* src/Wrapper.c:49-68 + src/CommandLineParser.cpp:49-105 (PHYLIP and FASTA only) */
void phylip_dna_matrin(const char *file_name, int n_file_format, Dataptr MSA);
/* build the DataStructure from rows already in memory (what read_file() fills) */
void matrix_from_rows(Dataptr MSA, long n, long m, const char *const *rows, const char *const *titles);

/* ---- tree operations: src/TreeOperations.c ---- */
TREESTACK_TREE_NODES *treealloc(Dataptr MSA, bool b_with_sitestate);                      /* :914-955 */
void treecopy(Dataptr MSA, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, bool b_with_sitestate); /* :737-773 */
void nodeclear(TREESTACK_TREE_NODES *tree, long brnch);                                   /* :49-59 */
void treeclear(Dataptr MSA, TREESTACK_TREE_NODES *tree);                                  /* :78-86 */
long getsister(const TREESTACK_TREE_NODES *tree, long branch);                            /* :657-676 */
void mutate_deterministic(Dataptr MSA, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root, long p, bool left); /* :105-158 */
void mutate_nni(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root); /* :160-209 */
void mutate_spr(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root); /* :236-335 */
void mutate_tbr(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *dest, const TREESTACK_TREE_NODES *src, long root); /* :337-541 */
/* :576-637; path (optional) receives the internal nodes between the two roots: the only
 * nodes whose subtree, hence state row, differs under the new root */
long lvb_reroot(Dataptr MSA, TREESTACK_TREE_NODES *tree, long oldroot, long newroot, bool b_with_sitestate,
                std::vector<long> *path = nullptr);
long arbreroot(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree, long oldroot, std::vector<long> *path = nullptr); /* :639-655 */
void PullRandomTree(Dataptr MSA, Uni &rng, TREESTACK_TREE_NODES *tree);                  /* :799-810 */
void SwapTrees(TREESTACK_TREE_NODES **t1, long *r1, TREESTACK_TREE_NODES **t2, long *r2); /* :1062-1080 */
void ss_init(Dataptr MSA, TREESTACK_TREE_NODES *tree);                                    /* :1500-1513 (flags only; leaf rows live on the device) */
void lvb_treeprint(Dataptr MSA, FILE *stream, const TREESTACK_TREE_NODES *tree, long root); /* :1144-1220 */
/* ---- proposals without treecopy (additions of this build; see tree.cpp:propose) ---- */
struct Proposal
{
    int op = 0;                                  /* 0 NNI, 1 SPR, 2 TBR */
    std::vector<long> node;                      /* nodes the operator touched */
    std::vector<TREESTACK_TREE_NODES> after;     /* their records in the proposed tree */
    std::vector<int64_t> dirty;                  /* internal nodes to recompute (flag 0 in the proposed tree) */
    std::vector<int64_t> mod_node, mod_left, mod_right; /* child links that differ from the current tree */
};
struct ProposalScratch
{
    std::vector<uint32_t> stamp;
    uint32_t cur = 0;
    std::vector<long> node;
    std::vector<TREESTACK_TREE_NODES> before;
    std::vector<Lvb_bit_length> flag_before;
};
void propose(Dataptr MSA, Uni &rng, int op, TREESTACK_TREE_NODES *current, long root, ProposalScratch &scratch,
             Proposal &out);
void apply_proposal(TREESTACK_TREE_NODES *current, const Proposal &p);
/* internal nodes whose flag word is 0 (the first loop of getplen, src/TreeEvaluation.c:191-202) */
void collect_dirty(const Dataptr MSA, const TREESTACK_TREE_NODES *tree, std::vector<int64_t> &out);
void mark_all_clean(Dataptr MSA, TREESTACK_TREE_NODES *tree);

/* ---- tree stack: src/Treestack.c, src/Hash.cpp.  Same push order and duplicate rule
 * (equal unrooted topology); the comparison key is a canonical text of the topology rooted
 * at taxon 0 instead of the reference's sorted clade sets (same equivalence classes). ---- */
struct TREESTACK_TREES
{
    std::vector<TREESTACK_TREE_NODES> tree; /* without state flags */
    long root = UNSET;
    std::string key;
};
struct TREESTACK
{
    std::vector<TREESTACK_TREES> stack;
    long next = 0;                       /* trees in use (ClearTreestack keeps the allocation) */
    std::vector<std::string> hash_keys;  /* Hash.cpp:56 static hashSet, per stack here */
};
TREESTACK CreateNewTreestack(void);                                                                   /* Treestack.c:163-173 */
long CountTreestack(const TREESTACK &s);                                                              /* :110-114 */
void ClearTreestack(TREESTACK *sp);                                                                   /* :513-518 */
long CompareTreeToTreestack(Dataptr MSA, TREESTACK *sp, const TREESTACK_TREE_NODES *tree, long root); /* :231-306 */
long CompareHashTreeToHashstack(Dataptr MSA, TREESTACK *sp, const TREESTACK_TREE_NODES *tree, long root); /* Hash.cpp:49-90 */
long PullTreefromTreestack(Dataptr MSA, TREESTACK_TREE_NODES *tree, long *root, TREESTACK *sp);       /* :352-374 */
int PrintTreestack(Dataptr MSA, TREESTACK *sp, FILE *outfp);                                          /* :376-414 */
std::string topology_key(Dataptr MSA, const TREESTACK_TREE_NODES *tree, long root);

/* ---- the device binding of one alignment: GetSoln's enc_mat + the trees Anneal allocates
 * (src/Solve.c:512-519, 194-197) ---- */
struct Device
{
    lvb_gpu_ctx *ctx = nullptr;
    Dataptr MSA = nullptr;
    int index = 0;              /* CUDA device */
    int64_t word_count = -1;    /* words of every row this context holds (site-block mode: fewer than nwords) */
    int64_t extra_launches = 0; /* launches of the additional contexts GetSolnBatched opened on this device */
    int64_t launches() const;
};
/* lvb_gpu_init + DNAToBinary + lvb_gpu_matrix_upload; crash() on any device error */
void device_open(Device &dev, int device_index, Dataptr MSA, int max_trees, int64_t word_begin = 0,
                 int64_t word_count = -1);
void device_open_encoded(Device &dev, int device_index, Dataptr MSA, Lvb_bit_length *const *rows, int max_trees,
                         int64_t word_begin = 0, int64_t word_count = -1);
void device_open_clone(Device &dev, int device_index, const Device &src, int max_trees);
void device_close(Device &dev);

/* ---- getplen: src/TreeEvaluation.c:49-270, same signature (the scratch pointers are
 * accepted and ignored).  `tree` is bound to a device tree on first use; its dirty flags are
 * turned into a dirty list, the device recomputes and stores those rows, flags become clean.
 * One launch and one 8-byte read-back per call: the compatible, unbatched route. ---- */
long getplen(Dataptr MSA, TREESTACK_TREE_NODES *tree, const Parameters &rc, long root, long *p_todo_arr,
             long *p_todo_arr_sum_changes, int *p_runs);
void alloc_memory_to_getplen(Dataptr MSA, long **p_todo_arr, long **p_todo_arr_sum_changes, int **p_runs); /* MemoryOperations.c:120-125 */
void free_memory_to_getplen(long **p_todo_arr, long **p_todo_arr_sum_changes, int **p_runs);                /* :127-132 */
/* which device getplen() talks to (set by GetSoln / tests) */
void getplen_bind_device(Device *dev);

/* ---- search: src/Solve.c, src/StartingTemperature.c ---- */
struct TrajectoryRow /* one line of the reference's changeAccepted.tsv (Solve.c:467-468) */
{
    long iter;
    int op; /* 0 NNI, 1 SPR, 2 TBR */
    int accepted;
    long length;
    double t;
    long stack;
};
struct SearchResult
{
    long final_length = 0;
    long rearrangements = 0; /* *current_iter */
    double start_temperature = 0;
    long initial_length = 0;
    long getplen_calls = 0;   /* trees scored (proposals + initial passes) */
    long launches_scored = 0; /* score launches */
    double seconds = 0;
    double prof[4] = {0, 0, 0, 0}; /* seconds in: proposal generation, commit calls, score calls, acceptance replay */
    long rounds = 0;
    int groups = 1;                /* device contexts / driver threads used */
    double setup_seconds = 0;      /* part of `seconds` spent opening the additional contexts */
};
struct Chain; /* one SA replicate: solve.cpp */

/* GetSoln for R independent replicates (chain r uses seed rc.seed + r), all driven in
 * lock-step batches on one device; stacks[r] receives chain r's best trees.  trajectory
 * (optional, chain 0 only) receives the changeAccepted.tsv rows. */
void GetSolnBatched(Dataptr MSA, const Parameters &rc, Device &dev, int replicates, std::vector<TREESTACK> &stacks,
                    std::vector<SearchResult> &results, std::vector<TrajectoryRow> *trajectory, FILE *tsv);

/* src/Solve.c:64-142 (switched off in the reference, Solve.c:571-572): NNI on every internal
 * branch, both sides, in index order, accepting every move that does not lengthen the tree and
 * is a new topology, until a whole sweep accepts nothing; returns the final length, the trees
 * of that length are on the stack.  One lvb_gpu_nni_neighbourhood call per accepted move
 * instead of one getplen per candidate; needs two free device trees. */
long deterministic_hillclimb(Dataptr MSA, Device &dev, TREESTACK *treestack_ptr, const TREESTACK_TREE_NODES *inittree,
                             const Parameters &rcstruct, long root, FILE *lenfp, long *current_iter, bool log_progress,
                             long *neighbourhood_calls = nullptr);

} // namespace lvbhost
#endif
