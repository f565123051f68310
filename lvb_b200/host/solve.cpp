/* solve.cpp — the simulated-annealing search (src/Solve.c: GetSoln, Anneal;
 * src/StartingTemperature.c) re-driven in batches for the device.
 *
 * The reference evaluates one proposal per getplen() call.  Here every launch scores a batch:
 *   - across chains: R independent replicates (seeds seed, seed+1, ...) advance in lock-step, and
 *   - along a chain: up to K proposals are generated speculatively from the SAME current tree,
 *     each assuming that the ones before it are rejected.
 * Speculation is exact, not approximate.  Between two proposals the reference touches the
 * random stream only through (a) the operator choice and the mutate_* draws, which do not
 * depend on the outcome of earlier proposals as long as the current tree is unchanged, and
 * (b) one uni() per WORSENING proposal whether or not it is accepted (src/Solve.c:356-376;
 * src/StartingTemperature.c:147-163).  A rejected proposal is by definition a worsening one,
 * so "the k earlier proposals were rejected" fixes the stream exactly.  The lengths come back,
 * acceptance is replayed in order with the recorded draws, and at the first acceptance the
 * chain resumes from the state snapshot taken at that proposal; the rest of the batch is
 * discarded.  All counters, the cooling schedule, operator probabilities and stop tests are
 * advanced by the same code in speculation and in replay (anneal_after / st_after).
 * Batches end where the reference changes the current tree or the device state must change:
 * re-rooting (every REROOT_INTERVAL proposals; at the start of every temperature step of
 * StartingTemperature), the end of a phase, an acceptance.
 *
 * Tree lengths are never computed here; they come from lvb_gpu_commit_batch /
 * lvb_gpu_score_batch (include/lvb_gpu.h).
 */
#include <cerrno>
#include <cfloat>
#include <chrono>
#include <thread>
#include <cmath>
#include <algorithm>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "lvb_host.h"

namespace lvbhost
{

/* src/MyMaths.c:91-120, 146-170, 203-244: libm with loud failure */
static double exp_wrapper(double x)
{
    const double val = exp(x);
    if (errno == EDOM)
        crash("internal error detected in function exp_wrapper():\ndomain error. x is %g", x);
    else if (val == HUGE_VAL || val == -HUGE_VAL)
        crash("internal error detected in function exp_wrapper():\nrange error. x is %g", x);
    else if (val == 0.0)
        crash("internal error detected in function exp_wrapper():\nunderflow. x is %g", x);
    return val;
}
static double log_wrapper(double x)
{
    if (x <= 0.0)
        crash("internal error detected in function log_wrapper():\ndomain error. x is %g", x);
    return log(x);
}
static double pow_wrapper(double x, double y)
{
    if (x < 0.0 || (x == 0.0 && y <= 0.0))
        crash("internal error detected in function pow_wrapper():\ndomain error. x is %g, y is %g", x, y);
    return pow(x, y);
}

enum Phase
{
    PH_ST_BEGIN, /* GetSoln: first random tree, StartingTemperature set-up   (Solve.c:537-541) */
    PH_ST_LOOP,  /* StartingTemperature's sampling loop                     (StartingTemperature.c:106-181) */
    PH_AN_BEGIN, /* GetSoln: second random tree, Anneal set-up               (Solve.c:543-545, 194-233) */
    PH_AN_LOOP,  /* Anneal's loop                                            (Solve.c:235-469) */
    PH_FINISH,   /* GetSoln after Anneal                                     (Solve.c:567-569) */
    PH_DONE
};
enum Outcome
{
    REJECT,
    ACCEPT_LE,   /* not longer than the current tree: accepted, no draw      (Solve.c:307) */
    ACCEPT_WORSE /* longer, accepted by the Metropolis draw                  (Solve.c:368) */
};

/* Everything the two loops carry from one proposal to the next, except the tree lengths
 * (known only after the device answered) and the trees themselves. */
struct SAState
{
    Phase phase = PH_ST_BEGIN;
    Uni rng;
    /* Anneal (Solve.c:159-182, 209-212) */
    long accepted = 0, failedcnt = 0, iter = 0, t_n = 0, proposed = 0, current_iter = 0;
    double t = 0, t0 = 0;
    double trops_counter[3] = {1, 1, 1}, trops_probs[3] = {0, 0, 0};
    long trops_id = 0;
    long w_changes_prop = 0, w_changes_acc = 0;
    int op = 0; /* operator of the proposal being handled: 0 NNI, 1 SPR, 2 TBR */
    /* StartingTemperature (StartingTemperature.c:81-86) */
    long st_iter = 0;
    int acc_pos_trans = 0, prop_pos_trans = 0;
};

/* What the device tree of a chain still has to be told before the next proposals are scored:
 * the rows to recompute, on the chain's current topology at flush time.  An acceptance, a
 * re-root and both together collapse into one request (the union of their dirty sets is
 * closed towards the root and covers every node whose subtree changed); a fresh random tree
 * replaces whatever was pending. */
struct CommitReq
{
    std::vector<int64_t> dirty;
    bool all = false;         /* every internal node */
    bool want_length = false; /* the chain waits for the length of this commit */
    bool used = false;
};

struct Chain
{
    Dataptr MSA = nullptr;
    Parameters rc;
    int id = 0;
    int dev_tree = -1;
    bool log_progress = false;
    TREESTACK *stack = nullptr;
    std::vector<TrajectoryRow> *trajectory = nullptr;
    FILE *tsv = nullptr;
    FILE *sumfp = nullptr;

    TREESTACK_TREE_NODES *cur = nullptr;
    long root = 0;
    long current_len = 0, best_len = 0;
    bool await_len = false;
    double tree_minimum_length = 0;
    double log_eps = 0, log_grad_geom = 0, log_t0 = 0;
    SAState S;

    int kmax = 32, k_now = 1;
    double ema_used = 0, ema_taken = 0; /* recent proposals consumed / accepted (speculation depth) */
    bool shared_device = false;         /* many chains per device: wasted proposals cost device time */
    struct Spec
    {
        SAState mid; /* state after the proposal was generated, before the acceptance draw */
        double u;    /* the draw a worsening proposal consumes */
        bool row_valid;
        TrajectoryRow row; /* log line if rejected (length filled in at replay) */
    };
    std::vector<Proposal> cand;
    /* programs compiled on the thread that generated the proposals (lvb_gpu_compile_schedule) */
    struct Compiled
    {
        std::vector<int32_t> instr;
        int64_t n_instr = 0, n_slots = 0;
        bool valid = false;
    };
    std::vector<Compiled> cprog; /* one per speculative proposal */
    Compiled ccommit;            /* the pending commit request, if it is not a whole-tree one */
    ProposalScratch scratch;
    std::vector<Spec> spec;
    int n_spec = 0;
    SAState spec_end;
    CommitReq commit;
    /* score AND keep (lvb_gpu_score_keep_*): the first proposal of the round is scored with its rows stored in the
     * spare rows of the device tree, and if it is the one accepted the device tree switches over to them
     * (lvb_gpu_accept_kept) instead of recomputing them in a commit.  Used while a proposal is more likely
     * accepted than not — the hot phase, where nearly every round ends with its first proposal accepted. */
    bool keep_ok = true;      /* LVB_GPU_KEEP=0 switches it off */
    bool keep_first = false;  /* this round's first proposal is kept */
    bool accept_kept = false; /* ... and was accepted: the driver tells the device */

    SearchResult res;

    ~Chain() { free(cur); }

    void init(Dataptr msa, const Parameters &prm, int chain_id, int device_tree, int kmax_)
    {
        MSA = msa;
        rc = prm;
        id = chain_id;
        dev_tree = device_tree;
        kmax = kmax_ > 0 ? kmax_ : 32;
        k_now = prm.speculation > 0 ? kmax : 1;
        cur = treealloc(MSA, true);
        cand.resize((size_t)kmax);
        cprog.resize((size_t)kmax);
        spec.resize((size_t)kmax);
        S.rng.rinit(rc.seed + chain_id); /* Main.c:104 */
        tree_minimum_length = (double)MSA->min_len_tree;
        log_eps = log_wrapper(LVB_EPS);
        log_grad_geom = log_wrapper(0.99);
    }

    /* the current tree with the links proposal pr rewired (pr == nullptr: the commit request) */
    lvb_gpu_topology topology_of(const Proposal *pr) const
    {
        lvb_gpu_topology tp = lvb_gpu_topology();
        tp.left = &cur[0].left;
        tp.right = &cur[0].right;
        tp.stride_bytes = (int64_t)sizeof(TREESTACK_TREE_NODES);
        tp.root = root;
        if (pr)
        {
            tp.dirty = pr->dirty.data();
            tp.n_dirty = (int64_t)pr->dirty.size();
            tp.mod_node = pr->mod_node.data();
            tp.mod_left = pr->mod_left.data();
            tp.mod_right = pr->mod_right.data();
            tp.n_mod = (int64_t)pr->mod_node.size();
        }
        else
        {
            tp.dirty = commit.dirty.data();
            tp.n_dirty = commit.all ? -1 : (int64_t)commit.dirty.size();
        }
        return tp;
    }

    void compile_into(Compiled &dst, const lvb_gpu_topology &tp, bool store)
    {
        const size_t cap = (size_t)tp.n_dirty + 2; /* one instruction per dirty node + the root step */
        if (dst.instr.size() < 4 * cap)
            dst.instr.resize(4 * cap);
        if (lvb_gpu_compile_schedule(MSA->n, &tp, store ? 1 : 0, dst.instr.data(), (int64_t)cap, &dst.n_instr, &dst.n_slots) != 0)
            crash("GPU error: %s", lvb_gpu_last_error());
        dst.valid = true;
    }

    /* end of advance(): the programs of this round, compiled on the thread that owns the chain */
    void compile_round()
    {
        ccommit.valid = false;
        if (commit.used && !commit.all)
            compile_into(ccommit, topology_of(nullptr), true);
        for (int k = 0; k < n_spec; k++)
            compile_into(cprog[(size_t)k], topology_of(&cand[(size_t)k]), false);
    }

    template <class It>
    void request_rows(It first, It last)
    {
        commit.used = true;
        if (!commit.all)
            commit.dirty.insert(commit.dirty.end(), first, last);
    }

    /* a new random start: PullRandomTree + ss_init (Solve.c:537-538, 543-544), all rows to compute */
    void fresh_tree()
    {
        PullRandomTree(MSA, S.rng, cur);
        ss_init(MSA, cur);
        root = 0;
        commit.used = true;
        commit.all = true;
        commit.want_length = true;
        commit.dirty.clear();
        await_len = true;
        mark_all_clean(MSA, cur);
        res.getplen_calls++;
    }

    /* arbreroot on the current tree (Solve.c:240; StartingTemperature.c:116-117).  The reference
     * leaves `current` dirty and pays a full pass per proposal until one is accepted; here the
     * rows of the nodes between the two roots — the only ones that change — are recomputed once. */
    void reroot_current()
    {
        std::vector<long> path;
        root = arbreroot(MSA, S.rng, cur, root, &path);
        mark_all_clean(MSA, cur);
        request_rows(path.begin(), path.end());
    }

    /* ---- the part of an Anneal iteration before getplen (Solve.c:237-298), without the
     * re-root: operator choice and mutate.  Returns the operator. ---- */
    int anneal_propose(SAState &W, Proposal &dst)
    {
        lvb_assert(W.t > DBL_EPSILON);
        const int a = rc.algorithm_selection;
        if (a == 2)
        {
            const long total = (long)(W.trops_counter[0] + W.trops_counter[1] + W.trops_counter[2]);
            for (int i = 0; i < 3; i++)
                W.trops_probs[i] = W.trops_counter[i] / total;
        }
        int op;
        if (a >= 1)
        {
            const double r = W.rng.uni();
            if (r < W.trops_probs[0])
                op = 0;
            else if (r < W.trops_probs[0] + W.trops_probs[1])
                op = 1;
            else
                op = 2;
            if (a == 2)
                W.trops_id = op;
        }
        else
            op = (W.iter & 1) ? 1 : 0;
        propose(MSA, W.rng, op, cur, root, scratch, dst); /* mutate_nni / _spr / _tbr, in place and undone */
        W.op = op;
        return op;
    }

    /* ---- the part of an Anneal iteration after getplen (Solve.c:307-468).  `real` is false
     * while speculating (outcome REJECT only; no tree, stack or length is touched).
     * Returns false when the loop ends. ---- */
    bool anneal_after(SAState &W, Outcome out, long len, const TREESTACK_TREE_NODES *proposed, bool real,
                      TrajectoryRow *row)
    {
        const int a = rc.algorithm_selection;
        int changeAcc = 0;
        if (out == ACCEPT_LE)
        {
            lvb_assert(real);
            if (len <= best_len)
            {
                if (len < best_len)
                    ClearTreestack(stack);
                if (CompareTreeToTreestack(MSA, stack, proposed, root) == 1)
                    W.accepted++;
            }
            current_len = len;
            if (len < best_len)
                best_len = len;
            if (a == 1)
                changeAcc = 1;
        }
        else
        {
            if (a == 2)
                W.w_changes_prop++;
            if (out == ACCEPT_WORSE)
            {
                lvb_assert(real);
                if (a == 2)
                    W.w_changes_acc++;
                current_len = len;
                if (a == 1)
                    changeAcc = 1;
            }
        }
        W.proposed++;

        bool dect = false;
        if (W.accepted >= MAXACCEPT_SLOW)
        {
            W.failedcnt = 0;
            dect = true;
        }
        else if (W.proposed >= MAXPROPOSE_SLOW)
        {
            W.failedcnt++;
            if (W.failedcnt >= MAXFAIL_SLOW && W.t < FROZEN_T)
                return false; /* frozen: end of cooling (Solve.c:389-399) */
            dect = true;
        }
        if (dect)
        {
            W.t_n++;
            if (rc.cooling_schedule == 0)
            {
                const double ln_t = ((double)W.t_n) * log_grad_geom + log_t0;
                if (ln_t < log_eps)
                    W.t = LVB_EPS;
                else
                    W.t = pow_wrapper(0.99, (double)W.t_n) * W.t0;
                if (a == 1)
                {
                    W.trops_probs[2] = W.t / W.t0;
                    W.trops_probs[1] = (1 - W.trops_probs[2]) / 2;
                    W.trops_probs[0] = W.trops_probs[1];
                }
            }
            else
            {
                W.t = W.t0 - (10 * LVB_EPS) * W.t_n;
                if (W.t < DBL_EPSILON || W.t <= LVB_EPS)
                    W.t = LVB_EPS;
            }
            W.proposed = 0;
            W.accepted = 0;
            if (a == 2)
            {
                W.w_changes_prop = 0;
                W.w_changes_acc = 0;
            }
        }
        W.iter++;
        if (rc.n_number_max_trees > 0 && stack->next >= rc.n_number_max_trees)
            return false;
        if (a == 2)
        {
            if (changeAcc == 1)
                W.trops_counter[W.trops_id]++;
            else
                for (int i = 0; i < 3; i++)
                    if (W.trops_id != i)
                        W.trops_counter[i] += 0.5;
        }
        if (row)
            *row = TrajectoryRow{W.iter, W.op, changeAcc, current_len, W.t, stack->next};
        if (rc.max_rearrangements > 0 && W.current_iter >= rc.max_rearrangements)
            return false; /* benchmark cut-off of this build; not in the reference */
        return true;
    }

    /* ---- StartingTemperature, after getplen (StartingTemperature.c:135-181).  Returns false when
     * the starting temperature is found; W.t0 then holds it. ---- */
    bool st_after(SAState &W, Outcome out, long len, bool real)
    {
        if (out == ACCEPT_LE)
        {
            lvb_assert(real);
            current_len = len;
        }
        else
        {
            W.prop_pos_trans++;
            if (out == ACCEPT_WORSE)
            {
                lvb_assert(real);
                current_len = len;
                W.acc_pos_trans++;
            }
        }
        W.st_iter++;
        if (W.st_iter <= 100) /* sample_size + 1 proposals per temperature (:111) */
            return true;
        const double increment_size = 0.00001;
        const double r_acc_to_prop = (double)W.acc_pos_trans / W.prop_pos_trans; /* 0/0 = NaN ends the loop (:171) */
        W.t += increment_size;
        W.st_iter = 0;
        W.prop_pos_trans = 0;
        W.acc_pos_trans = 0;
        if (W.t >= 1 || W.t <= 0)
        {
            W.t0 = 1; /* :175-176 */
            return false;
        }
        if (r_acc_to_prop <= 0.65)
            return true;
        W.t0 = W.t - increment_size; /* :193 */
        return false;
    }

    Outcome decide(const SAState &mid, double u, long len) const
    {
        const long change = len - current_len;
        double deltah = (tree_minimum_length / (double)current_len) - (tree_minimum_length / (double)len);
        if (deltah > 1.0)
            deltah = 1.0; /* MinimumTreeLength() problem with ambiguous sites (Solve.c:304-305) */
        if (change <= 0)
            return ACCEPT_LE;
        if (-deltah < mid.t * log_eps)
            return REJECT; /* pacc taken as 0; the draw is consumed all the same (Solve.c:356-364) */
        const double pacc = exp_wrapper(-deltah / mid.t);
        return u < pacc ? ACCEPT_WORSE : REJECT;
    }

    void emit(const TrajectoryRow &r)
    {
        static const char *const names[3] = {"NNI", "SPR", "TBR"};
        if (trajectory)
            trajectory->push_back(r);
        if (tsv)
            fprintf(tsv, "%ld\t%s\t%d\t%ld\t%lf\t%ld\n", r.iter, names[r.op], r.accepted, r.length, r.t * 10000, r.stack);
    }

    /* ---- one round, first half: bring the chain to the point where it needs lengths ---- */
    void advance()
    {
        n_spec = 0;
        keep_first = accept_kept = false;
        ccommit.valid = false;
        if (S.phase == PH_DONE)
            return;
        if (S.phase == PH_ST_BEGIN)
        {
            fresh_tree();
            S.t = LVB_EPS; /* StartingTemperature.c:76 */
            S.st_iter = 0;
            S.acc_pos_trans = S.prop_pos_trans = 0;
            S.phase = PH_ST_LOOP;
        }
        else if (S.phase == PH_AN_BEGIN)
        {
            fresh_tree();
            /* Anneal set-up, Solve.c:159-233 */
            S.accepted = S.failedcnt = S.iter = S.t_n = S.proposed = 0;
            S.t = S.t0;
            lvb_assert(((float)S.t >= (float)LVB_EPS) && (S.t <= 1.0));
            log_t0 = log_wrapper(S.t0);
            for (int i = 0; i < 3; i++)
            {
                S.trops_counter[i] = 1;
                S.trops_probs[i] = 0;
            }
            S.trops_id = 0;
            S.w_changes_prop = S.w_changes_acc = 0;
            CompareHashTreeToHashstack(MSA, stack, cur, root); /* init. tree initially best (Solve.c:207) */
            if (log_progress && S.current_iter == 0)
            {
                printf("\n--------------------------------------------------------");
                printf("\n  Temperature:   Rearrangement: TreeStack size: Length:\n");
                printf("--------------------------------------------------------\n");
            }
            if (tsv)
                fprintf(tsv, "Iteration\tAlgorithm\tAccepted\tLength\tTemperature\tTreestack Size\n");
            S.phase = PH_AN_LOOP;
        }
        else if (S.phase == PH_FINISH)
        {
            /* Solve.c:567-569: pop the last best tree and push it back through the hash path */
            TREESTACK_TREE_NODES *tmp = treealloc(MSA, true);
            long r = 0;
            PullTreefromTreestack(MSA, tmp, &r, stack);
            CompareHashTreeToHashstack(MSA, stack, tmp, r);
            free(tmp);
            res.final_length = best_len;
            res.rearrangements = S.current_iter;
            res.start_temperature = S.t0;
            S.phase = PH_DONE;
            return;
        }

        const bool annealing = S.phase == PH_AN_LOOP;
        /* the first proposal of the batch may begin with a re-root of the current tree */
        if (annealing)
        {
            S.current_iter++;
            if (S.current_iter % REROOT_INTERVAL == 0)
            {
                reroot_current();
                if (log_progress && S.current_iter % STAT_LOG_INTERVAL == 0 && !await_len)
                    printf("  %-15.8f%-15ld%-16ld%-15ld\n", S.t, S.current_iter, stack->next, current_len); /* lenlog */
            }
        }
        else if (S.st_iter % REROOT_INTERVAL == 0)
            reroot_current();

        SAState W = S;
        const int want = k_now;
        while (n_spec < want)
        {
            if (n_spec > 0)
            { /* later proposals: stop in front of a re-root */
                if (annealing)
                {
                    if ((W.current_iter + 1) % REROOT_INTERVAL == 0)
                        break;
                    W.current_iter++;
                }
                else if (W.st_iter % REROOT_INTERVAL == 0)
                    break;
            }
            Spec &sp = spec[(size_t)n_spec];
            Proposal &dst = cand[(size_t)n_spec];
            if (annealing)
                anneal_propose(W, dst);
            else
            {
                lvb_assert(W.t > DBL_EPSILON);
                W.op = (W.st_iter & 1) ? 1 : 0; /* StartingTemperature.c:123-126 */
                propose(MSA, W.rng, W.op, cur, root, scratch, dst);
            }
            sp.mid = W;
            sp.u = W.rng.uni();
            sp.row_valid = false;
            bool go_on;
            if (annealing)
            {
                go_on = anneal_after(W, REJECT, 0, nullptr, false, &sp.row);
                sp.row_valid = go_on; /* the iteration that ends the loop writes no line (Solve.c:394) */
                if (!go_on)
                    W.phase = PH_FINISH;
            }
            else
            {
                go_on = st_after(W, REJECT, 0, false);
                if (!go_on)
                    W.phase = PH_AN_BEGIN;
            }
            n_spec++;
            if (!go_on)
                break;
        }
        spec_end = W;
        res.getplen_calls += n_spec;
        keep_first = keep_ok && n_spec > 0 && (ema_taken + 0.25) / (ema_used + 0.5) >= 0.5;
        compile_round();
    }

    /* ---- one round, second half: lengths are known ---- */
    void resolve(const int64_t *commit_len, const int64_t *lens)
    {
        if (await_len)
        {
            lvb_assert(commit_len != nullptr);
            current_len = (long)*commit_len;
            await_len = false;
            if (S.phase == PH_AN_LOOP)
            {
                best_len = current_len; /* Solve.c:205 */
                res.initial_length = current_len;
                if (sumfp)
                    fprintf(sumfp, "%d\t%d\t%ld\t", 0, 0, current_len); /* Solve.c:558-559 */
            }
        }
        if (n_spec == 0)
            return;
        const bool annealing = S.phase == PH_AN_LOOP;
        int taken = -1;
        for (int k = 0; k < n_spec; k++)
        {
            Spec &sp = spec[(size_t)k];
            const long len = (long)lens[k];
            lvb_assert(len >= 1L);
            const Outcome out = decide(sp.mid, sp.u, len);
            if (out == REJECT)
            {
                if (sp.row_valid)
                {
                    sp.row.length = current_len;
                    emit(sp.row);
                }
                continue;
            }
            /* accepted: resume from the snapshot of this proposal */
            taken = k;
            S = sp.mid;
            if (out == ACCEPT_WORSE)
                (void)S.rng.uni();
            /* SwapTrees (Solve.c:323, 370): the proposal becomes the current tree; on the device
             * only its dirty rows are rewritten */
            apply_proposal(cur, cand[(size_t)k]);
            bool go_on;
            if (annealing)
            {
                TrajectoryRow row;
                go_on = anneal_after(S, out, len, cur, true, &row);
                if (go_on)
                    emit(row);
                else
                    S.phase = PH_FINISH;
            }
            else
            {
                go_on = st_after(S, out, len, true);
                if (!go_on)
                    S.phase = PH_AN_BEGIN;
            }
            if (go_on)
            {
                if (k == 0 && keep_first)
                    accept_kept = true; /* its rows are on the device already */
                else
                    request_rows(cand[(size_t)k].dirty.begin(), cand[(size_t)k].dirty.end());
            }
            break;
        }
        if (taken < 0)
            S = spec_end;
        if (S.phase == PH_AN_BEGIN && log_progress)
            printf("  SA Starting Temperature: %-.8f\n", S.t0);
        /* speculation depth for the next round.  With acceptance probability p per proposal a batch
         * of K uses (1 - (1-p)^K) / p of its proposals; the rest is scored for nothing.  p is
         * estimated from the recent rounds (exponentially forgotten counts) and K is a multiple of
         * the expected run of rejections 1/p: about two runs when this chain has the device to
         * itself (a round costs a launch and a read-back whatever K is, wasted proposals are
         * nearly free), under one run when many chains share the device (the device is the
         * bottleneck, a wasted proposal costs as much as a useful one).  Results do not depend
         * on K (tested). */
        ema_used = 0.9 * ema_used + (taken < 0 ? n_spec : taken + 1);
        ema_taken = 0.9 * ema_taken + (taken < 0 ? 0 : 1);
        if (rc.speculation > 0)
            k_now = kmax;
        else
        {
            const double p_acc = (ema_taken + 0.25) / (ema_used + 0.5);
            const double want = (shared_device ? 0.75 : 2.0) / p_acc;
            k_now = want >= kmax ? kmax : std::max(1, (int)(want + 0.5));
        }
    }
};

static void gpu_check(int rc_)
{
    if (rc_ != 0)
        crash("GPU error: %s", lvb_gpu_last_error());
}

/* src/Solve.c:481-609 for R chains at once */
/* One group of chains [r0, r1) on one device context, driven by the calling thread (its inner
 * loops over chains use `threads` OpenMP threads).  The group is worked as two alternating
 * halves: while the device scores the proposals of one half, the host replays the acceptances
 * of the other half, generates its next proposals and compiles their programs.  A round of a
 * half: advance every chain (speculative proposals), one commit launch for the chains whose
 * device tree must catch up, one score launch for all proposals (submitted, not waited for);
 * then, a half-round later, wait and replay.  Chains never interact, so the result of a chain
 * does not depend on which half or group it is in (tested). */
static void run_group(Dataptr MSA, Device &dev, std::vector<Chain> &chains, int r0, int r1, int threads,
                      long &score_launches, long &rounds, double prof[4])
{
    struct Half
    {
        int a = 0, b = 0;
        bool alive = true, inflight = false;
        int ticket = -1;
        std::vector<int> trees, owner;
        std::vector<lvb_gpu_topology> topos;
        std::vector<lvb_gpu_program_view> pre;
        std::vector<int64_t> lens, clen, accepts;
        std::vector<size_t> offsets;
        std::vector<uint8_t> keepf;
        bool kept_any = false;
    };
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto since = [](std::chrono::steady_clock::time_point a) {
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
    };
    /* site-block mode keeps the ranks in lock-step through the all-reduce: one half */
    const bool whole = dev.word_count >= 0 && dev.word_count != MSA->nwords;
    {
        const char *ke = getenv("LVB_GPU_KEEP");
        const bool keep_ok = !(ke && atoi(ke) == 0) && !whole;
        for (int r = r0; r < r1; r++)
            chains[(size_t)r].keep_ok = keep_ok;
    }
#ifdef _OPENMP
    /* this driver thread's team size, also seen by the library when it spreads the schedule
     * compile of a batch over threads (omp_get_max_threads there) */
    omp_set_num_threads(std::max(1, threads));
#endif
    const int nh = (r1 - r0 >= 2 && !whole) ? 2 : 1;
    Half H[2];
    for (int x = 0; x < nh; x++)
    {
        H[x].a = r0 + (int)((long)(r1 - r0) * x / nh);
        H[x].b = r0 + (int)((long)(r1 - r0) * (x + 1) / nh);
        H[x].clen.assign((size_t)(H[x].b - H[x].a), 0);
    }
    std::string failure;

    /* first part of a round: up to the submitted score launch */
    auto prep = [&](Half &h) {
        for (;;)
        {
            auto tp0 = now();
            /* chains are independent between device calls: spread them over the host threads (-p) */
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) if (threads > 1)
            for (int r = h.a; r < h.b; r++)
            {
                try
                {
                    chains[(size_t)r].advance();
                }
                catch (const LvbError &e)
                {
#pragma omp critical
                    failure = e.message;
                }
            }
            if (!failure.empty())
                crash("%s", failure.c_str());
            h.alive = false;
            for (int r = h.a; r < h.b; r++)
                h.alive = h.alive || chains[(size_t)r].S.phase != PH_DONE;
            prof[0] += since(tp0);
            rounds++;
            tp0 = now();
            /* what the device trees must catch up with (at most one request per chain), one launch */
            h.trees.clear();
            h.topos.clear();
            h.pre.clear();
            h.owner.clear();
            bool want = false;
            for (int r = h.a; r < h.b; r++)
            {
                Chain &c = chains[(size_t)r];
                if (!c.commit.used)
                    continue;
                h.trees.push_back(c.dev_tree);
                h.topos.push_back(c.topology_of(nullptr));
                /* compiled in advance() unless it is a whole-tree evaluation (cut into pieces by the library) */
                h.pre.push_back(c.ccommit.valid ? lvb_gpu_program_view{c.ccommit.instr.data(), c.ccommit.n_instr, c.ccommit.n_slots}
                                                : lvb_gpu_program_view{nullptr, 0, 0});
                h.owner.push_back(r);
                want = want || c.commit.want_length;
            }
            if (!h.trees.empty())
            {
                h.lens.resize(h.trees.size());
                gpu_check(lvb_gpu_commit_batch_pre(dev.ctx, (int64_t)h.trees.size(), h.trees.data(), h.topos.data(),
                                                   h.pre.data(), want ? h.lens.data() : nullptr));
                for (size_t i = 0; i < h.owner.size(); i++)
                {
                    Chain &c = chains[(size_t)h.owner[i]];
                    if (c.commit.want_length)
                        h.clen[(size_t)(h.owner[i] - h.a)] = h.lens[i];
                    c.commit.used = c.commit.all = c.commit.want_length = false;
                    c.commit.dirty.clear();
                }
            }
            prof[1] += since(tp0);
            if (!h.alive)
                return;
            tp0 = now();
            /* proposals of all chains of the half in one launch */
            h.trees.clear();
            h.topos.clear();
            h.pre.clear();
            h.offsets.resize((size_t)(h.b - h.a));
            for (int r = h.a; r < h.b; r++)
            {
                Chain &c = chains[(size_t)r];
                h.offsets[(size_t)(r - h.a)] = h.trees.size();
                for (int k = 0; k < c.n_spec; k++)
                {
                    const Chain::Compiled &cp = c.cprog[(size_t)k];
                    h.trees.push_back(c.dev_tree);
                    h.topos.push_back(c.topology_of(&c.cand[(size_t)k])); /* the current tree with the links this proposal rewired */
                    h.pre.push_back(lvb_gpu_program_view{cp.instr.data(), cp.n_instr, cp.n_slots});
                }
            }
            h.lens.resize(h.trees.size());
            h.keepf.assign(h.trees.size(), 0);
            h.kept_any = false;
            for (int r = h.a; r < h.b && !whole; r++)
                if (chains[(size_t)r].keep_first)
                {
                    h.keepf[h.offsets[(size_t)(r - h.a)]] = 1;
                    h.kept_any = true;
                }
            if (h.kept_any)
            {
                gpu_check(lvb_gpu_score_keep_submit_pre(dev.ctx, (int64_t)h.trees.size(), h.trees.data(), h.topos.data(),
                                                        h.pre.data(), h.keepf.data(), &h.ticket));
                score_launches++;
                h.inflight = true;
            }
            else if (!h.trees.empty())
            {
                gpu_check(lvb_gpu_score_batch_submit_pre(dev.ctx, (int64_t)h.trees.size(), h.trees.data(), h.topos.data(),
                                                         h.pre.data(), &h.ticket));
                score_launches++;
                h.inflight = true;
            }
            prof[2] += since(tp0);
            if (h.inflight)
                return;
            /* nothing to score (every live chain only changed phase): replay and go round again */
            auto tp1 = now();
            for (int r = h.a; r < h.b; r++)
                chains[(size_t)r].resolve(&h.clen[(size_t)(r - h.a)], nullptr);
            prof[3] += since(tp1);
        }
    };
    /* second part: the lengths are known */
    auto finish = [&](Half &h) {
        auto tp0 = now();
        gpu_check(lvb_gpu_batch_wait(dev.ctx, h.ticket, h.lens.data()));
        h.inflight = false;
        prof[2] += since(tp0);
        tp0 = now();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) if (threads > 1)
        for (int r = h.a; r < h.b; r++)
        {
            try
            {
                chains[(size_t)r].resolve(&h.clen[(size_t)(r - h.a)], h.lens.data() + h.offsets[(size_t)(r - h.a)]);
            }
            catch (const LvbError &e)
            {
#pragma omp critical
                failure = e.message;
            }
        }
        if (!failure.empty())
            crash("%s", failure.c_str());
        /* kept proposals that were accepted: the device trees switch over to their rows */
        h.accepts.clear();
        for (int r = h.a; r < h.b; r++)
            if (chains[(size_t)r].accept_kept)
            {
                lvb_assert(h.kept_any && h.keepf[h.offsets[(size_t)(r - h.a)]] == 1);
                h.accepts.push_back((int64_t)h.offsets[(size_t)(r - h.a)]);
                chains[(size_t)r].accept_kept = false;
            }
        if (!h.accepts.empty())
            gpu_check(lvb_gpu_accept_kept(dev.ctx, h.ticket, (int64_t)h.accepts.size(), h.accepts.data()));
        prof[3] += since(tp0);
    };
    for (int x = 0; x < nh; x++)
        prep(H[x]);
    for (int x = 0; H[0].inflight || H[nh - 1].inflight; x = (x + 1) % nh)
    {
        if (!H[x].inflight)
            continue;
        finish(H[x]);
        prep(H[x]);
    }
}

/* src/Solve.c:64-142, same loops and the same accept rule; the 2(n-3) getplen calls of one
 * sweep position onwards are answered by ONE lvb_gpu_nni_neighbourhood call, made again only
 * after a rearrangement was accepted (the lengths of the others are then stale). */
long deterministic_hillclimb(Dataptr MSA, Device &dev, TREESTACK *treestack_ptr, const TREESTACK_TREE_NODES *inittree,
                             const Parameters &rcstruct, long root, FILE *lenfp, long *current_iter, bool log_progress,
                             long *neighbourhood_calls)
{
    (void)rcstruct;
    const long n = MSA->n, number_of_internal_branches = MSA->numberofpossiblebranches - MSA->n;
    TREESTACK_TREE_NODES *p_current_tree = treealloc(MSA, true);
    TREESTACK_TREE_NODES *p_proposed_tree = treealloc(MSA, true);
    treecopy(MSA, p_current_tree, inittree, true);
    int dev_tree = -1, up_tree = -1;
    gpu_check(lvb_gpu_tree_create(dev.ctx, &dev_tree));
    gpu_check(lvb_gpu_tree_create(dev.ctx, &up_tree));
    const int64_t stride = (int64_t)sizeof(TREESTACK_TREE_NODES);
    auto topology_of = [&](const TREESTACK_TREE_NODES *t, const std::vector<int64_t> *dirty) {
        lvb_gpu_topology tp = lvb_gpu_topology();
        tp.left = &t[0].left;
        tp.right = &t[0].right;
        tp.stride_bytes = stride;
        tp.root = root;
        tp.dirty = dirty ? dirty->data() : nullptr;
        tp.n_dirty = dirty ? (int64_t)dirty->size() : -1;
        return tp;
    };
    int64_t len64 = 0;
    {
        const lvb_gpu_topology tp = topology_of(p_current_tree, nullptr); /* first evaluation: every node */
        gpu_check(lvb_gpu_tree_commit(dev.ctx, dev_tree, &tp, &len64));
    }
    mark_all_clean(MSA, p_current_tree);
    long current_tree_length = (long)len64;
    std::vector<int64_t> lens((size_t)(2 * number_of_internal_branches)), dirty;
    bool stale = true, newtree;
    long calls = 0;
    do
    {
        newtree = false;
        for (long i = 0; i < number_of_internal_branches; i++)
        {
            for (int j = 0; j < 2; j++)
            {
                if (stale)
                {
                    const lvb_gpu_topology tp = topology_of(p_current_tree, nullptr);
                    gpu_check(lvb_gpu_nni_neighbourhood(dev.ctx, dev_tree, up_tree, &tp, lens.data()));
                    stale = false;
                    calls++;
                }
                const long proposed_tree_length = (long)lens[(size_t)(2 * i + j)];
                lvb_assert(proposed_tree_length >= 1L);
                const long tree_length_change = proposed_tree_length - current_tree_length;
                if (tree_length_change <= 0)
                {
                    mutate_deterministic(MSA, p_proposed_tree, p_current_tree, root, n + i, j == 1);
                    if (tree_length_change < 0) /* very best so far */
                    {
                        ClearTreestack(treestack_ptr);
                        current_tree_length = proposed_tree_length;
                    }
                    if (CompareHashTreeToHashstack(MSA, treestack_ptr, p_proposed_tree, root) == 1)
                    {
                        newtree = true;
                        /* SwapTrees: the proposal becomes the current tree; on the device only the
                         * rows mutate_deterministic dirtied are rewritten */
                        collect_dirty(MSA, p_proposed_tree, dirty);
                        std::swap(p_current_tree, p_proposed_tree);
                        const lvb_gpu_topology tp = topology_of(p_current_tree, &dirty);
                        gpu_check(lvb_gpu_tree_commit(dev.ctx, dev_tree, &tp, nullptr));
                        mark_all_clean(MSA, p_current_tree);
                        stale = true;
                    }
                }
                if (log_progress && (*current_iter % STAT_LOG_INTERVAL) == 0 && lenfp)
                    fprintf(lenfp, "  %-15.8f%-15ld%-16ld%-15ld\n", 0.0, *current_iter, treestack_ptr->next, current_tree_length);
                *current_iter += 1;
            }
        }
    } while (newtree);
    gpu_check(lvb_gpu_tree_destroy(dev.ctx, dev_tree));
    gpu_check(lvb_gpu_tree_destroy(dev.ctx, up_tree));
    free(p_current_tree);
    free(p_proposed_tree);
    if (neighbourhood_calls)
        *neighbourhood_calls = calls;
    return current_tree_length;
}

/* GetSoln for R independent replicates.  The chains are split into G groups; every group has
 * its own device context (own stream, own copy of the leaf rows) and its own host thread, so
 * that one group's host work (proposing, replaying, compiling schedules) overlaps the other
 * groups' kernels.  Chain r uses seed rc.seed + r whatever the grouping: results do not depend
 * on G (tested). */
void GetSolnBatched(Dataptr MSA, const Parameters &rc, Device &dev, int replicates, std::vector<TREESTACK> &stacks,
                    std::vector<SearchResult> &results, std::vector<TrajectoryRow> *trajectory, FILE *tsv)
{
    const auto t_start = std::chrono::steady_clock::now();
    const int R = replicates;
    stacks.assign((size_t)R, TREESTACK());
    results.assign((size_t)R, SearchResult());
    std::vector<Chain> chains((size_t)R);
    FILE *sumfp = nullptr;
    if (rc.verbose)
    {
        sumfp = fopen("sum", "w");
        if (sumfp)
            fprintf(sumfp, "StartNo\tCycleNo\tCycInit\tCycBest\tCycTrees\n");
    }
    const int host_threads = std::max(1, rc.n_processors_available);
    /* Replicates are spread over rc.devices GPUs (device, device+1, ...: no communication between
     * them, BASELINE.json configs[4]) and, on each GPU, over driver groups of at least 16 chains, at
     * most 4: every group has its own context, stream and host thread (with its share of -p for the
     * per-chain work) and works its chains as two alternating halves.  Rounds of a group are bound
     * by the latency of its two launches (programs of up to ~300 dependent Fitch steps), so several
     * groups keep the device busier: 64 chains of 1000 x 200k on one B200 with a 16-core host, search
     * time without set-up 1.44 / 1.02 / 0.98 s with 1 / 2 / 4 groups once the programs are compiled
     * in the chains' own loop (profiles/r01i_sa_config5_groups.log; before that, when every device
     * call compiled its batch itself, one group was fastest: 1.60 / 1.95 / 2.1 s).  -G overrides. */
    const int ND = std::max(1, std::min(rc.devices, R));
    /* measured again in round 2 (profiles/r02j_sa_probe.log, LVB_GPU_PROFILE <= 1: timing every launch with events
     * had slowed the earlier runs): a round of a group costs its host thread two device calls and the wait for two
     * latency-bound kernels whatever the number of chains in it, so few chains go ONE per group — 8 chains of
     * 1000 x 200k: 0.49 / 0.29 / 0.24 s with 1 / 4 / 8 groups, 16 chains 0.35 / 0.33 s with 8 / 16 — and many
     * chains (the device is then busy) into 8 groups: 64 chains 0.88 / 0.81 / 0.89 s with 4 / 8 / 16 */
    const int chains_per_dev = std::max(1, R / ND);
    /* ... but never more groups than host threads: a driver thread spins while it waits for its lengths, and on the
     * 8-GPU box (32 cores, 4 per rank) 8 groups of one chain on 4 cores took 0.37 s where the 16-core single-GPU box
     * needs 0.19 s (profiles/r02o_bench_n8.json, r02n_sa_probe.log); groups of two chains work as two alternating
     * halves instead, one chain's host work under the other's kernel */
    const int by_threads = std::max(1, host_threads / ND);
    int per_dev = rc.groups > 0 ? std::max(1, rc.groups / ND) : std::min(by_threads, chains_per_dev <= 16 ? chains_per_dev : 8);
    per_dev = std::max(1, per_dev);
    int G = std::max(1, std::min(ND * per_dev, R));
    if (dev.word_count >= 0 && dev.word_count != MSA->nwords)
        G = 1; /* site-block mode: one context per rank, the all-reduce keeps the ranks in lock-step */
    std::vector<Device> extra((size_t)G);
    std::vector<int> first((size_t)G + 1);
    for (int g = 0; g <= G; g++)
        first[(size_t)g] = (int)((long)R * g / G);
    if (G > 1)
    { /* the extra contexts hold their own copy of the leaf rows, taken from the first context device
       * to device (no second DNAToBinary + upload); one thread per context: opening a device (CUDA
       * context, pool allocation) takes 0.1-0.3 s each */
        std::vector<std::thread> openers;
        for (int g = 1; g < G; g++)
            openers.emplace_back([&, g] {
                device_open_clone(extra[(size_t)g], dev.index + g % ND, dev, first[(size_t)g + 1] - first[(size_t)g]);
            });
        for (auto &t : openers)
            t.join();
    }
    const double setup_secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
    auto device_of = [&](int g) -> Device & { return g == 0 ? dev : extra[(size_t)g]; };
    for (int g = 0; g < G; g++)
        for (int r = first[(size_t)g]; r < first[(size_t)g + 1]; r++)
        {
            int t = -1;
            gpu_check(lvb_gpu_tree_create(device_of(g).ctx, &t));
            chains[(size_t)r].init(MSA, rc, r, t, rc.speculation);
            chains[(size_t)r].shared_device = R / ND >= 4;
            chains[(size_t)r].stack = &stacks[(size_t)r];
            if (r == 0)
            {
                chains[0].log_progress = true;
                chains[0].trajectory = trajectory;
                chains[0].tsv = tsv;
                chains[0].sumfp = sumfp;
            }
        }
    struct GroupStat
    {
        long score_launches = 0, rounds = 0;
        double prof[4] = {0, 0, 0, 0};
    };
    std::vector<GroupStat> stat((size_t)G);
    /* per-chain work of a round is tens of microseconds: beyond 16 threads a team costs more at its
     * barriers than it gains (measured on a 32-core box, 64 chains on one GPU: the replay loop
     * takes 3.2 s with a team of 32 against 0.35 s with 16; profiles/r01g_sa_config5_devices.log) */
    const int inner = std::max(1, std::min(16, host_threads / G));
    if (G == 1)
        run_group(MSA, dev, chains, 0, R, std::min(inner, R), stat[0].score_launches, stat[0].rounds, stat[0].prof);
    else
    {
        std::vector<std::thread> workers;
        for (int g = 0; g < G; g++)
            workers.emplace_back([&, g] {
                run_group(MSA, device_of(g), chains, first[(size_t)g], first[(size_t)g + 1],
                          std::min(inner, first[(size_t)g + 1] - first[(size_t)g]), stat[(size_t)g].score_launches,
                          stat[(size_t)g].rounds, stat[(size_t)g].prof);
            });
        for (auto &w : workers)
            w.join();
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
    for (int g = 0; g < G; g++)
    {
        for (int r = first[(size_t)g]; r < first[(size_t)g + 1]; r++)
        {
            results[(size_t)r] = chains[(size_t)r].res;
            results[(size_t)r].launches_scored = stat[(size_t)g].score_launches;
            results[(size_t)r].seconds = secs;
            results[(size_t)r].rounds = stat[(size_t)g].rounds;
            results[(size_t)r].groups = G;
            results[(size_t)r].setup_seconds = setup_secs;
            for (int i = 0; i < 4; i++)
                results[(size_t)r].prof[i] = stat[(size_t)g].prof[i];
            gpu_check(lvb_gpu_tree_destroy(device_of(g).ctx, chains[(size_t)r].dev_tree));
        }
        if (g > 0)
        {
            dev.extra_launches += extra[(size_t)g].launches();
            device_close(extra[(size_t)g]);
        }
    }
    if (sumfp)
    {
        /* Solve.c:579-585: the trees of this cycle, then the summary line */
        FILE *resfp = fopen("res_start0_cycle0", "w");
        long treec = 0;
        if (resfp)
        {
            treec = PrintTreestack(MSA, &stacks[0], resfp);
            fclose(resfp);
        }
        fprintf(sumfp, "%ld\t%ld\n", results[0].final_length, treec);
        fclose(sumfp);
    }
}

} // namespace lvbhost
