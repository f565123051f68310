/* oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY.
 *
 * C-linkage doorway into the UNMODIFIED reference (phylolvb/lvb 4.2), compiled
 * from /root/reference/src in place by oracle/Makefile into oracle/_ref/libLVB_ref.so.
 * It lets tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
 * call the reference's own functions (getplen, DNAToBinary, mutate_*, arbreroot,
 * PullRandomTree, uni, ...) on the same inputs as the CUDA path.  Nothing under
 * lvb_b200/ may load this library.
 *
 * Every function here is a thin call-through; no reference logic is restated in
 * this file.  Call order for the known-answer tests follows SURVEY.md Appendix B:
 * load -> matchange -> calc_distribution_processors -> rinit -> DNAToBinary ->
 * treealloc -> PullRandomTree -> ss_init -> getplen / mutate_* / arbreroot.
 */
#include "LVB.h"
#include "Hash.h"
#include "DataOperations.h"
#include "MemoryOperations.h"
#include "SearchParameters.h"

#include <time.h>

struct RefHandle
{
	Dataptr msa;
	Parameters rc;
	Lvb_bit_length **enc; /* DNAToBinary output, n rows */
	long *todo;
	long *sums;
	int *runs;
};

static void ref_finish_setup(RefHandle *h)
{
	matchange(h->msa, h->rc);					  /* DataOperations.c:309 */
	calc_distribution_processors(h->msa, h->rc); /* MemoryOperations.c:135 */
	h->enc = (Lvb_bit_length **)malloc(h->msa->n * sizeof(Lvb_bit_length *));
	for (long i = 0; i < h->msa->n; i++)
		h->enc[i] = (Lvb_bit_length *)malloc(h->msa->bytes);
	DNAToBinary(h->msa, h->enc); /* DataOperations.c:164 */
	alloc_memory_to_getplen(h->msa, &h->todo, &h->sums, &h->runs);
}

extern "C"
{

	void *ref_msa_load(const char *path, int format, int nthreads)
	{
		RefHandle *h = (RefHandle *)calloc(1, sizeof(RefHandle));
		defaults_params(&h->rc);
		strncpy(h->rc.file_name_in, path, LVB_FNAMSIZE - 1);
		h->rc.n_file_format = format;
		h->rc.n_processors_available = nthreads;
		h->rc.verbose = LVB_FALSE;
		lvb_initialize();
		h->msa = (Dataptr)calloc(1, sizeof(DataStructure));
		phylip_dna_matrin(h->rc.file_name_in, format, h->msa); /* Wrapper.c:49 */
		ref_finish_setup(h);
		return h;
	}

	/* Build the DataStructure in memory exactly as read_file() fills it
	 * (CommandLineParser.cpp:63-102), skipping the text parse. */
	void *ref_msa_from_rows(long n, long m, const char *const *rows, int nthreads)
	{
		RefHandle *h = (RefHandle *)calloc(1, sizeof(RefHandle));
		defaults_params(&h->rc);
		h->rc.n_processors_available = nthreads;
		h->rc.verbose = LVB_FALSE;
		lvb_initialize();
		h->msa = (Dataptr)calloc(1, sizeof(DataStructure));
		Dataptr d = h->msa;
		d->m = m;
		d->original_m = m;
		d->n = n;
		d->numberofpossiblebranches = 2 * n - 3;
		d->max_length_seq_name = 10;
		d->nsets = n - 3;
		d->mssz = n - 2;
		d->rowtitle = (char **)malloc(n * sizeof(char *));
		d->row = (char **)malloc(n * sizeof(char *));
		for (long i = 0; i < n; i++)
		{
			d->rowtitle[i] = (char *)malloc(16);
			snprintf(d->rowtitle[i], 16, "T%ld", i);
			d->row[i] = (char *)malloc(m + 1);
			memcpy(d->row[i], rows[i], m);
			d->row[i][m] = '\0';
		}
		ref_finish_setup(h);
		return h;
	}

	/* The same for an alignment that is ALREADY packed (DNAToBinary's layout, n x words_per_row(m)
	 * words): used to time getplen at 2000 x 1M, where making 2 GB of text and scanning it twice
	 * (matchange: constchar + MinimumTreeLength) would add minutes of set-up.  The fields matchange
	 * derives from m (DataOperations.c:336-340) are set by the reference's own functions; the row
	 * strings stay empty and min_len_tree is not computed (getplen reads neither). */
	void *ref_msa_from_packed(long n, long m, const Lvb_bit_length *enc, int nthreads)
	{
		RefHandle *h = (RefHandle *)calloc(1, sizeof(RefHandle));
		defaults_params(&h->rc);
		h->rc.n_processors_available = nthreads;
		h->rc.verbose = LVB_FALSE;
		lvb_initialize();
		h->msa = (Dataptr)calloc(1, sizeof(DataStructure));
		Dataptr d = h->msa;
		d->m = m;
		d->original_m = m;
		d->n = n;
		d->numberofpossiblebranches = 2 * n - 3;
		d->max_length_seq_name = 10;
		d->nsets = n - 3;
		d->mssz = n - 2;
		d->rowtitle = (char **)malloc(n * sizeof(char *));
		d->row = (char **)malloc(n * sizeof(char *));
		for (long i = 0; i < n; i++)
		{
			d->rowtitle[i] = (char *)calloc(16, 1);
			snprintf(d->rowtitle[i], 16, "T%ld", i);
			d->row[i] = (char *)calloc(1, 1);
		}
		d->bytes = bytes_per_row(d->m);
		d->nwords = words_per_row(d->m);
		d->tree_bytes = tree_bytes(d);
		d->tree_bytes_without_sitestate = tree_bytes_without_sitestate(d);
		d->min_len_tree = 0;
		calc_distribution_processors(d, h->rc); /* MemoryOperations.c:135 */
		h->enc = (Lvb_bit_length **)malloc(n * sizeof(Lvb_bit_length *));
		for (long i = 0; i < n; i++)
		{
			h->enc[i] = (Lvb_bit_length *)malloc(d->bytes);
			memcpy(h->enc[i], enc + i * d->nwords, d->bytes);
		}
		alloc_memory_to_getplen(d, &h->todo, &h->sums, &h->runs);
		return h;
	}

	void ref_msa_free(void *hv)
	{
		RefHandle *h = (RefHandle *)hv;
		free_memory_to_getplen(&h->todo, &h->sums, &h->runs);
		for (long i = 0; i < h->msa->n; i++)
			free(h->enc[i]);
		free(h->enc);
		rowfree(h->msa);
		free(h->msa);
		free(h);
	}

	/* out[0..9] = n, m, original_m, nwords, bytes, numberofpossiblebranches,
	 *             min_len_tree, n_threads_getplen, n_slice_size_getplen, tree_bytes */
	void ref_msa_dims(void *hv, long *out)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		out[0] = d->n;
		out[1] = d->m;
		out[2] = d->original_m;
		out[3] = d->nwords;
		out[4] = d->bytes;
		out[5] = d->numberofpossiblebranches;
		out[6] = d->min_len_tree;
		out[7] = d->n_threads_getplen;
		out[8] = d->n_slice_size_getplen;
		out[9] = d->tree_bytes;
	}

	/* re-plan the OpenMP team of getplen for another -p value (MemoryOperations.c:135) */
	void ref_msa_set_threads(void *hv, int nthreads)
	{
		RefHandle *h = (RefHandle *)hv;
		free_memory_to_getplen(&h->todo, &h->sums, &h->runs);
		h->rc.n_processors_available = nthreads;
		calc_distribution_processors(h->msa, h->rc);
		alloc_memory_to_getplen(h->msa, &h->todo, &h->sums, &h->runs);
	}

	/* filtered text row i (after matchange), m chars */
	void ref_msa_row(void *hv, long i, char *out)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		memcpy(out, d->row[i], d->m);
	}

	void ref_dna_to_binary(void *hv, uint64_t *out)
	{
		RefHandle *h = (RefHandle *)hv;
		for (long i = 0; i < h->msa->n; i++)
			memcpy(out + i * h->msa->nwords, h->enc[i], h->msa->bytes);
	}

	long ref_min_tree_length(void *hv) { return MinimumTreeLength(((RefHandle *)hv)->msa); }

	void ref_rinit(int seed) { rinit(seed); }
	double ref_uni(void) { return uni(); }
	long ref_randpint(long upper) { return randpint(upper); }

	void *ref_tree_alloc(void *hv) { return treealloc(((RefHandle *)hv)->msa, LVB_TRUE); }
	void ref_tree_free(void *t) { free(t); }

	void ref_pull_random_tree(void *hv, void *t)
	{
		PullRandomTree(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)t);
	}

	void ref_ss_init(void *hv, void *t)
	{
		RefHandle *h = (RefHandle *)hv;
		ss_init(h->msa, (TREESTACK_TREE_NODES *)t, h->enc);
	}

	long ref_getplen(void *hv, void *t, long root)
	{
		RefHandle *h = (RefHandle *)hv;
		return getplen(h->msa, (TREESTACK_TREE_NODES *)t, h->rc, root, h->todo, h->sums, h->runs);
	}

	void ref_treecopy(void *hv, void *dst, const void *src)
	{
		treecopy(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, LVB_TRUE);
	}

	void ref_mutate_nni(void *hv, void *dst, const void *src, long root)
	{
		mutate_nni(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, root);
	}
	void ref_mutate_spr(void *hv, void *dst, const void *src, long root)
	{
		mutate_spr(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, root);
	}
	void ref_mutate_tbr(void *hv, void *dst, const void *src, long root)
	{
		mutate_tbr(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, root);
	}
	void ref_mutate_deterministic(void *hv, void *dst, const void *src, long root, long p, int left)
	{
		mutate_deterministic(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)dst, (const TREESTACK_TREE_NODES *)src, root, p,
							 left ? LVB_TRUE : LVB_FALSE);
	}
	long ref_arbreroot(void *hv, void *t, long root)
	{
		return arbreroot(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)t, root);
	}
	long ref_lvb_reroot(void *hv, void *t, long oldroot, long newroot)
	{
		return lvb_reroot(((RefHandle *)hv)->msa, (TREESTACK_TREE_NODES *)t, oldroot, newroot, LVB_TRUE);
	}

	/* mark every internal node dirty (same store as TreeOperations.c:631-635) */
	void ref_tree_dirty_all(void *hv, void *tv)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		TREESTACK_TREE_NODES *t = (TREESTACK_TREE_NODES *)tv;
		for (long i = d->n; i < d->numberofpossiblebranches; i++)
			t[i].sitestate[0] = 0U;
	}

	void ref_tree_get(void *hv, const void *tv, long *parent, long *left, long *right, long *changes, int *dirty)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		const TREESTACK_TREE_NODES *t = (const TREESTACK_TREE_NODES *)tv;
		for (long i = 0; i < d->numberofpossiblebranches; i++)
		{
			parent[i] = t[i].parent;
			left[i] = t[i].left;
			right[i] = t[i].right;
			changes[i] = t[i].changes;
			dirty[i] = (t[i].sitestate[0] == 0U);
		}
	}

	/* install a topology (leaf rows must already be there via ref_ss_init) and dirty all internals */
	void ref_tree_set_topology(void *hv, void *tv, const long *parent, const long *left, const long *right)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		TREESTACK_TREE_NODES *t = (TREESTACK_TREE_NODES *)tv;
		for (long i = 0; i < d->numberofpossiblebranches; i++)
		{
			t[i].parent = parent[i];
			t[i].left = left[i];
			t[i].right = right[i];
			t[i].changes = (i < d->n) ? t[i].changes : 0;
			if (i >= d->n)
				t[i].sitestate[0] = 0U;
		}
	}

	/* all B state rows, nwords each, row-major */
	void ref_tree_rows(void *hv, const void *tv, uint64_t *out)
	{
		Dataptr d = ((RefHandle *)hv)->msa;
		const TREESTACK_TREE_NODES *t = (const TREESTACK_TREE_NODES *)tv;
		for (long i = 0; i < d->numberofpossiblebranches; i++)
			memcpy(out + i * d->nwords, t[i].sitestate, d->bytes);
	}

	double ref_starting_temperature(void *hv, const void *tv, long root)
	{
		RefHandle *h = (RefHandle *)hv;
		return StartingTemperature(h->msa, (const TREESTACK_TREE_NODES *)tv, h->rc, root, LVB_FALSE);
	}

	/* deterministic_hillclimb (Solve.c:64-142; switched off in the reference's own GetSoln) on a
	 * fresh tree stack.  out: final length, *current_iter, trees on the stack. */
	void ref_deterministic_hillclimb(void *hv, const void *tv, long root, const char *outtree_path, long *out)
	{
		RefHandle *h = (RefHandle *)hv;
		TREESTACK stack = CreateNewTreestack();
		long iter = 0;
		out[0] = deterministic_hillclimb(h->msa, &stack, (const TREESTACK_TREE_NODES *)tv, h->rc, root, stdout, &iter, LVB_FALSE);
		out[1] = iter;
		out[2] = stack.next;
		if (outtree_path)
		{
			FILE *fp = fopen(outtree_path, "w");
			if (fp)
			{
				PrintTreestack(h->msa, &stack, fp, LVB_FALSE);
				fclose(fp);
			}
		}
		FreeTreestackMemory(h->msa, &stack);
	}

	/* A sequence of pushes into one fresh tree stack: results[i] = what CompareTreeToTreestack (hash == 0,
	 * Treestack.c:231-306) or CompareHashTreeToHashstack (hash != 0, Hash.cpp:49-90) returns for trees[i]
	 * (1 = pushed as new, 0 = a duplicate of a topology already there); returns the final stack size.
	 * Used by tests/test_host_parity.py to pin the repository's own duplicate detection. */
	long ref_treestack_sequence(void *hv, long count, void *const *trees, const long *roots, int hash, long *results)
	{
		RefHandle *h = (RefHandle *)hv;
		TREESTACK stack = CreateNewTreestack();
		for (long i = 0; i < count; i++)
			results[i] = hash ? CompareHashTreeToHashstack(h->msa, &stack, (const TREESTACK_TREE_NODES *)trees[i], roots[i], LVB_FALSE)
			                  : CompareTreeToTreestack(h->msa, &stack, (const TREESTACK_TREE_NODES *)trees[i], roots[i], LVB_FALSE);
		const long size = stack.next;
		FreeTreestackMemory(h->msa, &stack);
		return size;
	}

	/* ---- timing legs for bench.py (cpu_baseline / --impl reference) ----
	 * Full passes: dirty every internal node, call getplen; returns seconds for `reps` calls
	 * and writes the last length to *len_out. */
	double ref_time_full_passes(void *hv, void *tv, long root, int reps, long *len_out)
	{
		RefHandle *h = (RefHandle *)hv;
		struct timespec a, b;
		long len = 0;
		clock_gettime(CLOCK_MONOTONIC, &a);
		for (int r = 0; r < reps; r++)
		{
			ref_tree_dirty_all(hv, tv);
			len = getplen(h->msa, (TREESTACK_TREE_NODES *)tv, h->rc, root, h->todo, h->sums, h->runs);
		}
		clock_gettime(CLOCK_MONOTONIC, &b);
		if (len_out)
			*len_out = len;
		return (b.tv_sec - a.tv_sec) + 1e-9 * (b.tv_nsec - a.tv_nsec);
	}

	/* Proposals the way Anneal/StartingTemperature make them (Solve.c:288-300):
	 * prop <- mutate_{nni|spr|tbr}(cur) then getplen(prop); kind: 0 nni, 1 spr, 2 tbr,
	 * 3 = alternate nni/spr by call index.  `cur` must be fully evaluated.  Returns total
	 * seconds; seconds spent inside getplen alone go to *getplen_s; the lengths to lens[]. */
	double ref_time_proposals(void *hv, void *curv, void *propv, long root, int kind, int count,
							  long *lens, double *getplen_s, long *dirty_total)
	{
		RefHandle *h = (RefHandle *)hv;
		Dataptr d = h->msa;
		TREESTACK_TREE_NODES *cur = (TREESTACK_TREE_NODES *)curv, *prop = (TREESTACK_TREE_NODES *)propv;
		struct timespec a, b, c;
		double total = 0, gp = 0;
		long dirty = 0;
		for (int k = 0; k < count; k++)
		{
			clock_gettime(CLOCK_MONOTONIC, &a);
			int op = (kind == 3) ? (k & 1) : kind;
			if (op == 0)
				mutate_nni(d, prop, cur, root);
			else if (op == 1)
				mutate_spr(d, prop, cur, root);
			else
				mutate_tbr(d, prop, cur, root);
			clock_gettime(CLOCK_MONOTONIC, &b);
			for (long i = d->n; i < d->numberofpossiblebranches; i++)
				dirty += (prop[i].sitestate[0] == 0U);
			long len = getplen(d, prop, h->rc, root, h->todo, h->sums, h->runs);
			clock_gettime(CLOCK_MONOTONIC, &c);
			if (lens)
				lens[k] = len;
			/* the dirty count loop sits between b and the getplen call; it is O(n) flag reads and
			 * is charged to getplen's side, which only makes the baseline slightly slower */
			total += (c.tv_sec - a.tv_sec) + 1e-9 * (c.tv_nsec - a.tv_nsec);
			gp += (c.tv_sec - b.tv_sec) + 1e-9 * (c.tv_nsec - b.tv_nsec);
		}
		if (getplen_s)
			*getplen_s = gp;
		if (dirty_total)
			*dirty_total = dirty;
		return total;
	}

} /* extern "C" */
