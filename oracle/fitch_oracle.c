/* oracle/fitch_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, single-threaded CPU restatement of the reference algorithm for the
 * Fitch tree-length hot path of phylolvb/lvb 4.2.  It exists to check the CUDA
 * path (lvb_b200/csrc) and is itself pinned against the real reference
 * (oracle/_ref/libLVB_ref.so, built from /root/reference by oracle/Makefile) and
 * against the golden vectors under tests/golden/ (tests/test_oracle.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
 * this library.  The product (lvb_b200/) never does: it has no CPU fallback.
 *
 * Reference locations restated here (paths relative to /root/reference):
 *   words_per_row ............. src/DataOperations.c:446-458
 *   DNAToBinary ............... src/DataOperations.c:164-249 (bit values src/LVB.h:73-76)
 *   constchar ................. src/DataOperations.c:272-307
 *   getstatev/MinimumTreeLength src/DataOperations.c:53-105
 *   Fitch word step ........... src/TreeEvaluation.c:219-230 (masks src/LVB.h:88-89)
 *   getplen, serial ........... src/TreeEvaluation.c:182-265
 *   getplen, OpenMP slices .... src/TreeEvaluation.c:64-181
 *   dirty flag ................ src/TreeOperations.c:49-59, 88-103, 631-635
 *
 * Data conventions (same as the reference): a tree has B = 2n-3 nodes; 0..n-1 are
 * leaves (taxon i), n..2n-4 internal; the root is a leaf index whose left/right are
 * its two children; rows is a B x nwords row-major array of 64-bit words holding 16
 * sites each, 4 bits per site; a node is "dirty" iff the first word of its row is 0.
 */
#include "fitch_oracle.h"

#include <stdlib.h>
#include <string.h>

#define SITES_PER_WORD 16
#define LOW3 0x7777777777777777ULL /* MASK_SEVEN, LVB.h:88 */
#define TOP1 0x8888888888888888ULL /* MASK_EIGHT, LVB.h:89 */

long oracle_words_per_row(long m)
{
    /* DataOperations.c:446-458: ceil(m / 16) */
    long w = m / SITES_PER_WORD;
    if (m % SITES_PER_WORD != 0)
        w++;
    return w;
}

/* IUPAC letter -> 4-bit state set, A=1 C=2 G=4 T=8 (LVB.h:73-76); 0 = not allowed.
 * DataOperations.c:192-232.  Upper case only, exactly like the reference ('O' and
 * lower case are rejected there too). */
static unsigned state_set_of(char base)
{
    switch (base)
    {
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 4;
    case 'T': case 'U': return 8;
    case 'Y': return 2 | 8;
    case 'R': return 1 | 4;
    case 'W': return 1 | 8;
    case 'S': return 2 | 4;
    case 'K': return 8 | 4;
    case 'M': return 2 | 1;
    case 'B': return 2 | 4 | 8;
    case 'D': return 1 | 4 | 8;
    case 'H': return 1 | 2 | 8;
    case 'V': return 1 | 2 | 4;
    case 'N': case 'X': case '?': case '-': return 15;
    default: return 0;
    }
}

/* out: n x words_per_row(m), row-major.  Site k of word j sits at bits 4k..4k+3
 * (DataOperations.c:244); sites past m are padded with 'N' = 15 (:187-188).
 * Returns 0, or -1 where the reference would crash("cannot read data MSA") (:235-241). */
int oracle_dna_to_binary(long n, long m, const char *const *rows, uint64_t *out)
{
    const long nwords = oracle_words_per_row(m);
    for (long i = 0; i < n; i++)
    {
        for (long j = 0; j < nwords; j++)
        {
            uint64_t word = 0;
            for (int k = 0; k < SITES_PER_WORD; k++)
            {
                const long site = j * SITES_PER_WORD + k;
                const unsigned s = (site >= m) ? 15u : state_set_of(rows[i][site]);
                if (s == 0)
                    return -1;
                word |= (uint64_t)s << (4 * k);
            }
            out[i * nwords + j] = word;
        }
    }
    return 0;
}

/* keep[k] = 1 where column k is NOT constant, i.e. some row differs from row 0 in the
 * literal character (DataOperations.c:285-296); returns the number of kept columns. */
long oracle_variable_columns(long n, long m, const char *const *rows, unsigned char *keep)
{
    long kept = 0;
    for (long k = 0; k < m; k++)
    {
        keep[k] = 0;
        for (long i = 1; i < n; i++)
        {
            if (rows[i][k] != rows[0][k])
            {
                keep[k] = 1;
                kept++;
                break;
            }
        }
    }
    return kept;
}

/* DataOperations.c:53-105.  Per column: distinct literal characters, not counting
 * '-', '?', 'N', 'X'; a column adds (distinct - 1), or MAXSTATES = 5 when more than 5
 * distinct characters are met.  Quirk kept: a column with no countable character adds
 * strlen("") - 1 = -1. */
long oracle_min_tree_length(long n, long m, const char *const *rows)
{
    long total = 0;
    for (long k = 0; k < m; k++)
    {
        char seen[8];
        int nseen = 0;
        int overflow = 0;
        seen[0] = '\0';
        for (long i = 0; i < n && !overflow; i++)
        {
            const char c = rows[i][k];
            if (memchr(seen, c, (size_t)nseen) == NULL)
            {
                if (c != '-' && c != '?' && c != 'N' && c != 'X')
                    seen[nseen++] = c;
                if (nseen > 5)
                    overflow = 1;
            }
        }
        total += overflow ? 5 : (long)nseen - 1;
    }
    return total;
}

/* One application of the Fitch step to 16 sites (TreeEvaluation.c:219-230).
 * Returns the number of sites whose two state sets do not intersect. */
int oracle_fitch_word(uint64_t x, uint64_t y, uint64_t *out)
{
    const uint64_t both = x & y;
    /* bit 3 of each nibble <- "nibble of `both` is non-zero" */
    uint64_t hit = (((both & LOW3) + LOW3) | both) & TOP1;
    const int changes = SITES_PER_WORD - __builtin_popcountll(hit);
    hit >>= 3;
    /* (hit + 7) ^ 8 per nibble: 0x0 where the sets met, 0xF where they did not */
    *out = both | ((x | y) & ((hit + LOW3) ^ TOP1));
    return changes;
}

static long root_step(long nwords, const uint64_t *l, const uint64_t *r, const uint64_t *rootrow, long from,
                      long to)
{
    /* TreeEvaluation.c:242-264: changes(l,r) + changes(fitch(l,r), root leaf); nothing stored */
    long ch = 0;
    (void)nwords;
    for (long j = from; j < to; j++)
    {
        uint64_t t, dummy;
        ch += oracle_fitch_word(l[j], r[j], &t);
        ch += oracle_fitch_word(t, rootrow[j], &dummy);
    }
    return ch;
}

/* getplen, serial path (TreeEvaluation.c:182-265).
 * Recomputes state rows and per-node change counts of dirty internal nodes in place,
 * sweeping the to-do list in index order until every node has both children clean;
 * clean nodes contribute their cached changes[]; then the root step.  Returns the tree
 * length; -1 if the dirty set can never resolve (malformed tree). */
long oracle_getplen(long n, long nwords, const long *left, const long *right, long root, uint64_t *rows,
                    long *changes)
{
    const long B = 2 * n - 3;
    long *todo = (long *)malloc((size_t)(B - n + 1) * sizeof(long));
    long ntodo = 0, done = 0, total = 0;

    for (long i = n; i < B; i++)
    {
        if (rows[i * nwords] == 0)
        {
            todo[ntodo++] = i;
            changes[i] = 0;
        }
        else
            total += changes[i];
    }
    while (done < ntodo)
    {
        long progressed = 0;
        for (long t = 0; t < ntodo; t++)
        {
            const long node = todo[t];
            uint64_t *dst = rows + node * nwords;
            if (dst[0] != 0)
                continue; /* already recomputed in an earlier sweep */
            const uint64_t *l = rows + left[node] * nwords;
            const uint64_t *r = rows + right[node] * nwords;
            if (l[0] == 0 || r[0] == 0)
                continue; /* a child is still dirty */
            for (long j = 0; j < nwords; j++)
            {
                const int ch = oracle_fitch_word(l[j], r[j], &dst[j]);
                changes[node] += ch;
                total += ch;
            }
            done++;
            progressed++;
        }
        if (!progressed)
        {
            free(todo);
            return -1;
        }
    }
    total += root_step(nwords, rows + left[root] * nwords, rows + right[root] * nwords, rows + root * nwords, 0,
                       nwords);
    free(todo);
    return total;
}

/* getplen, OpenMP path restated serially (TreeEvaluation.c:64-181): slice s owns words
 * [s*W, (s+1)*W) with W = nwords / nslices, the last slice also the remainder (:95-97);
 * every slice walks the whole to-do list with its own done flags (:104-110,130) and
 * leaves per-(node,slice) partial counts that are summed afterwards (:165-179).
 * Must give exactly what oracle_getplen gives. */
long oracle_getplen_sliced(long n, long nwords, const long *left, const long *right, long root, uint64_t *rows,
                           long *changes, int nslices)
{
    const long B = 2 * n - 3, I = B - n;
    if (nslices < 1 || nwords / nslices < 1)
        return -2;
    long *todo = (long *)malloc((size_t)(I + 1) * sizeof(long));
    long *partial = (long *)calloc((size_t)(I + 1) * (size_t)nslices, sizeof(long));
    unsigned char *clean = (unsigned char *)calloc((size_t)I * (size_t)nslices, 1);
    long ntodo = 0, total = 0;
    const long W = nwords / nslices;

    for (long i = n; i < B; i++)
    {
        if (rows[i * nwords] == 0)
            todo[ntodo++] = i;
        else
        {
            total += changes[i];
            for (int s = 0; s < nslices; s++)
                clean[(i - n) * nslices + s] = 1;
        }
    }
    for (int s = 0; s < nslices; s++)
    {
        const long from = W * s;
        const long to = (s == nslices - 1) ? nwords : W * (s + 1);
        long done = 0;
        while (done < ntodo)
        {
            long progressed = 0;
            for (long t = 0; t < ntodo; t++)
            {
                const long node = todo[t];
                if (clean[(node - n) * nslices + s])
                    continue;
                const long l = left[node], r = right[node];
                if (!((l < n || clean[(l - n) * nslices + s]) && (r < n || clean[(r - n) * nslices + s])))
                    continue;
                long ch = 0;
                for (long j = from; j < to; j++)
                    ch += oracle_fitch_word(rows[l * nwords + j], rows[r * nwords + j], &rows[node * nwords + j]);
                partial[t * nslices + s] = ch;
                clean[(node - n) * nslices + s] = 1;
                done++;
                progressed++;
            }
            if (!progressed)
            {
                free(todo); free(partial); free(clean);
                return -1;
            }
        }
        partial[ntodo * nslices + s] = root_step(nwords, rows + left[root] * nwords, rows + right[root] * nwords,
                                                 rows + root * nwords, from, to);
    }
    for (long t = 0; t < ntodo; t++)
    {
        long ch = 0;
        for (int s = 0; s < nslices; s++)
            ch += partial[t * nslices + s];
        changes[todo[t]] = ch;
        total += ch;
    }
    for (int s = 0; s < nslices; s++)
        total += partial[ntodo * nslices + s];
    free(todo); free(partial); free(clean);
    return total;
}

/* what lvb_reroot / ss_init do to the flags (TreeOperations.c:631-635, 1510-1511) */
void oracle_dirty_all(long n, long nwords, uint64_t *rows)
{
    for (long i = n; i < 2 * n - 3; i++)
        rows[i * nwords] = 0;
}
