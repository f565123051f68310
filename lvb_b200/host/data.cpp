/* data.cpp — alignment input, constant-column filter, minimum length and 4-bit packing.
 * Host-side, run once per search; mirrors src/DataOperations.c, src/Wrapper.c and the
 * PHYLIP/FASTA parts of src/MSAInput.cpp of the reference. */
#include <algorithm>
#include <cctype>
#include <cstring>
#include <fstream>
#include <sstream>

#include "lvb_host.h"

namespace lvbhost
{

/* src/DataOperations.c:446-458: 16 four-bit state sets per 64-bit word, rounded up */
long words_per_row(long m) { return (m >> 4) + ((m & 15) ? 1 : 0); }

/* src/DataOperations.c:460-467 */
long bytes_per_row(long m) { return words_per_row(m) * (long)sizeof(Lvb_bit_length); }

/* src/DataOperations.c:53-105.  Per column: the number of distinct LITERAL characters other
 * than '-', '?', 'N', 'X' (partial ambiguity letters count as states of their own), minus one;
 * a column with more than MAXSTATES such characters contributes MAXSTATES; a column with none
 * contributes -1 (the reference's strlen("") - 1). */
long MinimumTreeLength(const Dataptr MSA)
{
    long minlen = 0;
    for (long k = 0; k < MSA->m; ++k)
    {
        bool seen[256] = {false};
        long states = 0;
        bool too_many = false;
        for (long i = 0; i < MSA->n && !too_many; ++i)
        {
            const unsigned char ch = (unsigned char)MSA->row[i][k];
            if (seen[ch] || ch == '-' || ch == '?' || ch == 'N' || ch == 'X')
                continue;
            seen[ch] = true;
            if (++states > MAXSTATES)
                too_many = true;
        }
        minlen += too_many ? (long)MAXSTATES : states - 1;
    }
    return minlen;
}

static void set_shape(Dataptr MSA, long min_len_known)
{
    MSA->bytes = bytes_per_row(MSA->m);
    MSA->nwords = words_per_row(MSA->m);
    MSA->min_len_tree = min_len_known != UNSET ? min_len_known : MinimumTreeLength(MSA);
}

/* src/DataOperations.c:272-405: drop every column in which all rows hold the same character
 * (literal comparison with row 0), then set m, bytes, nwords and min_len_tree */
void matchange(Dataptr MSA, const Parameters &rc)
{
    std::vector<unsigned char> keep((size_t)MSA->m, 0);
    long kept = 0;
    for (long k = 0; k < MSA->m; ++k)
        for (long i = 1; i < MSA->n; ++i)
            if (MSA->row[i][k] != MSA->row[0][k])
            {
                keep[k] = 1;
                ++kept;
                break;
            }
    matchange_apply(MSA, rc, keep.data(), kept, UNSET);
}

/* the rest of matchange once the column scan is known (from the loop above or from
 * lvb_gpu_column_stats): logcut, cutcols, shape.  min_len_after: MinimumTreeLength of the
 * alignment after the cut if already known, else UNSET */
void matchange_apply(Dataptr MSA, const Parameters &rc, const unsigned char *keep, long kept, long min_len_after)
{
    if (rc.verbose)
    {
        printf("Constant columns excluded from analysis: ");
        if (kept == 0)
            printf(" none found.\n");
        else
        { /* logcut(), :407-444: 1-based numbers of the columns that go, eight per line */
            printf("\n");
            long on_line = 0;
            bool newline = false;
            for (long k = 0; k < MSA->m; ++k)
                if (!keep[k])
                {
                    printf("%ld", k + 1L);
                    if (++on_line == 8)
                    {
                        on_line = 0;
                        printf("\n");
                        newline = true;
                    }
                    else
                    {
                        printf("\t");
                        newline = false;
                    }
                }
            if (!newline)
                printf("\n");
        }
    }
    if (kept != MSA->m)
    {
        for (long i = 0; i < MSA->n; ++i)
        {
            std::string cut;
            cut.reserve((size_t)kept);
            const std::string &src = MSA->row[i];
            for (long k = 0; k < MSA->m; ++k)
                if (keep[k])
                    cut.push_back(src[k]);
            MSA->row[i].swap(cut);
        }
        MSA->m = kept;
    }
    if (MSA->m >= MIN_M)
        set_shape(MSA, min_len_after);
    if (MSA->m < MIN_M)
        crash("after constant columns are ignored, data MSA has\n"
              "%ld columns, which is less than LVB's lower limit of\n"
              "%ld columns.\n",
              MSA->m, (long)MIN_M);
    else if (rc.verbose)
        printf("\nIn total, %ld columns are excluded from the analysis\n\n", MSA->original_m - MSA->m);
}

/* 4-bit state set of a base symbol; 0 = not allowed (src/DataOperations.c:192-241; bit values
 * src/LVB.h:73-76) */
static unsigned state_of(char base)
{
    enum
    {
        A = 1,
        C = 2,
        G = 4,
        T = 8
    };
    switch (base)
    {
    case 'A': return A;
    case 'C': return C;
    case 'G': return G;
    case 'T': return T;
    case 'U': return T;
    case 'Y': return C | T;
    case 'R': return A | G;
    case 'W': return A | T;
    case 'S': return C | G;
    case 'K': return T | G;
    case 'M': return C | A;
    case 'B': return C | G | T;
    case 'D': return A | G | T;
    case 'H': return A | C | T;
    case 'V': return A | C | G;
    case 'N': case 'X': case '?': case '-': return A | C | G | T;
    default: return 0;
    }
}

/* src/DataOperations.c:164-249: site k of word j at bits 4k..4k+3, tail padded with 'N' */
void DNAToBinary(const Dataptr MSA, Lvb_bit_length **enc_mat)
{
    for (long i = 0; i < MSA->n; i++)
    {
        const char *text = MSA->row[i].data();
        for (long j = 0; j < MSA->nwords; j++)
        {
            Lvb_bit_length word = 0;
            for (long k = 0; k < 16; k++)
            {
                const long site = (j << 4) + k;
                const char base = site >= MSA->m ? 'N' : text[site];
                const Lvb_bit_length s = state_of(base);
                if (s == 0)
                {
                    fprintf(stderr, "bad base symbol in data MSA: %c\n", base);
                    crash("cannot read data MSA");
                }
                word |= s << (k << 2);
            }
            enc_mat[i][j] = word;
        }
    }
}

/* ---------------- input ---------------- */

static std::string trim(const std::string &s)
{
    size_t a = 0, b = s.size();
    while (a < b && (s[a] == ' ' || s[a] == '\n' || s[a] == '\r'))
        a++;
    while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\n' || s[b - 1] == '\r'))
        b--;
    return s.substr(a, b - a);
}

/* digits and blanks inside PHYLIP sequences are dropped (src/MSAInput.cpp:434-442) */
static void append_bases(std::string &dst, const std::string &chunk)
{
    for (char ch : chunk)
        if (!(ch >= '0' && ch <= '9') && ch != ' ' && ch != '\t')
            dst.push_back(ch);
}

static void finish_matrix(Dataptr MSA, std::vector<std::string> &names, std::vector<std::string> &seqs,
                          const char *file_name)
{
    static const char accept[] = "ACGTUYRWSKMBDHVNX?O-"; /* src/MSAInput.cpp:64 */
    if (seqs.size() < 2)
        crash("%s sequences were read from the file: %s", seqs.empty() ? "Zero" : "Only one", file_name);
    const size_t m = seqs[0].size();
    for (size_t i = 0; i < seqs.size(); i++)
    {
        if (seqs[i].size() != m)
            crash("Something wrong with the file.\nThe sequence lengths are different in the file: %s", file_name);
        for (char &ch : seqs[i])
        {
            ch = (char)toupper((unsigned char)ch);
            if (!strchr(accept, ch) || ch == '\0')
                crash("This char is not allowed (%c)\nThe char is in this line: %s", ch, names[i].c_str());
        }
    }
    MSA->n = (long)seqs.size();
    MSA->m = MSA->original_m = (long)m;
    MSA->numberofpossiblebranches = 2 * MSA->n - 3; /* brcnt(), src/CommandLineParser.cpp:129-132 */
    MSA->nsets = MSA->n - 3;
    MSA->mssz = MSA->n - 2;
    MSA->max_length_seq_name = 0;
    for (const auto &nm : names)
        MSA->max_length_seq_name = std::max<long>(MSA->max_length_seq_name, (long)nm.size());
    MSA->row.swap(seqs);
    MSA->rowtitle.swap(names);
}

static void read_phylip(const char *file_name, std::vector<std::string> &names, std::vector<std::string> &seqs)
{
    std::ifstream in(file_name);
    if (!in)
        crash("Failed to open alignment file: %s", file_name);
    std::string line;
    long n = 0, m = 0;
    std::vector<std::string> lines;
    while (std::getline(in, line))
    {
        line = trim(line);
        if (line.empty())
            continue;
        if (n == 0 && m == 0)
        {
            int a = 0, b = 0;
            if (sscanf(line.c_str(), "%d%d", &a, &b) == 2)
            {
                n = a;
                m = b;
            }
            continue;
        }
        lines.push_back(line);
    }
    if (n <= 0 || m <= 0)
        crash("Some problem reading the file. Please, check the file format.");
    const size_t name_len = 10; /* PHYLIP nmlngth, src/MSAInput.cpp:62 */
    auto split_named = [&](const std::string &l, std::string &nm, std::string &sq) {
        if (l.size() < name_len)
            crash("Failed to open alignment file: %s\nThis line has less than %d chars...", file_name, (int)name_len);
        nm = trim(l.substr(0, name_len));
        sq.clear();
        append_bases(sq, l.substr(name_len));
    };
    /* interleaved (also covers one line per taxon): n named lines, then blocks of n continuation lines */
    if (lines.size() >= (size_t)n && lines.size() % (size_t)n == 0)
    {
        std::vector<std::string> nm((size_t)n), sq((size_t)n);
        for (long i = 0; i < n; i++)
            split_named(lines[(size_t)i], nm[(size_t)i], sq[(size_t)i]);
        for (size_t k = (size_t)n; k < lines.size(); k++)
            append_bases(sq[k % (size_t)n], lines[k]);
        bool ok = true;
        for (long i = 0; i < n; i++)
            ok = ok && (long)sq[(size_t)i].size() == m;
        if (ok)
        {
            names.swap(nm);
            seqs.swap(sq);
            return;
        }
    }
    /* sequential: a named line, then continuation lines until m bases are read */
    {
        std::vector<std::string> nm, sq;
        size_t k = 0;
        while (k < lines.size() && (long)nm.size() < n)
        {
            std::string a, b;
            split_named(lines[k++], a, b);
            while ((long)b.size() < m && k < lines.size())
                append_bases(b, lines[k++]);
            nm.push_back(a);
            sq.push_back(b);
        }
        if ((long)nm.size() != n || k != lines.size())
            crash("The file has a different number of sequences.\nRead: %d\nIn the header: %ld", (int)nm.size(), n);
        for (long i = 0; i < n; i++)
            if ((long)sq[(size_t)i].size() != m)
                crash("This sequence %s has a different length %d from the one read in the header: %ld",
                      nm[(size_t)i].c_str(), (int)sq[(size_t)i].size(), m);
        names.swap(nm);
        seqs.swap(sq);
    }
}

static void read_fasta(const char *file_name, std::vector<std::string> &names, std::vector<std::string> &seqs)
{
    std::ifstream in(file_name);
    if (!in)
        crash("Failed to open alignment file: %s", file_name);
    std::string line;
    while (std::getline(in, line))
    {
        line = trim(line);
        if (line.empty())
            continue;
        if (line[0] == '>')
        {
            names.push_back(trim(line.substr(1)));
            seqs.emplace_back();
        }
        else if (!seqs.empty())
            append_bases(seqs.back(), line);
    }
}

/* src/Wrapper.c:49-68 */
void phylip_dna_matrin(const char *file_name, int n_file_format, Dataptr MSA)
{
    std::vector<std::string> names, seqs;
    if (n_file_format == FORMAT_PHYLIP)
        read_phylip(file_name, names, seqs);
    else if (n_file_format == FORMAT_FASTA)
        read_fasta(file_name, names, seqs);
    else
        crash("this build reads PHYLIP and FASTA alignments only (NEXUS and CLUSTAL input is outside the "
              "accelerated path)");
    finish_matrix(MSA, names, seqs, file_name);
    if (MSA->n < MIN_N)
        crash("The data matrix must have at least %ld sequences.", (long)MIN_N);
    else if (MSA->n > MAX_N)
        crash("The data matrix must have no more than %ld sequences.", (long)MAX_N);
    else if (MSA->m < MIN_M)
        crash("The data matrix must have at least %ld sites.", (long)MIN_M);
    else if (MSA->m > MAX_M)
        crash("The data matrix must have no more than %ld sites.", (long)MAX_M);
}

void matrix_from_rows(Dataptr MSA, long n, long m, const char *const *rows, const char *const *titles)
{
    std::vector<std::string> names, seqs;
    for (long i = 0; i < n; i++)
    {
        seqs.emplace_back(rows[i], (size_t)m);
        names.push_back(titles ? std::string(titles[i]) : "T" + std::to_string(i));
    }
    finish_matrix(MSA, names, seqs, "<memory>");
    if (MSA->n < MIN_N)
        crash("The data matrix must have at least %ld sequences.", (long)MIN_N);
}

} // namespace lvbhost
