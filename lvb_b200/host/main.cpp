/* main.cpp — the `lvb` command line on the B200 engine (src/Main.c:60-155, src/SearchParameters.c,
 * src/CommandLineParser.cpp:206-338 of the reference).  Same flags, defaults, file names and
 * result block; the lines the reference's tests look for ("Tree score:", "Topologies
 * recovered:", "PThreads:") are printed in the same form.  Additions of this build:
 *   -d <int>  CUDA device            -R <int>  independent SA replicates (seeds s, s+1, ...)
 *   -K <int>  proposals scored per launch and chain (0 = adaptive)
 *   -x <int>  stop after this many rearrangements (benchmarking)
 *   -G <int>  device contexts / driver threads the replicates are spread over (0 = auto)
 *   -D <int>  number of GPUs (device, device+1, ...) the replicates are spread over
 */
#include <ctime>
#include <getopt.h>
#include <omp.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

#include "lvb_host.h"

using namespace lvbhost;

static void usage()
{
    printf(" Usage: lvb -i <alignment> [options]\n\n");
    printf(" Help options: \n");
    printf("    Display program instructions      -h \n");
    printf("    Display program instructions      -? \n\n");
    printf(" General Preferences: \n");
    printf("   Verbose mode                       -v [ON|OFF]         Engage verbose mode; default: OFF\n\n");
    printf(" Input/Output Preferences: \n");
    printf("    Input file                        -i [FILE]           Input file name \n");
    printf("    Input file format                 -f [STRING]         'phylip'|'fasta' \n");
    printf("    Output file                       -o [FILE]           Output file name; default: 'outtree'\n\n");
    printf(" Search Origin Preferences: \n");
    printf("    Starting seed                     -s [VALUE]          Specify a starting seed; default: generated from system clock\n\n");
    printf(" Topological Rearrangement Preferences: \n");
    printf("     Rearrangement algorithm          -a [VALUE]          Iterations of NNI + SPR (0)\n"
           "                                                          TBR followed by iterations of NNI + SPR (1)\n"
           "                                                          Pseudo counts determining probability of NNI, SPR, or TBR iterations (2)\n\n");
    printf(" Simulated Annealing Preferences: \n");
    printf("    Cooling schedule                  -c [G|L]            SA cooling schedule; GEOMETRIC (G) or LINEAR (L); default: G \n\n");
    printf(" GPU Preferences: \n");
    printf("    CUDA device                       -d [VALUE]          default: 0\n");
    printf("    Independent SA replicates         -R [VALUE]          seeds s, s+1, ...; default: 1\n");
    printf("    Proposals per launch and chain    -K [VALUE]          default: adaptive\n");
    printf("    Driver threads / device contexts  -G [VALUE]          default: about one per 8 replicates, at most 4 per GPU\n");
    printf("    GPUs used for the replicates      -D [VALUE]          devices d, d+1, ...; default: 1\n");
    printf("    Maximum trees kept                -t [VALUE]\n\n");
    exit(0);
}

/* src/SearchParameters.c:52-99 + src/CommandLineParser.cpp:206-338 */
static void getparam(Parameters &p, int argc, char **argv)
{
    const time_t tim = time(nullptr);
    p.seed = (int)((unsigned long)tim % 900000001UL);
    p.n_processors_available = omp_get_max_threads(); /* SearchParameters.c:86: all cores unless -p */
    int c;
    opterr = 0;
    while ((c = getopt(argc, argv, "t:c:b:vs:i:o:f:a:p:N:SC:h?d:R:K:x:G:D:")) != -1)
    {
        switch (c)
        {
        case 'c':
            if (!strcmp(optarg, "g") || !strcmp(optarg, "G"))
                p.cooling_schedule = 0;
            else if (!strcmp(optarg, "l") || !strcmp(optarg, "L"))
                p.cooling_schedule = 1;
            else
                fprintf(stderr, "Unknown cooling schedule option\nPlease, choose between Geometric (g) or Linear (l).");
            break;
        case 'a':
            if (!strcmp(optarg, "0"))
                p.algorithm_selection = 0;
            else if (!strcmp(optarg, "1"))
                p.algorithm_selection = 1;
            else if (!strcmp(optarg, "2"))
                p.algorithm_selection = 2;
            else
                fprintf(stderr, "Unknown algorithm option\nPlease, choose between SN (0), SEQ-TNS (1), or PBS (2).");
            break;
        case 'v': p.verbose = 1; break;
        case 's': p.seed = atoi(optarg); break;
        case 'i': p.file_name_in = optarg; break;
        case 'o': p.file_name_out = optarg; break;
        case 'f':
            if (!strcmp(optarg, "phylip"))
                p.n_file_format = FORMAT_PHYLIP;
            else if (!strcmp(optarg, "fasta"))
                p.n_file_format = FORMAT_FASTA;
            else if (!strcmp(optarg, "nexus"))
                p.n_file_format = FORMAT_NEXUS;
            else if (!strcmp(optarg, "clustal"))
                p.n_file_format = FORMAT_CLUSTAL;
            else
                fprintf(stderr, "Unknown file format.");
            break;
        case 'p':
            p.n_processors_available = atoi(optarg);
            if (p.n_processors_available < 1)
                p.n_processors_available = omp_get_max_threads(); /* SearchParameters.c:86: all cores unless -p */
            break;
        case 't':
            p.n_number_max_trees = atoi(optarg);
            if (p.n_number_max_trees < 1)
                p.n_number_max_trees = 0;
            break;
        case 'd': p.device = atoi(optarg); break;
        case 'R': p.replicates = atoi(optarg) < 1 ? 1 : atoi(optarg); break;
        case 'K': p.speculation = atoi(optarg) < 0 ? 0 : atoi(optarg); break;
        case 'x': p.max_rearrangements = atol(optarg); break;
        case 'G': p.groups = atoi(optarg) < 0 ? 0 : atoi(optarg); break;
        case 'D': p.devices = atoi(optarg) < 1 ? 1 : atoi(optarg); break;
        default: usage();
        }
    }
}

/* src/SearchParameters.c:101-170 */
static void writeinf(const Parameters &p, Dataptr MSA, int argc, char **argv)
{
    printf("Executing: ' ");
    for (int i = 0; i < argc; ++i)
        printf("%s ", argv[i]);
    char buf[26];
    const time_t now = time(nullptr);
    strftime(buf, sizeof buf, "%H:%M", localtime(&now));
    printf("' at: %s\n\n", buf);
    printf("Analysis Properties: \n");
    printf("  Alignment:          '%s'\n", p.file_name_in.c_str());
    printf("  MSA format:          %s\n", p.n_file_format == FORMAT_PHYLIP ? "PHYLIP" : "FASTA");
    printf("  MSA size:            %ld x %ld\n", MSA->n, MSA->original_m);
    printf("  Seed:                %d\n", p.seed);
    printf("  Cooling schedule:    %s\n", p.cooling_schedule == 0 ? "GEOMETRIC" : "LINEAR");
    printf("  Algorithm: ");
    if (p.algorithm_selection == 0)
        printf("          0 (SN)\n");
    else if (p.algorithm_selection == 1)
        printf("          1 (SEQ-TNS)\n");
    else
        printf("          2 (PBS)\n");
    printf("\nParallelisation Properties: \n");
    printf("  PThreads:            %d\n", p.n_processors_available); /* host threads: chains are advanced in parallel; scoring runs on the GPU */
    printf("  GPU:                 device %d (+%d more), %d SA replicate(s)\n", p.device, p.devices - 1, p.replicates);
    printf("\n================================================================================\n");
    printf("\nInitialising search: \n");
}

int main(int argc, char **argv)
{
    Parameters rc;
    const clock_t start = clock();
    printf("LVB (B200 tree-length engine; search behaviour of LVB v.4.2)\n\n");
    getparam(rc, argc, argv);
    {
        char buf[26];
        const time_t now = time(nullptr);
        strftime(buf, sizeof buf, "%d/%m/%Y\n", localtime(&now));
        printf("Start date: %s\n", buf);
    }
    DataStructure MSA;
    phylip_dna_matrin(rc.file_name_in.c_str(), rc.n_file_format, &MSA);
    {
        /* the two text scans of matchange are O(n*m): on the device once the alignment is large
         * (LVB_GPU_MATCHANGE=1 / 0 forces the choice) */
        const char *mc = getenv("LVB_GPU_MATCHANGE");
        const bool on_device = mc ? atoi(mc) != 0 : (double)MSA.n * (double)MSA.m >= 64.0 * 1024 * 1024;
        if (on_device)
            matchange_device(&MSA, rc, rc.device);
        else
            matchange(&MSA, rc);
    }
    writeinf(rc, &MSA, argc, argv);
    if (rc.verbose)
        printf("MinimumTreeLength: %ld\n\n", MinimumTreeLength(&MSA));

    FILE *outtreefp = fopen(rc.file_name_out.c_str(), "w");
    if (!outtreefp)
        crash("cannot open file '%s'", rc.file_name_out.c_str());
    FILE *tsv = rc.verbose ? fopen("changeAccepted.tsv", "w") : nullptr;

    Device dev;
    device_open(dev, rc.device, &MSA, rc.replicates);
    std::vector<TREESTACK> stacks;
    std::vector<SearchResult> results;
    GetSolnBatched(&MSA, rc, dev, rc.replicates, stacks, results, nullptr, tsv);
    if (tsv)
        fclose(tsv);
    const long trees_output = PrintTreestack(&MSA, &stacks[0], outtreefp);
    fclose(outtreefp);
    for (int r = 1; r < rc.replicates; r++)
    {
        const std::string nm = rc.file_name_out + ".r" + std::to_string(r);
        FILE *fp = fopen(nm.c_str(), "w");
        if (!fp)
            crash("cannot open file '%s'", nm.c_str());
        PrintTreestack(&MSA, &stacks[(size_t)r], fp);
        fclose(fp);
    }
    printf("--------------------------------------------------------\n");
    const double secs = ((double)(clock() - start)) / CLOCKS_PER_SEC;

    /* src/LogFile.c:56-74 */
    {
        struct stat st;
        const bool exists = stat("logfile.tsv", &st) == 0;
        FILE *lf = fopen("logfile.tsv", "a+");
        if (lf)
        {
            if (!exists)
                fprintf(lf, "Implementation\tRearrangements\tTopologies\tScore\tRuntime\n");
            fprintf(lf, "%s\t%ld\t%ld\t%ld\t%.2lf\n", "B200", results[0].rearrangements, trees_output,
                    results[0].final_length, secs);
            fclose(lf);
        }
    }
    /* src/Print.cpp:117-130 */
    const double ci = (double)MinimumTreeLength(&MSA) / results[0].final_length;
    printf("\nSearch Complete\n");
    printf("\n================================================================================\n");
    printf("\nSearch Results:\n");
    printf("  Rearrangements evaluated: %ld\n", results[0].rearrangements);
    printf("  Topologies recovered:     %ld\n", trees_output);
    printf("  Tree score:               %ld\n", results[0].final_length);
    printf("  Consistency index:        %.2lf\n", ci);
    printf("  Homoplasy index:          %.2lf\n", 1 - ci);
    printf("  Total runtime (seconds):  %.2lf\n", secs);
    for (int r = 1; r < rc.replicates; r++)
        printf("  Replicate %d (seed %d): score %ld, %ld topologies, %ld rearrangements\n", r, rc.seed + r,
               results[(size_t)r].final_length, (long)stacks[(size_t)r].next, results[(size_t)r].rearrangements);
    long all_scored = 0;
    for (int r = 0; r < rc.replicates; r++)
        all_scored += results[(size_t)r].getplen_calls;
    printf("  Wall time of the search:  %.3lf s; trees scored: %ld in %ld launches; kernel launches: %ld\n",
           results[0].seconds, results[0].getplen_calls, results[0].launches_scored, (long)dev.launches());
    printf("  All replicates:           %ld trees scored, %.0f per second\n", all_scored,
           results[0].seconds > 0 ? all_scored / results[0].seconds : 0.0);
    printf("  Host time split (s):      propose %.3f, commit calls %.3f, score calls %.3f, replay %.3f; %ld rounds; %d group(s), set up in %.3f s\n",
           results[0].prof[0], results[0].prof[1], results[0].prof[2], results[0].prof[3], results[0].rounds,
           results[0].groups, results[0].setup_seconds);
    printf("\nAll topologies written to '%s'\n", rc.file_name_out.c_str());
    device_close(dev);
    return fflush(stdout) == 0 ? EXIT_SUCCESS : EXIT_FAILURE;
}
