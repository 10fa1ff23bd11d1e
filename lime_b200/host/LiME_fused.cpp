// LiME_fused -- the whole of LiME_paired.sh in one process: for every strand
// ClusterLCP + clusterAnalyze as ONE pass over <fasta>.lcp/.da/.ebwt (lime_cluster_similarity),
// clusterChoose on the device (lime_choose), then Classify over the device-resident strand rows --
// no .clrs / .res.bin / .res.pos round trips.  The intermediate files are written only with --keep,
// byte for byte what ClusterLCP / ClusterBWT_DA would have left behind.
//
//   LiME_fused N File_1 .. File_N output numReads numGenomes LineageFile readLen threads [alpha beta TaxRank] [--keep]
//
// N = 2 (single-end: F, RC) or 4 (paired-end: File1_F File1_RC File2_F File2_RC), the argument order and
// the defaults alpha=16 beta=0.25 TaxRank=1 are those of LiME_paired.sh:9-19.  While the GPU works on
// strand i the host reads the files of strand i+1 into page-locked memory.
// Build variants as in the reference Makefile (:13-14): -DEBWT=0|1, -DHIGHER=0|1.
#include <cstring>
#include <fstream>
#include <future>
#include <vector>
#include "host_util.h"
using namespace limehost;

#ifndef EBWT
#define EBWT 1
#endif
#ifndef HIGHER
#define HIGHER 0
#endif

static const char *USAGE =
    " N File_1 ... File_N output numReads numGenomes LineageFile readLen threads [alpha beta TaxRank] [--keep]";

struct StrandFiles {
    uint32_t *lcp = nullptr, *da = nullptr;
    uint8_t *bwt = nullptr;
    uint64_t n = 0;
};

static StrandFiles read_strand(const std::string &fasta)
{
    StrandFiles s;
    uint64_t bl = 0, bd = 0, bb = 0;
    s.lcp = static_cast<uint32_t *>(read_file_pinned(fasta + ".lcp", &bl));
    s.da = static_cast<uint32_t *>(read_file_pinned(fasta + ".da", &bd));
    s.n = bl / sizeof(uint32_t);
    if (bd / sizeof(uint32_t) < s.n) { std::cerr << "Error: " << fasta << ".da is shorter than " << fasta << ".lcp." << std::endl; exit(EXIT_FAILURE); }
#if EBWT
    s.bwt = static_cast<uint8_t *>(read_file_pinned(fasta + ".ebwt", &bb));
    if (bb < s.n) { std::cerr << "Error: " << fasta << ".ebwt is shorter than " << fasta << ".da." << std::endl; exit(EXIT_FAILURE); }
#endif
    (void)bb;
    return s;
}

static void free_strand(StrandFiles &s)
{
    lime_pinned_free(s.lcp); lime_pinned_free(s.da);
    if (s.bwt) lime_pinned_free(s.bwt);
    s = StrandFiles();
}

static void write_or_die(const std::string &path, const void *p, size_t bytes)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) die_open(path);
    if (bytes && fwrite(p, 1, bytes, f) != bytes) { std::cerr << "Error writing " << path << "." << std::endl; exit(EXIT_FAILURE); }
    fclose(f);
}

int main(int argc, char **argv)
{
    Timer total;
    bool keep = false;
    std::vector<char *> a;
    for (int i = 0; i < argc; i++) {
        if (i > 0 && strcmp(argv[i], "--keep") == 0) keep = true;
        else a.push_back(argv[i]);
    }
    const int ac = (int)a.size();
    if (ac < 2) { std::cerr << "Error usage " << a[0] << USAGE << std::endl; exit(1); }
    unsigned numFile = 0;
    sscanf(a[1], "%u", &numFile);
    if (numFile != 2 && numFile != 4) {                                     // Classify.cpp:334-338
        std::cerr << "Error usage " << a[0]
                  << ": the allowed number of input files is 2 (single-end reads), or 4 (paired-end reads)" << std::endl;
        exit(1);
    }
    if (ac != (int)numFile + 8 && ac != (int)numFile + 11) { std::cerr << "Error usage " << a[0] << USAGE << std::endl; exit(1); }
    std::vector<std::string> fasta;
    for (unsigned i = 0; i < numFile; i++) fasta.push_back(a[2 + i]);
    const std::string fileOutput = a[numFile + 2];
    uint32_t numReads = 0, numGenomes = 0, alpha = 16;                      // LiME_paired.sh:9-11
    unsigned char readLen = 0;                                              // uchar, ClusterBWT_DA.cpp:519-521
    float beta = 0.25f;
    int taxRank = 1, threads = 0;
    sscanf(a[numFile + 3], "%u", &numReads);
    sscanf(a[numFile + 4], "%u", &numGenomes);
    const std::string fileTaxID = a[numFile + 5];
    sscanf(a[numFile + 6], "%hhu", &readLen);
    sscanf(a[numFile + 7], "%d", &threads);
    if (ac == (int)numFile + 11) {
        sscanf(a[numFile + 8], "%u", &alpha);
        sscanf(a[numFile + 9], "%f", &beta);
        sscanf(a[numFile + 10], "%d", &taxRank);
    }
    if (taxRank > LIME_RANKS || taxRank < 0) {
        std::cerr << "Error usage: taxRank 0=Genome, 1=Species, 2=Genus, 3=Family, 4=Order, 5=Class, 6=Phylum." << std::endl;
        exit(1);
    }
    if (HIGHER && taxRank == 0) { std::cerr << "Error usage: the HIGHER variant needs taxRank >= 1." << std::endl; exit(1); }
    printf("Number of threads: %d (one CUDA device runs all steps)\n", threads);

    // lineage first: a bad taxonomy file should fail before any GPU work (Classify.cpp:378-398)
    std::cout << "Reading " << fileTaxID << std::endl;
    std::vector<uint32_t> v(numGenomes ? numGenomes : 1), m;
    if (HIGHER) m.assign((size_t)LIME_RANKS * (numGenomes ? numGenomes : 1), 0);
    const int64_t found = lime_host_fix_rank(fileTaxID.c_str(), taxRank, numGenomes, v.data(), numGenomes,
                                             HIGHER ? m.data() : nullptr);
    if (found < 0) { std::cerr << "Error opening file " << fileTaxID << ".\n"; exit(1); }
    if ((uint64_t)found != numGenomes) {
        std::cerr << "Number of taxIDs = " << found << " lower than genome number: poor taxonomy information to classify." << std::endl;
        exit(1);
    }

    lime_ctx *ctx = nullptr;
    int rc = lime_init(&ctx, device_from_env());
    if (rc) die_lime(nullptr, "lime_init", rc);
    alpha_limit_from_env(ctx, alpha);
    const uint32_t norm = (uint32_t)((int)readLen + 1 - (int)alpha);          // ClusterBWT_DA.cpp:555

    // ---- steps 1 + 2 per strand; the next strand's files are read while the GPU works ----------------
    std::future<StrandFiles> next = std::async(std::launch::async, read_strand, fasta[0]);
    for (unsigned i = 0; i < numFile; i++) {
        Timer t_s;
        StrandFiles s = next.get();
        if (i + 1 < numFile) next = std::async(std::launch::async, read_strand, fasta[i + 1]);
        rc = lime_load(ctx, s.lcp, s.da, s.bwt, s.n, 0, 0, 1, numReads, numGenomes, LIME_MEM_HOST);
        if (rc) die_lime(ctx, "lime_load", rc);
        uint64_t nClusters = 0, maxLen = 0;
        rc = lime_cluster_similarity(ctx, alpha, EBWT, keep ? 1 : 0, &nClusters, &maxLen);
        if (rc == LIME_ERR_MAXLEN) {                                        // ClusterBWT_DA.cpp:558-562
            std::cerr << "Error Usage: maximum cluster size is greater than sizeMaxBuf, please increase sizeMaxBuf in Tools.h" << std::endl;
            exit(1);
        }
        if (rc) die_lime(ctx, "lime_cluster_similarity", rc);
        std::cout << fasta[i] << ": clustering with alpha=" << alpha << " completed.\nTotal number of clusters: " << nClusters
                  << ".\nMaximum cluster size: " << maxLen << "." << std::endl;            // ClusterLCP.cpp:312
        uint64_t nRecords = 0;
        rc = lime_choose(ctx, (int)i, norm, beta, 0, numReads, &nRecords);
        if (rc) die_lime(ctx, "lime_choose", rc);
        std::cout << "Output (in bytes): " << nRecords << "\t" << sizeof(lime_pair_sim) << std::endl;  // ClusterBWT_DA.cpp:446
        if (keep) {
            // what ClusterLCP (:229-235, 294-310) and ClusterBWT_DA (:701-714) leave on disk
            std::vector<lime_cluster> cl(nClusters);
            rc = lime_clusters_copy(ctx, cl.data(), nClusters);
            if (rc) die_lime(ctx, "lime_clusters_copy", rc);
            write_or_die(fasta[i] + "." + std::to_string(alpha) + ".clrs", cl.data(), nClusters * sizeof(lime_cluster));
            const std::string fileaux = stem_of(fasta[i]) + ".out";
            FILE *aux = fopen(fileaux.c_str(), "w");
            if (!aux) die_open(fileaux);
            fwrite(&numReads, sizeof(uint32_t), 1, aux);
            fwrite(&numGenomes, sizeof(uint32_t), 1, aux);
            fwrite(&alpha, sizeof(uint32_t), 1, aux);
            fwrite(&maxLen, sizeof(uint64_t), 1, aux);
            fwrite(&nClusters, sizeof(uint64_t), 1, aux);
            fclose(aux);
            std::vector<lime_pair_sim> bin(nRecords);
            std::vector<uint64_t> pos(numReads ? numReads : 1);
            rc = lime_choose_copy(ctx, (int)i, bin.data(), pos.data());
            if (rc) die_lime(ctx, "lime_choose_copy", rc);
            write_or_die(fasta[i] + ".res.bin", bin.data(), nRecords * sizeof(lime_pair_sim));
            write_or_die(fasta[i] + ".res.pos", pos.data(), (size_t)numReads * sizeof(uint64_t));
        }
        free_strand(s);
        fprintf(stderr, "TIME strand %u: %.6lf\n", i, t_s.seconds());
    }

    // ---- step 3 over the device-resident rows (Classify.cpp:465-743) ---------------------------------
    std::cerr << "Start comparing..." << std::endl;
    std::vector<uint8_t> type(numReads ? numReads : 1);
    std::vector<uint32_t> tax(numReads ? numReads : 1);
    std::vector<float> sim(numReads ? numReads : 1);
    rc = lime_classify(ctx, (int)numFile, numGenomes, v.data(), HIGHER ? m.data() : nullptr, taxRank, HIGHER,
                       type.data(), tax.data(), sim.data());
    if (rc) die_lime(ctx, "lime_classify", rc);
    if (lime_host_write_csv(fileOutput.c_str(), numReads, type.data(), tax.data(), sim.data()) != 0)
        std::cerr << "ERROR: File Output not Open" << std::endl;
    uint64_t numC = 0, numA = 0, numNC = 0, numH = 0;
    for (uint32_t r = 0; r < numReads; r++) {
        if (type[r] == 'C') numC++; else if (type[r] == 'A') numA++; else if (type[r] == 'H') numH++; else numNC++;
    }
    std::cout << "Classification process at level " << taxRank << " completed.\nNumber of successfully classified reads: "
              << numC << "/" << numReads << ";" << std::endl;
#if HIGHER
    std::cout << "\tClassified at higher taxonomic ranks: " << numH << "." << std::endl;
#endif
    std::cout << "\tAmbiguously classified reads: " << numA << "." << std::endl;
    std::cout << "\tNot classified reads: " << numNC << "." << std::endl;
    fprintf(stdout, "Time: %.6lf\n", total.seconds());
    lime_destroy(ctx);
    return 0;
}
