// ClusterLCP -- drop-in for the reference executable (ClusterLCP.cpp:47-320): same argv, same
// input files (<fasta>.lcp, <fasta>.da), same outputs (<fasta>.<alpha>.clrs in ascending pStart order =
// the reference with threads=1, and the 28-byte <stem>.out), same summary lines.  The OpenMP scan
// (:139-283) is one call into the CUDA library.
#include <algorithm>
#include <vector>
#include "host_util.h"
using namespace limehost;

int main(int argc, char **argv)
{
    Timer total;
    if (argc != 6) {                                                         // :56-60
        std::cerr << "Error usage: " << argv[0] << " fileFasta numReads numGenomes alpha threads" << std::endl;
        exit(1);
    }
    const std::string fileFasta = argv[1];
    uint32_t numReads = 0, numGenomes = 0, alpha = 0;
    int threads = 0;
    sscanf(argv[2], "%u", &numReads);
    sscanf(argv[3], "%u", &numGenomes);
    sscanf(argv[4], "%u", &alpha);
    sscanf(argv[5], "%d", &threads);
    printf("Number of threads: %d (one CUDA device drives the scan)\n", threads);

    const std::string fileOutput = fileFasta + "." + std::to_string(alpha) + ".clrs";   // :90-92
    FILE *out = fopen(fileOutput.c_str(), "w");
    if (!out) { std::cerr << "Error opening " << fileOutput << "."; exit(1); }          // :95-98

    const std::string fnLCP = fileFasta + ".lcp", fnDA = fileFasta + ".da";
    std::cout << "\n\t" << fnLCP;
    uint64_t bl = 0, bd = 0;
    uint32_t *lcp = static_cast<uint32_t *>(read_file_pinned(fnLCP, &bl));
    std::cout << "\n\t" << fnDA;
    uint32_t *da = static_cast<uint32_t *>(read_file_pinned(fnDA, &bd));
    std::cout << std::endl;
    const uint64_t n = bl / sizeof(uint32_t);                                              // :126-127
    if (bd / sizeof(uint32_t) < n) { std::cerr << "Error: " << fnDA << " is shorter than " << fnLCP << "." << std::endl; exit(EXIT_FAILURE); }

    lime_ctx *ctx = nullptr;
    int rc = lime_init(&ctx, device_from_env());
    if (rc) die_lime(nullptr, "lime_init", rc);
    alpha_limit_from_env(ctx, alpha);
    rc = lime_load(ctx, lcp, da, nullptr, n, 0, 0, 1, numReads, numGenomes, LIME_MEM_HOST);
    if (rc) die_lime(ctx, "lime_load", rc);
    uint64_t nClusters = 0, maxLen = 0;
    rc = lime_cluster_detect(ctx, alpha, &nClusters, &maxLen);
    if (rc) die_lime(ctx, "lime_cluster_detect", rc);
    std::vector<lime_cluster> cl(nClusters);
    rc = lime_clusters_copy(ctx, cl.data(), nClusters);
    if (rc) die_lime(ctx, "lime_clusters_copy", rc);
    if (nClusters) fwrite(cl.data(), sizeof(lime_cluster), nClusters, out);              // :229-235
    fclose(out);

    // auxiliary file, field by field without padding (:294-310)
    const std::string fileaux = stem_of(fileFasta) + ".out";
    FILE *aux = fopen(fileaux.c_str(), "w");
    if (!aux) die_open(fileaux);
    fwrite(&numReads, sizeof(uint32_t), 1, aux);
    fwrite(&numGenomes, sizeof(uint32_t), 1, aux);
    fwrite(&alpha, sizeof(uint32_t), 1, aux);
    fwrite(&maxLen, sizeof(uint64_t), 1, aux);
    fwrite(&nClusters, sizeof(uint64_t), 1, aux);
    fclose(aux);

    std::cout << "Clustering process with alpha=" << alpha << " completed.\nTotal number of clusters: " << nClusters
              << ".\nMaximum cluster size: " << maxLen << "." << std::endl;            // :312
    fprintf(stdout, "Time: %.6lf\n", total.seconds());
    lime_destroy(ctx);
    lime_pinned_free(lcp);
    lime_pinned_free(da);
    return 0;
}
