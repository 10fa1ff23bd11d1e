// ClusterBWT_DA -- drop-in for the reference executable (ClusterBWT_DA.cpp:453-773, BIN=1): same
// argv, reads <stem>.out, <fasta>.<alpha>.clrs, <fasta>.da (and <fasta>.ebwt when built with EBWT=1,
// the Makefile variable of the reference), writes <fasta>.res.bin / <fasta>.res.pos.
// clusterAnalyze (:630-670) and clusterChoose (:737) are calls into the CUDA library.
#include <vector>
#include "host_util.h"
using namespace limehost;

#ifndef EBWT
#define EBWT 1
#endif
#ifndef BIN
#define BIN 1          // 0: text output <fasta>.res.txt (reference Makefile:12, ClusterBWT_DA.cpp:715-726)
#endif

int main(int argc, char **argv)
{
    Timer total;
    if (argc != 5) {                                                         // :496-500
        std::cerr << "Error usage " << argv[0] << " fileFasta readLen beta threads" << std::endl;
        exit(1);
    }
    const std::string fileFasta = argv[1];
    unsigned char readLen = 0;                                               // uchar like the reference (:519-521)
    float beta = 0.0f;
    int threads = 0;
    sscanf(argv[2], "%hhu", &readLen);
    sscanf(argv[3], "%f", &beta);
    sscanf(argv[4], "%d", &threads);
    printf("Number of threads: %d (one CUDA device runs the analysis)\n", threads);

    // auxiliary file written by ClusterLCP (:531-553)
    const std::string fileaux = stem_of(fileFasta) + ".out";
    FILE *aux = fopen(fileaux.c_str(), "rb");
    if (!aux) die_open(fileaux);
    uint32_t numRead = 0, numRef = 0, minLCP = 0;
    uint64_t maxLen = 0, nClusters = 0;
    size_t got = 0;
    got += fread(&numRead, sizeof(uint32_t), 1, aux);
    got += fread(&numRef, sizeof(uint32_t), 1, aux);
    got += fread(&minLCP, sizeof(uint32_t), 1, aux);
    got += fread(&maxLen, sizeof(uint64_t), 1, aux);
    got += fread(&nClusters, sizeof(uint64_t), 1, aux);
    fclose(aux);
    if (got != 5) { std::cerr << "Error reading " << fileaux << "." << std::endl; exit(EXIT_FAILURE); }
    std::cout << "numRead: " << numRead << ", numRef: " << numRef << ", minLCP: " << minLCP << ", nClusters: " << nClusters << std::endl;
    const uint32_t norm = (uint32_t)((int)readLen + 1 - (int)minLCP);         // :555
    if (maxLen > LIME_SIZE_MAX_BUF) {                                        // :558-562
        std::cerr << "Error Usage: maximum cluster size is " << maxLen
                  << " greater than sizeMaxBuf, please increase sizeMaxBuf in Tools.h" << std::endl;
        exit(1);
    }

    const std::string fnCluster = fileFasta + "." + std::to_string(minLCP) + ".clrs";   // :572-574
    const std::string fnDA = fileFasta + ".da", fnBWT = fileFasta + ".ebwt";
    uint64_t bc = 0, bd = 0, bb = 0;
    lime_cluster *cl = static_cast<lime_cluster *>(read_file_pinned(fnCluster, &bc));
    uint32_t *da = static_cast<uint32_t *>(read_file_pinned(fnDA, &bd));
    uint8_t *bwt = nullptr;
#if EBWT
    bwt = static_cast<uint8_t *>(read_file_pinned(fnBWT, &bb));
#endif
    const uint64_t n = bd / sizeof(uint32_t);
    if (bc / sizeof(lime_cluster) < nClusters) { std::cerr << "Error: " << fnCluster << " holds fewer than nClusters records." << std::endl; exit(EXIT_FAILURE); }
    if (EBWT && bb < n) { std::cerr << "Error: " << fnBWT << " is shorter than " << fnDA << "." << std::endl; exit(EXIT_FAILURE); }

    std::cerr << "Computing similarity arrays SimArray_i[1,numRead]..." << std::endl;     // :619
    Timer t_an;
    lime_ctx *ctx = nullptr;
    int rc = lime_init(&ctx, device_from_env());
    if (rc) die_lime(nullptr, "lime_init", rc);
    rc = lime_load(ctx, nullptr, da, bwt, n, 0, 0, 1, numRead, numRef, LIME_MEM_HOST);
    if (rc) die_lime(ctx, "lime_load", rc);
    rc = lime_clusters_set(ctx, cl, nClusters, maxLen);
    if (rc) die_lime(ctx, "lime_clusters_set", rc);
    rc = lime_similarity(ctx, EBWT);
    if (rc) die_lime(ctx, "lime_similarity", rc);
    fprintf(stderr, "TIME clusterAnalyze: %.6lf\n", t_an.seconds());           // :683-687

    const std::string fnF = fileFasta + ".res";                                // :701-714
#if BIN
    const std::string sResultPos = fnF + ".pos", sResultBin = fnF + ".bin";
    FILE *fdPos = fopen(sResultPos.c_str(), "wb");
    std::cerr << "Writing " << sResultPos << std::endl;
    FILE *fdBin = fopen(sResultBin.c_str(), "wb");
    std::cerr << "Writing " << sResultBin << std::endl;
    if (!fdPos) die_open(sResultPos);
    if (!fdBin) die_open(sResultBin);
#else
    const std::string sResultTxt = fnF + ".txt";
    {
        FILE *probe = fopen(sResultTxt.c_str(), "w");
        if (!probe) die_open(sResultTxt);
        fclose(probe);
    }
    std::cerr << "Writing " << sResultTxt << std::endl;
#endif
    Timer t_ch;
    uint64_t nRecords = 0;
    rc = lime_choose(ctx, 0, norm, beta, 0, numRead, &nRecords);
    if (rc) die_lime(ctx, "lime_choose", rc);
    std::vector<lime_pair_sim> bin(nRecords);
    std::vector<uint64_t> pos(numRead ? numRead : 1);
    rc = lime_choose_copy(ctx, 0, bin.data(), pos.data());
    if (rc) die_lime(ctx, "lime_choose_copy", rc);
#if BIN
    fwrite(bin.data(), sizeof(lime_pair_sim), nRecords, fdBin);
    if (numRead) fwrite(pos.data(), sizeof(uint64_t), numRead, fdPos);
    fclose(fdBin);
    fclose(fdPos);
    std::cout << "Output (in bytes): " << nRecords << "\t" << sizeof(lime_pair_sim) << std::endl;  // :446
#else
    if (lime_host_write_res_txt(sResultTxt.c_str(), bin.data(), nRecords, pos.data(), numRead) != 0) die_open(sResultTxt);
#endif
    fprintf(stdout, "Time: %.6lf\n", t_ch.seconds());                          // :744

    std::cout << "Cluster analysis completed with beta=" << beta << "." << std::endl;      // :757-758
    std::cout << "Number of clusters: " << nClusters << "." << std::endl;
    fprintf(stdout, "Time: %.6lf\n", total.seconds());
    lime_destroy(ctx);
    lime_pinned_free(cl);
    lime_pinned_free(da);
    if (bwt) lime_pinned_free(bwt);
    return 0;
}
