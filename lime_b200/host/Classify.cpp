// Classify -- drop-in for the reference executable (Classify.cpp:304-751, BIN=1): same argv, reads
// <in_i>.bin / <in_i>.pos of 2 or 4 strands and the lineage file, writes the CSV.  The per-read loop
// (:474-702) is one call into the CUDA library; FixRank (:33-89) and the print loop (:722-734) are
// the host helpers of lime_host.h.  Build with -DHIGHER=1 for the HIGHER variant (:262-302).
#include <fstream>
#include <vector>
#include "host_util.h"
using namespace limehost;

#ifndef HIGHER
#define HIGHER 0
#endif
#ifndef BIN
#define BIN 1          // 0: text inputs <in_i>.txt (reference Makefile:12, Classify.cpp:430-444)
#endif

static const char *USAGE = " N fileInput1 fileInput2 ... fileInputN numReads numGenomes fileOutput fileTaxo taxRank numThreads";

int main(int argc, char **argv)
{
    Timer total;
    if (argc < 2) { std::cerr << "Error usage " << argv[0] << USAGE << std::endl; exit(1); }      // :321-324
    unsigned numFile = 0;
    sscanf(argv[1], "%u", &numFile);
    if (argc != (int)numFile + 8) { std::cerr << "Error usage " << argv[0] << USAGE << std::endl; exit(1); }
    if (numFile != 2 && numFile != 4) {                                                             // :334-338
        std::cerr << "Error usage " << argv[0]
                  << ": the allowed number of input files is 2 (single-end reads), or 4 (paired-end reads)" << std::endl;
        exit(1);
    }
    std::vector<std::string> fileInput;
    for (unsigned i = 2; i < numFile + 2; i++) fileInput.push_back(argv[i]);
    uint32_t numReads = 0, numTarg = 0;
    sscanf(argv[numFile + 2], "%u", &numReads);
    sscanf(argv[numFile + 3], "%u", &numTarg);
    const std::string fileOutput = argv[numFile + 4], fileTaxID = argv[numFile + 5];
    int taxRank = 0, numThreads = 0;
    sscanf(argv[numFile + 6], "%d", &taxRank);
    if (taxRank > LIME_RANKS || taxRank < 0) {
        // the reference prints this and carries on into an out-of-range column (:358-359); we stop
        std::cerr << "Error usage: taxRank 0=Genome, 1=Species, 2=Genus, 3=Family, 4=Order, 5=Class, 6=Phylum." << std::endl;
        exit(1);
    }
    if (HIGHER && taxRank == 0) {   // the reference indexes m_corRef[-1] here (:62,78,674): unsupported
        std::cerr << "Error usage: the HIGHER variant needs taxRank >= 1." << std::endl;
        exit(1);
    }
    sscanf(argv[numFile + 7], "%d", &numThreads);
    printf("Number of threads: %d (one CUDA device classifies the reads)\n", numThreads);

    std::cout << "Reading " << fileTaxID << std::endl;                                              // :378
    {   // "Taxonomic rank: <header column>" (:51)
        std::ifstream f(fileTaxID.c_str());
        if (!f.is_open()) { std::cerr << "Error opening file " << fileTaxID << ".\n"; exit(1); }  // :39-42
        std::string field;
        for (int s = 0; s <= taxRank; s++) std::getline(f, field, s < 6 ? ';' : '\n');
        std::cout << "Taxonomic rank: " << field << std::endl;
    }
    std::vector<uint32_t> v(numTarg ? numTarg : 1), m;
    if (HIGHER) m.assign((size_t)LIME_RANKS * (numTarg ? numTarg : 1), 0);
    const int64_t found = lime_host_fix_rank(fileTaxID.c_str(), taxRank, numTarg, v.data(), numTarg,
                                             HIGHER ? m.data() : nullptr);
    if (found < 0) { std::cerr << "Error opening file " << fileTaxID << ".\n"; exit(1); }
    if ((uint64_t)found != numTarg) {                                                               // :395-398
        std::cerr << "Number of taxIDs = " << found << " lower than genome number: poor taxonomy information to classify." << std::endl;
        exit(1);
    }

    lime_ctx *ctx = nullptr;
    int rc = lime_init(&ctx, device_from_env());
    if (rc) die_lime(nullptr, "lime_init", rc);
    std::cout << "Reading files:";                                                                  // :401-445
    for (unsigned i = 0; i < numFile; i++) {
#if BIN
        const std::string sFileBin = fileInput[i] + ".bin", sFilePos = fileInput[i] + ".pos";
        std::cout << "\n\t" << sFileBin;
        std::cout << "\n\t" << sFilePos;
        uint64_t bb = 0, bp = 0;
        lime_pair_sim *bin = static_cast<lime_pair_sim *>(read_file_pinned(sFileBin, &bb));
        uint64_t *pos = static_cast<uint64_t *>(read_file_pinned(sFilePos, &bp));
        if (bp / sizeof(uint64_t) < numReads) { std::cerr << "\nError: " << sFilePos << " holds fewer than numReads entries." << std::endl; exit(EXIT_FAILURE); }
        rc = lime_slot_set(ctx, (int)i, bin, bb / sizeof(lime_pair_sim), pos, numReads);
        if (rc) die_lime(ctx, "lime_slot_set", rc);
        lime_pinned_free(bin);
        lime_pinned_free(pos);
#else
        // one text line per read; the same tokens the reference parses (Classify.cpp:128-153) become
        // the record images the classification kernel reads
        const std::string sFileTxt = fileInput[i] + ".txt";
        std::cout << "\n\t" << sFileTxt;
        std::vector<uint64_t> pos(numReads ? numReads : 1);
        const int64_t nRec = lime_host_read_res_txt(sFileTxt.c_str(), numReads, nullptr, 0, pos.data());
        if (nRec == -1) { std::cerr << "Error opening " << sFileTxt << std::endl; exit(1); }     // :439-442
        if (nRec < 0) { std::cerr << "\nError: " << sFileTxt << " holds fewer than numReads lines." << std::endl; exit(EXIT_FAILURE); }
        std::vector<lime_pair_sim> bin((size_t)nRec);
        lime_host_read_res_txt(sFileTxt.c_str(), numReads, bin.data(), (uint64_t)nRec, pos.data());
        rc = lime_slot_set(ctx, (int)i, bin.data(), (uint64_t)nRec, pos.data(), numReads);
        if (rc) die_lime(ctx, "lime_slot_set", rc);
#endif
    }
    std::cout << std::endl;

    std::cerr << "Start comparing..." << std::endl;                                                 // :465
    std::vector<uint8_t> type(numReads ? numReads : 1);
    std::vector<uint32_t> tax(numReads ? numReads : 1);
    std::vector<float> sim(numReads ? numReads : 1);
    rc = lime_classify(ctx, (int)numFile, numTarg, v.data(), HIGHER ? m.data() : nullptr, taxRank, HIGHER,
                       type.data(), tax.data(), sim.data());
    if (rc) die_lime(ctx, "lime_classify", rc);
    if (lime_host_write_csv(fileOutput.c_str(), numReads, type.data(), tax.data(), sim.data()) != 0)
        std::cerr << "ERROR: File Output not Open" << std::endl;                                   // :454

    uint64_t numC = 0, numA = 0, numNC = 0, numH = 0;
    for (uint32_t r = 0; r < numReads; r++) {
        if (type[r] == 'C') numC++; else if (type[r] == 'A') numA++; else if (type[r] == 'H') numH++; else numNC++;
    }
    std::cout << "Classification process at level " << taxRank << " completed.\nNumber of successfully classified reads: "
              << numC << "/" << numReads << ";" << std::endl;                                       // :738-743
#if HIGHER
    std::cout << "\tClassified at higher taxonomic ranks: " << numH << "." << std::endl;
#endif
    std::cout << "\tAmbiguously classified reads: " << numA << "." << std::endl;
    std::cout << "\tNot classified reads: " << numNC << "." << std::endl;
    fprintf(stdout, "Time: %.6lf\n", total.seconds());
    lime_destroy(ctx);
    return 0;
}
