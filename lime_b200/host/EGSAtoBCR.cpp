// EGSAtoBCR -- drop-in for the reference converter (EGSAtoBCR.cpp:17-103): same argv, reads
// <fastaFile>.<numSeq>.gesa (egsa: 13-byte records {u32 text, u32 suff, u32 lcp, u8 bwt}) and writes
// <fastaFile>.ebwt / .lcp / .da, the three arrays of the classification path.  Host only.
#include <vector>
#include "host_util.h"
using namespace limehost;

int main(int argc, char **argv)
{
    if (argc != 3) {                                                         // :19-23
        std::cerr << "Error usage: " << argv[0] << " fastaFile numSeq" << std::endl;
        exit(1);
    }
    const std::string fastaFile = argv[1], numSeq = argv[2];
    const std::string fnEGSA = fastaFile + "." + numSeq + ".gesa";
    const int64_t n = lime_host_read_gesa(fnEGSA.c_str(), nullptr, nullptr, nullptr, 0);
    if (n < 0) {                                                             // :33-37
        std::cerr << "Error opening " << fnEGSA << std::endl;
        printf("fopen failed, errno = %d\n", errno);
        exit(EXIT_FAILURE);
    }
    std::cerr << "file EGSA: " << fnEGSA << "." << std::endl;
    const std::string fnBWT = fastaFile + ".ebwt", fnLCP = fastaFile + ".lcp", fnDA = fastaFile + ".da";
    FILE *fb = fopen(fnBWT.c_str(), "wb");
    if (!fb) { std::cerr << "Error opening " << fnBWT << "." << std::endl; exit(EXIT_FAILURE); }
    FILE *fl = fopen(fnLCP.c_str(), "wb");
    if (!fl) { std::cerr << "Error opening " << fnLCP << "." << std::endl; exit(EXIT_FAILURE); }
    FILE *fd = fopen(fnDA.c_str(), "wb");
    if (!fd) { std::cerr << "Error opening " << fnDA << "." << std::endl; exit(EXIT_FAILURE); }
    std::cerr << "file ebwt : " << fnBWT << "." << std::endl;
    std::cerr << "file lcp: " << fnLCP << "." << std::endl;
    std::cerr << "file da: " << fnDA << "." << std::endl;
    std::vector<uint8_t> ebwt((size_t)n ? (size_t)n : 1);
    std::vector<uint32_t> lcp((size_t)n ? (size_t)n : 1), da((size_t)n ? (size_t)n : 1);
    lime_host_read_gesa(fnEGSA.c_str(), ebwt.data(), lcp.data(), da.data(), (uint64_t)n);
    if (n) {
        fwrite(ebwt.data(), 1, (size_t)n, fb);
        fwrite(lcp.data(), sizeof(uint32_t), (size_t)n, fl);
        fwrite(da.data(), sizeof(uint32_t), (size_t)n, fd);
    }
    fclose(fb); fclose(fl); fclose(fd);
    std::cerr << "The total number of elements is " << n << "\n";              // :99
    return 0;
}
