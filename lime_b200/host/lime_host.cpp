// lime_host.cpp -- lineage parsing and CSV output (host side of step 3), see include/lime_host.h.
#include "../../include/lime_host.h"
#include <array>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>
#include <cstring>

namespace {
using Row = std::array<std::string, 7>;

// one lineage row = six ';'-terminated fields and one '\n'-terminated field, read with the same
// stream primitives as Classify.cpp:48-50 so eof/fail behave identically on odd files
bool read_row(std::istream &in, Row &row)
{
    for (int s = 0; s < LIME_RANKS; s++) std::getline(in, row[s], ';');
    std::getline(in, row[6], '\n');
    return in.good();
}
} // namespace

extern "C" int64_t lime_host_fix_rank(const char *path, int tax_rank, uint32_t n_refs,
                                      uint32_t *v, uint64_t v_cap, uint32_t *m)
{
    std::ifstream in(path, std::ifstream::in);
    if (!in.is_open()) return -1;
    if (tax_rank < 0 || tax_rank > LIME_RANKS) return -2;
    if (m) std::memset(m, 0, sizeof(uint32_t) * LIME_RANKS * (size_t)n_refs);
    Row row;
    read_row(in, row);                       // header
    bool ok = read_row(in, row);             // first genome
    uint64_t found = 0, line = 0;
    while (ok) {
        if (tax_rank == 0) {
            if (found < v_cap) v[found] = (uint32_t)line;
            found++;
        } else if (!row[tax_rank].empty()) {
            if (found < v_cap) v[found] = (uint32_t)std::atoi(row[tax_rank].c_str());
            found++;
        }
        if (m && line < n_refs)
            for (int i = (tax_rank > 0 ? tax_rank - 1 : 0); i < LIME_RANKS; i++)
                if (!row[i + 1].empty()) m[(size_t)i * n_refs + line] = (uint32_t)std::atoi(row[i + 1].c_str());
        line++;
        ok = read_row(in, row);
    }
    return (int64_t)found;
}

extern "C" int lime_host_write_csv(const char *path, uint32_t n_reads,
                                   const uint8_t *type, const uint32_t *tax, const float *sim)
{
    std::ofstream out(path, std::ofstream::out);
    if (!out.is_open()) return -1;
    out << "C/U/A/H,IdSeqRead,TaxID,maxSim\n";
    std::string buf;
    buf.reserve(1 << 20);
    char tmp[64];
    for (uint32_t r = 0; r < n_reads; r++) {
        int n;
        if (type[r] == 'U')      n = std::snprintf(tmp, sizeof tmp, "U,%u,NA,0\n", r);
        else if (type[r] == 'A') n = std::snprintf(tmp, sizeof tmp, "A,%u,NA,0\n", r);
        else n = std::snprintf(tmp, sizeof tmp, "%c,%u,%u,%g\n", type[r] == 'C' ? 'C' : 'H', r, tax[r], (double)sim[r]);
        buf.append(tmp, (size_t)n);
        if (buf.size() > (1 << 20) - 128) { out.write(buf.data(), (std::streamsize)buf.size()); buf.clear(); }
    }
    out.write(buf.data(), (std::streamsize)buf.size());
    return out.good() ? 0 : -1;
}
