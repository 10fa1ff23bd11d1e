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

// ---- .gesa (egsa output) -> ebwt / lcp / da, EGSAtoBCR.cpp:9-15,72-91 --------------------------------
// record = { u32 text (document id), u32 suff, u32 lcp, u8 bwt }, 13 bytes, no padding, no header
extern "C" int64_t lime_host_read_gesa(const char *path, uint8_t *ebwt, uint32_t *lcp, uint32_t *da, uint64_t cap)
{
    FILE *f = std::fopen(path, "rb");
    if (!f) return -1;
    constexpr size_t REC = 13, BATCH = 1 << 16;
    std::string buf(REC * BATCH, '\0');
    uint64_t n = 0;
    size_t got;
    while ((got = std::fread(&buf[0], 1, REC * BATCH, f)) >= REC) {
        const size_t recs = got / REC;          // a truncated last record is dropped, as by the reference loop
        for (size_t i = 0; i < recs; i++, n++) {
            if (n >= cap) continue;             // keep counting: the caller sizes its arrays from the result
            const unsigned char *p = reinterpret_cast<const unsigned char *>(&buf[i * REC]);
            std::memcpy(&da[n], p, 4);
            std::memcpy(&lcp[n], p + 8, 4);
            ebwt[n] = p[12];
        }
        if (got % REC) break;
    }
    std::fclose(f);
    return (int64_t)n;
}

// ---- BIN=0: <fasta>.res.txt, one line per read (ClusterBWT_DA.cpp:414,424,440) -------------------------
extern "C" int lime_host_write_res_txt(const char *path, const lime_pair_sim *bin, uint64_t n_records,
                                       const uint64_t *pos, uint32_t n_reads)
{
    FILE *f = std::fopen(path, "w");
    if (!f) return -1;
    for (uint32_t r = 0; r < n_reads; r++) {
        const uint64_t h = pos[r];
        if (h) {
            if (h >= n_records || h + bin[h].idRef >= n_records) { std::fclose(f); return -2; }
            std::fprintf(f, "%.5f", bin[h].sim);
            for (uint32_t j = 1; j <= bin[h].idRef; j++) std::fprintf(f, "\t%u\t%.5f", bin[h + j].idRef, bin[h + j].sim);
        }
        std::fputc('\n', f);
    }
    return std::fclose(f) == 0 ? 0 : -1;
}

// The reverse (Classify.cpp:128-153 reads the same tokens with operator>>): images of .res.bin / .res.pos
// with the values the text holds.  Returns the number of records incl. the sentinel (it may exceed
// bin_cap: nothing is written past it), -1 if the file cannot be opened, -2 on fewer than n_reads lines.
extern "C" int64_t lime_host_read_res_txt(const char *path, uint32_t n_reads, lime_pair_sim *bin, uint64_t bin_cap,
                                          uint64_t *pos)
{
    std::ifstream in(path);
    if (!in.is_open()) return -1;
    uint64_t total = 1;
    if (bin_cap) { bin[0].sim = 0.0f; bin[0].idRef = 0; }
    std::string line;
    for (uint32_t r = 0; r < n_reads; r++) {
        if (!std::getline(in, line)) return -2;
        std::istringstream is(line);
        float mx;
        if (!(is >> mx)) { pos[r] = 0; continue; }
        const uint64_t h = total++;
        uint32_t count = 0, id;
        float s;
        while ((is >> id) && (is >> s)) {
            if (total < bin_cap) { bin[total].sim = s; bin[total].idRef = id; }
            total++; count++;
        }
        if (h < bin_cap) { bin[h].sim = mx; bin[h].idRef = count; }
        pos[r] = h;
    }
    return (int64_t)total;
}
