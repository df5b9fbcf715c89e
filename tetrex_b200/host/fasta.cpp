// fasta.cpp — FASTA/FASTQ reader with kseq_read's record semantics (include/kseq.h): name = header up to the
// first whitespace, comment = rest of the header line, seq = sequence lines concatenated without line ends.
// Input goes through zlib's gz layer like the reference's gzopen/gzread, so plain and gzip files both work.
#include <zlib.h>

#include "trx_host.hpp"

namespace trxh
{

std::vector<fasta_record> read_fasta(std::string const & path)
{
    gzFile f = gzopen(path.c_str(), "r");
    if (!f)
        throw error("Failed to open FASTA file: " + path); // include/query.h:119-122
    gzbuffer(f, 1 << 20);
    std::string data;
    {
        std::vector<char> buf(1 << 20);
        int n;
        while ((n = gzread(f, buf.data(), static_cast<unsigned>(buf.size()))) > 0)
            data.append(buf.data(), static_cast<size_t>(n));
        gzclose(f);
        if (n < 0)
            throw error("read error in " + path);
    }
    std::vector<fasta_record> out;
    size_t i = 0, n = data.size();
    auto line_end = [&](size_t p) {
        size_t e = data.find('\n', p);
        return e == std::string::npos ? n : e;
    };
    // skip to the first header
    while (i < n && data[i] != '>' && data[i] != '@')
        i = line_end(i) + 1;
    while (i < n)
    {
        char const marker = data[i];
        size_t e = line_end(i);
        size_t he = e;
        if (he > i && data[he - 1] == '\r')
            --he;
        fasta_record rec;
        size_t p = i + 1;
        size_t ws = p;
        while (ws < he && !isspace(static_cast<unsigned char>(data[ws])))
            ++ws;
        rec.name.assign(data, p, ws - p);
        if (ws < he)
            rec.comment.assign(data, ws + 1, he - ws - 1);
        i = e + 1;
        // sequence lines until the next header ('>' / '@' at line start) or, for FASTQ, the '+' line
        while (i < n && data[i] != '>' && data[i] != '+' && !(marker == '@' && data[i] == '@'))
        {
            e = line_end(i);
            for (size_t q = i; q < e; ++q)
            {
                unsigned char c = static_cast<unsigned char>(data[q]);
                if (isgraph(c)) // kseq keeps graphical characters only
                    rec.seq.push_back(static_cast<char>(c));
            }
            i = e + 1;
        }
        if (i < n && data[i] == '+')
        {
            // FASTQ: skip the '+' line and as many quality characters as there are bases
            i = line_end(i) + 1;
            size_t q = 0;
            while (i < n && q < rec.seq.size())
            {
                e = line_end(i);
                size_t le = e;
                if (le > i && data[le - 1] == '\r')
                    --le;
                q += le - i;
                i = e + 1;
            }
        }
        out.push_back(std::move(rec));
    }
    return out;
}

} // namespace trxh
