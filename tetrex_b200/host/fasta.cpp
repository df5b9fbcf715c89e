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
    // kseq_read (include/kseq.h:181-226) over the whole buffer, byte for byte:
    //  * a record starts at the next '>' or '@' — ANYWHERE in the stream when no header character is pending (file start,
    //    after a FASTQ record), else at the header character that ended the previous record's sequence;
    //  * name = up to the first isspace(); comment = rest of that line unless the name ended at '\n';
    //  * sequence: a line starting with '>', '+' or '@' ends it (whatever the record's own marker); empty lines are
    //    skipped; every other line is appended RAW (spaces, tabs and all) minus one trailing '\r' when the sequence is
    //    longer than one byte (ks_getuntil2's rule looks at the accumulated string, :145);
    //  * '+' starts a FASTQ quality block: the rest of that line is dropped, quality lines are read until they are as
    //    long as the sequence; a length mismatch (-2) ends the reading of the file, as the callers' `>= 0` loops do.
    std::vector<fasta_record> out;
    size_t i = 0, n = data.size();
    auto append_line = [&](std::string & str, size_t & pos) { // ks_getuntil2(KS_SEP_LINE, append)
        size_t e = data.find('\n', pos);
        if (e == std::string::npos)
            e = n;
        str.append(data, pos, e - pos);
        pos = e < n ? e + 1 : n;
        if (str.size() > 1 && str.back() == '\r')
            str.pop_back();
    };
    int last_char = 0;
    for (;;)
    {
        if (last_char == 0)
        {
            while (i < n && data[i] != '>' && data[i] != '@')
                ++i;
            if (i >= n)
                break;
            last_char = data[i++];
        }
        fasta_record rec;
        if (i >= n)
            break; // ks_getuntil on an exhausted stream: -1, normal exit
        size_t ws = i;
        while (ws < n && !isspace(static_cast<unsigned char>(data[ws])))
            ++ws;
        rec.name.assign(data, i, ws - i);
        int const delim = ws < n ? data[ws] : 0;
        i = ws < n ? ws + 1 : n;
        if (delim != '\n' && ws < n)
            append_line(rec.comment, i);
        int c = -1;
        while (i < n)
        {
            c = static_cast<unsigned char>(data[i++]);
            if (c == '>' || c == '+' || c == '@')
                break;
            if (c == '\n')
            {
                c = -1;
                continue;
            }
            rec.seq.push_back(static_cast<char>(c));
            append_line(rec.seq, i);
            c = -1;
        }
        last_char = (c == '>' || c == '@') ? c : last_char;
        if (c != '+')
        {
            if (c != '>' && c != '@')
                last_char = -1; // end of stream: this was the last record
            out.push_back(std::move(rec));
            if (last_char == -1)
                break;
            continue;
        }
        // FASTQ
        while (i < n && data[i] != '\n')
            ++i;
        if (i >= n)
            break; // -2: no quality string
        ++i;
        std::string qual;
        while (i < n && qual.size() < rec.seq.size())
            append_line(qual, i);
        last_char = 0;
        if (qual.size() != rec.seq.size())
            break; // -2: the callers stop reading this file here, the record is not delivered
        out.push_back(std::move(rec));
    }
    return out;
}

} // namespace trxh
