// main.cpp — `tetrex` command line (src/main.cpp, include/arg_parse.h): index / query / inspect with the
// reference's option letters.  Two extra developer sub-commands print intermediate results for the parity tests:
//   tetrex kgraph [-n] [-r murphy|li] -k K REGEX      postfix string + every node of the k-graph
//   tetrex bins [-a] [-g DINDEX] IDX REGEX            the candidate bins of process_query, one index per line
//   tetrex dindex-rewrite IN OUT                      DGram index file → image → file (host only)
//   tetrex hibf-canon IDX                             numbering-independent dump of a HIBF index file (host only)
// Unlike the reference (molecule_/ibf_ are frozen before parsing, so its -n and -i are dead) the
// flags take effect here.
#include <cstring>
#include <iostream>

#include "trx_host.hpp"

using namespace trxh;

namespace
{

struct arg_reader
{
    int argc;
    char ** argv;
    int pos;
    bool done() const { return pos >= argc; }
    std::string next(char const * what)
    {
        if (pos >= argc)
            throw error(std::string("missing value for ") + what);
        return argv[pos++];
    }
};

bool is_flag(std::string const & a, char s, char const * l)
{
    return (a.size() == 2 && a[0] == '-' && a[1] == s) || a == std::string("--") + l;
}

unsigned to_uint(std::string const & s, char const * what)
{
    try
    {
        size_t used = 0;
        unsigned long v = std::stoul(s, &used);
        if (used != s.size())
            throw 0;
        return static_cast<unsigned>(v);
    }
    catch (...)
    {
        throw error(std::string("Value parse failed for ") + what + ": " + s);
    }
}

int run_index(arg_reader a)
{
    index_arguments args;
    std::vector<std::string> positional;
    try
    {
        while (!a.done())
        {
            std::string s = a.next("argument");
            if (is_flag(s, 'k', "ksize")) args.k = static_cast<uint8_t>(to_uint(a.next("-k"), "-k"));
            else if (is_flag(s, 'p', "fpr")) args.fpr = std::stof(a.next("-p"));
            else if (is_flag(s, 'c', "hash_count")) args.hash_count = static_cast<uint8_t>(to_uint(a.next("-c"), "-c"));
            else if (is_flag(s, 't', "threads")) args.t = static_cast<uint8_t>(to_uint(a.next("-t"), "-t"));
            else if (is_flag(s, 'n', "nucleic_acid")) args.dna = true;
            else if (is_flag(s, 'i', "ibf")) args.idx = true;
            else if (is_flag(s, 'r', "reduce"))
            {
                args.reduction = a.next("-r");
                if (args.reduction != "murphy" && args.reduction != "li")
                    throw error("Validation failed for option -r/--reduce: Value " + args.reduction + " is not one of [murphy, li].");
            }
            else if (s.size() > 1 && s[0] == '-') throw error("Unknown option " + s + ".");
            else positional.push_back(s);
        }
        if (positional.size() < 2)
            throw error("Not enough positional arguments provided (Need at least 2).");
    }
    catch (std::exception const & e)
    {
        std::cerr << "[Indexing Parser Error] " << e.what() << "\n"; // src/main.cpp:25
        return 0;
    }
    if (!args.dna && args.k > 12)
    {
        std::cerr << "[Indexing Parser Error] Max kmer size for amino acids is 12" << "\n"; // src/main.cpp:30
        return 0;
    }
    args.ofile = positional[0];
    args.acid_libs.assign(positional.begin() + 1, positional.end());
    drive_index(args);
    return 0;
}

int run_query(arg_reader a)
{
    query_arguments args;
    std::vector<std::string> positional;
    bool read_file = false;
    try
    {
        while (!a.done())
        {
            std::string s = a.next("argument");
            if (is_flag(s, 'd', "draw")) args.draw = true;
            else if (is_flag(s, 'v', "verbose")) args.verbose = true;
            else if (is_flag(s, 'f', "file")) read_file = true;
            else if (is_flag(s, 'c', "conj")) args.conjunction = "on";
            else if (is_flag(s, 'a', "augment")) args.augment = true;
            else if (is_flag(s, 't', "threads")) args.t = static_cast<uint8_t>(to_uint(a.next("-t"), "-t"));
            else if (is_flag(s, 'o', "output")) args.ofile = a.next("-o");
            else if (is_flag(s, 'g', "gibf")) args.dibf = a.next("-g");
            else if (s == "--gpus") args.gpus = to_uint(a.next("--gpus"), "--gpus");
            else if (s.size() > 1 && s[0] == '-' && positional.size() < 1) throw error("Unknown option " + s + ".");
            else positional.push_back(s);
        }
        if (positional.empty())
            throw error("Not enough positional arguments provided (Need at least 1).");
    }
    catch (std::exception const & e)
    {
        std::cerr << "[Error TetRex Query module " << e.what() << "\n"; // src/main.cpp:47
        return 0;
    }
    args.idx = positional[0];
    args.regex = positional.size() > 1 ? positional[1] : "-";
    if (read_file)
    {
        args.read_file = args.regex;
        args.regex = "x"; // not read from stdin
    }
    try
    {
        drive_query(args);
    }
    catch (archive_error const & e)
    {
        std::cerr << "Filepath to (H)IBF Index not valid" << std::endl; // src/main.cpp:55-56
        std::cerr << e.what() << '\n';
    }
    return 0;
}

// run_dindex (src/main.cpp:95-109, initialize_dgram_parser include/arg_parse.h:111-123)
int run_dindex(arg_reader a)
{
    dindex_arguments args;
    std::vector<std::string> positional;
    try
    {
        while (!a.done())
        {
            std::string s = a.next("argument");
            if (is_flag(s, 'l', "lower")) args.min = to_uint(a.next("-l"), "-l");
            else if (is_flag(s, 'u', "upper")) args.max = to_uint(a.next("-u"), "-u");
            else if (is_flag(s, 'n', "nucleic_acid") || is_flag(s, 'i', "ibf")) {} // parsed and never read by drive_dindex
            else if (s.size() > 1 && s[0] == '-') throw error("Unknown option " + s + ".");
            else positional.push_back(s);
        }
        if (positional.size() < 2)
            throw error("Not enough positional arguments provided (Need at least 2).");
    }
    catch (std::exception const & e)
    {
        std::cerr << "[Error TetRex Dgram Indexing module " << e.what() << "\n"; // src/main.cpp:105
        return 0;
    }
    args.ofile = positional[0];
    args.acid_libs.assign(positional.begin() + 1, positional.end());
    drive_dindex(args);
    return 0;
}

int run_kgraph(arg_reader a)
{
    unsigned k = 3;
    bool dna = false;
    int reduction = TRX_RED_BASE;
    bool augment = false;
    std::string regex;
    while (!a.done())
    {
        std::string s = a.next("argument");
        if (s == "-k") k = to_uint(a.next("-k"), "-k");
        else if (s == "-n") dna = true;
        else if (s == "-a") augment = true;
        else if (s == "-r")
        {
            std::string r = a.next("-r");
            reduction = r == "murphy" ? TRX_RED_MURPHY : r == "li" ? TRX_RED_LI : TRX_RED_BASE;
        }
        else regex = s;
    }
    std::string rx = regex;
    if (!dna)
    {
        if (reduction != TRX_RED_BASE)
        {
            uint8_t aamap[256], redmap[256];
            trx_peptide_maps(reduction, aamap, redmap);
            rx = reduce_query_alphabet(rx, redmap);
        }
        rx = trim_regex(rx);
    }
    std::string postfix = translate(rx);
    std::cout << "postfix\t" << postfix << "\n";
    kgraph g = construct_kgraph(postfix, static_cast<uint8_t>(k), reduction != TRX_RED_BASE);
    if (augment)
        g.augment(); // -a: the graph the collector walks after OTFCollector::augment
    std::cout << "nodes\t" << g.size() << "\n";
    for (size_t i = 0; i < g.size(); ++i) // a Gap node's length rides in the second slot
        std::cout << i << "\t" << g.symbol[i] << "\t" << g.first[i] << "\t"
                  << (g.symbol[i] == Gap ? static_cast<long>(g.gap[i]) : static_cast<long>(g.second[i])) << "\t" << g.rank[i] << "\n";
    return 0;
}

int run_bins(arg_reader a)
{
    bool augment = false;
    std::string dibf_path, idx, regex;
    std::vector<std::string> positional;
    while (!a.done())
    {
        std::string s = a.next("argument");
        if (s == "-a" && positional.empty()) augment = true;
        else if (s == "-g" && positional.empty()) dibf_path = a.next("-g");
        else positional.push_back(s);
    }
    if (positional.size() != 2)
        throw error("usage: tetrex bins [-a] [-g DINDEX] IDX REGEX");
    idx = positional[0];
    regex = positional[1];
    trx_ctx_t * ctx = nullptr;
    check(trx_ctx_create(0, nullptr, &ctx));
    int rc = 0;
    try
    {
        DeviceIndex index(ctx);
        index.load(idx);
        DGramIndex dibf(ctx);
        if (!dibf_path.empty())
        {
            dibf.load(dibf_path);
            index.attach_dgram(&dibf);
        }
        collect_stats st;
        bitvector v = process_query(regex, index, &st, nullptr, augment);
        for (uint64_t b = 0; b < index.bin_count(); ++b)
            if (v[b / 64] >> (b % 64) & 1)
                std::cout << b << "\n";
        std::cerr << "probes " << st.probes << " ranks " << st.ranks << " items " << st.items << std::endl;
    }
    catch (std::exception const & e)
    {
        std::cerr << e.what() << std::endl;
        rc = 1;
    }
    trx_ctx_destroy(ctx);
    return rc;
}

// reads a DGram index file and writes it back (no device involved): the reader/writer round trip of the parity tests
int run_dindex_rewrite(arg_reader a)
{
    std::string in = a.next("input"), out = a.next("output");
    dgram_image img = read_dindex(in);
    write_dindex(img, out);
    std::cout << "min_gap\t" << img.min_gap << "\nmax_gap\t" << img.max_gap << "\nhash_count\t" << static_cast<int>(img.hash_count)
              << "\nfpr\t" << img.fpr << "\nbins\t" << img.bin_count << "\nbin_size\t" << img.ibf.bin_size << "\n";
    return 0;
}

// Canonical, numbering-independent description of a HIBF index file: a depth-first walk from the root that prints, for
// every IBF, its geometry and a hash of its bit matrix, and for every technical bin the user bin it belongs to or the
// IBF below it.  Two files with the same output answer every membership query identically (the reference numbers its
// IBFs in a shuffled order, lib/hibf/src/hierarchical_interleaved_bloom_filter.cpp:118-121, so files cannot be compared
// byte for byte).
void canon_ibf(index_image const & img, uint64_t id, std::string const & indent)
{
    ibf_image const & f = img.hibf.ibf_vector[id];
    uint64_t h = 1469598103934665603ull;
    for (uint64_t w : f.words)
        h = (h ^ w) * 1099511628211ull;
    std::cout << indent << "ibf bins " << f.bins << " technical " << f.technical_bins << " bin_size " << f.bin_size << " hash_funs "
              << f.hash_funs << " words " << f.words.size() << " fnv " << h << "\n";
    auto const & next = img.hibf.next_ibf_id[id];
    auto const & ub = img.hibf.ibf_bin_to_user_bin_id[id];
    for (size_t tb = 0; tb < next.size(); ++tb)
    {
        if (next[tb] != id)
        {
            std::cout << indent << " tb " << tb << " merged\n";
            canon_ibf(img, next[tb], indent + "  ");
        }
        else
            std::cout << indent << " tb " << tb << " user_bin " << static_cast<int64_t>(ub[tb]) << "\n";
    }
}

int run_hibf_canon(arg_reader a)
{
    index_image img = read_index(a.next("index"));
    if (!img.is_hibf)
        throw error("not a HIBF index");
    std::cout << "user_bins " << img.hibf.number_of_user_bins << " ibfs " << img.hibf.ibf_vector.size() << "\n";
    canon_ibf(img, 0, "");
    return 0;
}

void usage()
{
    std::cerr << "tetrex - regular expression search over (H)IBF-indexed sequence libraries (B200 build)\n"
                 "  tetrex index [-k K] [-p FPR] [-c HASHES] [-n] [-i] [-r murphy|li] OUT LIB...\n"
                 "  tetrex query [-v] [-f] [-c] [-a] [-g DINDEX] [-o FILE] [--gpus N] IDX REGEX\n"
                 "  tetrex dindex [-l LOWER] [-u UPPER] OUT LIB...\n"
                 "  tetrex inspect IDX\n";
}

} // namespace

int main(int argc, char ** argv)
{
    if (argc < 2)
    {
        usage();
        return 0;
    }
    std::string const cmd = argv[1];
    arg_reader a{argc, argv, 2};
    try
    {
        if (cmd == "index") return run_index(a);
        if (cmd == "query") return run_query(a);
        if (cmd == "inspect")
        {
            if (a.done())
            {
                std::cerr << "[Error TetRex Index Inspection module Not enough positional arguments provided (Need at least 1).\n";
                return 0;
            }
            drive_inspection(a.next("index"));
            return 0;
        }
        if (cmd == "kgraph") return run_kgraph(a);
        if (cmd == "bins") return run_bins(a);
        if (cmd == "dindex-rewrite") return run_dindex_rewrite(a);
        if (cmd == "hibf-canon") return run_hibf_canon(a);
        if (cmd == "dindex") return run_dindex(a);
        if (cmd == "model")
        {
            std::cerr << "[Error] sub-command '" << cmd << "' is outside this build (DESIGN.md, out of scope)\n";
            return 1;
        }
        usage();
        return cmd == "-h" || cmd == "--help" ? 0 : 1;
    }
    catch (std::exception const & e)
    {
        std::cerr << "[Error] " << e.what() << "\n"; // src/main.cpp:123
        return 1;
    }
}
