// regex_compile.cu — host-side compiler: verify regex → Glushkov position automaton (+ priority-ordered
// follow lists for leftmost-first).  Pure host code; lives in a .cu file only so that one toolchain
// builds the whole library.
//
// Semantics to reproduce: RE2 as iter_disk_search uses it (include/query.h:101-104,146-149;
// src/query.cpp:194-237; lib/re2/re2/re2.cc:428-437): the pattern is wrapped in one capture group,
// searched unanchored, POSIX ⇒ leftmost-longest, default ⇒ leftmost-first with greedy quantifiers.
#include <atomic>
#include <memory>
#include <stdexcept>

#include "common.cuh"
#include "regex.cuh"

namespace
{

struct charset
{
    uint64_t w[4]{0, 0, 0, 0};
    void set(unsigned c) { w[c >> 6] |= 1ULL << (c & 63); }
    bool has(unsigned c) const { return (w[c >> 6] >> (c & 63)) & 1; }
    void negate_ascii() // RE2 works on UTF-8: a negated class / '.' never matches a lone byte >= 0x80
    {
        w[0] = ~w[0];
        w[1] = ~w[1];
        w[2] = w[3] = 0;
    }
};

enum class nk { chr, cat, alt, star, plus, quest, empty };

struct node
{
    nk kind{nk::empty};
    charset cs{};
    std::unique_ptr<node> a, b;
};
using node_p = std::unique_ptr<node>;

node_p mk(nk k, node_p a = nullptr, node_p b = nullptr)
{
    auto n = std::make_unique<node>();
    n->kind = k;
    n->a = std::move(a);
    n->b = std::move(b);
    return n;
}

node_p clone(node const * n)
{
    if (!n)
        return nullptr;
    auto c = std::make_unique<node>();
    c->kind = n->kind;
    c->cs = n->cs;
    c->a = clone(n->a.get());
    c->b = clone(n->b.get());
    return c;
}

struct parse_error : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

size_t count_positions(node const * n)
{
    if (!n)
        return 0;
    if (n->kind == nk::chr)
        return 1;
    return count_positions(n->a.get()) + count_positions(n->b.get());
}

class parser
{
    std::string const & s;
    size_t i{0};

public:
    bool anchor_begin{false}, anchor_end{false};
    explicit parser(std::string const & p) : s(p) {}

    node_p parse()
    {
        if (i < s.size() && s[i] == '^')
        {
            anchor_begin = true;
            ++i;
        }
        node_p r = alternation(0);
        if (i != s.size())
            throw parse_error("unbalanced ')'");
        return r;
    }

private:
    bool eof() const { return i >= s.size(); }

    node_p alternation(int depth)
    {
        node_p left = concatenation(depth);
        while (!eof() && s[i] == '|')
        {
            ++i;
            node_p right = concatenation(depth);
            left = mk(nk::alt, std::move(left), std::move(right));
        }
        return left;
    }

    node_p concatenation(int depth)
    {
        node_p left;
        while (!eof() && s[i] != '|' && s[i] != ')')
        {
            if (s[i] == '$' && depth == 0 && i + 1 == s.size())
            {
                anchor_end = true;
                ++i;
                break;
            }
            node_p r = repetition(depth);
            left = left ? mk(nk::cat, std::move(left), std::move(r)) : std::move(r);
        }
        if (!left)
            left = mk(nk::empty);
        return left;
    }

    node_p repeat(node_p atom, long lo, long hi) // hi < 0 = unbounded
    {
        size_t const per = count_positions(atom.get());
        size_t const copies = static_cast<size_t>(hi < 0 ? (lo > 0 ? lo : 1) : hi);
        if (per * copies > TRX_REGEX_MAX_POSITIONS)
            throw parse_error("pattern needs more than " + std::to_string(TRX_REGEX_MAX_POSITIONS) +
                              " symbol positions after {m,n} expansion");
        // x{lo,hi} = x^lo (x(x(x)?)?)?   — the nesting RE2's simplifier produces (lib/re2/re2/simplify.cc)
        node_p out;
        auto append = [&](node_p n) { out = out ? mk(nk::cat, std::move(out), std::move(n)) : std::move(n); };
        if (hi < 0)
        {
            if (lo == 0)
                return mk(nk::star, std::move(atom));
            for (long c = 0; c + 1 < lo; ++c)
                append(clone(atom.get()));
            append(mk(nk::plus, std::move(atom)));
            return out;
        }
        if (hi == 0)
            return mk(nk::empty);
        for (long c = 0; c < lo; ++c)
            append(clone(atom.get()));
        if (hi > lo)
        {
            node_p opt;
            for (long c = 0; c < hi - lo; ++c)
            {
                node_p inner = clone(atom.get());
                if (opt)
                    inner = mk(nk::cat, std::move(inner), std::move(opt));
                opt = mk(nk::quest, std::move(inner));
            }
            append(std::move(opt));
        }
        return out;
    }

    node_p repetition(int depth)
    {
        node_p atom_node = atom(depth);
        while (!eof())
        {
            char c = s[i];
            if (c == '*' || c == '+' || c == '?')
            {
                ++i;
                if (!eof() && (s[i] == '?' || s[i] == '+' || s[i] == '*'))
                    throw parse_error("non-greedy / stacked quantifiers are not supported");
                atom_node = mk(c == '*' ? nk::star : c == '+' ? nk::plus : nk::quest, std::move(atom_node));
            }
            else if (c == '{')
            {
                size_t j = i + 1;
                auto number = [&](long & v)
                {
                    if (j >= s.size() || !isdigit(static_cast<unsigned char>(s[j])))
                        return false;
                    v = 0;
                    while (j < s.size() && isdigit(static_cast<unsigned char>(s[j])))
                    {
                        v = v * 10 + (s[j] - '0');
                        if (v > 1000)
                            throw parse_error("repeat count too large");
                        ++j;
                    }
                    return true;
                };
                long lo = 0, hi = 0;
                if (!number(lo))
                    throw parse_error("malformed {m,n}");
                if (j < s.size() && s[j] == ',')
                {
                    ++j;
                    if (!number(hi))
                        hi = -1;
                }
                else
                    hi = lo;
                if (j >= s.size() || s[j] != '}')
                    throw parse_error("malformed {m,n}");
                if (hi >= 0 && hi < lo)
                    throw parse_error("bad repetition operator: max < min");
                i = j + 1;
                if (!eof() && s[i] == '?')
                    throw parse_error("non-greedy quantifiers are not supported");
                atom_node = repeat(std::move(atom_node), lo, hi);
            }
            else
                break;
        }
        return atom_node;
    }

    unsigned escape()
    {
        if (eof())
            throw parse_error("trailing backslash");
        char c = s[i++];
        switch (c)
        {
        case 'n': return '\n';
        case 't': return '\t';
        case 'r': return '\r';
        case 'f': return '\f';
        case 'v': return '\v';
        case 'a': return '\a';
        default:
            if (isalnum(static_cast<unsigned char>(c)))
                throw parse_error(std::string("unsupported escape \\") + c);
            return static_cast<unsigned char>(c);
        }
    }

    node_p atom(int depth)
    {
        if (eof())
            throw parse_error("missing operand");
        char c = s[i];
        if (c == '(')
        {
            ++i;
            if (i + 1 < s.size() && s[i] == '?' && s[i + 1] == ':')
                i += 2;
            else if (!eof() && s[i] == '?')
                throw parse_error("(?...) groups other than (?:...) are not supported");
            node_p r = alternation(depth + 1);
            if (eof() || s[i] != ')')
                throw parse_error("missing ')'");
            ++i;
            return r;
        }
        if (c == '*' || c == '+' || c == '?' || c == '{')
            throw parse_error("quantifier without operand");
        if (c == '^' || c == '$')
            throw parse_error("'^' / '$' are supported only at the very start / end of the pattern");
        node_p n = mk(nk::chr);
        if (c == '.')
        {
            ++i;
            n->cs.set('\n');
            n->cs.negate_ascii();
            return n;
        }
        if (c == '[')
        {
            ++i;
            bool neg = false;
            if (!eof() && s[i] == '^')
            {
                neg = true;
                ++i;
            }
            bool first = true;
            while (true)
            {
                if (eof())
                    throw parse_error("missing ']'");
                if (s[i] == ']' && !first)
                {
                    ++i;
                    break;
                }
                first = false;
                if (s[i] == '[' && i + 1 < s.size() && s[i + 1] == ':')
                    throw parse_error("[:class:] is not supported");
                unsigned lo = static_cast<unsigned char>(s[i++]);
                if (lo == '\\')
                    lo = escape();
                unsigned hi = lo;
                if (i + 1 < s.size() && s[i] == '-' && s[i + 1] != ']')
                {
                    ++i;
                    hi = static_cast<unsigned char>(s[i++]);
                    if (hi == '\\')
                        hi = escape();
                    if (hi < lo)
                        throw parse_error("bad character class range");
                }
                if (hi >= 128)
                    throw parse_error("non-ASCII bytes in a pattern are not supported");
                for (unsigned x = lo; x <= hi; ++x)
                    n->cs.set(x);
            }
            if (neg)
                n->cs.negate_ascii();
            return n;
        }
        ++i;
        unsigned lit = static_cast<unsigned char>(c);
        if (c == '\\')
            lit = escape();
        if (lit >= 128)
            throw parse_error("non-ASCII bytes in a pattern are not supported");
        n->cs.set(lit);
        return n;
    }
};

} // namespace

namespace
{

// ---- prioritized Thompson program (Pike VM instruction set) ------------------------------------------
struct inst
{
    enum op_t { CHR, SPLIT, JMP, MATCH } op;
    int x{0}, y{0}; // SPLIT: x preferred over y; JMP: x; CHR: x = position id
};

struct program
{
    std::vector<inst> code;
    std::vector<charset> pos_class; // position id → symbol class
    std::vector<int> pos_pc;        // position id → pc of its CHR instruction

    int emit(inst::op_t op, int x = 0, int y = 0)
    {
        code.push_back(inst{op, x, y});
        return static_cast<int>(code.size()) - 1;
    }

    void gen(node const * n)
    {
        switch (n->kind)
        {
        case nk::empty: break;
        case nk::chr:
        {
            int id = static_cast<int>(pos_class.size());
            pos_class.push_back(n->cs);
            pos_pc.push_back(emit(inst::CHR, id));
            break;
        }
        case nk::cat:
            gen(n->a.get());
            gen(n->b.get());
            break;
        case nk::alt:
        {
            int sp = emit(inst::SPLIT);
            code[sp].x = static_cast<int>(code.size());
            gen(n->a.get());
            int j = emit(inst::JMP);
            code[sp].y = static_cast<int>(code.size());
            gen(n->b.get());
            code[j].x = static_cast<int>(code.size());
            break;
        }
        case nk::quest: // greedy: try the operand first
        {
            int sp = emit(inst::SPLIT);
            code[sp].x = static_cast<int>(code.size());
            gen(n->a.get());
            code[sp].y = static_cast<int>(code.size());
            break;
        }
        case nk::star:
        {
            int sp = emit(inst::SPLIT);
            code[sp].x = static_cast<int>(code.size());
            gen(n->a.get());
            emit(inst::JMP, sp);
            code[sp].y = static_cast<int>(code.size());
            break;
        }
        case nk::plus:
        {
            int start = static_cast<int>(code.size());
            gen(n->a.get());
            int sp = emit(inst::SPLIT, start);
            code[sp].y = static_cast<int>(code.size());
            break;
        }
        }
    }

    // priority-ordered epsilon closure of pc: positions (and kAccept) in the order a backtracking
    // matcher would try them; first visit of a pc wins, which also terminates empty loops.
    void closure(int pc, std::vector<char> & seen, std::vector<uint8_t> & out) const
    {
        std::vector<int> stack{pc};
        while (!stack.empty())
        {
            int p = stack.back();
            stack.pop_back();
            if (seen[p])
                continue;
            seen[p] = 1;
            inst const & in = code[p];
            switch (in.op)
            {
            case inst::CHR: out.push_back(static_cast<uint8_t>(in.x)); break;
            case inst::MATCH: out.push_back(trx::kAccept); break;
            case inst::JMP: stack.push_back(in.x); break;
            case inst::SPLIT:
                stack.push_back(in.y); // popped second
                stack.push_back(in.x); // popped first = preferred
                break;
            }
        }
    }
};

} // namespace

extern "C"
{

int trx_regex_compile(const char * pattern, int posix_longest, trx_regex_t ** out)
{
    if (!pattern || !out)
        return trx::set_error(TRX_EINVAL, "NULL argument");
    *out = nullptr;
    std::string const pat(pattern);
    program prog;
    bool anchor_begin = false, anchor_end = false;
    try
    {
        parser p(pat);
        node_p root = p.parse();
        anchor_begin = p.anchor_begin;
        anchor_end = p.anchor_end;
        if (count_positions(root.get()) > TRX_REGEX_MAX_POSITIONS)
            return trx::set_error(TRX_EREGEX, "pattern '%s' needs %zu symbol positions (max %d)", pattern,
                                  count_positions(root.get()), TRX_REGEX_MAX_POSITIONS);
        prog.gen(root.get());
        prog.emit(inst::MATCH);
    }
    catch (parse_error const & e)
    {
        return trx::set_error(TRX_EREGEX, "pattern '%s': %s", pattern, e.what());
    }
    int const m = static_cast<int>(prog.pos_class.size());
    if (m == 0)
        return trx::set_error(TRX_EREGEX, "pattern '%s' matches only the empty string (the reference's "
                                          "FindAndConsume loop never terminates on it)", pattern);
    auto re = std::make_unique<trx_regex>();
    re->pattern = pat;
    static std::atomic<uint64_t> next_serial{1};
    re->serial = next_serial.fetch_add(1);
    re->positions = m;
    re->words = m <= 64 ? 1 : (m <= 128 ? 2 : 4);
    re->posix_longest = posix_longest != 0;
    re->anchor_begin = anchor_begin;
    re->anchor_end = anchor_end;
    auto & mk = re->masks;
    auto setbit = [](uint64_t * w, int b) { w[b >> 6] |= 1ULL << (b & 63); };

    std::vector<char> seen(prog.code.size());
    re->first_list.clear();
    prog.closure(0, seen, re->first_list);
    for (uint8_t q : re->first_list)
    {
        if (q == trx::kAccept)
            return trx::set_error(TRX_EREGEX, "pattern '%s' can match the empty string (the reference's "
                                              "FindAndConsume loop never terminates on it)", pattern);
        setbit(mk.first, q);
    }
    re->follow_off.assign(m + 1, 0);
    for (int p = 0; p < m; ++p)
    {
        std::fill(seen.begin(), seen.end(), 0);
        std::vector<uint8_t> lst;
        prog.closure(prog.pos_pc[p] + 1, seen, lst);
        re->follow_off[p] = static_cast<uint16_t>(re->follow_list.size());
        for (uint8_t q : lst)
        {
            if (q == trx::kAccept)
                setbit(mk.last, p);
            else
                setbit(mk.follow[p], q);
            re->follow_list.push_back(q);
        }
    }
    re->follow_off[m] = static_cast<uint16_t>(re->follow_list.size());
    for (int p = 0; p < m; ++p)
        for (unsigned c = 0; c < 256; ++c)
            if (prog.pos_class[p].has(c))
                setbit(mk.B[c], p);
    *out = re.release();
    return TRX_OK;
}

void trx_regex_destroy(trx_regex_t * re)
{
    delete re;
}

int trx_regex_positions(const trx_regex_t * re)
{
    return re ? re->positions : 0;
}

// Host only, for output formatting: where the EARLIEST-ENDING match of `re` in text[0, len) ends (exclusive), -1 if there
// is none.  RE2::Set::Match reports the rules of a conjunction in the order in which their first matches complete
// (verify_fasta_set prints them in that order, src/query.cpp:318-339); the records concerned are the handful that K4
// found to match every rule.  One step of the position automaton per byte: S' = (Follow(S) | First) & B[c].
int64_t trx_regex_first_match_end(const trx_regex_t * re, const uint8_t * text, uint64_t len)
{
    if (!re || (!text && len))
        return -1;
    trx::regex_masks const & m = re->masks;
    uint64_t s[trx::kRegexWords] = {};
    for (uint64_t i = 0; i < len; ++i)
    {
        uint64_t nx[trx::kRegexWords] = {};
        for (int w = 0; w < trx::kRegexWords; ++w)
        {
            uint64_t bits = s[w];
            while (bits)
            {
                int const b = __builtin_ctzll(bits);
                bits &= bits - 1;
                for (int k = 0; k < trx::kRegexWords; ++k)
                    nx[k] |= m.follow[w * 64 + b][k];
            }
        }
        bool hit = false;
        for (int k = 0; k < trx::kRegexWords; ++k)
        {
            if (!re->anchor_begin || i == 0)
                nx[k] |= m.first[k];
            s[k] = nx[k] & m.B[text[i]][k];
            hit = hit || (s[k] & m.last[k]);
        }
        if (hit && (!re->anchor_end || i + 1 == len))
            return static_cast<int64_t>(i + 1);
    }
    return -1;
}

} // extern "C"
