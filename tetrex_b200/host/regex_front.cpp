// regex_front.cpp — the per-query host work that decides which k-mers exist: query pre-processing
// (src/query.cpp:78-155) and infix → postfix translation (include/utils.h:96-468, src/utils.cpp:3-15).
// Out-of-perf-scope host code kept for CLI parity, and deliberately the CLOSEST thing to the reference in this
// repository: the tokenizer's control flow follows PostfixConverter's and the error strings are the reference's word
// for word, because `tetrex query` must accept, reject and report exactly what the reference does; only the
// representation (one function, own token types) is this repo's.  Same token classes, same precedences
// (| = 1 < concat = 2 < * + ? {} = 3), '-' as the explicit concatenation operator, '.' and [^…] expanded over the
// 20 amino acids in the reference's orders (a required constant).
#include <algorithm>
#include <iostream>

#include "trx_host.hpp"

namespace trxh
{

namespace
{

enum class tk { chr, dot, star, plus, quest, pipe, lparen, rparen, quant, cls, concat };

struct token
{
    tk type;
    char ch{};
    int lo{}, hi{};
    bool exact{};
    bool negated{};
    std::vector<char> members;
};

bool is_operand_end(tk t)
{
    return t == tk::chr || t == tk::dot || t == tk::cls || t == tk::rparen || t == tk::star || t == tk::plus ||
           t == tk::quest || t == tk::quant;
}

bool is_operand_begin(tk t)
{
    return t == tk::chr || t == tk::dot || t == tk::cls || t == tk::lparen;
}

int precedence(tk t)
{
    switch (t)
    {
    case tk::pipe: return 1;
    case tk::concat: return 2;
    case tk::star:
    case tk::plus:
    case tk::quest:
    case tk::quant: return 3;
    default: return 0;
    }
}

std::vector<token> tokenize(std::string const & in)
{
    std::vector<token> out;
    size_t pos = 0;
    auto number = [&]() {
        int v = 0;
        while (pos < in.size() && isdigit(static_cast<unsigned char>(in[pos])))
            v = v * 10 + (in[pos++] - '0');
        return v;
    };
    while (pos < in.size())
    {
        char c = in[pos];
        switch (c)
        {
        case '.': out.push_back({tk::dot}); ++pos; break;
        case '*': out.push_back({tk::star}); ++pos; break;
        case '+': out.push_back({tk::plus}); ++pos; break;
        case '?': out.push_back({tk::quest}); ++pos; break;
        case '|': out.push_back({tk::pipe}); ++pos; break;
        case '(': out.push_back({tk::lparen}); ++pos; break;
        case ')': out.push_back({tk::rparen}); ++pos; break;
        case '[':
        {
            ++pos;
            if (pos >= in.size())
                throw error("Invalid character class: unexpected end of input");
            token t{tk::cls};
            if (in[pos] == '^')
            {
                t.negated = true;
                ++pos;
            }
            while (pos < in.size() && in[pos] != ']')
            {
                char m = in[pos];
                if (m == '\\')
                {
                    ++pos;
                    if (pos >= in.size())
                        throw error("Invalid escape in character class");
                    m = in[pos];
                    if (m == 'n') m = '\n';
                    else if (m == 't') m = '\t';
                    else if (m == 'r') m = '\r';
                }
                t.members.push_back(m);
                ++pos;
            }
            if (pos >= in.size())
                throw error("Invalid character class: missing closing ']'");
            ++pos;
            if (t.members.empty())
                throw error("Empty character class");
            out.push_back(std::move(t));
            break;
        }
        case '{':
        {
            ++pos;
            if (pos >= in.size() || !isdigit(static_cast<unsigned char>(in[pos])))
                throw error("Invalid quantifier: expected number after '{'");
            token t{tk::quant};
            t.lo = number();
            if (pos >= in.size())
                throw error("Invalid quantifier: unexpected end of input");
            if (in[pos] == '}')
            {
                ++pos;
                t.hi = t.lo;
                t.exact = true;
            }
            else if (in[pos] == ',')
            {
                ++pos;
                if (pos >= in.size())
                    throw error("Invalid quantifier: unexpected end after ','");
                if (in[pos] == '}')
                    throw error("Open-ended quantifiers {m,} not supported");
                if (!isdigit(static_cast<unsigned char>(in[pos])))
                    throw error("Invalid quantifier: expected number after ','");
                t.hi = number();
                if (pos >= in.size() || in[pos] != '}')
                    throw error("Invalid quantifier: expected '}' after max value");
                ++pos;
                if (t.lo > t.hi)
                    throw error("Invalid quantifier: min > max");
            }
            else
                throw error("Invalid quantifier: expected ',' or '}' after min value");
            out.push_back(std::move(t));
            break;
        }
        case '\\':
            ++pos;
            if (pos >= in.size())
                throw error("Invalid escape: end of input after '\\'");
            out.push_back({tk::chr, in[pos]});
            ++pos;
            break;
        default: out.push_back({tk::chr, c}); ++pos; break;
        }
    }
    return out;
}

// an operand in postfix form: a, or "ab|c|" for a class (first member, then "member|" for each further one)
std::string operand_postfix(token const & t)
{
    static std::string const amino = "ACDEFGHIKLMNPQRSTVWY";
    switch (t.type)
    {
    case tk::chr: return std::string(1, t.ch);
    case tk::dot: return "FQ|L|T|K|P|A|Y|R|N|H|G|E|C|I|V|D|W|S|M|"; // the reference's fixed wildcard order
    case tk::cls:
    {
        std::vector<char> members = t.members;
        if (t.negated)
        {
            std::sort(members.begin(), members.end());
            std::vector<char> rest;
            std::set_difference(amino.begin(), amino.end(), members.begin(), members.end(), std::back_inserter(rest));
            members = rest;
        }
        std::string s(1, members.at(0));
        for (size_t i = 1; i < members.size(); ++i)
        {
            s += members[i];
            s += '|';
        }
        return s;
    }
    default: return {};
    }
}

std::string operator_postfix(token const & t)
{
    switch (t.type)
    {
    case tk::star: return "*";
    case tk::plus: return "+";
    case tk::quest: return "?";
    case tk::pipe: return "|";
    case tk::concat: return "-";
    case tk::quant:
        return t.exact ? "{" + std::to_string(t.lo) + "}" : "{" + std::to_string(t.lo) + "," + std::to_string(t.hi) + "}";
    default: return {};
    }
}

} // namespace

std::string translate(std::string const & pattern)
{
    std::string result;
    try
    {
        std::vector<token> raw = tokenize(pattern);
        std::vector<token> toks;
        for (size_t i = 0; i < raw.size(); ++i)
        {
            if (i > 0 && is_operand_end(raw[i - 1].type) && is_operand_begin(raw[i].type))
                toks.push_back({tk::concat});
            toks.push_back(raw[i]);
        }
        std::vector<token> ops;
        for (token const & t : toks)
        {
            if (t.type == tk::chr || t.type == tk::dot || t.type == tk::cls)
                result += operand_postfix(t);
            else if (t.type == tk::lparen)
                ops.push_back(t);
            else if (t.type == tk::rparen)
            {
                while (!ops.empty() && ops.back().type != tk::lparen)
                {
                    result += operator_postfix(ops.back());
                    ops.pop_back();
                }
                if (!ops.empty())
                    ops.pop_back();
            }
            else
            {
                while (!ops.empty() && ops.back().type != tk::lparen && precedence(ops.back().type) >= precedence(t.type))
                {
                    result += operator_postfix(ops.back());
                    ops.pop_back();
                }
                ops.push_back(t);
            }
        }
        while (!ops.empty())
        {
            result += operator_postfix(ops.back());
            ops.pop_back();
        }
    }
    catch (std::exception const & e)
    {
        std::cout << "Error: " << e.what() << "\n---\n"; // src/utils.cpp:11-14 prints and carries on with ""
        return {};
    }
    return result;
}

namespace
{

// length of the "uninformative" token that starts at pos, 0 if it is informative (src/query.cpp:78-119)
size_t uninformative(std::string const & s, size_t pos)
{
    char c = s[pos];
    if (c == '^' || c == '$')
        return 1;
    if (c == '.')
    {
        if (pos + 1 < s.size() && (s[pos + 1] == '*' || s[pos + 1] == '+'))
            return 2;
        if (pos + 1 < s.size() && s[pos + 1] == '{')
        {
            size_t end = s.find('}', pos + 2);
            if (end != std::string::npos)
                return end - pos + 1;
        }
        return 1;
    }
    if (c == '[')
    {
        size_t end = s.find(']', pos + 1);
        if (end != std::string::npos)
        {
            std::string inside = s.substr(pos + 1, end - pos - 1);
            if (!inside.empty() && (inside[0] == '^' || inside == "."))
                return end - pos + 1;
            if (inside.find('-') != std::string::npos)
                return end - pos + 1;
        }
    }
    return 0;
}

} // namespace

std::string trim_regex(std::string const & regex)
{
    size_t start = 0, end = regex.size();
    while (start < end)
    {
        size_t len = uninformative(regex, start);
        if (len == 0)
            break;
        start += len;
    }
    // from the back the reference looks at one character at a time (so ".{2}" or "[^P]" at the end stay)
    while (end > start)
    {
        size_t len = uninformative(regex, end - 1);
        if (len == 0)
            break;
        end -= len;
    }
    if (start > end) // a front token may reach past `end`; substr would throw in the reference too
        return {};
    return regex.substr(start, end - start);
}

std::string reduce_query_alphabet(std::string const & regex, uint8_t const redmap[256])
{
    std::string out = regex;
    for (char & c : out)
        if (isalpha(static_cast<unsigned char>(c)))
            c = static_cast<char>(redmap[static_cast<unsigned char>(c)]);
    return out;
}

} // namespace trxh
