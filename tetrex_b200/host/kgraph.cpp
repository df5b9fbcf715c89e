// kgraph.cpp — postfix regex → linearised acyclic NFA ("k-graph") and its topological ranks.
//
// Behavioural restatement of construct_kgraph (src/construct_nfa.cpp:1-335) + update_arc_map
// (src/construction_tools.cpp:136-158) + OTFCollector::determine_top_sort (include/otf_collector.h:328-339).
// Which k-mers a query probes is decided here, so node numbering, arc order and the two-slot arc map follow
// the reference step for step (tests/test_host_cpu.py compares the result with the compiled reference
// node by node):  nodes are numbered in creation order; an arc from a non-split node overwrites both slots;
// a split node fills slot 1 then slot 2; '*' and '+' are unrolled to at most k-1 repetitions; {m,n} copies the
// operand; a quantifier whose operand is not the first thing on the stack concatenates immediately and eats the
// next '-' of the postfix string.
#include <algorithm>
#include <stack>

#include "trx_host.hpp"

namespace trxh
{

int kgraph::add_node(int sym)
{
    symbol.push_back(sym);
    first.push_back(-1);
    second.push_back(-1);
    return static_cast<int>(symbol.size()) - 1;
}

void kgraph::add_arc(int source, int target)
{
    arcs.emplace_back(source, target);
    int const sym = symbol[source];
    if (sym < Split || sym == Gap) // not a split: the same target in both slots (a later arc overwrites)
    {
        first[source] = target;
        second[source] = target;
    }
    else if (first[source] < 0) // first arc of a split: slot 1 (slot 2 stays invalid until the next arc)
    {
        first[source] = target;
    }
    else
    {
        second[source] = target;
    }
}

void kgraph::compute_ranks()
{
    // lemon::topologicalSort: DFS from every unreached node in descending id order, out-arcs newest first;
    // a node gets number N-1, N-2, … in the order it is LEFT.  The collector's rank equals that number (the write
    // map skips the node numbered 0 — the start node — which therefore keeps rank 0).
    int const n = static_cast<int>(symbol.size());
    std::vector<std::vector<int>> out(n); // arc ids per source, creation order
    for (int a = 0; a < static_cast<int>(arcs.size()); ++a)
        out[arcs[a].first].push_back(a);
    rank.assign(n, 0);
    std::vector<char> reached(n, 0);
    int num = n;
    for (int root = n - 1; root >= 0; --root)
    {
        if (reached[root])
            continue;
        reached[root] = 1;
        std::vector<std::pair<int, int>> stack; // (node, next out-arc position counted from the newest)
        stack.emplace_back(root, static_cast<int>(out[root].size()) - 1);
        while (!stack.empty())
        {
            auto & [node, pos] = stack.back();
            if (pos < 0)
            {
                rank[node] = --num;
                stack.pop_back();
                continue;
            }
            int const target = arcs[out[node][pos]].second;
            --pos;
            if (!reached[target])
            {
                reached[target] = 1;
                stack.emplace_back(target, static_cast<int>(out[target].size()) - 1);
            }
        }
    }
}

namespace
{

struct subgraph
{
    int start{-1}, end{-1};
};

// `reduced` selects construct_reduced_kgraph (src/construct_reduced_nfa.cpp), which the reference uses whenever the
// index has a reduced alphabet (include/query.h:239-240).  It creates symbol nodes lazily: a symbol only pushes a
// placeholder pair ("twins", start == end) and its character onto a side buffer; the node is created when an
// operator first needs it (twin_procedure), and an alternation of two placeholders with the same reduced character
// collapses into one placeholder (:146-154).  That code leans on unspecified behaviour in several places (empty
// std::stack::top(), stack underflow for {m} with m >= 2); those inputs raise trxh::error here.
struct builder
{
    kgraph g;
    std::stack<subgraph> stack;
    bool reduced{false};
    std::vector<int> buffer;

    static bool twins(subgraph const & s) { return s.start == s.end; }

    // twin_procedure (src/construct_reduced_nfa.cpp:107-112)
    void materialise(subgraph & s)
    {
        if (!reduced || !twins(s))
            return;
        if (buffer.empty())
            throw error("regular expression not supported on a reduced alphabet (symbol buffer underflow in the reference)");
        int sym = buffer.back();
        buffer.pop_back();
        int node = g.add_node(sym);
        s = {node, node};
    }

    subgraph pop()
    {
        if (stack.empty())
            throw error("malformed regular expression (operator without operand)");
        subgraph s = stack.top();
        stack.pop();
        return s;
    }

    // copy_subgraph (src/construct_nfa.cpp:4-71): duplicates every node that lies on a path start → end.  New
    // nodes are created in DESCENDING id order of the originals and arcs re-added in DESCENDING arc id order,
    // which is what lemon's NodeIt / ArcIt deliver.
    subgraph copy(subgraph const & sg)
    {
        if (sg.start == sg.end)
        {
            int sym;
            if (!reduced)
                sym = g.symbol[sg.start];
            else if (!buffer.empty())
                sym = buffer.back(); // src/construct_reduced_nfa.cpp:12 — whatever is on top of the buffer
            else
                throw error("regular expression not supported on a reduced alphabet (empty symbol buffer in the reference)");
            int twin = g.add_node(sym);
            return {twin, twin};
        }
        int const n = static_cast<int>(g.size());
        int const n_arcs = static_cast<int>(g.arcs.size());
        std::vector<std::vector<int>> fwd(n), bwd(n);
        for (auto const & [u, v] : g.arcs)
        {
            fwd[u].push_back(v);
            bwd[v].push_back(u);
        }
        auto reach = [&](int from, std::vector<std::vector<int>> const & adj) {
            std::vector<char> seen(n, 0);
            std::vector<int> todo{from};
            seen[from] = 1;
            while (!todo.empty())
            {
                int u = todo.back();
                todo.pop_back();
                for (int v : adj[u])
                    if (!seen[v])
                    {
                        seen[v] = 1;
                        todo.push_back(v);
                    }
            }
            return seen;
        };
        std::vector<char> from_s = reach(sg.start, fwd), to_t = reach(sg.end, bwd);
        std::vector<int> twin(n, -1);
        for (int u = n - 1; u >= 0; --u)
            if (from_s[u] && to_t[u])
                twin[u] = g.add_node(g.symbol[u]);
        for (int a = n_arcs - 1; a >= 0; --a)
        {
            auto [u, v] = g.arcs[a];
            if (twin[u] >= 0 && twin[v] >= 0)
                g.add_arc(twin[u], twin[v]);
        }
        return {twin[sg.start], twin[sg.end]};
    }

    void symbol(int c)
    {
        if (reduced)
        {
            buffer.push_back(c);
            stack.push({-1, -1}); // placeholder twins
            return;
        }
        int node = g.add_node(c);
        stack.push({node, node});
    }

    void concat()
    {
        subgraph b = pop(), a = pop();
        materialise(b);
        materialise(a);
        g.add_arc(a.end, b.start);
        stack.push({a.start, b.end});
    }

    void alternation()
    {
        subgraph b = pop(), a = pop();
        if (reduced)
        {
            size_t const n = buffer.size();
            bool const redundant = n >= 2 && buffer[n - 1] == buffer[n - 2];
            if (twins(a) && twins(b) && redundant)
            {
                int sym = buffer.back();
                buffer.pop_back();
                buffer.pop_back();
                symbol(sym);
                return;
            }
            materialise(a); // takes the top of the buffer, i.e. b's character when both are placeholders
            materialise(b);
        }
        int split = g.add_node(Split);
        g.add_arc(split, a.start);
        g.add_arc(split, b.start);
        int ghost = g.add_node(Ghost);
        g.add_arc(a.end, ghost);
        g.add_arc(b.end, ghost);
        stack.push({split, ghost});
    }

    void optional()
    {
        subgraph a = pop();
        materialise(a);
        int split = g.add_node(Split);
        g.add_arc(split, a.start);
        int ghost = g.add_node(Ghost);
        g.add_arc(split, ghost);
        g.add_arc(a.end, ghost);
        stack.push({split, ghost});
    }

    // '*' unrolled: ε, X, …, X^(k-1)   (src/construct_nfa.cpp:146-192)
    void kleene(unsigned k)
    {
        subgraph a = pop();
        materialise(a);
        int split = g.add_node(Split);
        g.add_arc(split, a.start);
        int ghost = g.add_node(Ghost);
        g.add_arc(split, ghost);
        int back = a.end;
        for (unsigned i = 1; i + 1 < k; ++i)
        {
            int inner = g.add_node(Split);
            g.add_arc(inner, ghost);
            subgraph c = copy(a);
            g.add_arc(back, inner);
            g.add_arc(inner, c.start);
            if (i + 2 == k)
            {
                g.add_arc(c.end, ghost);
                break;
            }
            back = c.end;
        }
        stack.push({split, ghost});
    }

    // '+' unrolled: X, …, X^(k-1)   (src/construct_nfa.cpp:195-223)
    void plus(unsigned k)
    {
        subgraph a = pop();
        materialise(a);
        int ghost = g.add_node(Ghost);
        int back = a.end;
        for (unsigned i = 1; i + 1 < k; ++i)
        {
            int inner = g.add_node(Split);
            subgraph c = copy(a);
            g.add_arc(back, inner);
            g.add_arc(inner, ghost);
            g.add_arc(inner, c.start);
            if (i + 2 == k)
            {
                g.add_arc(c.end, ghost);
                break;
            }
            back = c.end;
        }
        stack.push({a.start, ghost});
    }

    // {m,n} / {m}   (src/construct_nfa.cpp:226-262); returns true if the next '-' has to be skipped
    bool quantifier(size_t lo, size_t hi)
    {
        bool skip = false;
        if (lo == 0)
        {
            kleene(static_cast<unsigned>(hi + 1));
            if (stack.size() != 1)
            {
                concat();
                skip = true;
            }
            return skip;
        }
        if (stack.empty())
            throw error("malformed regular expression (quantifier without operand)");
        subgraph operand = stack.top();
        bool const was_twins = reduced && twins(operand);
        int const sym = was_twins && !buffer.empty() ? buffer.back() : 0;
        if (stack.size() != 1)
        {
            concat();
            if (was_twins)
                symbol(sym); // src/construct_reduced_nfa.cpp:289
            skip = true;
        }
        size_t const extra = hi == 0 ? 0 : hi - lo;
        for (size_t i = 1; i < lo; ++i)
        {
            if (!reduced)
                stack.push(copy(operand));
            else if (stack.size() < 2) // the reduced builder concatenates without pushing a copy (:293-300)
                throw error("quantifier {m} with m >= 2 is not supported on a reduced alphabet (stack underflow in the reference)");
            concat();
        }
        for (size_t i = 0; i < extra; ++i)
        {
            stack.push(copy(operand));
            optional();
            concat();
        }
        return skip;
    }
};

// parse_quant (src/construction_tools.cpp:4-18): {m} → (m, 0), {m,n} → (m, n)
std::pair<size_t, size_t> parse_quant(std::string const & postfix, size_t at)
{
    size_t comma = postfix.find(',', at);
    size_t end = postfix.find('}', at);
    if (end == std::string::npos)
        throw error("malformed quantifier in postfix expression");
    if (comma == std::string::npos || comma > end)
        return {static_cast<size_t>(std::stoul(postfix.substr(at + 1, end - at - 1))), 0};
    return {static_cast<size_t>(std::stoul(postfix.substr(at + 1, comma - at - 1))),
            static_cast<size_t>(std::stoul(postfix.substr(comma + 1, end - comma - 1)))};
}

} // namespace

kgraph construct_kgraph(std::string const & postfix, uint8_t k, bool reduced)
{
    if (postfix.empty())
        throw error("empty regular expression after pre-processing"); // the reference runs into stack.top() on an empty stack here
    builder b;
    b.reduced = reduced;
    int const start_node = b.g.add_node(Ghost);
    bool skip = false;
    for (size_t i = 0; i < postfix.size(); ++i)
    {
        int const sym = static_cast<unsigned char>(postfix[i]);
        if (isdigit(sym))
            continue;
        switch (sym)
        {
        case '-':
            if (skip)
            {
                skip = false;
                continue;
            }
            b.concat();
            break;
        case '|': b.alternation(); break;
        case '?': b.optional(); break;
        case '*': b.kleene(k); break;
        case '+': b.plus(k); break;
        case '{':
        {
            auto [lo, hi] = parse_quant(postfix, i);
            if (lo == 0 && hi == 1)
                b.optional();
            else
                skip = b.quantifier(lo, hi);
            break;
        }
        case '}':
        case ',': break;
        default: b.symbol(sym); break;
        }
    }
    // The reference caps whatever is on TOP of the stack (src/construct_nfa.cpp:316-333, construct_reduced_nfa.cpp:
    // 374-391; its `assert(stack.size() == 0)` is compiled out).  The reduced builder really gets there with operands
    // left over: `S{1}` after a prefix concatenates, pushes the symbol again and swallows the next '-' (:284-291), so
    // for "NLS{1}G" the graph that is walked is only "SG" — deterministic, and reproduced here node for node (the
    // orphaned prefix nodes stay in the graph and in the rank order, nothing points to them).
    if (b.stack.empty() || (!reduced && b.stack.size() != 1))
        throw error("malformed regular expression (operands left on the stack)");
    subgraph whole = b.stack.top();
    if (whole.start < 0)
        throw error("a single-symbol regular expression is not supported on a reduced alphabet");
    b.g.add_arc(start_node, whole.start);
    int const match = b.g.add_node(Match);
    b.g.add_arc(whole.end, match);
    b.g.compute_ranks();
    return std::move(b.g);
}

} // namespace trxh
