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

#include "rh_order.hpp"
#include "trx_host.hpp"

namespace trxh
{

int kgraph::add_node(int sym)
{
    symbol.push_back(sym);
    first.push_back(-1);
    second.push_back(-1);
    gap.push_back(0);
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
    int const n_arcs = static_cast<int>(arcs.size());
    // arc ids per source in creation order, as compressed rows: out[off[u] .. off[u + 1])
    std::vector<int> off(n + 1, 0), out(n_arcs), at;
    for (int a = 0; a < n_arcs; ++a)
        ++off[arcs[a].first + 1];
    for (int i = 0; i < n; ++i)
        off[i + 1] += off[i];
    at.assign(off.begin(), off.end() - 1);
    for (int a = 0; a < n_arcs; ++a)
        out[at[arcs[a].first]++] = a;
    rank.assign(n, 0);
    std::vector<char> reached(n, 0);
    std::vector<std::pair<int, int>> stack; // (node, next out-arc position counted from the newest)
    int num = n;
    for (int root = n - 1; root >= 0; --root)
    {
        if (reached[root])
            continue;
        reached[root] = 1;
        stack.clear();
        stack.emplace_back(root, off[root + 1] - off[root] - 1);
        while (!stack.empty())
        {
            auto & [node, pos] = stack.back();
            if (pos < 0)
            {
                rank[node] = --num;
                stack.pop_back();
                continue;
            }
            int const target = arcs[out[off[node] + pos]].second;
            --pos;
            if (!reached[target])
            {
                reached[target] = 1;
                stack.emplace_back(target, off[target + 1] - off[target] - 1);
            }
        }
    }
}

void kgraph::augment()
{
    if (catsites.empty())
        return;
    // merge_catsites (include/otf_collector.h:438-464): by rank of the entry node; a catsite that starts right after the
    // previous one ends is folded into it (gaps = all pairwise sums)
    std::vector<catsite> cats = catsites;
    std::sort(cats.begin(), cats.end(), [&](catsite const & a, catsite const & b) { return rank[a.cleavage_start] < rank[b.cleavage_start]; });
    {
        std::vector<catsite> merged;
        bool done = false;
        for (catsite const & c : cats)
        {
            if (merged.empty() || static_cast<size_t>(rank[c.cleavage_start]) - 1 != static_cast<size_t>(rank[merged.back().cleavage_end]))
            {
                merged.push_back(c);
                continue;
            }
            merged.back().cleavage_end = c.cleavage_end;
            rh_slots sums; // sumGaps (:426-436)
            std::vector<uint64_t> a, b;
            merged.back().gaps.for_each_key([&](uint64_t k) { a.push_back(k); });
            c.gaps.for_each_key([&](uint64_t k) { b.push_back(k); });
            for (uint64_t x : a)
                for (uint64_t y : b)
                    sums.insert_key(x + y);
            merged.back().gaps = sums;
            done = true;
        }
        if (done)
            cats = merged;
    }
    for (catsite & c : cats)
    {
        std::vector<uint64_t> gaps;
        c.gaps.for_each_key([&](uint64_t k) { gaps.push_back(k); });
        c.downstream = first[c.cleavage_end]; // complete(): arcs.find(id(cleavage_end_))->second.first
        auto add_gap = [&](uint64_t gapsize) {
            int const node = add_node(Gap);
            gap[node] = gapsize;
            add_arc(c.cleavage_site, node);
            add_arc(node, c.downstream);
        };
        if (gaps.size() == 1)
            add_gap(gaps[0]);
        else
        {
            // add_guard (:407-418): a Split in front, a Ghost behind; the gaps hang between them.  A split only has two
            // slots in the arc map, so of three or more gaps the walk sees the first and the LAST one added — as in
            // the reference
            int const split = add_node(Split);
            int const ghost = add_node(Ghost);
            add_arc(c.cleavage_site, split);
            add_arc(ghost, c.downstream);
            c.cleavage_site = split;
            c.downstream = ghost;
            for (uint64_t gsz : gaps)
                add_gap(gsz);
        }
    }
    compute_ranks();
}

namespace
{

// Subgraph's bookkeeping for catsite detection (include/construction_tools.h:74-146): number of paths through the
// subgraph and the set of their lengths.  `lengths` is a robin_hood::unordered_set in the reference and its ITERATION
// order is visible (it decides the order in which augment() creates Gap nodes, and which gap `*gaps.begin()` is), so it
// is kept in a table that places keys exactly like that container does (rh_order.hpp), filled by the same inserts in
// the same order.
struct subgraph
{
    int start{-1}, end{-1};
    size_t split_run_count{0};
    size_t paths{1};
    rh_slots lengths;

    static subgraph twins(int node) // {node, node, 0u, 1u, {1u}, Default}
    {
        subgraph s{node, node};
        s.lengths.insert_key(1);
        return s;
    }
    std::vector<uint64_t> length_list() const
    {
        std::vector<uint64_t> v;
        lengths.for_each_key([&](uint64_t k) { v.push_back(k); });
        return v;
    }
    void concat_info(subgraph const & a, subgraph const & b)
    {
        paths = a.paths * b.paths;
        std::vector<uint64_t> const x = a.length_list(), y = b.length_list();
        for (uint64_t l : x)
            for (uint64_t j : y)
                lengths.insert_key(l + j);
        split_run_count = std::max(a.split_run_count, b.split_run_count);
    }
    void union_info(subgraph const & a, subgraph const & b)
    {
        // std::set_union over the two sets' iteration orders (which are not sorted: the merge is what it is)
        std::vector<uint64_t> const x = a.length_list(), y = b.length_list();
        size_t i = 0, j = 0;
        while (i < x.size() && j < y.size())
        {
            if (x[i] < y[j])
                lengths.insert_key(x[i++]);
            else if (y[j] < x[i])
                lengths.insert_key(y[j++]);
            else
            {
                lengths.insert_key(x[i++]);
                ++j;
            }
        }
        for (; i < x.size(); ++i)
            lengths.insert_key(x[i]);
        for (; j < y.size(); ++j)
            lengths.insert_key(y[j]);
        split_run_count = a.split_run_count + b.split_run_count + 1;
        paths = a.paths + b.paths;
    }
    void option_info(subgraph const & a)
    {
        lengths.insert_key(0);
        for (uint64_t l : a.length_list())
            lengths.insert_key(l);
        split_run_count = a.split_run_count + 1;
        paths = a.paths + 1;
    }
    void kleene_info(subgraph const & a, uint8_t repeats)
    {
        std::vector<uint64_t> const x = a.length_list();
        for (size_t i = 0; i < repeats; ++i)
            for (uint64_t l : x)
                lengths.insert_key(i * l);
        split_run_count = a.split_run_count;
        paths = a.paths * repeats;
    }
    void copy_meta(subgraph const & a)
    {
        split_run_count = a.split_run_count;
        paths = a.paths;
        lengths = a.lengths;
    }
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
    std::vector<catsite> cats;

    static bool twins(subgraph const & s) { return s.start == s.end; }

    // detect_bad_graphs (src/construction_tools.cpp:161-180)
    void detect_bad_graphs(subgraph const & a, subgraph const & b, subgraph const & joined)
    {
        if (b.paths >= 15 || (joined.paths >= 690000u && b.start != b.end))
        {
            catsite c;
            c.cleavage_site = a.end;
            c.cleavage_start = b.start;
            c.cleavage_end = b.end;
            c.gaps = b.lengths;
            cats.push_back(c);
        }
    }

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
        s = subgraph::twins(node);
    }

    subgraph pop()
    {
        if (stack.empty())
            throw error("malformed regular expression (operator without operand)");
        subgraph s = std::move(stack.top());
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
            subgraph c{twin, twin};
            c.copy_meta(sg);
            return c;
        }
        int const n = static_cast<int>(g.size());
        int const n_arcs = static_cast<int>(g.arcs.size());
        // adjacency in compressed rows (three flat arrays per direction instead of a vector per node: this runs once per
        // repetition of every {m,n} and the graphs have a hundred nodes)
        auto reach = [&](int from, bool forward) {
            std::vector<int> off(n + 1, 0), adj(n_arcs), at;
            for (auto const & [u, v] : g.arcs)
                ++off[(forward ? u : v) + 1];
            for (int i = 0; i < n; ++i)
                off[i + 1] += off[i];
            at.assign(off.begin(), off.end() - 1);
            for (auto const & [u, v] : g.arcs)
                adj[at[forward ? u : v]++] = forward ? v : u;
            std::vector<char> seen(n, 0);
            std::vector<int> todo{from};
            seen[from] = 1;
            while (!todo.empty())
            {
                int u = todo.back();
                todo.pop_back();
                for (int e = off[u]; e < off[u + 1]; ++e)
                    if (!seen[adj[e]])
                    {
                        seen[adj[e]] = 1;
                        todo.push_back(adj[e]);
                    }
            }
            return seen;
        };
        std::vector<char> from_s = reach(sg.start, true), to_t = reach(sg.end, false);
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
        subgraph c{twin[sg.start], twin[sg.end]};
        c.copy_meta(sg);
        return c;
    }

    void symbol(int c)
    {
        if (reduced)
        {
            buffer.push_back(c);
            stack.push(subgraph::twins(-1)); // placeholder twins
            return;
        }
        int node = g.add_node(c);
        stack.push(subgraph::twins(node));
    }

    void concat()
    {
        subgraph b = pop(), a = pop();
        materialise(b);
        materialise(a);
        g.add_arc(a.end, b.start);
        subgraph joined{a.start, b.end};
        joined.concat_info(a, b);
        detect_bad_graphs(a, b, joined);
        stack.push(std::move(joined));
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
        subgraph u{split, ghost};
        u.union_info(a, b);
        stack.push(std::move(u));
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
        subgraph o{split, ghost};
        o.option_info(a);
        stack.push(std::move(o));
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
        subgraph ks{split, ghost};
        ks.kleene_info(a, static_cast<uint8_t>(k));
        stack.push(std::move(ks));
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
        stack.push(subgraph{a.start, ghost}); // no meta: Subgraph stack_subgraph{subgraph.start, ghost_node} (:221)
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
    b.g.catsites = std::move(b.cats);
    return std::move(b.g);
}

} // namespace trxh
