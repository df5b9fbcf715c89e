// regex.cuh — compiled form of a verify regex shared by the host compiler and the K4 scan kernels.
//
// The automaton is the Glushkov (position) automaton of the pattern: one state per symbol position
// after {m,n} expansion, at most TRX_REGEX_MAX_POSITIONS = 250 → one, two or four 64-bit words of state (the kernels are
// instantiated per width; position ids are bytes and 0xFF is ACCEPT, hence 250 and not 256).  All
// transitions into position i carry the same symbol class, so one step is
//      S' = (Follow(S) | inject·First) & B[c]          match ends here  ⇔  S' & Last ≠ 0
// which is the whole inner loop of the streaming pass.  For leftmost-first (Perl/RE2 default)
// resolution the same positions also carry priority-ORDERED first/follow lists (Pike-VM order,
// ACCEPT as an explicit list item).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

namespace trx
{

constexpr int kRegexWords = 4; // room for 256 positions
constexpr uint8_t kAccept = 0xFF;

struct regex_masks
{
    uint64_t first[kRegexWords];
    uint64_t last[kRegexWords];
    uint64_t B[256][kRegexWords];                       // symbol → positions it may occupy
    uint64_t follow[64 * kRegexWords][kRegexWords];     // position → follow set
};

} // namespace trx

namespace trx
{
// Subset automaton of the anchored (no re-injection) position automaton as one scan sees it (byte maps folded in):
// state 0 = dead, 1 = before the first byte, accepting states last.  usable = false: too many subsets, or the pattern
// is not POSIX leftmost-longest (leftmost-first needs thread priorities, which a DFA state does not carry).
struct dfa_table
{
    bool usable{false};
    uint32_t n_states{0}, n_classes{0}, first_accept{0};
    uint8_t cls[256]{};
    std::vector<uint16_t> trans; // n_states x n_classes
    std::vector<uint8_t> start_class; // [n_classes]: 1 = a byte of this class can begin a match
    // Reverse searcher over the same classes: read the text BACKWARDS with a new match end injected at every byte
    // (state 0 = nothing live); a state >= r_first_accept after consuming text[p] means "some match starts at p".
    // Only for patterns whose matches have a bounded length (max_len; no '*' / '+') and no anchors: a scan of one 128-byte
    // chunk warms up on the max_len - 1 bytes that follow it.
    bool r_usable{false};
    uint32_t r_states{0}, r_first_accept{0}, max_len{0};
    std::vector<uint16_t> r_trans; // r_states x n_classes
    // key of the cache below
    int reverse{0};
    uint64_t map_hash{0};
};
} // namespace trx

struct trx_regex
{
    int positions{0};
    int words{1};            // 1 if positions <= 64
    bool posix_longest{false};
    bool anchor_begin{false};
    bool anchor_end{false};
    trx::regex_masks masks{};
    // ordered lists, flattened (host copies; uploaded per scan)
    std::vector<uint8_t> first_list;
    std::vector<uint16_t> follow_off;
    std::vector<uint8_t> follow_list;
    std::string pattern;
    uint64_t serial{0}; // unique per compiled regex: key of the per-text filter-plan cache
    // determinised matchers built on first use, one per (strand, byte map) this regex was scanned with; a handle is used
    // by one host thread at a time (like every handle of this library)
    std::vector<trx::dfa_table> dfas;
};
