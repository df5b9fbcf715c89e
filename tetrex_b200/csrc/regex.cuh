// regex.cuh — compiled form of a verify regex shared by the host compiler and the K4 scan kernels.
//
// The automaton is the Glushkov (position) automaton of the pattern: one state per symbol position
// after {m,n} expansion, at most TRX_REGEX_MAX_POSITIONS = 128 → two 64-bit words of state.  All
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

constexpr int kRegexWords = 2; // 128 positions
constexpr uint8_t kAccept = 0xFF;

struct regex_masks
{
    uint64_t first[kRegexWords];
    uint64_t last[kRegexWords];
    uint64_t B[256][kRegexWords];                       // symbol → positions it may occupy
    uint64_t follow[64 * kRegexWords][kRegexWords];     // position → follow set
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
};
