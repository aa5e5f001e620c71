// Backtracking bytecode for patterns a DFA cannot decide (look-around, back-references,
// possessive quantifiers, atomic groups, MULTILINE anchors). Compiled on the host, executed by
// the device verifier on prefiltered candidate lines only; vm_match_host() runs the identical
// interpreter on the host for the CPU-only test tier.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include "rx.h"

namespace lpx {

struct VmInsn {
    uint8_t op;
    uint8_t flag;
    uint16_t aux;
    int32_t a;
    int32_t b;
};

struct VmProgram {
    std::vector<VmInsn> code;
    std::vector<uint32_t> set_off;   // per set: offset into set_ranges (pairs lo,hi), size n+1
    std::vector<uint32_t> set_ranges;
    uint32_t n_slots = 0;            // capture slots + loop marks
    uint32_t n_caps = 0;             // 2 * groups
    // Bytes a match can begin with (bit b of first[b >> 5]); all ones when the pattern can match the
    // empty string or uses back-references. find() skips start positions whose byte is not in the set:
    // an attempt there fails before it consumes anything, and without back-references a failed attempt
    // leaves nothing behind that a later one could observe.
    uint32_t first[8] = {~0u, ~0u, ~0u, ~0u, ~0u, ~0u, ~0u, ~0u};
};

// Throws RegexUnsupported for constructs the bytecode cannot express.
void build_vm(const ParsedRegex& rx, VmProgram& out);

// 1 match, 0 no match, -1 backtracking budget exhausted.
int vm_match_host(const VmProgram& p, const uint8_t* s, size_t n);
// interpreter steps that find() took on this input (tuning aid)
int vm_steps_host(const VmProgram& p, const uint8_t* s, size_t n);
// Matcher.start()/end() of the first find() as byte offsets (span[0], span[1]); return value as vm_match_host.
int vm_span_host(const VmProgram& p, const uint8_t* s, size_t n, uint32_t* span);

}  // namespace lpx
