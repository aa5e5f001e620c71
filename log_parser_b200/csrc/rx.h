// Host-side regex compiler of the B200 path: java.util.regex dialect -> AST -> byte-level NFA
// -> DFA tables (regular patterns) or backtracking bytecode (lookaround, back-references,
// possessive/atomic), plus required-literal extraction for the device prefilter.
//
// Replaces the per-request Pattern.compile block of the reference
// (AnalysisService.java:56-86; context regexes ContextAnalysisService.java:27-34): compilation is
// a pure function of the pattern string, so it runs once per pattern-set load on the host.
// What the device must reproduce is Matcher.find() as a boolean on one line
// (AnalysisService.java:93-95, ScoringService.java:281,300,330, ContextAnalysisService.java:64-79).
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace lpx {

// Mirrors java.util.regex.PatternSyntaxException (bad pattern -> HTTP 500 in the reference).
struct RegexSyntaxError : std::runtime_error {
    int index;
    RegexSyntaxError(const std::string& m, int i) : std::runtime_error(m), index(i) {}
};
// Legal Java syntax this build does not implement; reported, never silently approximated.
struct RegexUnsupported : std::runtime_error {
    explicit RegexUnsupported(const std::string& m) : std::runtime_error(m) {}
};

enum : int {
    RXF_UNIX_LINES = 0x01, RXF_ICASE = 0x02, RXF_COMMENTS = 0x04, RXF_MULTILINE = 0x08,
    RXF_DOTALL = 0x20, RXF_UNICODE_CASE = 0x40, RXF_UNICODE_CLASS = 0x100
};

struct CpRange { uint32_t lo, hi; };
using CpSet = std::vector<CpRange>;  // sorted, disjoint

enum class AstKind : uint8_t { Empty, Set, Cat, Alt, Repeat, Group, Assert, Atomic, Look, Backref };
enum class AssertKind : uint8_t {
    Begin,      // \A and non-MULTILINE ^
    End,        // \z ; also (?d)$ / (?d)\Z on a line that cannot contain \n
    Eol,        // non-MULTILINE $ and \Z : end, or before one final line terminator
    BolMulti,   // (?m)^   (backtracking engine only)
    EolMulti,   // (?m)$   (backtracking engine only)
    WordB, NotWordB
};

struct Ast {
    AstKind kind = AstKind::Empty;
    CpSet set;                       // Set
    int min = 0, max = 0, mode = 0;  // Repeat (max < 0: unbounded; mode 0 greedy 1 lazy 2 possessive)
    int group = -1;                  // Group capture index / Backref target
    bool icase = false;              // Backref
    bool unix_lines = false;         // BolMulti / EolMulti
    AssertKind assert_kind = AssertKind::Begin;
    bool neg = false, behind = false;  // Look
    int min_len = 0, max_len = 0;      // in code points (max_len < 0 unbounded)
    std::vector<std::unique_ptr<Ast>> kids;
};
using AstP = std::unique_ptr<Ast>;

struct ParsedRegex {
    AstP root;
    int n_groups = 1;
    bool regular = true;  // decidable by the DFA path (existence only)
};

ParsedRegex parse_java_regex(const std::string& utf8, int flags);

// ---------------------------------------------------------------- set helpers
void cpset_normalize(CpSet& s);
CpSet cpset_negate(const CpSet& s);
CpSet cpset_intersect(const CpSet& a, const CpSet& b);
bool cpset_has(const CpSet& s, uint32_t c);

// ---------------------------------------------------------------- DFA
// Symbols fed by the device walker: byte classes 0..n_byte_classes-1, then two virtual symbols:
//   SYM_TPOS  "the next code point is the line's single trailing terminator" (enables `$`)
//   SYM_EOT   end of line
// State 0 is DEAD (no match possible), state 1 is MATCH (absorbing).
struct Dfa {
    uint16_t n_classes = 0;          // including the two virtual symbols
    uint16_t start = 0;
    uint16_t start_ctx[3] = {0, 0, 0};   // anchored automata: start state when the walk begins at the line edge / after a non-word byte / after a word byte
    uint8_t classmap[256];
    std::vector<uint16_t> trans;     // n_states * n_classes
    uint32_t n_states() const { return n_classes ? (uint32_t)(trans.size() / n_classes) : 0; }
    bool match(const uint8_t* s, size_t n) const;  // host-side walk, identical to the device walker
    // anchored automata (build_dfa_anchored), `s[0..n)` is the line: does a match START at `pos` /
    // END at `pos`? Same loops as the device (k_confirm).
    bool match_from(const uint8_t* s, size_t n, size_t pos) const;
    bool match_ending_at(const uint8_t* s, size_t n, size_t pos) const;
};
constexpr uint16_t DFA_DEAD = 0, DFA_MATCH = 1;

// max_raw_states bounds subset construction, max_states the minimised table the device walks.
struct DfaLimits { uint32_t max_raw_states = 30000; uint32_t max_states = 8192; uint32_t max_nfa = 60000; };
// Returns false (and leaves `out` empty) if the pattern exceeds the limits.
bool build_dfa(const Ast& root, Dfa& out, const DfaLimits& lim = DfaLimits());
// Automaton for verifying a match at a known literal occurrence instead of walking the whole line:
// reverse = false: matches anchored at the walk's first byte; reverse = true: the reversed language,
// walked from the end of a match towards the line start (refused for patterns using `$`/\Z).
bool build_dfa_anchored(const Ast& root, bool reverse, Dfa& out, const DfaLimits& lim = DfaLimits());
struct Literal;
// True when every match accepted by the anchored automaton starts with (forward) / ends with (reverse)
// a member of `lits` — so checking the literal occurrences alone decides Matcher.find() on the line.
bool literals_cover(const Dfa& anchored, bool reverse, const std::vector<Literal>& lits);
// 0: verify by walking the whole line; 1 / 2: `out` is the forward / backward anchored automaton and the
// literal occurrences alone decide the line.
int choose_anchor(const Ast& root, const std::vector<Literal>& lits, Dfa& out, const DfaLimits& lim = DfaLimits());

// Product automaton of several regexes for the always-on scan (patterns without a usable literal):
// one walk of a line decides all of them. A state carries the set of regexes already matched
// (sticky); state_mask[trans[s][SYM_EOT]] is the line's result. Symbols as in Dfa.
struct MultiDfa {
    uint16_t n_classes = 0;
    uint16_t start = 0;
    uint8_t classmap[256];
    std::vector<uint16_t> trans;        // n_states * n_classes
    std::vector<uint32_t> state_mask;   // bit k: regex k of the group has matched
    std::vector<uint8_t> state_final;   // 1: no transition can change the state any more
    uint32_t n_states() const { return n_classes ? (uint32_t)(trans.size() / n_classes) : 0; }
    uint32_t run(const uint8_t* s, size_t n) const;   // host-side walk identical to the device's
};
bool build_multi_dfa(const std::vector<const Ast*>& roots, MultiDfa& out, uint32_t max_states, uint32_t max_nfa = 60000);

// ---------------------------------------------------------------- required literals
struct Literal {
    std::string bytes;  // UTF-8; if icase: ASCII letters are stored lower-case
    bool icase = false;
    bool lossy = false; // case-sensitive letters were folded into a case-insensitive literal: superset only
};
// A set L such that every match of the regex contains at least one member of L as a substring.
// Empty result = no usable literal (every line is a candidate).
std::vector<Literal> required_literals(const Ast& root, size_t min_len);

// "Literal with word-boundary conditions" form: the regex is an alternation of branches
//   [\b] [\w*] LITERAL-SET [\b]
// so that a line matches iff some occurrence of one of the literals satisfies its boundary
// conditions. Such patterns (the context regexes of CAS:27-34, plain-string secondaries, ...) are
// decided at the literal occurrence itself, without walking the line through an automaton.
struct LocalLiteral {
    Literal lit;
    bool left_boundary = false;    // \b required immediately before the literal
    bool right_boundary = false;   // \b required immediately after it
};
bool local_verify_form(const Ast& root, std::vector<LocalLiteral>& out, size_t min_len);

}  // namespace lpx
