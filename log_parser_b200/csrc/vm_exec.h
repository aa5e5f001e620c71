// Bounded backtracking interpreter shared by the device verifier (k_verify_vm) and the host-side
// test hook (vm_match_host). Decides Matcher.find() on one line for patterns the DFA path cannot
// represent (AnalysisService.java:93-95 with look-around, back-references, possessive/atomic
// constructs, MULTILINE anchors, or counting that explodes a DFA). Priority order is Java's:
// leftmost start, ordered alternation, greedy before lazy. Positions are byte offsets into UTF-8;
// sets are matched on decoded code points.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define LPX_HD __host__ __device__ __forceinline__
#else
#define LPX_HD inline
#endif

namespace lpx {

enum VmOp : uint8_t {
    VM_CHAR = 0,      // a = byte
    VM_SET,           // a = set index
    VM_SPLIT,         // a = preferred pc, b = other pc
    VM_JMP,           // a = pc
    VM_SAVE,          // a = slot            slots[a] = pos (undoable)
    VM_GROUP_END,     // a = group, b = slot holding the group's start
    VM_ASSERT,        // a = AssertKind, flag = unix_lines
    VM_BACKREF,       // a = group, flag = icase
    VM_GUARD,         // a = mark slot, b = exit pc: an iteration that consumed nothing leaves the loop
    VM_CURLY,         // a = set, b = min, flag = mode (0 greedy 1 lazy 2 possessive); next insn VM_DATA a = max (-1 unbounded)
    VM_DATA,
    VM_ATOMIC_BEGIN,
    VM_ATOMIC_END,
    VM_LOOK_BEGIN,    // flag bit0 = negative, bit1 = the condition is tested at the start of the match (a deferred leading look-around); a = pc after the matching VM_LOOK_END
    VM_LOOK_END,
    VM_BACK,          // a = code points to step back, b = slot with the position to step back from
    VM_POSEQ,         // a = slot: pos must equal slots[a]
    VM_ABORT_MARK,    // from here on, running out of alternatives ends the attempt at this start position (see build_vm)
    VM_MATCH
};

struct VmCode {   // same layout as lpx::VmInsn
    uint8_t op, flag;
    uint16_t aux;
    int32_t a, b;
};

struct VmView {
    const VmCode* code;
    const uint32_t* set_off;     // per set: offset into set_data; set_off[i+1] bounds it
    const uint32_t* set_data;    // per set: 4 words ASCII bitmap, then (lo,hi) pairs
    uint32_t n_slots;
    const uint32_t* first;       // 8 words: bytes a match can begin with, or nullptr (no skipping)
};

constexpr uint32_t VM_MAX_SLOTS = 64;
constexpr uint32_t VM_STACK_WORDS = 4096;     // backtracking stack per (line, regex) evaluation: 2048 entries
constexpr int VM_MAX_STEPS = 1 << 26;          // interpreter steps per (line, regex) evaluation
constexpr uint32_t VM_NONE = 0xFFFFFFFFu;

// stack entry kinds (top 4 bits of word 0)
enum : uint32_t { VK_CHOICE = 1, VK_UNDO, VK_BARRIER, VK_LOOKPOS, VK_LOOKNEG, VK_RANGE, VK_LAZY, VK_DATA, VK_ABORT };

LPX_HD uint32_t vm_decode(const uint8_t* s, uint32_t pos, uint32_t le, uint32_t& len) {
    const uint32_t c = s[pos];
    if (c < 0x80) { len = 1; return c; }
    const uint32_t need = c >= 0xF0 ? 3 : c >= 0xE0 ? 2 : c >= 0xC0 ? 1 : 0;
    if (need == 0 || pos + need >= le) { len = 1; return 0xFFFD; }   // stray continuation byte / truncated sequence
    uint32_t cp = need == 1 ? (c & 0x1F) : need == 2 ? (c & 0x0F) : (c & 0x07);
    for (uint32_t k = 1; k <= need; ++k) cp = (cp << 6) | (s[pos + k] & 0x3F);
    len = need + 1;
    return cp;
}
// start of the code point that ends just before `pos` (pos > ls)
LPX_HD uint32_t vm_prev_start(const uint8_t* s, uint32_t pos, uint32_t ls) {
    uint32_t p = pos - 1;
    while (p > ls && (s[p] & 0xC0) == 0x80 && pos - p < 4) --p;
    return p;
}
LPX_HD bool vm_set_has(const VmView& v, uint32_t set, uint32_t cp) {
    const uint32_t* d = v.set_data + v.set_off[set];
    if (cp < 128) return (d[cp >> 5] >> (cp & 31)) & 1u;
    const uint32_t n = (v.set_off[set + 1] - v.set_off[set] - 4) / 2;
    d += 4;
    for (uint32_t i = 0; i < n; ++i) {
        if (cp < d[2 * i]) return false;
        if (cp <= d[2 * i + 1]) return true;
    }
    return false;
}
LPX_HD bool vm_word(uint32_t c) { return (c >= '0' && c <= '9') || (c >= 'A' && c <= 'Z') || c == '_' || (c >= 'a' && c <= 'z'); }
LPX_HD bool vm_term(uint32_t c) { return c == '\n' || c == '\r' || c == 0x85 || c == 0x2028 || c == 0x2029; }

// ASCII fast path of the interpreter's byte loops (a `.*` inside a look-ahead walks the rest of the line at every start
// position; one dependent load per byte, plus the set's words, made that walk the whole cost of a C3 evaluation).
// m[0..3]: a 128-bit set of ASCII bytes. Returns how many bytes from p on are ASCII members of the set, looking at no
// more than `limit` bytes and not beyond `end`. On the device the text is read a word at a time.
LPX_HD bool vm_ascii_in(uint32_t c, uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
    const uint32_t w = (c & 64u) ? ((c & 32u) ? m3 : m2) : ((c & 32u) ? m1 : m0);
    return c < 128u && ((w >> (c & 31u)) & 1u);
}
LPX_HD uint32_t vm_run_ascii(const uint8_t* s, uint32_t p, uint32_t end, uint32_t limit, uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
    const uint32_t p0 = p;
    if (end - p > limit) end = p + limit;
#ifdef __CUDA_ARCH__
    while (p < end) {
        if ((p & 3u) == 0 && p + 4 <= end) {
            const uint32_t w = *reinterpret_cast<const uint32_t*>(s + p);
            uint32_t k = 0;
            if (vm_ascii_in(w & 0xFFu, m0, m1, m2, m3)) { k = 1;
                if (vm_ascii_in((w >> 8) & 0xFFu, m0, m1, m2, m3)) { k = 2;
                    if (vm_ascii_in((w >> 16) & 0xFFu, m0, m1, m2, m3)) { k = 3;
                        if (vm_ascii_in(w >> 24, m0, m1, m2, m3)) k = 4; } } }
            p += k;
            if (k < 4) break;
        } else {
            if (!vm_ascii_in(s[p], m0, m1, m2, m3)) break;
            ++p;
        }
    }
#else
    while (p < end && vm_ascii_in(s[p], m0, m1, m2, m3)) ++p;
#endif
    return p - p0;
}
// the same walking down: how many bytes before p (not below `lo`) are ASCII members of the set
LPX_HD uint32_t vm_run_ascii_back(const uint8_t* s, uint32_t p, uint32_t lo, uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
    const uint32_t p0 = p;
#ifdef __CUDA_ARCH__
    while (p > lo) {
        if ((p & 3u) == 0 && p >= lo + 4) {
            const uint32_t w = *reinterpret_cast<const uint32_t*>(s + p - 4);
            uint32_t k = 0;
            if (vm_ascii_in(w >> 24, m0, m1, m2, m3)) { k = 1;
                if (vm_ascii_in((w >> 16) & 0xFFu, m0, m1, m2, m3)) { k = 2;
                    if (vm_ascii_in((w >> 8) & 0xFFu, m0, m1, m2, m3)) { k = 3;
                        if (vm_ascii_in(w & 0xFFu, m0, m1, m2, m3)) k = 4; } } }
            p -= k;
            if (k < 4) break;
        } else {
            if (!vm_ascii_in(s[p - 1], m0, m1, m2, m3)) break;
            --p;
        }
    }
#else
    while (p > lo && vm_ascii_in(s[p - 1], m0, m1, m2, m3)) --p;
#endif
    return p0 - p;
}

LPX_HD bool vm_assert(int kind, bool unix_lines, const uint8_t* s, uint32_t ls, uint32_t le, uint32_t pos, long tpos) {
    switch (kind) {
        case 0: return pos == ls;                                   // Begin
        case 1: return pos == le;                                   // End
        case 2: return pos == le || (long)pos == tpos;              // Eol ($, \Z): end or before the final terminator
        case 3: {                                                   // BolMulti (Caret / UnixCaret)
            if (pos == le) return false;
            if (pos > ls) {
                uint32_t l; const uint32_t pc = vm_decode(s, vm_prev_start(s, pos, ls), le, l);
                if (unix_lines) return pc == '\n';
                if (!vm_term(pc)) return false;
                if (pc == '\r' && s[pos] == '\n') return false;
            }
            return true;
        }
        case 4: {                                                   // EolMulti (Dollar / UnixDollar, MULTILINE)
            if (pos == le) return true;
            uint32_t l; const uint32_t c = vm_decode(s, pos, le, l);
            if (unix_lines) return c == '\n';
            if (c == '\n') return !(pos > ls && s[pos - 1] == '\r');
            return c == '\r' || c == 0x85 || c == 0x2028 || c == 0x2029;
        }
        default: {                                                  // WordB (5) / NotWordB (6)
            const bool l = pos > ls && vm_word(s[pos - 1]);
            const bool r = pos < le && vm_word(s[pos]);
            return kind == 5 ? (l != r) : (l == r);
        }
    }
}

// Runs the program anchored at `start`. Returns 1 match, 0 no match, -1 budget exhausted.
// `stack` holds `cap` 32-bit words (anything indexable: a plain array on the host, a two-level stack on the device
// that starts small and asks vm_stack_grow for its deep part when it runs out); `steps` is decremented.
template <class Stack>
LPX_HD bool vm_stack_grow(Stack&, uint32_t&) { return false; }      // a plain array cannot grow
template <class Stack>
LPX_HD int vm_run(const VmView& v, const uint8_t* s, uint32_t ls, uint32_t le, uint32_t start, long tpos,
                  uint32_t* slots, Stack stack, uint32_t cap, int& steps, uint32_t* end_out = nullptr) {
    uint32_t sp = 0, pc = 0, pos = start;
    bool aborting = false;
#define VM_PUSH(w0, w1) do { if (sp + 2 > cap && !vm_stack_grow(stack, cap)) return -1; stack[sp++] = (w0); stack[sp++] = (w1); } while (0)
    for (;;) {
        if (--steps < 0) return -1;
        const VmCode in = v.code[pc];
        bool fail = false;
        switch (in.op) {
            case VM_CHAR:
                if (pos < le && s[pos] == (uint8_t)in.a) { ++pos; ++pc; } else fail = true;
                break;
            case VM_SET: {
                if (pos >= le) { fail = true; break; }
                uint32_t l; const uint32_t cp = vm_decode(s, pos, le, l);
                if (vm_set_has(v, (uint32_t)in.a, cp)) { pos += l; ++pc; } else fail = true;
                break;
            }
            case VM_SPLIT:
                VM_PUSH((VK_CHOICE << 28) | (uint32_t)in.b, pos);
                pc = (uint32_t)in.a;
                break;
            case VM_JMP: pc = (uint32_t)in.a; break;
            case VM_SAVE:
                VM_PUSH((VK_UNDO << 28) | (uint32_t)in.a, slots[in.a]);
                slots[in.a] = pos; ++pc;
                break;
            case VM_GROUP_END: {
                const uint32_t g = (uint32_t)in.a;
                VM_PUSH((VK_UNDO << 28) | (2 * g), slots[2 * g]);
                VM_PUSH((VK_UNDO << 28) | (2 * g + 1), slots[2 * g + 1]);
                slots[2 * g] = slots[in.b]; slots[2 * g + 1] = pos; ++pc;
                break;
            }
            case VM_ASSERT:
                if (vm_assert(in.a, in.flag != 0, s, ls, le, pos, tpos)) ++pc; else fail = true;
                break;
            case VM_BACKREF: {
                const uint32_t b = slots[2 * in.a], e = slots[2 * in.a + 1];
                if (b == VM_NONE || e == VM_NONE) { fail = true; break; }
                const uint32_t n = e - b;
                if (pos + n > le) { fail = true; break; }
                bool ok = true;
                for (uint32_t i = 0; i < n && ok; ++i) {
                    uint8_t x = s[b + i], y = s[pos + i];
                    if (x != y) {
                        if (in.flag) { if (x >= 'A' && x <= 'Z') x += 32; if (y >= 'A' && y <= 'Z') y += 32; }
                        ok = x == y && ((x >= 'a' && x <= 'z') || s[b + i] == s[pos + i]);
                    }
                }
                if (ok) { pos += n; ++pc; } else fail = true;
                break;
            }
            case VM_GUARD:
                pc = (pos == slots[in.a]) ? (uint32_t)in.b : pc + 1;
                break;
            case VM_CURLY: {
                const uint32_t set = (uint32_t)in.a;
                const int mn = in.b, mx = v.code[pc + 1].a;
                int cnt = 0;
                uint32_t p = pos, after_min = pos;
                const int first_stop = in.flag == 1 ? mn : (mx < 0 ? 0x7FFFFFFF : mx);
                const uint32_t* const sd = v.set_data + v.set_off[set];
                const uint32_t a0 = sd[0], a1 = sd[1], a2 = sd[2], a3 = sd[3];      // the set's ASCII part
                while (cnt < first_stop && p < le) {
                    // ASCII members in one go (up to the minimum first, so that after_min is exact), anything else one code point at a time
                    const int want = cnt < mn ? mn - cnt : first_stop - cnt;
                    const uint32_t k = vm_run_ascii(s, p, le, (uint32_t)want, a0, a1, a2, a3);
                    if (k) {
                        p += k; cnt += (int)k;
                        if (cnt == mn) after_min = p;
                        continue;
                    }
                    uint32_t l; const uint32_t cp = vm_decode(s, p, le, l);
                    if (cp < 128 || !vm_set_has(v, set, cp)) break;
                    p += l; ++cnt;
                    if (cnt == mn) after_min = p;
                }
                if (cnt < mn) { fail = true; break; }
                if (mn == 0) after_min = pos;
                if (in.flag == 0) {            // greedy: give back one code point at a time
                    if (p > after_min) { VM_PUSH((VK_DATA << 28), after_min); VM_PUSH((VK_RANGE << 28) | (pc + 2), p); }
                } else if (in.flag == 1) {     // lazy: take one more on demand
                    const int room = mx < 0 ? 0x7FFFFFFF : mx - cnt;
                    if (room > 0) { VM_PUSH((VK_DATA << 28) | set, (uint32_t)room); VM_PUSH((VK_LAZY << 28) | (pc + 2), p); }
                }
                pos = p; pc += 2;
                break;
            }
            case VM_ATOMIC_BEGIN: VM_PUSH((VK_BARRIER << 28), 0); ++pc; break;
            case VM_ATOMIC_END: {
                // commit: forget every alternative created since the barrier. Java runs the atom as a
                // stand-alone match, so what it captured is never rolled back either: the undo log
                // of the group is dropped with the choices.
                uint32_t r = sp;
                while (r > 0 && (stack[r - 2] >> 28) != VK_BARRIER) r -= 2;
                sp = r - 2; ++pc;
                break;
            }
            case VM_LOOK_BEGIN:
                if (in.flag & 1) VM_PUSH((VK_LOOKNEG << 28) | (uint32_t)in.a, pos);
                else VM_PUSH((VK_LOOKPOS << 28), pos);
                if (in.flag & 2) pos = start;
                ++pc;
                break;
            case VM_LOOK_END: {
                // the condition matched; like the atomic group, its captures stay whatever follows
                uint32_t r = sp;
                while (r > 0 && (stack[r - 2] >> 28) != VK_LOOKPOS && (stack[r - 2] >> 28) != VK_LOOKNEG) r -= 2;
                const uint32_t frame = r - 2;
                const bool negative = (stack[frame] >> 28) == VK_LOOKNEG;
                pos = stack[frame + 1];
                sp = frame;
                if (negative) fail = true; else ++pc;
                break;
            }
            case VM_BACK: {
                uint32_t p = slots[in.b];
                bool ok = true;
                for (int k = 0; k < in.a; ++k) {
                    if (p <= ls) { ok = false; break; }
                    p = vm_prev_start(s, p, ls);
                }
                if (ok) { pos = p; ++pc; } else fail = true;
                break;
            }
            case VM_POSEQ:
                if (pos == slots[in.a]) ++pc; else fail = true;
                break;
            case VM_ABORT_MARK: VM_PUSH((VK_ABORT << 28), 0); ++pc; break;
            case VM_MATCH: if (end_out) *end_out = pos; return 1;
            default: return -1;
        }
        if (!fail) continue;
        // ---- backtrack
        for (;;) {
            if (sp == 0) return 0;
            sp -= 2;
            const uint32_t w0 = stack[sp], w1 = stack[sp + 1];
            const uint32_t kind = w0 >> 28, arg = w0 & 0x0FFFFFFFu;
            if (kind == VK_UNDO) { slots[arg] = w1; continue; }
            if (kind == VK_ABORT) aborting = true;      // a deferred leading look-around said no: no alternative of what came before
            if (aborting) continue;                     // can change that; unwind (captures are restored), then the attempt fails
            if (kind == VK_CHOICE || kind == VK_LOOKNEG) { pc = arg; pos = w1; break; }
            if (kind == VK_RANGE) {
                const uint32_t lo = stack[sp - 1];
                uint32_t hi = vm_prev_start(s, w1, ls);
                // The continuation begins with a literal ASCII byte: give back straight to its previous
                // occurrence. Every position skipped would fail at that VM_CHAR at once, with no side
                // effect; an ASCII byte is always a code point boundary.
                if (v.code[arg].op == VM_CHAR) {
                    const uint32_t c = (uint8_t)v.code[arg].a;
                    if (c < 128 && hi > lo && s[hi] != c) {      // walk down over "any byte but c" a word at a time
                        const uint32_t bit = 1u << (c & 31u);
                        const uint32_t n0 = ~((c >> 5) == 0 ? bit : 0u), n1 = ~((c >> 5) == 1 ? bit : 0u), n2 = ~((c >> 5) == 2 ? bit : 0u), n3 = ~((c >> 5) == 3 ? bit : 0u);
                        for (;;) {
                            hi -= vm_run_ascii_back(s, hi + 1, lo + 1, n0, n1, n2, n3);        // bytes in (lo, hi] that are ASCII and not c
                            if (hi <= lo || s[hi] == c) break;
                            --hi;                                                               // a non-ASCII byte: step over it
                        }
                    } else while (hi > lo && s[hi] != c) --hi;
                }
                if (hi > lo) { stack[sp + 1] = hi; sp += 2; }   // keep the entry for the next give-back
                else sp -= 2;                                   // exhausted: drop the DATA word too
                pc = arg; pos = hi;
                break;
            }
            if (kind == VK_LAZY) {
                const uint32_t set = stack[sp - 2] & 0x0FFFFFFFu;
                uint32_t room = stack[sp - 1];
                if (w1 < le) {
                    uint32_t l; const uint32_t cp = vm_decode(s, w1, le, l);
                    if (vm_set_has(v, set, cp)) {
                        --room;
                        pos = w1 + l; pc = arg;
                        if (room > 0) { stack[sp - 1] = room; stack[sp + 1] = pos; sp += 2; }
                        else sp -= 2;
                        break;
                    }
                }
                sp -= 2;
                continue;
            }
            // VK_BARRIER, VK_LOOKPOS, VK_DATA: nothing to resume
        }
    }
#undef VM_PUSH
}

// Matcher.find(): try every start position (code point boundaries), leftmost first.
// span (optional, 2 words): Matcher.start() / Matcher.end() of that first match, as byte offsets into s.
template <class Stack>
LPX_HD int vm_find(const VmView& v, const uint8_t* s, uint32_t ls, uint32_t le, Stack stack, uint32_t cap, int max_steps,
                   uint32_t* span = nullptr, int* steps_used = nullptr) {
    long tpos = -1;
    const uint32_t n = le - ls;
    if (n >= 1 && s[le - 1] == '\r') tpos = (long)le - 1;
    else if (n >= 2 && s[le - 2] == 0xC2 && s[le - 1] == 0x85) tpos = (long)le - 2;
    else if (n >= 3 && s[le - 3] == 0xE2 && s[le - 2] == 0x80 && (s[le - 1] == 0xA8 || s[le - 1] == 0xA9)) tpos = (long)le - 3;
    uint32_t slots[VM_MAX_SLOTS];
    int steps = max_steps;
    uint32_t start = ls;
    for (uint32_t i = 0; i < v.n_slots; ++i) slots[i] = VM_NONE;   // once per find(), as Matcher.search() does
    uint32_t f[8];
    bool skip = v.first != nullptr;
    if (skip) {
        uint32_t all = ~0u;
        for (int i = 0; i < 8; ++i) { f[i] = v.first[i]; all &= f[i]; }
        skip = all != ~0u;
    }
    for (;;) {
        if (skip) {       // next start position whose byte a match can begin with (none: no match; such patterns are not nullable)
            for (;;) {      // ASCII bytes that no match begins with, a word at a time; other bytes one by one
                start += vm_run_ascii(s, start, le, le - start, ~f[0], ~f[1], ~f[2], ~f[3]);
                if (start >= le || ((f[s[start] >> 5] >> (s[start] & 31u)) & 1u)) break;
                ++start;
            }
            if (start >= le) return 0;
        }
        const int r = vm_run(v, s, ls, le, start, tpos, slots, stack, cap, steps, span ? span + 1 : nullptr);
        if (r > 0 && span) span[0] = start;
        if (steps_used) *steps_used = max_steps - steps;
        if (r != 0) return r;
        if (start >= le) return 0;
        uint32_t l;
        vm_decode(s, start, le, l);
        start += l;
    }
}

}  // namespace lpx
