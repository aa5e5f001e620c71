// AST -> backtracking bytecode for the patterns the DFA path cannot decide (vm.h / vm_exec.h).
// Emission order encodes java.util.regex priorities: first alternative first, greedy loops try
// the body before the exit, lazy loops the exit first; capturing groups publish start and end
// together at the group's tail (GroupHead/GroupTail); an empty loop iteration leaves the loop
// (Loop.match "i > beginIndex" test); possessive quantifiers are atomic greedy loops.
#include "vm.h"

#include <stdexcept>

#include "vm_exec.h"

namespace lpx {
namespace {

constexpr size_t MAX_CODE = 60000;

struct Emitter {
    VmProgram& p;
    uint32_t next_slot;
    explicit Emitter(VmProgram& prog, int n_groups) : p(prog), next_slot(2 * (uint32_t)n_groups) {}

    int pc() const { return (int)p.code.size(); }
    bool seek_start = false;     // the next look-around generated tests its condition at the start of the match
    int emit(uint8_t op, int a = 0, int b = 0, uint8_t flag = 0) {
        if (p.code.size() >= MAX_CODE) throw RegexUnsupported("pattern expands to too much backtracking code");
        VmInsn i{};
        i.op = op; i.flag = flag; i.aux = 0; i.a = a; i.b = b;
        p.code.push_back(i);
        return pc() - 1;
    }
    uint32_t slot() {
        if (next_slot >= VM_MAX_SLOTS) throw RegexUnsupported("pattern needs more than 64 capture/loop slots");
        return next_slot++;
    }
    int add_set(const CpSet& s) {
        uint32_t bitmap[4] = {0, 0, 0, 0};
        std::vector<uint32_t> hi;
        for (const auto& r : s) {
            for (uint32_t c = r.lo; c <= r.hi && c < 128; ++c) bitmap[c >> 5] |= 1u << (c & 31);
            if (r.hi >= 128) { hi.push_back(std::max<uint32_t>(r.lo, 128)); hi.push_back(r.hi); }
        }
        if (p.set_off.empty()) p.set_off.push_back(0);
        for (uint32_t w : bitmap) p.set_ranges.push_back(w);
        p.set_ranges.insert(p.set_ranges.end(), hi.begin(), hi.end());
        p.set_off.push_back((uint32_t)p.set_ranges.size());
        return (int)p.set_off.size() - 2;
    }

    static bool single_set(const Ast& a) { return a.kind == AstKind::Set; }

    void gen_set(const Ast& a) {
        if (a.set.size() == 1 && a.set[0].lo == a.set[0].hi && a.set[0].lo < 128) emit(VM_CHAR, (int)a.set[0].lo);
        else emit(VM_SET, add_set(a.set));
    }

    // Curly.match with Qtype.POSSESSIVE runs every iteration as a stand-alone match of the atom
    // (atom.match(...) against an accepting tail): an iteration, once matched, is never re-entered.
    void gen_body(const Ast& body, bool commit_each) {
        if (commit_each) emit(VM_ATOMIC_BEGIN);
        gen(body);
        if (commit_each) emit(VM_ATOMIC_END);
    }

    // X{lo,hi} for a general body; mode 0 greedy, 1 lazy (possessive: greedy + commit_each, wrapped by
    // the caller). Every iteration — mandatory ones included — is followed by Loop.match's test:
    // an iteration that consumed nothing goes straight on to whatever follows the loop.
    void gen_repeat(const Ast& body, int lo, int hi, int mode, bool commit_each = false) {
        const uint32_t mark = slot();
        std::vector<int> to_exit_b;     // instructions whose .b is the loop exit
        std::vector<int> to_exit_a;     // instructions whose .a is the loop exit
        auto iteration = [&]() {
            emit(VM_SAVE, (int)mark);
            gen_body(body, commit_each);
            to_exit_b.push_back(emit(VM_GUARD, (int)mark, 0));
        };
        for (int i = 0; i < lo; ++i) iteration();
        if (hi < 0) {
            const int loop = emit(VM_SPLIT);
            const int body_pc = pc();
            iteration();
            emit(VM_JMP, loop);
            if (mode == 1) { to_exit_a.push_back(loop); p.code[loop].b = body_pc; }
            else { p.code[loop].a = body_pc; to_exit_b.push_back(loop); }
        } else {
            for (int i = lo; i < hi; ++i) {
                const int sp = emit(VM_SPLIT);
                const int body_pc = pc();
                if (mode == 1) { to_exit_a.push_back(sp); p.code[sp].b = body_pc; }
                else { p.code[sp].a = body_pc; to_exit_b.push_back(sp); }
                iteration();
            }
        }
        const int exit_pc = pc();
        for (int i : to_exit_a) p.code[i].a = exit_pc;
        for (int i : to_exit_b) p.code[i].b = exit_pc;
    }

    void gen(const Ast& a) {
        switch (a.kind) {
            case AstKind::Empty: break;
            case AstKind::Set: gen_set(a); break;
            case AstKind::Cat: for (const auto& k : a.kids) gen(*k); break;
            case AstKind::Alt: {
                std::vector<int> jumps;
                for (size_t i = 0; i < a.kids.size(); ++i) {
                    if (i + 1 < a.kids.size()) {
                        const int sp = emit(VM_SPLIT);
                        p.code[sp].a = pc();
                        gen(*a.kids[i]);
                        jumps.push_back(emit(VM_JMP));
                        p.code[sp].b = pc();
                    } else gen(*a.kids[i]);
                }
                for (int j : jumps) p.code[j].a = pc();
                break;
            }
            case AstKind::Group: {
                if (a.group < 0) { gen(*a.kids[0]); break; }
                const uint32_t tmp = slot();
                emit(VM_SAVE, (int)tmp);
                gen(*a.kids[0]);
                emit(VM_GROUP_END, a.group, (int)tmp);
                break;
            }
            case AstKind::Repeat: {
                const Ast& body = *a.kids[0];
                if (single_set(body)) {
                    emit(VM_CURLY, add_set(body.set), a.min, (uint8_t)a.mode);
                    emit(VM_DATA, a.max);
                    break;
                }
                if (a.mode == 2) {
                    emit(VM_ATOMIC_BEGIN);
                    gen_repeat(body, a.min, a.max, 0, true);
                    emit(VM_ATOMIC_END);
                } else gen_repeat(body, a.min, a.max, a.mode);
                break;
            }
            case AstKind::Assert: emit(VM_ASSERT, (int)a.assert_kind, 0, a.unix_lines ? 1 : 0); break;
            case AstKind::Atomic:
                emit(VM_ATOMIC_BEGIN);
                gen(*a.kids[0]);
                emit(VM_ATOMIC_END);
                break;
            case AstKind::Backref: emit(VM_BACKREF, a.group, 0, a.icase ? 1 : 0); break;
            case AstKind::Look: {
                const Ast& body = *a.kids[0];
                const int begin = emit(VM_LOOK_BEGIN, 0, 0, (a.neg ? 1 : 0) | (seek_start ? 2 : 0));
                seek_start = false;
                if (!a.behind) {
                    gen(body);
                } else {
                    // Pattern.Behind tries the start positions from i-min down to i-max (code points)
                    if (body.max_len < 0) throw RegexSyntaxError("Look-behind group does not have an obvious maximum length", 0);
                    if (body.max_len - body.min_len > 64) throw RegexUnsupported("look-behind spanning more than 64 different lengths");
                    const uint32_t here = slot();
                    emit(VM_SAVE, (int)here);
                    std::vector<int> jumps;
                    for (int k = body.min_len; k <= body.max_len; ++k) {
                        int sp = -1;
                        if (k < body.max_len) { sp = emit(VM_SPLIT); p.code[sp].a = pc(); }
                        emit(VM_BACK, k, (int)here);
                        gen(body);
                        emit(VM_POSEQ, (int)here);
                        if (k < body.max_len) { jumps.push_back(emit(VM_JMP)); p.code[sp].b = pc(); }
                    }
                    for (int j : jumps) p.code[j].a = pc();
                }
                emit(VM_LOOK_END);
                p.code[begin].a = pc();
                break;
            }
        }
    }
};

}  // namespace

namespace {
struct FirstSet { uint32_t b[8] = {0, 0, 0, 0, 0, 0, 0, 0}; bool nullable = true; bool unknown = false; };
void fs_add(FirstSet& f, uint32_t byte) { f.b[byte >> 5] |= 1u << (byte & 31u); }
uint32_t lead_byte(uint32_t c) { return c < 0x80 ? c : c < 0x800 ? 0xC0 | (c >> 6) : c < 0x10000 ? 0xE0 | (c >> 12) : 0xF0 | (c >> 18); }
FirstSet first_set(const Ast& a) {
    FirstSet r;
    switch (a.kind) {
        case AstKind::Empty: case AstKind::Assert: case AstKind::Look: return r;          // zero-width
        case AstKind::Backref: r.unknown = true; return r;
        case AstKind::Set:
            r.nullable = false;
            for (const auto& x : a.set) {
                const uint32_t hi = std::min<uint32_t>(x.hi, 0x10FFFF);
                if (x.lo > hi) continue;
                // lead bytes are monotonic in the code point: every lead byte between the two ends can occur. 0x80..0xC1
                // lie between the ASCII and the multi-byte lead bytes and are none: with them in the set, the search for
                // the next start position would stop inside a character whose own lead byte no match begins with.
                for (uint32_t b = lead_byte(x.lo); b <= lead_byte(hi); ++b)
                    if (b < 0x80 || b >= 0xC2) fs_add(r, b);
            }
            return r;
        case AstKind::Group: case AstKind::Atomic: return first_set(*a.kids[0]);
        case AstKind::Cat:
            for (const auto& k : a.kids) {
                FirstSet f = first_set(*k);
                for (int i = 0; i < 8; ++i) r.b[i] |= f.b[i];
                r.unknown |= f.unknown;
                if (!f.nullable) { r.nullable = false; break; }
            }
            return r;
        case AstKind::Alt:
            r.nullable = false;
            for (const auto& k : a.kids) {
                FirstSet f = first_set(*k);
                for (int i = 0; i < 8; ++i) r.b[i] |= f.b[i];
                r.unknown |= f.unknown;
                r.nullable |= f.nullable;
            }
            return r;
        case AstKind::Repeat: {
            r = first_set(*a.kids[0]);
            if (a.min == 0) r.nullable = true;
            return r;
        }
    }
    r.unknown = true;
    return r;
}
bool has_backref(const Ast& a);
bool has_capture(const Ast& a) {
    if (a.kind == AstKind::Group && a.group >= 0) return true;
    for (const auto& k : a.kids) if (has_capture(*k)) return true;
    return false;
}
bool has_unbounded_repeat(const Ast& a) {
    if (a.kind == AstKind::Repeat && a.max < 0) return true;
    for (const auto& k : a.kids) if (has_unbounded_repeat(*k)) return true;
    return false;
}
bool deferrable_look(const Ast& a) {
    return a.kind == AstKind::Look && !a.behind && !has_capture(*a.kids[0]) && !has_backref(*a.kids[0]) && has_unbounded_repeat(*a.kids[0]);
}
bool has_backref(const Ast& a) {
    if (a.kind == AstKind::Backref) return true;
    for (const auto& k : a.kids) if (has_backref(*k)) return true;
    return false;
}
// Constructs that java.util.regex does not backtrack into: a capture made inside one survives the failure of what
// follows it (Pattern.Pos / Neg / Behind / Ques-Curly possessive / atomic groups do not restore the groups), so an
// attempt that fails at one start position can leave a group set for the attempt at the next one.
bool keeps_captures_on_failure(const Ast& a) {
    if (a.kind == AstKind::Look || a.kind == AstKind::Atomic || (a.kind == AstKind::Repeat && a.mode == 2)) return true;
    for (const auto& k : a.kids) if (keeps_captures_on_failure(*k)) return true;
    return false;
}
}  // namespace

void build_vm(const ParsedRegex& rx, VmProgram& out) {
    out = VmProgram();
    {
        const FirstSet f = first_set(*rx.root);
        // Start positions whose byte no match can begin with are skipped. With a back-reference that is only sound when
        // failed attempts leave no captures behind: groups are cleared once per find(), not per start position, and
        // GroupTail / GroupCurly restore them on failure, the constructs named above do not.
        const bool leaky = has_backref(*rx.root) && keeps_captures_on_failure(*rx.root);
        if (!f.nullable && !f.unknown && !leaky) for (int i = 0; i < 8; ++i) out.first[i] = f.b[i];
    }
    Emitter e(out, rx.n_groups);
    // Leading look-arounds with an unbounded walk inside ((?=.*\[tag\])phrase ...) are tested after the rest of the
    // pattern has matched instead of before it at every start position. A look-around is zero-width and, without
    // groups or back-references inside, neither changes nor reads anything the rest can see, so "condition at the
    // start AND rest from the start" gives the same answer, start and end in either order; only the cost moves:
    // the rest usually fails after a byte or two. When the deferred condition fails no other way of matching the
    // rest can help (the condition does not depend on it), so the attempt at this start position ends
    // (VM_ABORT_MARK), as it would have in the written order.
    size_t lead = 0;
    const Ast& root = *rx.root;
    if (root.kind == AstKind::Cat)
        while (lead < root.kids.size() && deferrable_look(*root.kids[lead])) ++lead;
    if (lead > 0 && lead < root.kids.size()) {
        for (size_t k = lead; k < root.kids.size(); ++k) e.gen(*root.kids[k]);
        e.emit(VM_ABORT_MARK);
        for (size_t k = 0; k < lead; ++k) { e.seek_start = true; e.gen(*root.kids[k]); }
    } else e.gen(root);
    e.emit(VM_MATCH);
    if (out.set_off.empty()) out.set_off.push_back(0);
    out.n_caps = 2 * (uint32_t)rx.n_groups;
    out.n_slots = e.next_slot;
    if (out.n_slots > VM_MAX_SLOTS) throw RegexUnsupported("pattern needs more than 64 capture/loop slots");
}

int vm_match_host(const VmProgram& p, const uint8_t* s, size_t n) {
    static_assert(sizeof(VmInsn) == sizeof(VmCode), "bytecode layouts must agree");
    VmView v;
    v.code = reinterpret_cast<const VmCode*>(p.code.data());
    v.set_off = p.set_off.data();
    v.set_data = p.set_ranges.data();
    v.n_slots = p.n_slots;
    v.first = p.first;
    std::vector<uint32_t> stack(VM_STACK_WORDS);
    return vm_find(v, s, 0, (uint32_t)n, stack.data(), (uint32_t)stack.size(), VM_MAX_STEPS);
}
int vm_steps_host(const VmProgram& p, const uint8_t* s, size_t n) {
    VmView v;
    v.code = reinterpret_cast<const VmCode*>(p.code.data());
    v.set_off = p.set_off.data();
    v.set_data = p.set_ranges.data();
    v.n_slots = p.n_slots;
    v.first = p.first;
    std::vector<uint32_t> stack(VM_STACK_WORDS);
    int used = 0;
    vm_find(v, s, 0, (uint32_t)n, stack.data(), (uint32_t)stack.size(), VM_MAX_STEPS, nullptr, &used);
    return used;
}
int vm_span_host(const VmProgram& p, const uint8_t* s, size_t n, uint32_t* span) {
    VmView v;
    v.code = reinterpret_cast<const VmCode*>(p.code.data());
    v.set_off = p.set_off.data();
    v.set_data = p.set_ranges.data();
    v.n_slots = p.n_slots;
    v.first = p.first;
    std::vector<uint32_t> stack(VM_STACK_WORDS);
    return vm_find(v, s, 0, (uint32_t)n, stack.data(), (uint32_t)stack.size(), VM_MAX_STEPS, span);
}

}  // namespace lpx
