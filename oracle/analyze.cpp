// ORACLE — TEST INFRASTRUCTURE ONLY. Not shipped, not linked by the product library.
//
// Line-by-line CPU restatement of the reference's hot path, in its own evaluation order
// (lazy, per-match, windowed regex finds — NOT the eager per-line restructuring the GPU
// path uses, so that the two derivations check each other):
//   AnalysisService.java:53        String.split("\\r?\\n")
//   AnalysisService.java:56-86     per-request regex compilation
//   AnalysisService.java:89-113    N x P primary scan, discovery order
//   AnalysisService.java:132-156   extractContext
//   ScoringService.java:63-112     7-factor product, left to right
//   ScoringService.java:123-151    chronological factor
//   ScoringService.java:161-190,315-347  proximity factor, closest secondary distance
//   ScoringService.java:199-305    temporal factor / sequences
//   ContextAnalysisService.java:27-34,46-144  context factor
//   FrequencyTrackingService.java:41-93       frequency penalty (deterministic model, below)
//
// Frequency model (PatternFrequency lives in the un-vendored com.redhat.podmortem:common
// 1.0-2112dab-SNAPSHOT, pom.xml:19,55-59; its source is absent => "parity unpinned"):
//   rate = (matches recorded inside the sliding window) / time-window-hours, with the count
//   read BEFORE the current match is recorded (SS:84-88); docs/SCORING_ALGORITHM.md:202 pins
//   only "seen 5 times in last hour, threshold 10 -> penalty 0.0". The caller passes the
//   in-window count at request start per pattern (prior_count).
#include <cmath>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>
#include <atomic>
#include <memory>

#include "jregex.hpp"
#include "oracle_api.h"

using orx::Regex;

namespace {

struct Lines {
    std::vector<uint64_t> start, end;  // byte offsets
    std::vector<int> cps;              // all lines decoded, concatenated
    std::vector<uint64_t> cp_off;      // per line offset into cps (size N+1)
};

// Pattern.split(input, 0) with the regex \r?\n  (AS:53)
void java_split(const uint8_t* s, size_t n, std::vector<uint64_t>& st, std::vector<uint64_t>& en) {
    st.clear(); en.clear();
    size_t index = 0;
    bool any = false;
    for (size_t i = 0; i < n; ++i) {
        if (s[i] != '\n') continue;
        size_t sep_start = (i > 0 && s[i - 1] == '\r' && i - 1 >= index) ? i - 1 : i;
        // String.split: a zero-length leading match never happens here (separator is >= 1 char)
        st.push_back(index); en.push_back(sep_start);
        index = i + 1;
        any = true;
    }
    if (!any) { st.push_back(0); en.push_back(n); return; }  // no match: [input]
    st.push_back(index); en.push_back(n);
    // limit == 0: strip all trailing empty strings
    while (!st.empty() && st.back() == en.back()) { st.pop_back(); en.pop_back(); }
}

struct Memo {
    std::vector<signed char> v;  // -1 unknown, 0/1
};

struct Rx {
    std::unique_ptr<Regex> re;
    Memo memo;
};

struct Env {
    const orc_cfg* cfg;
    const Lines* L;
    bool memo;
    uint64_t finds = 0;
    bool find(Rx& rx, uint64_t line) {
        if (memo) {
            signed char& c = rx.memo.v[line];
            if (c >= 0) return c != 0;
            bool r = rx.re->find_cp(L->cps.data() + L->cp_off[line], (int)(L->cp_off[line + 1] - L->cp_off[line]), nullptr, nullptr);
            c = r ? 1 : 0;
            return r;
        }
        ++finds;
        return rx.re->find_cp(L->cps.data() + L->cp_off[line], (int)(L->cp_off[line + 1] - L->cp_off[line]), nullptr, nullptr);
    }
};

struct CPat {
    const orc_pattern* p;
    Rx* primary;
    std::vector<Rx*> sec;
    std::vector<std::vector<Rx*>> seq;
    int freq_slot;  // -1: untracked
};

// SS:123-151
double chrono(const orc_cfg& c, uint64_t idx, uint64_t n) {
    double logPosition = (double)idx / (double)n;
    double f;
    if (logPosition <= c.early_bonus_threshold) {
        double bonusRange = c.max_early_bonus - 1.5;
        f = 1.5 + (c.early_bonus_threshold - logPosition) * (bonusRange / c.early_bonus_threshold);
    } else if (logPosition <= c.penalty_threshold) {
        double middleRange = c.penalty_threshold - c.early_bonus_threshold;
        f = 1.0 + (c.penalty_threshold - logPosition) * (0.5 / middleRange);
    } else {
        f = 0.5 + (1.0 - logPosition);
    }
    return f;
}

// SS:315-347
double closest_secondary(Env& e, Rx& rx, int window, int64_t primary, int64_t n) {
    int windowSize = std::min(e.cfg->max_window, window);
    int64_t start = std::max<int64_t>(0, primary - windowSize);
    int64_t end = std::min<int64_t>(n, primary + windowSize + 1);
    double closest = -1;
    for (int64_t line = start; line < end; ++line) {
        if (line == primary) continue;
        if (e.find(rx, (uint64_t)line)) {
            double d = (double)std::llabs(line - primary);
            if (closest < 0 || d < closest) closest = d;
        }
    }
    return closest;
}

// SS:161-190
double proximity(Env& e, CPat& cp, int64_t primary, int64_t n) {
    if (cp.sec.empty()) return 1.0;
    double total = 0.0;
    for (size_t s = 0; s < cp.sec.size(); ++s) {
        double d = closest_secondary(e, *cp.sec[s], cp.p->secondaries[s].proximity_window, primary, n);
        if (d >= 0) {
            double decay = std::exp(-d / e.cfg->decay_constant);
            total += cp.p->secondaries[s].weight * decay;
        }
    }
    return 1.0 + total;
}

// SS:272-286
bool near_primary(Env& e, Rx& rx, int64_t primary, int64_t n) {
    int windowSize = 5;
    int64_t start = std::max<int64_t>(0, primary - windowSize);
    int64_t end = std::min<int64_t>(n, primary + windowSize + 1);
    for (int64_t i = start; i < end; ++i)
        if (e.find(rx, (uint64_t)i)) return true;
    return false;
}
// SS:296-305
int64_t event_before(Env& e, Rx& rx, int64_t before) {
    for (int64_t i = before - 1; i >= 0; --i)
        if (e.find(rx, (uint64_t)i)) return i;
    return -1;
}
// SS:230-262
bool sequence_matched(Env& e, std::vector<Rx*>& events, int64_t primary, int64_t n) {
    if (events.empty()) return false;
    int64_t cur = 0;
    for (int i = (int)events.size() - 1; i >= 0; --i) {
        if (i == (int)events.size() - 1) {
            if (!near_primary(e, *events[i], primary, n)) return false;
            cur = primary;
        } else {
            int64_t f = event_before(e, *events[i], cur);
            if (f < 0) return false;
            cur = f;
        }
    }
    return true;
}
// SS:199-220
double temporal(Env& e, CPat& cp, int64_t primary, int64_t n) {
    if (cp.seq.empty()) return 1.0;
    double total = 0.0;
    for (size_t s = 0; s < cp.seq.size(); ++s)
        if (sequence_matched(e, cp.seq[s], primary, n)) total += cp.p->sequences[s].bonus_multiplier;
    return 1.0 + total;
}

struct CtxRx {
    Regex err{"\\b(ERROR|FATAL|CRITICAL|SEVERE)\\b", orx::F_CASE_INSENSITIVE};   // CAS:27-28
    Regex warn{"\\b(WARN|WARNING)\\b", orx::F_CASE_INSENSITIVE};                  // CAS:29-30
    Regex stack{"^\\s*at\\s+[\\w\\.\\$]+\\(.*\\)\\s*$"};                          // CAS:31-32
    Regex exc{"\\b\\w*Exception\\b|\\b\\w*Error\\b"};                              // CAS:33-34
};

// CAS:56-117 over an abstract list of lines; `flags(i)` returns bit0 ERR, bit1 WARN, bit2 STACK, bit3 EXC
template <class F>
double context_score(const orc_cfg& c, size_t total, F flags) {
    if (total == 0) return 1.0;
    double contextScore = 0.0;
    int errorLines = 0, stackTraceLines = 0;
    for (size_t i = 0; i < total; ++i) {
        int f = flags(i);
        if (f & 1) { errorLines++; contextScore += 0.4; }
        else if (f & 2) { contextScore += 0.2; }
        if (f & 4) { stackTraceLines++; contextScore += 0.1; }
        if (f & 8) { contextScore += 0.3; }
    }
    if (stackTraceLines > 0) contextScore += std::min(stackTraceLines * 0.1, 0.5);
    int totalLines = (int)total;
    if (totalLines > 10 && (stackTraceLines + errorLines) > totalLines * 0.7) contextScore *= 0.8;
    double factor = 1.0 + contextScore;
    if (factor > c.max_context_factor) factor = c.max_context_factor;
    return factor;
}

struct CtxMemo {
    std::vector<signed char> v;  // per line: -1 unknown else 4-bit flags
};

int ctx_flags_cp(CtxRx& cr, const int* t, int n) {
    int f = 0;
    // evaluation mirrors CAS:64-82 (WARN is only consulted when ERR did not match)
    if (cr.err.find_cp(t, n, nullptr, nullptr)) f |= 1;
    else if (cr.warn.find_cp(t, n, nullptr, nullptr)) f |= 2;
    if (cr.stack.find_cp(t, n, nullptr, nullptr)) f |= 4;
    if (cr.exc.find_cp(t, n, nullptr, nullptr)) f |= 8;
    return f;
}

// FTS:64-93
double freq_penalty(const orc_cfg& c, bool exists, double count) {
    if (!exists) return 0.0;
    double currentRate = count / (double)c.time_window_hours;
    if (currentRate <= c.frequency_threshold) return 0.0;
    double excess = currentRate - c.frequency_threshold;
    return std::min(c.max_penalty, excess / c.frequency_threshold);
}

bool blank_id(const char* id) {
    if (!id) return true;
    for (const char* p = id; *p; ++p)
        if ((unsigned char)*p > ' ') return false;  // String.trim(): chars <= U+0020
    return true;
}

struct Result {
    orc_result r;
    std::vector<uint32_t> line, pattern;
    std::vector<double> score, factors;
    std::vector<uint64_t> delta;
};

void do_analyze(const orc_cfg* cfg, const orc_pattern* pats, size_t n_pats, const uint8_t* log, size_t len,
                const uint64_t* prior, int options, Result& out) {
    // AS:53
    Lines L;
    java_split(log, len, L.start, L.end);
    uint64_t N = L.start.size();
    L.cp_off.resize(N + 1);
    {
        std::vector<int> tmp;
        L.cps.reserve(len);
        for (uint64_t i = 0; i < N; ++i) {
            L.cp_off[i] = L.cps.size();
            orx::decode_utf8(log + L.start[i], L.end[i] - L.start[i], tmp);
            L.cps.insert(L.cps.end(), tmp.begin(), tmp.end());
        }
        L.cp_off[N] = L.cps.size();
    }
    bool memo = (options & ORC_OPT_MEMO) != 0;
    bool first_opt = !(options & ORC_OPT_NO_FIRST_OPT);

    // AS:56-86 — every request recompiles every regex. With memo on, identical regex strings
    // share one compiled object + cache (pure function of the line, so results are unchanged).
    std::vector<std::unique_ptr<Rx>> pool;
    std::map<std::string, Rx*> dedupe;
    auto compile = [&](const char* rx) -> Rx* {
        if (memo) {
            auto it = dedupe.find(rx);
            if (it != dedupe.end()) return it->second;
        }
        std::unique_ptr<Rx> r(new Rx());
        r->re.reset(new Regex(rx, 0));
        r->re->set_first_opt(first_opt);
        if (memo) r->memo.v.assign(N, -1);
        Rx* raw = r.get();
        pool.push_back(std::move(r));
        if (memo) dedupe[rx] = raw;
        return raw;
    };
    std::vector<CPat> cps(n_pats);
    std::map<std::string, int> slots;
    std::vector<int> slot_first;
    for (size_t p = 0; p < n_pats; ++p) {
        CPat& c = cps[p];
        c.p = &pats[p];
        c.primary = compile(pats[p].primary_regex);
        for (uint32_t s = 0; s < pats[p].n_secondaries; ++s) c.sec.push_back(compile(pats[p].secondaries[s].regex));
        for (uint32_t s = 0; s < pats[p].n_sequences; ++s) {
            std::vector<Rx*> ev;
            for (uint32_t k = 0; k < pats[p].sequences[s].n_events; ++k) ev.push_back(compile(pats[p].sequences[s].events[k]));
            c.seq.push_back(std::move(ev));
        }
        if (blank_id(pats[p].id)) c.freq_slot = -1;
        else {
            auto it = slots.find(pats[p].id);
            if (it == slots.end()) { c.freq_slot = (int)slot_first.size(); slots[pats[p].id] = c.freq_slot; slot_first.push_back((int)p); }
            else c.freq_slot = it->second;
        }
    }
    std::vector<double> slot_count(slot_first.size(), 0.0);
    std::vector<char> slot_exists(slot_first.size(), 0);
    for (size_t s = 0; s < slot_first.size(); ++s) {
        uint64_t pc = prior ? prior[slot_first[s]] : 0;
        slot_count[s] = (double)pc;
        slot_exists[s] = pc > 0;
    }

    CtxRx cr;
    CtxMemo cm;
    if (memo) cm.v.assign(N, -1);
    Env env{cfg, &L, memo};
    out.delta.assign(n_pats, 0);

    auto line_flags = [&](uint64_t ln) -> int {
        if (memo) {
            signed char& c = cm.v[ln];
            if (c >= 0) return c;
            c = (signed char)ctx_flags_cp(cr, L.cps.data() + L.cp_off[ln], (int)(L.cp_off[ln + 1] - L.cp_off[ln]));
            return c;
        }
        return ctx_flags_cp(cr, L.cps.data() + L.cp_off[ln], (int)(L.cp_off[ln + 1] - L.cp_off[ln]));
    };

    // AS:89-113
    for (uint64_t i = 0; i < N; ++i) {
        for (size_t p = 0; p < n_pats; ++p) {
            CPat& c = cps[p];
            if (!env.find(*c.primary, i)) continue;
            // AS:132-156 context window
            int64_t b0, a1;
            if (c.p->has_context) {
                b0 = std::max<int64_t>(0, (int64_t)i - c.p->lines_before);
                a1 = std::min<int64_t>((int64_t)N, (int64_t)i + 1 + c.p->lines_after);
            } else { b0 = (int64_t)i; a1 = (int64_t)i + 1; }
            // SS:63-112
            double base = c.p->confidence;
            double sev = c.p->severity_multiplier;
            double ch = chrono(*cfg, i, N);
            double px = proximity(env, c, (int64_t)i, (int64_t)N);
            double tp = temporal(env, c, (int64_t)i, (int64_t)N);
            double cf = context_score(*cfg, (size_t)(a1 - b0), [&](size_t k) { return line_flags((uint64_t)b0 + k); });
            double pen = 0.0;
            if (c.freq_slot >= 0) {
                pen = freq_penalty(*cfg, slot_exists[c.freq_slot], slot_count[c.freq_slot]);
                slot_exists[c.freq_slot] = 1;      // FTS:46-48 computeIfAbsent
                slot_count[c.freq_slot] += 1.0;    // FTS:50
            }
            double score = base * sev * ch * px * tp * cf * (1.0 - pen);
            out.line.push_back((uint32_t)(i + 1));
            out.pattern.push_back((uint32_t)p);
            out.score.push_back(score);
            double fs[7] = {base, sev, ch, px, tp, cf, pen};
            out.factors.insert(out.factors.end(), fs, fs + 7);
            out.delta[p] += 1;
        }
    }
    out.r.n_lines = N;
    out.r.n_events = out.line.size();
    out.r.n_patterns = n_pats;
    out.r.line = out.line.data();
    out.r.pattern = out.pattern.data();
    out.r.score = out.score.data();
    out.r.factors = out.factors.data();
    out.r.freq_delta = out.delta.data();
}

void set_err(char* err, size_t errlen, const std::string& m) {
    if (!err || errlen == 0) return;
    size_t n = std::min(errlen - 1, m.size());
    memcpy(err, m.data(), n);
    err[n] = 0;
}

}  // namespace

extern "C" {

void* orc_regex_compile(const char* pattern, int flags, char* err, size_t errlen, int* err_kind) {
    if (err_kind) *err_kind = 0;
    try {
        return new Regex(pattern, flags);
    } catch (const orx::SyntaxError& e) {
        if (err_kind) *err_kind = 1;
        set_err(err, errlen, std::string(e.what()) + " near index " + std::to_string(e.index));
    } catch (const orx::Unsupported& e) {
        if (err_kind) *err_kind = 2;
        set_err(err, errlen, std::string("unsupported: ") + e.what());
    } catch (const std::exception& e) {
        if (err_kind) *err_kind = 3;
        set_err(err, errlen, e.what());
    }
    return nullptr;
}
void orc_regex_free(void* h) { delete static_cast<Regex*>(h); }
int orc_regex_find(void* h, const uint8_t* s, size_t n) { return static_cast<Regex*>(h)->find(s, n) ? 1 : 0; }
int orc_regex_find_span(void* h, const uint8_t* s, size_t n, int* start, int* end) {
    return static_cast<Regex*>(h)->find_span(s, n, start, end) ? 1 : 0;
}
void orc_regex_set_first_opt(void* h, int on) { static_cast<Regex*>(h)->set_first_opt(on != 0); }

uint64_t orc_split(const uint8_t* log, size_t len, uint64_t* starts, uint64_t* ends, uint64_t cap) {
    std::vector<uint64_t> st, en;
    java_split(log, len, st, en);
    for (uint64_t i = 0; i < st.size() && i < cap; ++i) { starts[i] = st[i]; ends[i] = en[i]; }
    return st.size();
}
double orc_chrono(const orc_cfg* cfg, uint64_t idx, uint64_t n) { return chrono(*cfg, idx, n); }
double orc_context_factor(const orc_cfg* cfg, const uint8_t* const* lines, const size_t* lens, size_t n) {
    CtxRx cr;
    std::vector<int> fl(n);
    std::vector<int> cps;
    for (size_t i = 0; i < n; ++i) {
        orx::decode_utf8(lines[i], lens[i], cps);
        fl[i] = ctx_flags_cp(cr, cps.data(), (int)cps.size());
    }
    return context_score(*cfg, n, [&](size_t k) { return fl[k]; });
}
double orc_frequency_penalty(const orc_cfg* cfg, int exists, double count) { return freq_penalty(*cfg, exists != 0, count); }
double orc_proximity_term(const orc_cfg* cfg, double weight, double distance) {
    return weight * std::exp(-distance / cfg->decay_constant);
}

int orc_analyze(const orc_cfg* cfg, const orc_pattern* pats, size_t n_pats, const uint8_t* log, size_t len,
                const uint64_t* prior_count, int options, orc_result** out, char* err, size_t errlen) {
    try {
        Result* r = new Result();
        try {
            do_analyze(cfg, pats, n_pats, log, len, prior_count, options, *r);
        } catch (...) { delete r; throw; }
        *out = &r->r;
        return 0;
    } catch (const orx::SyntaxError& e) {
        set_err(err, errlen, std::string("PatternSyntaxException: ") + e.what() + " near index " + std::to_string(e.index));
        return -2;
    } catch (const orx::Unsupported& e) {
        set_err(err, errlen, std::string("unsupported: ") + e.what());
        return -3;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return -1;
    }
}
void orc_result_free(orc_result* r) {
    if (!r) return;
    delete reinterpret_cast<Result*>(reinterpret_cast<char*>(r) - offsetof(Result, r));
}

int64_t orc_analyze_many(const orc_cfg* cfg, const orc_pattern* pats, size_t n_pats, const uint8_t* const* logs,
                         const size_t* lens, size_t n_logs, int threads, int options, uint64_t* total_lines,
                         double* checksum, char* err, size_t errlen) {
    if (threads < 1) threads = 1;
    std::atomic<size_t> next{0};
    std::atomic<int64_t> events{0};
    std::atomic<uint64_t> lines{0};
    std::vector<double> sums((size_t)threads, 0.0);
    std::atomic<int> failed{0};
    std::string first_err;
    auto work = [&](int tid) {
        for (;;) {
            size_t i = next.fetch_add(1);
            if (i >= n_logs) break;
            try {
                Result r;
                do_analyze(cfg, pats, n_pats, logs[i], lens[i], nullptr, options, r);
                events += (int64_t)r.line.size();
                lines += r.r.n_lines;
                for (double s : r.score) sums[tid] += s;
            } catch (const std::exception& e) {
                if (!failed.exchange(1)) first_err = e.what();
            }
        }
    };
    std::vector<std::thread> th;
    // big stacks: the matcher recurses per repetition on non-trivial loop bodies
    for (int t = 1; t < threads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    if (failed) { set_err(err, errlen, first_err); return -1; }
    if (total_lines) *total_lines = lines.load();
    if (checksum) { double s = 0; for (double x : sums) s += x; *checksum = s; }
    return events.load();
}

}  // extern "C"
