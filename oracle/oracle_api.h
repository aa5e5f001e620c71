/* ORACLE — TEST INFRASTRUCTURE ONLY. Not shipped, not linked by the product library.
 *
 * C interface of the CPU restatement of the reference hot path
 * (AnalysisService.java:50-156, ScoringService.java:63-347, ContextAnalysisService.java:46-144,
 * FrequencyTrackingService.java:41-93). Used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs as the checker / timed CPU baseline only.
 * Parity status: unpinned by the reference (no tests, no golden vectors upstream); pinned here
 * by SURVEY.md §8c hand-derived vectors and differential regex tests (see jregex.hpp).
 */
#ifndef ORACLE_API_H
#define ORACLE_API_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_cfg {
    double decay_constant;        /* scoring.proximity.decay-constant        SS:38 */
    int32_t max_window;           /* scoring.proximity.max-window            SS:41 */
    int32_t time_window_hours;    /* scoring.frequency.time-window-hours     FTS:33 */
    double early_bonus_threshold; /* scoring.chronological.early-bonus-threshold SS:44 */
    double max_early_bonus;       /* scoring.chronological.max-early-bonus   SS:47 */
    double penalty_threshold;     /* scoring.chronological.penalty-threshold SS:50 */
    double max_context_factor;    /* scoring.context.max-context-factor      CAS:24 */
    double frequency_threshold;   /* scoring.frequency.threshold             FTS:27 */
    double max_penalty;           /* scoring.frequency.max-penalty           FTS:30 */
} orc_cfg;

typedef struct orc_secondary {
    const char* regex;
    double weight;
    int32_t proximity_window;
} orc_secondary;

typedef struct orc_sequence {
    const char* const* events;
    uint32_t n_events;
    double bonus_multiplier;
} orc_sequence;

typedef struct orc_pattern {
    const char* id;               /* NULL / blank: not frequency-tracked (FTS:42,65) */
    const char* primary_regex;
    double confidence;
    double severity_multiplier;   /* resolved by the caller per SS:30-36,68-69 */
    int32_t has_context;          /* 0: contextExtraction == null (AS:137-139) */
    int32_t lines_before;
    int32_t lines_after;
    uint32_t n_secondaries;
    const orc_secondary* secondaries;
    uint32_t n_sequences;
    const orc_sequence* sequences;
} orc_pattern;

typedef struct orc_result {
    uint64_t n_lines;      /* Java split length */
    uint64_t n_events;
    uint32_t* line;        /* 1-based line numbers, discovery order (AS:101) */
    uint32_t* pattern;     /* flat pattern index */
    double* score;
    double* factors;       /* n_events x 7: base, sev, chrono, prox, temporal, ctx, penalty */
    uint64_t* freq_delta;  /* per pattern index: events of that pattern in this request */
    uint64_t n_patterns;
} orc_result;

enum { ORC_OPT_MEMO = 1, ORC_OPT_NO_FIRST_OPT = 2 };

/* ---- regex ---- */
void* orc_regex_compile(const char* pattern, int flags, char* err, size_t errlen, int* err_kind);
void orc_regex_free(void* h);
int orc_regex_find(void* h, const uint8_t* s, size_t n);
int orc_regex_find_span(void* h, const uint8_t* s, size_t n, int* start, int* end);
void orc_regex_set_first_opt(void* h, int on);

/* ---- pieces (for golden vectors) ---- */
/* String.split("\\r?\\n") AS:53; writes up to cap [start,end) byte offsets; returns N */
uint64_t orc_split(const uint8_t* log, size_t len, uint64_t* starts, uint64_t* ends, uint64_t cap);
double orc_chrono(const orc_cfg* cfg, uint64_t line_index0, uint64_t n_lines);        /* SS:123-151 */
double orc_context_factor(const orc_cfg* cfg, const uint8_t* const* lines, const size_t* lens, size_t n); /* CAS:46-117 */
double orc_frequency_penalty(const orc_cfg* cfg, int exists, double count);          /* FTS:64-93 */
double orc_proximity_term(const orc_cfg* cfg, double weight, double distance);       /* SS:177-178 */

/* ---- whole path ---- */
int orc_analyze(const orc_cfg* cfg, const orc_pattern* pats, size_t n_pats,
                const uint8_t* log, size_t len, const uint64_t* prior_count, int options,
                orc_result** out, char* err, size_t errlen);
void orc_result_free(orc_result* r);

/* Timing helper: n_logs independent requests spread over `threads` host threads
 * (the reference's only parallelism is one request per worker thread, AS:89-113).
 * Returns total events; frequency state is per request (prior = 0). */
int64_t orc_analyze_many(const orc_cfg* cfg, const orc_pattern* pats, size_t n_pats,
                         const uint8_t* const* logs, const size_t* lens, size_t n_logs,
                         int threads, int options, uint64_t* total_lines, double* checksum,
                         char* err, size_t errlen);

#ifdef __cplusplus
}
#endif
#endif
