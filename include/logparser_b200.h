/* liblogparser_b200.so — C ABI of the B200-native hot path of podmortem/log-parser.
 *
 * The reference (Java 21 / Quarkus) has no FFI today; this is the seam SURVEY.md §8b cuts into
 * AnalysisService.analyze(): everything between `data.getLogs().split("\\r?\\n")`
 * (AnalysisService.java:53) and the end of the N x P scan/score loop (AnalysisService.java:113)
 * runs behind these entry points. The JVM keeps Parse.java, PatternService.java, the
 * scoring.* configuration, FrequencyTrackingService's state/clock and the JSON assembly.
 * Java binds these symbols with Panama FFM; no JNI glue, no callbacks.
 *
 * Plain pointers and sizes only. All strings are NUL-terminated UTF-8. The library copies what
 * it needs from every input before returning. There is NO CPU fallback: every entry point that
 * computes fails with LP_ERR_CUDA when no CUDA device is usable.
 */
#ifndef LOGPARSER_B200_H
#define LOGPARSER_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LP_OK 0
#define LP_ERR_INVALID (-1)       /* bad argument                                              */
#define LP_ERR_SYNTAX (-2)        /* java.util.regex.PatternSyntaxException (AS:64,70,80)      */
#define LP_ERR_UNSUPPORTED (-3)   /* legal Java regex syntax this build does not implement     */
#define LP_ERR_CUDA (-4)          /* no device / CUDA runtime failure                          */
#define LP_ERR_LIMIT (-5)         /* input exceeds a documented limit                          */
#define LP_ERR_BACKTRACK (-6)     /* backtracking verifier ran out of its step/stack budget    */

typedef struct lp_ctx lp_ctx;
typedef struct lp_result lp_result;

/* The nine scoring.* keys (src/main/resources/application.properties:5-20; defaults
 * ScoringService.java:38-51, ContextAnalysisService.java:24, FrequencyTrackingService.java:27-34). */
typedef struct lp_scoring_cfg {
    double decay_constant;         /* scoring.proximity.decay-constant            10.0 */
    int32_t max_window;            /* scoring.proximity.max-window                100  */
    int32_t time_window_hours;     /* scoring.frequency.time-window-hours         1    */
    double early_bonus_threshold;  /* scoring.chronological.early-bonus-threshold 0.2  */
    double max_early_bonus;        /* scoring.chronological.max-early-bonus       2.5  */
    double penalty_threshold;      /* scoring.chronological.penalty-threshold     0.5  */
    double max_context_factor;     /* scoring.context.max-context-factor          2.5  */
    double frequency_threshold;    /* scoring.frequency.threshold                 10.0 */
    double max_penalty;            /* scoring.frequency.max-penalty               0.8  */
    /* Optional exp(-d/decay_constant) for d = 0..max_window computed by the caller (Math.exp on
     * the JVM) so proximity terms are bit-identical to ScoringService.java:177; NULL = libm exp. */
    const double* decay_lut;
} lp_scoring_cfg;

typedef struct lp_secondary_desc {      /* SecondaryPattern: SS:172-178, 319 */
    const char* regex;
    double weight;
    int32_t proximity_window;
} lp_secondary_desc;

typedef struct lp_sequence_desc {       /* SequencePattern / SequenceEvent: SS:207-258 */
    const char* const* event_regex;
    uint32_t n_events;
    double bonus_multiplier;
} lp_sequence_desc;

/* One Pattern of a PatternSet, flattened in (set, pattern) load order (AS:91-92). */
typedef struct lp_pattern_desc {
    const char* id;                 /* NULL/blank: not frequency-tracked (FTS:42-44,65-67)        */
    const char* primary_regex;      /* primaryPattern.regex (AS:62-65)                            */
    double confidence;              /* primaryPattern.confidence (SS:65)                          */
    double severity_multiplier;     /* SEVERITY_MULTIPLIERS lookup done by the caller (SS:30-36,68-69) */
    int32_t has_context;            /* 0: contextExtraction == null (AS:137-139)                  */
    int32_t lines_before;           /* contextExtraction.linesBefore (AS:142)                     */
    int32_t lines_after;            /* contextExtraction.linesAfter  (AS:148)                     */
    uint32_t n_secondaries;
    const lp_secondary_desc* secondaries;
    uint32_t n_sequences;
    const lp_sequence_desc* sequences;
} lp_pattern_desc;

/* Column views of one analysis result; pointers stay valid until lp_result_free(). */
typedef struct lp_result_view {
    uint64_t n_logs;
    uint64_t n_events;              /* all logs, grouped by log in input order                    */
    const uint64_t* log_n_lines;    /* [n_logs]  logLines.length per log (AS:53,170)              */
    const uint64_t* log_event_begin;/* [n_logs+1] events of log k are [begin[k], begin[k+1])      */
    const uint32_t* line;           /* [n_events] 1-based line number within its log (AS:101)     */
    const uint32_t* pattern;        /* [n_events] flat pattern index (AS:102)                     */
    const double* score;            /* [n_events] MatchedEvent.score (AS:106-107)                 */
    const uint32_t* rank;           /* [n_events] position when sorted by score desc, ties by     */
                                    /*   discovery order, within the event's log (extension)      */
    const double* factors;          /* [n_events*7] base,sev,chrono,prox,temporal,ctx,penalty; NULL unless LP_OPT_FACTORS */
    const uint64_t* freq_delta;     /* [n_patterns] events per pattern in this call (FTS:41-56)   */
    uint64_t n_patterns;
    const uint32_t* line_start;     /* [total_lines+1] byte offsets, NULL unless LP_OPT_LINE_OFFSETS */
    const uint32_t* line_end;       /* [total_lines]                                              */
    uint64_t total_lines;
    /* LP_OPT_SPANS (extension; the reference uses find() as a boolean, AS:95): Matcher.start() and
     * Matcher.end() of the first find() of the event's primary regex on its line, as UTF-8 byte
     * offsets from the start of the line (equal to Java's UTF-16 offsets on ASCII lines). */
    const uint32_t* span_start;     /* [n_events], NULL unless LP_OPT_SPANS                       */
    const uint32_t* span_end;       /* [n_events]                                                 */
    /* (line, regex) evaluations that the backtracking verifier abandoned after LP_VM_MAX_STEPS (2^26) steps or on
     * stack exhaustion; such a pair counts as "no match". java.util.regex would keep backtracking (or throw
     * StackOverflowError). 0 for every pattern set that stays on the automaton path. See LP_OPT_STRICT_BUDGET. */
    uint64_t budget_exceeded;
} lp_result_view;

enum {
    LP_OPT_FACTORS = 1,        /* also return the seven factors per event                        */
    LP_OPT_LINE_OFFSETS = 2,   /* also return line_start/line_end for context slicing (AS:132-156) */
    LP_OPT_NO_RANK = 4,        /* skip the ranking sort                                          */
    LP_OPT_SPANS = 8,          /* also return the match span of every event (backtracking engine) */
    LP_OPT_STRICT_BUDGET = 16  /* fail the call with LP_ERR_BACKTRACK instead of reporting budget_exceeded > 0 */
};

/* ---- host buffers --------------------------------------------------------------------------
 * No counterpart in the reference (its request body is a Java String, Parse.java:44). A log handed to
 * lp_analyze* in pageable memory is copied through a driver staging buffer while the calling thread
 * waits; from page-locked memory the copy runs asynchronously, in pieces, under the scan. The JVM
 * either takes its request buffer from lp_host_alloc (MemorySegment.reinterpret over the returned
 * address, lp_host_free when the buffer pool shrinks) or page-locks a segment it owns with
 * lp_host_register / lp_host_unregister. Both need a usable CUDA device. */
void* lp_host_alloc(size_t bytes);              /* NULL on failure, lp_last_error() says why */
void lp_host_free(void* p);
int lp_host_register(void* p, size_t bytes);
int lp_host_unregister(void* p);

/* ---- lifecycle ------------------------------------------------------------------------- */
/* Carries the scoring.* values into the library and binds it to one CUDA device
 * (replaces the @ConfigProperty injection of SS:38-51, CAS:24, FTS:27-34). NULL on failure.
 * device = -1: all visible devices, see lp_create_multi. */
lp_ctx* lp_create(const lp_scoring_cfg* cfg, int device);
/* One process driving several GPUs (the reference is ONE JVM): `lp_create(cfg, -1)` takes every visible device,
 * lp_create_multi the listed ones (a device may be listed more than once). lp_load_patterns then loads every device,
 * and lp_analyze_batch cuts a batch of two or more logs into contiguous blocks of logs, one per device, balanced by
 * bytes: each block is copied and scanned on a host thread of its own, the per-pattern event counts of the blocks meet
 * on the host (P x 8 bytes per device; FTS:41-93 is the only cross-request state), each block is scored with the
 * counts of the blocks in front of it, and ONE result comes back in batch order: the same events, scores, per-log
 * ranks and counts as on one device (LP_OPT_LINE_OFFSETS is not available there). Every other call runs on the first
 * device. One process per GPU with lp_comm_init + lp_analyze_batch_dist is the other way to use several GPUs. */
lp_ctx* lp_create_multi(const lp_scoring_cfg* cfg, const int* devices, size_t n_devices);
void lp_destroy(lp_ctx* ctx);
/* Thread-local message for the last failing call on this thread. */
const char* lp_last_error(void);
const char* lp_version(void);

/* Replaces the per-request compile block AS:56-86 (and the static context regexes CAS:27-34):
 * compiles every primary / secondary / sequence-event regex once and uploads the device
 * tables. On LP_ERR_SYNTAX / LP_ERR_UNSUPPORTED, *bad_pattern (if not NULL) receives the flat
 * index of the offending pattern. Not concurrent with lp_analyze* on the same ctx. */
int lp_load_patterns(lp_ctx* ctx, const lp_pattern_desc* patterns, size_t n_patterns, int64_t* bad_pattern);

/* ---- the hot path ------------------------------------------------------------------------ */
/* Replaces AS:53-113 for one request: split, scan, score. `log` is host memory (UTF-8 of
 * PodFailureData.logs, i.e. what String.getBytes(UTF_8) produces: well-formed. The reference matches on the decoded
 * String, where malformed input bytes have already become U+FFFD; here a malformed sequence matches no character
 * class, so a caller holding raw bytes replaces malformed sequences by EF BF BD first). prior_count[p] = matches of pattern p already inside the sliding
 * frequency window at request start (FTS state stays on the JVM); NULL = all zero.
 * FrequencyTrackingService keys its counts by pattern id (FTS:46-50): patterns that carry the same id
 * share one count. Pass the id's count in the entry of every pattern with that id (the library reads
 * the first one); freq_delta[] comes back per pattern, to be summed per id by the caller. */
int lp_analyze(lp_ctx* ctx, const uint8_t* log, size_t len, const uint64_t* prior_count,
               int options, lp_result** out);

/* Same, with the log already resident in device memory (`d_log` is a device pointer). */
int lp_analyze_device(lp_ctx* ctx, const uint8_t* d_log, size_t len, const uint64_t* prior_count,
                      int options, lp_result** out);

/* A batch of independent requests, processed as the reference would process them one after the
 * other in input order (frequency state carried from log k to log k+1). `logs` is the plain
 * concatenation of the logs; log k is bytes [offsets[k], offsets[k+1]). Total size < 4 GiB. */
int lp_analyze_batch(lp_ctx* ctx, const uint8_t* logs, const uint64_t* offsets, size_t n_logs,
                     const uint64_t* prior_count, int options, lp_result** out);
int lp_analyze_batch_device(lp_ctx* ctx, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs,
                            const uint64_t* prior_count, int options, lp_result** out);

/* Two-phase form for sharding a batch over several GPUs (one process per GPU): phase 1 scans
 * this rank's logs and returns its per-pattern event counts; the host exchanges counts between
 * ranks (NCCL all-gather) and passes the exclusive prefix over lower ranks (+ the window's
 * prior counts) to phase 2, which scores. lp_analyze_batch* == scan followed by score. */
int lp_batch_scan_device(lp_ctx* ctx, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs,
                         int options, uint64_t* counts_out /* [n_patterns] */);
int lp_batch_score(lp_ctx* ctx, const uint64_t* prior_count, lp_result** out);
/* For hosts that add the lower ranks' counts themselves: out[p] = sum of in[q] over the patterns q that carry
 * pattern p's id (FTS:46-50 keys the counts by id; 0 for untracked patterns). Pure host code. */
int lp_counts_by_id(const lp_ctx* ctx, const uint64_t* in /* [n_patterns] */, uint64_t* out /* [n_patterns] */);

/* The same two phases with the exchange kept on the device (no host round trip between them):
 * phase 1 only enqueues work on lp_stream() and writes [n_patterns counts, 1 status word] to
 * DEVICE memory `d_counts_out`; the host all-gathers those n_patterns + 1 words of every rank into
 * `d_all_counts` ([world][n_patterns + 1], rank order, device memory, ordered after lp_stream() — e.g.
 * ncclAllGather on that stream); phase 2 forms the exclusive prefix on the device, scores and
 * synchronises once. Returns LP_RETRY (not an error) when some rank's scratch estimates were too
 * small for its logs: every rank sees the same status words, so every rank repeats both phases. */
#define LP_RETRY 1
int lp_batch_scan_device_async(lp_ctx* ctx, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs,
                               int options, uint64_t* d_counts_out /* device, [n_patterns + 1] */);
int lp_batch_score_gathered(lp_ctx* ctx, const uint64_t* d_all_counts, uint32_t rank, uint32_t world,
                            const uint64_t* prior_count, lp_result** out);

/* The same batch step as ONE collective call: scan, ncclAllGather of the n_patterns + 1 words on
 * lp_stream() over NVLink, prefix and score, with the repeat-on-LP_RETRY loop inside. NCCL
 * (libnccl.so.2) is loaded at run time on the first lp_comm_* call; hosts that never call them need
 * no NCCL. Set-up: rank 0 calls lp_comm_unique_id() (128 bytes, ncclUniqueId) and hands the bytes to
 * the other ranks by whatever means the service has; every rank then calls lp_comm_init(). One
 * process per GPU; every rank calls lp_analyze_batch_dist_device once per step, in the same order.
 * The exchange is started as soon as this rank's match records are final and overlaps the rest of the
 * scan phase. As with any collective, an error return on one rank leaves the others waiting in their
 * next call: treat it as fatal for the communicator (lp_destroy on every rank, then start over). */
int lp_comm_unique_id(uint8_t* id128);
int lp_comm_init(lp_ctx* ctx, const uint8_t* id128, uint32_t rank, uint32_t world);
int lp_analyze_batch_dist_device(lp_ctx* ctx, const uint8_t* d_logs, const uint64_t* offsets, size_t n_logs,
                                 const uint64_t* prior_count, int options, lp_result** out);
/* The same collective step with this rank's logs still in HOST memory (page-locked for real overlap, see
 * lp_host_alloc): the host->device copy is cut in pieces on a second stream and each piece is scanned as it
 * lands, exactly as lp_analyze does for one request (Parse.java:44 hands the service a host String). */
int lp_analyze_batch_dist(lp_ctx* ctx, const uint8_t* logs, const uint64_t* offsets, size_t n_logs,
                          const uint64_t* prior_count, int options, lp_result** out);

/* ONE log sharded by line range over several GPUs (second mode; the reference walks
 * the whole log on one thread, AS:89-113). The host first removes the log's trailing separators
 * (the trailing empty strings String.split drops, AS:53), cuts the rest on line boundaries into
 * `world` contiguous pieces and gives every rank its piece plus up to lp_range_halo_lines() whole
 * lines on either side, WITHOUT the separator that ends the text (so the text ends with the content
 * of its last line). Halo lines feed only the windows of the rank's own lines:
 *   proximity  min(max-window, proximity_window) lines either side     SS:319-328
 *   sequence   5 lines either side for the last event                  SS:275-277
 *   context    lines_before / lines_after                              AS:142-150
 * Quantities that need the whole log travel in a per-rank summary of lp_range_summary_words()
 * uint64 words, exchanged by the host with one all-gather:
 *   [0] own line count -> global line numbers and the N of the chronological factor (SS:125)
 *   [1] n_patterns, [2..] per-pattern event counts -> exclusive prefix for FTS:64-93
 *   then, per sequence and per state its backward walk (SS:250-258, in forward form) can enter the
 *   shard in: the state on exit and the own line where the last-but-one event was reached.
 * lp_range_score composes them (ranks in text order) and scores this rank's lines; result line
 * numbers are global, `log_n_lines[0]` is the whole log's N, `rank[]` orders this shard's events
 * only (merge by score across shards on the host), freq_delta covers this shard's lines. */
uint32_t lp_range_halo_lines(const lp_ctx* ctx);
size_t lp_range_summary_words(const lp_ctx* ctx);
int lp_range_scan_device(lp_ctx* ctx, const uint8_t* d_text, size_t len, uint32_t halo_before_lines,
                         uint32_t halo_after_lines, int options, uint64_t* summary_out);
int lp_range_score(lp_ctx* ctx, const uint64_t* summaries /* [world][words] */, uint32_t rank, uint32_t world,
                   const uint64_t* prior_count, lp_result** out);
/* The same with the exchange kept on the device, over the communicator of lp_comm_init (one process per GPU; every
 * rank calls it once per step, ranks in text order): scan of this rank's piece, all-gather of the summaries over
 * NVLink, composition on the device (prior counts, global line numbers, sequence thresholds), scoring. Nothing
 * returns to the host between the scan and the events. Results as lp_range_scan_device + lp_range_score.
 * lp_analyze_range_dist takes the piece in host memory (page-locked memory from lp_host_alloc lets the copy overlap
 * the scan); `prior_count` is the window prior BEFORE the whole log, the same array on every rank. */
int lp_analyze_range_dist_device(lp_ctx* ctx, const uint8_t* d_text, size_t len, uint32_t halo_before_lines,
                                 uint32_t halo_after_lines, const uint64_t* prior_count, int options, lp_result** out);
int lp_analyze_range_dist(lp_ctx* ctx, const uint8_t* text, size_t len, uint32_t halo_before_lines,
                          uint32_t halo_after_lines, const uint64_t* prior_count, int options, lp_result** out);

/* No counterpart in the reference. Line splitting (the String.split step only) for hosts that may
 * not allocate GPU memory themselves or let a library do it on their behalf: processes with a
 * fixed device-memory plan, plugins living inside somebody else's allocator, callers for whom an
 * allocation-induced device synchronisation is unacceptable. Every other lp_* call grows a
 * scratch pool with cudaMalloc; this one allocates nothing, on the device or through any library:
 * its scratch is a fixed array inside the loaded module (about 128 KiB per GPU, held until the
 * process exits), which is why the input is limited to LP_NOALLOC_MAX_BYTES per call
 * (LP_ERR_LIMIT beyond that, with the reason in `err`).
 *   d_log         device pointer, 16-byte aligned, `len` bytes of UTF-8
 *   d_line_start  device, `capacity` entries: start offset of every line segment
 *   d_line_end    device, `capacity` entries: end offset (exclusive; a '\r' before '\n' is dropped)
 *   d_counts      device, 3 entries: [0] lines after Java's trailing-empty trimming (logLines.length),
 *                 [1] line segments written (newlines + 1), [2] non-zero if `capacity` was too small
 *   stream        cudaStream_t to run on (NULL = default stream). The call is synchronous: it reads the
 *                 overflow flag back and returns only after `stream` has drained; the calling thread's
 *                 current device is the same after the call as before
 *   err/errlen    optional buffer for a message when the return value is not LP_OK
 * The first call on each device looks up the scratch array's address, once. All calls on one device
 * share that array and it keeps the previous call's
 * contents (the call clears what it needs): calls that can overlap on the GPU corrupt each other,
 * so keep them on one stream or otherwise serialised. */
#define LP_NOALLOC_MAX_BYTES (1ull << 30)
int lp_split_lines_noalloc(int device, const uint8_t* d_log, size_t len, uint32_t* d_line_start, uint32_t* d_line_end,
                           size_t capacity, uint64_t* d_counts, void* stream, char* err, size_t errlen);

int lp_result_get(const lp_result* r, lp_result_view* view);
void lp_result_free(lp_result* r);

/* ---- measurement ------------------------------------------------------------------------- */
typedef struct lp_timings {
    float total_ms;        /* first to last kernel of the last lp_analyze* call (CUDA events)     */
    float scan_ms;         /* k_fscan: the one pass over the log (newline lists + literal prefilter), the dominant streaming kernel */
    float verify_ms;       /* line offsets (lines_ms) + candidate confirmation + DFA / backtracking verification */
    float score_ms;        /* sorting, event assembly, scoring, ranking                           */
    float h2d_ms;          /* host->device copy of the log (lp_analyze / lp_analyze_batch only)   */
    uint64_t kernel_launches;   /* kernels of this library launched by the last call             */
    uint64_t scan_bytes;        /* its algorithmic bytes: B read + 2 per newline + 4 per tile + 4 per candidate */
    uint64_t candidates;        /* gram-filter hits handed to confirmation                        */
    uint64_t verify_tasks;      /* (line, regex) pairs verified                                   */
    float dfa_scan_ms;          /* the always-on shared-memory product-DFA scan (part of verify_ms) */
    float lines_ms;             /* tile-count scan + line_start/line_end from the newline lists + Java trimming (part of verify_ms) */
    uint64_t dfa_scan_bytes;    /* its algorithmic bytes: B + 8(G+1) + G                          */
} lp_timings;
int lp_get_timings(const lp_ctx* ctx, lp_timings* out);
/* The cudaStream_t all kernels and copies of this context are issued on (for callers that order
 * their own work or CUDA events against it). */
void* lp_stream(const lp_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* LOGPARSER_B200_H */
