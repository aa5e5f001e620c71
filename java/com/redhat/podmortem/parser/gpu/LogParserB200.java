package com.redhat.podmortem.parser.gpu;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import com.redhat.podmortem.common.model.pattern.Pattern;
import com.redhat.podmortem.common.model.pattern.SecondaryPattern;
import com.redhat.podmortem.common.model.pattern.SequenceEvent;
import com.redhat.podmortem.common.model.pattern.SequencePattern;
import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemoryLayout;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.StructLayout;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;
import java.nio.charset.StandardCharsets;
import java.util.List;
import java.util.Map;
import java.util.regex.PatternSyntaxException;

/**
 * Panama FFM binding of liblogparser_b200.so (include/logparser_b200.h): the B200 path behind
 * {@code AnalysisService.analyze} (AnalysisService.java:53-113 of the reference). No JNI glue, no upcalls.
 *
 * <p>Java 21: {@code java.lang.foreign} is a preview API (JEP 442) - compile and run with
 * {@code --enable-preview}; final in 22 (JEP 454). JVM mode only. This file was written without a JDK at
 * hand (none exists in the authoring image); the identical call sequence, struct layouts and ownership
 * rules are exercised through ctypes by {@code log_parser_b200/binding.py} and the test-suite.
 *
 * <p>One instance per GPU. Calls on one instance are serialised by the library; requests are independent
 * of each other except for the frequency counts, which stay on the JVM ({@link FrequencyBridge}).
 */
public final class LogParserB200 implements AutoCloseable {

    public static final int OPT_FACTORS = 1, OPT_LINE_OFFSETS = 2, OPT_NO_RANK = 4, OPT_SPANS = 8, OPT_STRICT_BUDGET = 16;
    private static final int LP_ERR_SYNTAX = -2;

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB =
            SymbolLookup.libraryLookup(System.getProperty("logparser.b200.lib", "liblogparser_b200.so"), Arena.global());

    private static MethodHandle h(String name, FunctionDescriptor d) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
    }

    // lp_scoring_cfg: double, int, int, 6 doubles, pointer (72 bytes)
    static final StructLayout CFG = MemoryLayout.structLayout(
            JAVA_DOUBLE.withName("decay_constant"), JAVA_INT.withName("max_window"), JAVA_INT.withName("time_window_hours"),
            JAVA_DOUBLE.withName("early_bonus_threshold"), JAVA_DOUBLE.withName("max_early_bonus"),
            JAVA_DOUBLE.withName("penalty_threshold"), JAVA_DOUBLE.withName("max_context_factor"),
            JAVA_DOUBLE.withName("frequency_threshold"), JAVA_DOUBLE.withName("max_penalty"), ADDRESS.withName("decay_lut"));
    // lp_secondary_desc {const char* regex; double weight; int32 proximity_window;} (24 bytes)
    static final StructLayout SECONDARY = MemoryLayout.structLayout(ADDRESS, JAVA_DOUBLE, JAVA_INT, MemoryLayout.paddingLayout(4));
    // lp_sequence_desc {const char* const* event_regex; uint32 n_events; double bonus_multiplier;} (24 bytes)
    static final StructLayout SEQUENCE = MemoryLayout.structLayout(ADDRESS, JAVA_INT, MemoryLayout.paddingLayout(4), JAVA_DOUBLE);
    // lp_pattern_desc (72 bytes)
    static final StructLayout PATTERN = MemoryLayout.structLayout(
            ADDRESS.withName("id"), ADDRESS.withName("primary_regex"), JAVA_DOUBLE.withName("confidence"),
            JAVA_DOUBLE.withName("severity_multiplier"), JAVA_INT.withName("has_context"), JAVA_INT.withName("lines_before"),
            JAVA_INT.withName("lines_after"), JAVA_INT.withName("n_secondaries"), ADDRESS.withName("secondaries"),
            JAVA_INT.withName("n_sequences"), MemoryLayout.paddingLayout(4), ADDRESS.withName("sequences"));
    // lp_result_view (17 fields of 8 bytes)
    static final StructLayout VIEW = MemoryLayout.structLayout(
            JAVA_LONG.withName("n_logs"), JAVA_LONG.withName("n_events"), ADDRESS.withName("log_n_lines"),
            ADDRESS.withName("log_event_begin"), ADDRESS.withName("line"), ADDRESS.withName("pattern"), ADDRESS.withName("score"),
            ADDRESS.withName("rank"), ADDRESS.withName("factors"), ADDRESS.withName("freq_delta"), JAVA_LONG.withName("n_patterns"),
            ADDRESS.withName("line_start"), ADDRESS.withName("line_end"), JAVA_LONG.withName("total_lines"),
            ADDRESS.withName("span_start"), ADDRESS.withName("span_end"), JAVA_LONG.withName("budget_exceeded"));

    private static final MethodHandle CREATE = h("lp_create", FunctionDescriptor.of(ADDRESS, ADDRESS, JAVA_INT));
    private static final MethodHandle DESTROY = h("lp_destroy", FunctionDescriptor.ofVoid(ADDRESS));
    private static final MethodHandle LOAD = h("lp_load_patterns", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle ANALYZE =
            h("lp_analyze", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_INT, ADDRESS));
    private static final MethodHandle GET = h("lp_result_get", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle FREE = h("lp_result_free", FunctionDescriptor.ofVoid(ADDRESS));
    private static final MethodHandle ERR = h("lp_last_error", FunctionDescriptor.of(ADDRESS));
    private static final MethodHandle HOST_ALLOC = h("lp_host_alloc", FunctionDescriptor.of(ADDRESS, JAVA_LONG));
    private static final MethodHandle HOST_FREE = h("lp_host_free", FunctionDescriptor.ofVoid(ADDRESS));

    /** The nine scoring.* properties (application.properties:5-20). */
    public record ScoringProps(double decayConstant, int maxWindow, int timeWindowHours, double earlyBonusThreshold,
                               double maxEarlyBonus, double penaltyThreshold, double maxContextFactor,
                               double frequencyThreshold, double maxPenalty) {}

    /** Columns of one request's result, copied out of the library's page-locked buffer. */
    public record Columns(int totalLines, int[] line, int[] pattern, double[] score, int[] rank, long[] freqDelta, long budgetExceeded) {}

    // ScoringService.java:30-36
    private static final Map<String, Double> SEVERITY_MULTIPLIERS =
            Map.of("CRITICAL", 5.0, "HIGH", 3.0, "MEDIUM", 2.0, "LOW", 1.5, "INFO", 1.0);

    private final MemorySegment ctx;
    /** Page-locked request buffer of one calling thread, grown on demand (lp_host_alloc). */
    private static final class HostBuf {
        MemorySegment seg = MemorySegment.NULL;
        long cap = 0;
    }
    // one buffer per worker thread: lp_analyze calls from different threads run on different workers of the context
    // and overlap on the GPU, so the binding must not serialise them on a shared buffer
    private final java.util.Queue<HostBuf> allBufs = new java.util.concurrent.ConcurrentLinkedQueue<>();
    private final ThreadLocal<HostBuf> hostBufs = ThreadLocal.withInitial(() -> {
        HostBuf b = new HostBuf();
        allBufs.add(b);
        return b;
    });
    private int nPatterns = 0;

    /**
     * @param device CUDA device index, or -1 for every visible device: one context then owns a device context per GPU
     *               and lp_analyze_batch cuts a batch of requests into one block per GPU (single requests run on the first).
     */
    public LogParserB200(ScoringProps p, int device) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment cfg = a.allocate(CFG);
            cfg.set(JAVA_DOUBLE, 0, p.decayConstant());
            cfg.set(JAVA_INT, 8, p.maxWindow());
            cfg.set(JAVA_INT, 12, p.timeWindowHours());
            cfg.set(JAVA_DOUBLE, 16, p.earlyBonusThreshold());
            cfg.set(JAVA_DOUBLE, 24, p.maxEarlyBonus());
            cfg.set(JAVA_DOUBLE, 32, p.penaltyThreshold());
            cfg.set(JAVA_DOUBLE, 40, p.maxContextFactor());
            cfg.set(JAVA_DOUBLE, 48, p.frequencyThreshold());
            cfg.set(JAVA_DOUBLE, 56, p.maxPenalty());
            // Math.exp on the JVM, so that proximity terms are bit-identical to ScoringService.java:177
            MemorySegment lut = a.allocateArray(JAVA_DOUBLE, p.maxWindow() + 1L);
            for (int d = 0; d <= p.maxWindow(); d++) {
                lut.setAtIndex(JAVA_DOUBLE, d, Math.exp(-(double) d / p.decayConstant()));
            }
            cfg.set(ADDRESS, 64, lut);
            MemorySegment c = (MemorySegment) CREATE.invokeExact(cfg, device);
            if (c.equals(MemorySegment.NULL)) {
                throw new IllegalStateException("lp_create: " + lastError());
            }
            ctx = c;
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /**
     * Replaces the per-request compile block AnalysisService.java:56-86. {@code flat} is every pattern of every
     * pattern set in (set, pattern) list order, sets with {@code getPatterns() == null} skipped (AS:57-59); the
     * index into it is what comes back in {@link Columns#pattern}.
     */
    public void loadPatterns(List<Pattern> flat) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment descs = a.allocateArray(PATTERN, Math.max(1, flat.size()));
            for (int i = 0; i < flat.size(); i++) {
                Pattern p = flat.get(i);
                MemorySegment d = descs.asSlice(i * PATTERN.byteSize(), PATTERN.byteSize());
                d.set(ADDRESS, 0, p.getId() == null ? MemorySegment.NULL : a.allocateUtf8String(p.getId()));
                d.set(ADDRESS, 8, a.allocateUtf8String(p.getPrimaryPattern().getRegex()));
                d.set(JAVA_DOUBLE, 16, p.getPrimaryPattern().getConfidence());
                String sev = p.getSeverity() == null ? "" : p.getSeverity().toUpperCase();
                d.set(JAVA_DOUBLE, 24, SEVERITY_MULTIPLIERS.getOrDefault(sev, 1.0));        // SS:68-69
                boolean ctxRules = p.getContextExtraction() != null;                         // AS:137-139
                d.set(JAVA_INT, 32, ctxRules ? 1 : 0);
                d.set(JAVA_INT, 36, ctxRules ? p.getContextExtraction().getLinesBefore() : 0);
                d.set(JAVA_INT, 40, ctxRules ? p.getContextExtraction().getLinesAfter() : 0);
                List<SecondaryPattern> secs = p.getSecondaryPatterns() == null ? List.of() : p.getSecondaryPatterns();
                MemorySegment sa = a.allocateArray(SECONDARY, Math.max(1, secs.size()));
                for (int j = 0; j < secs.size(); j++) {
                    long o = j * SECONDARY.byteSize();
                    sa.set(ADDRESS, o, a.allocateUtf8String(secs.get(j).getRegex()));
                    sa.set(JAVA_DOUBLE, o + 8, secs.get(j).getWeight());
                    sa.set(JAVA_INT, o + 16, secs.get(j).getProximityWindow());
                }
                d.set(JAVA_INT, 44, secs.size());
                d.set(ADDRESS, 48, sa);
                List<SequencePattern> seqs = p.getSequencePatterns() == null ? List.of() : p.getSequencePatterns();
                MemorySegment qa = a.allocateArray(SEQUENCE, Math.max(1, seqs.size()));
                for (int j = 0; j < seqs.size(); j++) {
                    List<SequenceEvent> ev = seqs.get(j).getEvents() == null ? List.of() : seqs.get(j).getEvents();
                    MemorySegment ea = a.allocateArray(ADDRESS, Math.max(1, ev.size()));
                    for (int k = 0; k < ev.size(); k++) {
                        ea.setAtIndex(ADDRESS, k, a.allocateUtf8String(ev.get(k).getRegex()));
                    }
                    long o = j * SEQUENCE.byteSize();
                    qa.set(ADDRESS, o, ea);
                    qa.set(JAVA_INT, o + 8, ev.size());
                    qa.set(JAVA_DOUBLE, o + 16, seqs.get(j).getBonusMultiplier());
                }
                d.set(JAVA_INT, 56, seqs.size());
                d.set(ADDRESS, 64, qa);
            }
            MemorySegment bad = a.allocate(JAVA_LONG);
            int rc = (int) LOAD.invokeExact(ctx, descs, (long) flat.size(), bad);
            if (rc == LP_ERR_SYNTAX) {      // what AnalysisService.java:64,70,80 would have thrown on the first request
                int i = (int) bad.get(JAVA_LONG, 0);
                String rx = i >= 0 && i < flat.size() ? flat.get(i).getPrimaryPattern().getRegex() : "";
                throw new PatternSyntaxException(lastError(), rx, -1);
            }
            if (rc != 0) {
                throw new IllegalStateException("lp_load_patterns: " + lastError());
            }
            nPatterns = flat.size();
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /**
     * Replaces AnalysisService.java:53-113 for one request. {@code priorInWindow[p]} = matches of pattern p's id
     * inside the sliding frequency window at request start ({@link FrequencyBridge#priorCounts}); the library adds
     * the in-request rank itself, which is the read-before-record order of ScoringService.java:84-88.
     */
    public Columns analyze(String logs, long[] priorInWindow, int options) {
        byte[] utf8 = logs.getBytes(StandardCharsets.UTF_8);
        HostBuf hb = hostBufs.get();
        try (Arena a = Arena.ofConfined()) {
            // the request bytes go through a page-locked buffer owned by the library: the host->device copy then runs
            // asynchronously, in pieces, underneath the scan of the pieces that have already landed
            if (utf8.length > hb.cap) {
                if (!hb.seg.equals(MemorySegment.NULL)) {
                    HOST_FREE.invokeExact(hb.seg);
                }
                hb.cap = Math.max(1L << 20, utf8.length + (utf8.length >> 2));
                MemorySegment p = (MemorySegment) HOST_ALLOC.invokeExact(hb.cap);
                if (p.equals(MemorySegment.NULL)) {
                    hb.cap = 0;
                    hb.seg = MemorySegment.NULL;
                    throw new IllegalStateException("lp_host_alloc: " + lastError());
                }
                hb.seg = p.reinterpret(hb.cap);
            }
            MemorySegment.copy(utf8, 0, hb.seg, JAVA_BYTE, 0, utf8.length);
            MemorySegment prior = priorInWindow == null ? MemorySegment.NULL : a.allocateArray(JAVA_LONG, priorInWindow);
            MemorySegment out = a.allocate(ADDRESS);
            int rc = (int) ANALYZE.invokeExact(ctx, hb.seg, (long) utf8.length, prior, options, out);
            if (rc != 0) {
                throw new IllegalStateException("lp_analyze: " + lastError());
            }
            MemorySegment res = out.get(ADDRESS, 0);
            try {
                MemorySegment v = a.allocate(VIEW);
                int grc = (int) GET.invokeExact(res, v);
                if (grc != 0) {
                    throw new IllegalStateException("lp_result_get: " + lastError());
                }
                int m = (int) v.get(JAVA_LONG, 8);
                int total = (int) v.get(ADDRESS, 16).reinterpret(8).get(JAVA_LONG, 0);       // log_n_lines[0] = logLines.length
                int[] line = v.get(ADDRESS, 32).reinterpret(4L * m).toArray(JAVA_INT);
                int[] pat = v.get(ADDRESS, 40).reinterpret(4L * m).toArray(JAVA_INT);
                double[] score = v.get(ADDRESS, 48).reinterpret(8L * m).toArray(JAVA_DOUBLE);
                int[] rank = v.get(ADDRESS, 56).reinterpret(4L * m).toArray(JAVA_INT);
                long[] delta = v.get(ADDRESS, 72).reinterpret(8L * nPatterns).toArray(JAVA_LONG);
                long abandoned = v.get(JAVA_LONG, 128);      // evaluations the backtracking verifier gave up on (count as no match)
                return new Columns(total, line, pat, score, rank, delta, abandoned);
            } finally {
                FREE.invokeExact(res);
            }
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    private static String lastError() {
        try {
            return ((MemorySegment) ERR.invokeExact()).reinterpret(4096).getUtf8String(0);
        } catch (Throwable t) {
            return "lp_last_error failed: " + t;
        }
    }

    @Override
    public synchronized void close() {
        try {
            for (HostBuf b : allBufs) {
                if (!b.seg.equals(MemorySegment.NULL)) {
                    HOST_FREE.invokeExact(b.seg);
                    b.seg = MemorySegment.NULL;
                    b.cap = 0;
                }
            }
            DESTROY.invokeExact(ctx);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }
}
