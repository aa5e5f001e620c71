package com.redhat.podmortem.parser.gpu;

import com.redhat.podmortem.common.model.analysis.AnalysisMetadata;
import com.redhat.podmortem.common.model.analysis.AnalysisResult;
import com.redhat.podmortem.common.model.analysis.AnalysisSummary;
import com.redhat.podmortem.common.model.analysis.EventContext;
import com.redhat.podmortem.common.model.analysis.MatchedEvent;
import com.redhat.podmortem.common.model.kube.podmortem.PodFailureData;
import com.redhat.podmortem.common.model.pattern.ContextExtraction;
import com.redhat.podmortem.common.model.pattern.Pattern;
import com.redhat.podmortem.common.model.pattern.PatternSet;
import com.redhat.podmortem.parser.service.PatternService;
import jakarta.annotation.PostConstruct;
import jakarta.annotation.PreDestroy;
import jakarta.enterprise.context.ApplicationScoped;
import jakarta.inject.Inject;
import java.time.Instant;
import java.util.ArrayList;
import java.util.Arrays;
import java.util.Comparator;
import java.util.List;
import java.util.Map;
import java.util.UUID;
import java.util.function.Function;
import java.util.stream.Collectors;
import org.eclipse.microprofile.config.inject.ConfigProperty;

/**
 * Drop-in for {@code AnalysisService} (same {@code analyze(PodFailureData)} signature, called from Parse.java:53):
 * lines 53-113 of the reference - split, per-request regex compile, the N x P find() loop and calculateScore -
 * run on the GPU behind {@link LogParserB200}; context strings, metadata and summary are built here exactly as
 * AnalysisService.java:115-121,132-215 builds them. Uncompiled (no JDK in the authoring image).
 */
@ApplicationScoped
public class GpuAnalysisService {

    @Inject PatternService patternService;
    @Inject FrequencyBridge frequency;

    @ConfigProperty(name = "scoring.proximity.decay-constant", defaultValue = "10.0") double decayConstant;
    @ConfigProperty(name = "scoring.proximity.max-window", defaultValue = "100") int maxWindow;
    @ConfigProperty(name = "scoring.frequency.time-window-hours", defaultValue = "1") int timeWindowHours;
    @ConfigProperty(name = "scoring.chronological.early-bonus-threshold", defaultValue = "0.2") double earlyBonusThreshold;
    @ConfigProperty(name = "scoring.chronological.max-early-bonus", defaultValue = "2.5") double maxEarlyBonus;
    @ConfigProperty(name = "scoring.chronological.penalty-threshold", defaultValue = "0.5") double penaltyThreshold;
    @ConfigProperty(name = "scoring.context.max-context-factor", defaultValue = "2.5") double maxContextFactor;
    @ConfigProperty(name = "scoring.frequency.threshold", defaultValue = "10.0") double frequencyThreshold;
    @ConfigProperty(name = "scoring.frequency.max-penalty", defaultValue = "0.8") double maxPenalty;
    @ConfigProperty(name = "logparser.b200.device", defaultValue = "0") int device;

    private LogParserB200 gpu;
    private List<Pattern> flat = List.of();
    private List<PatternSet> loadedSets = List.of();

    @PostConstruct
    void start() {
        gpu = new LogParserB200(new LogParserB200.ScoringProps(decayConstant, maxWindow, timeWindowHours, earlyBonusThreshold,
                maxEarlyBonus, penaltyThreshold, maxContextFactor, frequencyThreshold, maxPenalty), device);
        reload();
    }

    /** Call again whenever PatternService reloads its directory (PatternService.java:45-69). */
    public synchronized void reload() {
        loadedSets = patternService.getPatternSets();
        List<Pattern> f = new ArrayList<>();
        for (PatternSet ps : loadedSets) {
            if (ps.getPatterns() == null) {
                continue;                                   // AS:57-59
            }
            f.addAll(ps.getPatterns());
        }
        gpu.loadPatterns(f);
        flat = f;
    }

    @PreDestroy
    void stop() {
        gpu.close();
    }

    public AnalysisResult analyze(PodFailureData data) {
        long startTime = System.currentTimeMillis();
        String logs = data.getLogs();
        List<Pattern> patterns;
        long[] prior;
        LogParserB200.Columns c;
        // Not synchronized: the reference lets concurrent requests interleave their reads and records of the frequency
        // state freely (FTS keeps a ConcurrentHashMap, AnalysisService is a shared singleton, AS:35). Here a request sees
        // the counts at its start plus its own earlier events, and requests from different threads overlap on the GPU.
        synchronized (this) {       // only the (pattern list, native plan) pair is read atomically against reload()
            patterns = flat;
        }
        prior = frequency.priorCounts(patterns);
        c = gpu.analyze(logs, prior, 0);
        frequency.apply(patterns, c.freqDelta());
        String[] logLines = logs.split("\\r?\\n");          // only for the context strings of the response (AS:132-156)
        List<MatchedEvent> foundEvents = new ArrayList<>(c.line().length);
        for (int k = 0; k < c.line().length; k++) {         // discovery order, as AS:100-109 appends them
            Pattern p = patterns.get(c.pattern()[k]);
            MatchedEvent e = new MatchedEvent();
            e.setLineNumber(c.line()[k]);
            e.setMatchedPattern(p);
            e.setContext(extractContext(logLines, c.line()[k] - 1, p.getContextExtraction()));
            e.setScore(c.score()[k]);
            foundEvents.add(e);
        }
        AnalysisResult result = new AnalysisResult();       // AS:115-121
        result.setAnalysisId(UUID.randomUUID().toString());
        result.setMetadata(buildMetadata(startTime, c.totalLines(), loadedSets));
        result.setEvents(foundEvents);
        result.setSummary(buildSummary(foundEvents));
        return result;
    }

    // AnalysisService.java:132-156
    private static EventContext extractContext(String[] allLines, int matchIndex, ContextExtraction rules) {
        EventContext context = new EventContext();
        context.setMatchedLine(allLines[matchIndex]);
        if (rules == null) {
            return context;
        }
        int beforeStart = Math.max(0, matchIndex - rules.getLinesBefore());
        context.setLinesBefore(Arrays.asList(Arrays.copyOfRange(allLines, beforeStart, matchIndex)));
        int afterEnd = Math.min(allLines.length, matchIndex + 1 + rules.getLinesAfter());
        context.setLinesAfter(Arrays.asList(Arrays.copyOfRange(allLines, matchIndex + 1, afterEnd)));
        return context;
    }

    // AnalysisService.java:166-180
    private static AnalysisMetadata buildMetadata(long startTime, int totalLines, List<PatternSet> patternSets) {
        AnalysisMetadata metadata = new AnalysisMetadata();
        metadata.setProcessingTimeMs(System.currentTimeMillis() - startTime);
        metadata.setTotalLines(totalLines);
        metadata.setAnalyzedAt(Instant.now().toString());
        metadata.setPatternsUsed(patternSets.stream().map(ps -> ps.getMetadata().getLibraryId()).collect(Collectors.toList()));
        return metadata;
    }

    // AnalysisService.java:188-215
    private static AnalysisSummary buildSummary(List<MatchedEvent> events) {
        AnalysisSummary summary = new AnalysisSummary();
        summary.setSignificantEvents(events.size());
        if (events.isEmpty()) {
            summary.setHighestSeverity("NONE");
            summary.setSeverityDistribution(Map.of());
            return summary;
        }
        Map<String, Long> distribution = events.stream()
                .map(e -> e.getMatchedPattern().getSeverity().toUpperCase())
                .collect(Collectors.groupingBy(Function.identity(), Collectors.counting()));
        summary.setSeverityDistribution(distribution);
        List<String> severityOrder = List.of("INFO", "LOW", "MEDIUM", "HIGH", "CRITICAL");
        summary.setHighestSeverity(events.stream()
                .map(e -> e.getMatchedPattern().getSeverity().toUpperCase())
                .max(Comparator.comparingInt(severityOrder::indexOf))
                .orElse("NONE"));
        return summary;
    }
}
