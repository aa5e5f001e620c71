package com.redhat.podmortem.parser.gpu;

import com.redhat.podmortem.common.model.analysis.PatternFrequency;
import com.redhat.podmortem.common.model.pattern.Pattern;
import com.redhat.podmortem.parser.service.FrequencyTrackingService;
import jakarta.enterprise.context.ApplicationScoped;
import jakarta.inject.Inject;
import java.util.List;
import java.util.Map;

/**
 * Keeps the frequency state where the reference keeps it - in {@link FrequencyTrackingService}'s
 * {@code Map<String, PatternFrequency>} with its sliding window and clock (FrequencyTrackingService.java:25,41-56)
 * - and moves only numbers across the FFM boundary:
 *
 * <ul>
 *   <li>before a request: {@link #priorCounts} snapshots, per flat pattern index, how many matches the
 *       pattern's id already has inside the window ({@code PatternFrequency.getCurrentCount()}, the value
 *       FTS:113 reports). The device adds the position of each event among the request's events of the same id,
 *       which reproduces the read-before-record order of ScoringService.java:84-88 for the whole request.
 *   <li>after a request: {@link #apply} records the request's matches ({@code freq_delta[p]} per pattern) through
 *       {@code recordPatternMatch}, so the admin API (getFrequencyStatistics / resetPatternFrequency /
 *       resetAllFrequencies, FTS:101-138) keeps working on the same map, untouched.
 * </ul>
 *
 * <p>The penalty formula itself (FTS:64-93: {@code rate <= threshold ? 0 : min(maxPenalty, (rate - threshold) /
 * threshold)}) runs on the device with {@code rate = count / timeWindowHours}; {@code getHourlyRate()} of the
 * un-vendored common library is not in the reference tree, see DESIGN.md.
 */
@ApplicationScoped
public class FrequencyBridge {

    @Inject FrequencyTrackingService fts;

    /** One entry per flat pattern: the in-window count of its id (0 for patterns without an id, FTS:65-67). */
    public long[] priorCounts(List<Pattern> flat) {
        long[] prior = new long[flat.size()];
        for (int p = 0; p < prior.length; p++) {
            String id = flat.get(p).getId();
            if (id == null || id.trim().isEmpty()) {
                continue;
            }
            PatternFrequency f = fts.getPatternFrequency(id);
            prior[p] = f == null ? 0 : f.getCurrentCount();
        }
        return prior;
    }

    /** FTS:41-56, once per event of the request (patterns that share an id add up in the one entry of that id). */
    public void apply(List<Pattern> flat, long[] freqDelta) {
        for (int p = 0; p < freqDelta.length; p++) {
            for (long c = freqDelta[p]; c > 0; c--) {
                fts.recordPatternMatch(flat.get(p).getId());
            }
        }
    }

    // ---- the admin surface of the reference, delegated unchanged (FTS:101-161)
    public Map<String, Integer> getFrequencyStatistics() {
        return fts.getFrequencyStatistics();
    }

    public void resetPatternFrequency(String patternId) {
        fts.resetPatternFrequency(patternId);
    }

    public void resetAllFrequencies() {
        fts.resetAllFrequencies();
    }
}
