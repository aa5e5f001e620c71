// ORACLE — TEST INFRASTRUCTURE ONLY. Never shipped, never on the product path.
//
// Single-file, dependency-free restatement of the reference hot path on the REAL java.util.regex,
// for boxes where a JDK exists (none in the authoring image, so this file has not been compiled;
// bench.py falls back to the C++ restatement when `java` is absent or this harness fails):
//   AnalysisService.java:53        split
//   AnalysisService.java:56-86     per-request Pattern.compile of every regex
//   AnalysisService.java:89-113    N x P scan in discovery order
//   ScoringService.java:63-347, ContextAnalysisService.java:46-144, FrequencyTrackingService.java:64-93
// Frequency model: rate = in-window count / time-window-hours (SURVEY.md §8a-F; PatternFrequency's
// source is absent). Per-match log.info (AS:96-99) is not emulated.
//
// Usage: java RefHarness.java <patterns.tsv> <log file> <warmup> <steps>
// patterns.tsv, one pattern per line, TAB-separated, Base64-encoded regex fields:
//   id  confidence  severityMultiplier  hasCtx  before  after  primaryB64  nSec {regexB64 weight window}*  nSeq {bonus nEv {regexB64}*}*
// Prints: lines=<N> events=<M> checksum=<sum of scores> ms_per_step=<t>
import java.nio.charset.StandardCharsets;
import java.nio.file.Files;
import java.nio.file.Paths;
import java.util.ArrayList;
import java.util.Base64;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.regex.Pattern;

public class RefHarness {
    static final double DECAY = 10.0, EARLY = 0.2, MAX_EARLY = 2.5, PENALTY = 0.5, MAX_CTX = 2.5, FREQ_THR = 10.0, MAX_PEN = 0.8;
    static final int MAX_WINDOW = 100, HOURS = 1;

    static final Pattern ERROR_PATTERN = Pattern.compile("\\b(ERROR|FATAL|CRITICAL|SEVERE)\\b", Pattern.CASE_INSENSITIVE);
    static final Pattern WARN_PATTERN = Pattern.compile("\\b(WARN|WARNING)\\b", Pattern.CASE_INSENSITIVE);
    static final Pattern STACK_TRACE_PATTERN = Pattern.compile("^\\s*at\\s+[\\w\\.\\$]+\\(.*\\)\\s*$");
    static final Pattern EXCEPTION_PATTERN = Pattern.compile("\\b\\w*Exception\\b|\\b\\w*Error\\b");

    static final class Sec { String regex; double weight; int window; Pattern re; }
    static final class Seq { double bonus; List<String> events = new ArrayList<>(); List<Pattern> res = new ArrayList<>(); }
    static final class Pat {
        String id, primary; double confidence, severity; boolean hasCtx; int before, after; Pattern re;
        List<Sec> secs = new ArrayList<>(); List<Seq> seqs = new ArrayList<>();
    }

    static String b64(String s) { return new String(Base64.getDecoder().decode(s), StandardCharsets.UTF_8); }

    static List<Pat> load(String path) throws Exception {
        List<Pat> out = new ArrayList<>();
        for (String line : Files.readAllLines(Paths.get(path), StandardCharsets.UTF_8)) {
            if (line.isEmpty()) continue;
            String[] f = line.split("\t");
            int k = 0;
            Pat p = new Pat();
            p.id = f[k++]; p.confidence = Double.parseDouble(f[k++]); p.severity = Double.parseDouble(f[k++]);
            p.hasCtx = f[k++].equals("1"); p.before = Integer.parseInt(f[k++]); p.after = Integer.parseInt(f[k++]);
            p.primary = b64(f[k++]);
            int ns = Integer.parseInt(f[k++]);
            for (int i = 0; i < ns; i++) { Sec s = new Sec(); s.regex = b64(f[k++]); s.weight = Double.parseDouble(f[k++]); s.window = Integer.parseInt(f[k++]); p.secs.add(s); }
            int nq = Integer.parseInt(f[k++]);
            for (int i = 0; i < nq; i++) {
                Seq q = new Seq(); q.bonus = Double.parseDouble(f[k++]);
                int ne = Integer.parseInt(f[k++]);
                for (int j = 0; j < ne; j++) q.events.add(b64(f[k++]));
                p.seqs.add(q);
            }
            out.add(p);
        }
        return out;
    }

    static double chrono(int idx, int n) {
        double pos = (double) idx / n;
        if (pos <= EARLY) return 1.5 + (EARLY - pos) * ((MAX_EARLY - 1.5) / EARLY);
        if (pos <= PENALTY) return 1.0 + (PENALTY - pos) * (0.5 / (PENALTY - EARLY));
        return 0.5 + (1.0 - pos);
    }

    static double closest(Sec s, int primary, String[] lines) {
        int w = Math.min(MAX_WINDOW, s.window);
        int start = Math.max(0, primary - w), end = Math.min(lines.length, primary + w + 1);
        double best = -1;
        for (int l = start; l < end; l++) {
            if (l == primary) continue;
            if (s.re.matcher(lines[l]).find()) { double d = Math.abs(l - primary); if (best < 0 || d < best) best = d; }
        }
        return best;
    }

    static boolean sequenceMatched(Seq q, int primary, String[] lines) {
        if (q.res.isEmpty()) return false;
        int cur = 0;
        for (int i = q.res.size() - 1; i >= 0; i--) {
            Pattern re = q.res.get(i);
            if (i == q.res.size() - 1) {
                boolean found = false;
                int start = Math.max(0, primary - 5), end = Math.min(lines.length, primary + 6);
                for (int l = start; l < end && !found; l++) found = re.matcher(lines[l]).find();
                if (!found) return false;
                cur = primary;
            } else {
                int f = -1;
                for (int l = cur - 1; l >= 0; l--) if (re.matcher(lines[l]).find()) { f = l; break; }
                if (f < 0) return false;
                cur = f;
            }
        }
        return true;
    }

    static double context(String[] lines, int from, int to) {
        if (to <= from) return 1.0;
        double score = 0.0; int err = 0, stack = 0;
        for (int l = from; l < to; l++) {
            String line = lines[l];
            if (ERROR_PATTERN.matcher(line).find()) { err++; score += 0.4; }
            else if (WARN_PATTERN.matcher(line).find()) { score += 0.2; }
            if (STACK_TRACE_PATTERN.matcher(line).find()) { stack++; score += 0.1; }
            if (EXCEPTION_PATTERN.matcher(line).find()) { score += 0.3; }
        }
        if (stack > 0) score += Math.min(stack * 0.1, 0.5);
        int total = to - from;
        if (total > 10 && (stack + err) > total * 0.7) score *= 0.8;
        double f = 1.0 + score;
        return f > MAX_CTX ? MAX_CTX : f;
    }

    static double[] analyze(List<Pat> pats, String logs) {
        String[] lines = logs.split("\\r?\\n");
        for (Pat p : pats) {                                   // AS:56-86, every request
            p.re = Pattern.compile(p.primary);
            for (Sec s : p.secs) s.re = Pattern.compile(s.regex);
            for (Seq q : p.seqs) { q.res.clear(); for (String e : q.events) q.res.add(Pattern.compile(e)); }
        }
        Map<String, Integer> counts = new HashMap<>();
        double checksum = 0; long events = 0;
        for (int i = 0; i < lines.length; i++) {
            String line = lines[i];
            for (Pat p : pats) {
                if (!p.re.matcher(line).find()) continue;
                double prox = 1.0;
                if (!p.secs.isEmpty()) {
                    double t = 0.0;
                    for (Sec s : p.secs) { double d = closest(s, i, lines); if (d >= 0) t += s.weight * Math.exp(-d / DECAY); }
                    prox = 1.0 + t;
                }
                double temporal = 1.0;
                if (!p.seqs.isEmpty()) {
                    double t = 0.0;
                    for (Seq q : p.seqs) if (sequenceMatched(q, i, lines)) t += q.bonus;
                    temporal = 1.0 + t;
                }
                int from = i, to = i + 1;
                if (p.hasCtx) { from = Math.max(0, i - p.before); to = Math.min(lines.length, i + 1 + p.after); }
                double ctx = context(lines, from, to);
                double pen = 0.0;
                boolean tracked = p.id != null && !p.id.trim().isEmpty();
                if (tracked) {
                    Integer c = counts.get(p.id);
                    if (c != null) {
                        double rate = (double) c / HOURS;
                        if (rate > FREQ_THR) pen = Math.min(MAX_PEN, (rate - FREQ_THR) / FREQ_THR);
                    }
                    counts.merge(p.id, 1, Integer::sum);
                }
                double score = p.confidence * p.severity * chrono(i, lines.length) * prox * temporal * ctx * (1.0 - pen);
                checksum += score; events++;
            }
        }
        return new double[] {lines.length, events, checksum};
    }

    public static void main(String[] a) throws Exception {
        List<Pat> pats = load(a[0]);
        String logs = new String(Files.readAllBytes(Paths.get(a[1])), StandardCharsets.UTF_8);
        int warm = Integer.parseInt(a[2]), steps = Integer.parseInt(a[3]);
        double[] r = null;
        for (int i = 0; i < warm; i++) r = analyze(pats, logs);
        long t0 = System.nanoTime();
        for (int i = 0; i < steps; i++) r = analyze(pats, logs);
        double ms = (System.nanoTime() - t0) / 1e6 / Math.max(1, steps);
        System.out.println("lines=" + (long) r[0] + " events=" + (long) r[1] + " checksum=" + r[2] + " ms_per_step=" + ms);
    }
}
