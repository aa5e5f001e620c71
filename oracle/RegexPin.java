// ORACLE — TEST INFRASTRUCTURE ONLY. Never shipped, never on the product path.
//
// Replays tests/golden/regex_pin_corpus.tsv.gz (written by tests/golden/make_regex_pin_corpus.py: what the C++
// restatement in oracle/jregex.cpp answers for ~5 000 (pattern, flags, text) cases) on the REAL java.util.regex, the
// engine behind AnalysisService.java:64,93-95 (Pattern.compile(regex) ... matcher(line).find()), and lists every case
// on which the JDK answers differently. The authoring image has no JDK, so this file has not been compiled or run
// there; on a box with JDK 21:
//     java oracle/RegexPin.java [tests/golden/regex_pin_corpus.tsv.gz]
// Exit status 0 = no unexpected disagreement. One class of disagreement is known and labelled as such: Pattern.Bound
// counts a non-spacing mark that follows a letter or digit as a word character, the restatement and the library test
// the ASCII word class only (DESIGN.md §3). A second one is labelled too: Pattern.Behind steps back in UTF-16 units
// unless the pattern itself holds a supplementary character (BehindS), the restatement steps back in code points, so a
// look-behind right after a character beyond the BMP can see half of it on the JDK.
//
// Line format, TAB separated: base64(pattern UTF-8), flags (Pattern.* bit values), base64(text UTF-8), found (0/1),
// start, end of the first find() in code points (-1 -1 when not found).
import java.io.BufferedReader;
import java.io.FileInputStream;
import java.io.InputStreamReader;
import java.nio.charset.StandardCharsets;
import java.util.Base64;
import java.util.regex.Matcher;
import java.util.regex.Pattern;
import java.util.regex.PatternSyntaxException;
import java.util.zip.GZIPInputStream;

public class RegexPin {
    public static void main(String[] args) throws Exception {
        String path = args.length > 0 ? args[0] : "tests/golden/regex_pin_corpus.tsv.gz";
        int cases = 0, same = 0, known = 0, unexpected = 0, rejected = 0;
        Base64.Decoder b64 = Base64.getDecoder();
        try (BufferedReader in = new BufferedReader(new InputStreamReader(new GZIPInputStream(new FileInputStream(path)), StandardCharsets.US_ASCII))) {
            String row;
            while ((row = in.readLine()) != null) {
                if (row.isEmpty()) continue;
                String[] f = row.split("\t");
                String regex = new String(b64.decode(f[0]), StandardCharsets.UTF_8);
                int flags = Integer.parseInt(f[1]);
                String text = new String(b64.decode(f[2]), StandardCharsets.UTF_8);
                boolean wantFound = f[3].equals("1");
                int wantStart = Integer.parseInt(f[4]), wantEnd = Integer.parseInt(f[5]);
                cases++;
                Pattern p;
                try {
                    p = Pattern.compile(regex, flags);
                } catch (PatternSyntaxException e) {
                    rejected++;
                    System.out.println("REJECTED BY THE JDK\t" + show(regex) + "\t" + e.getDescription());
                    continue;
                }
                Matcher m = p.matcher(text);
                boolean found = m.find();
                int start = found ? text.codePointCount(0, m.start()) : -1;
                int end = found ? text.codePointCount(0, m.end()) : -1;
                if (found == wantFound && start == wantStart && end == wantEnd) {
                    same++;
                    continue;
                }
                boolean mark = text.codePoints().anyMatch(c -> Character.getType(c) == Character.NON_SPACING_MARK);
                boolean boundary = regex.contains("\\b") || regex.contains("\\B");
                boolean unicodeClass = (flags & Pattern.UNICODE_CHARACTER_CLASS) != 0 || regex.contains("(?U");
                boolean beyondBmp = text.codePoints().anyMatch(Character::isSupplementaryCodePoint);
                boolean lookBehind = regex.contains("(?<=") || regex.contains("(?<!");
                String label;
                if (mark && boundary && !unicodeClass) {
                    known++;
                    label = "known (mark at a word boundary)";
                } else if (beyondBmp && lookBehind) {
                    known++;
                    label = "known (look-behind next to a character beyond the BMP)";
                } else {
                    unexpected++;
                    label = "UNEXPECTED";
                }
                System.out.println(label + "\t" + show(regex) + "\tflags=" + flags + "\t" + show(text) + "\tjdk=" + found + " [" + start + "," + end
                        + ")\trestatement=" + wantFound + " [" + wantStart + "," + wantEnd + ")");
            }
        }
        System.out.println("cases=" + cases + " same=" + same + " known_difference=" + known + " unexpected=" + unexpected + " rejected_by_jdk=" + rejected);
        System.exit(unexpected == 0 && rejected == 0 ? 0 : 1);
    }

    // printable form: non-ASCII and control characters as \\uXXXX
    static String show(String s) {
        StringBuilder b = new StringBuilder();
        for (int i = 0; i < s.length(); i++) {
            char c = s.charAt(i);
            if (c >= 0x20 && c < 0x7f) b.append(c);
            else b.append(String.format("\\u%04x", (int) c));
        }
        return b.toString();
    }
}
