"""Deterministic synthetic workloads for BASELINE.json's configs.

The reference ships no sample patterns or logs, so the shape is the one SURVEY §8d fixes: Java
pod logs (68 % `ts LEVEL [thread] logger - msg`, 30 % stack-trace lines, 2 % blank/odd lines incl.
a few \\r\\n), mean ~150 B/line, ASCII; pattern sets whose primaries are literal cores decorated
with classes / alternation / bounded repeats, 0-3 secondaries, 0-2 sequences; ~1 % of lines hit a
primary. Master seed 20261106.
"""
from __future__ import annotations

import random
from dataclasses import dataclass
from typing import Callable, List, Tuple

import numpy as np

from .model import PatternSpec, SecondaryPattern, SequencePattern

MASTER_SEED = 20261106

_W1 = ["Socket", "Connection", "Pool", "Session", "Channel", "Stream", "Lock", "Cache", "Index", "Shard", "Queue", "Topic",
       "Broker", "Replica", "Leader", "Quorum", "Lease", "Token", "Cert", "Tls", "Dns", "Route", "Proxy", "Gateway", "Volume",
       "Mount", "Disk", "Inode", "Heap", "Thread", "Executor", "Scheduler", "Batch", "Cursor", "Schema", "Migration", "Tenant",
       "Quota", "Rate", "Circuit"]
_W2 = ["Timeout", "Closed", "Refused", "Exhausted", "Corrupt", "Missing", "Conflict", "Overflow", "Starved", "Expired",
       "Rejected", "Unavailable", "Mismatch", "Interrupted", "Deadlock"]
_SUF = ["Exception", "Error", "Failure", "Fault"]
_PKG = ["com.acme.orders", "com.acme.billing", "com.acme.auth", "com.acme.search", "org.apache.kafka.clients", "io.netty.channel",
        "org.postgresql.core", "org.springframework.web", "com.zaxxer.hikari.pool", "io.grpc.internal", "redis.clients.jedis",
        "org.hibernate.engine", "com.acme.ingest.pipeline", "com.acme.gateway.filter", "java.util.concurrent", "java.net", "java.io"]
_CLS = ["Service", "Handler", "Worker", "Manager", "Client", "Controller", "Repository", "Listener", "Dispatcher", "Processor",
        "Factory", "Resolver", "Adapter", "Registry", "Monitor"]
_VERB = ["process", "handle", "execute", "invoke", "run", "doFilter", "commit", "flush", "poll", "send", "receive", "apply", "load"]
_PHRASES = [
    ("failed to acquire lock {w} after {n}ms", r"[Ff]ailed to acquire lock \w+ after \d+ ?ms"),
    ("connection to {w}:{p} refused", r"connection to [\w.-]+:\d{2,5} (refused|reset)"),
    ("connection to {w}:{p} reset", r"connection to [\w.-]+:\d{2,5} (refused|reset)"),
    ("liveness probe failed: HTTP probe failed with statuscode: 5{d}{d}", r"(liveness|readiness) probe failed: .*statuscode: 5\d\d"),
    ("readiness probe failed: HTTP probe failed with statuscode: 5{d}{d}", r"(liveness|readiness) probe failed: .*statuscode: 5\d\d"),
    ("container {w} terminated: OOMKilled exit code 137", r"OOMKilled|exit code (137|143)"),
    ("disk quota exceeded on volume {w}", r"disk quota exceeded on volume \S+"),
    ("no space left on device /var/lib/{w}", r"[Nn]o space left on device"),
    ("TLS handshake failed for peer {w}: certificate has expired", r"TLS handshake failed.*certificate (has )?expired"),
    ("could not resolve host {w}.svc.cluster.local", r"could not resolve host [\w.-]+\.svc\.cluster\.local"),
    ("leader election lost for partition {n}", r"leader election lost for partition \d+"),
    ("too many open files (ulimit {n})", r"[Tt]oo many open files"),
    ("checksum mismatch in segment {w}-{n}", r"checksum mismatch in segment [\w-]+"),
    ("replication lag {n}s exceeds threshold", r"replication lag \d+s exceeds threshold"),
    ("panic: runtime error: invalid memory address {w}", r"panic: runtime error: .*"),
]
_NORMAL = [
    "request {w} completed in {n}ms status=200", "heartbeat ok node={w} seq={n}", "cache hit ratio {d}{d}.{d}% for region {w}",
    "scheduled job {w} finished, {n} rows", "GET /api/v1/{w}/{n} HTTP/1.1\" 200 {n}", "acquired lease {w} ttl={n}s",
    "flushed {n} records to partition {d}", "user {w} authenticated via oidc", "compaction of segment {w}-{n} done",
    "rebalanced {d} shards across {d} nodes", "config reloaded from /etc/{w}/app.yaml", "gc pause {n}ms young gen",
    "opened connection {n} to {w}:{p}", "published event {w} id={n}", "retrying in {n}ms attempt {d}",
]
_SYMPTOMS = [
    ("retrying in {n}ms attempt {d}", r"retrying in \d+ms"), ("gc pause {n}ms young gen", r"gc pause \d+ms"),
    ("connection pool exhausted, {n} waiters", r"connection pool exhausted"), ("circuit breaker {w} opened", r"circuit breaker \w+ opened"),
    ("slow query took {n}ms: select * from {w}", r"slow query took \d+ms"), ("backpressure applied on stream {w}", r"backpressure applied"),
    ("thread pool {w} saturated queue={n}", r"thread pool \w+ saturated"), ("evicting {n} entries from cache {w}", r"evicting \d+ entries"),
    ("request {w} timed out after {n}ms", r"timed out after \d+ms"), ("memory usage {d}{d}% of limit", r"memory usage \d\d% of limit"),
    ("rate limit exceeded for tenant {w}", r"rate limit exceeded"), ("dropping message {n}: queue full", r"dropping message \d+"),
]
_PHASES = [
    ("Starting {w}Service on pod-{n}", r"Starting \w+Service"), ("Initializing datasource {w}", r"Initializing datasource"),
    ("Loaded {n} plugins", r"Loaded \d+ plugins"), ("Bound to port {p}", r"Bound to port \d+"),
    ("Joined cluster as member {w}", r"Joined cluster as member"), ("Began draining connections", r"Began draining connections"),
    ("Shutdown hook invoked", r"Shutdown hook invoked"), ("Received SIGTERM", r"Received SIGTERM"),
    ("Migration {n} applied", r"Migration \d+ applied"), ("Elected leader for term {n}", r"Elected leader for term \d+"),
]
_WORDS = ["alpha", "bravo", "cart", "delta", "edge", "feed", "gamma", "hub", "index", "jobs", "kappa", "ledger", "mesh", "nexus",
          "orders", "pay", "queue", "relay", "search", "trace", "users", "vault", "wallet", "xfer", "yield", "zone"]
_TAILS = [" trace_id={n}{n}{n} span={n} tenant={w}", " req={w}-{n}-{n} zone=eu-west-{d} attempt={d} elapsed={n}us",
          " correlation={n}{n} pod=svc-{w}-{n} node=ip-10-{d}{d}-{d}-{n}", " user={w} session={n}{n} bytes={n}", ""]
_SEV = ["CRITICAL", "HIGH", "MEDIUM", "LOW", "INFO"]


def _fill(t: str, rng: random.Random) -> str:
    out = []
    i = 0
    while i < len(t):
        if t[i] == "{" and i + 2 < len(t) and t[i + 2] == "}":
            k = t[i + 1]
            if k == "w":
                out.append(rng.choice(_WORDS) + str(rng.randint(1, 99)))
            elif k == "n":
                out.append(str(rng.randint(1, 99999)))
            elif k == "p":
                out.append(str(rng.randint(1024, 65535)))
            elif k == "d":
                out.append(str(rng.randint(0, 9)))
            i += 3
        else:
            out.append(t[i])
            i += 1
    return "".join(out)


@dataclass
class Workload:
    patterns: List[PatternSpec]
    renderers: List[Callable[[random.Random], str]]   # text that makes pattern i's primary match
    symptom_renderers: List[Callable[[random.Random], str]]
    phase_renderers: List[Callable[[random.Random], str]]
    nonregular: bool = False


def _exception_name(rng: random.Random, used: set) -> str:
    for _ in range(1000):
        n = rng.choice(_W1) + rng.choice(_W2) + rng.choice(_SUF)
        if n not in used:
            used.add(n)
            return n
    n = "Custom%dException" % len(used)
    used.add(n)
    return n


def make_patterns(n_patterns: int, seed: int, n_sets: int = 5, nonregular_frac: float = 0.0, scoring_heavy: bool = False,
                  noliteral: int = None) -> Workload:
    """P primaries in n_sets sets (flat (set, pattern) order). nonregular_frac > 0 builds config C3."""
    rng = random.Random(seed)
    used = set()
    pats: List[PatternSpec] = []
    rend: List[Callable[[random.Random], str]] = []
    sym = [(lambda r, t=t: _fill(t, r)) for t, _ in _SYMPTOMS]
    pha = [(lambda r, t=t: _fill(t, r)) for t, _ in _PHASES]
    if noliteral is None:
        noliteral = max(0, n_patterns // 100)
    for i in range(n_patterns):
        kind = rng.random()
        nonreg = rng.random() < nonregular_frac
        if i < noliteral:
            # no usable literal: every line is a candidate for these
            choice = i % 3
            if choice == 0:
                rx, text = r"\b\d{8,}ms\b", "operation took 1{d}{d}{n}00000ms"
            elif choice == 1:
                rx, text = r"^\s+\.\.\. \d+ more$", "\t... {d}{d} more"
            else:
                rx, text = r"\b[45]\d\d \d+ \d+\.\d{3}$", "upstream replied 5{d}{d} {n} {d}.{d}{d}{d}"
            r_ = (lambda r, t=text: _fill(t, r))
        elif kind < 0.45:
            name = _exception_name(rng, used)
            form = rng.randint(0, 3)
            if nonreg:
                form = 4 + rng.randint(0, 2)
            if form == 0:
                rx = r"\b" + name + r"\b"
            elif form == 1:
                rx = r"(?:java|javax|com|org|io|redis)\.[\w.]+\." + name
            elif form == 2:
                rx = name + r": .*(timed out|refused|closed|failed)"
            elif form == 3:
                rx = r"Caused by: \S*" + name
            elif form == 4:
                rx = r"(?<![\w.])" + name + r"(?!\w)"           # look-around
            elif form == 5:
                rx = name + r": (\w+) .*\b\1\b"                  # back-reference
            else:
                rx = r"\b" + name + r"(?>: \w++)"                # atomic + possessive
            pk = rng.choice(_PKG)
            tails = ["operation timed out", "peer refused", "stream closed", "write failed"]
            if form == 3:
                r_ = (lambda r, pk=pk, name=name, tails=tails: "Caused by: %s.%s: %s" % (pk, name, r.choice(tails)))
            elif form == 5:
                r_ = (lambda r, pk=pk, name=name: (lambda w: "%s.%s: %s failed for %s again" % (pk, name, w, w))(r.choice(_WORDS)))
            else:
                r_ = (lambda r, pk=pk, name=name, tails=tails: "%s.%s: %s" % (pk, name, r.choice(tails)))
        elif kind < 0.9:
            text, rx = rng.choice(_PHRASES)
            # specialise the phrase so that patterns differ: pin a word into both the text and the regex
            tag = rng.choice(_WORDS) + str(rng.randint(100, 999))
            text2 = text + " [" + tag + "]"
            rx2 = rx + r".*\[" + tag + r"\]"
            if nonreg:
                rx2 = r"(?=.*\[" + tag + r"\])" + rx
            rx = rx2
            r_ = (lambda r, t=text2: _fill(t, r))
        else:
            code = rng.randint(500, 599)
            tag = rng.choice(_WORDS)
            rx = r"HTTP/1\.[01]\" %d \d+ .*svc=%s\d*" % (code, tag)
            r_ = (lambda r, code=code, tag=tag: "POST /api/%s/%d HTTP/1.1\" %d %d upstream svc=%s%d"
                  % (r.choice(_WORDS), r.randint(1, 9999), code, r.randint(0, 9999), tag, r.randint(1, 9)))
        secs = []
        seqs = []
        if scoring_heavy:
            for _ in range(3):
                j = rng.randrange(len(_SYMPTOMS))
                secs.append(SecondaryPattern(_SYMPTOMS[j][1], round(rng.uniform(0.3, 0.8), 3), rng.randint(200, 1000)))
            for _ in range(2):
                k = rng.randint(2, 4)
                ev = [_PHASES[j][1] for j in rng.sample(range(len(_PHASES)), k)]
                seqs.append(SequencePattern(ev, round(rng.uniform(0.2, 0.5), 3)))
            before, after = 500, 500
        else:
            for _ in range(rng.randint(0, 3)):
                j = rng.randrange(len(_SYMPTOMS))
                secs.append(SecondaryPattern(_SYMPTOMS[j][1], round(rng.uniform(0.3, 0.8), 3), rng.randint(5, 50)))
            for _ in range(rng.randint(0, 2)):
                k = rng.randint(2, 4)
                ev = [_PHASES[j][1] for j in rng.sample(range(len(_PHASES)), k - 1)]
                # the last event of a sequence usually sits at the failure itself
                ev.append(rng.choice([_SYMPTOMS[rng.randrange(len(_SYMPTOMS))][1], r"\b(ERROR|FATAL)\b", _PHASES[rng.randrange(len(_PHASES))][1]]))
                seqs.append(SequencePattern(ev, round(rng.uniform(0.2, 0.5), 3)))
            before, after = rng.randint(0, 10), rng.randint(0, 20)
        has_ctx = rng.random() < 0.9
        pats.append(PatternSpec(
            id="pat-%04d" % i, name="pattern %d" % i, primary_regex=rx, confidence=round(rng.uniform(0.3, 1.0), 3),
            severity=rng.choice(_SEV) if rng.random() < 0.95 else "unknown", set_index=i * n_sets // max(1, n_patterns),
            library_id="lib-%d" % (i * n_sets // max(1, n_patterns)),
            lines_before=before if has_ctx else None, lines_after=after if has_ctx else None,
            secondaries=secs, sequences=seqs))
        rend.append(r_)
    return Workload(pats, rend, sym, pha, nonregular=nonregular_frac > 0)


def _prefix(rng: random.Random, level: str) -> str:
    return "2026-11-%02dT%02d:%02d:%02d.%03dZ %-5s [%s-%d] %s.%s%s - " % (
        rng.randint(1, 28), rng.randint(0, 23), rng.randint(0, 59), rng.randint(0, 59), rng.randint(0, 999), level,
        rng.choice(["http-nio-8080-exec", "kafka-consumer", "scheduler", "grpc-default-executor", "main"]), rng.randint(1, 64),
        rng.choice(_PKG), rng.choice(_W1), rng.choice(_CLS))


def _frames(rng: random.Random, n: int) -> List[str]:
    out = []
    for _ in range(n):
        cls = rng.choice(_W1) + rng.choice(_CLS)
        out.append("\tat %s.%s.%s(%s.java:%d)" % (rng.choice(_PKG), cls, rng.choice(_VERB), cls, rng.randint(10, 999)))
    return out


def make_log(wl: Workload, n_lines: int, seed: int, hit_rate: float = 0.01, symptom_rate: float = 0.03,
             phase_rate: float = 0.002, pool_blocks: int = 6000, crlf_rate: float = 0.002) -> bytes:
    """One pod log of exactly n_lines lines (Java split semantics), '\\n'-terminated."""
    rng = random.Random(seed)
    nrng = np.random.default_rng(seed)
    P = len(wl.patterns)
    blocks: List[Tuple[bytes, int]] = []   # (text without trailing newline, number of lines)

    def level():
        x = rng.random()
        return "INFO" if x < 0.78 else "DEBUG" if x < 0.88 else "WARN" if x < 0.96 else "ERROR"

    # Zipf-ish popularity of patterns so a few dominate, as real incident logs do
    weights = np.array([1.0 / (1 + (i % 97)) for i in range(P)]) if P else np.array([1.0])
    weights = weights / weights.sum()
    n_trace_blocks = max(1, int(pool_blocks * 0.012 / (0.012 + 0.68 + 0.02) + 0.5))
    n_normal = pool_blocks
    for b in range(n_normal):
        x = rng.random()
        if x < hit_rate and P:
            i = int(nrng.choice(P, p=weights))
            msg = wl.renderers[i](rng)
            if msg.startswith("\t") or msg.startswith("Caused by"):
                line = msg
            else:
                line = _prefix(rng, rng.choice(["ERROR", "WARN", "ERROR", "FATAL"])) + msg
        elif x < hit_rate + symptom_rate:
            line = _prefix(rng, rng.choice(["WARN", "INFO"])) + rng.choice(wl.symptom_renderers)(rng)
        elif x < hit_rate + symptom_rate + phase_rate:
            line = _prefix(rng, "INFO") + rng.choice(wl.phase_renderers)(rng)
        elif x < hit_rate + symptom_rate + phase_rate + 0.02:
            line = rng.choice(["", "", "   ", "---", "====== restart ======", "\r", "plain text without a prefix"])
        else:
            line = _prefix(rng, level()) + _fill(rng.choice(_NORMAL), rng) + _fill(rng.choice(_TAILS), rng)
        if rng.random() < crlf_rate:
            line = line + "\r"   # becomes \r\n when joined: the \r?\n separator eats it
        blocks.append((line.encode("ascii"), 1))
    for b in range(n_trace_blocks * 4):
        pk = rng.choice(_PKG)
        if rng.random() < 0.35 and P:
            i = int(nrng.choice(P, p=weights))
            head = wl.renderers[i](rng)
            if head.startswith("\t"):
                head = "%s.%s%s: boom" % (pk, rng.choice(_W1), "Exception")
        else:
            head = "%s.%s%s%s: %s" % (pk, rng.choice(_W1), rng.choice(_W2), rng.choice(["Exception", "Error"]), _fill("request {w} failed after {n}ms", rng))
        lines = [_prefix(rng, "ERROR") + "Unhandled exception in " + rng.choice(_VERB), head] + _frames(rng, rng.randint(8, 40))
        if rng.random() < 0.5:
            lines.append("Caused by: %s.%s%sException: %s" % (rng.choice(_PKG), rng.choice(_W1), rng.choice(_W2), _fill("cause {w}", rng)))
            lines += _frames(rng, rng.randint(3, 12))
            lines.append("\t... %d more" % rng.randint(3, 40))
        blocks.append(("\n".join(lines).encode("ascii"), len(lines)))
    single = [b for b in blocks if b[1] == 1]
    traces = [b for b in blocks if b[1] > 1]
    avg_trace = sum(t[1] for t in traces) / len(traces)
    p_trace = 0.30 / avg_trace / (0.30 / avg_trace + 0.70)
    out: List[bytes] = []
    total = 0
    # draw in chunks with numpy for speed
    while total < n_lines:
        k = max(1024, (n_lines - total))
        is_trace = nrng.random(k) < p_trace
        si = nrng.integers(0, len(single), k)
        ti = nrng.integers(0, len(traces), k)
        for j in range(k):
            if is_trace[j]:
                t, c = traces[ti[j]]
                if total + c > n_lines:
                    continue
            else:
                t, c = single[si[j]]
            out.append(t)
            total += c
            if total >= n_lines:
                break
    # Java's split drops trailing empty strings: make sure the last line is not empty
    while out and out[-1] in (b"", ):
        out[-1] = b"end of log"
    return b"\n".join(out) + b"\n"


# ---------------------------------------------------------------------------- named configs
def config_c1(seed: int = MASTER_SEED ^ 1):
    wl = make_patterns(50, seed, n_sets=5, noliteral=1)
    return wl, make_log(wl, 10_000, seed ^ 0x5151, pool_blocks=3000)


def config_c2(n_lines: int = 1_000_000, n_patterns: int = 500, seed: int = MASTER_SEED ^ 2):
    wl = make_patterns(n_patterns, seed, n_sets=5)
    return wl, make_log(wl, n_lines, seed ^ 0x5252)


def config_c3(n_lines: int = 1_000_000, n_patterns: int = 500, seed: int = MASTER_SEED ^ 3):
    wl = make_patterns(n_patterns, seed, n_sets=5, nonregular_frac=0.12)
    return wl, make_log(wl, n_lines, seed ^ 0x5353)


def config_c4(n_lines: int = 1_000_000, n_patterns: int = 500, seed: int = MASTER_SEED ^ 4):
    wl = make_patterns(n_patterns, seed, n_sets=5, scoring_heavy=True, noliteral=0)
    return wl, make_log(wl, n_lines, seed ^ 0x5454, hit_rate=0.05, symptom_rate=0.4, phase_rate=0.01)


def config_c5_logs(n_logs: int, n_lines: int = 20_000, n_patterns: int = 500, seed: int = MASTER_SEED ^ 5, distinct: int = 16):
    """Batch of pod logs; `distinct` different logs are generated and tiled (a pool of
    distinct logs instead of a 30 GB host set)."""
    wl = make_patterns(n_patterns, seed, n_sets=5)
    pool = [make_log(wl, n_lines, seed ^ (0x5500 + i), pool_blocks=4000) for i in range(min(distinct, n_logs))]
    return wl, [pool[i % len(pool)] for i in range(n_logs)]
