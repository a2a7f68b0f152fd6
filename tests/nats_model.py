"""NATS subject matching, restated for the tests (third-party semantics: nats.go v1.31.0 / nats-server 2.10, not
vendored in the reference; the reference only ever subscribes to chat.<id>.tokens and the wildcards chat.*.tokens /
chat.*.control -- src/sse-adapter/sse_handler.go:133, docs/security-inspection-patterns.md:110,268).  PARITY UNPINNED.
Subjects are '.'-separated tokens; '*' matches exactly one token, '>' as the last token matches one or more tokens;
a subject with an empty token matches nothing."""


def subject_matches(pattern: bytes, subject: bytes) -> bool:
    p, s = pattern.split(b"."), subject.split(b".")
    if any(t == b"" for t in s):
        return False
    for i, tok in enumerate(p):
        if tok == b">":
            return i == len(p) - 1 and len(s) > i
        if i >= len(s):
            return False
        if tok != b"*" and tok != s[i]:
            return False
    return len(p) == len(s)


def interested(subs, subject):
    """subs: list of (subject pattern bytes, subscriber id) -> sorted list of ids whose subscription matches."""
    return sorted(i for pat, i in subs if subject_matches(pat, subject))
