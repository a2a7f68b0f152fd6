"""The reference's two illustrative pattern sets, quoted verbatim, as (kind, action, text) triples.
kind 0 = literal (ASCII case-insensitive), 1 = regex subset; action 1 redact, 2 drop, 3 alert."""
LIT, RE = 0, 1
REDACT, DROP, ALERT = 1, 2, 3

# docs/security-inspection-patterns.md:205-209 (regexes, Action::Redact) and :229-234 (literals, Action::Alert)
SET_A = [
    (RE, REDACT, rb"\b\d{3}-\d{2}-\d{4}\b"),
    (RE, REDACT, rb"\b\d{16}\b"),
    (RE, REDACT, rb"\b[A-Za-z0-9._%+-]+@[A-Za-z0-9.-]+\.[A-Z|a-z]{2,}\b"),
    (LIT, ALERT, b"ignore previous instructions"),
    (LIT, ALERT, b"disregard the above"),
    (LIT, ALERT, b"new instructions:"),
    (LIT, ALERT, b"system prompt:"),
]
# src/spin-functions/nats-subscriber/src/lib.rs:59-64 ("redact"), :79-84 ("drop")
SET_B = [
    (LIT, REDACT, b"password"), (LIT, REDACT, b"secret"), (LIT, REDACT, b"api_key"), (LIT, REDACT, b"credit_card"),
    (LIT, DROP, b"ignore previous"), (LIT, DROP, b"disregard above"), (LIT, DROP, b"new instructions"), (LIT, DROP, b"system prompt"),
]
ACTION_NAME = {0: "allow", 1: "redact", 2: "drop", 3: "alert"}
