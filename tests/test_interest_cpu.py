"""CPU test of the interest table: host builder + lookup (compiled for the host as a test double) against the NATS
subject-matching model."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import hostcheck_lib as hc  # noqa: E402
import interest_case  # noqa: E402
import nats_model  # noqa: E402


def test_nats_subject_rules():
    m = nats_model.subject_matches
    assert m(b"chat.a.tokens", b"chat.a.tokens") and not m(b"chat.a.tokens", b"chat.b.tokens")
    assert m(b"chat.*.tokens", b"chat.a.tokens") and not m(b"chat.*.tokens", b"chat.a.b.tokens")
    assert m(b"chat.>", b"chat.a.tokens") and not m(b"chat.>", b"chat") and not m(b"chat.>.x", b"chat.a.x")
    assert not m(b"chat.*.tokens", b"chat..tokens") and not m(b"chat.*", b"chat.a.tokens")


import pytest  # noqa: E402


@pytest.mark.parametrize("generic", [True, False])
def test_lookup_matches_model(oracle, generic):
    lib = hc.load()
    deltas, live, payloads, subjects = interest_case.build(generic_wildcard=generic)
    arena, off = oracle.pack(payloads)
    hit, status, subs, wild_ids = hc.interest(lib, deltas, arena, off)
    assert interest_case.check(hit, status, subs, wild_ids, live, subjects) > 1000
