"""A seeded interest-table scenario shared by the CPU and GPU tests (BASELINE config 3 in miniature):
literal subscriptions chat.<id>.tokens with a skewed fan-out, two wildcard consumers, some unsubscribes,
ids that need JSON escapes, subjects nobody listens to, subscriptions that can never match this path."""
import random

import make_golden as mg
import nats_model


def build(n_conv=400, seed=17, generic_wildcard=True):
    rng = random.Random(seed)
    ids = ["conv-%d" % i for i in range(n_conv)] + ["we\"ird\\id", "d.o.t", "sp ace", "café", "", "x" * 200, "a*", ">"]
    deltas, live = [], []
    sid = 1
    for c in ids:
        r = rng.random()
        fan = 0 if r < 0.1 else (1 if r < 0.8 else rng.randrange(2, 40))
        for _ in range(fan):
            subj = b"chat." + c.encode("utf-8") + b".tokens"
            deltas.append((0, sid, subj))
            live.append((subj, sid))
            sid += 1
    wilds = [b"chat.*.tokens", b"chat.*.control", b"other.subject", b"chat.conv-3.tokens.extra"]
    if generic_wildcard:                      # a wildcard with a literal in the id position forces the token-by-token matcher
        wilds.append(b"chat.conv-7.>")
    else:
        wilds += [b"chat.>", b"*.*.*", b"chat.*.tokens.>"]
    for subj in wilds:
        deltas.append((0, sid, subj))
        live.append((subj, sid))
        sid += 1
    for _ in range(60):                                   # unsubscribe a few, one of them twice
        subj, i = live.pop(rng.randrange(len(live)))
        deltas.append((1, i, subj))
    deltas.append((1, 999999, b"chat.nobody.tokens"))
    payloads, subjects = [], []
    for k in range(3000):
        c = rng.choice(ids) if rng.random() < 0.9 else "unknown-%d" % k
        if rng.random() < 0.2:
            p = b'{"token":"t","conversation_id":' + mg.json_str_variants(rng, c) + b',"sequence":%d}' % k
        else:
            p = mg.canon(c, " tok", k, False, 5)
        payloads.append(p)
        subjects.append(b"chat." + c.encode("utf-8") + b".tokens")
    payloads += [b"{}x", b'{"token":5}']
    subjects += [None, None]
    return deltas, live, payloads, subjects


def check(hit, status, subs, wild_ids, live, subjects):
    """hit/status from the implementation under test; live = subscriptions after all deltas."""
    n_int = 0
    for i, subj in enumerate(subjects):
        if subj is None:
            assert int(status[i]) in (1, 2)
            continue
        want = nats_model.interested(live, subj)
        off, cnt, mask = (int(x) for x in hit[i])
        got = sorted(list(subs[off:off + cnt]) + [int(wild_ids[w]) for w in range(len(wild_ids)) if mask >> w & 1])
        assert got == want, (i, subj, got, want)
        assert (int(status[i]) & 0xF) == (0 if want else 5), (i, subj)
        n_int += bool(want)
    return n_int
