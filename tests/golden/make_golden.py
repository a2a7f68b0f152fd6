#!/usr/bin/env python3
"""Generates tests/golden/adapter_vectors.json.gz from the pure-Python model (tests/pymodel.py).

Run from the repo root:  python tests/golden/make_golden.py
Deterministic (seeded).  The fixtures are what the C oracle (-m "not gpu") and the CUDA
path (-m gpu) are compared with; the hand-typed vectors live in hand_vectors.py.
No reference implementation can be run here (Go/Rust toolchains absent), so these are
outputs of the independent Python restatement, not of the reference: PARITY UNPINNED.
"""
import gzip
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import pymodel  # noqa: E402

WORDS = ("The quick brown fox jumps over the lazy dog.\nThis is a sample response from a large language model "
         "that simulates streaming tokens.\nEach word is sent as a separate token with timestamps for latency "
         "measurement.\nNATS provides excellent fan-out capabilities for distributing these tokens to edge "
         "locations.\nThe SSE adapter converts NATS messages into Server-Sent Events for browser clients.").split(" ")
assert len(WORDS) == 56

KEYS = ["conversation_id", "token", "sequence", "done", "timestamp"]


def canon(cid, tok, seq, done, ts):
    return pymodel.marshal({"cid": cid, "tok": tok, "seq": seq, "done": done, "ts": ts})


def rand_text(rng, n, spicy):
    out = []
    for _ in range(n):
        r = rng.random()
        if r < spicy * 0.3:
            out.append(rng.choice('"\\\n\r\t<>&\b\f\x00\x1f\x7f/'))
        elif r < spicy * 0.6:
            out.append(rng.choice(["é", "€", "😀", "\u2028", "\u2029", "�", "K", "ſ", "中"]))
        else:
            out.append(chr(rng.randrange(0x20, 0x7F)))
    return "".join(out)


def json_str_variants(rng, s):
    """A JSON string literal for s using a random mix of raw bytes and escapes."""
    out = bytearray(b'"')
    for ch in s:
        o = ord(ch)
        mode = rng.random()
        if ch in '"\\' or o < 0x20:
            short = {'"': b'\\"', "\\": b"\\\\", "\n": b"\\n", "\r": b"\\r", "\t": b"\\t", "\b": b"\\b", "\f": b"\\f"}
            if ch in short and mode < 0.6:
                out += short[ch]
            else:
                out += (b"\\u%04x" if mode < 0.8 else b"\\u%04X") % o
        elif mode < 0.12:
            if o < 0x10000:
                out += (b"\\u%04x" if rng.random() < 0.5 else b"\\u%04X") % o
            else:
                v = o - 0x10000
                out += b"\\u%04x\\u%04x" % (0xD800 + (v >> 10), 0xDC00 + (v & 0x3FF))
        elif ch == "/" and mode < 0.5:
            out += b"\\/"
        else:
            out += ch.encode("utf-8")
    out += b'"'
    return bytes(out)


def key_variant(rng, k):
    r = rng.random()
    if r < 0.6:
        return k
    if r < 0.75:
        return k.upper()
    if r < 0.85:
        return "".join(c.upper() if rng.random() < 0.5 else c for c in k)
    if r < 0.92:
        return k.replace("s", "ſ", 1).replace("k", "K", 1)
    if r < 0.96:
        return k + "x"
    return k[:-1]


def ws(rng):
    return rng.choice(["", "", "", " ", "  ", "\n", "\t", "\r\n "])


def rand_unknown_value(rng, depth=0):
    r = rng.random()
    if depth > 3 or r < 0.4:
        return rng.choice(["1", "-0", "1.5e+3", "true", "false", "null", '"x\\n"', '"\\ud800"', "0.0", "123456789012345678901234567890"])
    if r < 0.7:
        return "[" + ",".join(rand_unknown_value(rng, depth + 1) for _ in range(rng.randrange(0, 4))) + "]"
    return "{" + ",".join('"k%d":%s' % (i, rand_unknown_value(rng, depth + 1)) for i in range(rng.randrange(0, 3))) + "}"


def structured(rng):
    """A mostly-valid, non-canonical document."""
    fields = []
    for k in KEYS:
        if rng.random() < 0.85:
            fields.append(k)
    for _ in range(rng.randrange(0, 3)):
        fields.append(rng.choice(KEYS))          # duplicates
    for _ in range(rng.randrange(0, 3)):
        fields.append("?unknown")
    rng.shuffle(fields)
    parts = []
    for k in fields:
        if k == "?unknown":
            parts.append(ws(rng).encode() + json_str_variants(rng, rand_text(rng, rng.randrange(0, 6), 0.2)) + ws(rng).encode()
                         + b":" + ws(rng).encode() + rand_unknown_value(rng).encode())
            continue
        name = json_str_variants(rng, key_variant(rng, k)) if rng.random() < 0.3 else b'"' + key_variant(rng, k).encode() + b'"'
        r = rng.random()
        if r < 0.06:
            val = b"null"
        elif r < 0.12:
            val = rng.choice([b"1", b'"s"', b"true", b"[]", b"{}", b"1.0", b"1e2", b"-", b"9223372036854775808", b"-9223372036854775809"])
        elif k in ("conversation_id", "token"):
            val = json_str_variants(rng, rand_text(rng, rng.randrange(0, 20), rng.choice([0, 0.1, 0.5])))
        elif k in ("sequence", "timestamp"):
            val = str(rng.choice([0, 1, -1, rng.randrange(1, 2000), rng.randrange(-(1 << 63), 1 << 63), (1 << 63) - 1, -(1 << 63)])).encode()
            if rng.random() < 0.05:
                val = b"-0"
        else:
            val = rng.choice([b"true", b"false"])
        parts.append(ws(rng).encode() + name + ws(rng).encode() + b":" + ws(rng).encode() + val + ws(rng).encode())
    return ws(rng).encode() + b"{" + b",".join(parts) + b"}" + ws(rng).encode()


def mutate(rng, b):
    b = bytearray(b)
    for _ in range(rng.randrange(1, 4)):
        if not b:
            break
        r = rng.random()
        i = rng.randrange(len(b))
        if r < 0.35:
            b[i] = rng.randrange(256)
        elif r < 0.55:
            del b[i]
        elif r < 0.75:
            b.insert(i, rng.choice(b'"\\{}[],:\x00\xff\xc0\xe2\xf0\x80 u'))
        elif r < 0.9:
            del b[i:]
        else:
            b[i:i] = rng.choice([b"\xed\xa0\x80", b"\xf4\x90\x80\x80", b"\xe2\x82", b"\xc0\xaf", b"\xf0\x9f\x98", b"\\ud800", b"\\udfff\\ud800"])
    return bytes(b)


def build():
    rng = random.Random(0x5EED60)
    cases = []

    def add(name, payload, start_seq=0):
        cases.append((name, bytes(payload), start_seq))

    # config-0 shaped canonical records (demo/load-generator/main.go:215-221, fixed clock)
    T = 1700000000000000000
    for i in range(3):
        for j in list(range(0, 60)) + [510, 511]:
            add("loadgen_%d_%d" % (i, j), canon("loadtest-%d-%d" % (T, i), WORDS[j % 56], j + 1, j == 511, T + j * 50_000_000 + i))
    # canonical with every interesting single character in the token / id
    for o in list(range(0, 0x80)) + [0x80, 0xFF, 0x7FF, 0x800, 0xFFFD, 0xFFFF, 0x2027, 0x2028, 0x2029, 0x202A, 0x10000, 0x10FFFF, 0x017F, 0x212A]:
        add("canon_char_%04x" % o, canon("c" + chr(o), "a" + chr(o) + "b", 5, False, 17))
    # canonical with random content
    for k in range(200):
        add("canon_rand_%d" % k, canon(rand_text(rng, rng.randrange(0, 40), 0.1), rand_text(rng, rng.randrange(0, 24), rng.choice([0, 0.2, 0.6])),
                                     rng.choice([0, 1, 9, 10, 99, 100, 999, 1000, -1, rng.randrange(-(1 << 63), 1 << 63)]),
                                     rng.random() < 0.1, rng.choice([0, T, -T, (1 << 63) - 1, -(1 << 63)])))
    # every digit count of the two integers (the CUDA canonical path parses them word-wise)
    for k in range(1, 20):
        for v in (10 ** (k - 1), 10 ** k - 1, int("1234567890123456789"[:k]), int("9876543210987654321"[:k])):
            for sign in (1, -1):
                if abs(v) < (1 << 63):
                    add("digits_seq_%d_%d" % (k, sign * v), canon("c", "t", sign * v, False, 7))
                    add("digits_ts_%d_%d" % (k, sign * v), canon("c", "t", 7, False, sign * v))
    # near-canonical: one byte changed / inserted / removed
    base = canon("7f3c2b1a-1111-4222-8333-944455556666", " hello", 42, False, T)
    for i in range(len(base)):
        add("near_del_%d" % i, base[:i] + base[i + 1:])
    for i in range(0, len(base), 3):
        add("near_ins_sp_%d" % i, base[:i] + b" " + base[i:])
        add("near_flip_%d" % i, base[:i] + bytes([base[i] ^ 0x20]) + base[i + 1:])
    for tail in (b" ", b"\n", b"}", b"x", b",", b"\x00"):
        add("near_tail_%r" % tail, base + tail)
    # canonical shape but non-canonical escapes / numbers
    for tok in (b"\\/", b"\\u003C", b"\\u0041", b"\\b", b"\\f", b"\\u000a", b"\\u0022", b"\\u005c", b"\\ud83d\\ude00", b"\\ud800",
                b"\\udc00", b"\xff", b"\xc3", b"\xe2\x80\xa8", b"\\u2028", b"\\u2029", b"<", b">", b"&", b"\\u003c", b"\\u003e",
                b"\\u0026", b"\\u0000", b"\\u001f", b"\\u001F", b"\\u007f", b"\x7f", b"\\n\\r\\t\\\"\\\\"):
        add("esc_tok_%r" % tok, b'{"conversation_id":"c","token":"' + tok + b'","sequence":1,"done":false,"timestamp":2}')
        add("esc_cid_%r" % tok, b'{"conversation_id":"' + tok + b'","token":"t","sequence":1,"done":true,"timestamp":2}')
    for num in (b"-0", b"00", b"01", b"1.0", b"1e0", b"1E5", b"-", b"+1", b"0x1", b"9223372036854775807", b"9223372036854775808",
                b"-9223372036854775808", b"-9223372036854775809", b"99999999999999999999", b"1.", b".5", b"1e", b"1e+", b"0.0e-0"):
        add("num_seq_%r" % num, b'{"conversation_id":"c","token":"t","sequence":' + num + b',"done":false,"timestamp":2}')
        add("num_ts_%r" % num, b'{"conversation_id":"c","token":"t","sequence":1,"done":false,"timestamp":' + num + b'}')
    for lit in (b"true", b"false", b"null", b"True", b"tru", b"truee", b"nul", b"1", b'"true"', b"[]"):
        add("lit_done_%r" % lit, b'{"conversation_id":"c","token":"t","sequence":1,"done":' + lit + b',"timestamp":2}')
    # top-level kinds, nesting depth
    for doc in (b"null", b" null ", b"nul", b"nulll", b"true", b"0", b'""', b"[]", b"{}", b"{} ", b" {}", b"{}{}", b"[", b"]", b"{",
                b"}", b'{"a"}', b'{"a":}', b'{"a":1,}', b"{,}", b'{"a":1 "b":2}', b"[1,]", b"[,1]", b'{"a":[}', b'{"a":{]}',
                b'{1:2}', b"{'a':1}", b'{"a":1}//', b"\xef\xbb\xbf{}", b"\x00", b'{"token":"abc', b'{"token":"abc\\', b'{"token":"\\u12"}'):
        add("doc_%r" % doc, doc)
    for depth in (1, 2, 100, 9998, 9999, 10000):
        # the top-level object is level 1, so the deepest level is depth+1
        add("depth_arr_%d" % depth, b'{"x":' + b"[" * depth + b"]" * depth + b',"sequence":3}')
        add("depth_obj_%d" % depth, b'{"x":' + b'{"a":' * depth + b"1" + b"}" * depth + b',"sequence":3}')
    add("depth_top_arr_10000", b"[" * 10000 + b"]" * 10000)
    add("depth_top_arr_10001", b"[" * 10001 + b"]" * 10001)
    add("depth_mixed_mismatch", b'{"x":' + b"[{\"a\":" * 50 + b"1" + b"]}" * 50 + b"}")
    # long strings (cross tile / smem limits on the GPU side)
    for n in (255, 256, 257, 1000, 4096, 20000, 70000):
        add("long_tok_%d" % n, canon("c", "x" * n, 1, False, 2))
        add("long_tok_esc_%d" % n, canon("c", "<\n" * (n // 2), 1, False, 2))
        add("long_cid_%d" % n, canon("i" * n, "t", 1, True, 2))
    add("long_unknown", b'{"pad":"' + b"p" * 50000 + b'","token":"t"}')
    add("long_ws", b" " * 30000 + b'{"token":"t"}' + b"\n" * 30000)
    # sequence filter (sse_handler.go:93-96, :142-145)
    for start, seq in ((4, 3), (4, 4), (4, 5), (1, 0), (1, 1), (1, 2), (0, -5), (-7, -9), ((1 << 63) - 1, (1 << 63) - 1), (5, -(1 << 63))):
        add("filter_%d_%d" % (start, seq), canon("c", "t", seq, False, 0), start)
    add("filter_on_noncanonical", b'{"sequence": 2}', 3)
    add("filter_on_bad_json", b'{"sequence": 2', 3)
    # structured and mutated fuzz
    for k in range(700):
        add("struct_%d" % k, structured(rng))
    for k in range(500):
        src = rng.choice([base, structured(rng), canon(rand_text(rng, 8, 0.3), rand_text(rng, 8, 0.5), k, False, T)])
        add("mut_%d" % k, mutate(rng, src))
    return cases


def main():
    cases = build()
    out = []
    for name, payload, start in cases:
        st, fr = pymodel.adapter(payload, start)
        bst, subj, js = pymodel.bridge(payload)
        st122, fr122 = pymodel.adapter(payload, start, go122=True)
        rec = {"name": name, "in": payload.hex(), "start": start, "status": st, "frame": fr.hex(),
               "bstatus": bst, "subject": subj.hex(), "json": js.hex()}
        if fr122 != fr:
            rec["frame122"] = fr122.hex()
        out.append(rec)
    path = os.path.join(HERE, "adapter_vectors.json.gz")
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        f.write(json.dumps(out, separators=(",", ":")).encode())
    print("wrote %d cases -> %s (%d bytes)" % (len(out), path, os.path.getsize(path)))


if __name__ == "__main__":
    main()
