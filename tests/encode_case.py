"""Producer-side encode cases shared by the CPU and GPU tests: messages as (cid bytes, token bytes, seq, done, ts),
an independent Python model of json.Marshal(TokenMessage) over RAW Go strings (invalid UTF-8 included), and the
hand-written known answers.  No reference test covers publishToken (llm-stream-proxy/main.go:230-245), so these
expectations are written from the Go language spec / encoding/json source, like tests/golden/hand_vectors.py."""
import random

import numpy as np

import pymodel

# (cid, token, seq, done, ts) -> expected json.Marshal output, written out by hand
HAND = [
    ((b"abc", b" Hello", 1, False, 1700000000000000000),
     b'{"conversation_id":"abc","token":" Hello","sequence":1,"done":false,"timestamp":1700000000000000000}'),
    ((b"", b"", 0, True, 0), b'{"conversation_id":"","token":"","sequence":0,"done":true,"timestamp":0}'),
    ((b"c", b'"q"\\', -5, False, -1), b'{"conversation_id":"c","token":"\\"q\\"\\\\","sequence":-5,"done":false,"timestamp":-1}'),
    ((b"c", b"<b>&</b>", 2, False, 3), b'{"conversation_id":"c","token":"\\u003cb\\u003e\\u0026\\u003c/b\\u003e","sequence":2,"done":false,"timestamp":3}'),
    ((b"c", b"\n\r\t\x00\x1f\x7f", 2, False, 3), b'{"conversation_id":"c","token":"\\n\\r\\t\\u0000\\u001f\x7f","sequence":2,"done":false,"timestamp":3}'),
    ((b"c", b"\x08\x0c", 2, False, 3), b'{"conversation_id":"c","token":"\\u0008\\u000c","sequence":2,"done":false,"timestamp":3}'),
    ((b"c", "café 世界 \U0001f600".encode(), 9, False, 1), b'{"conversation_id":"c","token":"' + "café 世界 \U0001f600".encode() + b'","sequence":9,"done":false,"timestamp":1}'),
    ((b"c", "a b c".encode(), 9, False, 1), b'{"conversation_id":"c","token":"a\\u2028b\\u2029c","sequence":9,"done":false,"timestamp":1}'),
    # invalid UTF-8: a lone continuation byte, a truncated 3-byte sequence (two bytes, each replaced), an overlong, a surrogate
    ((b"c", b"\x80", 1, False, 1), b'{"conversation_id":"c","token":"\\ufffd","sequence":1,"done":false,"timestamp":1}'),
    ((b"c", b"x\xe4\xb8", 1, False, 1), b'{"conversation_id":"c","token":"x\\ufffd\\ufffd","sequence":1,"done":false,"timestamp":1}'),
    ((b"c", b"\xc0\xaf", 1, False, 1), b'{"conversation_id":"c","token":"\\ufffd\\ufffd","sequence":1,"done":false,"timestamp":1}'),
    ((b"c", b"\xed\xa0\x80", 1, False, 1), b'{"conversation_id":"c","token":"\\ufffd\\ufffd\\ufffd","sequence":1,"done":false,"timestamp":1}'),
    # a correctly encoded U+FFFD stays as it is
    ((b"c", b"\xef\xbf\xbd", 1, False, 1), b'{"conversation_id":"c","token":"\xef\xbf\xbd","sequence":1,"done":false,"timestamp":1}'),
    ((b"i\xffd", b"t", 9223372036854775807, True, -9223372036854775808),
     b'{"conversation_id":"i\\ufffdd","token":"t","sequence":9223372036854775807,"done":true,"timestamp":-9223372036854775808}'),
]
HAND_GO122 = [
    ((b"c", b"\x08\x0c\x0b", 2, False, 3), b'{"conversation_id":"c","token":"\\b\\f\\u000b","sequence":2,"done":false,"timestamp":3}'),
]
# channel = "chat." + raw id + ".tokens" (main.go:240): raw bytes, no escaping
HAND_CHANNEL = [
    ((b"abc", b"t", 1, False, 2), b"chat.abc.tokens", b'{"conversation_id":"abc","token":"t","sequence":1,"done":false,"timestamp":2}'),
    ((b'a"b', b"t", 1, False, 2), b'chat.a"b.tokens', b'{"conversation_id":"a\\"b","token":"t","sequence":1,"done":false,"timestamp":2}'),
    ((b"", b"t", 1, False, 2), b"chat..tokens", b'{"conversation_id":"","token":"t","sequence":1,"done":false,"timestamp":2}'),
]


def go_decode(raw):
    """RAW Go string bytes -> list of (char, was_invalid_byte) the way `for _, r := range s` / utf8.DecodeRune walks it."""
    out, i = [], 0
    while i < len(raw):
        ch, k = pymodel.decode_one_rune(raw, i)
        invalid = k == 1 and raw[i] >= 0x80
        out.append((ch, invalid))
        i += k
    return out


def model_string(raw, go122=False):
    out = bytearray(b'"')
    for ch, invalid in go_decode(raw):
        out += b"\\ufffd" if invalid else pymodel.go_string(ch, go122)[1:-1]
    return bytes(out + b'"')


def model_encode(msg, go122=False):
    cid, tok, seq, done, ts = msg
    return (b'{"conversation_id":' + model_string(cid, go122) + b',"token":' + model_string(tok, go122) + b',"sequence":%d' % seq
            + b',"done":' + (b"true" if done else b"false") + b',"timestamp":%d}' % ts)


def model_channel(msg):
    return b"chat." + msg[0] + b".tokens"


_ALPHABET = [b"a", b"Z", b" ", b"the", b" quick", b'"', b"\\", b"<", b">", b"&", b"\n", b"\t", b"\r", b"\x00", b"\x08", b"\x0c", b"\x1f",
             b"\x7f", "é".encode(), "世".encode(), "\U0001f600".encode(), " ".encode(), " ".encode(), "�".encode(),
             b"\x80", b"\xbf", b"\xc0", b"\xc1", b"\xc2", b"\xe0\x80", b"\xe4\xb8", b"\xed\xa0\x80", b"\xf0\x9f", b"\xf4\x90\x80\x80",
             b"\xf5", b"\xff", b"\xe2\x80", b"\xe2\x80\xa7", b"\xe2\x80\xaa"]
_INTS = [0, 1, -1, 9, 10, 99, 100, 12345678, 123456789, 2**31 - 1, 2**31, -2**31, 2**53, 2**63 - 1, -2**63, 1700000000123456789]


def random_messages(n, seed=5):
    """a mix of ordinary LLM-like messages and hostile strings of every length 0..40."""
    rng = random.Random(seed)
    msgs = []
    for i in range(n):
        if rng.random() < 0.5:
            cid = b"conv-%d" % rng.randrange(100000)
            tok = rng.choice([b" the", b" of", b"Hello", b",", b" world", b".", b"\n", b" caf\xc3\xa9", b""])
            msgs.append((cid, tok, i + 1, rng.random() < 0.02, 1700000000000000000 + i * 1000003))
        else:
            cid = b"".join(rng.choice(_ALPHABET) for _ in range(rng.randrange(0, 6)))
            tok = b"".join(rng.choice(_ALPHABET) for _ in range(rng.randrange(0, 24)))
            msgs.append((cid, tok, rng.choice(_INTS), rng.random() < 0.5, rng.choice(_INTS)))
    return msgs


def columns(msgs):
    """list of messages -> (cid arena, cid_off, tok arena, tok_off, seq, done, ts) numpy columns (arenas padded)."""
    n = len(msgs)
    cid_off = np.zeros(n + 1, dtype=np.uint32)
    tok_off = np.zeros(n + 1, dtype=np.uint32)
    if n:
        cid_off[1:] = np.cumsum([len(m[0]) for m in msgs])
        tok_off[1:] = np.cumsum([len(m[1]) for m in msgs])
    cid = np.frombuffer(b"".join(m[0] for m in msgs) + b"\0" * 32, dtype=np.uint8).copy()
    tok = np.frombuffer(b"".join(m[1] for m in msgs) + b"\0" * 32, dtype=np.uint8).copy()
    seq = np.array([m[2] for m in msgs], dtype=np.int64)
    done = np.array([1 if m[3] else 0 for m in msgs], dtype=np.uint8)
    ts = np.array([m[4] for m in msgs], dtype=np.int64)
    return cid, cid_off, tok, tok_off, seq, done, ts


# SSEGPU_OP_RESP: the RESP2 command go-redis writes for rdb.Publish(ctx, "chat.<id>.tokens", payload), typed out by hand
HAND_RESP = [
    ((b"abc", b"t", 1, False, 2),
     b'*3\r\n$7\r\npublish\r\n$15\r\nchat.abc.tokens\r\n$77\r\n{"conversation_id":"abc","token":"t","sequence":1,"done":false,"timestamp":2}\r\n'),
    ((b"", b"", 0, True, 0), b'*3\r\n$7\r\npublish\r\n$12\r\nchat..tokens\r\n$72\r\n{"conversation_id":"","token":"","sequence":0,"done":true,"timestamp":0}\r\n'),
    ((b"a\r\nb", b"<", 10, False, 5),
     b'*3\r\n$7\r\npublish\r\n$16\r\nchat.a\r\nb.tokens\r\n$86\r\n{"conversation_id":"a\\r\\nb","token":"\\u003c","sequence":10,"done":false,"timestamp":5}\r\n'),
]


def model_resp(msg, go122=False):
    chan, payload = model_channel(msg), model_encode(msg, go122)
    return b"*3\r\n$7\r\npublish\r\n$%d\r\n%s\r\n$%d\r\n%s\r\n" % (len(chan), chan, len(payload), payload)
