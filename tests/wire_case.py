"""Seeded socket-byte scenarios for the wire splitter tests (row N4): per connection a stream of NATS server->client
protocol lines or RESP pushes, cut at a random point (the tail stays in the socket buffer) or damaged (protocol violation).
Shared by the CPU and GPU tests.  Hand vectors are typed from the protocol specifications (NATS client protocol, RESP2/3);
PARITY UNPINNED: nats.go / go-redis are not vendored in the reference and it has no test for this."""
import ctypes as C
import random

import numpy as np

import make_golden as mg

# (protocol, socket bytes, expected deliveries [(subject/channel, sid, payload)], consumed, status, pings).  `consumed` is written
# as the length of the part of the stream that forms complete protocol messages, so that nobody has to count bytes by hand.
def _v(proto, complete, rest, msgs, status=0, pings=0):
    return (proto, complete + rest, msgs, len(complete), status, pings)


HAND = [
    _v(0, b"MSG chat.c1.tokens 7 5\r\nhello\r\n", b"", [(b"chat.c1.tokens", 7, b"hello")]),
    # reply-to field, a PING between messages, lower-case verb, tab / runs of blanks as separators, empty payload
    _v(0, b"MSG chat.c1.tokens 7 _INBOX.1 5\r\nhello\r\nPING\r\nmsg\ta.b  12   0\r\n\r\n", b"", [(b"chat.c1.tokens", 7, b"hello"), (b"a.b", 12, b"")], 0, 1),
    # control lines are skipped, -ERR is noted; the last message's payload has not arrived yet: it stays in the buffer
    _v(0, b"INFO {\"server_id\":\"x\"}\r\n+OK\r\nPONG\r\n-ERR 'Unknown Protocol Operation'\r\n", b"MSG s 1 3\r\nab", [], 2, 0),
    # HMSG: 12 header bytes of 15 total -> the payload is the last 3
    _v(0, b"HMSG s 9 12 15\r\nNATS/1.0\r\n\r\nabc\r\n", b"", [(b"s", 9, b"abc")]),
    # after the payload everything up to the closing LF is skipped (nats.go MSG_END)
    _v(0, b"MSG s 1 3\r\nabcXYZ\nPING\r\n", b"", [(b"s", 1, b"abc")], 0, 1),
    _v(0, b"", b"MSG s x 3\r\nabc\r\n", [], 1),                               # sid is not a number
    _v(0, b"PING\r\n", b"BOGUS\r\nPING\r\n", [], 1, 1),                      # unknown verb: stop there
    _v(0, b"", b"MSG s 1 2 3 4 5\r\n", [], 1),                                   # too many arguments
    _v(0, b"", b"MSG s 1 3", [], 0),                                               # control line itself incomplete
    _v(1, b"*3\r\n$7\r\nmessage\r\n$14\r\nchat.c1.tokens\r\n$5\r\nhello\r\n", b"", [(b"chat.c1.tokens", 0, b"hello")]),
    # pmessage carries the pattern in front of the channel; subscribe acknowledgements and simple strings are skipped
    _v(1, b"*4\r\n$8\r\npmessage\r\n$13\r\nchat.*.tokens\r\n$14\r\nchat.c1.tokens\r\n$2\r\n{}\r\n*3\r\n$10\r\npsubscribe\r\n$1\r\nx\r\n:1\r\n+PONG\r\n", b"",
       [(b"chat.c1.tokens", 0, b"{}")]),
    # RESP3 push type; empty payload; the second push is incomplete and stays
    _v(1, b">3\r\n$7\r\nmessage\r\n$1\r\nc\r\n$0\r\n\r\n", b"*3\r\n$7\r\nmessage\r\n$1\r\nc\r\n$5\r\nhel", [(b"c", 0, b"")]),
    _v(1, b"*3\r\n$7\r\nMessage\r\n$1\r\nc\r\n$1\r\nx\r\n", b"", []),                       # go-redis matches the kind exactly
    _v(1, b"*2\r\n*1\r\n$1\r\na\r\n$-1\r\n_\r\n#t\r\n,1.5\r\n", b"", []),            # nesting, null bulk, RESP3 scalars
    _v(1, b"", b"*3\n$7\r\nmessage\r\n", [], 1),                              # a header line that ends in LF without CR
    _v(1, b"", b"*3\r\n$7\r\nmessage\n", [], 0),                                # the bulk's CRLF has not arrived: incomplete, not an error
    _v(1, b"", b"?what\r\n", [], 1),
    _v(1, b"", b"$3\r\nabcd\r\n", [], 1),                                        # bulk longer than announced
]


def nats_item(rng, k):
    r = rng.random()
    if r < 0.75:
        cid = "conv-%d" % rng.randrange(1000)
        payload = mg.canon(cid, mg.rand_text(rng, rng.randrange(0, 12), 0.2), k, rng.random() < 0.05, 1700000000000000000 + k) if rng.random() < 0.9 \
            else bytes(rng.randrange(256) for _ in range(rng.randrange(0, 300)))
        subj = b"chat." + cid.encode() + b".tokens"
        sid = rng.randrange(1, 100000)
        if rng.random() < 0.1:
            hdr = b"NATS/1.0\r\nK: v\r\n\r\n"
            line = b"HMSG %s %d %s%d %d\r\n" % (subj, sid, b"_INBOX.x " if rng.random() < 0.3 else b"", len(hdr), len(hdr) + len(payload))
            return line + hdr + payload + b"\r\n", (subj, sid, payload)
        sep = rng.choice([b" ", b"  ", b"\t"])
        verb = rng.choice([b"MSG", b"MSG", b"msg", b"Msg"])
        reply = (b"_INBOX.%d" % k + sep) if rng.random() < 0.2 else b""
        return verb + sep + subj + sep + b"%d" % sid + sep + reply + b"%d" % len(payload) + rng.choice([b"\r\n", b"\n"]) + payload + b"\r\n", (subj, sid, payload)
    return rng.choice([b"PING\r\n", b"PONG\r\n", b"+OK\r\n", b"-ERR 'Slow Consumer'\r\n", b"INFO {\"a\":1}\r\n"]), None


def bulk(b):
    return b"$%d\r\n%s\r\n" % (len(b), b)


def resp_item(rng, k):
    r = rng.random()
    if r < 0.8:
        cid = "conv-%d" % rng.randrange(1000)
        payload = mg.canon(cid, mg.rand_text(rng, rng.randrange(0, 12), 0.2), k, False, 5) if rng.random() < 0.9 \
            else bytes(rng.randrange(256) for _ in range(rng.randrange(0, 200)))
        chan = b"llm:tokens:" + cid.encode()
        if rng.random() < 0.5:
            return rng.choice([b"*4\r\n", b">4\r\n"]) + bulk(b"pmessage") + bulk(b"llm:tokens:*") + bulk(chan) + bulk(payload), (chan, 0, payload)
        return rng.choice([b"*3\r\n", b">3\r\n"]) + bulk(b"message") + bulk(chan) + bulk(payload), (chan, 0, payload)
    return rng.choice([b"*3\r\n" + bulk(b"psubscribe") + bulk(b"llm:tokens:*") + b":1\r\n", b"+PONG\r\n", b"*2\r\n" + bulk(b"pong") + bulk(b""),
                       b":42\r\n", b"$-1\r\n", b"*0\r\n", b"*2\r\n*2\r\n:1\r\n:2\r\n$1\r\nz\r\n"]), None


def build(proto, n_conn, seed):
    """-> (buf u8, buf_off u32[n_conn+1], expected deliveries per connection, expected full-parse flags)"""
    rng = random.Random(seed)
    chunks, want = [], []
    for c in range(n_conn):
        stream, msgs = b"", []
        for k in range(rng.randrange(0, 9)):
            item, m = (resp_item if proto else nats_item)(rng, k)
            stream += item
            if m:
                msgs.append(m)
        mode = rng.random()
        if mode < 0.25 and stream:
            stream = stream[: rng.randrange(len(stream))]            # cut anywhere: whatever is complete must still come out
            msgs = None                                               # (the oracle says which ones)
        elif mode < 0.32:
            stream += rng.choice([b"BOGUS 1 2\r\n", b"MSG a b c\r\n"] if proto == 0 else [b"!x\r\n", b"*x\r\n", b"$2\r\nabc\r\n"])
            msgs = None
        chunks.append(stream)
        want.append(msgs)
    off = np.zeros(n_conn + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in chunks], dtype=np.uint64).astype(np.uint32)
    buf = np.frombuffer(b"".join(chunks) + b"\0" * 64, dtype=np.uint8).copy()
    return buf, off, want


class OMsg(C.Structure):
    _fields_ = [("payload_off", C.c_uint32), ("payload_len", C.c_uint32), ("subj_off", C.c_uint32), ("subj_len", C.c_uint32),
                ("sid", C.c_uint64), ("conn", C.c_uint32), ("pad", C.c_uint32)]


class OConn(C.Structure):
    _fields_ = [("n_msgs", C.c_uint32), ("payload_bytes", C.c_uint32), ("consumed", C.c_uint32), ("status", C.c_uint32), ("pings", C.c_uint32)]


def oracle_split(lib, proto, buf, off):
    """oracle/sse_wire_oracle.c -> ([(conn, subject bytes, sid, payload bytes)], [(n_msgs, consumed, status, pings)])"""
    u8p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)
    lib.sseo_wire_split.restype = C.c_uint32
    lib.sseo_wire_split.argtypes = [C.c_uint32, u8p, u32p, C.c_uint32, C.POINTER(C.POINTER(OMsg)), u32p, C.POINTER(OConn)]
    lib.sseo_free.argtypes = [C.c_void_p]
    n_conn = len(off) - 1
    conns = (OConn * max(n_conn, 1))()
    pm = C.POINTER(OMsg)()
    n = lib.sseo_wire_split(proto, buf.ctypes.data_as(u8p), off.ctypes.data_as(u32p), n_conn, C.byref(pm), None, conns)
    raw = buf.tobytes()
    msgs = [(pm[i].conn, raw[pm[i].subj_off:pm[i].subj_off + pm[i].subj_len], int(pm[i].sid), raw[pm[i].payload_off:pm[i].payload_off + pm[i].payload_len])
            for i in range(n)]
    lib.sseo_free(pm)
    return msgs, [(conns[c].n_msgs, conns[c].consumed, conns[c].status, conns[c].pings) for c in range(n_conn)]
