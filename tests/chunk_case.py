"""Seeded vLLM stream lines for the chunk decoder tests (row N3, decode half; reference src/llm-stream-proxy/main.go:188-226).
Hand vectors typed from the Go code and encoding/json's rules; a small Python model (json.loads + Go's typing rules) is a
third witness for the shapes it can express.  PARITY UNPINNED: the reference has no test for streamFromLLM."""
import ctypes as C
import json
import random

import numpy as np

import make_golden as mg

DONE = (b"[DONE]", 1)
# (line bytes, status, [(token bytes, done)])
HAND = [
    (b'data: {"id":"c1","object":"chat.completion.chunk","choices":[{"index":0,"delta":{"content":"Hello"},"logprobs":null,"finish_reason":null}]}\n',
     1, [(b"Hello", 0)]),
    (b'data: {"choices":[{"index":0,"delta":{"role":"assistant","content":""},"finish_reason":null}]}\r\n', 1, []),       # empty content: nothing published
    (b'data: {"choices":[{"index":0,"delta":{},"finish_reason":"stop"}]}\n', 1 | 0x100, [DONE]),
    (b'data: {"choices":[{"delta":{"content":"bye"},"finish_reason":"length"}]}\n', 1 | 0x100, [(b"bye", 0), DONE]),      # token first, then the sentinel
    (b"data: [DONE]\n", 2 | 0x100, [DONE]),
    (b"\xc2\xa0 \tdata: [DONE] \xe2\x80\xa8\r\n", 2 | 0x100, [DONE]),                                                     # TrimSpace is Unicode-aware
    (b"data:[DONE]\n", 0, []),                                                                                           # the prefix is "data: " with the blank
    (b"\n", 0, []), (b": keep-alive\n", 0, []), (b"event: x\n", 0, []), (b"data: \n", 0, []),
    (b"data: {not json}\n", 0, []),
    (b'data: {"choices":[{"delta":{"content":7}}]}\n', 0, []),                                                            # type error: Unmarshal fails, line skipped
    (b'data: {"choices":{"delta":{"content":"x"}}}\n', 0, []),
    (b'data: [1,2]\n', 0, []),
    (b'data: null\n', 1, []),
    (b'data: {"CHOICES":[{"Delta":{"CONTENT":"a\\u0062\\n"},"Fini\xc5\xbfh_rea\xc5\xbfon":null}]}\n', 1, [(b"ab\n", 0)]),   # folded keys (U+017F for s)
    (b'data: {"choices":[{"delta":{"content":"one"}},{"delta":{"content":"two"},"finish_reason":"stop"},{"delta":{"content":"never"}}]}\n',
     1 | 0x100, [(b"one", 0), (b"two", 0), DONE]),                                                                         # the loop returns at the first finish_reason
    (b'data: {"choices":[{"delta":{"content":"a"},"delta":{"x":1}},null,{"delta":null,"finish_reason":"stop","finish_reason":null}]}\n',
     1, [(b"a", 0)]),                                                                                                     # duplicate delta merges, null no-ops, null pointer wins
    (b'data: {"choices":[{"delta":{"content":"a"}},{"delta":{"content":"b"}}],"choices":[{"finish_reason":"stop"}]}\n',
     1 | 0x100, [(b"a", 0), DONE]),                                                                                        # second "choices" decodes INTO element 0, then cuts the slice to 1
    (b'data: {"choices":[{"delta":{"content":"\xff\\ud800x"}}]}\n', 1, [(b"\xef\xbf\xbd\xef\xbf\xbdx", 0)]),
    (b'data: {"choices":[' + b",".join([b'{"delta":{"content":"t"}}'] * 17) + b']}\n', 1 | 0x200, []),                      # product limit, told apart from a skip
]


class Obj(list):
    """a JSON object as its (key, value) pairs in input order (duplicates kept); a JSON array stays a plain list"""


def model_line(line):
    """Go semantics through CPython's json (no duplicate-`choices` stale-slice emulation: callers avoid those)."""
    s = line.decode("utf-8", "surrogateescape").strip(" \t\n\v\f\r\x85\xa0                　")
    if not s.startswith("data: "):
        return 0, []
    data = s[6:]
    if data == "[DONE]":
        return 2 | 0x100, [DONE]
    try:
        doc = json.loads(data, object_pairs_hook=Obj)
    except Exception:
        return 0, []

    def fold(k):
        return k.upper().replace("ſ", "S").replace("K", "K") if isinstance(k, str) else k

    if doc is None:
        return 1, []
    if not isinstance(doc, Obj):
        return 0, []
    err = [False]
    choices = None
    for k, v in doc:
        if fold(k) != "CHOICES":
            continue
        if v is None:
            choices = None
            continue
        if not isinstance(v, list) or isinstance(v, Obj):
            err[0] = True
            continue
        old = choices or []
        new = []
        for i, el in enumerate(v):
            c = dict(old[i]) if i < len(old) else {"content": "", "fin": False}
            if el is None:
                pass
            elif not isinstance(el, Obj):
                err[0] = True
            else:
                for k2, v2 in el:
                    if fold(k2) == "DELTA":
                        if v2 is None:
                            continue
                        if not isinstance(v2, Obj):
                            err[0] = True
                            continue
                        for k3, v3 in v2:
                            if fold(k3) == "CONTENT":
                                if isinstance(v3, str):
                                    c["content"] = v3
                                elif v3 is not None:
                                    err[0] = True
                    elif fold(k2) == "FINISH_REASON":
                        if isinstance(v2, str):
                            c["fin"] = True
                        elif v2 is None:
                            c["fin"] = False
                        else:
                            err[0] = True
            new.append(c)
        choices = new
    if err[0]:
        return 0, []
    st, msgs = 1, []
    for c in choices or []:
        if c["content"]:
            msgs.append((c["content"].encode("utf-8", "surrogatepass"), 0))
        if c["fin"]:
            msgs.append(DONE)
            st |= 0x100
            break
    return st, msgs


def rand_line(rng, k):
    r = rng.random()
    if r < 0.55:
        tok = mg.rand_text(rng, rng.randrange(0, 9), 0.2)
        fin = rng.random() < 0.05
        d = {"id": "cmpl-%d" % k, "object": "chat.completion.chunk", "created": 1700000000 + k, "model": "mistral",
             "choices": [{"index": 0, "delta": {"content": tok} if rng.random() < 0.9 else {"role": "assistant", "content": ""},
                          "logprobs": None, "finish_reason": "stop" if fin else None}]}
        body = json.dumps(d, ensure_ascii=rng.random() < 0.5, separators=(",", ":") if rng.random() < 0.7 else (", ", ": ")).encode("utf-8", "surrogatepass")
        return b"data: " + body + rng.choice([b"\n", b"\r\n", b" \n"])
    if r < 0.65:
        return rng.choice([b"\n", b"\r\n", b": ping\n", b"data: [DONE]\n", b"data: [DONE]\r\n", b"event: error\n", b"data:{}\n", b"  data: {}\n"])
    if r < 0.8:                                       # several choices, odd keys, nulls
        n = rng.randrange(0, 5)
        els = []
        for _ in range(n):
            parts = []
            if rng.random() < 0.8:
                c = rng.choice([b'"x"', b'""', b'null', b'"\\u00e9\\t"', b'"' + mg.rand_text(rng, 3, 0.3).encode("utf-8", "surrogatepass").replace(b"\\", b"\\\\").replace(b'"', b'\\"') + b'"'])
                parts.append(rng.choice([b'"delta"', b'"Delta"', b'"DELTA"']) + b":{" + rng.choice([b'"content"', b'"Content"']) + b":" + c + b"}")
            if rng.random() < 0.5:
                parts.append(rng.choice([b'"finish_reason"', b'"FINISH_REASON"']) + b":" + rng.choice([b"null", b'"stop"', b'"length"']))
            if rng.random() < 0.3:
                parts.append(b'"extra":[1,{"a":[]}]')
            rng.shuffle(parts)
            els.append(b"{" + b",".join(parts) + b"}" if rng.random() < 0.9 else b"null")
        return b'data: {"choices":[' + b",".join(els) + b']}\n'
    if r < 0.9:                                       # type errors and syntax errors
        return b"data: " + rng.choice([b'{"choices":[{"delta":{"content":1}}]}', b'{"choices":[{"delta":[]}]}', b'{"choices":[{"finish_reason":0}]}',
                                       b'{"choices":[3]}', b'{"choices":"x"}', b'"str"', b'7', b'{"choices":[{"delta":{"content":"x"}}]', b'{"choices":[}',
                                       b'{"choices":[{"delta":{"content":"a\x01"}}]}', b'{"a":01}']) + b"\n"
    return b"data: " + mg.structured(rng) + b"\n"      # TokenMessage-shaped objects: valid JSON without choices -> a chunk with no events


def pack(lines):
    off = np.zeros(len(lines) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in lines], dtype=np.uint64).astype(np.uint32)
    return np.frombuffer(b"".join(lines) + b"\0" * 64, dtype=np.uint8).copy(), off


def run_c(fn, buf, off):
    """fn = oracle's sseo_chunk_decode or the test double's hc_chunk_decode -> (line status list, [(line, token, done)])"""
    u8p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)
    fn.restype = C.c_uint32
    fn.argtypes = [u8p, u32p, C.c_uint32, u8p, u32p, u8p, u32p, u32p]
    n = len(off) - 1
    cap = int(off[n]) // 8 + n + 16
    tok = np.zeros(3 * int(off[n]) + 6 * n + 64, dtype=np.uint8)
    tok_off, done, msg_line, st = np.zeros(cap + 1, np.uint32), np.zeros(cap, np.uint8), np.zeros(cap, np.uint32), np.zeros(n, np.uint32)
    m = fn(buf.ctypes.data_as(u8p), off.ctypes.data_as(u32p), n, tok.ctypes.data_as(u8p), tok_off.ctypes.data_as(u32p), done.ctypes.data_as(u8p),
           msg_line.ctypes.data_as(u32p), st.ctypes.data_as(u32p))
    assert m != 0xFFFFFFFF
    raw = tok.tobytes()
    return st.tolist(), [(int(msg_line[k]), raw[int(tok_off[k]):int(tok_off[k + 1])], int(done[k])) for k in range(m)]
