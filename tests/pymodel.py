"""Second, independently written restatement of the reference hot path, in pure Python.

Test infrastructure only.  It exists so that the C oracle (oracle/sse_oracle.c) is not
the only statement of the semantics: the two are written in different styles (this one
builds a parse tree by recursive descent and leans on Python's strict UTF-8 codec; the C
one is Go-shaped: validate, walk, unquote, append) and tests/golden/make_golden.py uses
THIS model to generate the committed fixtures the C oracle and the CUDA path are checked
against.  PARITY UNPINNED against real Go (no toolchain here) -- see oracle/sse_oracle.h.

Reference call sites restated:
  decode   src/sse-adapter/sse_handler.go:136-140, :317-321; src/redis-nats-bridge/main.go:153-157
  filter   src/sse-adapter/sse_handler.go:93-96, :142-145
  encode   src/sse-adapter/sse_handler.go:211, :427; src/redis-nats-bridge/main.go:161
  frame    src/sse-adapter/sse_handler.go:212-214, :428-430
  subject  src/redis-nats-bridge/main.go:160
"""
import sys

OK, ERR_SYNTAX, ERR_TYPE, FILTERED, INSPECT_DROP = 0, 1, 2, 3, 4
DONE = 0x100
MAX_DEPTH = 10000
WS = b" \t\r\n"

sys.setrecursionlimit(max(sys.getrecursionlimit(), 60000))


class SyntaxErr(Exception):
    pass


class _P:
    """Recursive-descent JSON parser over bytes; keeps strings/numbers as raw literals."""

    def __init__(self, b):
        self.b = b
        self.i = 0

    def ws(self):
        b, i = self.b, self.i
        while i < len(b) and b[i] in WS:
            i += 1
        self.i = i

    def peek(self):
        if self.i >= len(self.b):
            raise SyntaxErr("eof")
        return self.b[self.i]

    def value(self, depth):
        self.ws()
        c = self.peek()
        if c == 0x7B:  # {
            if depth + 1 > MAX_DEPTH:
                raise SyntaxErr("depth")
            self.i += 1
            items = []
            self.ws()
            if self.peek() == 0x7D:
                self.i += 1
                return ("obj", items)
            while True:
                self.ws()
                if self.peek() != 0x22:
                    raise SyntaxErr("key")
                k = self.string()
                self.ws()
                if self.peek() != 0x3A:
                    raise SyntaxErr("colon")
                self.i += 1
                v = self.value(depth + 1)
                items.append((k, v))
                self.ws()
                c = self.peek()
                self.i += 1
                if c == 0x2C:
                    continue
                if c == 0x7D:
                    return ("obj", items)
                raise SyntaxErr("obj sep")
        if c == 0x5B:  # [
            if depth + 1 > MAX_DEPTH:
                raise SyntaxErr("depth")
            self.i += 1
            items = []
            self.ws()
            if self.peek() == 0x5D:
                self.i += 1
                return ("arr", items)
            while True:
                items.append(self.value(depth + 1))
                self.ws()
                c = self.peek()
                self.i += 1
                if c == 0x2C:
                    continue
                if c == 0x5D:
                    return ("arr", items)
                raise SyntaxErr("arr sep")
        if c == 0x22:
            return ("str", self.string())
        if c == 0x2D or 0x30 <= c <= 0x39:
            return ("num", self.number())
        for lit, node in ((b"true", ("bool", True)), (b"false", ("bool", False)), (b"null", ("null", None))):
            if c == lit[0]:
                if self.b[self.i:self.i + len(lit)] != lit:
                    raise SyntaxErr("literal")
                self.i += len(lit)
                return node
        raise SyntaxErr("value")

    def string(self):
        b = self.b
        i = self.i + 1
        start = i
        while True:
            if i >= len(b):
                raise SyntaxErr("eof in string")
            c = b[i]
            if c == 0x22:
                self.i = i + 1
                return bytes(b[start:i])
            if c < 0x20:
                raise SyntaxErr("ctl in string")
            if c == 0x5C:
                i += 1
                if i >= len(b):
                    raise SyntaxErr("eof in esc")
                e = b[i]
                if e in b'"\\/bfnrt':
                    pass
                elif e == 0x75:
                    h = b[i + 1:i + 5]
                    if len(h) != 4 or any(x not in b"0123456789abcdefABCDEF" for x in h):
                        raise SyntaxErr("bad \\u")
                    i += 4
                else:
                    raise SyntaxErr("bad esc")
            i += 1

    def number(self):
        b, i = self.b, self.i
        s = i
        n = len(b)

        def digits(j):
            k = j
            while k < n and 0x30 <= b[k] <= 0x39:
                k += 1
            if k == j:
                raise SyntaxErr("digits")
            return k
        if b[i] == 0x2D:
            i += 1
            if i >= n:
                raise SyntaxErr("lone minus")
        if b[i] == 0x30:
            i += 1
        elif 0x31 <= b[i] <= 0x39:
            i = digits(i)
        else:
            raise SyntaxErr("num")
        if i < n and b[i] == 0x2E:
            i = digits(i + 1)
        if i < n and b[i] in b"eE":
            i += 1
            if i < n and b[i] in b"+-":
                i += 1
            i = digits(i)
        self.i = i
        return bytes(b[s:i])


def parse(b):
    p = _P(b)
    v = p.value(0)
    p.ws()
    if p.i != len(b):
        raise SyntaxErr("trailing")
    return v


def decode_one_rune(b, i):
    """Go utf8.DecodeRune: (char, size) or (U+FFFD, 1) -- via Python's strict codec."""
    for k in (1, 2, 3, 4):
        try:
            s = b[i:i + k].decode("utf-8")
        except UnicodeDecodeError:
            continue
        if len(s) == 1 and len(b[i:i + k]) == k:
            return s, k
    return "\ufffd", 1


def unquote(raw):
    """Go unquoteBytes on the bytes between the quotes -> str of Unicode scalars."""
    out = []
    i, n = 0, len(raw)
    simple = {0x22: '"', 0x5C: "\\", 0x2F: "/", 0x62: "\b", 0x66: "\f", 0x6E: "\n", 0x72: "\r", 0x74: "\t"}

    def u4(j):
        if raw[j:j + 2] != b"\\u" or j + 6 > n:
            return -1
        return int(raw[j + 2:j + 6], 16)
    while i < n:
        c = raw[i]
        if c == 0x5C:
            e = raw[i + 1]
            if e == 0x75:
                r = u4(i)
                i += 6
                if 0xD800 <= r < 0xE000:
                    r2 = u4(i)
                    if r < 0xDC00 and 0xDC00 <= r2 < 0xE000:
                        out.append(chr(0x10000 + ((r - 0xD800) << 10) + (r2 - 0xDC00)))
                        i += 6
                        continue
                    r = 0xFFFD
                out.append(chr(r))
            else:
                out.append(simple[e])
                i += 2
        elif c < 0x80:
            out.append(chr(c))
            i += 1
        else:
            ch, k = decode_one_rune(raw, i)
            out.append(ch)
            i += k
    return "".join(out)


def fold_key(s):
    """Go foldName: canonical representative under simple case folding (ASCII-relevant part)."""
    out = []
    for ch in s:
        if "a" <= ch <= "z":
            ch = ch.upper()
        elif ch == "\u212a":  # KELVIN SIGN
            ch = "K"
        elif ch == "\u017f":  # LATIN SMALL LETTER LONG S
            ch = "S"
        out.append(ch)
    return "".join(out)


FIELDS = {"CONVERSATION_ID": "cid", "TOKEN": "tok", "SEQUENCE": "seq", "DONE": "done", "TIMESTAMP": "ts"}


def unmarshal(b):
    """-> (status, msg dict).  msg strings are Python str of scalars."""
    msg = {"cid": "", "tok": "", "seq": 0, "done": False, "ts": 0}
    try:
        tree = parse(bytes(b))
    except (SyntaxErr, IndexError):
        return ERR_SYNTAX, msg
    kind, val = tree
    if kind == "null":
        return OK, msg
    if kind != "obj":
        return ERR_TYPE, msg
    bad = False
    for kraw, (vk, vv) in val:
        f = FIELDS.get(fold_key(unquote(kraw)))
        if f is None or vk == "null":
            continue
        if f in ("cid", "tok"):
            if vk == "str":
                msg[f] = unquote(vv)
            else:
                bad = True
        elif f in ("seq", "ts"):
            ok = False
            if vk == "num":
                t = vv.decode()
                body = t[1:] if t.startswith("-") else t
                if body.isdigit():
                    x = int(t)
                    if -(1 << 63) <= x < (1 << 63):
                        msg[f] = x
                        ok = True
            bad = bad or not ok
        else:
            if vk == "bool":
                msg["done"] = vv
            else:
                bad = True
    return (ERR_TYPE if bad else OK), msg


def go_string(s, go122=False):
    """Go appendString with escapeHTML=true on a str of scalars -> bytes."""
    out = bytearray(b'"')
    for ch in s:
        o = ord(ch)
        if ch in '"\\':
            out += b"\\" + ch.encode()
        elif ch == "\n":
            out += b"\\n"
        elif ch == "\r":
            out += b"\\r"
        elif ch == "\t":
            out += b"\\t"
        elif go122 and ch == "\b":
            out += b"\\b"
        elif go122 and ch == "\f":
            out += b"\\f"
        elif o < 0x20 or ch in "<>&":
            out += b"\\u%04x" % o
        elif o in (0x2028, 0x2029):
            out += b"\\u%04x" % o
        else:
            out += ch.encode("utf-8")
    out += b'"'
    return bytes(out)


def marshal(m, go122=False):
    return (b'{"conversation_id":' + go_string(m["cid"], go122) + b',"token":' + go_string(m["tok"], go122)
            + b',"sequence":' + str(m["seq"]).encode() + b',"done":' + (b"true" if m["done"] else b"false")
            + b',"timestamp":' + str(m["ts"]).encode() + b"}")


def frame(m, go122=False):
    return b"event: token\nid: " + str(m["seq"]).encode() + b"\ndata: " + marshal(m, go122) + b"\n\n"


def adapter(payload, start_seq=0, go122=False):
    """-> (status word, frame bytes)."""
    st, m = unmarshal(payload)
    if st == OK and start_seq > 0 and m["seq"] <= start_seq:
        st = FILTERED
    if st != OK:
        return st, b""
    return (OK | (DONE if m["done"] else 0)), frame(m, go122)


def bridge(payload, go122=False):
    """-> (status word, subject bytes, json bytes)."""
    st, m = unmarshal(payload)
    if st != OK:
        return st, b"", b""
    return (OK | (DONE if m["done"] else 0)), b"chat." + m["cid"].encode("utf-8") + b".tokens", marshal(m, go122)


def start_sequence(last_event_id):
    """fmt.Sscanf(id, "%d", &s); s++ (sse_handler.go:92-96)."""
    if not last_event_id:
        return 0
    s = last_event_id
    i = 0
    while i < len(s) and (s[i] in " \t\v\f" or (s[i] == "\r" and s[i + 1:i + 2] != "\n")):
        i += 1
    j = i
    if j < len(s) and s[j] in "+-":
        j += 1
    k = j
    while k < len(s) and (s[k].isascii() and s[k].isdigit() or s[k] == "_"):
        k += 1
    v = 0
    if k > j and "_" not in s[i:k]:
        x = int(s[i:k])
        if -(1 << 63) <= x < (1 << 63):
            v = x
    v += 1
    if v >= (1 << 63):
        v -= 1 << 64
    return v
