"""ctypes binding of oracle/libsse_oracle.so -- test infrastructure only."""
import ctypes as C
import gzip
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
u8p, u32p, u64p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int64))


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.sseo_adapter_batch.restype = C.c_int
        lib.sseo_adapter_batch.argtypes = [u8p, u32p, C.c_uint32, i64p, C.c_int, u8p, C.c_uint64, u64p, u32p]
        lib.sseo_bridge_batch.restype = C.c_int
        lib.sseo_bridge_batch.argtypes = [u8p, u32p, C.c_uint32, C.c_int, u8p, C.c_uint64, u64p, u32p, u32p]
        lib.sseo_bench_adapter.restype = C.c_double
        lib.sseo_bench_adapter.argtypes = [u8p, u32p, C.c_uint32, C.c_int, C.c_int, u64p, u64p]
        lib.sseo_start_sequence.restype = C.c_int64
        lib.sseo_start_sequence.argtypes = [C.c_char_p]
        lib.sseo_fnv1a64.restype = C.c_uint64
        lib.sseo_fnv1a64.argtypes = [u8p, C.c_size_t]

    def encode(self, cols, go122=False, with_channel=False):
        """cols = tests.encode_case.columns(...) -> (list of record bytes, chan_len)"""
        cid, cid_off, tok, tok_off, seq, done, ts = cols
        n = len(seq)
        cap = (int(cid_off[n]) * 7 + int(tok_off[n]) * 6) + 128 * n + 64
        out = np.empty(cap, dtype=np.uint8)
        off = np.empty(n + 1, dtype=np.uint64)
        chan = np.empty(n, dtype=np.uint32)
        self.lib.sseo_encode_batch.restype = C.c_size_t
        self.lib.sseo_encode_batch.argtypes = [u8p, u32p, u8p, u32p, i64p, u8p, i64p, C.c_uint32, C.c_int, C.c_int, u8p, C.c_size_t, u64p, u32p]
        w = self.lib.sseo_encode_batch(_ptr(cid, u8p), _ptr(cid_off, u32p), _ptr(tok, u8p), _ptr(tok_off, u32p), _ptr(seq, i64p),
                                       _ptr(done, u8p), _ptr(ts, i64p), n, int(go122), int(with_channel), _ptr(out, u8p), cap,
                                       _ptr(off, u64p), _ptr(chan, u32p))
        assert w == int(off[n])
        return [out[int(off[i]):int(off[i + 1])].tobytes() for i in range(n)], chan

    def inspector(self, patterns):
        """patterns: list of (kind, action, text-bytes) -> OracleInspector"""
        return OracleInspector(self.lib, patterns)

    @staticmethod
    def pack(payloads):
        """list[bytes] -> (arena u8, offsets u32[n+1])."""
        off = np.zeros(len(payloads) + 1, dtype=np.uint32)
        if payloads:
            off[1:] = np.cumsum([len(p) for p in payloads], dtype=np.uint64).astype(np.uint32)
        arena = np.frombuffer(b"".join(payloads) + b"\0" * 64, dtype=np.uint8).copy()
        return arena, off

    def adapter(self, arena, off, start_seq=None, go122=False):
        n = len(off) - 1
        cap = int(off[n]) * 6 + 160 * n + 64
        out = np.empty(cap, dtype=np.uint8)
        out_off = np.empty(n + 1, dtype=np.uint64)
        status = np.empty(n, dtype=np.uint32)
        ss = None if start_seq is None else np.ascontiguousarray(start_seq, dtype=np.int64)
        rc = self.lib.sseo_adapter_batch(_ptr(arena, u8p), _ptr(off, u32p), n, _ptr(ss, i64p), int(go122),
                                         _ptr(out, u8p), cap, _ptr(out_off, u64p), _ptr(status, u32p))
        assert rc == 0
        return out[: int(out_off[n])], out_off, status

    def bridge(self, arena, off, go122=False):
        n = len(off) - 1
        cap = int(off[n]) * 9 + 200 * n + 64
        out = np.empty(cap, dtype=np.uint8)
        out_off = np.empty(n + 1, dtype=np.uint64)
        subj_len = np.empty(n, dtype=np.uint32)
        status = np.empty(n, dtype=np.uint32)
        rc = self.lib.sseo_bridge_batch(_ptr(arena, u8p), _ptr(off, u32p), n, int(go122), _ptr(out, u8p), cap,
                                        _ptr(out_off, u64p), _ptr(subj_len, u32p), _ptr(status, u32p))
        assert rc == 0
        return out[: int(out_off[n])], out_off, subj_len, status

    def decode_columns(self, arena, off):
        """payload arena -> the encoder's columns (cid, cid_off, tok, tok_off, seq, done, ts); every payload must decode."""
        n = len(off) - 1
        cap = int(off[n]) + 64
        cid, tok = np.zeros(cap, dtype=np.uint8), np.zeros(cap, dtype=np.uint8)
        cid_off, tok_off = np.empty(n + 1, dtype=np.uint32), np.empty(n + 1, dtype=np.uint32)
        seq, ts, done = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64), np.empty(n, dtype=np.uint8)
        self.lib.sseo_decode_columns.restype = C.c_int64
        self.lib.sseo_decode_columns.argtypes = [u8p, u32p, C.c_uint32, u8p, C.c_size_t, u32p, u8p, C.c_size_t, u32p, i64p, u8p, i64p]
        rc = self.lib.sseo_decode_columns(_ptr(arena, u8p), _ptr(off, u32p), n, _ptr(cid, u8p), cap - 32, _ptr(cid_off, u32p), _ptr(tok, u8p), cap - 32,
                                          _ptr(tok_off, u32p), _ptr(seq, i64p), _ptr(done, u8p), _ptr(ts, i64p))
        assert rc == 0, rc
        return cid[: int(cid_off[n]) + 32].copy(), cid_off, tok[: int(tok_off[n]) + 32].copy(), tok_off, seq, done, ts

    def bridge_pub(self, arena, off, go122=False):
        n = len(off) - 1
        cap = int(off[n]) * 9 + 260 * n + 64
        out = np.empty(cap, dtype=np.uint8)
        out_off = np.empty(n + 1, dtype=np.uint64)
        subj_len = np.empty(n, dtype=np.uint32)
        status = np.empty(n, dtype=np.uint32)
        self.lib.sseo_bridge_pub_batch.restype = C.c_int
        self.lib.sseo_bridge_pub_batch.argtypes = [u8p, u32p, C.c_uint32, C.c_int, u8p, C.c_uint64, u64p, u32p, u32p]
        rc = self.lib.sseo_bridge_pub_batch(_ptr(arena, u8p), _ptr(off, u32p), n, int(go122), _ptr(out, u8p), cap,
                                            _ptr(out_off, u64p), _ptr(subj_len, u32p), _ptr(status, u32p))
        assert rc == 0
        return out[: int(out_off[n])], out_off, subj_len, status

    def encode_resp(self, cols, go122=False):
        cid, cid_off, tok, tok_off, seq, done, ts = cols
        n = len(seq)
        cap = (int(cid_off[n]) * 8 + int(tok_off[n]) * 6) + 256 * n + 64
        out = np.empty(cap, dtype=np.uint8)
        off = np.empty(n + 1, dtype=np.uint64)
        chan = np.empty(n, dtype=np.uint32)
        self.lib.sseo_encode_resp_batch.restype = C.c_size_t
        self.lib.sseo_encode_resp_batch.argtypes = [u8p, u32p, u8p, u32p, i64p, u8p, i64p, C.c_uint32, C.c_int, u8p, C.c_size_t, u64p, u32p]
        w = self.lib.sseo_encode_resp_batch(_ptr(cid, u8p), _ptr(cid_off, u32p), _ptr(tok, u8p), _ptr(tok_off, u32p), _ptr(seq, i64p),
                                            _ptr(done, u8p), _ptr(ts, i64p), n, int(go122), _ptr(out, u8p), cap, _ptr(off, u64p), _ptr(chan, u32p))
        assert w == int(off[n])
        return [out[int(off[i]):int(off[i + 1])].tobytes() for i in range(n)], chan

    def bench_adapter(self, arena, off, threads, reps=1):
        n = len(off) - 1
        frames, nbytes = C.c_uint64(0), C.c_uint64(0)
        sec = self.lib.sseo_bench_adapter(_ptr(arena, u8p), _ptr(off, u32p), n, threads, reps, C.byref(frames), C.byref(nbytes))
        return sec, frames.value, nbytes.value

    def bench_prepared(self, arena, off, threads):
        """-> run() -> (seconds, frames, out_bytes) for one timed pass with warm, pre-touched buffers; close() frees them."""
        lib = self.lib
        lib.sseo_bench_new.restype = C.c_void_p
        lib.sseo_bench_new.argtypes = [u8p, u32p, C.c_uint32, C.c_int]
        lib.sseo_bench_run.restype = C.c_double
        lib.sseo_bench_run.argtypes = [C.c_void_p, u64p, u64p]
        lib.sseo_bench_free.argtypes = [C.c_void_p]
        h = lib.sseo_bench_new(_ptr(arena, u8p), _ptr(off, u32p), len(off) - 1, threads)
        assert h, "oracle: out of memory"

        class Prepared:
            def run(self_):
                frames, nbytes = C.c_uint64(0), C.c_uint64(0)
                sec = lib.sseo_bench_run(h, C.byref(frames), C.byref(nbytes))
                return sec, frames.value, nbytes.value

            def close(self_):
                lib.sseo_bench_free(h)
        return Prepared()

    def start_sequence(self, s):
        return self.lib.sseo_start_sequence(s.encode() if s is not None else None)

    def fnv1a64(self, b):
        a = np.frombuffer(bytes(b) + b"\0", dtype=np.uint8)
        return self.lib.sseo_fnv1a64(_ptr(a, u8p), len(b))


class _Pat(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("action", C.c_uint32), ("text", C.c_char_p)]


class OracleInspector:
    """Brute-force matcher of oracle/sse_inspect_oracle.c."""

    def __init__(self, lib, patterns):
        self.lib = lib
        lib.sseo_inspect_new.restype = C.c_void_p
        lib.sseo_inspect_new.argtypes = [C.c_void_p, C.c_uint32]
        lib.sseo_inspect_free.argtypes = [C.c_void_p]
        i32p = C.POINTER(C.c_int32)
        lib.sseo_inspect_token.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, u32p, i32p, i32p]
        lib.sseo_inspect_stream.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, u32p, i32p, i32p]
        self._keep = (_Pat * len(patterns))(*[_Pat(k, a, t) for k, a, t in patterns])
        self.h = lib.sseo_inspect_new(self._keep, len(patterns))
        assert self.h, "oracle rejected the pattern set"

    def token(self, text):
        v, b, e = C.c_uint32(0), C.c_int32(0), C.c_int32(0)
        self.lib.sseo_inspect_token(self.h, text, len(text), C.byref(v), C.byref(b), C.byref(e))
        return v.value, b.value, e.value

    def stream(self, text, s, t, done):
        v, b, e = C.c_uint32(0), C.c_int32(0), C.c_int32(0)
        self.lib.sseo_inspect_stream(self.h, text, s, t, int(done), C.byref(v), C.byref(b), C.byref(e))
        return v.value, b.value, e.value

    def __del__(self):
        try:
            self.lib.sseo_inspect_free(self.h)
        except Exception:
            pass


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "libsse_oracle.so"])


def load():
    so = os.path.join(ORACLE_DIR, "libsse_oracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".c", ".h"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        build()
    return Oracle(C.CDLL(so))


def load_golden():
    path = os.path.join(ROOT, "tests", "golden", "adapter_vectors.json.gz")
    with gzip.open(path, "rb") as f:
        return json.loads(f.read())
