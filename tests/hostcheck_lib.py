"""ctypes binding of tests/hostcheck/libhostcheck.so: the product's per-record device logic
(csrc/sse_record.cuh) compiled for the CPU as a TEST DOUBLE.  Never used by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HC_DIR = os.path.join(ROOT, "tests", "hostcheck")
CSRC = os.path.join(ROOT, "distributed-sse-for-llm-response_b200", "csrc")
u8p, u32p, u64p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int64))
OP_BRIDGE, OP_GO122, OP_NO_FAST, OP_NATS_PUB, OP_RESP = 1, 2, 4, 128, 256


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def load():
    so = os.path.join(HC_DIR, "libhostcheck.so")
    srcs = [os.path.join(HC_DIR, "hostcheck.cpp"), os.path.join(CSRC, "sse_record.cuh"), os.path.join(CSRC, "sse_inspect.cuh"),
            os.path.join(CSRC, "inspect_compile.hpp"), os.path.join(CSRC, "sse_interest.cuh"), os.path.join(CSRC, "interest_table.hpp"),
            os.path.join(CSRC, "sse_wire.cuh"), os.path.join(CSRC, "sse_chunk.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared", "-I", CSRC, "-o", so, srcs[0]])
    lib = C.CDLL(so)
    lib.hc_batch.restype = C.c_int
    lib.hc_batch.argtypes = [u8p, u32p, C.c_uint32, i64p, C.c_uint32, u8p, C.c_uint64, u64p, u32p, u32p, u32p]
    lib.hc_inspect.restype = C.c_int
    lib.hc_inspect.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, u8p, u32p, C.c_uint32, u32p, C.c_uint32, u32p,
                               C.POINTER(C.c_int32), u32p, u32p]
    lib.hc_encode.restype = C.c_int
    lib.hc_encode.argtypes = [u8p, u32p, u8p, u32p, i64p, u8p, i64p, C.c_uint32, C.c_uint32, u8p, C.c_uint64, u64p, u32p]
    return lib


def encode(lib, cols, op=0):
    """cols = tests.encode_case.columns(...) -> (list of record bytes, chan_len)"""
    cid, cid_off, tok, tok_off, seq, done, ts = cols
    n = len(seq)
    cap = (int(cid_off[n]) * 8 + int(tok_off[n]) * 6) + 256 * n + 64
    out = np.zeros(cap, dtype=np.uint8)
    out_off = np.empty(n + 1, dtype=np.uint64)
    chan = np.empty(n, dtype=np.uint32)
    rc = lib.hc_encode(_ptr(cid, u8p), _ptr(cid_off, u32p), _ptr(tok, u8p), _ptr(tok_off, u32p), _ptr(seq, i64p), _ptr(done, u8p),
                       _ptr(ts, i64p), n, op, _ptr(out, u8p), cap, _ptr(out_off, u64p), _ptr(chan, u32p))
    assert rc == 0, rc
    return [out[int(out_off[i]):int(out_off[i + 1])].tobytes() for i in range(n)], chan


class HcDelta(C.Structure):
    _fields_ = [("op", C.c_uint32), ("sub_id", C.c_uint32), ("subject", C.c_char_p), ("subject_len", C.c_uint32)]


def interest(lib, deltas, arena, off):
    """deltas: list of (op, sub_id, subject bytes) -> (hit[n,3], status, subs, wild_ids)"""
    lib.hc_interest.restype = C.c_int
    lib.hc_interest.argtypes = [C.c_void_p, C.c_uint32, u8p, u32p, C.c_uint32, u32p, u32p, u32p, C.c_uint32, u32p, u32p, u32p]
    n = len(off) - 1
    arr = (HcDelta * len(deltas))(*[HcDelta(o, i, s, len(s)) for o, i, s in deltas])
    hit = np.zeros(3 * n, dtype=np.uint32)
    status = np.zeros(n, dtype=np.uint32)
    subs = np.zeros(64 * len(deltas) + 65536, dtype=np.uint32)     # the arena is append-only between rebuilds
    wild = np.zeros(64, dtype=np.uint32)
    ns, nw = C.c_uint32(0), C.c_uint32(0)
    rc = lib.hc_interest(arr, len(deltas), _ptr(arena, u8p), _ptr(off, u32p), n, _ptr(hit, u32p), _ptr(status, u32p),
                         _ptr(subs, u32p), len(subs), C.byref(ns), _ptr(wild, u32p), C.byref(nw))
    assert rc == 0, rc
    return hit.reshape(n, 3), status, subs[: ns.value], wild[: nw.value]


class HcPattern(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("action", C.c_uint32), ("text", C.c_char_p)]


def inspect(lib, patterns, mode, arena, off, conv_slot=None, conv_slots=0):
    """patterns: list of (kind, action, text-bytes).  -> (verdict, match[n,2], status, n_states)"""
    n = len(off) - 1
    arr = (HcPattern * len(patterns))(*[HcPattern(k, a, t) for k, a, t in patterns])
    verdict = np.zeros(n, dtype=np.uint32)
    match = np.zeros(2 * n, dtype=np.int32)
    status = np.zeros(n, dtype=np.uint32)
    ns = C.c_uint32(0)
    cs = None if conv_slot is None else np.ascontiguousarray(conv_slot, dtype=np.uint32)
    rc = lib.hc_inspect(arr, len(patterns), mode, _ptr(arena, u8p), _ptr(off, u32p), n, _ptr(cs, u32p), conv_slots,
                        _ptr(verdict, u32p), match.ctypes.data_as(C.POINTER(C.c_int32)), _ptr(status, u32p), C.byref(ns))
    assert rc == 0, rc
    return verdict, match.reshape(n, 2), status, ns.value


def run(lib, arena, off, start_seq=None, op=0):
    n = len(off) - 1
    cap = int(off[n]) * 9 + 260 * n + 64
    out = np.zeros(cap, dtype=np.uint8)
    out_off = np.empty(n + 1, dtype=np.uint64)
    subj = np.empty(n, dtype=np.uint32)
    status = np.empty(n, dtype=np.uint32)
    nfast = C.c_uint32(0)
    ss = None if start_seq is None else np.ascontiguousarray(start_seq, dtype=np.int64)
    rc = lib.hc_batch(_ptr(arena, u8p), _ptr(off, u32p), n, _ptr(ss, i64p), op, _ptr(out, u8p), cap,
                      _ptr(out_off, u64p), _ptr(subj, u32p), _ptr(status, u32p), C.byref(nfast))
    assert rc == 0, rc
    return out[: int(out_off[n])], out_off, subj, status, nfast.value
