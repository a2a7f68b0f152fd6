"""Synthetic token streams (BASELINE.json configs) -- ctypes front of host/loadgen.cpp.

config 0: demo/load-generator content (loadtest ids, the 56 sample words), fixed clock
config 1: UUID ids + Mistral-7B-like token lengths
Records come in arrival order (token index major, conversation minor).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SEED_IDS, SEED_TOK = 0x5EED0001, 0x5EED0002
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "libsseload.so")
        if not os.path.exists(path):
            raise ImportError("libsseload.so is not built: run `python __graft_entry__.py build`")
        lib = C.CDLL(path)
        lib.sseload_generate.restype = C.c_uint64
        lib.sseload_generate.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64,
                                         C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        lib.sseload_shard_members.restype = C.c_uint32
        lib.sseload_shard_members.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32]
        lib.sseload_conversation_id.restype = C.c_uint32
        lib.sseload_conversation_id.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_char_p]
        _LIB = lib
    return _LIB


def shard_members(config, n_total, rank, world, seed_ids=SEED_IDS):
    """Conversation indices in [0, n_total) that the host routes to GPU `rank` of `world` (fnv1a64(id) % world)."""
    out = np.empty(n_total, dtype=np.uint32)
    m = _lib().sseload_shard_members(config, seed_ids, n_total, rank, world, out.ctypes.data, n_total)
    return out[:m].copy()


def generate(config, n_conv, n_tok, conv_first=0, seed_ids=SEED_IDS, seed_tok=SEED_TOK, threads=None, arena=None,
             conv_map=None):
    """-> (arena uint8[total+64], off uint32[n+1]).  `arena` may be a preallocated uint8 array (e.g. a pinned view);
    `conv_map` (uint32 indices, e.g. from shard_members) replaces the contiguous range [conv_first, conv_first+n_conv)."""
    lib = _lib()
    cm = None
    if conv_map is not None:
        conv_map = np.ascontiguousarray(conv_map, dtype=np.uint32)
        n_conv = len(conv_map)
        cm = conv_map.ctypes.data
    threads = threads or min(32, os.cpu_count() or 1)
    n = n_conv * n_tok
    off = np.empty(n + 1, dtype=np.uint32)
    total = lib.sseload_generate(config, n_conv, n_tok, conv_first, cm, seed_ids, seed_tok, None, 0, off.ctypes.data, threads)
    if total == 0 and n:
        raise ValueError("workload exceeds the 4 GiB arena limit")
    if arena is None:
        arena = np.zeros(total + 64, dtype=np.uint8)
    assert arena.size >= total + 16
    got = lib.sseload_generate(config, n_conv, n_tok, conv_first, cm, seed_ids, seed_tok, arena.ctypes.data, arena.size,
                               off.ctypes.data, threads)
    assert got == total
    return arena, off


def generate_columns(config, n_conv, n_tok, conv_first=0, seed_ids=SEED_IDS, seed_tok=SEED_TOK, conv_map=None):
    """The messages of generate() before marshalling, as the producer-side encoder takes them:
    -> (cid uint8[], cid_off uint32[n+1], tok uint8[], tok_off uint32[n+1], seq int64[n], done uint8[n], ts int64[n]);
    both byte arenas carry 32 bytes of padding."""
    lib = _lib()
    cm = None
    if conv_map is not None:
        conv_map = np.ascontiguousarray(conv_map, dtype=np.uint32)
        n_conv = len(conv_map)
        cm = conv_map.ctypes.data
    n = n_conv * n_tok
    ct, tt = C.c_uint64(0), C.c_uint64(0)
    f = lib.sseload_generate_columns
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p,
                  C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    if f(config, n_conv, n_tok, conv_first, cm, seed_ids, seed_tok, None, 0, None, None, 0, None, None, None, None, C.byref(ct), C.byref(tt)) != 0:
        raise ValueError("workload exceeds the 4 GiB arena limit")
    cid, tok = np.zeros(ct.value + 32, dtype=np.uint8), np.zeros(tt.value + 32, dtype=np.uint8)
    cid_off, tok_off = np.empty(n + 1, dtype=np.uint32), np.empty(n + 1, dtype=np.uint32)
    seq, ts, done = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64), np.empty(n, dtype=np.uint8)
    rc = f(config, n_conv, n_tok, conv_first, cm, seed_ids, seed_tok, cid.ctypes.data, ct.value, cid_off.ctypes.data, tok.ctypes.data, tt.value,
           tok_off.ctypes.data, seq.ctypes.data, done.ctypes.data, ts.ctypes.data, None, None)
    assert rc == 0, rc
    return cid, cid_off, tok, tok_off, seq, done, ts


def conversation_id(config, i, seed_ids=SEED_IDS):
    buf = C.create_string_buffer(64)
    n = _lib().sseload_conversation_id(config, seed_ids, i, buf)
    return buf.raw[:n]
