"""ctypes binding of the C ABI in include/ssegpu.h.  Mirrors the header one to one."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

OP_ADAPTER, OP_BRIDGE, OP_GO122, OP_NO_FAST, OP_ORDERED, OP_INSPECT, OP_INSPECT_INLINE, OP_MATCH = 0, 1, 2, 4, 8, 16, 32, 64
OP_NATS_PUB, OP_RESP, OP_HEADER_ONLY = 128, 256, 512
ST_NO_INTEREST = 5
INSPECT_PER_TOKEN, INSPECT_STREAMING = 0, 1
KIND_LITERAL, KIND_REGEX = 0, 1
WIRE_NATS, WIRE_RESP = 0, 1
ACT_ALLOW, ACT_REDACT, ACT_DROP, ACT_ALERT = 0, 1, 2, 3
VERDICT_BEGIN_UNKNOWN = 0x10000
ST_OK, ST_ERR_SYNTAX, ST_ERR_TYPE, ST_FILTERED, ST_INSPECT_DROP = 0, 1, 2, 3, 4
ST_CODE_MASK, ST_DONE, ST_REDACTED, ST_HEADER_ONLY = 0xF, 0x100, 0x200, 0x400
E_OVERFLOW, E_AGAIN, E_CONVSLOT, E_PATTERN = -6, -4, -9, -10

u8p, u32p, u64p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int64))


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_records", C.c_uint32), ("max_in_bytes", C.c_uint64),
                ("max_out_bytes", C.c_uint64), ("ring_slots", C.c_uint32), ("conv_slots", C.c_uint32),
                ("flags", C.c_uint32)]


class Pattern(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("action", C.c_uint32), ("text", C.c_char_p)]


class SubDelta(C.Structure):
    _fields_ = [("op", C.c_uint32), ("sub_id", C.c_uint32), ("subject", C.c_char_p), ("subject_len", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("records", "frames", "bytes_in", "bytes_out", "err_syntax", "err_type",
                                          "filtered", "done", "fast_path", "batches")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class SseGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("ssegpu error %d: %s" % (code, msg))
        self.code = code


def lib_path():
    return os.environ.get("SSEGPU_LIB") or os.path.join(HERE, "libssegpu.so")      # SSEGPU_LIB: experimental builds only


_LIB = None


def load_library():
    """Loads libssegpu.so; raises (never falls back) when it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError("libssegpu.so is not built (%s): run `python __graft_entry__.py build`; "
                          "there is no CPU fallback for this path" % path)
    lib = C.CDLL(path)
    vp = C.c_void_p
    lib.ssegpu_init.restype = C.c_int
    lib.ssegpu_init.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.ssegpu_destroy.restype = None
    lib.ssegpu_destroy.argtypes = [vp]
    lib.ssegpu_device_count.restype = C.c_int
    lib.ssegpu_strerror.restype = C.c_char_p
    lib.ssegpu_strerror.argtypes = [C.c_int]
    lib.ssegpu_last_cuda_error.restype = C.c_char_p
    lib.ssegpu_abi_version.restype = C.c_uint32
    lib.ssegpu_shard.restype = C.c_uint32
    lib.ssegpu_shard.argtypes = [C.c_char_p, C.c_size_t, C.c_uint32]
    lib.ssegpu_start_sequence.restype = C.c_int64
    lib.ssegpu_start_sequence.argtypes = [C.c_char_p]
    lib.ssegpu_batch_acquire.restype = C.c_int
    lib.ssegpu_batch_acquire.argtypes = [vp, C.POINTER(vp)]
    for name, rt in (("ssegpu_in_bytes", u8p), ("ssegpu_in_off", u32p), ("ssegpu_start_seq", i64p), ("ssegpu_conv_slot", u32p),
                     ("ssegpu_out_bytes", u8p), ("ssegpu_out_off", u64p), ("ssegpu_out_len", u32p), ("ssegpu_status", u32p), ("ssegpu_subj_len", u32p)):
        f = getattr(lib, name)
        f.restype = rt
        f.argtypes = [vp]
    lib.ssegpu_batch_submit.restype = C.c_int
    lib.ssegpu_batch_submit.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_int]
    lib.ssegpu_batch_wait.restype = C.c_int
    lib.ssegpu_batch_wait.argtypes = [vp, vp, C.c_uint32]
    lib.ssegpu_batch_release.restype = C.c_int
    lib.ssegpu_batch_release.argtypes = [vp, vp]
    lib.ssegpu_counters_get.restype = C.c_int
    lib.ssegpu_counters_get.argtypes = [vp, C.POINTER(Counters)]
    lib.ssegpu_frame_launch.restype = C.c_int
    lib.ssegpu_frame_launch.argtypes = [vp, vp, vp, vp, C.c_uint32, C.c_uint32, vp, C.c_uint64, vp, vp, vp, vp, vp]
    lib.ssegpu_frame_launch_result.restype = C.c_int
    lib.ssegpu_frame_launch_result.argtypes = [vp, C.POINTER(Counters)]
    lib.ssegpu_launch_count.restype = C.c_uint64
    lib.ssegpu_launch_count.argtypes = [vp]
    lib.ssegpu_encode_launch.restype = C.c_int
    lib.ssegpu_encode_launch.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, C.c_uint32, C.c_uint32, vp, C.c_uint64, vp, vp, vp, vp]
    lib.ssegpu_encode_launch_result.restype = C.c_int
    lib.ssegpu_encode_launch_result.argtypes = [vp, C.POINTER(C.c_uint64)]
    lib.ssegpu_interest_apply.restype = C.c_int
    lib.ssegpu_interest_apply.argtypes = [vp, C.POINTER(SubDelta), C.c_uint32]
    lib.ssegpu_interest_subs.restype = u32p
    lib.ssegpu_interest_subs.argtypes = [vp, u32p]
    lib.ssegpu_interest_wild_ids.restype = u32p
    lib.ssegpu_interest_wild_ids.argtypes = [vp, u32p]
    lib.ssegpu_interest_version.restype = C.c_uint64
    lib.ssegpu_interest_version.argtypes = [vp]
    lib.ssegpu_batch_subs.restype = u32p
    lib.ssegpu_batch_subs.argtypes = [vp, u32p]
    lib.ssegpu_batch_wild_ids.restype = u32p
    lib.ssegpu_batch_wild_ids.argtypes = [vp, u32p]
    lib.ssegpu_batch_table_version.restype = C.c_uint64
    lib.ssegpu_batch_table_version.argtypes = [vp]
    lib.ssegpu_sub_span.restype = u32p
    lib.ssegpu_sub_span.argtypes = [vp]
    lib.ssegpu_interest_bind.restype = C.c_int
    lib.ssegpu_interest_bind.argtypes = [vp, vp]
    lib.ssegpu_patterns_load.restype = C.c_int
    lib.ssegpu_patterns_load.argtypes = [vp, C.POINTER(Pattern), C.c_uint32, C.c_uint32]
    lib.ssegpu_conv_reset.restype = C.c_int
    lib.ssegpu_conv_reset.argtypes = [vp]
    lib.ssegpu_inspect_bind.restype = C.c_int
    lib.ssegpu_inspect_bind.argtypes = [vp, vp, vp, vp]
    lib.ssegpu_verdict.restype = u32p
    lib.ssegpu_verdict.argtypes = [vp]
    lib.ssegpu_match.restype = C.POINTER(C.c_int32)
    lib.ssegpu_match.argtypes = [vp]
    lib.ssegpu_split_launch.restype = C.c_int
    lib.ssegpu_split_launch.argtypes = [vp, C.c_uint32, vp, vp, C.c_uint32, vp, C.c_uint64, vp, C.c_uint32, vp, vp, vp, vp, vp, vp, vp]
    lib.ssegpu_split_result.restype = C.c_int
    lib.ssegpu_split_result.argtypes = [vp, u32p, u64p]
    lib.ssegpu_fanout_launch.restype = C.c_int
    lib.ssegpu_fanout_launch.argtypes = [vp, vp, C.c_uint32, vp, vp, C.c_uint64, vp, vp]
    lib.ssegpu_fanout_result.restype = C.c_int
    lib.ssegpu_fanout_result.argtypes = [vp, u64p]
    lib.ssegpu_chunk_decode_launch.restype = C.c_int
    lib.ssegpu_chunk_decode_launch.argtypes = [vp, vp, vp, C.c_uint32, vp, C.c_uint64, vp, vp, vp, C.c_uint32, vp, vp]
    lib.ssegpu_chunk_decode_result.restype = C.c_int
    lib.ssegpu_chunk_decode_result.argtypes = [vp, u32p, u64p]
    lib.ssegpu_frame_launch_static.restype = C.c_int
    lib.ssegpu_frame_launch_static.argtypes = [C.c_int, vp, vp, vp, C.c_uint32, C.c_uint32, vp, C.c_uint64, vp, vp, vp, vp, vp]
    lib.ssegpu_static_last_error.restype = C.c_char_p
    _LIB = lib
    return lib


def shard(conversation_id: bytes, n_gpus: int) -> int:
    return load_library().ssegpu_shard(conversation_id, len(conversation_id), n_gpus)


def start_sequence(last_event_id):
    return load_library().ssegpu_start_sequence(None if last_event_id is None else last_event_id.encode())


STATIC_MAX_RECORDS = 1 << 20


def frame_launch_static(device, d_in, d_in_off, d_start_seq, n, op, d_out, out_cap, d_out_off, d_out_len, d_subj_len,
                        d_status, stream=0):
    """ssegpu_frame_launch_static: context-free launch that allocates nothing (built-in scratch array, at most
    STATIC_MAX_RECORDS records).  Pointer arguments are raw device addresses.  Overflow shows as out_off[n] > out_cap."""
    lib = load_library()
    rc = lib.ssegpu_frame_launch_static(device, d_in, d_in_off, d_start_seq or None, n, op, d_out, out_cap, d_out_off,
                                        d_out_len, d_subj_len or None, d_status, stream or None)
    if rc != 0:
        raise SseGpuError(rc, "%s [%s]" % (lib.ssegpu_strerror(rc).decode(), lib.ssegpu_static_last_error().decode()))


class Batch:
    """One pinned ring slot.  The numpy arrays are views of library-owned pinned memory."""

    def __init__(self, owner, handle):
        self.owner, self.h = owner, handle
        lib, cfg = owner.lib, owner.cfg
        n = cfg.max_records
        self.in_bytes = np.ctypeslib.as_array(lib.ssegpu_in_bytes(handle), shape=(cfg.max_in_bytes + 16,))
        self.in_off = np.ctypeslib.as_array(lib.ssegpu_in_off(handle), shape=(n + 1,))
        self.start_seq = np.ctypeslib.as_array(lib.ssegpu_start_seq(handle), shape=(n,))
        self.conv_slot = np.ctypeslib.as_array(lib.ssegpu_conv_slot(handle), shape=(n,))
        self.n = 0

    def fill(self, arena, off, start_seq=None, conv_slot=None):
        n = len(off) - 1
        total = int(off[n])
        self.in_bytes[:total] = arena[:total]
        self.in_off[: n + 1] = off
        if start_seq is not None:
            self.start_seq[:n] = start_seq
        if conv_slot is not None:
            self.conv_slot[:n] = conv_slot
        self.n = n
        return n

    def submit(self, n=None, op=OP_ADAPTER, use_filter=False):
        n = self.n if n is None else n
        self.n = n
        self.owner._check(self.owner.lib.ssegpu_batch_submit(self.owner.ctx, self.h, n, op, int(use_filter)))

    def wait(self, timeout_us=0):
        self.owner._check(self.owner.lib.ssegpu_batch_wait(self.owner.ctx, self.h, timeout_us))
        lib, n = self.owner.lib, self.n
        self.out_off = np.ctypeslib.as_array(lib.ssegpu_out_off(self.h), shape=(n + 1,))
        self.out_len = np.ctypeslib.as_array(lib.ssegpu_out_len(self.h), shape=(max(n, 1),))[:n]
        self.status = np.ctypeslib.as_array(lib.ssegpu_status(self.h), shape=(max(n, 1),))[:n]
        self.subj_len = np.ctypeslib.as_array(lib.ssegpu_subj_len(self.h), shape=(max(n, 1),))[:n]
        self.sub_span = np.ctypeslib.as_array(lib.ssegpu_sub_span(self.h), shape=(3 * max(n, 1),))[:3 * n].reshape(n, 3)
        self.verdict = np.ctypeslib.as_array(lib.ssegpu_verdict(self.h), shape=(max(n, 1),))[:n]
        self.match = np.ctypeslib.as_array(lib.ssegpu_match(self.h), shape=(2 * max(n, 1),))[:2 * n].reshape(n, 2)
        total = int(self.out_off[n])
        self.out = np.ctypeslib.as_array(lib.ssegpu_out_bytes(self.h), shape=(max(total, 1),))[:total]
        return self

    def subs(self):
        """subscriber-id arena of the table generation this batch was matched against (copy)"""
        n = C.c_uint32(0)
        p = self.owner.lib.ssegpu_batch_subs(self.h, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy() if n.value else np.zeros(0, np.uint32)

    def wild_ids(self):
        n = C.c_uint32(0)
        p = self.owner.lib.ssegpu_batch_wild_ids(self.h, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy() if n.value else np.zeros(0, np.uint32)

    def table_version(self):
        return int(self.owner.lib.ssegpu_batch_table_version(self.h))

    def release(self):
        self.owner._check(self.owner.lib.ssegpu_batch_release(self.owner.ctx, self.h))


class SseGpu:
    """One context per GPU (ssegpu_init / ssegpu_destroy)."""

    def __init__(self, device=0, max_records=1 << 16, max_in_bytes=1 << 24, max_out_bytes=0, ring_slots=0, conv_slots=0):
        self.lib = load_library()
        self.cfg = Config(device, max_records, max_in_bytes, max_out_bytes, ring_slots, conv_slots, 0)
        ctx = C.c_void_p()
        rc = self.lib.ssegpu_init(C.byref(self.cfg), C.byref(ctx))
        if rc != 0:
            raise SseGpuError(rc, "%s [%s]" % (self.lib.ssegpu_strerror(rc).decode(), self.lib.ssegpu_last_cuda_error().decode()))
        self.ctx = ctx

    def _check(self, rc):
        if rc != 0:
            raise SseGpuError(rc, "%s [%s]" % (self.lib.ssegpu_strerror(rc).decode(), self.lib.ssegpu_last_cuda_error().decode()))

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.ssegpu_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_patterns(self, patterns, mode=INSPECT_PER_TOKEN):
        """patterns: list of (kind, action, text-bytes) -- ssegpu_patterns_load."""
        arr = (Pattern * len(patterns))(*[Pattern(k, a, t) for k, a, t in patterns])
        self._check(self.lib.ssegpu_patterns_load(self.ctx, arr, len(patterns), mode))

    def interest_apply(self, deltas):
        """deltas: list of (op 0 add / 1 del, subscriber id, subject bytes) -- ssegpu_interest_apply."""
        arr = (SubDelta * max(len(deltas), 1))(*[SubDelta(o, i, s, len(s)) for o, i, s in deltas])
        self._keep_deltas = arr
        self._check(self.lib.ssegpu_interest_apply(self.ctx, arr, len(deltas)))

    def interest_subs(self):
        n = C.c_uint32(0)
        p = self.lib.ssegpu_interest_subs(self.ctx, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy() if n.value else np.zeros(0, np.uint32)

    def interest_wild_ids(self):
        n = C.c_uint32(0)
        p = self.lib.ssegpu_interest_wild_ids(self.ctx, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy() if n.value else np.zeros(0, np.uint32)

    def interest_version(self):
        return int(self.lib.ssegpu_interest_version(self.ctx))

    def interest_bind(self, d_sub_span):
        self._check(self.lib.ssegpu_interest_bind(self.ctx, d_sub_span))

    def conv_reset(self):
        self._check(self.lib.ssegpu_conv_reset(self.ctx))

    def inspect_bind(self, d_conv_slot, d_verdict, d_match):
        self._check(self.lib.ssegpu_inspect_bind(self.ctx, d_conv_slot or None, d_verdict, d_match))

    def acquire(self):
        h = C.c_void_p()
        self._check(self.lib.ssegpu_batch_acquire(self.ctx, C.byref(h)))
        return Batch(self, h)

    def counters(self):
        c = Counters()
        self._check(self.lib.ssegpu_counters_get(self.ctx, C.byref(c)))
        return c.as_dict()

    def frame_launch(self, d_in, d_in_off, d_start_seq, n, op, d_out, out_cap, d_out_off, d_out_len, d_subj_len, d_status,
                     stream=0):
        """All pointer arguments are raw device addresses (ints); asynchronous on `stream`."""
        self._check(self.lib.ssegpu_frame_launch(self.ctx, d_in, d_in_off, d_start_seq or None, n, op, d_out, out_cap, d_out_off,
                                                 d_out_len, d_subj_len or None, d_status, stream or None))

    def launch_result(self):
        c = Counters()
        rc = self.lib.ssegpu_frame_launch_result(self.ctx, C.byref(c))
        if rc not in (0, E_OVERFLOW):
            self._check(rc)
        return rc, c.as_dict()

    def encode_launch(self, d_cid, d_cid_off, d_tok, d_tok_off, d_seq, d_done, d_ts, n, op, d_out, out_cap, d_out_off, d_out_len,
                      d_chan_len=0, stream=0):
        """Producer-side batch encode (publishToken, llm-stream-proxy/main.go:230-245); raw device addresses, asynchronous."""
        self._check(self.lib.ssegpu_encode_launch(self.ctx, d_cid or None, d_cid_off or None, d_tok or None, d_tok_off or None,
                                                  d_seq or None, d_done or None, d_ts or None, n, op, d_out or None, out_cap, d_out_off,
                                                  d_out_len or None, d_chan_len or None, stream or None))

    def encode_result(self):
        """-> (rc, bytes used); call after the stream has been synchronised."""
        used = C.c_uint64(0)
        rc = self.lib.ssegpu_encode_launch_result(self.ctx, C.byref(used))
        if rc not in (0, E_OVERFLOW):
            self._check(rc)
        return rc, used.value

    def split_launch(self, proto, d_buf, d_buf_off, n_conn, d_arena, arena_cap, d_in_off, max_msgs, d_msg_conn, d_msg_sid, d_msg_subj,
                     d_conn_consumed, d_conn_status, d_conn_pings, stream=0):
        """ssegpu_split_launch: NATS MSG / RESP push bytes of many connections -> packed payload arena; raw device addresses"""
        self._check(self.lib.ssegpu_split_launch(self.ctx, proto, d_buf or None, d_buf_off, n_conn, d_arena, arena_cap, d_in_off, max_msgs,
                                                 d_msg_conn or None, d_msg_sid or None, d_msg_subj or None, d_conn_consumed or None,
                                                 d_conn_status or None, d_conn_pings or None, stream or None))

    def split_result(self):
        n, used = C.c_uint32(0), C.c_uint64(0)
        rc = self.lib.ssegpu_split_result(self.ctx, C.byref(n), C.byref(used))
        if rc not in (0, E_OVERFLOW):
            self._check(rc)
        return rc, n.value, used.value

    def fanout_launch(self, d_sub_span, n, d_pair_sub, d_pair_rec, max_pairs, d_rec_off, stream=0):
        """ssegpu_fanout_launch: spans of OP_MATCH -> flat (subscriber id, record) pairs; raw device addresses"""
        self._check(self.lib.ssegpu_fanout_launch(self.ctx, d_sub_span or None, n, d_pair_sub or None, d_pair_rec or None, max_pairs, d_rec_off, stream or None))

    def fanout_result(self):
        n = C.c_uint64(0)
        rc = self.lib.ssegpu_fanout_result(self.ctx, C.byref(n))
        if rc not in (0, E_OVERFLOW):
            self._check(rc)
        return rc, n.value

    def chunk_decode_launch(self, d_lines, d_line_off, n_lines, d_tok, tok_cap, d_tok_off, d_done, d_msg_line, max_msgs, d_line_status, stream=0):
        """ssegpu_chunk_decode_launch: vLLM stream lines -> the encoder's token columns; raw device addresses"""
        self._check(self.lib.ssegpu_chunk_decode_launch(self.ctx, d_lines or None, d_line_off, n_lines, d_tok, tok_cap, d_tok_off, d_done or None,
                                                        d_msg_line or None, max_msgs, d_line_status or None, stream or None))

    def chunk_decode_result(self):
        n, used = C.c_uint32(0), C.c_uint64(0)
        rc = self.lib.ssegpu_chunk_decode_result(self.ctx, C.byref(n), C.byref(used))
        if rc not in (0, E_OVERFLOW):
            self._check(rc)
        return rc, n.value, used.value

    def launch_count(self):
        return int(self.lib.ssegpu_launch_count(self.ctx))
