"""ctypes front of libssehost.so (host/adapter_host.cpp): the host-side mirror of the SSE Adapter's per-connection
flow -- connect (Last-Event-ID), NATS-callback append, batched GPU pass, one coalesced write per connection."""
import ctypes as C
import os

from .binding import SseGpuError, load_library

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        load_library()
        path = os.path.join(HERE, "libssehost.so")
        if not os.path.exists(path):
            raise ImportError("libssehost.so is not built: run `python __graft_entry__.py build`")
        lib = C.CDLL(path)
        vp = C.c_void_p
        lib.ssehost_new.restype = vp
        lib.ssehost_new.argtypes = [vp, C.c_uint32, C.c_uint64]
        lib.ssehost_free.argtypes = [vp]
        lib.ssehost_connect.argtypes = [vp, C.c_uint64, C.c_char_p, C.c_char_p, C.c_int]
        lib.ssehost_disconnect.argtypes = [vp, C.c_uint64]
        lib.ssehost_keepalive.argtypes = [vp, C.c_uint64]
        lib.ssehost_on_message.argtypes = [vp, C.c_uint64, C.c_char_p, C.c_uint32]
        lib.ssehost_flush.argtypes = [vp]
        lib.ssehost_take.argtypes = [vp, C.c_uint64, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
        lib.ssehost_stats.argtypes = [vp, C.POINTER(C.c_uint64)]
        lib.ssehost_set_header_only.argtypes = [vp, C.c_int]
        _LIB = lib
    return _LIB


class AdapterHost:
    def __init__(self, engine, max_records, max_in_bytes):
        self.lib, self.engine = _lib(), engine
        self.h = self.lib.ssehost_new(engine.ctx, max_records, max_in_bytes)
        if not self.h:
            raise SseGpuError(-1, "ssehost_new failed")
        self._buf = C.create_string_buffer(1 << 20)

    def _check(self, rc):
        if rc != 0:
            raise SseGpuError(rc, self.engine.lib.ssegpu_strerror(rc).decode())

    def set_header_only(self, on=True):
        """SSEGPU_OP_HEADER_ONLY egress: the GPU returns frame headers only, the coalescing loop adds the payload it still holds"""
        self._check(self.lib.ssehost_set_header_only(self.h, int(on)))

    def connect(self, conn_id, conversation_id, last_event_id=None, chat=False):
        self._check(self.lib.ssehost_connect(self.h, conn_id, conversation_id, last_event_id, int(chat)))

    def disconnect(self, conn_id):
        self._check(self.lib.ssehost_disconnect(self.h, conn_id))

    def keepalive(self, conn_id):
        self._check(self.lib.ssehost_keepalive(self.h, conn_id))

    def on_message(self, conn_id, payload):
        rc = self.lib.ssehost_on_message(self.h, conn_id, payload, len(payload))
        if rc == -4:                     # batch full: flush and retry, as the batcher goroutine does
            self.flush()
            rc = self.lib.ssehost_on_message(self.h, conn_id, payload, len(payload))
        self._check(rc)

    def flush(self):
        self._check(self.lib.ssehost_flush(self.h))

    def take(self, conn_id):
        """-> (bytes written by this Write+Flush, handler returned?)"""
        n, closed = C.c_uint64(0), C.c_int(0)
        self._check(self.lib.ssehost_take(self.h, conn_id, self._buf, len(self._buf), C.byref(n), C.byref(closed)))
        if n.value > len(self._buf):
            self._buf = C.create_string_buffer(int(n.value) * 2)
            self._check(self.lib.ssehost_take(self.h, conn_id, self._buf, len(self._buf), C.byref(n), C.byref(closed)))
        return self._buf.raw[: n.value], bool(closed.value)

    def stats(self):
        a = (C.c_uint64 * 4)()
        self._check(self.lib.ssehost_stats(self.h, a))
        return dict(delivered=a[0], dropped=a[1], filtered=a[2], after_close=a[3])

    def close(self):
        if self.h:
            self.lib.ssehost_free(self.h)
            self.h = None
