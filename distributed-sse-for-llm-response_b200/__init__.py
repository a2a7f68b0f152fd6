"""B200-native SSE hot path: thin ctypes binding of libssegpu.so (include/ssegpu.h).

The package directory name contains hyphens, so it is loaded by file path
(`__graft_entry__.load_package()`), not by a plain import statement.

There is no CPU fallback anywhere in this package: if libssegpu.so is missing, or no
sm_100 device is present, construction fails loudly.
"""
from .binding import (  # noqa: F401
    ACT_ALERT, ACT_ALLOW, ACT_DROP, ACT_REDACT, INSPECT_PER_TOKEN, INSPECT_STREAMING, KIND_LITERAL, KIND_REGEX, OP_INSPECT,
    OP_INSPECT_INLINE, OP_MATCH, ST_NO_INTEREST, VERDICT_BEGIN_UNKNOWN,
    OP_ADAPTER, OP_BRIDGE, OP_GO122, OP_HEADER_ONLY, OP_NATS_PUB, OP_NO_FAST, OP_ORDERED, OP_RESP, ST_HEADER_ONLY, ST_CODE_MASK, ST_DONE, ST_ERR_SYNTAX, ST_ERR_TYPE, ST_FILTERED, ST_OK,
    STATIC_MAX_RECORDS, WIRE_NATS, WIRE_RESP, Batch, SseGpu, SseGpuError, frame_launch_static, lib_path, load_library, shard, start_sequence,
)
from . import workload  # noqa: F401
from . import sharding  # noqa: F401
from . import host  # noqa: F401
