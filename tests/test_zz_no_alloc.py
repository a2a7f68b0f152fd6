"""GPU test of ssegpu_frame_launch_static, the entry point for callers that may not allocate GPU memory (a service
sharing its GPU with an inference server that already owns nearly all of HBM, or a latency-critical loop that cannot
take cudaMalloc's device-wide synchronisation).  It keeps the kernel's scratch in a fixed array inside the library
instead of a cudaMalloc'ed block, so the things worth checking are: it computes exactly what the regular entry point
(ssegpu_frame_launch, scratch allocated by ssegpu_init) computes; it does so twice in a row, i.e. the shared array
really is cleared between calls; and it refuses batches larger than the array was sized for.

Collected last on purpose (file name): a run that stops at the first failure has then already run everything else.
Called through the C ABI (ctypes).  Comparison is bit-exact, like every other parity test here."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise  # noqa: E402

pytestmark = pytest.mark.gpu


def _device_buffers(torch, arena, off, cap):
    n = len(off) - 1
    return dict(
        d_in=torch.from_numpy(arena).cuda(),
        d_off=torch.from_numpy(off.view(np.int32)).cuda(),
        d_out=torch.zeros(cap, dtype=torch.uint8, device="cuda"),
        d_out_off=torch.zeros(n + 1, dtype=torch.int64, device="cuda"),
        d_out_len=torch.zeros(max(n, 1), dtype=torch.int32, device="cuda"),
        d_status=torch.zeros(max(n, 1), dtype=torch.int32, device="cuda"),
    )


def _result(buf, n, ordered):
    out_len = buf["d_out_len"].cpu().numpy().astype(np.uint32)[:n]
    out, out_off = normalise(buf["d_out"].cpu().numpy(), buf["d_out_off"].cpu().numpy().astype(np.uint64), out_len, ordered)
    return out, out_off, out_len, buf["d_status"].cpu().numpy().astype(np.uint32)[:n]


def _workloads(pkg, oracle):
    import make_golden as mg
    a1, o1 = pkg.workload.generate(1, 700, 9)                       # canonical, config-1 shape, 6300 records
    rows = [a1[o1[i]:o1[i + 1]].tobytes() for i in range(len(o1) - 1)]
    rows[5:5] = [b'{ "done": true, "token":"x<y>", "extra":[1,{"a":null}], "sequence": 7 , "conversation_id":"c","timestamp":1}',
                 b'{"sequence":1.0}', b"null", b'{"token":"\xff\\ud800"}', b"{}x", b"",
                 mg.canon("c", "<\n&" * 9000, 3, True, 4)]            # the last one overflows the shared-memory tile caps
    first = oracle.pack(rows)
    second = pkg.workload.generate(0, 300, 40)                       # a different batch, different size
    return [first, second]


def test_static_entry_matches_regular_entry_twice_in_a_row(oracle):
    import torch
    pkg = load_package()
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    eng = pkg.SseGpu(device=0, max_records=1 << 16, max_in_bytes=8 << 20)     # regular entry point: scratch from ssegpu_init
    try:
        work = _workloads(pkg, oracle)
        calls = list(zip(work, (pkg.OP_ADAPTER, pkg.OP_ADAPTER | pkg.OP_ORDERED))) + [(work[0], pkg.OP_BRIDGE), (work[1], pkg.OP_BRIDGE | pkg.OP_NATS_PUB)]
        for call, ((arena, off), op) in enumerate(calls):
            n = len(off) - 1
            cap = int(off[n]) * 7 + 128 * n + 4096
            ref = _device_buffers(torch, arena, off, cap)
            got = _device_buffers(torch, arena, off, cap)
            ref["d_subj"] = torch.zeros(n, dtype=torch.int32, device="cuda")
            got["d_subj"] = torch.zeros(n, dtype=torch.int32, device="cuda")
            eng.frame_launch(ref["d_in"].data_ptr(), ref["d_off"].data_ptr(), 0, n, op, ref["d_out"].data_ptr(), cap,
                             ref["d_out_off"].data_ptr(), ref["d_out_len"].data_ptr(), ref["d_subj"].data_ptr(), ref["d_status"].data_ptr(),
                             stream.cuda_stream)
            stream.synchronize()
            # the entry point under test: no context, no allocation
            pkg.frame_launch_static(0, got["d_in"].data_ptr(), got["d_off"].data_ptr(), 0, n, op, got["d_out"].data_ptr(), cap,
                                    got["d_out_off"].data_ptr(), got["d_out_len"].data_ptr(), got["d_subj"].data_ptr(), got["d_status"].data_ptr(),
                                    stream.cuda_stream)
            stream.synchronize()
            ordered = bool(op & pkg.OP_ORDERED)
            r_out, r_off, r_len, r_status = _result(ref, n, ordered)
            g_out, g_off, g_len, g_status = _result(got, n, ordered)
            assert np.array_equal(g_status, r_status), "call %d: status words differ" % call
            assert np.array_equal(g_len, r_len), "call %d: frame lengths differ" % call
            assert np.array_equal(g_off, r_off) and np.array_equal(g_out, r_out), "call %d: frame bytes differ" % call
            assert np.array_equal(got["d_subj"].cpu().numpy(), ref["d_subj"].cpu().numpy()), "call %d: subject lengths differ" % call
            # and both agree with the CPU oracle
            if op & pkg.OP_NATS_PUB:
                want_out, want_off, want_subj, want_status = oracle.bridge_pub(arena, off)
            elif op & pkg.OP_BRIDGE:
                want_out, want_off, want_subj, want_status = oracle.bridge(arena, off)
            else:
                want_out, want_off, want_status = oracle.adapter(arena, off)
            assert np.array_equal(g_status, want_status) and np.array_equal(g_out, want_out) and np.array_equal(g_off, want_off)
            assert int(got["d_out_off"][n].item()) <= cap
    finally:
        eng.close()


def test_static_entry_rejects_oversize_and_reports_overflow(oracle):
    import torch
    pkg = load_package()
    with pytest.raises(pkg.SseGpuError) as e:
        pkg.frame_launch_static(0, 256, 256, 0, pkg.STATIC_MAX_RECORDS + 1, 0, 256, 1 << 20, 256, 256, 0, 256, 0)
    assert e.value.code == -5 and "SSEGPU_STATIC_MAX_RECORDS" in str(e.value)
    # inspection and the interest table live in a context: this context-free entry must refuse them, not ignore them
    for op in (pkg.OP_INSPECT, pkg.OP_INSPECT_INLINE, pkg.OP_BRIDGE | pkg.OP_MATCH):
        with pytest.raises(pkg.SseGpuError) as e:
            pkg.frame_launch_static(0, 256, 256, 0, 16, op, 256, 1 << 20, 256, 256, 0, 256, 0)
        assert e.value.code == -1
    with pytest.raises(pkg.SseGpuError) as e:           # d_out off the 16-byte grid
        pkg.frame_launch_static(0, 256, 256, 0, 16, 0, 264, 1 << 20, 256, 256, 0, 256, 0)
    assert e.value.code == -1
    # an output arena that is too small is visible as out_off[n] > out_cap, nothing is written past out_cap
    arena, off = pkg.workload.generate(1, 500, 4)
    n = len(off) - 1
    cap = 100_000
    buf = _device_buffers(torch, arena, off, cap + 4096)
    buf["d_out"].fill_(0xAB)
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    pkg.frame_launch_static(0, buf["d_in"].data_ptr(), buf["d_off"].data_ptr(), 0, n, 0, buf["d_out"].data_ptr(), cap,
                            buf["d_out_off"].data_ptr(), buf["d_out_len"].data_ptr(), 0, buf["d_status"].data_ptr(),
                            stream.cuda_stream)
    stream.synchronize()
    assert int(buf["d_out_off"][n].item()) > cap
    assert bool((buf["d_out"][cap:] == 0xAB).all())
    # offsets that are not monotonic: records are clamped into the arena and bit 63 of out_off[n] says so
    bad = off.copy()
    bad[7] = bad[9] + 5
    buf = _device_buffers(torch, arena, bad, 1 << 20)
    pkg.frame_launch_static(0, buf["d_in"].data_ptr(), buf["d_off"].data_ptr(), 0, n, 0, buf["d_out"].data_ptr(), 1 << 20,
                            buf["d_out_off"].data_ptr(), buf["d_out_len"].data_ptr(), 0, buf["d_status"].data_ptr(), stream.cuda_stream)
    stream.synchronize()
    # (torch reads the word as int64: the flag is the sign bit)
    assert int(buf["d_out_off"][n].item()) < 0
    buf = _device_buffers(torch, arena, off, 1 << 20)
    pkg.frame_launch_static(0, buf["d_in"].data_ptr(), buf["d_off"].data_ptr(), 0, n, 0, buf["d_out"].data_ptr(), 1 << 20,
                            buf["d_out_off"].data_ptr(), buf["d_out_len"].data_ptr(), 0, buf["d_status"].data_ptr(), stream.cuda_stream)
    stream.synchronize()
    assert int(buf["d_out_off"][n].item()) >= 0
