"""GPU parity tests of the wire splitter (row N4, receiving side) through the C ABI: NATS MSG / RESP push bytes of
thousands of connections -> packed payload arena, against the CPU oracle; then the arena goes straight into the frame
launch and the frames must be what the oracle frames from the same payloads (socket bytes in, SSE frames out, all on the GPU)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import wire_case  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402
from gpu_util import normalise, slices  # noqa: E402

pytestmark = pytest.mark.gpu


def split_on_gpu(pkg, eng, torch, proto, buf, off, max_msgs, arena_cap):
    n_conn = len(off) - 1
    d_buf = torch.from_numpy(buf).cuda()
    d_off = torch.from_numpy(off.view(np.int32)).cuda()
    z = lambda m, dt: torch.zeros(max(m, 1), dtype=dt, device="cuda")                      # noqa: E731
    d = dict(arena=z(arena_cap, torch.uint8), in_off=z(max_msgs + 1, torch.int32), conn=z(max_msgs, torch.int32), sid=z(max_msgs, torch.int64),
             subj=z(2 * max_msgs, torch.int32), consumed=z(n_conn, torch.int32), status=z(n_conn, torch.int32), pings=z(n_conn, torch.int32))
    eng.split_launch(proto, d_buf.data_ptr(), d_off.data_ptr(), n_conn, d["arena"].data_ptr(), arena_cap, d["in_off"].data_ptr(), max_msgs,
                     d["conn"].data_ptr(), d["sid"].data_ptr(), d["subj"].data_ptr(), d["consumed"].data_ptr(), d["status"].data_ptr(),
                     d["pings"].data_ptr())
    torch.cuda.synchronize()
    rc, n, used = eng.split_result()
    return rc, n, used, d


@pytest.mark.parametrize("proto", [0, 1])
def test_split_matches_oracle_and_feeds_the_frame_kernel(oracle, proto):
    import torch
    pkg = load_package()
    buf, off, _ = wire_case.build(proto, 20000, 77 + proto)
    msgs, conns = wire_case.oracle_split(oracle.lib, proto, buf, off)
    eng = pkg.SseGpu(device=0, max_records=1 << 18, max_in_bytes=64 << 20)
    try:
        rc, n, used, d = split_on_gpu(pkg, eng, torch, proto, buf, off, 1 << 18, 64 << 20)
        assert rc == 0 and n == len(msgs) and used == sum(len(m[3]) for m in msgs)
        in_off = d["in_off"].cpu().numpy().astype(np.uint32)[: n + 1]
        arena = d["arena"].cpu().numpy()
        raw = buf.tobytes()
        conn, sid, subj = d["conn"].cpu().numpy(), d["sid"].cpu().numpy(), d["subj"].cpu().numpy().reshape(-1, 2)
        got = [(int(conn[i]), raw[int(subj[i, 0]):int(subj[i, 0]) + int(subj[i, 1])], int(sid[i]), arena[int(in_off[i]):int(in_off[i + 1])].tobytes())
               for i in range(n)]
        assert got == msgs
        n_conn = len(off) - 1
        got_c = list(zip(np.bincount(conn[:n], minlength=n_conn).tolist(), d["consumed"].cpu().numpy().tolist(), d["status"].cpu().numpy().tolist(),
                         d["pings"].cpu().numpy().tolist()))
        assert got_c == conns
        # socket bytes -> frames without leaving the GPU: the arena + offsets are the frame launch's input as they are
        cap = int(used) * 7 + 64 * n + 4096
        z = lambda m, dt: torch.zeros(max(m, 1), dtype=dt, device="cuda")                  # noqa: E731
        d_out, d_out_off, d_out_len, d_status = z(cap, torch.uint8), z(n + 1, torch.int64), z(n, torch.int32), z(n, torch.int32)
        eng.frame_launch(d["arena"].data_ptr(), d["in_off"].data_ptr(), 0, n, 0, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr(), 0,
                         d_status.data_ptr())
        torch.cuda.synchronize()
        out, out_off = normalise(d_out.cpu().numpy(), d_out_off.cpu().numpy().astype(np.uint64), d_out_len.cpu().numpy().astype(np.uint32)[:n], False)
        p_arena, p_off = oracle.pack([m[3] for m in msgs])
        want_out, want_off, want_status = oracle.adapter(p_arena, p_off)
        assert np.array_equal(d_status.cpu().numpy().astype(np.uint32)[:n], want_status)
        assert slices(out, out_off) == slices(want_out, want_off)
        # capacity too small: reported, nothing written past the limits
        rc, n2, _, _ = split_on_gpu(pkg, eng, torch, proto, buf, off, 100, 64 << 20)
        assert rc == -6 and n2 == len(msgs)
    finally:
        eng.close()


def test_hand_vectors_on_gpu(oracle):
    import torch
    pkg = load_package()
    eng = pkg.SseGpu(device=0, max_records=4096, max_in_bytes=1 << 20)
    try:
        for proto in (0, 1):
            vecs = [v for v in wire_case.HAND if v[0] == proto]
            off = np.zeros(len(vecs) + 1, dtype=np.uint32)
            off[1:] = np.cumsum([len(v[1]) for v in vecs])
            buf = np.frombuffer(b"".join(v[1] for v in vecs) + b"\0" * 64, dtype=np.uint8).copy()
            rc, n, used, d = split_on_gpu(pkg, eng, torch, proto, buf, off, 4096, 1 << 20)
            assert rc == 0
            in_off, arena = d["in_off"].cpu().numpy().astype(np.uint32), d["arena"].cpu().numpy()
            conn, sid, subj = d["conn"].cpu().numpy(), d["sid"].cpu().numpy(), d["subj"].cpu().numpy().reshape(-1, 2)
            raw = buf.tobytes()
            k = 0
            for c, (_, data, want_msgs, consumed, status, pings) in enumerate(vecs):
                got = []
                while k < n and int(conn[k]) == c:
                    got.append((raw[int(subj[k, 0]):int(subj[k, 0]) + int(subj[k, 1])], int(sid[k]), arena[int(in_off[k]):int(in_off[k + 1])].tobytes()))
                    k += 1
                assert got == want_msgs, (proto, c)
                assert (int(d["consumed"][c]), int(d["status"][c]), int(d["pings"][c])) == (consumed, status, pings), (proto, c)
            assert k == n
    finally:
        eng.close()
