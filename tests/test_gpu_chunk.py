"""GPU parity tests of the vLLM stream-chunk decoder (row N3, decode half) through the C ABI, against the CPU oracle; then
its token columns go straight into the encoder (ssegpu_encode_launch) and the payloads must be what the proxy would
publish -- json.Marshal(TokenMessage{...}) of src/llm-stream-proxy/main.go:230-239 -- for the same lines."""
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import chunk_case  # noqa: E402
import pymodel  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pytestmark = pytest.mark.gpu


def decode_on_gpu(eng, torch, buf, off, max_msgs, tok_cap):
    n = len(off) - 1
    d_lines = torch.from_numpy(buf).cuda()
    d_off = torch.from_numpy(off.view(np.int32)).cuda()
    z = lambda m, dt: torch.zeros(max(m, 1), dtype=dt, device="cuda")                      # noqa: E731
    d = dict(tok=z(tok_cap, torch.uint8), tok_off=z(max_msgs + 1, torch.int32), done=z(max_msgs, torch.uint8), line=z(max_msgs, torch.int32),
             status=z(n, torch.int32))
    eng.chunk_decode_launch(d_lines.data_ptr(), d_off.data_ptr(), n, d["tok"].data_ptr(), tok_cap, d["tok_off"].data_ptr(), d["done"].data_ptr(),
                            d["line"].data_ptr(), max_msgs, d["status"].data_ptr())
    torch.cuda.synchronize()
    rc, m, used = eng.chunk_decode_result()
    return rc, m, used, d


def test_chunk_decode_matches_oracle_and_feeds_the_encoder(oracle):
    import torch
    pkg = load_package()
    rng = random.Random(23)
    lines = [v[0] for v in chunk_case.HAND if not v[1] & 0x200] + [chunk_case.rand_line(rng, k) for k in range(60000)]
    buf, off = chunk_case.pack(lines)
    st_o, msgs_o = chunk_case.run_c(oracle.lib.sseo_chunk_decode, buf, off)
    eng = pkg.SseGpu(device=0, max_records=1 << 17, max_in_bytes=64 << 20)
    try:
        rc, m, used, d = decode_on_gpu(eng, torch, buf, off, 1 << 17, 32 << 20)
        assert rc == 0 and m == len(msgs_o)
        assert d["status"].cpu().numpy().astype(np.uint32).tolist() == st_o
        tok_off = d["tok_off"].cpu().numpy().astype(np.uint32)[: m + 1]
        tok = d["tok"].cpu().numpy()
        raw = tok.tobytes()
        done, line = d["done"].cpu().numpy(), d["line"].cpu().numpy()
        got = [(int(line[k]), raw[int(tok_off[k]):int(tok_off[k + 1])], int(done[k])) for k in range(m)]
        assert got == msgs_o
        assert used == int(tok_off[m])
        # the host's part of publishToken: conversation id, sequence and timestamp per message; the token columns stay on the GPU
        cids = [b"chat-%05d" % (ln % 977) for ln, _, _ in got]
        cid_off = np.zeros(m + 1, dtype=np.uint32)
        cid_off[1:] = np.cumsum([len(c) for c in cids])
        cid = np.frombuffer(b"".join(cids) + b"\0" * 64, dtype=np.uint8).copy()
        seq = np.arange(1, m + 1, dtype=np.int64)
        ts = (1700000000000000000 + np.arange(m)).astype(np.int64)
        d_cid, d_cid_off = torch.from_numpy(cid).cuda(), torch.from_numpy(cid_off.view(np.int32)).cuda()
        d_seq, d_ts = torch.from_numpy(seq).cuda(), torch.from_numpy(ts).cuda()
        cap = int(used) * 6 + 160 * m + 4096
        d_out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        d_out_off, d_out_len = torch.zeros(m + 1, dtype=torch.int64, device="cuda"), torch.zeros(m, dtype=torch.int32, device="cuda")
        eng.encode_launch(d_cid.data_ptr(), d_cid_off.data_ptr(), d["tok"].data_ptr(), d["tok_off"].data_ptr(), d_seq.data_ptr(), d["done"].data_ptr(),
                          d_ts.data_ptr(), m, 0, d_out.data_ptr(), cap, d_out_off.data_ptr(), d_out_len.data_ptr())
        torch.cuda.synchronize()
        rc2, _ = eng.encode_result()
        assert rc2 == 0
        out, o_off, o_len = d_out.cpu().numpy().tobytes(), d_out_off.cpu().numpy(), d_out_len.cpu().numpy()
        for k in range(0, m, 13):
            want = pymodel.marshal({"cid": cids[k].decode(), "tok": got[k][1].decode("utf-8"), "seq": int(seq[k]), "done": bool(got[k][2]), "ts": int(ts[k])})
            assert out[int(o_off[k]):int(o_off[k]) + int(o_len[k])] == want, k
        # the product's limit is reported, not hidden: 17 choices in one line
        big = [v[0] for v in chunk_case.HAND if v[1] & 0x200]
        b2, o2 = chunk_case.pack(big)
        rc, m2, _, d2 = decode_on_gpu(eng, torch, b2, o2, 64, 4096)
        assert rc == 0 and m2 == 0 and int(d2["status"][0]) == (1 | 0x200)
        # capacity too small
        rc, _, _, _ = decode_on_gpu(eng, torch, buf, off, 10, 32 << 20)
        assert rc == -6
    finally:
        eng.close()
