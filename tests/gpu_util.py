"""Helpers for the GPU parity tests: run a batch through the C ABI and normalise the result.

By default the library places frames in the arena in completion order: the canonical records of a tile (128 consecutive
records) leave as one 16-byte aligned slice, records that needed the general decoder are framed afterwards, a warp's
worth per 16-byte aligned slice.  Every frame is out[out_off[i] : out_off[i]+out_len[i]].  normalise() checks exactly
that -- the frames are disjoint, they tile the arena from 0 to out_off[n] with gaps of at most 15 bytes, every gap byte
is zero -- and rebuilds the record-ordered arena, which must then equal the oracle's byte for byte.
With OP_ORDERED the arena must equal the oracle's without any normalisation."""
import numpy as np

TILE = 128
OP_ORDERED = 8


def normalise(out, out_off, out_len, ordered):
    n = len(out_len)
    lens = out_len.astype(np.int64)
    total = int(lens.sum())
    arena_bytes = int(out_off[n])
    want_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=want_off[1:].view(np.int64))
    if ordered:
        assert arena_bytes == total, "ordered mode: out_off[n] is not the sum of the record lengths"
        assert np.array_equal(out_off[: n + 1], want_off), "ordered mode: offsets are not the prefix sum of the lengths"
        return out[:total], want_off
    off = out_off[:n].astype(np.int64)
    nz = np.nonzero(lens)[0]
    if len(nz) == 0:
        assert arena_bytes == 0
        return out[:0], want_off
    order = nz[np.argsort(off[nz], kind="stable")]
    starts, ends = off[order], off[order] + lens[order]
    gaps = starts[1:] - ends[:-1]
    assert starts[0] == 0, "the arena does not start with a frame"
    assert (gaps >= 0).all(), "frames overlap"
    assert (gaps < 16).all(), "a gap between frames is longer than an alignment pad"
    assert arena_bytes % 16 == 0 and 0 <= arena_bytes - ends[-1] < 16, "out_off[n] is not the arena size"
    cover = np.zeros(arena_bytes + 1, dtype=np.int32)
    cover[starts] += 1
    cover[ends] -= 1
    inside = np.cumsum(cover[:arena_bytes]) > 0
    assert not out[:arena_bytes][~inside].any(), "pad bytes between frames are not zero"
    # record-ordered arena: byte k of record i comes from out[off[i] + k]
    src = np.repeat(off - want_off[:n].astype(np.int64), lens) + np.arange(total, dtype=np.int64)
    return out[src], want_off


def run_batch(eng, arena, off, start=None, op=0):
    """-> (record-ordered arena, prefix offsets, status, subj_len) after checking the placement invariants."""
    b = eng.acquire()
    b.fill(arena, off, start)
    b.submit(len(off) - 1, op, use_filter=start is not None)
    b.wait()
    out, out_off = normalise(b.out.copy(), b.out_off.copy(), b.out_len.copy(), bool(op & OP_ORDERED))
    res = (out, out_off, b.status.copy(), b.subj_len.copy())
    b.release()
    return res


def slices(out, out_off):
    raw = out.tobytes()
    return [raw[int(out_off[i]):int(out_off[i + 1])] for i in range(len(out_off) - 1)]
