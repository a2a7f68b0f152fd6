"""Helpers for the GPU parity tests: run a batch through the C ABI and normalise the result.

By default the library places tiles (128 consecutive records) in the arena in completion order, each tile 16-byte
aligned with at most 15 zero bytes after its last frame; every frame is out[out_off[i] : out_off[i]+out_len[i]].
normalise() checks exactly that (alignment, no overlap, no gap beyond the pad, zeroed pads, record order inside a tile)
and rebuilds the record-ordered arena, which must then equal the oracle's byte for byte.
With OP_ORDERED the arena must equal the oracle's without any normalisation."""
import numpy as np

TILE = 128
OP_ORDERED = 8


def normalise(out, out_off, out_len, ordered):
    n = len(out_len)
    lens = out_len.astype(np.uint64)
    total = int(lens.sum())
    arena_bytes = int(out_off[n])
    want_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=want_off[1:])
    if ordered:
        assert arena_bytes == total, "ordered mode: out_off[n] is not the sum of the record lengths"
        assert np.array_equal(out_off[: n + 1], want_off), "ordered mode: offsets are not the prefix sum of the lengths"
        return out[:total], want_off
    off = out_off[:n].astype(np.uint64)
    arena = np.empty(total, dtype=np.uint8)
    starts, sizes = [], []
    for t0 in range(0, n, TILE):
        t1 = min(n, t0 + TILE)
        # record order and contiguity inside the tile
        assert np.array_equal(off[t0 + 1:t1], off[t0:t1 - 1] + lens[t0:t1 - 1]), "records of a tile are not contiguous"
        size = int(want_off[t1] - want_off[t0])
        starts.append(int(off[t0]))
        sizes.append(size)
        arena[int(want_off[t0]):int(want_off[t1])] = out[int(off[t0]):int(off[t0]) + size]
    # the tiles partition the arena: 16-byte aligned, no overlap, only zero pad bytes in between
    order = np.argsort(np.array(starts, dtype=np.int64), kind="stable")
    pos = 0
    for k in order:
        if sizes[k] == 0:
            continue
        assert starts[k] % 16 == 0 and starts[k] == pos, "tiles overlap, are misaligned or leave a gap in the arena"
        end = starts[k] + sizes[k]
        pos = (end + 15) & ~15
        assert not out[end:pos].any(), "pad bytes after a tile are not zero"
    assert pos == arena_bytes, "out_off[n] is not the arena size"
    return arena, want_off


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
