"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/ssegpu.h declares, agrees with
the oracle's status constants, and refuses to work without a GPU (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest

from __graft_entry__ import ROOT, load_package


@pytest.fixture(scope="module")
def pkg():
    import __graft_entry__ as g
    g.build()
    return load_package()


def _header():
    return open(os.path.join(ROOT, "include", "ssegpu.h")).read()


def test_every_declared_symbol_is_exported(pkg):
    lib = pkg.load_library()
    names = sorted(set(re.findall(r"\b(ssegpu_[a-z0-9_]+)\s*\(", _header())))
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libssegpu.so does not export " + n


def test_status_constants_match_oracle_header(pkg):
    h = _header()
    o = open(os.path.join(ROOT, "oracle", "sse_oracle.h")).read()
    for name in ("OK", "ERR_SYNTAX", "ERR_TYPE", "FILTERED", "INSPECT_DROP", "NO_INTEREST", "CODE_MASK", "DONE", "REDACTED"):
        a = re.search(r"#define SSEGPU_ST_%s\s+(\w+)" % name, h).group(1)
        b = re.search(r"#define SSEO_%s\s+(\w+)" % name, o).group(1)
        assert int(a.rstrip("u"), 0) == int(b.rstrip("u"), 0), name
    assert pkg.ST_DONE == 0x100 and pkg.ST_ERR_TYPE == 2


def test_host_helpers_match_hand_vectors(pkg, oracle):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import hand_vectors as hv
    for s, want in hv.START_SEQUENCE:
        assert pkg.start_sequence(s) == want, s
    for cid in (b"", b"a", b"7f3c2b1a-1111-4222-8333-944455556666"):
        for n in (1, 2, 8):
            assert pkg.shard(cid, n) == oracle.fnv1a64(cid) % n


def test_no_cpu_fallback_without_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.SseGpuError) as e:
        pkg.SseGpu(device=0, max_records=16, max_in_bytes=4096)
    assert e.value.code == -2          # SSEGPU_E_CUDA


def test_product_does_not_reference_oracle():
    """The shipped sources never include, link or call anything under oracle/."""
    pk = os.path.join(ROOT, "distributed-sse-for-llm-response_b200")
    for d, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".cpp", ".py")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                assert "sse_oracle" not in txt and "sseo_" not in txt and "oracle_lib" not in txt, f


def test_workload_shapes(pkg, oracle):
    arena, off = pkg.workload.generate(0, 10, 60)
    n = len(off) - 1
    assert n == 600
    first = arena[off[0]:off[1]].tobytes()
    assert first == (b'{"conversation_id":"loadtest-1700000000000000000-0","token":"The","sequence":1,"done":false,'
                     b'"timestamp":1700000000000000000}')
    # arrival order: record r = token r // n_conv of conversation r % n_conv; word 9 carries a newline
    rec = arena[off[8 * 10 + 3]:off[8 * 10 + 4]].tobytes()
    assert b'"loadtest-1700000000000000000-3"' in rec and b'"token":"dog.\\nThis","sequence":9' in rec
    last = arena[off[n - 1]:off[n]].tobytes()
    assert b'"done":true' in last
    arena, off = pkg.workload.generate(1, 2000, 50)
    out, out_off, status = oracle.adapter(arena, off)
    assert (status & 0xF == 0).all()
    n = len(off) - 1
    assert 125 < off[n] / n < 140 and 150 < out_off[n] / n < 170
    # deterministic and sliceable by conversation range
    a2, o2 = pkg.workload.generate(1, 1000, 50, conv_first=1000)
    assert a2[o2[0]:o2[1]].tobytes() == arena[off[1000]:off[1001]].tobytes()
