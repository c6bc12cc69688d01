"""`.gg` -> device fast path against streams written by the reference itself.

tests/golden/gg/*.gg are grove streams the reference serialized (format 0.3) after building each grove through every
insert path; *.script holds the reference's own answers for queries on the grove IT restored from that stream
(`tests/cpp/ref_parity.bin --write-gg-fixtures tests/golden/gg`, see tests/golden/make_gg_fixtures.sh).  The reader
takes the leaf chain out of the blocks without building a tree.

* CPU (`-m "not gpu"`): the restated format parses, the chains it yields give the reference's answers through the flat
  oracle, the directory counts agree.
* GPU: the same through gg_index_upload_sorted + the CUDA kernels.
The post-load mutation lines of the scripts (A / I) are replayed by the C++ facade test (tests/cpp/grove_tests.cpp).
"""
import os

import numpy as np
import pytest

from genogrove_b200 import gg_file

HERE = os.path.dirname(os.path.abspath(__file__))
GG = os.path.join(HERE, "golden", "gg")
CASES = [
    ("interval_int_o3", "interval", "<i", 3),
    ("interval_int_o32", "interval", "<i", 32),
    ("interval_str_o5", "interval", "str", 5),
    ("gcoord_int_o4", "genomic_coordinate", "<i", 4),
    ("gcoord_str_o16", "genomic_coordinate", "str", 16),
]


def _script_queries(stem):
    """Q lines before the first mutation: (index, strand, start, end, [payload strings])."""
    out = []
    with open(os.path.join(GG, stem + ".script")) as f:
        for line in f:
            line = line.rstrip("\n")
            if not line:
                continue
            if line[0] != "Q":
                break
            head, _, tail = line.partition("|")
            _, index, strand, s, e = head.split()
            # payloads are separated by single spaces; an empty string payload shows up as an empty token
            want = tail[1:].split(" ") if tail[1:] else []
            out.append((index, strand, int(s), int(e), want))
    return out


def _flat_inputs(stream):
    names, chrom, start, end, strand, payloads = gg_file.flatten(stream)
    coff = np.concatenate([[0], np.cumsum(np.bincount(chrom, minlength=max(1, len(names))))]).astype(np.uint64)
    return names, chrom, start, end, strand, payloads, coff


def _batch(names, queries):
    qs = [q for q in queries if q[0] != "*"]          # '*' walks the reference's hash map: the C++ facade test covers it
    ids = {n: i for i, n in enumerate(names)}
    qc = np.array([ids.get(q[0], 0xFFFFFFFF) for q in qs], dtype=np.uint32)
    a = np.array([q[2] for q in qs], dtype=np.uint64)
    b = np.array([q[3] for q in qs], dtype=np.uint64)
    st = np.frombuffer("".join(q[1] for q in qs).encode(), dtype=np.uint8)
    return qs, qc, a, b, st


def _compare(qs, qc, offsets, ranks, payloads):
    assert len(qs) > 50
    total = 0
    for i, q in enumerate(qs):
        got = [str(payloads[qc[i]][r]) for r in ranks[int(offsets[i]):int(offsets[i + 1])]] if qc[i] != 0xFFFFFFFF else []
        assert got == q[4], (q[:4], got[:8], q[4][:8])
        total += len(got)
    assert total > 0


@pytest.mark.parametrize("stem,key,payload,order", CASES)
def test_stream_parses_and_matches_the_reference_through_the_oracle(oracle, stem, key, payload, order):
    stream = gg_file.read_gg(os.path.join(GG, stem + ".gg"), key, payload)
    assert stream.order == order and not stream.cli_wrapper
    assert sum(len(c.start) for c in stream.chains) == stream.leaf_keys
    assert [c.name for c in stream.chains if c.name == "gone"] == []      # an emptied index is not written
    for c in stream.chains:
        assert 0 < c.rightmost_leaf_keys < order
        assert np.all(c.start <= c.end)
    names, chrom, start, end, strand, payloads, coff = _flat_inputs(stream)
    qs, qc, a, b, st = _batch(names, _script_queries(stem))
    if strand is None:
        offsets, ranks = oracle.flat_intersect_batch(start, end, coff, qc, a, b)
    else:
        offsets, ranks = oracle.flat_intersect_batch_stranded(start, end, strand, coff, qc, a, b, st)
    _compare(qs, qc, offsets, ranks, payloads)


def test_wrapper_and_malformed_streams():
    with open(os.path.join(GG, "interval_int_o3.gg"), "rb") as f:
        good = f.read()
    wrapped = b"GROV\x00\x03\x00\x01\x00\x02\x00\x00" + good
    s = gg_file.read_gg(wrapped)
    assert s.cli_wrapper and s.payload_type == 2 and s.order == 3
    for bad in (b"", good[: len(good) // 2], b"GGB\x02" + good[4:], b"GROV\x00\x02\x00\x01\x00\x02\x00\x00" + good):
        with pytest.raises(Exception):
            gg_file.read_gg(bad)


@pytest.mark.gpu
@pytest.mark.parametrize("stem,key,payload,order", CASES)
def test_stream_to_device_matches_the_reference(stem, key, payload, order):
    idx, names, payloads = gg_file.load_gg(os.path.join(GG, stem + ".gg"), key, payload)
    qs, qc, a, b, st = _batch(names, _script_queries(stem))
    if key == "interval":
        offsets, ranks = idx.intersect_batch(qc, a, b)
    else:
        offsets, ranks = idx.intersect_batch_stranded(qc, a, b, st)
    _compare(qs, qc, offsets, ranks, payloads)
    idx.close()
