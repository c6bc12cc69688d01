"""tools/gg_isec.bin — the batched `genogrove intersect` for BED files.
Known answers restate tests/cli/intersect-test.cpp:86-118 with the content of tests/cli/data/{target,query}.bed;
the larger case is checked against the flat oracle (hits printed in leaf-chain order, queries in file order)."""
import os
import subprocess

import numpy as np
import pytest

import oracle as O
from helpers import chrom_offsets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tools", "gg_isec.bin")


def test_isec_binary_is_built():
    assert os.path.exists(BIN), "run __graft_entry__.build()"


def _run(args, expect_rc=0):
    r = subprocess.run([BIN] + args, capture_output=True, text=True, timeout=600)
    assert r.returncode == expect_rc, r.stderr
    return r


@pytest.mark.gpu
def test_isec_reference_cli_known_answer(tmp_path):
    t = tmp_path / "target.bed"
    q = tmp_path / "query.bed"
    t.write_text("chr1\t100\t500\nchr1\t600\t900\nchr2\t200\t400\n")
    q.write_text("chr1\t150\t250\nchr1\t850\t950\nchr2\t500\t600\n")
    r = _run(["-q", str(q), "-t", str(t)])
    assert r.stdout == "chr1\t100\t500\nchr1\t600\t900\n"  # chr2:500-600 does not overlap chr2:200-400
    out = tmp_path / "o.bed"
    _run(["-q", str(q), "-t", str(t), "-o", str(out), "-k", "5"])
    assert out.read_text() == r.stdout
    # half-open BED: [100,500) and [500,600) do not overlap; [499,500) does
    q.write_text("# comment\ntrack name=x\nchr1\t500\t600\nchr1\t499\t500\nchrUn\t0\t10\n")
    assert _run(["-q", str(q), "-t", str(t)]).stdout == "chr1\t100\t500\n"


@pytest.mark.gpu
def test_isec_errors(tmp_path):
    t = tmp_path / "t.bed"
    t.write_text("chr1\t100\t500\n")
    bad = tmp_path / "bad.bed"
    bad.write_text("chr1\t10\t10\n")
    assert "Start coordinate is greater than or equal" in _run(["-q", str(bad), "-t", str(t)], 1).stderr
    bad.write_text("chr1\tx\t10\n")
    assert "Invalid coordinate format" in _run(["-q", str(bad), "-t", str(t)], 1).stderr
    assert "order must be at least 3" in _run(["-q", str(t), "-t", str(t), "-k", "2"], 1).stderr
    assert "Failed to open" in _run(["-q", str(tmp_path / "missing.bed"), "-t", str(t)], 1).stderr


@pytest.mark.gpu
@pytest.mark.parametrize("sorted_target", [False, True])
def test_isec_matches_oracle(tmp_path, sorted_target):
    rng = np.random.default_rng(11 + sorted_target)
    n, nq, n_chrom = 20000, 30000, 5
    chrom = rng.integers(0, n_chrom, n).astype(np.uint32)
    s = rng.integers(0, 500_000, n).astype(np.uint64)
    e = s + rng.integers(1, 3000, n).astype(np.uint64)  # half-open end
    # distinct (chrom, start, end): the print does not tell equal records apart anyway
    key = np.unique(np.stack([chrom, s, e], 1), axis=0)
    rng.shuffle(key)
    chrom, s, e = key[:, 0].astype(np.uint32), key[:, 1], key[:, 2]
    if sorted_target:
        o = np.lexsort((e, s, chrom))
        chrom, s, e = chrom[o], s[o], e[o]
    names = [f"chr{c + 1}" for c in range(n_chrom)]
    t = tmp_path / "t.bed"
    t.write_text("".join(f"{names[c]}\t{a}\t{b}\n" for c, a, b in zip(chrom, s, e)))
    qc = rng.integers(0, n_chrom + 1, nq)
    qs = rng.integers(0, 510_000, nq).astype(np.uint64)
    qe = qs + rng.integers(1, 2000, nq).astype(np.uint64)
    qn = names + ["chrMissing"]
    q = tmp_path / "q.bed"
    q.write_text("".join(f"{qn[c]}\t{a}\t{b}\n" for c, a, b in zip(qc, qs, qe)))
    got = _run(["-q", str(q), "-t", str(t)] + (["-s"] if sorted_target else [])).stdout
    # oracle: closed coordinates [start, end - 1]; chain order = (start, end)
    perm = np.lexsort((e, s, chrom))
    coff = chrom_offsets(chrom, n_chrom)
    qcc = np.where(qc < n_chrom, qc, O.NO_CHROM).astype(np.uint32)
    off, ranks = O.flat_intersect_batch(s[perm], e[perm] - np.uint64(1), coff, qcc, qs, qe - np.uint64(1))
    lines = []
    for k in range(nq):
        if qc[k] >= n_chrom:
            continue
        base = int(coff[qc[k]])
        for r in ranks[int(off[k]):int(off[k + 1])]:
            i = perm[base + int(r)]
            lines.append(f"{names[chrom[i]]}\t{s[i]}\t{e[i]}\n")
    assert got == "".join(lines)


GFF_TARGET = ("##gff-version 3\nchr1\ttest\tgene\t101\t500\t.\t+\t.\tID=gene1\nchr1\ttest\tgene\t601\t900\t.\t-\t.\tID=gene2\n"
              "chr2\ttest\tgene\t201\t400\t.\t+\t.\tID=gene3\n")
GFF_QUERY = ("##gff-version 3\nchr1\ttest\texon\t151\t250\t.\t+\t.\tID=q1;Parent=gene1\n"
             "chr1\ttest\texon\t851\t950\t.\t-\t.\tID=q2;Parent=gene2\nchr2\ttest\texon\t501\t600\t.\t+\t.\tID=q3\n")


@pytest.mark.gpu
def test_isec_gff_and_cross_type_known_answers(tmp_path):
    """tests/cli/intersect-e2e-test.cpp:371-385 (GFF vs GFF), :431-443 (BED query vs GFF target), :445-461 (GFF query vs
    BED target), with the content of tests/cli/data/{target,query}.{gff,bed}."""
    tg, qg = tmp_path / "target.gff", tmp_path / "query.gff"
    tb, qb = tmp_path / "target.bed", tmp_path / "query.bed"
    tg.write_text(GFF_TARGET)
    qg.write_text(GFF_QUERY)
    tb.write_text("chr1\t100\t500\nchr1\t600\t900\nchr2\t200\t400\n")
    qb.write_text("chr1\t150\t250\nchr1\t850\t950\nchr2\t500\t600\n")
    assert _run(["-q", str(qg), "-t", str(tg)]).stdout == "chr1\t101\t500\nchr1\t601\t900\n"
    assert _run(["-q", str(qb), "-t", str(tg)]).stdout == "chr1\t101\t500\nchr1\t601\t900\n"
    assert _run(["-q", str(qg), "-t", str(tb)]).stdout == "chr1\t100\t500\nchr1\t600\t900\n"
    # 1-based inclusive vs 0-based half-open: GFF [501,600] == BED [500,600): touches BED target [100,500)? no.
    qg.write_text("chr1\t.\tx\t500\t500\t.\t+\t.\tID=a\nchr1\t.\tx\t501\t501\t.\t+\t.\tID=b\n")
    assert _run(["-q", str(qg), "-t", str(tb)]).stdout == "chr1\t100\t500\n"  # base 500 (1-based) = 0-based 499: inside [100,500)
    qg.write_text("chr1\t.\tx\t0\t5\t.\t+\t.\tID=a\n")
    assert "Invalid GFF coordinates" in _run(["-q", str(qg), "-t", str(tb)], 1).stderr


@pytest.mark.gpu
def test_isec_sam_queries_follow_the_reader_coordinate_rule(tmp_path):
    """Read-shaped (text SAM) queries: start = POS - 1, end = start + reference bases the CIGAR consumes, closed
    interval(start, end - 1) ONLY when the record consumes reference (include/genogrove/io/bam_reader.hpp:244-248,283-287;
    src/io/bam_reader.cpp:261-282).  Records restate tests/io/data/test.sam and tests/io/bamfile-test.cpp:676-765."""
    t = tmp_path / "target.bed"
    # targets placed to separate the cases: one base past each expected end must NOT be hit
    t.write_text("chr1\t999\t1000\nchr1\t1049\t1050\nchr1\t1050\t1051\n"      # read001: 50M at POS 1000 -> [999, 1049)
                 "chr1\t2039\t2040\nchr1\t2040\t2041\n"                        # read002: 30M10I10M -> 40 ref bases: [1999, 2039)
                 "chr2\t5053\t5054\nchr2\t5054\t5055\n"                        # read003: 25M5D25M -> 55 ref bases: [4999, 5054)
                 "chr1\t0\t10\nchr1\t499\t500\nchr1\t999\t1000\n"              # unmapped / hard-clip-only / soft-clip-only ask nothing
                 "chr3\t100\t400\n")                                           # read006: 10S20M100N20M5H -> 140 ref bases: [199, 339)
    q = tmp_path / "reads.sam"
    q.write_text("@HD\tVN:1.6\tSO:coordinate\n@SQ\tSN:chr1\tLN:248956422\n"
                 "read001\t99\tchr1\t1000\t60\t50M\t=\t1200\t250\tACGT\tIIII\n"
                 "read002\t0\tchr1\t2000\t30\t30M10I10M\t*\t0\t0\tACGT\tIIII\n"
                 "read003\t16\tchr2\t5000\t45\t25M5D25M\t*\t0\t0\tACGT\tIIII\n"
                 "read004\t4\t*\t0\t0\t*\t*\t0\t0\tNNNN\t!!!!\n"
                 "read_hardclip\t256\tchr1\t500\t60\t100H\t*\t0\t0\t*\t*\n"
                 "read_softclip\t0\tchr1\t1000\t60\t100S\t*\t0\t0\tAAAA\t!!!!\n"
                 "read006\t0\tchr3\t200\t60\t10S20M100N20M5H\t*\t0\t0\tAAAA\t!!!!\n")
    r = _run(["-q", str(q), "-t", str(t)])
    assert r.stdout == ("chr1\t999\t1000\nchr1\t999\t1000\n"   # read001 hits both [999,1000) targets (equal keys), not [1049,1050)
                        # read002 covers [1999, 2039): neither [2039,2040) nor [2040,2041)
                        "chr2\t5053\t5054\n"                   # last base 5053 of read003, not 5054
                        "chr3\t100\t400\n")
    # SAM on the target side: hits print RNAME, start, end (0-based half-open)
    q2 = tmp_path / "q.bed"
    q2.write_text("chr1\t1040\t1060\nchr3\t338\t339\nchr3\t339\t340\n")
    assert _run(["-q", str(q2), "-t", str(q)]).stdout == "chr1\t999\t1049\nchr3\t199\t339\n"
    bad = tmp_path / "bad.sam"
    bad.write_text("r\t0\tchr1\t10\t60\t5Q\t*\t0\t0\tA\t!\n")
    assert "Invalid CIGAR" in _run(["-q", str(bad), "-t", str(t)], 1).stderr
