"""Reader for genogrove's serialized groves (stream format 0.3) that goes straight to leaf-chain order.

Python twin of include/genogrove_b200/gg_loader.hpp for the harness: parse the directory, inflate each node block,
keep only what the device snapshot needs (leaf keys in chain order, per index) and upload them with
gg_index_upload_sorted.  No node objects, no pointer tree.  The layout is restated from the reference's writer:
grove_serialize.ipp:267-345 (directory + length-prefixed zlib blocks), node.hpp:527-571 (node block),
src/data_type/interval.cpp:18-24 and genomic_coordinate.cpp:15-22 (keys), serialization_traits.hpp:118-160
(payloads), io/gg_format.hpp:24-40 (the CLI's 12-byte "GROV" wrapper).
"""
from __future__ import annotations

import struct
import zlib
from dataclasses import dataclass, field

import numpy as np

NO_BLOCK = 0xFFFFFFFF


@dataclass
class GgChain:
    """One index of the stream: keys in leaf-chain order."""
    name: str
    start: np.ndarray
    end: np.ndarray
    strand: np.ndarray | None      # uint8 ASCII, genomic_coordinate keys only
    payload: list
    rightmost_leaf_keys: int       # key count of the last leaf (what a later sorted append needs)


@dataclass
class GgStream:
    order: int
    cli_wrapper: bool
    payload_type: int
    num_blocks: int
    ext_block_begin: int
    leaf_keys: int
    external_keys: int
    chains: list = field(default_factory=list)   # directory order = the reference's root-map insertion order


def _payload_reader(payload):
    """payload: None (void), 'str', a struct format such as '<i', or callable(buf, pos) -> (value, new_pos)."""
    if payload is None:
        return None
    if callable(payload):
        return payload
    if payload == "str":
        def rd(buf, pos):
            (n,) = struct.unpack_from("<I", buf, pos)
            pos += 4
            if pos + n > len(buf):
                raise ValueError("string length exceeds the block")
            return bytes(buf[pos:pos + n]).decode("utf-8", "surrogateescape"), pos + n
        return rd
    st = struct.Struct(payload)

    def rd(buf, pos):
        return st.unpack_from(buf, pos)[0], pos + st.size
    return rd


def read_gg(data, key: str = "interval", payload="<i") -> GgStream:
    """Parse a format-0.3 grove stream (bytes or a path).  key: 'interval' | 'genomic_coordinate'."""
    if not isinstance(data, (bytes, bytearray, memoryview)):
        with open(data, "rb") as f:
            data = f.read()
    buf = memoryview(data)
    pos = 0
    cli, ptype = False, 0
    if len(buf) >= 12 and bytes(buf[:4]) == b"GROV":
        if buf[4] != 0 or buf[5] != 3:
            raise ValueError("unsupported .gg format version (this reader handles 0.3)")
        cli, ptype, pos = True, buf[9], 12
    if bytes(buf[pos:pos + 4]) != b"GGB\x03":
        raise ValueError("bad magic (not a format 0.3 grove stream)")
    pos += 4
    order, n_idx = struct.unpack_from("<iI", buf, pos)
    pos += 8
    roots = []
    for _ in range(n_idx):
        (ln,) = struct.unpack_from("<I", buf, pos)
        pos += 4
        name = bytes(buf[pos:pos + ln]).decode()
        pos += ln
        (rid,) = struct.unpack_from("<I", buf, pos)
        pos += 4
        roots.append((name, rid))
    n_blocks, ext_begin, leaf_keys, ext_keys = struct.unpack_from("<IIQQ", buf, pos)
    pos += 24
    stranded = key == "genomic_coordinate"
    rd = _payload_reader(payload)
    key_fmt = struct.Struct("<cQQ" if stranded else "<QQ")
    # per block: (is_leaf, link, keys) — link = next leaf (leaf) or first child (internal)
    blocks = []
    for b in range(ext_begin):
        (clen,) = struct.unpack_from("<Q", buf, pos)
        pos += 8
        if pos + clen > len(buf):
            raise ValueError("compressed block length exceeds remaining stream")
        raw = zlib.decompress(buf[pos:pos + clen])
        pos += clen
        (packed,) = struct.unpack_from("<I", raw, 0)
        p = 4
        leaf, nk = bool(packed >> 31), packed & 0x7FFFFFFF
        if nk >= order:
            raise ValueError("num_keys exceeds order")
        keys = []
        for _ in range(nk):
            kv = key_fmt.unpack_from(raw, p)
            p += key_fmt.size
            pv = None
            if rd is not None:
                pv, p = rd(raw, p)
            if leaf:
                keys.append((kv, pv))
        if leaf:
            (link,) = struct.unpack_from("<I", raw, p)
        else:
            nch, link = struct.unpack_from("<II", raw, p)
            if nch != nk + 1:
                raise ValueError("child count does not match key count")
        blocks.append((leaf, link, keys))
    out = GgStream(order, cli, ptype, n_blocks, ext_begin, leaf_keys, ext_keys)
    total = 0
    for name, rid in roots:
        b = rid
        while not blocks[b][0]:
            b = blocks[b][1]
        s, e, st, pl, last = [], [], [], [], 0
        seen = 0
        while b != NO_BLOCK:
            leaf, nxt, keys = blocks[b]
            if not leaf or seen > ext_begin:
                raise ValueError("leaf chain leaves the node blocks")
            seen += 1
            for kv, pv in keys:
                if stranded:
                    st.append(kv[0][0])
                    s.append(kv[1])
                    e.append(kv[2])
                else:
                    s.append(kv[0])
                    e.append(kv[1])
                pl.append(pv)
            last = len(keys)
            b = nxt
        total += len(s)
        out.chains.append(GgChain(name, np.asarray(s, dtype=np.uint64), np.asarray(e, dtype=np.uint64),
                                  np.asarray(st, dtype=np.uint8) if stranded else None, pl, last))
    if total != leaf_keys:
        raise ValueError("leaf chains do not cover every leaf key")
    return out


def flatten(stream: GgStream):
    """(names sorted -> chromosome ids, chrom, start, end, strand|None, payload lists per id) for upload_sorted."""
    order = sorted(range(len(stream.chains)), key=lambda i: stream.chains[i].name)
    names = [stream.chains[i].name for i in order]
    chrom = np.concatenate([np.full(len(stream.chains[i].start), k, dtype=np.uint32) for k, i in enumerate(order)]) \
        if order else np.zeros(0, np.uint32)
    start = np.concatenate([stream.chains[i].start for i in order]) if order else np.zeros(0, np.uint64)
    end = np.concatenate([stream.chains[i].end for i in order]) if order else np.zeros(0, np.uint64)
    strand = None
    if order and stream.chains[0].strand is not None:
        strand = np.concatenate([stream.chains[i].strand for i in order])
    payloads = [stream.chains[i].payload for i in order]
    return names, chrom, start, end, strand, payloads


def load_gg(data, key: str = "interval", payload="<i", device: int = 0):
    """Stream -> DeviceIndex (gg_index_upload_sorted [+ gg_index_set_strands]).  Returns (index, names, payloads)."""
    from .index import DeviceIndex
    stream = read_gg(data, key, payload)
    names, chrom, start, end, strand, payloads = flatten(stream)
    idx = DeviceIndex(device)
    idx.upload_sorted(chrom, start, end, max(1, len(names)))
    if strand is not None:
        idx.set_strands(strand)
    return idx, names, payloads
