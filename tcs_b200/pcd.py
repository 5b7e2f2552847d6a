"""PCD (Point Cloud Data v0.7) reader/writer: ascii, binary and binary_compressed (LZF).

Support code for the fixtures and the headless CLI: the reference loads its
scenes through pcl::io (src/io.cpp:32-85), which is not in this image, and all of data/*.pcd are
``DATA binary_compressed`` (data/test1.pcd:11).  The format is the public PCD v0.7 one: a text header,
then for binary_compressed two uint32 (compressed size, uncompressed size) followed by an LZF stream whose
decompressed bytes hold the fields as struct-of-arrays (all x, then all y, ...).
"""
from __future__ import annotations

import struct
from typing import Dict, Tuple

import numpy as np

_TYPES = {("F", 4): np.float32, ("F", 8): np.float64, ("U", 1): np.uint8, ("U", 2): np.uint16,
          ("U", 4): np.uint32, ("I", 1): np.int8, ("I", 2): np.int16, ("I", 4): np.int32,
          ("U", 8): np.uint64, ("I", 8): np.int64}


def lzf_decompress(src: bytes, out_len: int) -> bytes:
    """Decompress one LZF block (Marc Lehmann's liblzf format, as used by PCL's lzfDecompress)."""
    out = bytearray(out_len)
    ip, op, n = 0, 0, len(src)
    while ip < n:
        ctrl = src[ip]
        ip += 1
        if ctrl < 32:                      # literal run of ctrl+1 bytes
            ctrl += 1
            out[op:op + ctrl] = src[ip:ip + ctrl]
            ip += ctrl
            op += ctrl
        else:                              # back reference
            length = ctrl >> 5
            ref = op - ((ctrl & 0x1F) << 8) - 1
            if length == 7:
                length += src[ip]
                ip += 1
            ref -= src[ip]
            ip += 1
            length += 2
            if ref < 0 or op + length > out_len:
                raise ValueError("corrupt LZF stream")
            if ref + length <= op:         # non-overlapping: slice copy
                out[op:op + length] = out[ref:ref + length]
                op += length
            else:                          # overlapping run: replicate the period
                period = op - ref
                chunk = bytes(out[ref:op])
                reps = length // period + 1
                out[op:op + length] = (chunk * reps)[:length]
                op += length
    if op != out_len:
        raise ValueError("LZF stream ended early (%d of %d bytes)" % (op, out_len))
    return bytes(out)


def lzf_compress_store(raw: bytes) -> bytes:
    """A valid LZF stream made of literal runs only (no back references) - enough for a writer."""
    out = bytearray()
    for i in range(0, len(raw), 32):
        chunk = raw[i:i + 32]
        out.append(len(chunk) - 1)
        out += chunk
    return bytes(out)


def read_pcd(path: str) -> Tuple[Dict[str, np.ndarray], Dict[str, object]]:
    """Return ({field: 1-D array}, header dict).  Multi-count fields become 2-D (n, count)."""
    with open(path, "rb") as f:
        blob = f.read()
    hdr: Dict[str, object] = {}
    pos = 0
    while True:
        end = blob.index(b"\n", pos)
        line = blob[pos:end].decode("ascii", "replace").strip()
        pos = end + 1
        if not line or line.startswith("#"):
            continue
        key, _, rest = line.partition(" ")
        hdr[key] = rest.split()
        if key == "DATA":
            break
    fields = hdr["FIELDS"]
    sizes = [int(s) for s in hdr["SIZE"]]
    types = hdr["TYPE"]
    counts = [int(c) for c in hdr.get("COUNT", ["1"] * len(fields))]
    n = int(hdr["POINTS"][0]) if "POINTS" in hdr else int(hdr["WIDTH"][0]) * int(hdr["HEIGHT"][0])
    mode = hdr["DATA"][0]
    dts = [np.dtype(_TYPES[(t, s)]) for t, s in zip(types, sizes)]
    out: Dict[str, np.ndarray] = {}
    if mode == "ascii":
        rows = np.loadtxt(blob[pos:].decode("ascii").splitlines()[:n], dtype=np.float64, ndmin=2)  # a graph file has an edge tail
        col = 0
        for name, dt, c in zip(fields, dts, counts):
            a = rows[:, col:col + c]
            col += c
            a = a.astype(dt)
            out[name] = a[:, 0] if c == 1 else a
    elif mode == "binary":
        rec = np.dtype([(nm, dt, (c,)) if c > 1 else (nm, dt) for nm, dt, c in zip(fields, dts, counts)])
        arr = np.frombuffer(blob, dtype=rec, count=n, offset=pos)
        for name in fields:
            out[name] = np.ascontiguousarray(arr[name])
    elif mode == "binary_compressed":
        csize, usize = struct.unpack_from("<II", blob, pos)
        raw = lzf_decompress(blob[pos + 8:pos + 8 + csize], usize)
        off = 0
        for name, dt, c in zip(fields, dts, counts):
            nbytes = n * c * dt.itemsize
            a = np.frombuffer(raw, dtype=dt, count=n * c, offset=off).copy()
            off += nbytes
            out[name] = a if c == 1 else a.reshape(n, c)
    else:
        raise ValueError("unknown PCD DATA mode %r" % mode)
    hdr["n"] = n
    return out, hdr


def write_pcd(path: str, fields: Dict[str, np.ndarray], width: int = 0, height: int = 1,
              mode: str = "binary_compressed") -> None:
    """Write 1-D fields (float32 / uint32 / int32 ...) as PCD v0.7."""
    names = list(fields)
    n = len(fields[names[0]])
    if not width:
        width, height = n, 1
    tchar = {"f": "F", "u": "U", "i": "I"}
    head = ["# .PCD v0.7 - Point Cloud Data file format", "VERSION 0.7", "FIELDS " + " ".join(names),
            "SIZE " + " ".join(str(fields[k].dtype.itemsize) for k in names),
            "TYPE " + " ".join(tchar[fields[k].dtype.kind] for k in names),
            "COUNT " + " ".join("1" for _ in names), "WIDTH %d" % width, "HEIGHT %d" % height,
            "VIEWPOINT 0 0 0 1 0 0 0", "POINTS %d" % n, "DATA " + mode]
    with open(path, "wb") as f:
        f.write(("\n".join(head) + "\n").encode("ascii"))
        if mode == "binary_compressed":
            raw = b"".join(np.ascontiguousarray(fields[k]).tobytes() for k in names)
            comp = lzf_compress_store(raw)
            f.write(struct.pack("<II", len(comp), len(raw)))
            f.write(comp)
        elif mode == "binary":
            rec = np.dtype([(k, fields[k].dtype) for k in names])
            arr = np.empty(n, dtype=rec)
            for k in names:
                arr[k] = fields[k]
            f.write(arr.tobytes())
        elif mode == "ascii":
            cols = [fields[k] for k in names]
            for i in range(n):
                f.write((" ".join(repr(c[i].item()) for c in cols) + "\n").encode("ascii"))
        else:
            raise ValueError(mode)
