"""Reader/writer for the ".flat" scenario container (include/sumob200_flatfile.h) -- numpy views, no copies."""
import struct

import numpy as np

_DTYPES = {0: np.uint8, 1: np.int32, 2: np.int64, 3: np.float64, 4: np.uint64}
_CODES = {np.dtype(v): k for k, v in _DTYPES.items()}


def load_flat(path):
    """returns {name: np.ndarray}"""
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:8] != b"SB2FLAT1":
        raise ValueError("%s: not a .flat file" % path)
    (n,) = struct.unpack_from("<I", buf, 8)
    off = 12
    out = {}
    for _ in range(n):
        name = buf[off:off + 40].split(b"\0", 1)[0].decode()
        dtype, esize, count = struct.unpack_from("<IIQ", buf, off + 40)
        off += 56
        nbytes = count * esize
        out[name] = np.frombuffer(buf, dtype=_DTYPES[dtype], count=count, offset=off)
        off += nbytes + (8 - nbytes % 8) % 8
    return out


def save_flat(path, arrays):
    with open(path, "wb") as f:
        f.write(b"SB2FLAT1")
        f.write(struct.pack("<I", len(arrays)))
        for name in sorted(arrays):
            a = np.ascontiguousarray(arrays[name])
            f.write(name.encode().ljust(40, b"\0")[:40])
            f.write(struct.pack("<IIQ", _CODES[a.dtype], a.dtype.itemsize, a.size))
            raw = a.tobytes()
            f.write(raw)
            f.write(b"\0" * ((8 - len(raw) % 8) % 8))


def names(arrays, prefix):
    """decode a name table (<prefix>_name_begin / <prefix>_names) into a list of str"""
    b = arrays[prefix + "_name_begin"]
    raw = arrays[prefix + "_names"].tobytes()
    return [raw[b[i]:b[i + 1]].decode() for i in range(len(b) - 1)]


def load_trace(path):
    """reference trace written by integration/sumo_flatten.cpp --trace-out: list of (t_ms, ids, lanes, pos, speed)"""
    raw = np.fromfile(path, dtype=np.int64)
    steps = []
    off = 0
    while off < raw.size:
        t, n = int(raw[off]), int(raw[off + 1])
        off += 2
        if off + 4 * n > raw.size:
            break  # truncated last record (writer was interrupted)
        rec = raw[off:off + 4 * n].reshape(n, 4)
        off += 4 * n
        steps.append((t, rec[:, 0].copy(), rec[:, 1].copy(), rec[:, 2].copy().view(np.float64), rec[:, 3].copy().view(np.float64)))
    return steps
