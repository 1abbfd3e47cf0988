"""TEST INFRASTRUCTURE (oracle): numpy / pure-Python restatement of the PCD v0.7 wire format of
`pcl::PointCloud<LidarSlam::LidarPoint>` as the reference writes and reads it through PCL
(reference call sites: PointCloudStorage.h:85-115 savePointCloudToPCD -> pcl::io::savePCDFile{ASCII,
Binary,BinaryCompressed}; Slam.cxx:531 pcl::io::loadPCDFile; registered fields LidarPoint.h:68-77).

PCL (README min 1.8, CI image 1.10) is not vendored under /root/reference: the format is restated
from its published description (pointclouds.org "The PCD file format", liblzf's stream grammar).
Parity unpinned against a PCL binary (none available here); pinned instead by a hand-written
golden file (tests/golden/lidar_points_*.pcd, produced by oracle/pcd_py.py and checked by eye
against the format description) and by cross round-trips with the C++ implementation.

Only tests/ may import this module."""
import struct

import numpy as np

# (name, numpy dtype, PCD SIZE, PCD TYPE) in POINT_CLOUD_REGISTER_POINT_STRUCT order
FIELDS = (("x", "<f4", 4, "F"), ("y", "<f4", 4, "F"), ("z", "<f4", 4, "F"), ("time", "<f8", 8, "F"),
          ("intensity", "<f4", 4, "F"), ("laser_id", "<u2", 2, "U"), ("device_id", "u1", 1, "U"), ("label", "u1", 1, "U"))
PACKED_DTYPE = np.dtype([(n, d) for n, d, _, _ in FIELDS])   # 28 bytes, no padding
assert PACKED_DTYPE.itemsize == 28


def header(n):
    return ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\n"
            "FIELDS " + " ".join(f[0] for f in FIELDS) + "\n"
            "SIZE " + " ".join(str(f[2]) for f in FIELDS) + "\n"
            "TYPE " + " ".join(f[3] for f in FIELDS) + "\n"
            "COUNT " + " ".join("1" for _ in FIELDS) + "\n"
            "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\n" % (n, n))


# ---- LZF (liblzf stream grammar) -------------------------------------------------
def lzf_decompress(data, raw_size):
    out = bytearray()
    ip, n = 0, len(data)
    while ip < n:
        c = data[ip]
        ip += 1
        if c < 32:                      # literal run of c + 1 bytes
            out += data[ip:ip + c + 1]
            ip += c + 1
        else:                           # back reference
            ln = c >> 5
            if ln == 7:
                ln += data[ip]
                ip += 1
            dist = ((c & 31) << 8 | data[ip]) + 1
            ip += 1
            ln += 2
            if dist > len(out):
                raise ValueError("LZF back reference before the start of the output")
            for _ in range(ln):         # may overlap its own output
                out.append(out[-dist])
    if len(out) != raw_size:
        raise ValueError("LZF stream yields %d bytes, header says %d" % (len(out), raw_size))
    return bytes(out)


def lzf_compress(data):
    """Greedy dictionary coder emitting the liblzf grammar (any valid stream is acceptable to a reader)."""
    out = bytearray()
    n = len(data)
    last = {}
    ip = lit = 0

    def flush(end):
        nonlocal lit
        while lit < end:
            run = min(32, end - lit)
            out.append(run - 1)
            out.extend(data[lit:lit + run])
            lit += run

    while ip + 2 < n:
        key = data[ip:ip + 3]
        ref = last.get(key)
        last[key] = ip
        if ref is not None and ip - ref <= 8192:
            ln = 3
            mx = min(264, n - ip)
            while ln < mx and data[ref + ln] == data[ip + ln]:
                ln += 1
            flush(ip)
            off, l2 = ip - ref - 1, ln - 2
            if l2 < 7:
                out.append((l2 << 5) | (off >> 8))
            else:
                out.append((7 << 5) | (off >> 8))
                out.append(l2 - 7)
            out.append(off & 0xFF)
            ip += ln
            lit = ip
        else:
            ip += 1
    flush(n)
    return bytes(out)


# ---- writer / reader ---------------------------------------------------------------
def _real(v):
    return "nan" if v != v else "%.8g" % v          # std::ostream, precision 8, default float field


def pack(points):
    p = np.zeros(len(points), dtype=PACKED_DTYPE)
    for name, _, _, _ in FIELDS:
        p[name] = points[name]
    return p


def write(path, points, fmt):
    """fmt: 0 ascii, 1 binary, 2 binary_compressed."""
    n = len(points)
    body = header(n).encode()
    if fmt == 0:
        lines = []
        for q in points:
            lines.append(" ".join([_real(float(q["x"])), _real(float(q["y"])), _real(float(q["z"])), _real(float(q["time"])),
                                   _real(float(q["intensity"])), str(int(q["laser_id"])), str(int(q["device_id"])),
                                   str(int(q["label"]))]))
        body += b"DATA ascii\n" + ("\n".join(lines) + "\n").encode()
    elif fmt == 1:
        body += b"DATA binary\n" + pack(points).tobytes()
    else:
        p = pack(points)
        soa = b"".join(np.ascontiguousarray(p[name]).tobytes() for name, _, _, _ in FIELDS)
        comp = lzf_compress(soa)
        body += b"DATA binary_compressed\n" + struct.pack("<II", len(comp), len(soa)) + comp
    with open(path, "wb") as f:
        f.write(body)


def read(path, point_dtype):
    """Returns a structured array of `point_dtype` (the 32-byte LidarPoint layout); fields absent from the
    file keep LidarPoint's default constructor values (0, w = 1)."""
    buf = open(path, "rb").read()
    pos = 0
    meta = {}
    while True:
        eol = buf.index(b"\n", pos)
        line = buf[pos:eol].decode().strip()
        pos = eol + 1
        if not line or line.startswith("#"):
            continue
        key, *vals = line.split()
        meta[key] = vals
        if key == "DATA":
            break
    names = meta["FIELDS"]
    sizes = [int(v) for v in meta["SIZE"]]
    types = meta["TYPE"]
    counts = [int(v) for v in meta.get("COUNT", ["1"] * len(names))]
    n = int(meta["POINTS"][0]) if "POINTS" in meta else int(meta["WIDTH"][0]) * int(meta.get("HEIGHT", ["1"])[0])
    known = {f[0]: f for f in FIELDS}
    out = np.zeros(n, dtype=point_dtype)
    out["w"] = 1.0
    file_dtype = []
    for i, (nm, sz, ty, ct) in enumerate(zip(names, sizes, types, counts)):
        k = known.get(nm)
        ok = k is not None and ct == 1 and sz == k[2] and ty == k[3]
        file_dtype.append((nm if ok else "_skip%d" % i, k[1] if ok else "V%d" % (sz * ct)))
    kind = meta["DATA"][0]
    if kind == "ascii":
        rows = [r.split() for r in buf[pos:].decode().splitlines() if r.strip()]
        assert len(rows) >= n
        col = 0
        for (nm, _), ct in zip(file_dtype, counts):
            if not nm.startswith("_skip"):
                vals = [r[col] for r in rows[:n]]
                if known[nm][3] == "F":
                    out[nm] = np.array([float(v) for v in vals]).astype(known[nm][1])
                else:
                    out[nm] = np.array([int(v) for v in vals])
            col += ct
    elif kind == "binary":
        dt = np.dtype(file_dtype)
        arr = np.frombuffer(buf, dtype=dt, count=n, offset=pos)
        for nm, _ in file_dtype:
            if not nm.startswith("_skip"):
                out[nm] = arr[nm]
    elif kind == "binary_compressed":
        csize, rsize = struct.unpack_from("<II", buf, pos)
        raw = lzf_decompress(buf[pos + 8:pos + 8 + csize], rsize)
        o = 0
        for nm, d in file_dtype:
            w = np.dtype(d).itemsize
            if not nm.startswith("_skip"):
                out[nm] = np.frombuffer(raw, dtype=d, count=n, offset=o)
            o += w * n
    else:
        raise ValueError("unknown DATA " + kind)
    return out
