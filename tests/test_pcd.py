"""PCD wire format of the keypoint maps (reference PointCloudStorage.h:85-115,
Slam.cxx:504-543): the C++ reader / writer behind lsb_pcd_write / lsb_pcd_read against the
independent numpy restatement (oracle/pcd_py.py), a hand-written known-answer file and the
committed golden files.  Host-only byte work: runs without a GPU."""
import os
import struct

import numpy as np
import pytest

from lidarslam_b200 import capi
from oracle import pcd_py

HERE = os.path.dirname(os.path.abspath(__file__))
FORMATS = (capi.PCD_ASCII, capi.PCD_BINARY, capi.PCD_BINARY_COMPRESSED)


def _cloud(n, seed=0, with_nan=False):
    r = np.random.default_rng(seed)
    p = np.zeros(n, dtype=capi.POINT_DTYPE)
    p["x"], p["y"], p["z"] = (r.normal(0, 40, n).astype(np.float32) for _ in range(3))
    p["w"] = 1.0
    p["time"] = 1.7e9 + r.uniform(0, 1e4, n)          # maps hold absolute wall-clock stamps (Slam.cxx:537)
    p["intensity"] = r.uniform(0, 255, n).astype(np.float32)
    p["laser_id"] = r.integers(0, 128, n)
    p["device_id"] = r.integers(0, 3, n)
    p["label"] = r.integers(0, 2, n)
    if with_nan and n > 2:
        p["x"][1] = np.nan
        p["intensity"][2] = np.nan
    return p


def _ascii_image(p):
    """What survives the ascii format: 8 significant digits (std::ostream precision 8), also for the double `time`."""
    q = p.copy()
    for f in ("x", "y", "z", "intensity"):
        q[f] = np.array([float("%.8g" % v) for v in p[f].astype(np.float64)]).astype(np.float32)
    q["time"] = np.array([float("%.8g" % v) for v in p["time"]])
    return q


def _same(a, b):
    return len(a) == len(b) and a.tobytes() == b.tobytes()


# a hand-written known answer: every byte of this file follows from the format description
KAT_POINTS = [(1.5, -2.25, 0.125, 0.5, 100.0, 7, 1, 0), (1e-3, 123456.789, -0.0, 1700000000.25, 42.5, 65535, 255, 1)]
KAT_ASCII = (b"# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\n"
             b"FIELDS x y z time intensity laser_id device_id label\nSIZE 4 4 4 8 4 2 1 1\nTYPE F F F F F U U U\n"
             b"COUNT 1 1 1 1 1 1 1 1\nWIDTH 2\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 2\nDATA ascii\n"
             b"1.5 -2.25 0.125 0.5 100 7 1 0\n"
             b"0.001 123456.79 -0 1.7e+09 42.5 65535 255 1\n")


def _kat_cloud():
    p = np.zeros(len(KAT_POINTS), dtype=capi.POINT_DTYPE)
    p["w"] = 1.0
    for i, v in enumerate(KAT_POINTS):
        for f, x in zip(("x", "y", "z", "time", "intensity", "laser_id", "device_id", "label"), v):
            p[f][i] = x
    return p


def test_known_answer_ascii_and_binary(built, tmp_path):
    p = _kat_cloud()
    path = str(tmp_path / "kat.pcd")
    assert capi.pcd_write(path, p, capi.PCD_ASCII) == 0
    assert open(path, "rb").read() == KAT_ASCII
    assert capi.pcd_write(path, p, capi.PCD_BINARY) == 0
    raw = open(path, "rb").read()
    head = KAT_ASCII[:KAT_ASCII.index(b"DATA")] + b"DATA binary\n"
    assert raw.startswith(head) and len(raw) == len(head) + 2 * 28
    rec = struct.pack("<fffdfHBB", *KAT_POINTS[1])                 # fields packed back to back, no padding
    assert raw[len(head) + 28:] == rec


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("n", [1, 2, 33, 1000])
def test_cxx_writer_against_python_reader_and_writer(built, tmp_path, fmt, n):
    p = _cloud(n, seed=n, with_nan=True)
    a, b = str(tmp_path / "a.pcd"), str(tmp_path / "b.pcd")
    assert capi.pcd_write(a, p, fmt) == 0
    pcd_py.write(b, p, fmt)
    want = _ascii_image(p) if fmt == capi.PCD_ASCII else p
    assert _same(pcd_py.read(a, capi.POINT_DTYPE), want)            # C++ writer  -> numpy reader
    assert _same(capi.pcd_read(b), want)                           # numpy writer -> C++ reader
    assert _same(capi.pcd_read(a), want)
    if fmt != capi.PCD_BINARY_COMPRESSED:                           # ascii and binary are canonical byte for byte
        assert open(a, "rb").read() == open(b, "rb").read()
    else:                                                          # LZF streams may differ, headers and raw size may not
        ra, rb = open(a, "rb").read(), open(b, "rb").read()
        i = ra.index(b"DATA binary_compressed\n") + len(b"DATA binary_compressed\n")
        assert ra[:i] == rb[:i] and ra[i + 4:i + 8] == rb[i + 4:i + 8] == struct.pack("<I", 28 * n)


def test_lzf_long_runs_and_incompressible(built, tmp_path):
    # constant fields give matches longer than the 264-byte maximum and self-overlapping copies
    n = 5000
    p = np.zeros(n, dtype=capi.POINT_DTYPE)
    p["w"] = 1.0
    p["x"] = np.arange(n, dtype=np.float32)
    path = str(tmp_path / "runs.pcd")
    assert capi.pcd_write(path, p, capi.PCD_BINARY_COMPRESSED) == 0
    assert os.path.getsize(path) < 28 * n // 3
    assert _same(pcd_py.read(path, capi.POINT_DTYPE), p) and _same(capi.pcd_read(path), p)
    # incompressible payload: the stream grows (literal runs), both sides still agree
    q = _cloud(3000, seed=5)
    q["x"] = np.frombuffer(np.random.default_rng(1).bytes(4 * 3000), dtype=np.float32)
    q["x"][np.isnan(q["x"])] = 0
    assert capi.pcd_write(path, q, capi.PCD_BINARY_COMPRESSED) == 0
    assert _same(pcd_py.read(path, capi.POINT_DTYPE), q) and _same(capi.pcd_read(path), q)


def test_reader_accepts_foreign_layouts(built, tmp_path):
    """Other field order, `_` padding, an unknown field and a field of the wrong type (ignored like PCL's
    'Failed to find match for field'); missing fields keep LidarPoint's default constructor values."""
    n = 5
    dt = np.dtype([("intensity", "<f4"), ("_", "V4"), ("z", "<f4"), ("y", "<f4"), ("x", "<f4"), ("ring", "<u2"), ("label", "<f4")])
    a = np.zeros(n, dtype=dt)
    a["intensity"], a["x"], a["y"], a["z"], a["ring"], a["label"] = np.arange(n) + 0.5, np.arange(n), -np.arange(n), 7, 3, 1
    head = ("# comment\nVERSION .7\nFIELDS intensity _ z y x ring label\nSIZE 4 1 4 4 4 2 4\nTYPE F U F F F U F\n"
            "COUNT 1 4 1 1 1 1 1\nWIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary\n" % (n, n)).encode()
    path = str(tmp_path / "foreign.pcd")
    open(path, "wb").write(head + a.tobytes())
    for got in (capi.pcd_read(path), pcd_py.read(path, capi.POINT_DTYPE)):
        assert np.array_equal(got["x"], a["x"]) and np.array_equal(got["y"], a["y"]) and np.array_equal(got["z"], a["z"])
        assert np.array_equal(got["intensity"], a["intensity"])
        assert not got["label"].any() and not got["laser_id"].any() and not got["time"].any() and (got["w"] == 1).all()
    # the same cloud as ascii, WIDTH x HEIGHT without POINTS, CRLF line ends
    rows = "".join("%g 0 0 0 0 %g %g %g %d %g\r\n" % (r["intensity"], r["z"], r["y"], r["x"], r["ring"], r["label"]) for r in a)
    head = head.decode().replace("POINTS %d\n" % n, "").replace("binary", "ascii").replace("\n", "\r\n")
    open(path, "wb").write((head + rows).encode())
    got = capi.pcd_read(path)
    assert np.array_equal(got["x"], a["x"]) and np.array_equal(got["intensity"], a["intensity"]) and (got["z"] == 7).all()


def test_error_codes(built, tmp_path):
    p = _cloud(4)
    assert capi.pcd_write(str(tmp_path / "e.pcd"), p[:0], capi.PCD_BINARY) == -3         # PointCloudStorage.h:91-92
    assert capi.pcd_write(str(tmp_path / "e.pcd"), p, 7) == -4                            # PointCloudStorage.h:111-113
    assert capi.pcd_write(str(tmp_path / "no_such_dir" / "e.pcd"), p, capi.PCD_BINARY) == -1
    with pytest.raises(RuntimeError):
        capi.pcd_read(str(tmp_path / "missing.pcd"))
    path = str(tmp_path / "trunc.pcd")
    assert capi.pcd_write(path, p, capi.PCD_BINARY) == 0
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-5])
    with pytest.raises(RuntimeError, match="shorter"):
        capi.pcd_read(path)
    # a header that promises far more points than the file holds is refused before anything is allocated
    head = raw[:raw.index(b"DATA")].replace(b"POINTS 4", b"POINTS 4000000000000").replace(b"WIDTH 4", b"WIDTH 4000000000000")
    for kind in (b"binary", b"ascii", b"binary_compressed"):
        open(path, "wb").write(head + b"DATA " + kind + b"\n" + raw[-4 * 28:])
        with pytest.raises(RuntimeError):
            capi.pcd_read(path)
    assert capi.pcd_write(path, p, capi.PCD_BINARY_COMPRESSED) == 0
    raw = bytearray(open(path, "rb").read())
    raw[-1] ^= 0xFF
    raw[-2] ^= 0xFF
    open(path, "wb").write(bytes(raw) + b"\xff")
    try:                                                                                  # corrupt stream: error or different bytes, never a crash
        capi.pcd_read(path)
    except RuntimeError:
        pass


@pytest.mark.parametrize("name", ["lidar_points_ascii.pcd", "lidar_points_binary.pcd", "lidar_points_binary_compressed.pcd"])
def test_golden_files(built, name):
    """Committed files (tests/golden/make_golden_pcd.py): both readers must keep decoding them to the committed cloud."""
    want = np.load(os.path.join(HERE, "golden", "lidar_points.npy"))
    path = os.path.join(HERE, "golden", name)
    if "ascii" in name:
        want = _ascii_image(want)
    assert _same(capi.pcd_read(path), want)
    assert _same(pcd_py.read(path, capi.POINT_DTYPE), want)
