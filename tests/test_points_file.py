"""The `pointsfile` of the integrator (multipolesubsurface.cpp:176-181, surfacepoints.cpp:335-347) in its two forms:
the reference's raw SurfacePoint[44 B] dump -- byte for byte the reference's own points -- and the tagged variant (32-byte header:
magic, version, record size, count, byte-order mark) that lets a reader refuse a truncated, foreign or unrelated file.  CPU only:
the reader and writer are host code (host/skin_render.cpp); the device path that renders from a points file is
tests/test_pbrt_frontend.py::test_tessellatesurfacepoints_renderer_and_pointsfile."""
import os
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ref_points():
    raw = np.load(os.path.join(ROOT, "tests", "golden", "skin_tiny.npz"))["surface_points"]       # the reference's points, as bytes
    return np.frombuffer(raw.tobytes(), np.uint8)


def test_raw_form_is_the_references_dump(pkg, ref_points, tmp_path):
    pts = np.frombuffer(ref_points.tobytes(), pkg.skc.DT_SURFACE_POINT)
    p = str(tmp_path / "pts.bin")
    pkg.write_points_file(p, pts)
    assert np.array_equal(np.fromfile(p, np.uint8), ref_points)                  # nothing added, nothing reordered
    back = pkg.read_points_file(p)
    assert back.tobytes() == pts.tobytes()
    with open(p, "ab") as f:                                                     # a trailing partial record is ignored (ReadBinaryFile)
        f.write(b"\x01\x02\x03")
    assert pkg.read_points_file(p).tobytes() == pts.tobytes()


def test_tagged_form_round_trip_and_refusals(pkg, ref_points, tmp_path):
    pts = np.frombuffer(ref_points.tobytes(), pkg.skc.DT_SURFACE_POINT)[:1000]
    p = str(tmp_path / "pts.skp")
    pkg.write_points_file(p, pts, versioned=True)
    blob = open(p, "rb").read()
    assert len(blob) == 32 + 44 * len(pts) and blob[:8] == b"SKINPTS\x01"
    version, record, count, order = struct.unpack("<IIQI", blob[8:28])
    assert (version, record, count, order) == (1, 44, len(pts), 0x01020304)
    assert blob[32:] == pts.tobytes()
    assert pkg.read_points_file(p).tobytes() == pts.tobytes()
    for name, bad, word in (("short.skp", blob[:-44], "truncated"), ("long.skp", blob + b"\0" * 44, "truncated"),
                            ("v2.skp", blob[:8] + struct.pack("<I", 2) + blob[12:], "version"),
                            ("swapped.skp", blob[:24] + struct.pack(">I", 0x01020304) + blob[28:], "byte order")):
        q = str(tmp_path / name)
        open(q, "wb").write(bad)
        with pytest.raises(RuntimeError) as e:
            pkg.read_points_file(q)
        assert word in str(e.value), (name, str(e.value))
    # an empty tagged file is a valid file with no points
    pkg.write_points_file(str(tmp_path / "empty.skp"), pts[:0], versioned=True)
    assert len(pkg.read_points_file(str(tmp_path / "empty.skp"))) == 0
    with pytest.raises(RuntimeError):
        pkg.read_points_file(str(tmp_path / "nothere.skp"))
