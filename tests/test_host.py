"""CPU tests of the product's host side (no GPU): the C-ABI library loads and exports every symbol
include/vr.h declares; the input surfaces kept from the app (camera, projection, gradient, NRRD
loader) give the reference's bits; partition helpers; error behaviour without a device."""
import os
import re
import struct

import numpy as np
import pytest

from common import bits
from oracle import pyoracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "host_surfaces.npz")


def test_library_exports_every_declared_symbol(vr):
    hdr = open(os.path.join(ROOT, "include", "vr.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(vr_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 45
    from volumetric_renderer_b200 import _lib
    missing = [n for n in sorted(declared) if not hasattr(vr.lib, n)]
    assert not missing, f"declared in include/vr.h but not exported: {missing}"
    unbound = [n for n in sorted(declared) if n not in _lib.SIGNATURES]
    assert not unbound, f"declared but not bound in _lib.py: {unbound}"
    assert vr.lib.vr_abi_version() == 2


def test_no_cpu_fallback_without_device(vr):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(vr.VrError) as e:
        vr.OffscreenPass(64, 64)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "volumetric-renderer_b200")
    for dp, _, fns in os.walk(pkg):
        if "build" in dp:
            continue
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "oracle/" not in txt and "pyoracle" not in txt and "vr_oracle" not in txt, os.path.join(dp, fn)


def test_camera_projection_gradient_match_golden_from_reference(vr):
    g = np.load(HOST_GOLDEN)
    cam = vr.Camera()
    for m, want in zip([(0, 0), (4, 0), (-120, 80), (33.3, -17.7), (400, 5)], g["cameras"]):
        cam.rotate(m)
        got = np.concatenate([cam.get_view(), cam.get_position()])
        assert np.array_equal(bits(got), bits(want))
    assert np.array_equal(bits(vr.projection(512, 512)), bits(g["proj_512"]))
    assert np.array_equal(bits(vr.projection(1920, 1080)), bits(g["proj_1080"]))
    assert np.array_equal(bits(vr.projection(100, 100)), bits(g["proj_100"]))
    from common import vr_gradient
    assert np.array_equal(vr.Gradient().discretize(256), g["lut_A"])
    assert np.array_equal(vr_gradient(vr, "B").discretize(256), g["lut_B"])
    assert np.array_equal(vr_gradient(vr, "C").discretize(256), g["lut_C"])
    assert np.array_equal(vr_gradient(vr, "B").discretize(7), g["lut_B_7"])


def test_camera_matches_reference_build_on_random_input(vr, ref):
    rng = np.random.default_rng(3)
    a, b = vr.Camera(), ref.camera()
    for _ in range(200):
        d = tuple(float(x) for x in rng.uniform(-500, 500, 2))
        a.rotate(d)
        b.rotate(*d)
        z = float(rng.uniform(-1, 1))
        a.zoom(z)
        b.zoom(z)
    assert np.array_equal(bits(a.get_view()), bits(b.view))
    assert np.array_equal(bits(a.get_position()), bits(b.position))


def test_gradient_matches_reference_build_on_random_markers(vr, ref):
    rng = np.random.default_rng(5)
    for _ in range(20):
        nc, na = int(rng.integers(1, 6)), int(rng.integers(1, 6))
        cl, al = np.sort(rng.uniform(0, 1, nc)), np.sort(rng.uniform(0, 1, na))
        cm = [(float(l), *[float(x) for x in rng.uniform(-0.2, 1.2, 3)]) for l in cl]
        am = [(float(l), float(rng.uniform(-0.2, 1.2))) for l in al]
        got = vr.Gradient([(m[0], m[1:]) for m in cm], am).discretize(64)
        assert np.array_equal(got, ref.discretize(cm, am, count=64))


# ---- NRRD loader -------------------------------------------------------------------------------

NRRD_TYPES = [("int8", np.int8), ("uchar", np.uint8), ("short", np.int16), ("ushort", np.uint16), ("int", np.int32),
              ("uint", np.uint32), ("longlong", np.int64), ("ulonglong", np.uint64), ("float", np.float32),
              ("double", np.float64)]


def write_nrrd(path, arr, tname, encoding="raw", endian="little", detached=False, extra=""):
    nz, ny, nx = arr.shape
    hdr = f"NRRD0004\n# test file\ntype: {tname}\ndimension: 3\nsizes: {nx} {ny} {nz}\n{extra}encoding: {encoding}\n"
    if arr.dtype.itemsize > 1 and encoding != "ascii":
        hdr += f"endian: {endian}\n"
    data = arr.astype(arr.dtype.newbyteorder("<" if endian == "little" else ">"))
    if encoding == "raw":
        payload = data.tobytes()
    elif encoding == "ascii":
        payload = ("\n".join(" ".join(repr(x.item()) for x in row) for row in arr.reshape(-1, nx)) + "\n").encode()
    else:
        payload = data.tobytes().hex().encode()
    if detached:
        open(str(path) + ".raw", "wb").write(payload)
        hdr += f"data file: {os.path.basename(str(path))}.raw\n"
        open(path, "wb").write(hdr.encode())
    else:
        open(path, "wb").write(hdr.encode() + b"\n" + payload)


@pytest.mark.parametrize("tname,dt", NRRD_TYPES)
def test_nrrd_reader_matches_reference_parser_all_types(vr, ref, tmp_path, tname, dt):
    rng = np.random.default_rng(11)
    info = np.iinfo(dt) if np.issubdtype(dt, np.integer) else None
    arr = (rng.integers(max(info.min, -2**40), min(info.max, 2**40), (3, 4, 5)).astype(dt) if info
           else rng.normal(0, 100, (3, 4, 5)).astype(dt))
    p = tmp_path / f"v_{tname}.nrrd"
    write_nrrd(p, arr, tname)
    got = vr.read_nrrd(str(p))
    want, mn, mx = ref.load_nrrd(str(p))  # NrrdFileParser::parse + NrrdIO
    assert got.shape == want.shape == (3, 4, 5)
    if dt in (np.uint8, np.uint16):
        assert got.dtype == dt and np.array_equal(got, arr)  # native on the device
    else:
        assert got.dtype == np.float32
    assert np.array_equal(got.astype(np.float32), want)      # same floats as the reference's widening
    assert float(got.min()) == mn and float(got.max()) == mx


@pytest.mark.parametrize("encoding,endian,detached", [("raw", "big", False), ("ascii", "little", False),
                                                      ("hex", "little", False), ("hex", "big", False),
                                                      ("raw", "little", True)])
def test_nrrd_reader_encodings(vr, ref, tmp_path, encoding, endian, detached):
    arr = np.arange(2 * 3 * 4, dtype=np.uint16).reshape(2, 3, 4) * 1000
    p = tmp_path / ("v.nhdr" if detached else "v.nrrd")
    write_nrrd(p, arr, "ushort", encoding, endian, detached, extra="spacings: 1 2 3\nkey:=value\n")
    got = vr.read_nrrd(str(p))
    want, _, _ = ref.load_nrrd(str(p))
    assert np.array_equal(got, arr) and np.array_equal(got.astype(np.float32), want)


@pytest.mark.parametrize("encoding", ["raw", "ascii", "hex"])
@pytest.mark.parametrize("tname", ["uchar", "double"])
def test_nrrd_header_claiming_a_huge_volume_is_an_error_not_a_crash(vr, tmp_path, encoding, tname):
    """sizes come from the header alone.  A tiny file claiming 65536^3 voxels (or sizes whose product wraps
    64 bits) must fail with a status -- the reference reports 'Failed to read file' -- not allocate or terminate."""
    for sizes in ("65536 65536 65536", "4294967295 4294967295 4294967295", "2048 2048 2048"):
        p = tmp_path / f"huge_{encoding}_{tname}_{sizes.split()[0]}.nrrd"
        open(p, "wb").write(f"NRRD0004\ntype: {tname}\ndimension: 3\nsizes: {sizes}\nencoding: {encoding}\nendian: little\n\n".encode()
                            + b"0 1 2 3 4 5 6 7\n")
        with pytest.raises(vr.VrError) as e:
            vr.read_nrrd(str(p))
        assert "Failed to read file" in str(e.value)


def test_nrrd_reader_rejects_what_the_reference_rejects(vr, ref, tmp_path):
    p = tmp_path / "two_d.nrrd"
    open(p, "wb").write(b"NRRD0004\ntype: uchar\ndimension: 2\nsizes: 4 4\nencoding: raw\n\n" + bytes(16))
    with pytest.raises(vr.VrError) as e:
        vr.read_nrrd(str(p))
    assert "Invalid file properties" in str(e.value)  # nrrd_file_parser.cpp:28-30
    with pytest.raises(RuntimeError):
        ref.load_nrrd(str(p))
    with pytest.raises(vr.VrError):
        vr.read_nrrd(str(tmp_path / "missing.nrrd"))
    q = tmp_path / "short.nrrd"
    open(q, "wb").write(b"NRRD0004\ntype: uchar\ndimension: 3\nsizes: 4 4 4\nencoding: raw\n\n" + bytes(10))
    with pytest.raises(vr.VrError):
        vr.read_nrrd(str(q))
    z = tmp_path / "gz.nrrd"
    open(z, "wb").write(b"NRRD0004\ntype: uchar\ndimension: 3\nsizes: 2 2 2\nencoding: gzip\n\n" + bytes(8))
    with pytest.raises(vr.VrError):
        vr.read_nrrd(str(z))


# ---- partition helpers -----------------------------------------------------------------------------

_B = ["type: uchar", "dimension: 3", "sizes: 2 2 2", "encoding: raw"]
_P8 = bytes(range(1, 9))
NRRD_HEADER_CASES = [
    # (name, header lines after the magic, payload, magic): the product reader must accept exactly what NrrdIO accepts
    ("baseline", _B, _P8, "NRRD0004"),
    ("upper-case key and values", ["TYPE: UCHAR", "Dimension: 3", "sizes: 2 2 2", "encoding: RAW"], _P8, "NRRD0004"),
    ("unknown field", _B + ["foo: bar"], _P8, "NRRD0004"),
    ("line without colon", _B + ["garbage line"], _P8, "NRRD0004"),
    ("colon without space", ["type:uchar", "dimension:3", "sizes:2 2 2", "encoding:raw"], _P8, "NRRD0004"),
    ("space before the colon", ["type : uchar"] + _B[1:], _P8, "NRRD0004"),
    ("leading space", [" type: uchar"] + _B[1:], _P8, "NRRD0004"),
    ("trailing space in a value", ["type: uchar "] + _B[1:], _P8, "NRRD0004"),
    ("sizes before dimension", ["type: uchar", "sizes: 2 2 2", "dimension: 3", "encoding: raw"], _P8, "NRRD0004"),
    ("spacings before dimension", ["type: uchar", "spacings: 1 1 1", "dimension: 3", "sizes: 2 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("field given twice", _B + ["type: uchar"], _P8, "NRRD0004"),
    ("alias given twice", _B + ["old min: 0", "oldmin: 0"], _P8, "NRRD0004"),
    ("content twice", _B + ["content: a", "content: b"], _P8, "NRRD0004"),
    ("empty content", _B + ["content:"], _P8, "NRRD0004"),
    ("key/value pairs, comments", ["# c", "a:=b", ":=x"] + _B + ["k k:=v v", "# d"], _P8, "NRRD0001"),
    ("free-text fields", _B + ["number: abc", "min: abc", "max: x", "old max: 1.5", "sample units: HU"], _P8, "NRRD0002"),
    ("old min must be a number", _B + ["old min: x"], _P8, "NRRD0004"),
    ("aliases", _B + ["centerings: cell node ???", "axismins: 0 0 0", "axismaxs: 1 1 1", "lineskip: 0", "byteskip: 0", "sampleunits: mm"], _P8, "NRRD0004"),
    ("space origin before space", _B + ["space origin: (0,0,0)", "space: right-anterior-superior"], _P8, "NRRD0004"),
    ("space directions before space", _B + ["space directions: (1,0,0) (0,1,0) (0,0,1)", "space: RAS"], _P8, "NRRD0004"),
    ("space, origin, directions", _B + ["space: left-posterior-superior", "space origin: (0,0,0)", "space directions: none (0,1,0) (0,0,1)"], _P8, "NRRD0004"),
    ("space and space dimension", _B + ["space: RAS", "space dimension: 3"], _P8, "NRRD0004"),
    ("space dimension 2", _B + ["space dimension: 2", "space directions: (1,0) (0,1) (0,0)"], _P8, "NRRD0004"),
    ("direction with too few components", _B + ["space: RAS", "space directions: (1,0) (0,1) (0,0)"], _P8, "NRRD0004"),
    ("too few directions", _B + ["space: RAS", "space directions: (1,0,0) (0,1,0)"], _P8, "NRRD0004"),
    ("directions and spacings", _B + ["space: RAS", "space directions: (1,0,0) (0,1,0) (0,0,1)", "spacings: 1 1 1"], _P8, "NRRD0004"),
    ("space and spacings", _B + ["space: RAS", "spacings: 1 1 1"], _P8, "NRRD0004"),
    ("space with per-axis fields", _B + ["space: LPS", 'labels: "x" "y" "z"', 'units: "mm" "mm" "mm"', "axis mins: 0 0 0", "thicknesses: 1 1 1"], _P8, "NRRD0004"),
    ("directions and axis mins", _B + ["space: LPS", "space directions: (1,0,0) (0,1,0) (0,0,1)", "axis mins: 0 0 0"], _P8, "NRRD0004"),
    ("directions and axis maxs", _B + ["space: LPS", "space directions: (1,0,0) (0,1,0) (0,0,1)", "axis maxs: 0 0 0"], _P8, "NRRD0004"),
    ("directions and units", _B + ["space: LPS", "space directions: none (0,1,0) (0,0,1)", 'units: "mm" "mm" "mm"'], _P8, "NRRD0004"),
    ("directions and thicknesses", _B + ["space: LPS", "space directions: (1,0,0) (0,1,0) (0,0,1)", "thicknesses: 1 1 1"], _P8, "NRRD0004"),
    ("unknown space", _B + ["space: foo"], _P8, "NRRD0004"),
    ("space units without space", _B + ['space units: "mm" "mm" "mm"'], _P8, "NRRD0004"),
    ("measurement frame without space", _B + ["measurement frame: (1,0,0) (0,1,0) (0,0,1)"], _P8, "NRRD0004"),
    ("measurement frame", _B + ["space: RAS", "measurement frame: (1,0,0) (0,1,0) (0,0,1)"], _P8, "NRRD0004"),
    ("magic 0006", _B, _P8, "NRRD0006"),
    ("magic 0000", _B, _P8, "NRRD0000"),
    ("no encoding", _B[:3], _P8, "NRRD0004"),
    ("no type", _B[1:], _P8, "NRRD0004"),
    ("ushort raw without endian", ["type: ushort", "dimension: 3", "sizes: 2 2 1", "encoding: raw"], _P8, "NRRD0004"),
    ("ushort hex without endian", ["type: ushort", "dimension: 3", "sizes: 2 2 1", "encoding: hex"], b"0001000200030004\n", "NRRD0004"),
    ("ushort ascii without endian", ["type: ushort", "dimension: 3", "sizes: 2 2 2", "encoding: ascii"], b"1 2 3 4 5 6 7 8\n", "NRRD0004"),
    ("endian for uchar", _B + ["endian: little"], _P8, "NRRD0004"),
    ("bad endian", ["type: ushort", "dimension: 3", "sizes: 2 2 1", "encoding: raw", "endian: middle"], _P8, "NRRD0004"),
    ("two sizes", ["type: uchar", "dimension: 3", "sizes: 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("four sizes", ["type: uchar", "dimension: 3", "sizes: 2 2 2 1", "encoding: raw"], _P8, "NRRD0004"),
    ("zero size", ["type: uchar", "dimension: 3", "sizes: 2 0 2", "encoding: raw"], _P8, "NRRD0004"),
    ("negative size", ["type: uchar", "dimension: 3", "sizes: 2 -2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("signed size", ["type: uchar", "dimension: 3", "sizes: +2 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("fractional size and dimension", ["type: uchar", "dimension: 3.0", "sizes: 2.0 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("tab-separated sizes", ["type: uchar", "dimension: 03", "sizes: 2\t2\t2", "encoding: raw"], _P8, "NRRD0004"),
    ("dimension 17", ["type: uchar", "dimension: 17", "sizes: 2 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("dimension 0", ["type: uchar", "dimension: 0", "sizes: 2 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("two spacings", _B + ["spacings: 1 1"], _P8, "NRRD0004"),
    ("four spacings", _B + ["spacings: 1 1 1 1"], _P8, "NRRD0004"),
    ("nan spacing", _B + ["spacings: nan 1 1"], _P8, "NRRD0004"),
    ("non-numeric spacings", _B + ["spacings: a b c"], _P8, "NRRD0004"),
    ("two thicknesses", _B + ["thicknesses: 1 1"], _P8, "NRRD0004"),
    ("two axis mins", _B + ["axis mins: 0 0"], _P8, "NRRD0004"),
    ("unknown kind", _B + ["kinds: foo domain domain"], _P8, "NRRD0004"),
    ("kinds that fit", _B + ["kinds: 2-vector complex Domain"], _P8, "NRRD0004"),
    ("kind that does not fit", _B + ["kinds: 3-vector domain domain"], _P8, "NRRD0004"),
    ("stub needs size 1", _B + ["kinds: stub domain domain"], _P8, "NRRD0004"),
    ("unknown center", _B + ["centers: foo cell cell"], _P8, "NRRD0004"),
    ("two labels", _B + ['labels: "x" "y"'], _P8, "NRRD0004"),
    ("four labels", _B + ['labels: "x" "y" "z" "w"'], _P8, "NRRD0004"),
    ("unquoted labels", _B + ["labels: x y z"], _P8, "NRRD0004"),
    ("three labels and units", _B + ['labels: "x" "y y" "z"', 'units: "mm" "mm" "mm"'], _P8, "NRRD0004"),
    ("block type", ["type: block", "dimension: 3", "sizes: 2 2 2", "encoding: raw"], _P8, "NRRD0004"),
    ("block size for uchar", _B + ["block size: 4"], _P8, "NRRD0004"),
    ("gzip", _B[:3] + ["encoding: gzip"], _P8, "NRRD0004"),
    ("bzip2", _B[:3] + ["encoding: bzip2"], _P8, "NRRD0004"),
    ("negative line skip", _B + ["line skip: -1"], _P8, "NRRD0004"),
    ("byte skip -1", _B + ["byte skip: -1"], bytes(10) + _P8, "NRRD0004"),
    ("extra payload", _B, bytes([1]) * 20, "NRRD0004"),
    ("short payload", _B, bytes([1]) * 5, "NRRD0004"),
    ("missing data file", _B + ["data file: nothere.raw"], _P8, "NRRD0004"),
    ("ascii too few", _B[:3] + ["encoding: ascii"], b"1 2 3\n", "NRRD0004"),
    ("ascii too many", _B[:3] + ["encoding: ascii"], b"1 2 3 4 5 6 7 8 9 10\n", "NRRD0004"),
    ("ascii not a number", _B[:3] + ["encoding: ascii"], b"1 2 x 4 5 6 7 8\n", "NRRD0004"),
    ("ascii out of range, negative, fractional", _B[:3] + ["encoding: txt"], b"1 300 -3 4.5 5 6 7 8\n", "NRRD0004"),
    ("hex odd", _B[:3] + ["encoding: hex"], b"010203040506070\n", "NRRD0004"),
    ("hex not hex", _B[:3] + ["encoding: hex"], b"01020304050607zz\n", "NRRD0004"),
    ("hex with white space", _B[:3] + ["encoding: hex"], b"01 02 03 04\n05 06 07 08\n", "NRRD0004"),
]


@pytest.mark.parametrize("case", NRRD_HEADER_CASES, ids=[c[0] for c in NRRD_HEADER_CASES])
def test_nrrd_reader_accepts_exactly_what_nrrdio_accepts(vr, ref, tmp_path, case):
    """Header rules of NrrdIO (what NrrdFileParser::parse, nrrd_file_parser.cpp:23-30, lets through): field
    syntax, known fields, no field twice, ordering, per-axis counts, names, required fields, payload forms."""
    _, lines, payload, magic = case
    p = tmp_path / "h.nrrd"
    p.write_bytes((magic + "\n" + "\n".join(lines) + "\n\n").encode() + payload)
    try:
        want = ref.load_nrrd(str(p))[0]
    except RuntimeError:
        want = None
    try:
        got = vr.read_nrrd(str(p))
    except vr.VrError:
        got = None
    assert (got is None) == (want is None), "accepted" if got is not None else "rejected"
    if want is not None:
        assert got.shape == want.shape and np.array_equal(got.astype(np.float32), want)


def test_nrrd_reader_against_nrrdio_on_random_headers(vr, ref, tmp_path):
    """Seeded random headers -- every type name, encoding, version, optional fields in valid and invalid orders,
    both line endings: same verdict as the reference parser and, where it reads the file, the same values."""
    import random
    rng = random.Random(20260411)
    types = {"uchar": np.uint8, "unsigned char": np.uint8, "uint8": np.uint8, "uint8_t": np.uint8, "signed char": np.int8,
             "int8": np.int8, "short": np.int16, "int16": np.int16, "short int": np.int16, "ushort": np.uint16,
             "unsigned short": np.uint16, "uint16_t": np.uint16, "int": np.int32, "signed int": np.int32, "int32": np.int32,
             "uint": np.uint32, "unsigned int": np.uint32, "longlong": np.int64, "long long": np.int64, "int64": np.int64,
             "ulonglong": np.uint64, "unsigned long long": np.uint64, "uint64": np.uint64, "float": np.float32, "double": np.float64}
    optional = ["spacings: 1 2 3", "kinds: domain domain domain", "centers: cell cell cell", "content: hello world", "# a comment",
                "space: left-posterior-superior", 'labels: "x" "y" "z"', 'units: "mm" "mm" "mm"', "min: 0", "max: 5", "old min: 0",
                "thicknesses: 1 1 1", "axis mins: 0 0 0", "sample units: HU", "line skip: 0", "byte skip: 0", "key:=value", "a b:=c d",
                "space directions: (1,0,0) (0,1,0) (0,0,1)", "space origin: (0,0,0)"]
    accepted = rejected = 0
    for it in range(160):
        tname = rng.choice(list(types))
        dt = types[tname]
        nx, ny, nz = rng.randint(1, 5), rng.randint(1, 4), rng.randint(1, 3)
        arr = np.random.default_rng(it).integers(0, 100, (nz, ny, nx)).astype(dt)
        enc, endian = rng.choice(["raw", "ascii", "txt", "text", "hex"]), rng.choice(["little", "big"])
        fields = [f"type: {tname}", f"sizes: {nx} {ny} {nz}", f"encoding: {enc}"]
        if (arr.dtype.itemsize > 1 and enc in ("raw", "hex")) or rng.random() < 0.3:
            fields.append(f"endian: {endian}")
        fields += [o for o in optional if rng.random() < 0.2]
        rng.shuffle(fields)
        if rng.random() < 0.75:   # mostly valid: dimension first, space before the space vectors, no spacings with directions
            fields = ["dimension: 3"] + [f for f in fields if f.startswith("space:")] + [f for f in fields if not f.startswith("space:")]
            if any(f.startswith("space directions") for f in fields):
                fields = [f for f in fields if not f.startswith("spacings")]
            if not any(f.startswith("space:") for f in fields):
                fields = [f for f in fields if not f.startswith("space ")]
        else:
            fields.insert(rng.randint(0, len(fields)), "dimension: 3")
        eol = rng.choice(["\n", "\r\n"])
        data = arr.astype(arr.dtype.newbyteorder("<" if endian == "little" else ">"))
        payload = (data.tobytes() if enc == "raw" else data.tobytes().hex().encode() if enc == "hex" else
                   ("\n".join(" ".join(repr(x.item()) for x in row) for row in arr.reshape(-1, nx)) + "\n").encode())
        p = tmp_path / f"r{it}.nrrd"
        p.write_bytes((rng.choice(["NRRD0001", "NRRD0003", "NRRD0004", "NRRD0005"]) + eol + eol.join(fields) + eol + eol).encode() + payload)
        try:
            want = ref.load_nrrd(str(p))[0]
        except RuntimeError:
            want = None
        try:
            got = vr.read_nrrd(str(p))
        except vr.VrError:
            got = None
        assert (got is None) == (want is None), (it, fields)
        if want is not None:
            accepted += 1
            assert got.shape == want.shape and np.array_equal(got.astype(np.float32), want), (it, fields)
        else:
            rejected += 1
    assert accepted >= 60 and rejected >= 15, (accepted, rejected)


def write_csv_stack(tmp_path, vol, fmt="%.9g", eol="\n", last_eol=True, name="s"):
    paths = []
    for z, sl in enumerate(vol):
        text = eol.join(",".join(fmt % v for v in row) for row in sl) + (eol if last_eol else "")
        p = tmp_path / f"{name}{z:03d}.csv"
        p.write_bytes(text.encode())
        paths.append(str(p))
    return paths


@pytest.mark.parametrize("eol,last_eol", [("\n", True), ("\n", False), ("\r\n", True)])
def test_csv_reader_matches_reference_parser(vr, ref, tmp_path, eol, last_eol):
    """src/data/csv_file_parser.cpp:14-50 compiled as it lies (oracle/_ref) against the product's reader."""
    rng = np.random.default_rng(5)
    vol = rng.normal(3.0, 50.0, (4, 5, 7)).astype(np.float32)
    vol[1, 2, 3] = 1e-3
    paths = write_csv_stack(tmp_path, vol, eol=eol, last_eol=last_eol)
    got = vr.read_csv(paths)
    want, mn, mx = ref.load_csv(paths)
    assert got.data.shape == want.shape == vol.shape
    assert np.array_equal(bits(got.data), bits(want))
    assert np.array_equal(bits(got.data), bits(vol))  # %.9g round-trips a float
    assert (got.min, got.max) == (mn, mx)
    # all-positive data: the reference's minimum is still 0 (value-initialised Dataset)
    pos = np.abs(vol) + 1.0
    q = write_csv_stack(tmp_path, pos, name="p")
    g2, (w2, mn2, mx2) = vr.read_csv(q), ref.load_csv(q)
    assert g2.min == mn2 == 0.0 and g2.max == mx2 == float(pos.max())


def test_csv_reader_field_and_line_rules(vr, ref, tmp_path):
    """Trailing separators, blanks, exponents, inf/nan, trailing characters: whatever std::getline + std::stof accept."""
    a, b = tmp_path / "a.csv", tmp_path / "b.csv"
    a.write_bytes(b" 1.5, -2e3 ,0x10,\n7abc,nan,inf,\n")
    b.write_bytes(b"1,2,3\n4,5,6")
    got = vr.read_csv([str(a), str(b)])
    want, mn, mx = ref.load_csv([str(a), str(b)])
    assert got.data.shape == want.shape == (2, 2, 3)
    assert np.array_equal(bits(got.data), bits(want))
    assert np.array_equal(bits(np.float32([got.min, got.max])), bits(np.float32([mn, mx])))
    assert got.data[0, 0].tolist() == [1.5, -2000.0, 16.0] and got.data[0, 1, 0] == 7.0 and np.isnan(got.data[0, 1, 1])
    # no files at all / an unreadable file: zero extents, no error (the reference never checks the stream)
    for paths in ([], [str(tmp_path / "missing.csv")]):
        g, (w, _, _) = vr.read_csv(paths), ref.load_csv(paths)
        assert g.data.shape == w.shape and g.data.size == 0


@pytest.mark.parametrize("files,message", [
    ([b"1,2,3\n4,5\n"], "Inconsistant dimensions"),               # ragged line
    ([b"1,2\n3,4\n", b"1,2\n"], "Inconsistant dimensions"),       # slice with fewer lines
    ([b"1,2\n3,4\n", b"1,2\n3,4\n5,6\n"], "Inconsistant dimensions"),
    ([b"1,2\n\n3,4\n"], "Inconsistant dimensions"),               # blank line = zero fields
    ([b"1,,3\n"], "stof"),                                         # empty field
    ([b"1,x,3\n"], "stof"),
    ([b"1,1e99,3\n"], "stof"),                                     # out of float range
    ([b"1,2\r\n\r\n"], "stof"),                                   # a line holding only a carriage return
])
def test_csv_reader_rejects_what_the_reference_rejects(vr, ref, tmp_path, files, message):
    paths = []
    for i, content in enumerate(files):
        p = tmp_path / f"e{i}.csv"
        p.write_bytes(content)
        paths.append(str(p))
    with pytest.raises(RuntimeError) as want:
        ref.load_csv(paths)
    with pytest.raises(vr.VrError) as got:
        vr.read_csv(paths)
    assert message in str(want.value) and message in str(got.value)
    assert str(want.value) in str(got.value)  # the text the reference's error popup would show


def test_csv_reader_against_reference_parser_on_random_text(vr, ref, tmp_path):
    """Seeded random slice stacks -- grids of plausible fields with noise, and plain random text over the CSV
    alphabet: same verdict, same error text, same bits and the same min / max as the reference parser."""
    import random
    rng = random.Random(4711)
    alphabet = "0123456789" * 3 + ".,,,,-+eE \n\n\n\r\tx" + "infa"
    tokens = ["1", "-2.5", "3e2", " 4", "5 ", "0x1p3", "nan", "inf", "1e-50", "7abc", ".5", "-.5e1", "+3"]
    accepted = 0
    for it in range(400):
        paths = []
        for z in range(rng.randint(1, 3)):
            if it % 2:
                nx, ny = 1 + it % 4, 1 + (it // 4) % 3
                rows = [",".join(rng.choice(tokens) for _ in range(nx)) for _ in range(ny)]
                text = rng.choice(["\n", "\r\n"]).join(rows) + rng.choice(["", "\n", "\n\n", ","])
                if rng.random() < 0.15:
                    text = text.replace(",", ",,", 1)
            else:
                text = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 30)))
            p = tmp_path / f"f{it}_{z}.csv"
            p.write_bytes(text.encode())
            paths.append(str(p))
        try:
            want, werr = ref.load_csv(paths), None
        except RuntimeError as e:
            want, werr = None, str(e)
        try:
            got, gerr = vr.read_csv(paths), None
        except vr.VrError as e:
            got, gerr = None, str(e)
        assert (got is None) == (want is None), (it, paths)
        if want is None:
            assert werr in gerr
        else:
            accepted += 1
            assert got.data.shape == want[0].shape and np.array_equal(bits(got.data), bits(want[0]))
            assert np.array_equal(bits(np.float32([got.min, got.max])), bits(np.float32([want[1], want[2]])))
    assert 40 <= accepted <= 360


def test_partition_boxes_tile_the_grid(vr):
    for grid, full in [((2, 2, 2), (64, 64, 64)), ((3, 2, 1), (50, 33, 7)), ((1, 1, 1), (5, 5, 5)), ((2, 1, 1), (7, 3, 3))]:
        seen = np.zeros(full[::-1], np.int32)
        for b in range(grid[0] * grid[1] * grid[2]):
            org, dims = vr.partition_box(grid, full, b)
            seen[org[2]:org[2] + dims[2], org[1]:org[1] + dims[1], org[0]:org[0] + dims[0]] += 1
        assert (seen == 1).all()
    with pytest.raises(vr.VrError):
        vr.partition_box((2, 2, 2), (64, 64, 64), 8)


def test_partition_order_is_a_valid_front_to_back_order(vr):
    """For random cameras and random rays from them: boxes are visited in an order consistent with out[]."""
    rng = np.random.default_rng(17)
    for grid in [(2, 2, 2), (3, 2, 1), (4, 1, 2)]:
        full = (60, 60, 60)
        nb = grid[0] * grid[1] * grid[2]
        for _ in range(20):
            cam = rng.uniform(-3, 3, 3).astype(np.float32)
            order = vr.partition_order(grid, full, cam).tolist()
            assert sorted(order) == list(range(nb))
            rank_of = {b: k for k, b in enumerate(order)}
            for _ in range(50):
                target = rng.uniform(-0.5, 0.5, 3)
                d = target - cam
                ts = np.linspace(0, 2, 400)
                pts = cam[None, :] + ts[:, None] * d[None, :] + 0.5  # texture space
                inside = np.all((pts >= 0) & (pts < 1), axis=1)
                visited = []
                for q in pts[inside]:
                    idx = [max(k for k in range(grid[a]) if q[a] >= (k * full[a] // grid[a]) / full[a]) for a in range(3)]
                    b = idx[0] + grid[0] * (idx[1] + grid[1] * idx[2])
                    if not visited or visited[-1] != b:
                        visited.append(b)
                ranks = [rank_of[b] for b in visited]
                assert ranks == sorted(ranks), (grid, cam, visited, order)


# ---- C++ host side over the ABI -------------------------------------------------------------------

def test_cpp_mirror_builds_and_fails_loudly_without_a_device(vr):
    """host/offscreen_pass.hpp (the C++ mirror of the reference's classes) compiles against include/vr.h
    and links to the library; without a GPU the pass constructor throws like the reference would."""
    import subprocess
    import torch
    pkg = os.path.join(ROOT, "volumetric-renderer_b200")
    subprocess.run(["make", "-s", "-C", pkg, os.path.join(pkg, "host", "example_orbit"), os.path.join(pkg, "host", "example_group")], check=True)
    exe = os.path.join(pkg, "host", "example_orbit")
    assert os.path.exists(exe) and os.path.exists(os.path.join(pkg, "host", "example_group"))
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "/tmp/_vr_none.ppm", "32", "32", "8", "1"], capture_output=True, text=True)
        assert r.returncode == 1 and "no CPU fallback" in r.stderr


def test_partition_balance_planes(vr):
    """vr_partition_balance / vr_partition_stored_box / vr_partition_box_planes (host only): a camera on an axis of symmetry
    leaves the planes of the other axes in the middle, the plane of its own axis moves TOWARDS the camera (the near half
    holds more samples), never further than the margin; margins that would let two planes meet are refused."""
    grid, full = (2, 2, 2), (256, 256, 256)
    pl = vr.partition_balance(grid, full, (0.0, -3.0, 0.0), (40, 40, 40))
    assert list(pl) == [128, pl[1], 128] and 128 - 40 <= pl[1] < 128
    assert pl[1] == 107                                   # 1/r^2 between 2.5 and 3.5: the median sits 0.083 of the edge nearer
    pl = vr.partition_balance(grid, full, (3.0, 0.0, 0.0), (10, 10, 10))
    assert list(pl) == [138, 128, 128]                    # clamped to the margin
    assert list(vr.partition_balance(grid, full, (0.0, -3.0, 0.0), (0, 0, 0))) == [128, 128, 128]
    far = vr.partition_balance(grid, full, (0.0, -300.0, 0.0), (40, 40, 40))
    assert abs(int(far[1]) - 128) <= 1                    # parallel projection in the limit: nothing to balance
    g3 = (3, 1, 2)
    pl3 = vr.partition_balance(g3, (300, 64, 128), (-2.0, -2.0, 1.0), (20, 0, 9))
    assert len(pl3) == 3 and pl3[0] < pl3[1] and abs(int(pl3[0]) - 100) <= 20 and abs(int(pl3[1]) - 200) <= 20 and abs(int(pl3[2]) - 64) <= 9
    boxes = [vr.partition_box_planes(g3, (300, 64, 128), pl3, i) for i in range(6)]
    vol = np.zeros((128, 64, 300), np.int32)
    for o, d in boxes:
        vol[o[2]:o[2] + d[2], o[1]:o[1] + d[1], o[0]:o[0] + d[0]] += 1
    assert (vol == 1).all()
    assert vr.partition_box_planes(g3, (300, 64, 128), None, 4) == vr.partition_box(g3, (300, 64, 128), 4)
    for i in range(6):
        so, sd = vr.partition_stored_box(g3, (300, 64, 128), i, (20, 0, 9))
        ro, rd = vr.partition_box(g3, (300, 64, 128), i)
        for a in range(3):
            assert so[a] == max(ro[a] - (20, 0, 9)[a], 0) if ro[a] > 0 else so[a] == 0
            assert so[a] + sd[a] == (ro[a] + rd[a] + (20, 0, 9)[a] if ro[a] + rd[a] < (300, 64, 128)[a] else (300, 64, 128)[a])
    with pytest.raises(vr.VrError):
        vr.partition_stored_box(grid, (16, 16, 16), 0, (4, 4, 4))     # 2 * 4 + 1 > 16 / 2
    # the visiting order follows the planes: a camera inside the cube between the regular and the moved plane
    o_reg = vr.partition_order((2, 1, 1), (256, 1, 1), (0.05, 0.0, 0.0))
    o_mov = vr.partition_order((2, 1, 1), (256, 1, 1), (0.05, 0.0, 0.0), planes=[160])
    assert list(o_reg) == [1, 0] and list(o_mov) == [0, 1]


def test_nrrd_spacings_are_read_but_not_applied_by_themselves(vr, tmp_path):
    """the reference ignores `spacings` / `space directions` (nrrd_file_parser.cpp:32-33); the reader reports
    them (NaN where the header states none) for a caller who wants them honoured (vr_set_volume_spacing)."""
    body = bytes(2 * 3 * 4)
    p = tmp_path / "a.nrrd"
    p.write_bytes(b"NRRD0004\ntype: uchar\ndimension: 3\nsizes: 2 3 4\nspacings: 0.5 1.25 3\nencoding: raw\n\n" + body)
    assert list(vr.nrrd_spacing(str(p))) == [0.5, 1.25, 3.0]
    p.write_bytes(b"NRRD0004\ntype: uchar\ndimension: 3\nsizes: 2 3 4\nencoding: raw\n\n" + body)
    assert np.isnan(vr.nrrd_spacing(str(p))).all()
    p.write_bytes(b"NRRD0004\ntype: uchar\ndimension: 3\nspace dimension: 3\nsizes: 2 3 4\n"
                  b"space directions: (3,4,0) (0,0.5,0) (0,0,-2)\nencoding: raw\n\n" + body)
    assert list(vr.nrrd_spacing(str(p))) == [5.0, 0.5, 2.0]
    p.write_bytes(b"NRRD0004\ntype: uchar\ndimension: 3\nsizes: 2 3 4\nspacings: nan 2 nan\nencoding: raw\n\n" + body)
    s = vr.nrrd_spacing(str(p))
    assert np.isnan(s[0]) and s[1] == 2.0 and np.isnan(s[2])
    with pytest.raises(vr.VrError):
        vr.nrrd_spacing(str(tmp_path / "missing.nrrd"))


def test_sort_last_merge_region(vr):
    """vr_sort_last_region (host only): the first frame into an image slot merges the whole frame; later ones the bounding
    rectangle of this frame's and the previous frame's sub-box rectangles; empty rectangles are ignored."""
    slot = np.zeros(4, np.uint32)
    a = [(100, 50, 300, 200), (250, 150, 400, 260), (0, 0, 0, 0)]
    assert vr.sort_last_region(a, 640, 480, True, slot) == (0, 0, 640, 480)
    assert tuple(slot) == (100, 50, 400, 260)
    b = [(120, 40, 200, 100), (130, 60, 220, 300)]
    assert vr.sort_last_region(b, 640, 480, False, slot) == (100, 40, 400, 300)     # union with what the slot held
    assert tuple(slot) == (120, 40, 220, 300)
    assert vr.sort_last_region([(5, 5, 5, 9)], 640, 480, False, slot) == (120, 40, 220, 300)   # nothing drawn now: restore the old region
    assert tuple(slot) == (0, 0, 0, 0)
    assert vr.sort_last_region([(0, 0, 0, 0)], 640, 480, False, slot) == (0, 0, 0, 0)          # nothing before, nothing now
    assert vr.sort_last_region(b, 640, 480, False, slot) == (120, 40, 220, 300)


def test_abi_header_is_plain_c(tmp_path):
    """include/vr.h is the drop-in boundary: it must compile as C99 on its own (no C++, no CUDA, no torch types in it)."""
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "vr.h"\nint main(void) { vr_options o; vr_frame_info f; vr_sort_last_wiring w; (void)o; (void)f; (void)w; return VR_ABI_VERSION == 2 ? 0 : 1; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    import re
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "vr.h")).read(), flags=re.S)   # declarations only
    for word in ("torch", "cudaStream_t", "std::", "Tensor", "#include <cuda"):
        assert word not in hdr, word
