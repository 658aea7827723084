"""CPU tests of the host side: the C-ABI library loads and exports every symbol the
header declares, refuses to compute without a GPU (no CPU fallback), and its file
formats / path rules are byte-compatible with the reference (golden .cbook / .seq
bytes written by the reference itself, tests/golden/make_golden.py)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def test_library_exports_every_declared_symbol(built_lib):
    from ecoz2rs_b200 import vq
    with open(os.path.join(ROOT, "include", "ecoz2_vq_b200.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(ecoz2b200_[a-z0-9_]+|ecoz2_vq_[a-z_]+|vq_(?:learn|learn_using_base_codebook|quantize|classify)|(?<=double )quantize)\s*\(", header))
    declared -= {"ecoz2b200_stage_cb", "ecoz2b200_pass_cb"}
    assert declared == set(vq.EXPORTED_SYMBOLS), declared ^ set(vq.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(built_lib, name), name


def test_no_cpu_fallback(built_lib):
    """Without a CUDA device the engine cannot be created and says why."""
    from ecoz2rs_b200 import vq
    if vq.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(vq.VqError, match="no CUDA device"):
        vq.VqEngine(36)
    # the drop-in entry points fail with the CUDA error code instead of computing on the CPU
    import tempfile
    from ecoz2rs_b200 import formats
    with tempfile.TemporaryDirectory() as tmp:
        prd = os.path.join(tmp, "TRAIN", "c", "x.prd")
        formats.write_prd(prd, np.ones((4, 13)), "c")
        cwd = os.getcwd()
        os.chdir(tmp)
        try:
            assert vq.vq_learn(12, 0.05, "c", [prd]) == 3
        finally:
            os.chdir(cwd)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under ecoz2rs_b200/ may reference it."""
    pkg = os.path.join(ROOT, "ecoz2rs_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for fn in files:
            if fn.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                with open(os.path.join(dirpath, fn), errors="ignore") as f:
                    src = f.read()
                assert "import oracle" not in src and "from oracle" not in src, fn
                assert "vq_oracle" not in src and "oracle/_ref" not in src, fn
    # the dev / measurement scripts do not use it either (checkers that need it live in tests/tools/)
    for fn in os.listdir(os.path.join(ROOT, "scripts")):
        if fn.endswith((".py", ".sh")):
            with open(os.path.join(ROOT, "scripts", fn), errors="ignore") as f:
                src = f.read()
            assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, fn


def test_formats_roundtrip_and_reference_bytes(built_lib, tmp_path):
    from ecoz2rs_b200 import formats
    L = built_lib
    g = np.load(os.path.join(GOLD, "learn_P36.npz"))
    refl = g["refl_4"]
    ref_bytes = g["cbook_bytes_stage2"].tobytes()
    # C++ writer == the reference's cb_save bytes
    p = str(tmp_path / "a.cbook").encode()
    assert L.ecoz2b200_cbook_save(p, b"gold", refl.ctypes.data_as(C.POINTER(C.c_double)), 4, 36) == 0
    assert open(p, "rb").read() == ref_bytes
    # numpy writer too
    formats.write_cbook(str(tmp_path / "b.cbook"), refl, "gold")
    assert open(tmp_path / "b.cbook", "rb").read() == ref_bytes
    # C++ reader
    Pp, M = C.c_int(), C.c_int()
    cls = C.create_string_buffer(96)
    assert L.ecoz2b200_cbook_info(p, C.byref(Pp), C.byref(M), cls) == 0
    assert (Pp.value, M.value, cls.value) == (36, 4, b"gold")
    out = np.empty((4, 37))
    assert L.ecoz2b200_cbook_read(p, out.ctypes.data_as(C.POINTER(C.c_double)), out.size) == 0
    assert np.array_equal(out, refl)
    # .seq: C++ writer == bytes the reference's vq.quantize wrote
    q = np.load(os.path.join(GOLD, "quantize_P36.npz"))
    seq = np.ascontiguousarray(q["seq"], np.uint16)
    sp = str(tmp_path / "x.seq").encode()
    assert L.ecoz2b200_seq_save(sp, b"gold", seq.ctypes.data_as(C.POINTER(C.c_uint16)), len(seq),
                                q["refl"].shape[0]) == 0
    assert open(sp, "rb").read() == q["seq_file_bytes"].tobytes()
    # .prd: numpy writer -> C++ reader
    x = g["x"][:17]
    pp = str(tmp_path / "TRAIN" / "cls" / "f.prd")
    formats.write_prd(pp, x, "cls")
    T = C.c_int()
    assert L.ecoz2b200_prd_info(pp.encode(), C.byref(Pp), C.byref(T), cls) == 0
    assert (Pp.value, T.value, cls.value) == (36, 17, b"cls")
    back = np.empty_like(x)
    assert L.ecoz2b200_prd_read(pp.encode(), back.ctypes.data_as(C.POINTER(C.c_double)), back.size) == 0
    assert np.array_equal(back, x)
    # errors: missing file, wrong ident, too small a buffer
    assert L.ecoz2b200_prd_info(b"/nonexistent.prd", C.byref(Pp), C.byref(T), cls) != 0
    assert L.ecoz2b200_prd_info(p, C.byref(Pp), C.byref(T), cls) != 0
    assert L.ecoz2b200_prd_read(pp.encode(), back.ctypes.data_as(C.POINTER(C.c_double)), 5) != 0


def test_class_name_header_padding(tmp_path):
    """96-byte class field: name, NUL, then '_' padding (fileutil.c:162-182)."""
    from ecoz2rs_b200 import formats
    p = tmp_path / "c.cbook"
    formats.write_cbook(str(p), np.zeros((1, 3)), "ab")
    raw = open(p, "rb").read()
    assert raw[:16] == b"<codebook>" + b"\0" * 6
    assert raw[16:16 + 96] == b"ab\0" + b"_" * 93
    assert raw[112:120] == np.array([2, 1], "<i4").tobytes()
    assert len(raw) == 120 + 3 * 8


def test_search_kernels_are_tcgen05_code(built_lib):
    """The shipped library holds sm_100a machine code whose search kernels issue tcgen05 MMAs
    (UTCHMMA), read their accumulators from tensor memory (LDTM), stage operands with bulk copies
    (UBLKCP) and hand over through tensor-core barriers (UTCBAR) -- the mnemonics
    /opt/skills/guides/B200_PROFILING.md names as the proof.  No GPU needed: cuobjdump reads the
    .so the tests load."""
    import shutil
    import subprocess
    from ecoz2rs_b200 import vq
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not installed")
    elf = subprocess.run([exe, "-lelf", vq.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "sm_100a" in elf
    sass = subprocess.run([exe, "-sass", "-fun", "vq_assign_tc2", vq.LIB_PATH], capture_output=True, text=True)
    if sass.returncode != 0 or "UTCHMMA" not in sass.stdout:      # (-fun wants mangled names on some versions)
        sass = subprocess.run([exe, "-sass", vq.LIB_PATH], capture_output=True, text=True, check=True)
    for op in ("UTCHMMA", "LDTM", "UBLKCP", "UTCBAR"):
        assert sass.stdout.count(op) > 0, op


def test_path_helpers_match_reference_fileutil(built_lib, tmp_path, monkeypatch):
    """get_class_name / get_output_filename / camext (ecoz2/src/utl/fileutil.c:14-63, 97-119) decide
    where every `.cbook` and `.seq` lands.  Golden outputs of the reference's own functions
    (tests/golden/make_golden_paths.py through oracle/_ref): no parent directory, dots in directory
    names (camext takes the LAST dot of the whole path), extensions shorter than two characters,
    empty / dot-only new extensions, doubled slashes, class names longer than the 96-byte field --
    and the directories the call creates."""
    import json
    L = built_lib
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "path_helpers.json")))
    monkeypatch.chdir(tmp_path)
    for path, want in gold["class_name"]:
        buf = C.create_string_buffer(b"\xff" * 95, 96)
        assert L.ecoz2b200_class_name_of(path.encode(), buf) == 0
        assert buf.value.decode() == want, path
    for path, base, ext, want, made in gold["output_filename"]:
        out = C.create_string_buffer(4096)
        assert L.ecoz2b200_output_filename(path.encode(), base.encode(), ext.encode(), out, 4096) == 0
        assert out.value.decode() == want, (path, base, ext)
        assert os.path.isdir(os.path.dirname(want)) == made, want
    small = C.create_string_buffer(8)
    assert L.ecoz2b200_output_filename(b"a/b/c.prd", b"sequences", b".seq", small, 8) != 0


def test_min_dists_csv_writer_matches_printf(built_lib, tmp_path):
    """The parallel "<cbook>.min_dists.csv" writer reproduces fprintf("%d,%f\\n")
    of reference report.c:118-120 byte for byte: random magnitudes, exact ties of the 6th decimal
    (k/128, k/2^20: round-half-to-even), negatives, subnormals, huge values, inf and nan."""
    import struct
    from ecoz2rs_b200 import vq
    rng = np.random.default_rng(7)
    n = 300_000
    md = np.concatenate([
        rng.random(n) * 10.0 ** rng.integers(-9, 6, size=n),
        rng.integers(0, 2 ** 24, size=5000) / 128.0,          # x 10^6 ends in .5 or .0: ties
        rng.integers(0, 2 ** 30, size=5000) / 2.0 ** 20,
        -rng.random(2000),
        np.array([0.0, -0.0, 5e-324, 1e-310, 0.5e-6, 1.5e-6, 2.5e-6, 0.9999995, 0.99999949999,
                  1e15, 123456789012345.678, 9.2e18, 1e19, 1e300, np.inf, -np.inf, np.nan]),
        # 317-character lines, more of them than a formatter thread's line buffer was sized for
        np.full(40_000, -1.7976931348623157e308), rng.random(30_000) * 1e300,
    ])
    cw = rng.integers(0, 65536, size=md.shape[0]).astype(np.int32)
    cw[:3] = [0, -1, 2 ** 31 - 1]
    path = tmp_path / "x.min_dists.csv"
    vq.write_min_dists_csv(str(path), cw, md)
    got = path.read_text().split("\n")
    assert got[0] == "codeword,minDistortion" and got[-1] == "" and len(got) == md.shape[0] + 2
    want = ["%d,%f" % (int(c), float(x)) for c, x in zip(cw, md)]
    bad = [i for i, (a, b) in enumerate(zip(got[1:-1], want)) if a != b]
    assert not bad, f"{len(bad)} lines differ, first: {got[1 + bad[0]]!r} vs {want[bad[0]]!r} ({md[bad[0]]!r})"


def test_cli_orders_files_like_rust_pathbuf(tmp_path):
    """`list.sort()` on Vec<PathBuf> (src/utl/mod.rs:222,244) compares path COMPONENTS: "a/x.prd" comes
    before "a-b/x.prd" although '-' < '/' as bytes.  The order is the frame order of `vq learn`."""
    import subprocess
    import __graft_entry__ as g
    g.build()
    exe = os.path.join(ROOT, "ecoz2rs_b200", "csrc", "ecoz2")
    names = ["a/x.prd", "a-b/x.prd", "a/b/y.prd", "a.b/x.prd", "a/x-1.prd", "a/x.1.prd", "B/z.prd", "a+/q.prd"]
    for n in names:
        p = tmp_path / "predictors" / n
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_bytes(b"")
    r = subprocess.run([exe, "vq", "files", "--predictors", "predictors"], cwd=tmp_path, capture_output=True,
                       text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    got = [ln[len("predictors/"):] for ln in r.stdout.split()]
    want = sorted(names, key=lambda s: [c.encode() for c in s.split("/")])
    assert got == want
    assert got.index("a/x.prd") < got.index("a-b/x.prd") and sorted(names) != want
