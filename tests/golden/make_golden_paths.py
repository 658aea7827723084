"""Golden vectors of the reference's path helpers (ecoz2/src/utl/fileutil.c:14-63, 97-119), produced
by calling get_class_name / get_output_filename of the reference itself (oracle/_ref, compiled from
/root/reference/ecoz2/src by oracle/Makefile).  Run in the build container:
    python tests/golden/make_golden_paths.py      ->  tests/golden/path_helpers.json
"""
import ctypes as C
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbuild  # noqa: E402

PATHS = [
    "a/b/c.prd", "c.prd", "b/c.prd", "/abs/dir/cls/x.y.prd", "dir.v2/cls/name", "x/y/z.p", "x/y/z.",
    "x/y/.hidden", "x/y/noext", "data/predictors/TRAIN/B/from_MARS_20161221_000046_SongSession_16kHz_HPF5Hz.wav__1.prd",
    "p/" + "k" * 120 + "/f.prd", "./cls/f.prd", "cls//f.prd", "a/b/c.tar.gz", "a/b/",
]
EXTS = [".seq", "seq", "", ".", ".cbook"]
BASES = ["sequences", "sequences/TRAIN/M0256", "codebooks"]


def main():
    L = refbuild.lib(256)
    L.get_class_name.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    L.get_output_filename.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p]
    out = {"class_name": [], "output_filename": []}
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            for p in PATHS:
                buf = C.create_string_buffer(96)
                L.get_class_name(p.encode(), buf, 96)
                out["class_name"].append([p, buf.value.decode()])
            for p in PATHS:
                for e in EXTS:
                    for b in BASES[:1] if e not in (".seq", ".cbook") else BASES:
                        src = C.create_string_buffer(p.encode(), 2048)
                        res = C.create_string_buffer(8192)
                        L.get_output_filename(src, b.encode(), e.encode(), res)
                        made = os.path.isdir(os.path.dirname(res.value.decode()))
                        out["output_filename"].append([p, b, e, res.value.decode(), made])
        finally:
            os.chdir(cwd)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "path_helpers.json"), "w") as f:
        json.dump(out, f, indent=0)
    print(len(out["class_name"]), "class names,", len(out["output_filename"]), "output names")


if __name__ == "__main__":
    main()
