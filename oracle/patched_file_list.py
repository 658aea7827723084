"""TEST INFRASTRUCTURE ONLY.  Applies integration/build.rs.patch to a scratch copy of the reference's
build.rs (/root/reference/build.rs:8-78) and prints the C files the PATCHED build script would hand to
cc::Build, one absolute path per line -- the static `ecoz2_lib` of the Rust crate minus the VQ hot path.
Used by oracle/Makefile (target `linkcheck`) to show that this list plus libecoz2_vq_b200.so links."""
import os
import re
import shutil
import subprocess
import sys
import tempfile

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
PATCH = os.path.join(os.path.dirname(HERE), "integration", "build.rs.patch")


def patched_build_rs() -> str:
    with tempfile.TemporaryDirectory() as td:
        shutil.copy(os.path.join(REF_ROOT, "build.rs"), os.path.join(td, "build.rs"))
        subprocess.run(["patch", "-p1", "-s", "-i", PATCH], cwd=td, check=True)
        with open(os.path.join(td, "build.rs")) as f:
            return f.read()


def file_list(text: str):
    out = []
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("//"):
            continue
        m = re.match(r'"(ecoz2/src/[^"]+\.c)",', s)
        if m:
            out.append(os.path.join(REF_ROOT, m.group(1)))
    return out


if __name__ == "__main__":
    if len(sys.argv) > 1:
        REF_ROOT = sys.argv[1]
    print("\n".join(file_list(patched_build_rs())))
