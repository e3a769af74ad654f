"""Build recipe for the checker libraries (TEST INFRASTRUCTURE ONLY).

* oracle/libibp_oracle.so         — the C++ restatement (oracle/pafprocess_oracle.cpp)
* oracle/_ref/libpafprocess_ref.so — the UNMODIFIED reference utils/pafprocess/pafprocess.cpp compiled where it
  lies under /root/reference plus the extern "C" forwarders of oracle/ref_shim.cpp.  Built only when the
  reference tree is present (this container); the GPU box uses the prebuilt file that travels with the snapshot.

Flags follow SURVEY.md §8(c): g++ -O2 -std=c++14 -fPIC -shared -ffp-contract=off, no -march.
Nothing is copied from the reference; outputs are git-ignored.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC_DIR = "/root/reference/utils/pafprocess"
CXXFLAGS = ["-O2", "-std=c++14", "-fPIC", "-shared", "-ffp-contract=off"]

ORACLE_SO = os.path.join(HERE, "libibp_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpafprocess_ref.so")


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def build_oracle(force=False):
    src = [os.path.join(HERE, "pafprocess_oracle.cpp")]
    if force or not _newer(ORACLE_SO, src):
        subprocess.check_call(["g++"] + CXXFLAGS + src + ["-o", ORACLE_SO])
    return ORACLE_SO


def build_ref(force=False):
    """Returns the path of the reference .so, or None when it can neither be built nor found."""
    ref_cpp = os.path.join(REF_SRC_DIR, "pafprocess.cpp")
    shim = os.path.join(HERE, "ref_shim.cpp")
    if os.path.exists(ref_cpp):
        if force or not _newer(REF_SO, [ref_cpp, shim]):
            os.makedirs(os.path.dirname(REF_SO), exist_ok=True)
            subprocess.check_call(["g++"] + CXXFLAGS + ["-I", REF_SRC_DIR, ref_cpp, shim, "-o", REF_SO])
        return REF_SO
    return REF_SO if os.path.exists(REF_SO) else None


if __name__ == "__main__":
    print(build_oracle(force="--force" in sys.argv))
    print(build_ref(force="--force" in sys.argv))
