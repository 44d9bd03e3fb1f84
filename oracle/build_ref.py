"""Build the UNMODIFIED reference extension ``remixt/bpmodel.pyx`` into ``oracle/_ref/``.

TEST INFRASTRUCTURE ONLY.  Nothing under ``remixt_b200/`` may import this.

The reference source is compiled *where it lies* under ``/root/reference`` (read-only):
Cython writes the generated C++ into ``oracle/_ref/`` (git-ignored) and g++ produces
``oracle/_ref/remixt/bpmodel.<abi>.so``.  No reference source is copied into the repo.

The reference's own ``setup.py`` is bypassed (it hard-pins numpy==2.4.4, ``setup.py:7-14``);
the one extension on the hot path is built directly with the interpreter's default CFLAGS
(-O2 on this image; reference ``setup.py:34-35`` only adds ``-g``).

Usage:  python oracle/build_ref.py [--force]
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("REMIXT_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
PKG = os.path.join(OUT, "remixt")


def ref_so_path():
    return os.path.join(PKG, "bpmodel" + sysconfig.get_config_var("EXT_SUFFIX"))


def build(force=False, verbose=True):
    """Returns the path of the built extension, or None when the reference tree is absent."""
    so = ref_so_path()
    if os.path.exists(so) and not force:
        return so
    pyx = os.path.join(REF_ROOT, "remixt", "bpmodel.pyx")
    if not os.path.exists(pyx):
        return None
    import numpy
    os.makedirs(PKG, exist_ok=True)
    cpp = os.path.join(OUT, "bpmodel.cpp")
    cmd = [sys.executable, "-m", "cython", "--cplus", "-3", "-o", cpp, pyx]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    cflags = (sysconfig.get_config_var("CFLAGS") or "-O2").split()
    cmd = (["g++", "-shared", "-fPIC", "-g"] + cflags +
           ["-I" + sysconfig.get_paths()["include"], "-I" + numpy.get_include(),
            "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
            cpp, "-o", so])
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    with open(os.path.join(OUT, "BUILD_INFO.txt"), "w") as f:
        f.write("source: %s\ncython: %s\ncflags: %s\n" % (
            pyx, __import__("Cython").__version__, " ".join(cflags)))
    return so


if __name__ == "__main__":
    p = build(force="--force" in sys.argv)
    print("reference extension:", p)
