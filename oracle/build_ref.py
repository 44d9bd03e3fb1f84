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


# the reference's own Python modules on the fit path: staged verbatim NEXT TO the compiled extension in
# the git-ignored oracle/_ref/ so that the reference arm of bench.py and the drop-in test run the UNMODIFIED
# reference driver on the GPU box too, where /root/reference does not exist.  Build output, never
# committed, never imported by remixt_b200/.
PY_FILES = ["cn_model.py", "likelihood.py", "utils.py", "config.py", "defaults.py", "_version.py",
            "analysis/__init__.py", "analysis/pipeline.py", "analysis/experiment.py", "analysis/readdepth.py"]


def stage_py(verbose=True):
    """Copies PY_FILES from the reference tree into oracle/_ref/remixt/ (no-op without the reference tree)."""
    import shutil
    src_pkg = os.path.join(REF_ROOT, "remixt")
    if not os.path.isdir(src_pkg):
        return False
    for rel in PY_FILES:
        src, dst = os.path.join(src_pkg, rel), os.path.join(PKG, rel)
        if not os.path.exists(src):
            continue
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or open(src, "rb").read() != open(dst, "rb").read():
            shutil.copyfile(src, dst)
            if verbose:
                print("staged", rel)
    return True


def build(force=False, verbose=True):
    """Returns the path of the built extension, or None when the reference tree is absent."""
    so = ref_so_path()
    if os.path.exists(so) and not force:
        stage_py(verbose)
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
    stage_py(verbose)
    return so


if __name__ == "__main__":
    p = build(force="--force" in sys.argv)
    print("reference extension:", p)
