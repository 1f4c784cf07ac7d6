"""
Recipe that compiles the reference's two Cython extensions FROM WHERE THEY LIE
under /root/reference into oracle/_ref/ (git-ignored; travels to the GPU box).
Test infrastructure only.  No reference source is copied into the repository:
Cython reads the .pyx/.pxd in place and only the derived shared objects are
kept.  Reference build recipe being replaced: source/core/setup.py:28-45.

Usage: python oracle/build_ref.py [--reference /root/reference]
Returns exit code 0 and prints the built files, or prints why it could not.
"""
import argparse
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODULES = ("common_functions", "perturbation_functions")


def build(reference="/root/reference", verbose=True):
    src_dir = os.path.join(reference, "source", "core", "functions")
    if not os.path.isdir(src_dir):
        if verbose:
            print("reference tree not present at", reference,
                  "- keeping whatever is already in", OUT)
        return False
    import numpy
    os.makedirs(OUT, exist_ok=True)
    suffix = sysconfig.get_config_var("EXT_SUFFIX")
    py_inc = sysconfig.get_paths()["include"]
    built = []
    for mod in MODULES:
        pyx = os.path.join(src_dir, mod + ".pyx")
        c_file = os.path.join(OUT, mod + ".c")
        so = os.path.join(OUT, mod + suffix)
        if os.path.exists(so) and os.path.getmtime(so) >= os.path.getmtime(pyx):
            built.append(so)
            continue
        subprocess.check_call([sys.executable, "-m", "cython", "-3", pyx,
                               "-o", c_file])
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-w",
                               "-I" + py_inc, "-I" + numpy.get_include(),
                               c_file, "-o", so])
        os.remove(c_file)  # derived from reference source: do not keep
        built.append(so)
    if verbose:
        print("\n".join(built))
    return True


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    a = ap.parse_args()
    sys.exit(0 if build(a.reference) else 1)
