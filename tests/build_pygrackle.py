#!/usr/bin/env python
"""Build the REFERENCE's own Python binding (pygrackle: grackle_wrapper.pyx + grackle_defs.pxd,
src/python/pygrackle/) against THIS repository's headers and CUDA library - test infrastructure.

    python tests/build_pygrackle.py          ->  tests/bin/pygrackle_build/pygrackle/

The reference sources are used where they lie (/root/reference, dev container only): the .pyx is
cythonized and compiled unmodified with -I include/ and linked to grackle_b200/lib/
libgrackle_b200.so; the pure-Python modules the wrapper needs at run time (fluid_container,
utilities.{misc,physical_constants,units,convenience}) are placed beside it in the SCRATCH build
directory, which is git-ignored (tests/bin/) but travels to the GPU box like every other build
product.  The package's own __init__.py is not used: it imports yt_fields / evolve, which need yt
(not in this image); the scratch __init__ exposes FluidContainer and chemistry_data only.
Nothing of this is part of the product; tests/test_gpu_pygrackle.py drives the CUDA path through it.
"""
import os
import shutil
import subprocess
import sys
import sysconfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/src/python/pygrackle"
OUT = os.path.join(ROOT, "tests", "bin", "pygrackle_build")
PKG = os.path.join(OUT, "pygrackle")


def build():
    if not os.path.isdir(REF):
        print("build_pygrackle: %s absent; keeping any prebuilt wrapper" % REF)
        return os.path.isdir(PKG)
    import numpy
    from Cython.Build import cythonize  # noqa: F401  (fails early if Cython is missing)
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    os.makedirs(os.path.join(PKG, "utilities"))
    # run-time Python modules of the binding, as they are
    for rel in ("fluid_container.py", "utilities/__init__.py", "utilities/misc.py", "utilities/physical_constants.py",
                "utilities/units.py", "utilities/convenience.py"):
        src = os.path.join(REF, rel)
        if os.path.exists(src):
            shutil.copy(src, os.path.join(PKG, rel))
        else:
            open(os.path.join(PKG, rel), "w").close()
    with open(os.path.join(PKG, "__init__.py"), "w") as f:
        f.write("# scratch package: the reference's wrapper built against grackle-b200 (tests/build_pygrackle.py)\n"
                "import numpy as _np\n"
                "if not hasattr(_np, 'VisibleDeprecationWarning'):      # the reference's utilities/misc.py predates numpy 2\n"
                "    _np.VisibleDeprecationWarning = _np.exceptions.VisibleDeprecationWarning\n"
                "from .fluid_container import FluidContainer  # noqa: F401\n"
                "from .grackle_wrapper import chemistry_data  # noqa: F401\n")
    # cythonize a copy of the two Cython sources inside the scratch dir (cython writes its .c beside the .pyx)
    for name in ("grackle_wrapper.pyx", "grackle_defs.pxd"):
        shutil.copy(os.path.join(REF, name), os.path.join(PKG, name))
    subprocess.run([sys.executable, "-m", "cython", "-3", "--module-name", "pygrackle.grackle_wrapper",
                    os.path.join(PKG, "grackle_wrapper.pyx")], check=True, cwd=OUT)
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    so = os.path.join(PKG, "grackle_wrapper" + ext)
    libdir = os.path.join(ROOT, "grackle_b200", "lib")
    cmd = ["gcc", "-shared", "-fPIC", "-O1", "-w", "-DCONFIG_BFLOAT_8", "-DLARGE_INTS",
           "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
           "-I" + os.path.join(ROOT, "include"), "-I" + numpy.get_include(), "-I" + sysconfig.get_paths()["include"],
           os.path.join(PKG, "grackle_wrapper.c"), "-o", so,
           "-L" + libdir, "-lgrackle_b200", "-Wl,-rpath,$ORIGIN/../../../../grackle_b200/lib"]
    subprocess.run(cmd, check=True)
    for junk in ("grackle_wrapper.c", "grackle_wrapper.pyx", "grackle_defs.pxd"):
        os.remove(os.path.join(PKG, junk))
    print("build_pygrackle: built", so)
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
