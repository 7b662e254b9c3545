#!/usr/bin/env python
"""Compile and link the REFERENCE's own example programs, unchanged, against this repository's
headers and CUDA library (what src/python/tests/test_code_examples.py:41 does upstream with the
reference's library) - test infrastructure.

    python tests/build_examples.py      ->  tests/bin/examples/<name>[_omp]

The sources are compiled where they lie (/root/reference/src/example, dev container only).  Every
example is built with and without -fopenmp: chemistry_data carries an `omp_nthreads` tail member
under _OPENMP, so sizeof(chemistry_data) differs between the two kinds of caller and both must
work against the one library.  cxx_omp_example.C includes <omp.h> and is built with
-fopenmp only.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/src/example"
OUT = os.path.join(ROOT, "tests", "bin", "examples")
EXAMPLES = [("c_example.c", "gcc", (False, True)), ("cxx_example.C", "g++", (False, True)),
            ("cxx_grid_example.C", "g++", (False, True)), ("cxx_omp_example.C", "g++", (True,))]


def command(src, compiler, omp, exe):
    libdir = os.path.join(ROOT, "grackle_b200", "lib")
    return ([compiler, "-O1", "-w", "-DCONFIG_BFLOAT_8"] + (["-fopenmp"] if omp else []) +
            ["-I" + os.path.join(ROOT, "include"), os.path.join(REF, src), "-o", exe,
             "-L" + libdir, "-lgrackle_b200", "-lm", "-Wl,-rpath,$ORIGIN/../../../grackle_b200/lib"])


def build(out=OUT):
    """returns {exe name: path}; raises CalledProcessError when an example does not compile or link"""
    if not os.path.isdir(REF):
        print("build_examples: %s absent; keeping any prebuilt examples" % REF)
        return {}
    os.makedirs(out, exist_ok=True)
    built = {}
    for src, compiler, kinds in EXAMPLES:
        for omp in kinds:
            name = os.path.splitext(src)[0] + ("_omp" if omp and len(kinds) > 1 else "")
            exe = os.path.join(out, name)
            subprocess.run(command(src, compiler, omp, exe), check=True)
            built[name] = exe
    return built


if __name__ == "__main__":
    for k, v in build().items():
        print(k, "->", v)
