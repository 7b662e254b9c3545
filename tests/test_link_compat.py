"""Link compatibility with the reference's own callers (upstream:
src/python/tests/test_code_examples.py:41 builds every example).  CPU part: the examples compile and
link unchanged against include/ + libgrackle_b200.so, with and without -fopenmp, and the header
gives both kinds of caller the struct sizes the reference's header gives them.  GPU part: the
reference's c_example / cxx_example binaries RUN on the CUDA path (the table they open by relative
path is the synthetic HM2012-schema file).
"""
import os

import numpy as np
import shutil
import subprocess

import pytest

import build_examples

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.isdir(build_examples.REF)

_SIZES = r"""
#include <stdio.h>
#include <grackle.h>
int main(void) { printf("%zu %zu %zu\n", sizeof(chemistry_data), sizeof(grackle_field_data), sizeof(code_units)); return 0; }
"""


@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not available")
def test_reference_examples_compile_and_link_unchanged(tmp_path):
    built = build_examples.build(str(tmp_path))
    assert set(built) == {"c_example", "c_example_omp", "cxx_example", "cxx_example_omp", "cxx_grid_example",
                          "cxx_grid_example_omp", "cxx_omp_example"}
    # (the binaries carry an rpath relative to tests/bin/examples; in this scratch directory the loader is told)
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "grackle_b200", "lib"))
    for name, exe in built.items():
        out = subprocess.run(["ldd", exe], capture_output=True, text=True, env=env).stdout
        assert "libgrackle_b200.so" in out and "not found" not in out, (name, out)
        # without a GPU (here) and without their data file the programs must fail cleanly, not crash
        rc = subprocess.run([exe], cwd=str(tmp_path), capture_output=True, text=True, timeout=120, env=env)
        assert rc.returncode in (0, 1), (name, rc.returncode, rc.stderr[-300:])


@pytest.mark.skipif(not HAVE_REF, reason="reference checkout not available")
@pytest.mark.parametrize("omp", [False, True])
def test_struct_sizes_equal_the_reference_header_for_both_kinds_of_caller(tmp_path, omp):
    src = tmp_path / "sizes.c"
    src.write_text(_SIZES)
    sizes = []
    for inc in ([os.path.join(ROOT, "include")],
                [os.path.join(ROOT, "oracle", "_ref"), "/root/reference/src/clib"]):      # _ref holds grackle_float.h
        exe = tmp_path / ("sizes_%d" % len(sizes))
        subprocess.run(["gcc", "-DLINUX"] + (["-fopenmp"] if omp else []) + ["-I" + d for d in inc] +
                       [str(src), "-o", str(exe)], check=True)
        sizes.append(subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split())
    assert sizes[0] == sizes[1], sizes


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c_example", "c_example_omp", "cxx_example"])
def test_reference_example_runs_on_the_cuda_path(tmp_path, synth_table, name):
    exe = os.path.join(build_examples.OUT, name)
    if not os.path.exists(exe):
        pytest.skip("examples not built (python tests/build_examples.py, dev container)")
    # the examples open "../../input/CloudyData_UVB=HM2012.h5" relative to their working directory
    (tmp_path / "input").mkdir()
    shutil.copy(synth_table, str(tmp_path / "input" / "CloudyData_UVB=HM2012.h5"))
    cwd = tmp_path / "src" / "example"
    cwd.mkdir(parents=True)
    run = subprocess.run([exe], cwd=str(cwd), capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, run.stdout[-500:] + run.stderr[-500:]
    # the examples report on stderr: "Cooling time = ... s.", "Temperature = ... K.", "Pressure = ...",
    # "gamma = ...", "dust_temperature = ... K." (c_example.c:197-244, cxx_example.C:204-251)
    out = run.stdout + run.stderr
    vals = {}
    for ln in out.splitlines():
        for key in ("Cooling time", "Temperature", "Pressure", "gamma", "dust_temperature"):
            if ln.strip().startswith(key + " ="):
                vals[key] = float(ln.split("=")[1].split()[0].rstrip("."))
    assert set(vals) == {"Cooling time", "Temperature", "Pressure", "gamma", "dust_temperature"}, out[-1500:]
    assert all(np.isfinite(v) for v in vals.values()), vals
    # the example's cell: 1 cm^-3 gas at ~1000 K after one solve step, dust_chemistry = 1
    assert 1.0 < vals["gamma"] < 1.7 and 1.0 < vals["Temperature"] < 1e5 and vals["Pressure"] > 0.0, vals
    assert vals["Cooling time"] != 0.0 and 2.7 < vals["dust_temperature"] < 1e3, vals
