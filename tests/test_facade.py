"""The C++14 facade (libear_b200/include/ear/*.hpp): libear's class names and signatures over the
C ABI.  CPU: it compiles as C++14 with -Wall -Wextra -Werror, libear's own example program
compiles against it unmodified (when /root/reference is present) and, without a device, fails
loudly.  GPU: the in-process self test (tests/cpp/facade_test.cpp) and the Objects golden vectors
through GainCalculatorObjects::calculate<double>."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "facade_test.cpp")
BUILD = os.path.join(ROOT, "tests", "cpp", "build")
EXE = os.path.join(BUILD, "facade_test")
FLAGS = ["-std=c++14", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "libear_b200", "include"),
         "-I", os.path.join(ROOT, "include")]
LINK = ["-L", os.path.join(ROOT, "libear_b200"), "-learb200", "-Wl,-rpath," + os.path.join(ROOT, "libear_b200")]


def build_exe():
    deps = [SRC, os.path.join(ROOT, "include", "earb200.h")]
    inc = os.path.join(ROOT, "libear_b200", "include", "ear_b200")
    deps += [os.path.join(inc, f) for f in os.listdir(inc)]
    deps.append(os.path.join(ROOT, "libear_b200", "libearb200.so"))
    if os.path.exists(EXE) and all(os.path.getmtime(EXE) >= os.path.getmtime(d) for d in deps):
        return EXE
    os.makedirs(BUILD, exist_ok=True)
    subprocess.run(["g++"] + FLAGS + [SRC] + LINK + ["-o", EXE], check=True)
    return EXE


def test_facade_compiles_as_cxx14():
    build_exe()


def test_reference_example_compiles_unmodified(tmp_path):
    example = "/root/reference/examples/objects_gains.cpp"
    if not os.path.exists(example):
        pytest.skip("reference tree not present")
    subprocess.run(["g++"] + FLAGS + [example] + LINK + ["-o", str(tmp_path / "objects_gains")], check=True)


def test_facade_fails_loudly_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([build_exe()], capture_output=True, text=True)
    assert r.returncode != 0 and "no CUDA device" in r.stdout + r.stderr


@pytest.mark.gpu
def test_facade_selftest():
    r = subprocess.run([build_exe()], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "facade selftest: ok" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["0+5+0", "9+10+3"])
def test_facade_objects_golden(layout):
    g = np.load(os.path.join(ROOT, "tests", "golden", "objects.npz"))
    rows = np.stack(list(g[f"{layout}/in"]), axis=1)[:200]
    text = "\n".join(" ".join(repr(float(v)) for v in row) for row in rows) + "\n"
    r = subprocess.run([build_exe(), "objects", layout], input=text, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = np.array([[float(v) for v in line.split()] for line in r.stdout.strip().splitlines()])
    n_out = out.shape[1] // 2
    tol = 1e-12 if layout == "0+5+0" else 1e-5
    assert np.abs(out[:, :n_out] - g[f"{layout}/direct64"][:200]).max() < tol
    assert np.abs(out[:, n_out:] - g[f"{layout}/diffuse64"][:200]).max() < tol
