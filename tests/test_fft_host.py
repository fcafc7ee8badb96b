"""The per-thread FFT code of the CUDA kernel (whisper_rs_b200/csrc/logmel_fft.cuh), compiled for the host
with g++ and checked against a float64 DFT -- no GPU needed."""
import os
import subprocess
import sys

from conftest import ROOT


def test_fft_routines_against_float64_dft(tmp_path):
    exe = str(tmp_path / "fft_check")
    subprocess.run(["g++", "-O2", "-o", exe, os.path.join(ROOT, "tests", "host", "fft_check.cpp")], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    sys.stdout.write(r.stdout)
    assert r.returncode == 0, r.stdout + r.stderr
