"""The per-thread selection and geometry code of the projection kernel (whisper_rs_b200/csrc/decode_proj_select.cuh),
compiled for the host with g++ and checked against brute force -- mask ranges, the branch-free best-K insertion with
floors and ties, the K-th largest of the group maxima, vocab tile bounds, compact row counts.  No GPU needed."""
import os
import subprocess
import sys

from conftest import ROOT


def test_projection_select_code_against_brute_force(tmp_path):
    exe = str(tmp_path / "proj_select_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "host", "proj_select_check.cpp")], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    sys.stdout.write(r.stdout)
    assert r.returncode == 0, r.stdout + r.stderr
