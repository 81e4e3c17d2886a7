"""Host builds of header-only pieces of the library, run as small native programs (tests/native/*.cpp):
  * csrc/word_acc.h — the per-thread bit accumulator and sequence walk of the pack kernels — against a bit-by-bit
    MSB-first writer (BitBuffer::WriteBits, BitBuffer.cpp:78-110; writeGolombCode, :228-241) and the slow definition of the walk;
  * csrc/host_pool.h — exceptions thrown by jobs reach the caller, never std::terminate, the pool stays usable;
  * csrc/spin_team.h — the decoder's spinning helper team under oversubscription: no index of a job is ever lost."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_and_run(tmp_path, name, flags, token):
    if not shutil.which("g++"):
        pytest.skip("no g++")
    exe = str(tmp_path / name)
    subprocess.run(["g++", "-std=c++20", "-O1", *flags, "-o", exe, os.path.join(ROOT, "tests", "native", name + ".cpp")], check=True, timeout=300)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and token in r.stdout, r.stdout + r.stderr


def test_word_accumulator_equals_bitwise_writer(tmp_path):
    build_and_run(tmp_path, "word_acc_test", ["-fsanitize=address,undefined"], "word_acc OK")


def test_host_pool_survives_throwing_jobs(tmp_path):
    build_and_run(tmp_path, "host_pool_test", ["-pthread"], "host_pool OK")


def test_spin_team_oversubscribed(tmp_path):
    build_and_run(tmp_path, "spin_team_test", ["-pthread"], "spin_team OK")
