"""csrc/word_acc.h — the per-thread bit accumulator of the MNENC_PACK=2 pack kernel — built for the host and checked against a
bit-by-bit MSB-first writer (BitBuffer::WriteBits, BitBuffer.cpp:78-110; writeGolombCode, :228-241) on random codes."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_word_accumulator_equals_bitwise_writer(tmp_path):
    if not shutil.which("g++"):
        pytest.skip("no g++")
    exe = str(tmp_path / "word_acc_test")
    subprocess.run(["g++", "-std=c++20", "-O1", "-fsanitize=address,undefined", "-o", exe,
                    os.path.join(ROOT, "tests", "native", "word_acc_test.cpp")], check=True, timeout=300)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "word_acc OK" in r.stdout, r.stdout + r.stderr
