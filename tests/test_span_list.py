"""CPU: the free list of the exchange's receive arena (ballista_b200/csrc/span_list.h: offers the largest free span,
takes blocks from its start, coalesces on release) against a byte map, 200 k random takes / gives."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_span_list_against_a_byte_map(tmp_path):
    exe = str(tmp_path / "span_list_test")
    src = os.path.join(ROOT, "tests", "cpp", "span_list_test.cpp")
    inc = os.path.join(ROOT, "ballista_b200", "csrc")
    subprocess.run(["g++", "-std=c++17", "-O2", "-I", inc, src, "-o", exe, "-pthread"], check=True)
    p = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "span_list: all ok" in p.stdout, p.stdout + p.stderr
