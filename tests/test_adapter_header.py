"""include/mgb200_dealii.h (the adapters a maintainer drops beside the reference's include/operator.h:11-557 and
multigrid_throughput.cc:847-883, 1600-1604, 1800-1817) is real C++: it is compiled here against stub declarations of the
deal.II members it uses (tests/stubs/deal.II; deal.II itself is not installed) with every adapter class instantiated."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_adapter_header_compiles_against_the_stub_declarations():
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "stubs"),
           os.path.join(ROOT, "tests", "stubs", "adapter_check.cc")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
