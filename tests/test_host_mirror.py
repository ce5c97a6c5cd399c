"""Builds and runs tests/cpp/test_host_mirror.cpp: the reference's serving/speed unit tests replayed in
C++ against the host-side mirror (oryx_b200/host), which reaches the GPU only through the C ABI."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "oryx_b200", "host")
EXE = os.path.join(ROOT, "tests", "cpp", "test_host_mirror")


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oryx_b200", "csrc"), "-j", str(os.cpu_count() or 4)])
    subprocess.check_call(["make", "-s", "-C", HOST])
    src = os.path.join(ROOT, "tests", "cpp", "test_host_mirror.cpp")
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(src), os.path.getmtime(os.path.join(ROOT, "oryx_b200", "liboryx_host.so"))):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-I" + HOST, "-I" + os.path.join(ROOT, "include"), src, "-o", EXE,
                               "-L" + os.path.join(ROOT, "oryx_b200"), "-loryx_host", "-loryx_b200",
                               "-Wl,-rpath," + os.path.join(ROOT, "oryx_b200"), "-pthread"])


def test_host_mirror_builds_and_links():
    """CPU check: the C++ mirror compiles against include/oryx_b200.h and links to the library."""
    _build()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_host_mirror_replays_reference_tests():
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0, r.stdout[-3000:]
