"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol that
include/oryx_b200.h declares, and fails loudly (no CPU fallback) when no CUDA device is usable."""
import os
import re

import pytest

import oryx_b200
from oryx_b200 import native
from als_harness import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "oryx_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\bint32_t\s+(oryx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = native.lib()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/oryx_b200.h but not exported"
    # and the binding covers exactly the header
    assert sorted(native.SIGNATURES) == declared


def test_lsh_plan_goldens_through_c_abi():
    # LocalitySensitiveHashTest.java:36-70,120-124 via oryx_lsh_plan (host integer arithmetic)
    for rate, cores, hashes, bits in load_golden("lsh")["plan"]["cases"]:
        assert oryx_b200.lsh_plan(rate, cores) == (hashes, bits), (rate, cores)


def test_bad_arguments_are_rejected():
    with pytest.raises(oryx_b200.OryxError) as ei:
        oryx_b200.lsh_plan(0.0, 4)  # Preconditions.checkArgument(sampleRate > 0.0 ...) ALSServingModel.java:99
    assert ei.value.status == native.E_INVALID
    with pytest.raises(oryx_b200.OryxError) as ei:
        oryx_b200.Model(0)  # features > 0, ALSServingModel.java:98
    assert ei.value.status == native.E_INVALID


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(oryx_b200.OryxError) as ei:
        oryx_b200.Model(8)
    assert ei.value.status == native.E_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "oryx_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "oryx_oracle" not in text and "liboryx_oracle" not in text, f
