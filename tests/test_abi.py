"""The C-ABI library loads and exports every symbol include/vo2w.h declares (no GPU needed)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from visual_onoma_to_wave_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "vo2w.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vo2w_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = header_functions()
    assert len(names) >= 14
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/vo2w.h but not exported"
    assert sorted(_lib.SYMBOLS) == names, "ctypes binding and header disagree"
    assert _lib.lib().vo2w_version() >= 100


def test_header_is_plain_c():
    src = open(os.path.join(ROOT, "include", "vo2w.h")).read()
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    assert "torch" not in code.lower() and "at::" not in code and "std::" not in code
    assert set(re.findall(r"#include\s*[<\"]([^>\"]+)", code)) == {"stddef.h", "stdint.h"}
    assert 'extern "C"' in src


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_compute_fails_loudly_without_a_device():
    lib = _lib.lib()
    assert lib.vo2w_device_check() == _lib.E_NODEVICE
    cfg = _lib.GenConfig()
    cfg.num_upsamples, cfg.num_kernels, cfg.num_mels, cfg.upsample_initial_channel = 4, 3, 80, 512
    w = np.zeros(16, dtype=np.float32)
    h = C.c_void_p()
    rc = lib.vo2w_generator_create(C.byref(cfg), w.ctypes.data, w.size, _lib.PREC_BF16, C.byref(h))
    assert rc == _lib.E_NODEVICE
    with pytest.raises(RuntimeError, match="no CUDA device|not sm_100"):
        _lib.check(rc)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_python_mirrors_refuse_cpu_tensors():
    import visual_onoma_to_wave_b200 as v
    from visual_onoma_to_wave_b200.synthetic import V1_CONFIG
    g = v.Generator(v.AttrDict(V1_CONFIG))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        g(torch.zeros(1, 80, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        v.LengthRegulator()(torch.zeros(1, 2, 4), torch.ones(1, 2), None)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "visual_onoma_to_wave_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f
