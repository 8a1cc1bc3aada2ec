"""The C-ABI boundary (include/tfode_batch.h) without a GPU: the library loads, exports every declared symbol,
struct layouts agree with the ctypes mirror, presets agree, and compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from tf_ode_b200 import abi, engine, make_config

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "tfode_batch.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tfode_\w+)\s*\(", src)))


def test_header_declares_what_the_host_binds():
    assert set(header_functions()) == set(engine.EXPORTS)


def test_library_exports_every_declared_symbol():
    L = engine.load_library()
    for name in header_functions():
        assert hasattr(L, name), name


def test_struct_sizes_match_compiled_library():
    L = engine.load_library()
    assert L.tfode_sizeof_config() == C.sizeof(abi.BatchConfig)
    assert L.tfode_sizeof_state() == C.sizeof(abi.EnvState)
    import oracle_lib
    O = oracle_lib.lib()
    assert O.orc_sizeof_config() == C.sizeof(abi.BatchConfig) and O.orc_sizeof_state() == C.sizeof(abi.EnvState)


@pytest.mark.parametrize("preset", [0, 1, 2, 3])
def test_presets_agree_between_c_and_python(preset):
    L = engine.load_library()
    c = abi.BatchConfig()
    assert L.tfode_batch_config_default(C.byref(c), preset) == 0
    p = make_config(preset, 1)
    for name, _ in abi.BatchConfig._fields_:
        if name in ("hf_heights", "hf_pos"):
            continue
        assert getattr(c, name) == getattr(p, name), name
    assert L.tfode_batch_config_default(C.byref(c), 9) == -1


def test_invalid_arguments_are_rejected_before_touching_the_device():
    L = engine.load_library()
    h = C.c_void_p()
    cfg = make_config(abi.PRESET_TRAINING, 0)
    assert L.tfode_batch_create(C.byref(cfg), C.byref(h)) == -1 and b"n_envs" in L.tfode_last_error(None)
    cfg = make_config(abi.PRESET_TRAINING, 4)
    cfg.abi_version = 99
    assert L.tfode_batch_create(C.byref(cfg), C.byref(h)) == -1 and b"abi_version" in L.tfode_last_error(None)
    cfg = make_config(abi.PRESET_MT_RANDOM, 4, scene=abi.SCENE_HEIGHTFIELD)
    assert L.tfode_batch_create(C.byref(cfg), C.byref(h)) == -1 and b"heightfield" in L.tfode_last_error(None)
    assert L.tfode_batch_step(None, 1) == -1


def test_no_cpu_fallback_create_fails_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(engine.TfodeError):
        engine.RoverBatch(make_config(abi.PRESET_TRAINING, 4))


def test_product_package_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tf_ode_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in txt and "liboracle" not in txt and "rover_oracle" not in txt, f
