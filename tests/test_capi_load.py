"""CPU-side checks of the drop-in boundary: libsvo_gpu.so loads and exports every symbol include/svo_gpu.h declares;
argument validation that needs no device."""
import ctypes as C
import os
import re

import pytest

from svo_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return capi.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "svo_gpu.h")).read()
    declared = set(re.findall(r"SVO_GPU_API\s+[\w\s\*]+?\b(svo_gpu_\w+)\s*\(", hdr))
    assert len(declared) >= 18
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None


def test_abi_version(lib):
    assert lib.svo_gpu_abi_version() == 1


def test_struct_layouts_match_header():
    # svo_gpu_step_params: 12 ints/floats + 2 doubles with natural alignment
    assert C.sizeof(capi.StepParams) == 80
    assert capi.StepParams.epsilon.offset == 48 and capi.StepParams.min_eig_threshold.offset == 64
    assert C.sizeof(capi.StepOutputs) == 120 and capi.StepParams.carry_tracks.offset == 72
    # svo_gpu_keyframe_params: 9 ints/floats, pad, double, 2 ints, double, float, pad
    assert C.sizeof(capi.KeyframeParams) == 72
    assert capi.KeyframeParams.epsilon.offset == 40 and capi.KeyframeParams.min_eig_threshold.offset == 56
    assert capi.KeyframeParams.y_threshold.offset == 64


def test_create_without_gpu_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.SvoGpuError) as e:
        capi.Context()
    assert e.value.code in (capi.E_CUDA, capi.E_ARG)       # no device: an error, never a CPU fallback


def test_null_context_is_rejected(lib):
    assert lib.svo_gpu_upload(None, 0, None, 1, 1, 1) == capi.E_ARG
    assert lib.svo_gpu_step_reset(None) == capi.E_ARG
    assert lib.svo_gpu_filter_tracks(None, None, None, None, 0, -1.0, None, None, None, 0, None) == capi.E_ARG
    assert lib.svo_gpu_keyframe(None, 0, 1, None, 0, None, None, None, 0, None, None) == capi.E_ARG
    assert lib.svo_gpu_destroy(None) == 0


def test_lk_kernel_does_not_spill(lib):
    """The 21x21 LK kernel runs at its 96-register cap (20 one-warp CTAs per SM); a change that makes ptxas spill costs ~5 %
    (measured: 1.02 -> 1.08 ms per 128 k tracks for a 16-byte spill), so the build is checked for it."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    out = subprocess.run(["cuobjdump", "-res-usage", capi.LIB_PATH], capture_output=True, text=True).stdout
    lines = out.splitlines()
    hits = [lines[i + 1] for i, l in enumerate(lines) if "klt_run_kernelILi21ELi21" in l and i + 1 < len(lines)]
    assert hits, "21x21 LK kernel not found in the library"
    for h in hits:
        assert "STACK:0" in h and "REG:96" in h, h
