"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/spmm_b200.h declares, validates arguments, and FAILS LOUDLY without a GPU (no CPU
fallback).  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, random_csr


def _header_symbols():
    syms = set()
    for name in sorted(os.listdir(os.path.join(ROOT, "include"))):      # spmm_b200.h, spmm_b200_pipeline.h
        if not name.endswith(".h"):
            continue
        src = open(os.path.join(ROOT, "include", name)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        syms |= set(re.findall(r"\b(spmm_b200_[a-z_0-9]+)\s*\(", src))
    return sorted(syms)


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load_library()
    syms = _header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/*.h but not exported"
    assert sorted(pkg.EXPORTS) == syms


def test_headers_are_plain_c99():
    """The boundary is a C ABI: every header under include/ compiles as C99 (no C++, no torch, no CUDA types)."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("gcc not on PATH")
    for name in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if name.endswith(".h"):
            r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                                os.path.join(ROOT, "include", name)], capture_output=True, text=True)
            assert r.returncode == 0, r.stderr
            code = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", name)).read(), flags=re.S)
            assert "torch" not in code and "cudaStream_t" not in code and "std::" not in code


def test_library_is_in_tree_and_sm100(pkg):
    assert os.path.realpath(pkg.LIB_PATH).startswith(os.path.realpath(ROOT))
    blob = open(pkg.LIB_PATH, "rb").read()
    assert b"sm_100a" in blob or b"sm_100" in blob


def test_strerror_and_codes(pkg):
    lib = pkg.load_library()
    assert lib.spmm_b200_strerror(0) == b"ok"
    for code in range(-6, 0):
        assert lib.spmm_b200_strerror(code) not in (b"ok", b"unknown error")
    assert lib.spmm_b200_strerror(-99) == b"unknown error"


def test_argument_validation(pkg):
    lib = pkg.load_library()
    ia, ja, a = random_csr(10, 10, 2.0, 0)
    h = C.c_void_p()
    rp, ci, va = ia.ctypes.data, ja.ctypes.data, a.ctypes.data
    assert lib.spmm_b200_create(None, rp, ci, va, 10, 10, len(ja), 1, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create(C.byref(h), None, ci, va, 10, 10, len(ja), 1, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create(C.byref(h), rp, ci, va, -1, 10, len(ja), 1, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create(C.byref(h), rp, ci, va, 10, 10, len(ja), 0, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create(C.byref(h), rp, ci, va, 10, 10, len(ja), 1, 0, 17, 0) == -1
    assert lib.spmm_b200_create(C.byref(h), rp, ci, va, 10, 10, 2 ** 31, 1, 0, pkg.AUTO, 0) == -5
    assert lib.spmm_b200_run(None, None, None, 1, 1) == -1
    assert lib.spmm_b200_partition(rp, 10, len(ja), 0, pkg.PART_NNZ, None, 0) == -1
    assert lib.spmm_b200_rows(None) == -1
    lib.spmm_b200_destroy(None)   # must be a no-op
    # pipeline entry points (include/spmm_b200_pipeline.h)
    assert lib.spmm_b200_pipeline_create(None, rp, ci, va, 10, len(ja), 8, 0) == -1
    assert lib.spmm_b200_pipeline_create(C.byref(h), None, ci, va, 10, len(ja), 8, 0) == -1
    assert lib.spmm_b200_pipeline_create(C.byref(h), rp, ci, va, 10, len(ja), 0, 0) == -1
    assert lib.spmm_b200_pipeline_create(C.byref(h), rp, ci, va, 10, 2 ** 31, 8, 0) == -5
    assert lib.spmm_b200_pipeline_sddmm(None, None, 0) == -1
    assert lib.spmm_b200_pipeline_final(None, None) == -1
    assert lib.spmm_b200_pipeline_attention(None, None, None, 0, 0) == -1
    assert lib.spmm_b200_pipeline_launch_count(None) == 0
    lib.spmm_b200_pipeline_destroy(None)
    # the "_f" pipeline entry points
    af = a.astype(np.float32)
    assert lib.spmm_b200_pipeline_create_f32(None, rp, ci, af.ctypes.data, 10, len(ja), 8, 0) == -1
    assert lib.spmm_b200_pipeline_create_f32(C.byref(h), rp, None, af.ctypes.data, 10, len(ja), 8, 0) == -1
    assert lib.spmm_b200_pipeline_create_f32(C.byref(h), rp, ci, af.ctypes.data, 10, len(ja), 0, 0) == -1
    assert lib.spmm_b200_pipeline_create_f32(C.byref(h), rp, ci, af.ctypes.data, 10, 2 ** 31, 8, 0) == -5
    assert lib.spmm_b200_pipeline_sddmm_f32(None, None, 0) == -1
    assert lib.spmm_b200_pipeline_final_f32(None, None) == -1
    assert lib.spmm_b200_pipeline_softmax_f32(None, None, 1) == -1
    assert lib.spmm_b200_pipeline_launch_count_f32(None) == 0
    lib.spmm_b200_pipeline_destroy_f32(None)


def test_argument_validation_round2_entry_points(pkg):
    """Multi-GPU, option, timeline and fp32-variant entry points reject bad arguments before touching a device."""
    lib = pkg.load_library()
    ia, ja, a = random_csr(10, 10, 2.0, 0)
    h = C.c_void_p()
    rp, ci, va = ia.ctypes.data, ja.ctypes.data, a.ctypes.data
    # single-process multi-GPU
    assert lib.spmm_b200_create_multi(None, rp, ci, va, 10, 10, len(ja), 4, 1, None, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), None, ci, va, 10, 10, len(ja), 4, 1, None, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), rp, ci, va, 10, 10, len(ja), 4, 0, None, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), rp, ci, va, 10, 10, len(ja), 4, 1000, None, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), rp, ci, va, 10, 10, len(ja), 4, 1, None, 17, 0) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), rp, ci, va, 10, 10, len(ja), 4, 1, None, pkg.AUTO, 9) == -1
    assert lib.spmm_b200_create_multi(C.byref(h), rp, ci, va, 10, 10, 2 ** 31, 4, 1, None, pkg.AUTO, 0) == -5
    assert h.value is None
    assert lib.spmm_b200_run_multi(None, None, None, 4, 1) == -1
    assert lib.spmm_b200_set_b_multi(None, None, 4) == -1
    assert lib.spmm_b200_run_resident_multi(None, 4) == -1
    assert lib.spmm_b200_fetch_c_multi(None, None, 4) == -1
    assert lib.spmm_b200_multi_ngpus(None) == -1
    assert lib.spmm_b200_multi_shard(None, 0) is None
    lib.spmm_b200_destroy_multi(None)
    # per-handle options and the slice timeline
    assert lib.spmm_b200_set_option(None, b"TICKET", 1) == -1
    assert lib.spmm_b200_set_option_f32(None, b"TICKET", 1) == -1
    assert lib.spmm_b200_pipeline_set_option(None, b"RESIDENT", 1) == -1
    assert lib.spmm_b200_last_timeline(None, None, 0) == -1
    # fp32: variant range includes the fp32-only variants, nothing beyond
    af = a.astype(np.float32)
    assert lib.spmm_b200_create_f32(None, rp, ci, af.ctypes.data, 10, 10, len(ja), 4, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_create_f32(C.byref(h), rp, ci, af.ctypes.data, 10, 10, len(ja), 4, 0, 99, 0) == -1
    assert lib.spmm_b200_create_f32(C.byref(h), rp, ci, af.ctypes.data, 10, 10, len(ja), 0, 0, pkg.AUTO, 0) == -1
    assert lib.spmm_b200_variant_f32(None) == -1
    # introspection added with the row-band pipeline / window upload / download routing
    assert lib.spmm_b200_last_rowband(None) == -1
    assert lib.spmm_b200_last_transfer_bytes(None, None, None) == -1
    assert lib.spmm_b200_col_window(None, None, None) == -1
    assert lib.spmm_b200_multi_routes(None, None, 0) == -1
    assert lib.spmm_b200_run_routed(None, None, None, 4, 0, 0, None, None, None, 0, C.cast(None, pkg.XFER_FN), None) == -1
    # the allocation-free entry point checks arguments and the scratch bound before it touches a device
    fake = C.c_void_p(4096)
    assert lib.spmm_b200_spmm_merge_noalloc(None, fake, fake, 10, 10, 10, fake, fake, 32, None) == -1
    assert lib.spmm_b200_spmm_merge_noalloc(fake, fake, fake, 10, 10, 10, fake, fake, 3, None) == -6
    assert lib.spmm_b200_spmm_merge_noalloc(fake, fake, fake, 10, 10, 2 ** 31, fake, fake, 32, None) == -5
    assert lib.spmm_b200_spmm_merge_noalloc(fake, fake, fake, 20_000_000, 10, 20_000_000, fake, fake, 128, None) == -10
    assert b"scratch" in lib.spmm_b200_strerror(-10) and b"needs" in lib.spmm_b200_last_cuda_error()
    # the nnz-carry split needs an output array like the other modes
    assert lib.spmm_b200_partition(rp, 10, len(ja), 4, pkg.PART_NNZ_CARRY, None, 0) == -1


def test_no_cpu_fallback(pkg):
    """Without a usable device construction raises; with one this test is vacuous."""
    lib = pkg.load_library()
    if lib.spmm_b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    ia, ja, a = random_csr(10, 10, 2.0, 0)
    with pytest.raises(pkg.SpmmB200Error) as e:
        pkg.csr_to_format(ia, ja, a, 10, 10, len(ja), 4)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(pkg.SpmmB200Error):
        pkg.partition(ia, 10, len(ja), 4, pkg.PART_NNZ)


def test_host_wrapper_type_checks(pkg):
    ia, ja, a = random_csr(10, 10, 2.0, 0)
    with pytest.raises(TypeError):
        pkg.csr_to_format(ia.astype(np.int64), ja, a, 10, 10, len(ja), 4)
    with pytest.raises(TypeError):
        pkg.csr_to_format(ia, ja, a.astype(np.float32), 10, 10, len(ja), 4)


def test_product_does_not_touch_the_oracle():
    """The product tree may not import, link or execute anything under oracle/."""
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "spmm-research_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(d, f), errors="ignore").read()
                for line in txt.splitlines():
                    code = line.split("//")[0].split("#")[0]
                    if re.search(r"oracle[/.]|liboracle|_ref/lib", code):
                        bad.append((f, line.strip()))
    assert not bad, bad


def test_reference_harness_dropin_fails_loudly_without_gpu(pkg):
    """The reference's unchanged harness linked with our plug-in TU: without a GPU it must print
    the error on stdout and exit(1) (reference convention), never compute on the CPU."""
    exe = os.path.join(ROOT, "oracle", "_ref", "spmm_b200_d.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/spmm_b200_d.exe not built")
    if pkg.load_library().spmm_b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    import subprocess
    env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="1", USE_DLCM_MATRICES="0", NUM_COLS="4",
               COOLDOWN="0", RAPL_REGISTERS="0", OMP_NUM_THREADS="2")
    r = subprocess.run([exe] + "2000 2000 5 1 normal random 1 0 0 0 14".split(), env=env, capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stdout
