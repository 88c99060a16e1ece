"""Host-side code of the B200 path under AddressSanitizer + UndefinedBehaviorSanitizer (SURVEY.md section 5;
compute-sanitizer is closed on the GPU pool, so the host side gets the coverage a sanitizer can give here):
csrc/plan.hpp, csrc/fileio.hpp, csrc/blockmath.hpp (g++) and the control state machine of csrc/kernels.cuh compiled for
the host (nvcc host pass).  A sanitizer report aborts the binary; the tests need no GPU."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SAN = ["-g", "-O1", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-fno-omit-frame-pointer"]
ENV = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")


def test_plan_fileio_blockmath_under_asan_ubsan(tmp_path):
    exe = str(tmp_path / "host_sanitize")
    r = subprocess.run(["g++", "-std=c++17", *SAN, "-pthread", os.path.join(ROOT, "tests", "sanitize", "host_sanitize.cpp"),
                        os.path.join(ROOT, "tests", "blockmath", "sim.cpp"), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=dict(ENV, SANITIZE_TMP=str(tmp_path)))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "no sanitizer report" in r.stdout


def test_control_state_machine_under_asan_ubsan(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "statemachine_sanitize")
    r = subprocess.run([nvcc, "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared",
                        "-Xcompiler", "-g,-O1,-fsanitize=address,-fsanitize=undefined,-fno-sanitize-recover=all,-fno-omit-frame-pointer",
                        os.path.join(ROOT, "tests", "statemachine", "sim.cu"),
                        os.path.join(ROOT, "tests", "sanitize", "statemachine_main.cpp"), "-o", exe, "-lasan", "-lubsan"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    # ASan must not intercept the CUDA runtime's start-up probing on a box without a driver: no CUDA call is made
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600,
                       env=dict(ENV, ASAN_OPTIONS="detect_leaks=0:protect_shadow_gap=0"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "no sanitizer report" in r.stdout


def test_group_meeting_point_under_thread_sanitizer(tmp_path):
    """csrc/group.hpp (barrier / exchange / abort / reset of the host threads of a single-process GPU group) under
    ThreadSanitizer: 2, 3 and 8 threads, 2000 exchange rounds each, a failing rank, a reset."""
    cuda_inc = "/usr/local/cuda/include"
    if not os.path.isdir(cuda_inc):
        pytest.skip("CUDA headers not available")
    exe = str(tmp_path / "group_tsan")
    r = subprocess.run(["g++", "-std=c++17", "-g", "-O1", "-fsanitize=thread", "-pthread", f"-I{cuda_inc}",
                        os.path.join(ROOT, "tests", "sanitize", "group_tsan.cpp"), "-o", exe, "-ldl"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=dict(os.environ, TSAN_OPTIONS="halt_on_error=1"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "no sanitizer report" in r.stdout
