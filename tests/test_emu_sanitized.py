"""compute-sanitizer is closed on this GPU pool, so the memory-safety pass SURVEY.md section 5 asks for runs on the host:
the per-thread bodies of the search / build / stats kernels (the same source the device compiles, tests/emu) under
AddressSanitizer + UndefinedBehaviorSanitizer over the golden fixtures.  One subprocess (ASan must be loaded first)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_kernel_bodies_are_clean_under_asan_and_ubsan():
    libasan = subprocess.run(["/usr/bin/g++", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(libasan) or not os.path.exists(libasan):
        import pytest

        pytest.skip("libasan.so not found")
    env = dict(os.environ, FIN_EMU_ASAN="1", LD_PRELOAD=libasan, ASAN_OPTIONS="detect_leaks=0:halt_on_error=1:abort_on_error=0",
               UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_kernel_logic_cpu.py"), "-x", "-q", "-p", "no:cacheprovider",
                        "-k", "reference_output or reference_files or long_reads or stats_mode_bodies_match_reference"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "AddressSanitizer" not in tail and "runtime error" not in tail, tail
