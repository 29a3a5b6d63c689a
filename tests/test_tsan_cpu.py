"""ThreadSanitizer over the CPU-side native code (the reference's CI compiles nothing and runs no sanitizer).

`tests/native/tsan_host.cpp` runs the shared-memory transport's ranks as threads of one process over ONE mapping of the
segment — so TSan sees every cross-rank access — through all-reduce (with idle ranks), the fused step, the elastic
average, broadcasts from rotating roots and concurrent AsyncEA syncs (lock-free and queued), then the batch loader's
worker threads.  Zero reports expected.  A second build with the barrier's release store weakened to relaxed must be
caught: that is the check that the harness can see a protocol bug at all."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = [os.path.join(ROOT, "tests", "native", "tsan_host.cpp"), os.path.join(ROOT, "csrc", "host_transport.cpp"),
       os.path.join(ROOT, "csrc", "loader.cpp")]
FLAGS = ["-std=c++17", "-O1", "-g", "-fsanitize=thread", "-DDLH_NO_CLONES", "-ffp-contract=off", "-pthread"]


def _build(tmp_path, sources, name):
    cxx = shutil.which("g++") or shutil.which("clang++")
    if cxx is None:
        pytest.skip("no C++ compiler")
    exe = str(tmp_path / name)
    res = subprocess.run([cxx, *FLAGS, "-o", exe, *sources, "-lrt"], capture_output=True, text=True)
    if res.returncode != 0 and "tsan" in (res.stderr + res.stdout).lower():
        pytest.skip("ThreadSanitizer runtime not available: " + res.stderr[-200:])
    assert res.returncode == 0, res.stderr[-2000:]
    return exe


def _run(exe, world, rounds):
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66")
    return subprocess.run([exe, str(world), str(rounds)], capture_output=True, text=True, env=env, timeout=600)


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world", [2, 4])
def test_host_transport_and_loader_are_race_free_under_tsan(tmp_path, world):
    exe = _build(tmp_path, SRC, "tsan_host")
    res = _run(exe, world, 12)
    assert "WARNING: ThreadSanitizer" not in res.stderr, res.stderr[-4000:]
    assert res.returncode == 0, (res.stdout + res.stderr)[-2000:]
    assert "0 failures" in res.stdout


@pytest.mark.timeout(900)
def test_the_harness_catches_a_weakened_release(tmp_path):
    broken = tmp_path / "host_transport_broken.cpp"
    text = open(SRC[1]).read()
    needle = ".store(word, std::memory_order_release)"
    assert text.count(needle) == 1
    text = text.replace(needle, ".store(word, std::memory_order_relaxed)").replace('"../include/', f'"{ROOT}/include/')
    broken.write_text(text)
    exe = _build(tmp_path, [SRC[0], str(broken), SRC[2]], "tsan_broken")
    res = _run(exe, 4, 6)
    assert "WARNING: ThreadSanitizer: data race" in res.stderr, "a relaxed flag store must show up as a data race"
