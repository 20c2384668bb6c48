"""The C ABI from plain C: tests/c_abi/consumer.c is compiled as strict C99 against include/pe_rotation.h, linked with
libpe_rotation.so and the CUDA runtime only, and run - without a GPU for the header / linkage / error-convention
checks, with one for the known-answer tests of the orthogonalisation (no Python or torch in that process)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "poseestimation_b200")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


@pytest.fixture(scope="module")
def consumer(tmp_path_factory):
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    if not os.path.exists(os.path.join(LIBDIR, "libpe_rotation.so")):
        import __graft_entry__
        __graft_entry__.build()
    exe = str(tmp_path_factory.mktemp("c_abi") / "consumer")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c_abi", "consumer.c"), "-L", LIBDIR, "-lpe_rotation",
           "-isystem", os.path.join(CUDA, "include"), "-L", os.path.join(CUDA, "lib64"), "-lcudart", "-lm",
           f"-Wl,-rpath,{LIBDIR}", f"-Wl,-rpath,{os.path.join(CUDA, 'lib64')}", "-o", exe]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return exe


def test_header_is_c99_and_library_links_from_c(consumer):
    res = subprocess.run([consumer], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "cpu checks ok" in res.stdout


@pytest.mark.gpu
def test_known_answers_from_a_c_host_program(consumer):
    res = subprocess.run([consumer, "gpu"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "gpu checks ok" in res.stdout
