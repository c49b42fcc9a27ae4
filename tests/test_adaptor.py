"""The C++ adaptor header (include/sba_b200_adaptor.hpp: ProjectorModule::addProjectionCostFunction / SteppedProcess protocol
over the C ABI) compiled into the reference's simplePnP scenario (tests/test_sba_solver/main.cpp:61-236)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "adaptor", "adaptor_pnp.cpp")
LIBDIR = os.path.join(ROOT, "steviapp_b200")


def build(tmp_path):
    exe = str(tmp_path / "adaptor_pnp")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
                           "-L", LIBDIR, "-lsba_b200", f"-Wl,-rpath,{LIBDIR}"])
    return exe


def test_adaptor_compiles_links_and_fails_loudly_without_a_device(tmp_path):
    import torch
    exe = build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stdout + r.stderr
    else:   # no CPU fallback: init() returns false and the error names the missing device
        assert r.returncode == 3, r.stdout + r.stderr
        assert "init failed" in r.stderr


@pytest.mark.gpu
def test_adaptor_simple_pnp_on_the_gpu(tmp_path):
    r = subprocess.run([build(tmp_path)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "adaptor_pnp: ok" in r.stdout
