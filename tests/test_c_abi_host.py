"""the C ABI driven by a compiled host that is not Python — tests/c_abi_host.c walks the call
sequence of the reference's renderer (construct, force-field upload, tick per frame, read back, reconstruct), the same
sequence rust/src/simulation_cuda.rs binds (no Rust toolchain exists on this image, so gcc stands in)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gpu_fluid_simulation_b200")


def _build(tmp_path):
    exe = str(tmp_path / "c_abi_host")
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    subprocess.run([cc, "-O1", "-Wall", "-Wextra", "-Werror", "-std=c99", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "c_abi_host.c"),
                    "-L", PKG, "-lfluid_b200", f"-Wl,-rpath,{PKG}", "-o", exe], check=True, capture_output=True, text=True)
    return exe


def test_header_compiles_as_c99_and_links(fluid, tmp_path):
    """no GPU needed: the header is plain C and the library exports everything the host program references."""
    _build(tmp_path)


@pytest.mark.gpu
def test_c_host_walks_the_renderer_call_sequence(fluid, tmp_path):
    exe = _build(tmp_path)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "c_abi_host ok" in res.stdout
