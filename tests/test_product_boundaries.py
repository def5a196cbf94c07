"""The product never routes through the oracle or any CPU fallback: nothing under ptype_b200/ or
include/ mentions oracle/, and the package refuses to load without its CUDA library."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_product_does_not_touch_the_oracle():
    offenders = []
    for path in list((ROOT / "ptype_b200").rglob("*")) + list((ROOT / "include").rglob("*")) + \
            list((ROOT / "go").rglob("*")):
        if path.is_file() and path.suffix in {".py", ".cu", ".cpp", ".h", ".hpp", ".go", ""} and path.name != "Makefile":
            text = path.read_text(errors="ignore")
            if "oracle" in text.lower():
                offenders.append(str(path.relative_to(ROOT)))
    assert not offenders, offenders


def test_package_fails_loudly_without_the_cuda_library(tmp_path):
    env = dict(os.environ, PTYPE_B200_LIB=str(tmp_path / "missing.so"))
    out = subprocess.run([sys.executable, "-c", "import ptype_b200"], cwd=ROOT, env=env,
                         capture_output=True, text=True)
    assert out.returncode != 0 and "no CPU fallback" in out.stderr


def test_abi_rejects_bad_arguments():
    import ctypes as C
    import ptype_b200 as p
    from ptype_b200 import _abi
    lib = _abi.lib
    assert lib.ptype_b200_strerror(_abi.E_NO_IMAGE).decode() == "no registry image installed"
    assert lib.ptype_b200_open(None, 0, None) == _abi.E_INVAL
    assert lib.ptype_b200_dispatch(None, 0, None, 10, None, 0) == _abi.E_INVAL
    assert lib.ptype_b200_publish_image(None, None, 0) == _abi.E_INVAL
    assert lib.ptype_b200_image_validate(b"x" * 100, 100) == _abi.E_BAD_IMAGE
    assert lib.ptype_b200_builder_new(b"a" * 200, 3, C.byref(C.c_void_p())) == _abi.E_LIMIT
    assert lib.ptype_b200_reply_checksum_noalloc(0, None, 1, 0, None, None) == _abi.E_INVAL
    assert lib.ptype_b200_call_batch(None, 0, 0, 0, None, 0, None, None, 2, 0, None, 0) == _abi.E_INVAL
