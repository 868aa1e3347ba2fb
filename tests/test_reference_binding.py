"""The reference's OWN Python binding and its OWN pytest suite, unmodified, on top of libslimg_b200.

bindings/python/slimg/_types.py imports the UniFFI-generated `slimg._lowlevel`; here that name resolves to
slimg_b200._lowlevel, nothing else changes.  The files are byte-for-byte copies staged by oracle/stage_ref_binding.py
into oracle/_ref/pybinding/ (the GPU box has no /root/reference); MANIFEST.json carries their hashes.

Collected from the reference suite: test_operations.py (crop / extend pixels, resize dimensions, argument
validation), test_convert.py (pipeline + encode), test_format.py, test_decode.py, test_optimize.py.  Deselected, with
the reason: the JXL case (test_to_jxl) -- libjxl's entropy coder is out of scope and has no
stand-in on this box.
"""
import hashlib
import json
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGED = os.path.join(ROOT, "oracle", "_ref", "pybinding")

# a conftest placed NEXT TO the staged suite's own conftest cannot exist (same file name), so the module swap happens
# in a plugin passed with -p
PLUGIN = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, os.environ["SLIMG_B200_ROOT"])
    import slimg_b200
    slimg_b200.install_as_slimg_lowlevel()      # `from slimg import _lowlevel` now finds libslimg_b200's binding
    sys.path.insert(0, os.environ["SLIMG_REF_BINDING"])
    import slimg                                 # the reference's package, unmodified
    assert slimg._types._lowlevel is sys.modules["slimg_b200._lowlevel"]
''')


def _staged() -> bool:
    return os.path.exists(os.path.join(STAGED, "MANIFEST.json"))


def test_staged_copy_is_unmodified():
    if not _staged():
        pytest.skip("oracle/_ref/pybinding not staged (python oracle/stage_ref_binding.py where /root/reference exists)")
    with open(os.path.join(STAGED, "MANIFEST.json")) as fh:
        manifest = json.load(fh)["sha256"]
    assert "slimg/_types.py" in manifest and "tests/test_operations.py" in manifest
    for rel, digest in manifest.items():
        with open(os.path.join(STAGED, rel), "rb") as fh:
            assert hashlib.sha256(fh.read()).hexdigest() == digest, rel
        ref = os.path.join("/root/reference/bindings/python", rel)
        if os.path.exists(ref):          # in the build container: identical to the reference's file
            with open(ref, "rb") as fh:
                assert hashlib.sha256(fh.read()).hexdigest() == digest, rel


def _run_reference_suite(tmp_path, extra):
    plugin_dir = tmp_path / "plugin"
    plugin_dir.mkdir()
    (plugin_dir / "slimg_b200_swap.py").write_text(PLUGIN)
    env = dict(os.environ, SLIMG_B200_ROOT=ROOT, SLIMG_REF_BINDING=STAGED,
               PYTHONPATH=os.pathsep.join([str(plugin_dir), os.environ.get("PYTHONPATH", "")]))
    cmd = [sys.executable, "-m", "pytest", "-q", "-p", "slimg_b200_swap", "-p", "no:cacheprovider",
           "--rootdir", os.path.join(STAGED, "tests"), os.path.join(STAGED, "tests")] + extra
    return subprocess.run(cmd, env=env, capture_output=True, text=True, cwd=str(tmp_path))


def test_reference_wrapper_validation_cases_without_gpu(tmp_path):
    """Argument validation of the unmodified wrapper needs no device: ValueError cases of test_operations.py."""
    if not _staged():
        pytest.skip("oracle/_ref/pybinding not staged")
    res = _run_reference_suite(tmp_path, ["-k", "(both_args or no_args or fill_channel or fill_invalid or fill_wrong or "
                                          "no_mode or multiple_modes or quality_ or unknown_format_string or TestFormat) "
                                          "and not optimize", "--durations=3"])
    assert res.returncode == 0, res.stdout + res.stderr
    assert " passed" in res.stdout and "failed" not in res.stdout


@pytest.mark.gpu
def test_reference_binding_suite_unmodified(tmp_path):
    if not _staged():
        pytest.skip("oracle/_ref/pybinding not staged")
    res = _run_reference_suite(tmp_path, ["--deselect", "test_convert.py::TestConvert::test_to_jxl", "-rA"])
    sys.stdout.write(res.stdout[-6000:])
    assert res.returncode == 0, res.stdout + res.stderr
    import re
    m = re.search(r"(\d+) passed", res.stdout)
    assert m and int(m.group(1)) >= 60, res.stdout[-2000:]
    assert "failed" not in res.stdout.splitlines()[-1]
