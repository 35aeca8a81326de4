"""The reference arm of bench.py (baseline/): the copied HeterPS sources are byte-identical to
/root/reference, the arm prints a well-formed line on a box without GPUs, and it never touches
the product package."""
import hashlib
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference/paddle/fluid/framework/fleet/heter_ps")


@pytest.mark.skipif(not REF.exists(), reason="reference tree not present on this box")
def test_reference_sources_are_unmodified_copies():
    sys.path.insert(0, str(ROOT / "baseline"))
    import build_ref
    man = build_ref.copy_sources()
    assert len(man) >= 20
    for f, digest in man.items():
        assert hashlib.sha256((REF / f).read_bytes()).hexdigest() == digest, f


def test_reference_arm_does_not_import_the_product():
    src = (ROOT / "baseline" / "reference_arm.py").read_text()
    assert "import paddlebox_b200" not in src and "from paddlebox_b200" not in src
    harness = (ROOT / "baseline" / "harness" / "ref_harness.cu").read_text()
    assert "pbx_" not in harness


def test_reference_arm_line_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU box: the arm would run the real benchmark")
    if not (ROOT / "baseline" / "_ref" / "libref_heterps.so").exists() and not REF.exists():
        pytest.skip("no library and no reference tree")
    env = dict(os.environ, RANK="0")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line
