"""Drop-in boundary test against the reference's OWN caller stack (CPU; skipped where
/root/reference is not mounted, e.g. on the GPU box)."""
import json
import os
import subprocess
import sys

import pytest

from conftest import REFERENCE_SRC, ROOT, gpu_available

pytestmark = pytest.mark.skipif(not os.path.isdir(REFERENCE_SRC),
                                reason="/root/reference not available on this box")


def _run(backend):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_ref_integration_driver.py"),
                          "--backend", backend], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-3000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


def test_store_drops_into_reference_executor():
    res = _run("native" if gpu_available() else "oracle")
    assert res["ok"], res
    assert res["responses"] == 70 and res["docs_per_response"] == 10 and res["unloaded"]
