"""The reference's OWN caller stack, unmodified, driving the CUDA store on the B200:
``RetrievalExecutor`` + ``BatchScheduler`` + ``ComponentRegistry`` + ``create_component``
(/root/reference/src/pipeline/services/retrieval/api.py:337-349,376-425,
services/gateway/batch_scheduler.py:237-306, component_registry.py:31-83,
component_factory.py:63-104) imported from the copy ``__graft_entry__.build()`` stages under
``baseline/_ref/`` (untracked; it travels to the GPU box with the snapshot).

70 requests go through the real scheduler (two "full" flushes of 32 and one "timeout" flush), the
executor calls ``registry.get("faiss_store").search(...)`` from its worker thread, and every
response's doc ids and scores are compared with the CPU oracle.  Single-device store and two-shard
group (``devices=[0, 0]`` on the one-GPU box), the reference's result cache off and on.
``tests/test_gpu_caller_contract.py`` restates the same caller by hand; it stays as the fallback for
a box without the staged copy."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

STAGED = os.path.join(ROOT, "baseline", "_ref", "src", "pipeline")
HAVE_REF = os.path.isdir("/root/reference/src/pipeline") or os.path.isdir(STAGED)


def _run(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_ref_integration_driver.py"),
                          "--backend", "native", *extra], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-4000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.skipif(not HAVE_REF, reason="reference not staged: run __graft_entry__.build() where /root/reference is mounted")
@pytest.mark.parametrize("devices", ["none", "0,0"])
@pytest.mark.parametrize("cache", ["off", "on"])
def test_reference_executor_drives_the_cuda_store(devices, cache):
    res = _run("--devices", devices, "--cache", cache, "--rows", "50000", "--dim", "128")
    assert res["ok"], res
    assert res["backend"] == "native" and res["store"] == "B200FlatIPStore"
    assert res["group"] == (devices != "none")
    assert res["responses"] == (140 if cache == "on" else 70) and res["docs_per_response"] == 10
    assert res["largest_batch"] == 32 and res["search_calls"] >= 3      # full + timeout flushes
    assert res["ids_ok"] and res["max_score_err"] < 1e-4 and res["unloaded"]
