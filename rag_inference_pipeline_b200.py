"""Import shim: the package directory is named ``rag-inference-pipeline_b200`` (with hyphens, as
the task layout prescribes), which Python cannot import by name.  ``import
rag_inference_pipeline_b200`` loads that directory as a regular package under this name."""

import importlib.util
import pathlib
import sys

_dir = pathlib.Path(__file__).resolve().parent / "rag-inference-pipeline_b200"
_spec = importlib.util.spec_from_file_location(
    __name__, _dir / "__init__.py", submodule_search_locations=[str(_dir)])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
