"""Import the UNMODIFIED reference package (vllm_omni, installed by scripts/install_reference.sh into the
git-ignored baseline/_ref, which travels to the GPU box) inside a test process and its subprocesses. Third-party
dependencies of its import chain that this image does not have are replaced by the small stand-ins under
tests/stubs/ (STUBBED lists them; a stand-in is only used when the real package is missing). Nothing under
vllm_omni_b200/ imports this."""
from __future__ import annotations

import importlib.util
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
REF = ROOT / "baseline" / "_ref"
STUBS = ROOT / "tests" / "stubs"
STUBBED = ("omegaconf", "aenum")


def available() -> bool:
    return (REF / "vllm_omni" / "__init__.py").exists()


def setup() -> list[str]:
    """Make `import vllm_omni` resolve to baseline/_ref here and in child processes (vLLM inspects model classes
    in a subprocess); returns the stand-ins in use."""
    if not available():
        raise RuntimeError("baseline/_ref is missing: run scripts/install_reference.sh where /root/reference exists")
    missing = [n for n in STUBBED if importlib.util.find_spec(n) is None or str(STUBS) in str(importlib.util.find_spec(n).origin)]
    paths = [str(REF), str(ROOT)] + ([str(STUBS)] if missing else [])
    for p in reversed(paths):
        if p not in sys.path:
            sys.path.insert(0, p)
    old = os.environ.get("PYTHONPATH", "")
    os.environ["PYTHONPATH"] = os.pathsep.join([p for p in paths if p not in old.split(os.pathsep)] + ([old] if old else []))
    return missing
