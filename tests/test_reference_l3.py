"""the reference's own L3 Python files run unmodified on this repo's
`MinkowskiEngineBackend._C` (drop-in at the pybind boundary, pybind/extern.hpp:517-806).
A subprocess puts the staged reference package first on sys.path (tests/ref_l3_driver.py)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT
from harness import stage

pytestmark = pytest.mark.skipif(not stage.is_staged() and not stage.stage(), reason="reference files not staged")


def _run(dev):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_l3_driver.py"), dev, ROOT],
                       capture_output=True, text=True, timeout=600, cwd="/tmp")
    assert r.returncode == 0 and "REF-L3-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_reference_l3_python_on_oracle_backend_cpu():
    _run("cpu-oracle")


@pytest.mark.gpu
def test_reference_l3_python_on_the_engine():
    _run("cuda")
