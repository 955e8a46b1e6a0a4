"""The adapter (integration/dash_adapter.h) and the in-process example
compile against the UNMODIFIED reference headers.  Only where
/root/reference exists (the build container); CPU only, compile without
linking or running."""
import os
import subprocess

import pytest

from helpers import ROOT

REF = "/root/reference/dash/include"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sources not present")
def test_inprocess_example_compiles(tmp_path):
    obj = str(tmp_path / "example.o")
    cmd = ["/usr/local/cuda/bin/nvcc", "-x", "cu", "-std=c++17", "-w", "-c", "-Xcompiler", "-maes,-mrdrnd,-msse4.2,-fopenmp",
           "-I" + REF, "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "integration"),
           os.path.join(ROOT, "integration", "example_inprocess.cpp"), "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr[-3000:]
    assert os.path.getsize(obj) > 0
