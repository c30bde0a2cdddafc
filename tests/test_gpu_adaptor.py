"""-m gpu: the reference-side binding, compiled. oracle/_ref/adaptor_test is include/nvbio_b200_adaptor.h built by nvcc against the
reference's OWN headers (oracle/Makefile target `adaptor`, in the container that has /root/reference): it drives libnvbio_b200.so
through the reference's types and call shapes -- io::FMIndexDataCore, ConcatenatedStringSet, aln::make_*_aligner,
aln::batch_alignment_score(..., DeviceB200Scheduler, ...), batch_banded_alignment_score<15>, the tracebacks, match / locate --
and compares every result with the reference's host code in the same binary."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "adaptor_test")


def test_adaptor_compiled_against_the_reference_headers():
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/adaptor_test not built (needs /root/reference at build time)")
    r = subprocess.run([EXE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "adaptor test ok" in r.stdout
    assert r.stdout.count(" ok ") >= 28 and "FAIL" not in r.stdout
