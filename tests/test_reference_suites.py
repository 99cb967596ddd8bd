"""The reference's OWN CuTest suites for the hot path, compiled from /root/reference where they
lie (oracle/Makefile target `cutest`) and linked against the binding: every getAlignedPairs* / poa_realign /
computeForwardProbability / poa_polish call of tests/pairwiseAlignerTest.c (23 tests), tests/polisherTest.c (20 tests)
and tests/stitchingTest.c (3 tests) then runs on the GPU through libmargin_b200.so.  The binding's own CuTest suite
(oracle/b200_tests.c: test_poa_augment_example restated for b200_poa_augment, b200_poa_augment vs poa_augment A/B on
aligner output, the stThrowNew exception path, bubble-graph batching A/B) is linked into the same binary."""
import os
import re
import subprocess
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
CWD = os.path.join(REF, "data", "build")


def _run(binary, suites, timeout):
    exe = os.path.join(REF, binary)
    if not (os.path.exists(exe) and os.path.isdir(CWD)):
        pytest.skip("oracle/_ref/%s not built (needs /root/reference at build time)" % binary)
    t0 = time.time()
    r = subprocess.run([exe, "critical"] + suites, cwd=CWD, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=timeout)
    dt = time.time() - t0
    tail = r.stdout[-2500:]
    m = re.search(r"OK \((\d+) tests?\)", r.stdout)
    assert r.returncode == 0 and m, tail
    print("%s %s: OK (%s tests) in %.1f s" % (binary, " ".join(suites), m.group(1), dt))
    return int(m.group(1)), dt


def test_reference_pairwise_suite_passes_unmodified():
    """CPU: pins that the suite binary itself is sound (the unmodified reference passes its own tests here)."""
    n, _ = _run("cutest_ref", ["pairwise"], 600)
    assert n == 23


def test_binding_suite_on_the_cpu_stand_in():
    """CPU: the HOST side of the binding (batch assembly, b200_poa_augment vs the reference's poa_augment, the
    deferred bubble-graph queue, stThrowNew) against oracle/stub_b200.c, a test-only stand-in for the CUDA library
    built on the oracle.  The same suite runs against the real library in test_binding_suite (GPU)."""
    n, _ = _run("cutest_stub", ["b200"], 900)
    assert n == 6


@pytest.mark.gpu
def test_reference_pairwise_suite_on_the_binding():
    n, _ = _run("cutest_b200", ["pairwise"], 900)
    assert n == 23


@pytest.mark.gpu
def test_reference_polisher_and_stitching_suites_on_the_binding():
    n, _ = _run("cutest_b200", ["polisher", "stitching"], 2400)
    assert n == 23


@pytest.mark.gpu
def test_binding_suite():
    n, _ = _run("cutest_b200", ["b200"], 900)
    assert n == 6
