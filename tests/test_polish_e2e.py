"""End-to-end drop-in check (SURVEY.md §8 C-4 iv): the reference's own `margin polish` CLI, with the seven functions of
compat/pairwiseAligner_b200.c linked in place of the reference's (oracle/Makefile target `polish`), must write a
polished FASTA that is byte-identical to the unmodified reference's on the bundled ONT region
(tests/data/realData/HG002.r94g360.chr20_59M_100k.bam, r94-g360 params).  The expected digest was produced here by
oracle/_ref/margin_ref (the unmodified reference, 44 s on 8 cores) and equals the one recorded in SURVEY.md §6."""
import hashlib
import os
import subprocess
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
EXPECTED_FASTA_MD5 = "7feefc269d7d43b18c387a21b23e1ef5"


def _polish(binary, out_prefix, threads, env=None):
    data = os.path.join(REF, "data")
    cmd = [os.path.join(REF, binary), "polish", os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam"),
           os.path.join(data, "hg38.chr20_59M_100k.fa"),
           os.path.join(data, "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
           "-t", str(threads), "-o", out_prefix, "-a", "info"]
    t0 = time.time()
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=1500)
    dt = time.time() - t0
    assert r.returncode == 0, r.stdout[-3000:]
    with open(out_prefix + ".fa", "rb") as fh:
        return hashlib.md5(fh.read()).hexdigest(), dt, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["tiles", "sequential", "tiles_f32"])
def test_margin_polish_with_b200_aligner_is_byte_identical(tmp_path, mode):
    if not os.path.exists(os.path.join(REF, "margin_b200cli")):
        pytest.skip("oracle/_ref/margin_b200cli not built (needs /root/reference at build time)")
    env = dict(os.environ, MARGIN_B200_MODE=mode.split("_")[0])
    if mode.endswith("f32"):  # the FP32 fast mode: posteriors differ by ~1e-5, the consensus must not
        env["MARGIN_B200_PRECISION"] = "f32"
    md5, dt, log = _polish("margin_b200cli", str(tmp_path / "b200"), 2, env)
    stages = [l.strip() for l in log.splitlines() if "Took" in l]
    print("margin_b200cli (%s): %.1f s wall\n%s" % (mode, dt, "\n".join(stages[:12])))
    assert md5 == EXPECTED_FASTA_MD5


# `margin polish --diploid` (polish.c:676: phasing bubble graphs, then one polish per haplotype) on the same region; the
# digests are the unmodified reference's (oracle/_ref/margin_ref -2, 43 s on 8 cores of the GPU box,
# profiles/r01_v13_diploid.log).
EXPECTED_DIPLOID_MD5 = {"hap1": "5d467b84be719d49a0bf18a8caaca4ae", "hap2": "8b8440e24cf2ad7aa411afe6e4fc86bb"}


@pytest.mark.gpu
def test_diploid_polish_with_b200_aligner_is_byte_identical(tmp_path):
    """The diploid path builds its bubble graphs from polish.c, not from poa_polish: the binding's wrapper of
    bubbleGraph_constructFromPoaAndVCF must batch those forward matrices too and leave both haplotype FASTAs as the
    reference writes them."""
    if not os.path.exists(os.path.join(REF, "margin_b200cli")):
        pytest.skip("oracle/_ref/margin_b200cli not built (needs /root/reference at build time)")
    data = os.path.join(REF, "data")
    prefix = str(tmp_path / "dip")
    cmd = [os.path.join(REF, "margin_b200cli"), "polish", os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam"),
           os.path.join(data, "hg38.chr20_59M_100k.fa"),
           os.path.join(data, "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
           "-t", "8", "-o", prefix, "-2", "-a", "info"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    batched = [l for l in r.stdout.splitlines() if "margin_b200 bubble graph (phasing)" in l]
    assert batched, "the phasing bubble graphs did not go through the batching wrapper"
    for hap, want in EXPECTED_DIPLOID_MD5.items():
        with open("%s.fa.%s" % (prefix, hap), "rb") as fh:
            assert hashlib.md5(fh.read()).hexdigest() == want, hap


def test_reference_cli_reproduces_expected_digest(tmp_path):
    """CPU: the unmodified reference built by oracle/Makefile writes the expected FASTA (pins the digest itself)."""
    if not os.path.exists(os.path.join(REF, "margin_ref")):
        pytest.skip("oracle/_ref/margin_ref not built")
    if os.environ.get("MARGIN_B200_SLOW_TESTS") != "1":
        pytest.skip("44 s of CPU; set MARGIN_B200_SLOW_TESTS=1")
    md5, dt, _ = _polish("margin_ref", str(tmp_path / "ref"), 8)
    assert md5 == EXPECTED_FASTA_MD5


@pytest.mark.gpu
def test_every_batch_of_the_polish_run_matches_the_oracle(tmp_path, monkeypatch):
    """Real-data parity below the FASTA (SURVEY.md §8 C-4 iii on config C1): every posterior batch `margin polish`
    sends to the library on the bundled region (2 x 314 alignments, reads up to 69 k symbols, long indels, plus the
    stitcher's overlap alignment) is dumped by the binding and replayed here; all tuple lists must have the oracle's
    coordinates and weights within one unit of 1e-7 (every tile seam is guarded), on the plain and on the pipelined path."""
    import glob
    import sys
    import numpy as np
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import compare_lists, load_dump
    from margin_b200 import abi
    from margin_b200.aligner import PairHmmAligner
    from oracle.oracle_py import Oracle
    if not os.path.exists(os.path.join(REF, "margin_b200cli")):
        pytest.skip("oracle/_ref/margin_b200cli not built (needs /root/reference at build time)")
    env = dict(os.environ, MARGIN_B200_DUMP=str(tmp_path / "d"))
    md5, _, _ = _polish("margin_b200cli", str(tmp_path / "b200"), 2, env)
    assert md5 == EXPECTED_FASTA_MD5
    dumps = sorted(glob.glob(str(tmp_path / "d.*.bin")))
    assert len(dumps) >= 2
    orc = Oracle()
    n_aln = 0
    for path in dumps:
        items, models, params, want = load_dump(path)
        ref, _ = orc.posteriors_batch(models, params, abi.Batch(items), want, threads=os.cpu_count() or 1)
        for env_set in (dict(MARGIN_B200_PIPELINE="0"), dict(MARGIN_B200_PIPELINE_FORCE="1", MARGIN_B200_PIPELINE_MIN_BYTES="0")):
            for k in ("MARGIN_B200_PIPELINE", "MARGIN_B200_PIPELINE_FORCE", "MARGIN_B200_PIPELINE_MIN_BYTES"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env_set.items():
                monkeypatch.setenv(k, v)
            al = PairHmmAligner(models, params, device=0)
            al.set_pair_format(abi.PAIRS_PACKED64)
            got = al.posteriors_batch(abi.Batch(items, compact=True), bool(want))
            al.close()
            for i in range(len(items)):
                try:
                    compare_lists(got[i], ref[i], tol_w=1)
                except AssertionError as e:
                    raise AssertionError("%s alignment %d (%s): %s" % (os.path.basename(path), i, env_set, e))
        n_aln += len(items)
    print("%d batches, %d alignments replayed against the oracle" % (len(dumps), n_aln))


# Synthetic diploid region with reads that margin sets aside and assigns to haplotypes afterwards (MAPQ below the filter
# threshold): exercises bubbleGraph_partitionFilteredReads (impl/bubbleGraph.c:1500, polish.c:821), which the binding
# wraps with its record / replay batching.  Digests of both haplotype FASTAs from the unmodified reference
# (oracle/_ref/margin_ref -2 -t 1 in this container; 41 filtered reads: 18 to hap1, 23 to hap2).
EXPECTED_FILTERED_DIPLOID_MD5 = ("21485d0dde6ca98ff9307926e0729b81", "840c07e181bc34c1dbe5b44835ea26a9")


def _filtered_diploid_run(tmp_path, binary):
    import sys
    sys.path.insert(0, ROOT)
    from margin_b200 import synth
    if not (os.path.exists(os.path.join(REF, binary)) and os.path.exists(os.path.join(REF, "sam2bam"))):
        pytest.skip("oracle/_ref/%s not built (needs /root/reference at build time)" % binary)
    fa, sam, _, _ = synth.write_polish_dataset(str(tmp_path), 7, 30000, 40, 8000, synth.ONT, 0.01, het=synth.HET, low_mapq_frac=0.25)
    bam = str(tmp_path / "reads.bam")
    subprocess.run([os.path.join(REF, "sam2bam"), sam, bam], check=True, stderr=subprocess.DEVNULL)
    prefix = str(tmp_path / "out")
    cmd = [os.path.join(REF, binary), "polish", bam, fa,
           os.path.join(REF, "data", "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
           "-t", "1", "-2", "-o", prefix, "-a", "info"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:]
    got = tuple(hashlib.md5(open("%s.fa.%s" % (prefix, h), "rb").read()).hexdigest() for h in ("hap1", "hap2"))
    return got, r.stdout


def test_filtered_reads_partition_on_the_cpu_stand_in(tmp_path):
    """CPU: host side of the record / replay wrapper against oracle/stub_b200.c (test-only stand-in for the CUDA library)."""
    got, log = _filtered_diploid_run(tmp_path, "margin_stubcli")
    assert "margin_b200 partition of filtered reads: 465 forward matrices in one batch" in log
    assert "Of 41 filtered reads: 18 (0.44) were hap1, 23 (0.56) were hap2" in log
    assert got == EXPECTED_FILTERED_DIPLOID_MD5


@pytest.mark.gpu
def test_filtered_reads_partition_is_batched_and_identical(tmp_path):
    got, log = _filtered_diploid_run(tmp_path, "margin_b200cli")
    assert "margin_b200 partition of filtered reads: 465 forward matrices in one batch" in log
    assert got == EXPECTED_FILTERED_DIPLOID_MD5


# Two overlapping chunks polished by two threads at once: batch assembly, the graph built from slabs (observations, node-ordered
# inserts / deletes), deferred bubble-graph likelihoods, repeat counts and the stitcher, host side only.  Digest of the polished
# FASTA from the unmodified reference (oracle/_ref/margin_ref -t 2 in this container).
EXPECTED_TWO_CHUNK_MD5 = "abb95603c7cfc7b48999defb7b3f1464"


def _two_chunk_run(tmp_path, binary):
    import sys
    sys.path.insert(0, ROOT)
    from margin_b200 import synth
    if not (os.path.exists(os.path.join(REF, binary)) and os.path.exists(os.path.join(REF, "sam2bam"))):
        pytest.skip("oracle/_ref/%s not built (needs /root/reference at build time)" % binary)
    fa, sam, _, _ = synth.write_polish_dataset(str(tmp_path), 11, 130000, 12, 12000, synth.ONT, 0.01)
    bam = str(tmp_path / "reads.bam")
    subprocess.run([os.path.join(REF, "sam2bam"), sam, bam], check=True, stderr=subprocess.DEVNULL)
    prefix = str(tmp_path / "out")
    cmd = [os.path.join(REF, binary), "polish", bam, fa,
           os.path.join(REF, "data", "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
           "-t", "2", "-o", prefix, "-a", "info"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:]
    return hashlib.md5(open(prefix + ".fa", "rb").read()).hexdigest(), r.stdout


def test_two_chunks_two_threads_on_the_cpu_stand_in(tmp_path):
    """CPU: the binding's host side on a multi-chunk, multi-thread run against oracle/stub_b200.c, FASTA byte-identical to
    the unmodified reference's."""
    got, log = _two_chunk_run(tmp_path, "margin_stubcli")
    assert "Merging results from 2 chunks" in log and "node-ordered commit" in log
    assert got == EXPECTED_TWO_CHUNK_MD5


@pytest.mark.gpu
def test_two_chunks_two_threads_are_byte_identical(tmp_path):
    got, log = _two_chunk_run(tmp_path, "margin_b200cli")
    assert "Merging results from 2 chunks" in log
    assert got == EXPECTED_TWO_CHUNK_MD5


# `margin phase` on the bundled region and its VCF (tests/data/realData in the reference, copied by oracle/Makefile): the
# bubble graph comes from bubbleGraph_constructFromVCFAndBamChunkReadVcfEntrySubstrings (impl/bubbleGraph.c:1338), which the
# binding wraps so that its 6 334 forward likelihoods run as one batch.  -p 10000: no down-sampling (lp_solve is stubbed out
# of the oracle build).  Digests of the phased VCF and the haplotagged BAM from the unmodified reference in this container.
EXPECTED_PHASE_MD5 = ("4be6af00df7d7c95408bf98f1762c29d", "7c843561f21834b233db361e59e9e1eb")


def _phase_run(tmp_path, binary):
    data = os.path.join(REF, "data")
    vcf = os.path.join(data, "HG002.r94g360.chr20_59M_100k.vcf")
    if not (os.path.exists(os.path.join(REF, binary)) and os.path.exists(vcf)):
        pytest.skip("oracle/_ref/%s or the bundled VCF not there (needs /root/reference at build time)" % binary)
    prefix = str(tmp_path / "out")
    cmd = [os.path.join(REF, binary), "phase", os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam"),
           os.path.join(data, "hg38.chr20_59M_100k.fa"), vcf,
           os.path.join(data, "params", "phase", "allParams.phase_vcf.ont.json"), "-t", "2", "-p", "10000", "-a", "info", "-o", prefix]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:]
    return tuple(hashlib.md5(open(prefix + ext, "rb").read()).hexdigest() for ext in (".phased.vcf", ".haplotagged.bam")), r.stdout


def test_margin_phase_on_the_cpu_stand_in(tmp_path):
    got, log = _phase_run(tmp_path, "margin_stubcli")
    assert "margin_b200 bubble graph (VCF read substrings): 6334 forward matrices in one batch" in log
    assert got == EXPECTED_PHASE_MD5


@pytest.mark.gpu
def test_margin_phase_is_batched_and_identical(tmp_path):
    got, log = _phase_run(tmp_path, "margin_b200cli")
    assert "margin_b200 bubble graph (VCF read substrings): 6334 forward matrices in one batch" in log
    assert got == EXPECTED_PHASE_MD5


# `margin polish --diploid --vcf VCF --onlyVcfAlleles --skipOutputFasta` (no-RLE params): the phasing bubble graphs come from
# bubbleGraph_constructFromPoaAndVCFOnlyVCFAllele (impl/bubbleGraph.c:1126), the third constructor the binding wraps.
# Digest of the haplotagged BAM from the unmodified reference in this container.
EXPECTED_VCF_ALLELES_BAM_MD5 = "7c843561f21834b233db361e59e9e1eb"


def _vcf_alleles_run(tmp_path, binary):
    data = os.path.join(REF, "data")
    vcf = os.path.join(data, "HG002.r94g360.chr20_59M_100k.vcf")
    if not (os.path.exists(os.path.join(REF, binary)) and os.path.exists(vcf)):
        pytest.skip("oracle/_ref/%s or the bundled VCF not there (needs /root/reference at build time)" % binary)
    prefix = str(tmp_path / "out")
    cmd = [os.path.join(REF, binary), "polish", os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam"),
           os.path.join(data, "hg38.chr20_59M_100k.fa"), os.path.join(data, "params", "polish", "misc", "allParams.no_rle.json"),
           "-t", "2", "-p", "10000", "-2", "-v", vcf, "-A", "-T", "-a", "info", "-o", prefix]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:]
    return hashlib.md5(open(prefix + ".haplotagged.bam", "rb").read()).hexdigest(), r.stdout


def test_vcf_allele_graphs_on_the_cpu_stand_in(tmp_path):
    got, log = _vcf_alleles_run(tmp_path, "margin_stubcli")
    assert "margin_b200 bubble graph (VCF alleles): 4190 forward matrices in one batch" in log
    assert got == EXPECTED_VCF_ALLELES_BAM_MD5


@pytest.mark.gpu
def test_vcf_allele_graphs_are_batched_and_identical(tmp_path):
    got, log = _vcf_alleles_run(tmp_path, "margin_b200cli")
    assert "margin_b200 bubble graph (VCF alleles): 4190 forward matrices in one batch" in log
    assert got == EXPECTED_VCF_ALLELES_BAM_MD5
