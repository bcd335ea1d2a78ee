import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def has_ref():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libcm_ref.so"))


needs_ref = pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")

# contig size that splits the small synthetic genomes into several packed contigs ("rounds")
T0_CONTIG_SIZE = 250_000


@pytest.fixture(scope="session")
def t0(tmp_path_factory):
    """T0 synthetic dataset (2 chromosomes -> 2 packed contigs)."""
    from circminer_b200 import synth
    d = tmp_path_factory.mktemp("t0")
    return synth.generate(synth.CONFIGS["T0"], str(d))


@pytest.fixture(scope="session")
def t0_packed(t0):
    from circminer_b200 import annotation, synth
    chroms = [(synth.chrom_name(i), synth.codes_to_ascii(c)) for i, c in enumerate(t0["chroms"])]
    contigs, info = annotation.pack_genome(chroms, T0_CONTIG_SIZE)
    flat = annotation.flatten(t0["gtf"], info, [len(c) for c in contigs], 300)
    return dict(contigs=contigs, info=info, flat=flat)


@pytest.fixture(scope="session")
def ref_t0(t0):
    """The unmodified reference (oracle/_ref) opened on T0.  Process-global singletons: one per session."""
    if not has_ref():
        pytest.skip("oracle/_ref not built")
    from oracle.refharness import RefHarness, default_params
    h = RefHarness(default_params())
    h.build_index(t0["fa"], contig_size=T0_CONTIG_SIZE)
    h.open(t0["fa"], t0["gtf"], os.path.join(os.path.dirname(t0["fa"]), "ref"))
    h.rounds_done = 0
    return h


def ascii_reads(d):
    from circminer_b200 import synth
    return synth.codes_to_ascii(d["r1"]), synth.codes_to_ascii(d["r2"])
