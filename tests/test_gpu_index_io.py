"""cm_gpu_write_index / cm_gpu_load_index_file against the reference's own index builder: the file written from the
device-resident index must equal the one `circminer --index` writes (compact `-m`: byte for byte; full: every byte the
reference initialises), and an index read back from the reference's file must hold the same buckets as the device
build from the contig text."""
import filecmp
import os

import numpy as np
import pytest

from conftest import has_ref
from test_rounds import build_ref_index
import refbinary

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("kmer,compact", [(20, False), (22, False), (14, True), (20, True)])
def test_written_index_is_the_reference_file(tmp_path, kmer, compact):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    w = workload.prepare("T0", str(tmp_path / "t0"), contig_size=250_000)
    build_ref_index(w["fa"], 250_000, kmer=kmer, compact=compact)
    ref_file = w["fa"] + ".packed.fa.index"
    m = Mapper(default_params(kmer=kmer))
    workload.load_into(m, w)
    mine = str(tmp_path / "mine.index")
    m.write_index(mine, len(w["contigs"]), compact=compact)
    assert os.path.getsize(mine) == os.path.getsize(ref_file)
    if compact:
        assert filecmp.cmp(mine, ref_file, shallow=False), "index file differs from the reference's"
    else:
        # every byte the reference initialises (its table is malloc'ed: the checksum field of the bucket headers and the slots
        # behind a bucket's last entry are never written and never read, src/mrsfast/HashTable.c:346, :824-839)
        refbinary.assert_same_index(mine, ref_file)
    # and back: the reference's file -> device index == device build from text
    m2 = Mapper(default_params(kmer=kmer))
    assert m2.load_index_file(ref_file) == len(w["contigs"])
    rng = np.random.default_rng(5)
    for c in range(len(w["contigs"])):
        assert m.contig_stats(c) == m2.contig_stats(c)
        code = {65: 0, 67: 1, 71: 2, 84: 3}
        seq = w["contigs"][c]
        for p in rng.integers(0, len(seq) - 14, size=200):
            win = seq[p:p + 14]
            if any(int(b) not in code for b in win):
                continue
            hv = 0
            for b in win:
                hv = hv * 4 + code[int(b)]
            a, b_ = m.read_bucket(c, hv), m2.read_bucket(c, hv)
            assert a[2] == b_[2] and (a[0] == b_[0]).all() and (a[1] == b_[1]).all()
    m.close()
    m2.close()
