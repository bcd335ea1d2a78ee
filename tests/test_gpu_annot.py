"""Device annotation look-ups against the reference's GTFParser: get_upper_bound, get_location_overlap_ind,
get_shift and the intronic mask on 10^6 random positions per contig plus every position within 3 bases of an exon-interval
border, two (mlen, rlen) settings -- the calls chaining (src/chain.cpp:73-301) and extension (src/extend.cpp) make."""
import os

import numpy as np
import pytest

from conftest import has_ref
from parity import live_dump

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_device_lookups_equal_gtfparser(tmp_path):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000, n_pairs=200)
    dump = live_dump(str(tmp_path / "lk.npz"), workload="T1", contig_size=3_500_000, pairs=200,
                     lookups=1_000_000, no_stage_dumps=True)
    m = Mapper(default_params())
    workload.load_into(m, w)
    total = 0
    for c in range(len(w["contigs"])):
        pos = dump[f"c{c}_lookup_pos"]
        total += len(pos)
        for tag, (ml, rl) in (("a", (20, 130)), ("b", (20, 30))):
            got = m.annot_lookup(c, pos, ml, rl)
            for k, v in got.items():
                want = dump[f"c{c}_lookup_{tag}_{k}"]
                bad = np.nonzero(v != want)[0]
                assert len(bad) == 0, f"contig {c} ({ml},{rl}) {k}: {len(bad)} differ, first pos {pos[bad[:5]]}: {v[bad[:5]]} vs {want[bad[:5]]}"
        assert (got["ol_iv"] >= 0).sum() > 1000 and (got["loc_iv"] >= 0).sum() > 1000 and got["intronic"].sum() > 1000
    assert total > 1_500_000
    m.close()
