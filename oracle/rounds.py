"""TEST / BENCH INFRASTRUCTURE ONLY.  The reference's round bookkeeping around FilterRead::process_read when it is driven
through oracle/_ref/libcm_ref.so instead of the binary: the skip rule of map_reads (src/circminer.cpp:386-390) and the
MatchedRead a pair carries into the next round after the remain-FASTQ header round trip (src/filter.cpp:413-455 ->
src/fastq_parser.cpp:203-269), restated on the numpy cm_mread rows."""
from __future__ import annotations

import numpy as np

MAPPED = (0, 1, 2, 3, 4, 5, 7)


def skip_rule(scan_level: int, state, mr, len_sum):
    if scan_level == 0:
        return state == 0
    if scan_level == 1:
        return (state == 0) & (mr["gm_compatible"] != 0) & (mr["ed_r1"] + mr["ed_r2"] == 0) & \
               ((mr["mlen_r1"] + mr["mlen_r2"]).astype(np.int64) == len_sum)
    return np.zeros(len(state), bool)


def header_round_trip(mr, max_ed: int = 4):
    from .refharness import fresh_mreads
    out = mr.copy()
    lost = ~np.isin(mr["type"], MAPPED)
    f = fresh_mreads(int(lost.sum()), max_ed)
    f["type"] = mr["type"][lost]
    out[lost] = f
    return out
