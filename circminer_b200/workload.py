"""Synthetic workloads of BASELINE.json's configs, prepared for the C-ABI: packed contigs
(GenomePacker::pack_genome, src/genome.cpp:96-145), the flat annotation of every contig and the
read batch as fixed-stride ASCII rows.  Used by bench.py, smoke() and the parity tests; nothing
here touches oracle/."""
from __future__ import annotations

import os

import numpy as np

from . import annotation, synth


def prepare(name: str, outdir: str, contig_size: int = annotation.DEF_CONTIG_SIZE, n_pairs: int | None = None,
            max_read_len: int = 300, write_reads: bool = False, seed_shift: int = 0) -> dict:
    """Generate config `name` (synth.CONFIGS) under outdir.  seed_shift changes only the read sample
    (rank-specific shards of the same reference)."""
    cfg = synth.CONFIGS[name]
    if n_pairs is not None or seed_shift:
        from dataclasses import replace
        cfg = replace(cfg, n_pairs=n_pairs or cfg.n_pairs, seed_reads=cfg.seed_reads + seed_shift)
    d = synth.generate(cfg, outdir, write_reads=write_reads)
    chroms = [(synth.chrom_name(i), synth.codes_to_ascii(c)) for i, c in enumerate(d["chroms"])]
    contigs, info = annotation.pack_genome(chroms, contig_size)
    flat = annotation.flatten(d["gtf"], info, [len(c) for c in contigs], max_read_len)
    d.update(contigs=contigs, info=info, flat=flat,
             s1=synth.codes_to_ascii(d["r1"]), s2=synth.codes_to_ascii(d["r2"]))
    return d


def load_into(mapper, w: dict) -> None:
    """Build every contig's index on the device and upload its annotation."""
    for c, seq in enumerate(w["contigs"]):
        mapper.build_contig(c, seq)
        mapper.load_annotation(c, w["flat"][c])
