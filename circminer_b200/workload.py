"""Synthetic workloads of BASELINE.json's configs, prepared for the C-ABI: packed contigs
(GenomePacker::pack_genome, src/genome.cpp:96-145), the flat annotation of every contig and the
read batch as fixed-stride ASCII rows.  Used by bench.py, smoke() and the parity tests; nothing
here touches oracle/.

C1 / C2 / T* come from synth.py; the GRCh38-scale family C3 / C4 / C5 (and their 1/50 versions C3s ...)
from synth_scale.py.  The scale family caches its reference (code array, packed contigs, flat annotation)
under `<outdir>/../<family>_ref`, guarded by a file lock, so that the ranks of one box and the back-to-back
invocations of a scaling run build it once."""
from __future__ import annotations

import fcntl
import json
import os

import numpy as np

from . import annotation, synth, synth_scale


def is_scale(name: str) -> bool:
    return name in synth_scale.CONFIGS


def config(name: str):
    return synth_scale.CONFIGS[name] if is_scale(name) else synth.CONFIGS[name]


def chrom_names(name: str, n: int):
    return list(config(name).chrom_names) if is_scale(name) else [synth.chrom_name(i) for i in range(n)]


def params_for(name: str):
    """cm_params of the config (k, scan level ... )."""
    from .api import default_params
    return default_params(**(config(name).params if is_scale(name) else {}))


def prepare(name: str, outdir: str, contig_size: int | None = None, n_pairs: int | None = None,
            max_read_len: int = 300, write_reads: bool = False, seed_shift: int = 0, need_fasta: bool = False) -> dict:
    """Generate config `name` under outdir.  seed_shift changes only the read sample
    (rank-specific shards of the same reference)."""
    if is_scale(name):
        return _prepare_scale(name, outdir, n_pairs, seed_shift, max_read_len, write_reads, need_fasta)
    cfg = synth.CONFIGS[name]
    if n_pairs is not None or seed_shift:
        from dataclasses import replace
        cfg = replace(cfg, n_pairs=n_pairs or cfg.n_pairs, seed_reads=cfg.seed_reads + seed_shift)
    d = synth.generate(cfg, outdir, write_reads=write_reads)
    chroms = [(synth.chrom_name(i), synth.codes_to_ascii(c)) for i, c in enumerate(d["chroms"])]
    contigs, info = annotation.pack_genome(chroms, contig_size or annotation.DEF_CONTIG_SIZE)
    flat = annotation.flatten(d["gtf"], info, [len(c) for c in contigs], max_read_len)
    d.update(contigs=contigs, info=info, flat=flat, chr_names=[r[0] for r in info],
             s1=synth.codes_to_ascii(d["r1"]), s2=synth.codes_to_ascii(d["r2"]))
    return d


_FLAT_U4 = ("iv_spos", "iv_epos", "iv_max_end", "iv_min_end", "iv_max_next_exon", "iv_seg_off", "seg_start", "seg_end",
            "seg_next_exon_beg", "seg_gene_id", "seg_tid_off", "tid", "t2s_off", "gene_start", "gene_end", "chr_shift")


def _prepare_scale(name, outdir, n_pairs, seed_shift, max_read_len, write_reads, need_fasta):
    cfg = synth_scale.CONFIGS[name]
    fam = "grch38" + ("s" if cfg.scale != 1.0 else "")
    shared = os.path.join(os.path.dirname(os.path.abspath(outdir)), fam + "_ref")
    os.makedirs(shared, exist_ok=True)
    os.makedirs(outdir, exist_ok=True)
    with open(os.path.join(shared, ".lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        ref = synth_scale.reference(cfg, shared, need_fasta=need_fasta)
        mark = os.path.join(shared, f".packed_{max_read_len}")
        if not os.path.exists(mark):
            chroms = [(nm, synth.codes_to_ascii(np.asarray(c))) for nm, c in zip(ref["names"], ref["chroms"])]
            contigs, info = annotation.pack_genome(chroms, cfg.contig_size)
            del chroms
            flat = annotation.flatten(ref["gtf"], info, [len(c) for c in contigs], max_read_len)
            for c, seq in enumerate(contigs):
                np.save(os.path.join(shared, f"contig{c}.npy"), seq)
                np.savez(os.path.join(shared, f"flat{c}_{max_read_len}.npz"), **flat[c])
            json.dump([list(r) for r in info], open(os.path.join(shared, "info.json"), "w"))
            with open(ref["fa"] + ".packed.fa.index.info", "w") as f:   # GenomePacker::pack_genome's side file
                for nm, cid, s, e in info:
                    f.write(f"{cid}\t{s}\t{e}\t{nm}\n")
            del contigs, flat
            open(mark, "w").close()
        fcntl.flock(lk, fcntl.LOCK_UN)
    info = [tuple(r) for r in json.load(open(os.path.join(shared, "info.json")))]
    n_contigs = max(r[1] for r in info)
    contigs = [np.load(os.path.join(shared, f"contig{c}.npy"), mmap_mode="r") for c in range(n_contigs)]
    flat = []
    for c in range(n_contigs):
        z = np.load(os.path.join(shared, f"flat{c}_{max_read_len}.npz"))
        flat.append({k: (z[k] if z[k].ndim else z[k].item()) for k in z.files})
    r1, r2, kind = synth_scale.reads(cfg, ref, n_pairs=n_pairs, seed_shift=seed_shift)
    d = dict(ref)
    d.update(contigs=contigs, info=info, flat=flat, chr_names=[r[0] for r in info], r1=r1, r2=r2, kind=kind,
             s1=synth.codes_to_ascii(r1), s2=synth.codes_to_ascii(r2), shared=shared)
    if write_reads:
        d["fq1"] = os.path.join(outdir, "reads_1.fastq")
        d["fq2"] = os.path.join(outdir, "reads_2.fastq")
        synth.write_fastq(d["fq1"], r1, 1)
        synth.write_fastq(d["fq2"], r2, 2)
    return d


def load_into(mapper, w: dict) -> None:
    """Build every contig's index on the device and upload its annotation."""
    for c, seq in enumerate(w["contigs"]):
        mapper.build_contig(c, np.ascontiguousarray(seq))
        mapper.load_annotation(c, w["flat"][c])
