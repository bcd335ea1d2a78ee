"""Whole-binary parity of the drop-in: the reference's own driver with its mapping threads routed to
libcm_gpu.so (oracle/_ref/circminer_gpu = unmodified reference objects + integration/gpu_map_reads.cpp)
against the unmodified reference binary, same inputs, two packed contigs (two rounds with the MatchedRead
state carried through the remain FASTQ headers) and stage 2 (circRNA calling) run by the reference code on
both sides.  Every output file must be byte-identical: mapping.pam, the per-round remain FASTQs,
candidates.pam and circ_report."""
import filecmp
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import has_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "circminer")
GPU = os.path.join(ROOT, "oracle", "_ref", "circminer_gpu")

pytestmark = pytest.mark.gpu


def _run(binary, w, out, env=None, extra=()):
    cmd = [binary, "-r", w["fa"], "-g", w["gtf"], "-1", w["fq1"], "-2", w["fq2"], "-k", "20", "-t", "1", "-P", "-z", "-o", out, *extra]
    e = dict(os.environ, LC_ALL="C")
    e.update(env or {})
    p = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=1500)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return p


@pytest.mark.skipif(not (has_ref() and os.path.exists(GPU)), reason="oracle/_ref/circminer_gpu not built")
@pytest.mark.parametrize("scan,devices", [("0", "0"), ("2", "0"), ("0", "multi")])
def test_dropin_binary_outputs_are_byte_identical(tmp_path, scan, devices):
    """devices "multi": every visible GPU, or -- on a one-GPU box -- device 0 twice (two root contexts with their own copy
    of the index): batches dealt round-robin over four contexts, files still the reference's."""
    if devices == "multi":
        import torch
        devices = "all" if torch.cuda.device_count() > 1 else "0,0"
    from circminer_b200 import workload
    d = str(tmp_path / "t1")
    w = workload.prepare("T1", d, contig_size=3_500_000, write_reads=True)
    # the reference's index with two packed contigs (its CLI always packs at 1.1 Gbp)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0, %r)\n"
                    "from oracle.refharness import RefHarness, default_params\n"
                    "RefHarness(default_params()).build_index(%r, contig_size=3500000)" % (ROOT, w["fa"])],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    info = open(w["fa"] + ".packed.fa.index.info").read().split("\n")
    assert len({ln.split()[0] for ln in info if ln.strip()}) >= 2, "expected at least two packed contigs"
    ref_out, gpu_out = str(tmp_path / "ref"), str(tmp_path / "gpu")
    _run(REF, w, ref_out, extra=("-a", scan))
    pg = _run(GPU, w, gpu_out, env={"CM_GPU_BATCH": "4096" if devices == "0" else "1500", "CM_GPU_DEVICES": devices, "CM_GPU_VERBOSE": "1"},
              extra=("-a", scan))
    # stage 2's single-split pairs went through the device (integration/gpu_circ_detect.cpp), not through the reference's loop
    import re
    st2 = [tuple(int(x) for x in m) for m in re.findall(r"stage 2, contig \d+: (\d+) single-split pairs of \d+ in one batch, (\d+) check_split_map "
                                                        r"attempts; (\d+) double-split pairs, (\d+) events", pg.stderr)]
    assert st2 and sum(a for a, _, _, _ in st2) > 50 and sum(b for _, b, _, _ in st2) > 20 and sum(c for _, _, c, _ in st2) > 5, pg.stderr[-1500:]
    ref_files = sorted(os.path.basename(f)[3:] for f in glob.glob(ref_out + "*") if os.path.isfile(f))
    gpu_files = sorted(os.path.basename(f)[3:] for f in glob.glob(gpu_out + "*") if os.path.isfile(f))
    assert ref_files == gpu_files and any(f.endswith("circ_report") for f in ref_files), (ref_files, gpu_files)
    assert any("remain_R1" in f for f in ref_files)
    for suffix in ref_files:
        a, b = ref_out + suffix, gpu_out + suffix
        assert filecmp.cmp(a, b, shallow=False), f"{suffix} differs ({os.path.getsize(a)} vs {os.path.getsize(b)} bytes)"
    assert os.path.getsize(ref_out + ".mapping.pam") > 100000
    assert os.path.getsize(ref_out + ".circ_report") > 0


@pytest.mark.skipif(not (has_ref() and os.path.exists(GPU)), reason="oracle/_ref/circminer_gpu not built")
def test_dropin_annotation_cache_file(tmp_path):
    """The drop-in flattens the annotation once and keeps it as files: the second run uploads the files and
    writes the same outputs; a changed -l does not reuse them."""
    from circminer_b200 import api, workload
    d = str(tmp_path / "t1")
    w = workload.prepare("T1", d, contig_size=3_500_000, write_reads=True, n_pairs=5000)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0, %r)\n"
                    "from oracle.refharness import RefHarness, default_params\n"
                    "RefHarness(default_params()).build_index(%r, contig_size=3500000)" % (ROOT, w["fa"])],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    cache = str(tmp_path / "cache")
    os.makedirs(cache)
    env = {"CM_GPU_ANNOT_CACHE": cache, "CM_GPU_VERBOSE": "1", "CM_GPU_BATCH": "2000"}
    p1 = _run(GPU, w, str(tmp_path / "a"), env=env)
    files = sorted(os.listdir(cache))
    assert len(files) == len(w["contigs"]) and all(".cmannot." in f and f.endswith(".l300") for f in files), files
    assert "written to" in p1.stderr and "annotation of contig 0 from" not in p1.stderr
    # the files hold what the Python flattening of the same GTF gives (tests/test_cpu_parity.py checks that one against the reference)
    for c, f in enumerate(files):
        got = api.read_annotation_file(os.path.join(cache, f))
        for k in ("iv_spos", "iv_epos", "iv_seg_off", "seg_start", "seg_end", "seg_tid_off", "tid", "t2s", "gene_start", "gene_end", "chr_shift"):
            assert (got[k] == np.asarray(w["flat"][c][k])).all(), (c, k)
        nw = min(len(got["near_border"]), len(w["flat"][c]["near_border"])) - 1
        assert (got["near_border"][:nw] == w["flat"][c]["near_border"][:nw]).all() and (got["intronic"][:nw] == w["flat"][c]["intronic"][:nw]).all()
    p2 = _run(GPU, w, str(tmp_path / "b"), env=env)
    assert "annotation of contig 0 from" in p2.stderr and "written to" not in p2.stderr
    for suffix in (".mapping.pam", ".candidates.pam", ".circ_report", "_1_remain_R1.fastq", "_2_remain_R2.fastq"):
        assert filecmp.cmp(str(tmp_path / "a") + suffix, str(tmp_path / "b") + suffix, shallow=False), suffix
    p3 = _run(GPU, w, str(tmp_path / "c"), env=env, extra=("-l", "200"))
    assert "annotation of contig 0 from" not in p3.stderr and any(f.endswith(".l200") for f in os.listdir(cache))


@pytest.mark.skipif(not (has_ref() and os.path.exists(GPU)), reason="oracle/_ref/circminer_gpu not built")
def test_dropin_stage2_switches(tmp_path):
    """Stage 2 of the drop-in three ways -- both per-read paths on the device (default), the double-split pairs through the
    reference's call_circ (CM_GPU_STAGE2_DOUBLE=0: the contig text is unpacked lazily for them), the reference's whole do_process
    (CM_GPU_STAGE2=0) -- and the reference binary: candidates.pam and circ_report are the same bytes every time."""
    from circminer_b200 import workload
    d = str(tmp_path / "t1")
    w = workload.prepare("T1", d, contig_size=3_500_000, write_reads=True)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0, %r)\n"
                    "from oracle.refharness import RefHarness, default_params\n"
                    "RefHarness(default_params()).build_index(%r, contig_size=3500000)" % (ROOT, w["fa"])],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    outs = {"ref": str(tmp_path / "ref")}
    _run(REF, w, outs["ref"])
    for tag, env in (("dev", {}), ("dbl0", {"CM_GPU_STAGE2_DOUBLE": "0"}), ("st0", {"CM_GPU_STAGE2": "0"})):
        outs[tag] = str(tmp_path / tag)
        p = _run(GPU, w, outs[tag], env=dict(env, CM_GPU_VERBOSE="1"))
        assert ("stage 2, contig" in p.stderr) == (tag != "st0")
        if tag == "dbl0":
            assert "0 double-split pairs" in p.stderr
    for tag in ("dev", "dbl0", "st0"):
        for suffix in (".candidates.pam", ".circ_report"):
            assert filecmp.cmp(outs["ref"] + suffix, outs[tag] + suffix, shallow=False), (tag, suffix)
    assert os.path.getsize(outs["ref"] + ".candidates.pam") > 10000
