"""Host side of the drop-in (integration/host_feed.h: block reads, parallel FASTQ parse, header tokens -> cm_mread,
PAM / remain-FASTQ text in input order) checked on the CPU: oracle/_ref/circminer_feedtest is the drop-in's shim with a
stand-in for the C-ABI that answers each batch with the reference's own process_read (tests/hostfeed/feed_mock.cpp), so
every byte that differs from the reference binary is a host-feed bug.  Two packed contigs = two rounds, the second one
parsing the 23-token headers the first one wrote.  Stage 2 of the drop-in (integration/gpu_circ_detect.cpp) runs here as well,
with the batch entries answered by the host build of the device code, and in one test the mapping batches go through that build too.
The GPU counterpart is tests/test_gpu_dropin.py."""
import filecmp
import glob
import os
import subprocess
import sys

import pytest

from conftest import has_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "circminer")
FEED = os.path.join(ROOT, "oracle", "_ref", "circminer_feedtest")


def _run(binary, w, out, threads, env=None, extra=()):
    cmd = [binary, "-r", w["fa"], "-g", w["gtf"], "-1", w["fq1"], "-2", w["fq2"], "-k", "20", "-t", str(threads), "-P", "-z",
           "-o", out, *extra]
    e = dict(os.environ, LC_ALL="C")
    e.update(env or {})
    p = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=1500)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return p.stdout + p.stderr


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    from circminer_b200 import workload
    d = str(tmp_path_factory.mktemp("feed") / "t1")
    w = workload.prepare("T1", d, contig_size=3_500_000, n_pairs=6000, write_reads=True)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0, %r)\n"
                    "from oracle.refharness import RefHarness, default_params\n"
                    "RefHarness(default_params()).build_index(%r, contig_size=3500000)" % (ROOT, w["fa"])],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    return w


@pytest.mark.skipif(not (has_ref() and os.path.exists(FEED)), reason="oracle/_ref/circminer_feedtest not built")
@pytest.mark.parametrize("scan,threads,batch,devices,per", [("0", 4, 1000, "0", 2), ("2", 3, 257, "0,0,0", 2), ("0", 1, 262144, "0", 1),
                                                            ("0", 2, 500, "0,0", 1)])
def test_host_feed_outputs_are_byte_identical(dataset, tmp_path, scan, threads, batch, devices, per):
    """devices / per: contexts the batches are dealt to (CM_GPU_DEVICES x CM_GPU_CTX_PER_DEVICE: 2, 6, 1, 2 batches in flight)."""
    w = dataset
    ref_out, feed_out = str(tmp_path / "ref"), str(tmp_path / "feed")
    _run(REF, w, ref_out, 1, extra=("-a", scan))
    _run(FEED, w, feed_out, threads, env={"CM_GPU_BATCH": str(batch), "CM_GPU_DEVICES": devices, "CM_GPU_CTX_PER_DEVICE": str(per), "CM_GPU_STAGE2": "0"},
         extra=("-a", scan))
    ref_files = sorted(os.path.basename(f)[3:] for f in glob.glob(ref_out + "*") if os.path.isfile(f))
    feed_files = sorted(os.path.basename(f)[4:] for f in glob.glob(feed_out + "*") if os.path.isfile(f))
    assert ref_files == feed_files and any(f.endswith("circ_report") for f in ref_files), (ref_files, feed_files)
    assert any("remain_R1" in f for f in ref_files)
    for suffix in ref_files:
        a, b = ref_out + suffix, feed_out + suffix
        assert filecmp.cmp(a, b, shallow=False), f"{suffix} differs ({os.path.getsize(a)} vs {os.path.getsize(b)} bytes)"
    assert os.path.getsize(ref_out + ".mapping.pam") > 10000


_LOAD_CHECK = r"""
import ctypes as C, sys, zlib
sys.path.insert(0, %(root)r)
from oracle.refharness import RefHarness, default_params
h = RefHarness(default_params(), threads=2)
h.build_index(%(fa)r, contig_size=3500000, compact=%(compact)r)
h.open(%(fa)r, %(gtf)r, %(tmp)r)
L = h.lib
libc = C.CDLL(None)
libc.ftell.argtypes = [C.c_void_p]; libc.ftell.restype = C.c_long
g = lambda name, t: t.in_dll(L, name)
index_file = (%(fa)r + ".packed.fa.index").encode()   # already open: initLoadingHashTableMeta rewinds it
def state(flag):
    n = g("_ih_crefGenLen", C.c_int).value
    buf = C.string_at(g("_ih_crefGen", C.c_void_p).value, n * 8)
    return (flag, C.string_at(g("_ih_refGenName", C.c_void_p).value), g("_ih_refGenOff", C.c_int).value, g("_ih_refGenLen", C.c_int).value,
            zlib.crc32(buf), libc.ftell(g("_ih_fp", C.c_void_p).value))
def walk(fn):
    assert L.initLoadingHashTableMeta(index_file) == 1
    out, t = [], C.c_double()
    while True:
        flag = fn(C.byref(t))
        out.append(state(flag))
        if not flag or len(out) > 50:
            return out
a = walk(L.loadHashTable)
b = walk(L.loadCompressedRefGenome)
print(len(a), a == b, g("loadFullHashTable", C.c_int).value)
assert len(a) >= 2 and a == b, (a, b)
"""


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("compact", [False, True])
def test_genome_only_load_leaves_the_state_of_the_full_load(tmp_path, compact):
    """The drop-in renames mapping()'s loadHashTable to the reference's own loadCompressedRefGenome (integration/gpu_shim.h,
    cm_shim_loadHashTable): record by record of a full and of a compact index, both leave the same return value, contig name, offset,
    length, packed contig and file position -- everything the drop-in reads after the call."""
    from circminer_b200 import workload
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000, n_pairs=10, write_reads=True)
    code = _LOAD_CHECK % dict(root=ROOT, fa=w["fa"], gtf=w["gtf"], tmp=str(tmp_path / "ref"), compact=compact)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    n, same, full = p.stdout.split()[-3:]
    assert int(n) >= 2 and same == "True" and int(full) == (0 if compact else 1), p.stdout[-500:]


@pytest.mark.skipif(not (has_ref() and os.path.exists(FEED)), reason="oracle/_ref/circminer_feedtest not built")
@pytest.mark.parametrize("variant", ["smallcap", "double0", "ragged"])   # plain runs and scan level 2: test_whole_dropin_over_the_host_build
def test_stage2_of_the_dropin_on_the_cpu(dataset, tmp_path, variant):
    """The drop-in's do_process (integration/gpu_circ_detect.cpp: batches per contig, lines and CircRes in file order, pairs handed back
    to the reference's call_circ) with the batch entries answered by the host build of the device code (CM_MOCK_HOSTSIM): candidates.pam
    and circ_report byte for byte against the reference binary.  "smallcap": the host build with tiny fixed-size lists, so that some
    pairs come back as type -2 and are redone by the reference's own code in their place -- the files must not change.  "double0":
    CM_GPU_STAGE2_DOUBLE=0.  "scan2": scan level 2 (-a 2), which leaves stage 2 more candidates.  "ragged": reads of different
    lengths."""
    import re
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim_api import build
    so = "libhostsim_smallcap.so" if variant == "smallcap" else "libhostsim.so"
    build(so)
    w = dataset
    ref_out, feed_out = str(tmp_path / "ref"), str(tmp_path / "feed")
    extra = ("-a", "2") if variant == "scan2" else ()
    if variant == "ragged":                          # 70 % of the reads cut to a random length of 60 .. 150, each mate on its own
        import random
        rnd = random.Random(5)
        w = dict(w)
        for key in ("fq1", "fq2"):
            dst = str(tmp_path / (key + ".fastq"))
            with open(dataset[key]) as f, open(dst, "w") as g:
                while True:
                    h = f.readline()
                    if not h:
                        break
                    seq, plus, qual = f.readline().rstrip("\n"), f.readline(), f.readline().rstrip("\n")
                    n = rnd.randint(60, len(seq)) if rnd.random() < 0.7 else len(seq)
                    g.write(h + seq[:n] + "\n" + plus + qual[:n] + "\n")
            w[key] = dst
    _run(REF, w, ref_out, 1, extra=extra)
    cmd = [FEED, "-r", w["fa"], "-g", w["gtf"], "-1", w["fq1"], "-2", w["fq2"], "-k", "20", "-t", "2", "-P", "-z", "-o", feed_out, *extra]
    env = dict(os.environ, LC_ALL="C", CM_GPU_BATCH="2000", CM_GPU_VERBOSE="1", CM_MOCK_HOSTSIM=os.path.join(ROOT, "tests", "hostsim", so))
    if variant == "double0":
        env["CM_GPU_STAGE2_DOUBLE"] = "0"           # double-split pairs through the reference's call_circ, between the batched ones
    p = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=1500)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    for suffix in (".candidates.pam", ".circ_report", ".mapping.pam"):
        assert filecmp.cmp(ref_out + suffix, feed_out + suffix, shallow=False), f"{suffix} differs"
    assert os.path.getsize(ref_out + ".candidates.pam") > 1000
    log = p.stdout + p.stderr
    batches = re.findall(r"stage 2, contig \d+: (\d+) single-split pairs of \d+ in one batch, \d+ check_split_map attempts; (\d+) double-split", log)
    assert batches and sum(int(a) for a, _ in batches) > 50 and (sum(int(b) for _, b in batches) > 5) == (variant != "double0"), log[-1500:]
    # the annotation's two bit vectors went over as words of the std::bitsets (the stand-in compares every bit with operator[])
    assert "bit vectors copied by words" in log and "read bit by bit" not in log, log[-1500:]
    redone = int(re.search(r"\((\d+) of them redone after a device list overflowed\)", log).group(1))
    assert (redone > 0) == (variant == "smallcap"), (variant, redone)


@pytest.mark.skipif(not (has_ref() and os.path.exists(FEED)), reason="oracle/_ref/circminer_feedtest not built")
@pytest.mark.parametrize("case", ["scan0", "scan2", "dirty", "params"])
def test_whole_dropin_over_the_host_build(dataset, tmp_path, case):
    """Both stages of the drop-in binary with the device code's host build behind the C-ABI stand-in (CM_MOCK_STAGE1=hostsim: the mapping
    batches too, not the reference's process_read): the product's logic -- shim, host feed, device functions, stage-2 orchestration --
    inside the reference's own driver on a box without a GPU.  Every output file byte for byte against the reference binary; "dirty":
    reads with runs of N, lower-case stretches and lengths of 22 .. 45 bases among the others; "params": maximum edit distance 6, soft
    clip 10, band 5, seed limit 50, template length 800, intron 500 000, chain length 12, read length 200."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim_api import build
    build("libhostsim.so")
    w = dataset
    scan = "2" if case == "scan2" else "0"
    if case == "dirty":                              # runs of N, lower-case stretches, reads of 22 .. 45 bases
        import random
        rnd = random.Random(11)
        w = dict(w)
        for key in ("fq1", "fq2"):
            dst = str(tmp_path / (key + ".fastq"))
            with open(dataset[key]) as f, open(dst, "w") as g:
                while True:
                    h = f.readline()
                    if not h:
                        break
                    seq, plus, qual = f.readline().rstrip("\n"), f.readline(), f.readline().rstrip("\n")
                    r = rnd.random()
                    if r < 0.06:
                        a, n = rnd.randint(0, len(seq) - 6), rnd.randint(1, 5)
                        seq = seq[:a] + "N" * n + seq[a + n:]
                    elif r < 0.12:
                        a = rnd.randint(0, len(seq) - 30)
                        seq = seq[:a] + seq[a:a + 30].lower() + seq[a + 30:]
                    elif r < 0.18:
                        n = rnd.randint(22, 45)
                        seq, qual = seq[:n], qual[:n]
                    g.write(h + seq + "\n" + plus + qual + "\n")
            w[key] = dst
    ref_out, feed_out = str(tmp_path / "ref"), str(tmp_path / "feed")
    extra = ("-a", scan)
    if case == "params":                             # every mapping parameter off its default (cm_params as the shim fills it)
        extra += ("-e", "6", "-c", "10", "-w", "5", "-S", "50", "-T", "800", "-I", "500000", "-C", "12", "-l", "200")
    _run(REF, w, ref_out, 1, extra=extra)
    log = _run(FEED, w, feed_out, 3, env={"CM_GPU_BATCH": "1500", "CM_GPU_DEVICES": "0,0", "CM_MOCK_STAGE1": "hostsim",
                                          "CM_MOCK_HOSTSIM": os.path.join(ROOT, "tests", "hostsim", "libhostsim.so")}, extra=extra)
    assert "the mapping batches too" in log, log[-800:]
    ref_files = sorted(os.path.basename(f)[3:] for f in glob.glob(ref_out + "*") if os.path.isfile(f))
    feed_files = sorted(os.path.basename(f)[4:] for f in glob.glob(feed_out + "*") if os.path.isfile(f))
    assert ref_files == feed_files and any(f.endswith("circ_report") for f in ref_files), (ref_files, feed_files)
    for suffix in ref_files:
        a, b = ref_out + suffix, feed_out + suffix
        assert filecmp.cmp(a, b, shallow=False), f"{suffix} differs ({os.path.getsize(a)} vs {os.path.getsize(b)} bytes)"
    assert os.path.getsize(ref_out + ".mapping.pam") > 10000 and os.path.getsize(ref_out + ".candidates.pam") > 1000


@pytest.mark.skipif(not (has_ref() and os.path.exists(FEED)), reason="oracle/_ref/circminer_feedtest not built")
def test_whole_dropin_over_the_host_build_k14_compact(tmp_path):
    """C5's shape in small (T2: 2x250, k = 14 on a compact index, two packed contigs) through both stages of the drop-in over the host
    build: other seed geometry, longer reads and stage-2 windows than the k = 20 cases.  Every output file byte for byte."""
    from circminer_b200 import workload
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim_api import build
    build("libhostsim.so")
    w = workload.prepare("T2", str(tmp_path / "t2"), contig_size=1_600_000, write_reads=True)
    subprocess.run([sys.executable, "-c",
                    "import sys; sys.path.insert(0, %r)\n"
                    "from oracle.refharness import RefHarness, default_params\n"
                    "RefHarness(default_params(kmer=14)).build_index(%r, contig_size=1600000, compact=True)" % (ROOT, w["fa"])],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900)
    ref_out, feed_out = str(tmp_path / "ref"), str(tmp_path / "feed")

    def run(binary, out, threads, env=None):
        cmd = [binary, "-r", w["fa"], "-g", w["gtf"], "-1", w["fq1"], "-2", w["fq2"], "-k", "14", "-t", str(threads), "-P", "-z", "-o", out]
        p = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, LC_ALL="C", **(env or {})), timeout=1500)
        assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
        return p.stdout + p.stderr
    run(REF, ref_out, 1)
    log = run(FEED, feed_out, 3, {"CM_GPU_BATCH": "1000", "CM_MOCK_STAGE1": "hostsim", "CM_MOCK_HOSTSIM": os.path.join(ROOT, "tests", "hostsim", "libhostsim.so")})
    assert "the mapping batches too" in log and "Genome index type: Compact" in log, log[-800:]
    ref_files = sorted(os.path.basename(f)[3:] for f in glob.glob(ref_out + "*") if os.path.isfile(f))
    feed_files = sorted(os.path.basename(f)[4:] for f in glob.glob(feed_out + "*") if os.path.isfile(f))
    assert ref_files == feed_files and any(f.endswith("circ_report") for f in ref_files), (ref_files, feed_files)
    for suffix in ref_files:
        assert filecmp.cmp(ref_out + suffix, feed_out + suffix, shallow=False), f"{suffix} differs"
    assert os.path.getsize(ref_out + ".candidates.pam") > 1000
