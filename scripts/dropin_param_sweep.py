"""One parameter setting of the reference's CLI through the whole drop-in on the CPU (T1, 6 000 pairs): oracle/_ref/circminer_feedtest with
both stages answered by the host build of the device code (tests/hostfeed/feed_mock.cpp) against the reference binary, every output file
byte for byte.  The index must exist (the first call builds it).  Test infrastructure: executes oracle/_ref.
python scripts/dropin_param_sweep.py -w 7 -e 8 -c 12"""
import os, subprocess, sys, filecmp, glob
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from circminer_b200 import workload
K = os.environ.get('CM_CHECK_K', '20')               # k-mer size: its own data directory, the index is built at that k
D = os.environ.get('CM_CHECK_DIR', '/tmp/cm_param_sweep') + ('' if K == '20' else '_k' + K)
w = workload.prepare("T1", os.path.join(D, 't1'), contig_size=3_500_000, n_pairs=6000, write_reads=True)
if not os.path.exists(w['fa'] + '.packed.fa.index'):
    subprocess.run([sys.executable, '-c', 'import sys; sys.path.insert(0, %r)\nfrom oracle.refharness import RefHarness, default_params\nRefHarness(default_params(kmer=%s)).build_index(%r, contig_size=3500000)' % (ROOT, K, w['fa'])], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
extra=sys.argv[1:]
def run(binary,out,threads,env=None):
    cmd=[binary,"-r",w["fa"],"-g",w["gtf"],"-1",w["fq1"],"-2",w["fq2"],"-k",K,"-t",str(threads),"-P","-z","-o",out,*extra]
    e=dict(os.environ,LC_ALL="C"); e.update(env or {})
    p=subprocess.run(cmd,capture_output=True,text=True,env=e)
    return p.returncode, p.stdout+p.stderr
os.makedirs(D + '/pp',exist_ok=True)
for f in glob.glob(D + '/pp/*'): os.remove(f)
r1,_=run(ROOT+"/oracle/_ref/circminer",D + '/pp/ref',1)
r2,log=run(ROOT+"/oracle/_ref/circminer_feedtest",D + '/pp/feed',3,{"CM_GPU_BATCH":"1000","CM_GPU_VERBOSE":"1","CM_MOCK_STAGE1":"hostsim","CM_MOCK_HOSTSIM":ROOT+"/tests/hostsim/libhostsim.so"})
ref_files=sorted(os.path.basename(f)[3:] for f in glob.glob(D + '/pp/ref*') if os.path.isfile(f))
bad=[s for s in ref_files if not (os.path.exists(D + '/pp/feed'+s) and filecmp.cmp(D + '/pp/ref'+s,D + '/pp/feed'+s,shallow=False))]
st2=sum(int(x.split(' single-split')[0].split(': ')[-1]) for x in log.splitlines() if 'stage 2, contig' in x)
print(' '.join(extra), '| rc', r1, r2, '| stage-2 singles', st2, '| differs', bad, '| cand bytes', os.path.getsize(D + '/pp/ref.candidates.pam') if r1==0 else None)
if r2: print(log[-600:])
