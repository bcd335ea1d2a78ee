"""check_split_map on a repeat-rich workload (C3s = C3 at 1/50), CPU only: every call the reference's single-split path makes on
the pairs stage 1 keeps is recorded by the reference harness and replayed through the host build of the device code
(tests/hostsim); prints whether all outputs agree and the branch counters of the classification (CM_CIRC_TRACE).
Test infrastructure: uses oracle/.  python scripts/circ_cover.py [pairs]"""
import sys, numpy as np, os, ctypes as C, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from circminer_b200 import workload
from oracle.cpu_baseline import ensure_ref_index
from oracle.refharness import RefHarness, fresh_mreads
from hostsim_api import HostSim
d = os.environ.get('CM_COVER_DIR', '/tmp/cm_circ_cover'); os.makedirs(d,exist_ok=True)
npairs=int(sys.argv[1]) if len(sys.argv)>1 else 20000
w = workload.prepare("C3s", d+"/w", n_pairs=npairs)
p=workload.params_for("C3s")
t0=time.time(); print(ensure_ref_index(w,"C3s",p.kmer), time.time()-t0, flush=True)
h=RefHarness(p); h.open(w["fa"], w["gtf"], d+"/ref")
s1,s2=w["s1"],w["s2"]; n,L=s1.shape; lens=np.full(n,L,np.uint16)
hs = HostSim(p)
for c, seq in enumerate(w["contigs"]):
    hs.build_contig(c, seq); hs.load_annotation(c, w["flat"][c])
more=1; tot=0
while more:
    more=h.next_contig(); c=h.contig
    mr1,_=h.process_pe(s1,s2,lens,lens,fresh_mreads(n,p.max_ed))
    rec=h.circ_single(s1,s2,lens,lens,mr1)
    nn=len(rec['ret']); tot+=nn
    if not nn: continue
    got=hs.circ_check(c, rec)
    ok=[(got['ret']==rec['ret']).all(), (got['cr']==rec['cr']).all(), (got['r1']==rec['out_r1']).all(), (got['r2']==rec['out_r2']).all(), (got['pm']==rec['out_pm']).all()]
    print('contig',c,'calls',nn,'equal',ok, np.unique(rec['ret'],return_counts=True), flush=True)
    np.savez(d+f'/rec{c}.npz', **rec)
tr=(C.c_long*16)(); hs.lib.hs_circ_trace(tr); print('total',tot,'trace',list(tr))
os._exit(0)
