"""The drop-in binary's two stages on a box without a GPU, at C3s scale (C3 at 1/50, repeat-rich): oracle/_ref/circminer_feedtest =
the drop-in's shim and do_process over a stand-in for libcm_gpu.so (mapping answered by the reference's process_read, stage 2 by the
host build of the device code, tests/hostfeed/feed_mock.cpp) against the reference binary: candidates.pam and circ_report byte for
byte, mapping.pam as a multiset of lines.  CM_CHECK_HOSTSIM=libhostsim_smallcap.so: the host build with tiny list caps (pairs handed
back to the reference's call_circ).  Test infrastructure: executes oracle/_ref.  python scripts/dropin_cpu_check.py [pairs]"""
import os, subprocess, sys, time, filecmp, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from circminer_b200 import workload
from oracle.cpu_baseline import ensure_ref_index
npairs=int(sys.argv[1]) if len(sys.argv)>1 else 50000
d = os.environ.get('CM_CHECK_DIR', '/tmp/c3s_feed'); os.makedirs(d,exist_ok=True)
w = workload.prepare("C3s", d+"/w", n_pairs=npairs, write_reads=True)
p=workload.params_for("C3s")
t0=time.time(); print(ensure_ref_index(w,"C3s",p.kmer), round(time.time()-t0,1), flush=True)
print({k:(v if isinstance(v,str) else type(v)) for k,v in w.items() if k in ('fa','gtf','fq1','fq2')}, flush=True)
def run(binary,out,env=None,threads=8):
    cmd=[binary,"-r",w["fa"],"-g",w["gtf"],"-1",w["fq1"],"-2",w["fq2"],"-k","20","-t",str(threads),"-P","-o",out]
    e=dict(os.environ,LC_ALL="C"); e.update(env or {})
    t0=time.time(); pr=subprocess.run(cmd,capture_output=True,text=True,env=e); 
    print(os.path.basename(binary), 'rc',pr.returncode, round(time.time()-t0,1),'s', flush=True)
    if pr.returncode: print((pr.stdout+pr.stderr)[-1500:])
    return pr.stdout+pr.stderr
run(ROOT+"/oracle/_ref/circminer", d+"/ref")
stage1 = {"CM_MOCK_STAGE1": "hostsim"} if os.environ.get("CM_CHECK_STAGE1") else {}   # the mapping batches through the host build too
log=run(ROOT+"/oracle/_ref/circminer_feedtest", d+"/feed", env={**stage1, "CM_GPU_BATCH":"20000","CM_GPU_VERBOSE":"1","CM_MOCK_HOSTSIM":ROOT+"/tests/hostsim/"+os.environ.get("CM_CHECK_HOSTSIM","libhostsim.so")})
for ln in log.splitlines():
    if 'stage 2' in ln: print(ln[:300])
for suf in (".candidates.pam",".circ_report"):
    print(suf, filecmp.cmp(d+"/ref"+suf, d+"/feed"+suf, shallow=False), os.path.getsize(d+"/ref"+suf))
a=sorted(open(d+"/ref.mapping.pam","rb").readlines()); b=sorted(open(d+"/feed.mapping.pam","rb").readlines()); print('mapping.pam same multiset', a==b, len(a))
