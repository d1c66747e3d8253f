import sys, json, numpy as np
sys.path.insert(0,'/root/repo')
import bench, __graft_entry__ as ge, torch
fec=ge.load_package().fec
n=296*4
eng=fec.FecEngine(device=0,harq_slots=n,max_batch_cbs=n)
llr=bench.make_llrs(n,-2.0,1)
for early in (False,True):
    d=fec.make_descs(n); bench.fill_descs(fec,d,n,early)
    out=np.zeros((n,1056),np.uint8)
    res=eng.run_host(d,llr,out); res=eng.run_host(d,llr,out)
    r=res['reserved']
    print('early',early,'setup',(r&255).mean()*256,'dematch',((r>>8)&255).mean()*1024,'decode',((r>>16)&255).mean()*4096,'tail',((r>>24)&255).mean()*256,'iters',res['nof_ldpc_iterations'].mean())
