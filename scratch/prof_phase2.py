import sys, numpy as np
sys.path.insert(0,'/root/repo')
import bench, __graft_entry__ as ge
fec=ge.load_package().fec
n=296*4
eng=fec.FecEngine(device=0,harq_slots=n,max_batch_cbs=n)
llr=bench.make_llrs(n,-2.0,1)
d=fec.make_descs(n); bench.fill_descs(fec,d,n,False)
out=np.zeros((n,1056),np.uint8)
res=eng.run_host(d,llr,out); res=eng.run_host(d,llr,out)
r=res['reserved']
print('hard clks',(r&0xffff).mean()*16,'crc clks',(r>>16).mean()*16)
