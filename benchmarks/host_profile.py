import sys, json, time
sys.path.insert(0,'/root/repo')
import numpy as np, torch
import openbnsl_b200
from openbnsl_b200 import networks as nw
L=openbnsl_b200.openbnsllib
net=nw.alarm(); codes=net.sample(100000,0); cards=[int(codes[:,v].max())+1 for v in range(net.n)]
df=L.base.DataframeWrapper.from_codes(codes, net.names, cards)
test=L.citest.GSquare(0.05)
L.profile_enable(True); df.to_device()
for i in range(5): L.structure_learning.pc(df,test,4,True)
L.profile_read(True)
t0=time.perf_counter()
R=20
acc={}
for i in range(R):
    L.structure_learning.pc(df,test,4,True)
    st=L.structure_learning.last_pc_stats()
    for k in ('enumerate_ms','evaluate_ms','resolve_ms'): acc[k]=acc.get(k,0)+st[k]
dt=(time.perf_counter()-t0)/R*1e3
p=L.profile_read(True)
print('step ms',dt, {k:v/R for k,v in acc.items()}, 'prep',p['host_prep_ms']/R,'wait',p['host_wait_ms']/R,'kernel',p['count_kernel_ms']/R, st['tests_per_depth'], st['batches_per_depth'])
