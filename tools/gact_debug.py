import sys
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, longreadmapper_b200 as lm, oracle_lib
from test_gpu_parity import random_pairs
g=lm.sim_genome(11,200000)
ix=oracle_lib.OracleIndex.build(['c'],[g.tobytes()])
m=oracle_lib.gpu_mapper(lm,ix,0)
qs,ts=random_pairs(5,4000,100,250,0.1)
score,ops=m.cigar_align_batch(qs,ts,m.params(flags=8))
bad=0
for i,(q,t) in enumerate(zip(qs,ts)):
    es,eops,_=oracle_lib.gact(np.frombuffer(q,np.uint8),np.frombuffer(t,np.uint8))
    if es!=score[i] or not np.array_equal(eops,ops[i]):
        bad+=1
        if bad<=3:
            print('pair',i,'len',len(q),len(t),'score',es,score[i],'n',len(eops),len(ops[i]))
            print(' exp',''.join(map(str,eops)))
            print(' got',''.join(map(str,ops[i])))
print('bad',bad,'of',len(qs))
