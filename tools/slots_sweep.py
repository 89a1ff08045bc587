import sys; sys.path.insert(0,'/root/repo')
import bench
from hivemind_b200 import weights as W
from hivemind_b200.engine import Engine
B=512
pk,mv,offs=bench.workload(B)
eng=Engine(0,B,4)
assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0))
for w in range(4): assert eng.evaluatePositions(pk,mv,offs,w) is not None
for n in (1,2,3,4):
    eng.benchResidentMulti(20,n)
    ms,l=eng.benchResidentMulti(200,n)
    print(n,'slots', B*200/(ms*1e-3), 'pos/s', ms/200,'ms/batch')
