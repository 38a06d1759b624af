import sys, os, time, threading, hashlib
sys.path.insert(0, os.getcwd())
import nodexa_chain_core_b200 as nx
TAG=b"NodexaPW"
def header(nonce):
    base = hashlib.sha512(b"nodexa-equihash-bench").digest() * 3
    return base[:108] + nonce.to_bytes(32, "little")
for n,k,total in ((200,9,192),(144,5,48)):
    for P in (1,2,3,4):
        solvers=[nx.equihash.Solver(n,k,TAG) for _ in range(P)]
        for s in solvers: s.solve(header(10**6))
        counts=[0]*P
        def work(i):
            for nonce in range(i,total,P): counts[i]+=len(solvers[i].solve(header(nonce)))
        t0=time.perf_counter()
        th=[threading.Thread(target=work,args=(i,)) for i in range(P)]
        [x.start() for x in th]; [x.join() for x in th]
        dt=time.perf_counter()-t0
        print((n,k),"streams",P,"solutions",sum(counts),"Sol/s",round(sum(counts)/dt,1),"ms/nonce",round(dt/total*1e3,3), flush=True)
        for s in solvers: s.close()
