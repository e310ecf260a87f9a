# experiment: two contexts on one GPU, each polishing half of the batch from its own host thread
import sys, time, threading
sys.path.insert(0, '.')
import bench
from sviper_b200 import api
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
vs = bench.make_batch(N, seed=1234)
one = api.Context(0)
pb = one.prepare(vs, True)
for _ in range(3): one.run_prepared(pb)
t = time.perf_counter()
for _ in range(10): one.run_prepared(pb)
t1 = (time.perf_counter() - t) / 10
print("one context: %.2f ms -> %.0f variants/s (e2e)" % (t1 * 1e3, N / t1))
for K in (2, 3, 4):
    ctxs = [api.Context(0) for _ in range(K)]
    parts = [vs[i::K] for i in range(K)]
    pbs = [c.prepare(p, True) for c, p in zip(ctxs, parts)]
    def work(c, p, n):
        for _ in range(n): c.run_prepared(p)
    def run(n):
        th = [threading.Thread(target=work, args=(c, p, n)) for c, p in zip(ctxs, pbs)]
        for x in th: x.start()
        for x in th: x.join()
    run(3)
    t = time.perf_counter(); run(10); t2 = (time.perf_counter() - t) / 10
    print("%d contexts: %.2f ms -> %.0f variants/s (e2e)" % (K, t2 * 1e3, N / t2))
    for c in ctxs: c.close()
