"""Developer aid: fixed cost of the sharded k-NN's kernels without any peer (world = 1) against the plain scan, per shard size."""
import sys, torch
sys.path.insert(0, '.')
from mam3slam_b200 import ORBmatcher
from mam3slam_b200.matcher import KnnDatabase, KnnExchange
NQ, K = 2000, 2
g = torch.Generator(device="cuda").manual_seed(1)
db = torch.randint(0, 256, (1_000_000, 32), dtype=torch.uint8, device="cuda", generator=g)
q = db[torch.randint(0, 1_000_000, (NQ,), device="cuda", generator=g)].clone()
m = ORBmatcher(0.6, True)
x = KnnExchange(m, 0, 1, NQ, K)
od = torch.empty((NQ, K), dtype=torch.int16, device="cuda"); oi = torch.empty((NQ, K), dtype=torch.int32, device="cuda")
ts = torch.cuda.Stream(); torch.cuda.set_stream(ts)
st = ts.cuda_stream
def timed(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
for n in (1_000_000, 500_000, 250_000, 125_000):
    D = KnnDatabase(m, None, device_ptr=db.data_ptr(), n=n)
    a = timed(lambda: m.knn_device(D, q.data_ptr(), NQ, K, od.data_ptr(), oi.data_ptr(), st))
    b = timed(lambda: m.knn_sharded(x, D, q.data_ptr(), NQ, K, od.data_ptr(), oi.data_ptr(), st))
    print(f"rows {n}: scan+merge {a*1e3:.1f} us, sharded(world=1) {b*1e3:.1f} us, ideal from 1M = {n/1e6:.3f} x")
    D.close()
