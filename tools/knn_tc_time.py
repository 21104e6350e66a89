"""Developer aid: 1 M x 2000 k-NN, POPC kernel against the tensor-core kernel, same box; with ORBM_TC_STATS the per-role and
per-CTA timeline of one tensor-core call (the library reads a device pointer for its counters from that variable)."""
import os, sys, torch
sys.path.insert(0, '.')
stats = torch.zeros(16 + 4 * 4096, dtype=torch.int64, device="cuda")
want_stats = "ORBM_TC_STATS" in os.environ
if want_stats:
    os.environ["ORBM_TC_STATS"] = str(stats.data_ptr())
from mam3slam_b200 import ORBmatcher
from mam3slam_b200.matcher import KnnDatabase
NQ, K = 2000, 2
g = torch.Generator(device="cuda").manual_seed(1)
db = torch.randint(0, 256, (1_000_000, 32), dtype=torch.uint8, device="cuda", generator=g)
q = db[torch.randint(0, 1_000_000, (NQ,), device="cuda", generator=g)].clone(); q[:, 0] ^= 1
m = ORBmatcher(0.6, True)
ts = torch.cuda.Stream(); torch.cuda.set_stream(ts); st = ts.cuda_stream
def timed(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
for n in (1_000_000, 125_000):
    D = KnnDatabase(m, None, device_ptr=db.data_ptr(), n=n)
    res = []
    for fn in (m.knn_device, m.knn_device_tc):
        od = torch.empty((NQ, K), dtype=torch.int16, device="cuda"); oi = torch.empty((NQ, K), dtype=torch.int32, device="cuda")
        t = timed(lambda: fn(D, q.data_ptr(), NQ, K, od.data_ptr(), oi.data_ptr(), st))
        res.append((t, od.clone(), oi.clone()))
    same = torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2])
    print(f"rows {n}: popc kernel {res[0][0]*1e3:.1f} us, tensor-core kernel {res[1][0]*1e3:.1f} us, x{res[0][0]/res[1][0]:.2f}, identical {same}, "
          f"{n*NQ/res[1][0]/1e6:.0f} G evaluations/s")
    D.close()
if want_stats:
    D = KnnDatabase(m, None, device_ptr=db.data_ptr(), n=1_000_000)
    torch.cuda.synchronize(); stats.zero_(); torch.cuda.synchronize()
    od = torch.empty((NQ, K), dtype=torch.int16, device="cuda"); oi = torch.empty((NQ, K), dtype=torch.int32, device="cuda")
    m.knn_device_tc(D, q.data_ptr(), NQ, K, od.data_ptr(), oi.data_ptr(), st); torch.cuda.synchronize()
    v = stats.cpu().numpy().astype('uint64')
    n = max(int(v[6]), 1)
    print(f"per tile (CTA 0,0; {n} tiles): issue waits dEmpty {v[0]/n:.0f} bFull {v[1]/n:.0f} | expansion waits bEmpty {v[2]/n:.0f} works {v[3]/n:.0f} | epilogue waits dFull {v[4]/n:.0f} works {v[5]/n:.0f} cycles")
    ncta = int(v[11]); rec = v[16:16 + 4 * ncta].reshape(ncta, 4)
    t0 = int(rec[:, 1].min())
    import numpy as np
    order = np.argsort(rec[:, 1])
    print("ctas", ncta, "distinct SMs", len(set(rec[:, 0].tolist())), "span us", (int(rec[:, 2].max()) - t0) / 1e3)
    for i in list(order[:6]) + list(order[-6:]):
        print(f"  cta {i:3d} sm {int(rec[i,0]):3d} start {(int(rec[i,1]) - t0)/1e3:8.1f} us  end {(int(rec[i,2]) - t0)/1e3:8.1f} us")
    starts = (rec[:, 1].astype('int64') - t0) / 1e3
    print("start histogram (us):", np.histogram(starts, bins=8)[0].tolist(), np.histogram(starts, bins=8)[1].round(0).tolist())
