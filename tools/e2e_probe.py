import sys, time, os
sys.path.insert(0, os.getcwd())
import bench
import lara_b200 as lb
import numpy as np
w, files = bench.make_workload("c3", 0)
store = bench.load_store(files, w["fmt"])
p = lb.Parameters()
flat = store.flat()
n = len(store)
ia, ib = np.triu_indices(n, k=1)
ia = ia.astype(np.int32); ib = ib.astype(np.int32)
ctx = lb.Context(p, device_ids=[0])
for k in range(9):
    t0 = time.perf_counter()
    ctx.set_sequences(flat)
    t1 = time.perf_counter()
    res = ctx.solve_pairs(ia, ib)
    t2 = time.perf_counter()
    st = ctx.stats()
    ctx.free(res)
    t3 = time.perf_counter()
    sys.stderr.write("CALL %d set %.1f solve %.1f free %.1f | ms_total %.1f device %.1f h2d %.1f d2h %.1f setup %.1f\n" % (
        k, 1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), st.ms_total, st.ms_device, st.ms_h2d, st.ms_d2h, st.ms_setup))
ctx.close()
