import sys, time
sys.argv=["x"]; sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import numpy as np
import bench
n = int(__import__('os').environ.get("N", "2048"))
t0=time.time()
free = bench.cpu_compiled("firstgenrepeater_clifford", n, None)
t1=time.time()
print("uncensored: nodes", free["nodes"], "trace_s %.1f"%free["trace_s"], "total %.1f"%(t1-t0), "complete", free["complete"], "censored", free["censored"])
cen = bench.cpu_compiled("firstgenrepeater_clifford", n, 3)
print("censored: nodes", cen["nodes"], "trace_s %.1f"%cen["trace_s"], "complete", cen["complete"], "censored", cen["censored"])
rf, rc = free["rows"], cen["rows"]
print(rf.shape)
mf = rf[:,1]/np.maximum(rf[:,0],1); mc = rc[:,1]/np.maximum(rc[:,0],1)
d = np.abs(mf-mc)
print("max bias", d.max(), "at row", d.argmax(), "end rows", mf[-2:], mc[-2:], "alive", rf[-2:,0], rc[-2:,0])
