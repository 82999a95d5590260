import sys,time; sys.path.insert(0,'.')
import numpy as np
import stage_indexorganized_b200 as sg
from oracle import pyoracle
import bench
sg.init(0)
t,kind=bench.build_host_table(10_000_000, 1_000_000)
for i in range(2):
    t0=time.time(); s=sg.Snapshot.from_tree(t.root,4,1000); print("from_tree",time.time()-t0, flush=True); s.close()
