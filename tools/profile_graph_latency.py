"""ncu target: one-query calls of the graph engine (latency form of K5) on 1M x 128, ef = 40.
ncu --set full --clock-control none --import-source on -k regex:graph_search_(cta|inline) --launch-skip 60 -c 2 -o gpurun_out/r2_lat python tools/profile_graph_latency.py [inline 0|1]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
inl = int(sys.argv[1]) if len(sys.argv) > 1 else 1
base, q = datasets.make("sift", 1_000_000, 256, 128, variant="lowrank")
e = expann_b200.GraphEngine(); e.set_option("graph_inline", inl); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
for i in range(80):
    e.query_k_batch(q[i:i + 1], 10)
e.close()
