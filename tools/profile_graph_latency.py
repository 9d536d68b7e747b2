"""ncu target: one-query calls of the graph engine (latency form of K5) on 1M x 128, ef = 40.
ncu --set full --clock-control none --import-source on -k regex:graph_search_cta --launch-skip 60 -c 2 -o gpurun_out/r1_cta python tools/profile_graph_latency.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import expann_b200
from expann_b200 import datasets
base, q = datasets.make("sift", 1_000_000, 256, 128, variant="lowrank")
e = expann_b200.GraphEngine(); e.store_many_vectors(base); e.build_graph(M=16, M0=32, K=64); e.build(); e.set_ef_search(40)
for i in range(80):
    e.query_k_batch(q[i:i + 1], 10)
e.close()
