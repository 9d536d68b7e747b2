"""SASS mnemonic counts of the hot kernels in expann_b200/libexpann_gpu.so -> profiles/r2_sass_summary.txt
(evidence that the contraction / tile-movement kernels are tcgen05 + TMEM + TMA code, checkable without rebuilding)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "expann_b200", "libexpann_gpu.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs, cur = {}, None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
    elif cur is not None:
        funcs[cur].append(line)
want = ["tc_filter_pair_kernel", "scan_topk_f32_kernel", "tc_filter2_kernel", "graph_search_kernel", "graph_search_cta_kernel",
        "scan_topk_u8_kernel", "rerank_kernel", "merge_topk_kernel", "graph_search_inline_kernel", "graph_search_blocks_kernel"]
key = re.compile(r"^(UTC[A-Z0-9]*MMA|UTMALDG|UTMAPF|LDTM|UTCBAR|SYNCS|FFMA2|FADD2|IDP|FMNMX3|VIMNMX3|LDG|STG|LDS|STS|ATOMG|ATOMS|REDG|ELECT|UBLKPF|UBLKCP|CREDUX|REDUX|BAR)")
names = subprocess.run(["c++filt"], input="\n".join(funcs), capture_output=True, text=True).stdout.splitlines()
rows = []
for (name, lines), dn in zip(funcs.items(), names):
    if not any(w in dn for w in want):
        continue
    cnt, n = collections.Counter(), 0
    for ln in lines:
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            n += 1
            cnt[m.group(1)] += 1
    rows.append((dn, n, {k: v for k, v in cnt.items() if key.match(k)}))
rows.sort()
path = os.path.join(ROOT, "profiles", "r2_sass_summary.txt")
with open(path, "w") as f:
    f.write("SASS mnemonic counts per kernel of expann_b200/libexpann_gpu.so (cuobjdump -sass, sm_100a), the round-2 build.\n"
            "tcgen05 = UTCHMMA (kind::f16) / UTCIMMA (kind::i8) [.2CTA = cta_group::2], TMEM loads = LDTM, TMA = UTMALDG, commit = UTCBAR,\n"
            "mbarrier = SYNCS, packed fp32 = FFMA2 / FADD2, dp4a = IDP.4A, cp.async.bulk.prefetch.L2 = UBLKPF, warp reduce = CREDUX. Regenerate: python tools/sass_summary.py (this file is its output).\n\n")
    for dn, n, sel in rows:
        f.write(f"{dn}\n  instructions: {n}\n  " + ", ".join(f"{k} x{v}" for k, v in sorted(sel.items())) + "\n\n")
print("wrote", path, len(rows), "kernels")
