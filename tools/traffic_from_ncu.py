#!/usr/bin/env python
"""DRAM traffic per launch of each kernel class from the `ncu --set full` capture of one block
(tools/run.sh <tag> ncufull: patch GEMM, LN, qkv, attention, proj, LN, fc1, GELU, fc2), weighted by how
often each launch occurs in one DeiT-S forward.  usage: traffic_from_ncu.py raw.csv > profiles/r2_traffic.json"""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
def to_bytes(v, u):
    return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
seq = [("gemm_patch", "patch", 1), ("layernorm_requant", "norm1", 12.5), ("gemm_qkv", "qkv", 12),
       ("attention_i8", "attention", 12), ("gemm_proj", "proj", 12), ("layernorm_requant", "norm2", 12.5),
       ("gemm_fc1", "fc1", 12), ("gelu_lut_apply", "gelu", 12), ("gemm_fc2", "fc2", 12)]
assert len(rows) - 2 == len(seq), "expected the 9-kernel capture of tools/run.sh ncufull"
out = {}
for r, (cls, what, per_step) in zip(rows[2:], seq):
    b = to_bytes(r[ri], units[ri]) + to_bytes(r[wi], units[wi])
    o = out.setdefault(cls, {"bytes_per_step": 0.0, "launches_per_step": 0.0, "launches": {}})
    o["bytes_per_step"] += b * per_step
    o["launches_per_step"] += per_step
    o["launches"][what] = {"kernel": r[ki].split("(")[0].replace("void ", ""), "dram_bytes": b, "per_step": per_step}
for o in out.values():
    o["traffic_per_launch"] = o["bytes_per_step"] / o["launches_per_step"]
out["_source"] = "ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum, one capture per launch " \
                 "(cold L2: ncu flushes caches between replays); the classifier-head GEMM (1 launch, < 1 MB) is not in the capture"
print(json.dumps(out, indent=1))
