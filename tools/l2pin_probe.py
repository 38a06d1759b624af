import os, sys, time
sys.path.insert(0, os.getcwd())
import torch
import nodexa_chain_core_b200 as nx
for epoch in (448, 0):
    ctx = nx.EpochContext(epoch, 0)
    buf = torch.empty(ctx.dag_bytes, dtype=torch.uint8, device="cuda:0")
    ts = []
    for _ in range(4):
        ctx.generate_range(0, ctx.dag_num_items, buf.data_ptr()); ts.append(round(ctx.last_timing()[0], 2))
    print("epoch", epoch, "NX_NO_L2_PIN=", os.environ.get("NX_NO_L2_PIN"), "build", ctx.build_times_ms(), "regen ms", ts, "checksum", ctx.checksum())
    ctx.close()
