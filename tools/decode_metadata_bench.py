"""Decode-metadata step: DeviceDecodeState.step (one launch on persistent device tables) against the
reference's way -- Python lists -> torch.tensor(..., device=cuda) every step (model_runner.py:344-377, restated by the
oracle) -- for the BASELINE batch shapes.  Latency-bound integer work: reported as microseconds per step."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import paged_kv_oracle as O  # noqa: E402  (the reference-style CPU leg only)
from vllm_splitwise_b200.decode_metadata import DeviceDecodeState  # noqa: E402


def run(name, batch, ctx, bs=16, iters=200):
    nblk = (ctx + 64 + bs - 1) // bs
    st = DeviceDecodeState(batch, nblk, bs, "cuda")
    tables = [list(range(i * nblk, (i + 1) * nblk)) for i in range(batch)]
    for i in range(batch):
        st.add_sequence(i, tables[i], ctx - 1)
    B = st.set_batch(list(range(batch)))
    st.step(B)
    torch.cuda.synchronize()
    # device-timed: the kernel alone, back to back
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        st.seq_lens.fill_(ctx - 1)
        st.step(B)
    e1.record()
    torch.cuda.synchronize()
    dev_us = e0.elapsed_time(e1) / iters * 1000
    # host wall clock per step, outputs resident on the device at the end (what the model runner needs)
    t0 = time.perf_counter()
    for _ in range(iters):
        st.step(B)
    torch.cuda.synchronize()
    ours_us = (time.perf_counter() - t0) / iters * 1e6
    lens = [ctx] * batch
    cur = [t[:(ctx + bs - 1) // bs] for t in tables]
    t0 = time.perf_counter()
    for _ in range(max(10, iters // 10)):
        outs = O.prepare_decode(lens, cur, bs)
        outs = [o.cuda() for o in outs]
    torch.cuda.synchronize()
    ref_us = (time.perf_counter() - t0) / max(10, iters // 10) * 1e6
    out_bytes = batch * (8 + 8 + 4 + nblk * 4)
    print(json.dumps({"case": name, "workload": f"batch {batch}, ctx {ctx}, block {bs}: positions + slot_mapping + "
                      f"context_lens + block_tables [{batch},{nblk}]", "output_bytes": out_bytes,
                      "device_step_us_incl_len_reset_fill": round(dev_us, 2), "host_wall_us_per_step": round(ours_us, 2),
                      "reference_style_lists_to_device_us_per_step": round(ref_us, 1),
                      "speedup_vs_reference_style": round(ref_us / ours_us, 1), "bound": "latency",
                      "h2d_bytes_per_step": 0, "reference_h2d_bytes_per_step": out_bytes}), flush=True)


if __name__ == "__main__":
    run("c2_batch64_ctx4096", 64, 4096)
    run("c3_batch256_ctx4096", 256, 4096)
    run("c4_batch128_ctx8192", 128, 8192)
