#!/usr/bin/env python
"""bench.py -- the paged-KV decode hot path on B200, one JSON line per run.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3tp8|c3|c4|long|...] [--version v1|v2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...        # the reference's CPU path (oracle port) on the host cores
    python bench.py --workload n1 [...]         # SURVEY 8f N1, prefix prefill: TFLOP/s against the tensor peak

Step       = one layer-call of the decode path over one batch: reshape_and_cache of the step's new
             token per sequence, then paged_attention (v1 signature by default) over the whole context.
             Consecutive steps cycle over `--layers` distinct layer caches, each far larger than L2,
             with distinct physical blocks per (sequence, position), so nothing is served from L2.
Workload   = BASELINE.json configs[1] (C2: Llama-2-7B decode shape, 32 heads, head 128, block 16,
             batch 64, ctx 4096, fp16) per GPU.  N > 1: every rank holds the same per-GPU shape -- the
             KV-head shard of an N-times-wider model (weak scaling; KV-head sharding has no collective,
             SURVEY 8e); block tables are replicated.  `--workload c3tp8` etc. run the other configs.
value      = algorithmic GB/s of the whole job with inputs resident in HBM (bytes formula: BASELINE.md
             section 3), device-timed with CUDA events, max over ranks.
e2e        = the same metric through the repo's composed decode step (decode_step.PipelinedDecodeStep over
             ops.decode_step): every step copies its packed q/k/v rows from PINNED HOST memory to the device,
             builds positions / slot_mapping / context_lens / block_tables on the device, runs RoPE fused into
             the cache write and paged attention, and copies the attention output back to pinned host memory;
             copies and kernels of neighbouring steps overlap (two steps in flight), the host waits for every
             step's result.  `e2e.sync_each_step` is the same step as ONE captured graph (H2D -> kernels -> D2H)
             with a host synchronise after every replay (nothing overlaps).  The context grows by one token per
             step, as in a real decode loop; bytes are counted per step from the actual context length.
roofline   = attention kernel only: algorithmic bytes per launch / mean launch duration (CUDA events on
             the launching stream inside the timed region) against MEASURED_PEAKS.json hbm_gbs.
cpu_baseline = the oracle port of the reference's torch CPU path (tests/kernels/test_attention.py:58-111)
             on the host cores, on a bounded sample of the same workload.
extra      = the rest of BASELINE.json's metric in the same run (each block carries its own roofline / e2e):
             N = 1: c3tp8, c3 (unsharded), c4_v1, c4_v2, refbench, long_v2 / long_v1 (ctx 32768);
             N >= 2: c3_strong (the FIXED C3 problem split by KV head over the N ranks), kv_transfer (ranks
             0..N/2-1 -> N/2..N-1 through KVCacheCommManager over CUDA IPC; N = 8 is C5), allreduce (N4).
"""
import argparse
import json
import math
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "paged_attention_decode_hbm_throughput"
UNIT = "GB/s"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback, used only if MEASURED_PEAKS.json is absent

# name -> per-GPU shape.  Hq/Hkv are heads ON ONE GPU.
WORKLOADS = {
    "c2": dict(desc="C2 Llama-2-7B decode: 32 q / 32 kv heads, head 128, block 16, batch 64, ctx 4096, fp16",
               S=64, Hq=32, Hkv=32, D=128, bs=16, ctx=4096, dtype="float16", kv="auto"),
    "c3tp8": dict(desc="C3 Llama-2-70B GQA tp=8 shard: 8 q / 1 kv head, head 128, block 16, batch 256, ctx 4096, bf16",
                  S=256, Hq=8, Hkv=1, D=128, bs=16, ctx=4096, dtype="bfloat16", kv="auto"),
    "c3tp4": dict(desc="C3 Llama-2-70B GQA tp=4 shard: 16 q / 2 kv heads, batch 256, ctx 4096, bf16",
                  S=256, Hq=16, Hkv=2, D=128, bs=16, ctx=4096, dtype="bfloat16", kv="auto"),
    "c3tp2": dict(desc="C3 Llama-2-70B GQA tp=2 shard: 32 q / 4 kv heads, batch 256, ctx 4096, bf16",
                  S=256, Hq=32, Hkv=4, D=128, bs=16, ctx=4096, dtype="bfloat16", kv="auto"),
    "c3": dict(desc="C3 Llama-2-70B GQA unsharded: 64 q / 8 kv heads, head 128, block 16, batch 256, ctx 4096, bf16",
               S=256, Hq=64, Hkv=8, D=128, bs=16, ctx=4096, dtype="bfloat16", kv="auto"),
    "c4": dict(desc="C4 Mistral-7B GQA: 32 q / 8 kv heads, head 128, block 16, batch 128, ctx 8192, bf16 query, fp8_e5m2 KV",
               S=128, Hq=32, Hkv=8, D=128, bs=16, ctx=8192, dtype="bfloat16", kv="fp8_e5m2"),
    "refbench": dict(desc="reference benchmark default (benchmarks/kernels/benchmark_paged_attention.py): 64 q / 8 kv heads, "
                          "head 128, block 16, batch 8, ctx 4096, fp16",
                     S=8, Hq=64, Hkv=8, D=128, bs=16, ctx=4096, dtype="float16", kv="auto"),
    "long": dict(desc="long context: 32 q / 8 kv heads, head 128, block 16, batch 16, ctx 32768, bf16 (the op wrapper picks "
                      "v2 above 8192 tokens, paged_attn.py:62-63)",
                 S=16, Hq=32, Hkv=8, D=128, bs=16, ctx=32768, dtype="bfloat16", kv="auto"),
    "small": dict(desc="small batch: 32 q / 8 kv heads, head 128, block 16, batch 4, ctx 8192, bf16 query, fp8_e5m2 KV",
                  S=4, Hq=32, Hkv=8, D=128, bs=16, ctx=8192, dtype="bfloat16", kv="fp8_e5m2"),
    "c1": dict(desc="C1 OPT-125m shape: 12 heads, head 64, block 16, batch 8, ctx 512, fp32",
               S=8, Hq=12, Hkv=12, D=64, bs=16, ctx=512, dtype="float32", kv="auto"),
}

# what the N = 1 line carries besides the headline workload: (key, workload, version)
EXTRA_N1 = [("c3tp8", "c3tp8", "v1"), ("c3", "c3", "v1"), ("c4_v1", "c4", "v1"), ("c4_v2", "c4", "v2"),
            ("refbench", "refbench", "v1"), ("long_v2", "long", "v2"), ("long_v1", "long", "v1")]

# dram__bytes_read.sum + dram__bytes_write.sum per attention launch, taken from the committed `ncu --set full`
# captures of the same commands (mean of the captured launches) -- NOT measured by this run; the file is named in
# roofline.traffic_source.  Workloads without a capture report traffic null.
NCU_TRAFFIC = {
    "c2": (4_295_645_000 + 4_476_672, "profiles/r05_c2_v1_attn_ncu_full.txt"),
    "c4": (2_148_934_000 + 3_824_640, "profiles/r05_c4_v1_attn_ncu_full.txt"),
    "c3tp8": (537_801_216 + 6_942_464, "profiles/r05_c3tp8_v1_attn_ncu_full.txt"),
}


def elt_bytes(name: str) -> int:
    return {"float32": 4, "float16": 2, "bfloat16": 2}[name]


def attention_bytes(w, ctx=None) -> int:
    """Algorithmic bytes of one paged_attention call (SURVEY 8d / BASELINE.md section 3):
    K+V once per KV head + q in + out + block table + context_lens."""
    ctx = w["ctx"] if ctx is None else ctx
    cache_elt = 1 if w["kv"] == "fp8_e5m2" else elt_bytes(w["dtype"])
    kv = w["S"] * ctx * w["Hkv"] * w["D"] * 2 * cache_elt
    qo = 2 * w["S"] * w["Hq"] * w["D"] * elt_bytes(w["dtype"])
    tables = w["S"] * math.ceil(ctx / w["bs"]) * 4 + w["S"] * 4
    return kv + qo + tables


def reshape_and_cache_bytes(w) -> int:
    """N tokens x Hkv x D x (K,V) x (read T + write cache elt) + slot_mapping."""
    cache_elt = 1 if w["kv"] == "fp8_e5m2" else elt_bytes(w["dtype"])
    return w["S"] * w["Hkv"] * w["D"] * 2 * (elt_bytes(w["dtype"]) + cache_elt) + w["S"] * 8


def kernel_name(w) -> str:
    """Which consumer the library's dispatch picks (csrc/paged_attention.cu plan_tile)."""
    grouped = w["Hq"] != w["Hkv"] and w["dtype"] != "float32" and w["bs"] in (16, 32)
    return "paged_attn_kernel<MmaImpl> (persistent, mma.sync consumer)" if grouped else \
        "paged_attn_kernel<SimtImpl> (persistent, SIMT consumer)"


def step_bytes(w, ctx=None) -> int:
    return attention_bytes(w, ctx) + reshape_and_cache_bytes(w)


def workload_config(w, version, layers, world=1, per_gpu_note=True):
    """The `config` object: the same keys for this arm and for `--impl reference` (the driver compares them)."""
    return {"workload": w["desc"] + (" per GPU (KV-head shard of an N-times-wider model)" if world > 1 and per_gpu_note else ""),
            "version": version, "layers_cycled": layers,
            "l2": "inputs larger than L2: each layer cache is %.2f GB, distinct physical blocks, %d layers cycled"
                  % (attention_bytes(w) / 1e9, layers),
            "step": "reshape_and_cache + paged_attention_%s for one layer" % version}


def max_over_ranks(values, device, world):
    """Element-wise MAX of a list of floats over all ranks (identity when world == 1)."""
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def dist_env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback 6.65 TB/s (B200_PROFILING.md); MEASURED_PEAKS.json absent"


# ---------------------------------------------------------------------------------------------
# clocks under load (NVML in a thread; nvidia-smi as a fallback)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index: int, period_s: float = 0.01):
        self.index, self.period = index, period_s
        self.samples, self.mask, self.max_mhz, self.power = [], 0, None, []
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except ValueError:
                    phys = index
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _loop(self):
        n = self._nvml
        while not self._stop.is_set():
            try:
                self.samples.append(int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
                self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                self.power.append(n.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        if not self.samples:  # NVML unavailable or region shorter than one period: one nvidia-smi reading
            try:
                import subprocess
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=20)
                a, b = out.stdout.strip().split(",")
                self.samples, self.max_mhz = [int(a)], int(b)
            except Exception:
                pass
        reasons = [name for bit, name in self.REASONS.items() if self.mask & bit]
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(self.samples),
                "power_w_max": round(max(self.power), 1) if self.power else None}


# ---------------------------------------------------------------------------------------------
# workload construction (device)
# ---------------------------------------------------------------------------------------------
def build_device_workload(w, num_layers, device, seed, spare_blocks=0):
    """Layer caches by the reference fixture recipe (vllm/utils.py:366-392), distinct physical blocks per
    (sequence, position) in a random permutation (SURVEY 8d rule 1).  `spare_blocks` extra blocks per sequence
    back the tokens the e2e decode loop appends."""
    import torch
    from vllm_splitwise_b200.utils import create_kv_caches_with_random
    dt = getattr(torch, w["dtype"])
    S, ctx, bs = w["S"], w["ctx"], w["bs"]
    nblk = math.ceil(ctx / bs)
    ncol = nblk + spare_blocks
    NB = S * ncol
    g = torch.Generator().manual_seed(seed)
    table = torch.randperm(NB, generator=g).view(S, ncol).to(torch.int32)
    bt = table[:, :nblk].contiguous()
    kcs, vcs = create_kv_caches_with_random(NB, bs, num_layers, w["Hkv"], w["D"], w["kv"], dt, seed, device)
    scale = float(w["D"] ** -0.5)
    host = dict(
        q=torch.empty(S, w["Hq"], w["D"], dtype=dt).uniform_(-scale, scale, generator=g).pin_memory(),
        key=torch.empty(S, w["Hkv"], w["D"], dtype=dt).uniform_(-scale, scale, generator=g).pin_memory(),
        value=torch.empty(S, w["Hkv"], w["D"], dtype=dt).uniform_(-scale, scale, generator=g).pin_memory(),
        bt=bt.pin_memory(),
        cl=torch.full((S,), ctx, dtype=torch.int32).pin_memory(),
        # the step's new token is the last one of the context: slot = table[(ctx-1)//bs]*bs + (ctx-1)%bs
        slots=(bt[:, (ctx - 1) // bs].long() * bs + (ctx - 1) % bs).pin_memory(),
    )
    dev = {k: v.to(device) for k, v in host.items()}
    return dict(kcs=kcs, vcs=vcs, host=host, dev=dev, scale=scale, NB=NB, table=table)


# ---------------------------------------------------------------------------------------------
# one workload on this rank: device-resident timing + end to end.  Returns per-rank times; the caller takes the
# max over ranks.
# ---------------------------------------------------------------------------------------------
def measure_workload(w, version, steps, warmup, layers, device, seed, barrier, e2e=True, clocks_index=None):
    import torch
    from vllm_splitwise_b200 import _lib
    from vllm_splitwise_b200._C import cache_ops, ops
    lib = _lib.load()
    total_e2e_steps = warmup + steps
    spare = (total_e2e_steps + w["bs"] - 1) // w["bs"] + 1 if e2e else 0  # blocks the e2e decode loops append into
    wl = build_device_workload(w, layers, device, seed, spare_blocks=spare)
    d = wl["dev"]
    S, Hq, D = w["S"], w["Hq"], w["D"]
    P = math.ceil(w["ctx"] / 512)
    out = torch.empty_like(d["q"])
    if version == "v2":
        tmp = torch.empty(S, Hq, P, D, dtype=out.dtype, device=device)
        es = torch.empty(S, Hq, P, dtype=torch.float32, device=device)
        ml = torch.empty_like(es)

    def attention(layer):
        if version == "v1":
            ops.paged_attention_v1(out, d["q"], wl["kcs"][layer], wl["vcs"][layer], w["Hkv"], wl["scale"], d["bt"],
                                   d["cl"], w["bs"], w["ctx"], None, w["kv"])
        else:
            ops.paged_attention_v2(out, es, ml, tmp, d["q"], wl["kcs"][layer], wl["vcs"][layer], w["Hkv"], wl["scale"],
                                   d["bt"], d["cl"], w["bs"], w["ctx"], None, w["kv"])

    def step(i, ev=None):
        layer = i % layers
        cache_ops.reshape_and_cache(d["key"], d["value"], wl["kcs"][layer], wl["vcs"][layer], d["slots"], w["kv"])
        if ev is not None:
            ev[0].record()
        attention(layer)
        if ev is not None:
            ev[1].record()

    # ---- device-resident timing ----
    for i in range(warmup):
        step(i)
    barrier()
    attn_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t_beg, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(clocks_index) if clocks_index is not None else None
    launches0 = lib.b200_kv_launch_count()
    if sampler:
        sampler.start()
    barrier()
    t_beg.record()
    for i in range(steps):
        step(i, attn_ev[i])
    t_end.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    launches = lib.b200_kv_launch_count() - launches0
    res = dict(total_ms=t_beg.elapsed_time(t_end), attn_ms=sum(a.elapsed_time(b) for a, b in attn_ev) / steps,
               launches=int(launches), clocks=clocks, steps=steps)
    if not e2e:
        return res

    # ---- end to end through the composed decode step, host buffers ----
    from vllm_splitwise_b200.decode_metadata import DeviceDecodeState
    from vllm_splitwise_b200.decode_step import DecodeStepRunner, PipelinedDecodeStep
    from vllm_splitwise_b200.utils import rotary_cos_sin_cache
    dt = getattr(torch, w["dtype"])
    ctx, bs = w["ctx"], w["bs"]
    ncol = wl["table"].shape[1]
    max_ctx = ctx + total_e2e_steps + 1
    state = DeviceDecodeState(S, ncol, bs, device, max_batch=S)
    state.block_tables.copy_(wl["table"])
    state.set_batch(list(range(S)))
    cos_sin = rotary_cos_sin_cache(D, max_ctx + 8, 10000.0, dt, device)
    runner = DecodeStepRunner(state, Hq, w["Hkv"], D, wl["scale"], dt, w["kv"], cos_sin, True, None)
    kv_caches = list(zip(wl["kcs"], wl["vcs"]))
    width = (Hq + 2 * w["Hkv"]) * D
    qkv_rows = torch.empty(S, width, dtype=dt).uniform_(-wl["scale"], wl["scale"])
    h2d_bytes = qkv_rows.numel() * qkv_rows.element_size()
    d2h_bytes = S * Hq * D * qkv_rows.element_size()
    first_ctx = ctx  # context length of the first TIMED step of each e2e loop

    # (a) one graph per layer cache: H2D -> metadata -> RoPE + cache write -> attention -> D2H; sync after every replay
    caps = [runner.capture(S, [kv_caches[l]], max_ctx) for l in range(layers)]
    for c in caps:
        c.qkv_host[0].copy_(qkv_rows)
    state.seq_lens.fill_(first_ctx - 1 - warmup)  # the step advances it before use: first timed context == ctx
    torch.cuda.synchronize()
    for i in range(warmup):
        caps[i % layers].replay()
        caps[i % layers].synchronize()
    barrier()
    e_beg, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_beg.record()
    for i in range(steps):
        c = caps[i % layers]
        c.replay()
        c.synchronize()  # the caller consumes the step's result on the host
    e_end.record()
    barrier()
    res["e2e_sync_ms"] = e_beg.elapsed_time(e_end)
    assert int(state.seq_lens[0].item()) == first_ctx - 1 + steps

    # (b) pipelined: copies of neighbouring steps overlap the kernels, two steps in flight
    pipe = PipelinedDecodeStep(runner, S, kv_caches, max_ctx, depth=2)
    for s in pipe.slots:
        s.qkv_host.copy_(qkv_rows)
    state.seq_lens.fill_(first_ctx - 1 - warmup)
    torch.cuda.synchronize()
    for i in range(warmup):
        pipe.submit(i, i % layers)
        pipe.result(i)
    barrier()
    t0 = time.perf_counter()
    e_beg.record()
    for i in range(steps):
        pipe.submit(warmup + i, (warmup + i) % layers)
        if i > 0:
            pipe.result(warmup + i - 1)  # the host consumes every step's output, one step behind the submission
    pipe.result(warmup + steps - 1)
    wall_ms = (time.perf_counter() - t0) * 1e3
    e_end.record()
    barrier()
    # device-side span of the pipelined loop; the host-side wall clock is reported next to it
    res["e2e_pipe_ms"] = max(e_beg.elapsed_time(e_end), wall_ms)
    res["e2e_bytes"] = sum(step_bytes(w, first_ctx + i) for i in range(steps))
    res["h2d_bytes"], res["d2h_bytes"] = h2d_bytes, d2h_bytes
    return res


def summarize(w, version, layers, res, world, peak, peak_src, name=None):
    """JSON block of one measured workload (times in `res` are already the max over ranks)."""
    steps = res["steps"]
    job_bytes = step_bytes(w) * steps * world
    value = job_bytes / (res["total_ms"] * 1e-3) / 1e9
    achieved = attention_bytes(w) / (res["attn_ms"] * 1e-3) / 1e9
    traffic, traffic_src = NCU_TRAFFIC.get(name, (None, None)) if version == "v1" else (None, None)
    block = {
        "value": round(value, 2), "unit": UNIT, "ms_per_step": round(res["total_ms"] / steps, 5), "steps": steps,
        "config": workload_config(w, version, layers, world),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": kernel_name(w), "kernel_ms": round(res["attn_ms"], 5),
                     "algorithmic_bytes_per_launch": attention_bytes(w), "peak_source": peak_src,
                     "frac_of_8TBps_spec": round(achieved / 8000.0, 4)},
        "gpu_launches": res["launches"],
    }
    if "e2e_pipe_ms" in res:
        e2e_val = res["e2e_bytes"] * world / (res["e2e_pipe_ms"] * 1e-3) / 1e9
        sync_val = res["e2e_bytes"] * world / (res["e2e_sync_ms"] * 1e-3) / 1e9
        block["e2e"] = {"value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": res["h2d_bytes"],
                        "d2h_bytes_per_step": res["d2h_bytes"], "ms_per_step": round(res["e2e_pipe_ms"] / steps, 5),
                        "frac_of_device_value": round(e2e_val / value, 4),
                        "through": "decode_step.PipelinedDecodeStep: per step H2D of the packed qkv rows (pinned) -> "
                                   "device-side metadata -> RoPE + cache write -> paged attention -> D2H of the output; "
                                   "two steps in flight, the host waits for every result",
                        "sync_each_step": {"value": round(sync_val, 2),
                                           "ms_per_step": round(res["e2e_sync_ms"] / steps, 5),
                                           "frac_of_device_value": round(sync_val / value, 4),
                                           "through": "DecodeStepRunner.capture: ONE CUDA graph per step (H2D, 3 kernels, "
                                                      "D2H), host synchronise after every replay"}}
    return block


# ---------------------------------------------------------------------------------------------
# CPU baseline (oracle port of the reference torch path) -- rank 0 only
# ---------------------------------------------------------------------------------------------
def cpu_sample_inputs(w, n_seqs, seed=0):
    """CPU tensors of the first n_seqs sequences of the workload shape (same recipe, compact cache)."""
    import torch
    dt = getattr(torch, w["dtype"])
    nblk = math.ceil(w["ctx"] / w["bs"])
    NB = n_seqs * nblk
    g = torch.Generator().manual_seed(seed)
    scale = float(w["D"] ** -0.5)
    cache_elt = 1 if w["kv"] == "fp8_e5m2" else elt_bytes(w["dtype"])
    x = 16 // cache_elt
    kc = torch.empty(NB, w["Hkv"], w["D"] // x, w["bs"], x, dtype=torch.float32).uniform_(-scale, scale, generator=g)
    vc = torch.empty(NB, w["Hkv"], w["D"], w["bs"], dtype=torch.float32).uniform_(-scale, scale, generator=g)
    if w["kv"] == "fp8_e5m2":
        from oracle import paged_kv_oracle as O
        kc, vc = O.e5m2_to_tensor(O.e5m2_from_tensor(kc), dt), O.e5m2_to_tensor(O.e5m2_from_tensor(vc), dt)
    else:
        kc, vc = kc.to(dt), vc.to(dt)
    q = torch.empty(n_seqs, w["Hq"], w["D"], dtype=dt).uniform_(-scale, scale, generator=g)
    bt = torch.randperm(NB, generator=g).view(n_seqs, nblk).to(torch.int32)
    cl = torch.full((n_seqs,), w["ctx"], dtype=torch.int32)
    return q, kc, vc, bt, cl, scale


def time_cpu_oracle(w, n_seqs, repeats=1):
    """Seconds per pass of the oracle over n_seqs sequences (all torch threads)."""
    import torch
    from oracle import paged_kv_oracle as O  # the one place bench.py executes oracle/: the CPU baseline
    q, kc, vc, bt, cl, scale = cpu_sample_inputs(w, n_seqs)
    out = torch.empty_like(q)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        O.single_query_cached_kv_attention(out, q, w["Hq"] // w["Hkv"], kc, vc, bt, cl, scale, None)
        best = min(best, time.perf_counter() - t0)
    return best


def sample_bytes(w, n_seqs) -> int:
    return attention_bytes(dict(w, S=n_seqs))


def cpu_baseline(w, budget_s=15.0):
    import torch
    t1 = time_cpu_oracle(w, 1)
    n = max(1, min(w["S"], int(budget_s / max(t1, 1e-3))))
    t = time_cpu_oracle(w, n) if n > 1 else t1
    return {"value": round(sample_bytes(w, n) / t / 1e9, 3), "unit": UNIT, "cores": torch.get_num_threads(),
            "kind": "port",
            "sample": f"attention over the first {n} of {w['S']} sequences (ctx {w['ctx']}, {w['Hq']}q/{w['Hkv']}kv heads), "
                      f"{t:.2f} s; oracle port of tests/kernels/test_attention.py:58-111 (vectorised gather), torch CPU"}


# ---------------------------------------------------------------------------------------------
# the reference arm: CPU only, rank 0 only
# ---------------------------------------------------------------------------------------------
def run_reference(args, w):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
    t1 = time_cpu_oracle(w, 1)
    # bounded sample per step so that (steps + warmup) passes end within a few minutes
    total_budget = 150.0
    n = max(1, min(w["S"], int(total_budget / (args.steps + args.warmup) / max(t1, 1e-3))))
    from oracle import paged_kv_oracle as O
    q, kc, vc, bt, cl, scale = cpu_sample_inputs(w, n)
    out = torch.empty_like(q)
    for _ in range(args.warmup):
        O.single_query_cached_kv_attention(out, q, w["Hq"] // w["Hkv"], kc, vc, bt, cl, scale, None)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.single_query_cached_kv_attention(out, q, w["Hq"] // w["Hkv"], kc, vc, bt, cl, scale, None)
    dt = time.perf_counter() - t0
    value = sample_bytes(w, n) * args.steps / dt / 1e9
    sample = (f"each step = attention over {n} of {w['S']} sequences (ctx {w['ctx']}); oracle port of the reference's "
              f"torch CPU path (tests/kernels/test_attention.py:58-111), vectorised gather")
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": w["dtype"], "data": "synthetic",
            # the same config object as the b200 arm prints (the bounded CPU sample is described in cpu_baseline)
            "config": workload_config(w, args.version, args.layers, max(1, args.gpus)),
            "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample},
            "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# N >= 2 extras
# ---------------------------------------------------------------------------------------------
def extra_c3_strong(args, rank, world, device, barrier, peak, peak_src, steps):
    """The FIXED C3 problem (S=256, 64 q / 8 kv heads, ctx 4096, bf16) split by KV head over the `world` ranks
    (sharding.shard_heads = the reference's partition, vllm/config.py:288-296), no data-path collective; and the
    unsharded problem on one GPU in the same run, for the strong-scaling efficiency T1 / (N * TN)."""
    from vllm_splitwise_b200.sharding import shard_heads
    full = WORKLOADS["c3"]
    sh = shard_heads(full["Hq"], full["Hkv"], world, rank)
    shard = dict(full, Hq=sh.num_q, Hkv=sh.num_kv,
                 desc=f"C3 Llama-2-70B GQA, KV heads sharded over tp={world}: {sh.num_q} q / {sh.num_kv} kv heads per GPU, "
                      f"head 128, block 16, batch 256, ctx 4096, bf16")
    r = measure_workload(shard, "v1", steps, args.warmup, args.layers, device, 1000 + rank, barrier)
    keys = ["total_ms", "attn_ms", "e2e_sync_ms", "e2e_pipe_ms"]
    for k, v in zip(keys, max_over_ranks([r[k] for k in keys], device, world)):
        r[k] = v
    one = measure_workload(full, "v1", steps, args.warmup, args.layers, device, 2000 + rank, barrier)
    for k, v in zip(keys, max_over_ranks([one[k] for k in keys], device, world)):
        one[k] = v
    if rank != 0:
        return None
    # whole-job numbers: the N shards together are one C3 layer-call
    block = summarize(shard, "v1", args.layers, r, world, peak, peak_src)
    block["config"]["workload"] = shard["desc"]
    block["scaling"] = "strong"
    block["unsharded_on_one_gpu"] = {"ms_per_step": round(one["total_ms"] / steps, 5),
                                     "e2e_ms_per_step": round(one["e2e_pipe_ms"] / steps, 5),
                                     "value": round(step_bytes(full) * steps / (one["total_ms"] * 1e-3) / 1e9, 2),
                                     "note": "the same C3 problem, all 64 q / 8 kv heads on one GPU, measured by every "
                                             "rank in this run (max over ranks)"}
    block["efficiency_vs_one_gpu"] = {"device": round(one["total_ms"] / (world * r["total_ms"]), 4),
                                      "e2e": round(one["e2e_pipe_ms"] / (world * r["e2e_pipe_ms"]), 4),
                                      "formula": "T(1 GPU, unsharded) / (N * T(N GPUs, sharded)), max over ranks"}
    return block


def extra_kv_transfer(args, rank, world, device):
    """C5: prompt ranks 0..N/2-1 -> token ranks N/2..N-1 through KVCacheCommManager over CUDA IPC."""
    import torch.distributed as dist
    from tools import kv_transfer_bench
    boot = dist.new_group(backend="gloo")
    # library-default put kernel and CTA budget; per-layer launches and 4 layers merged per launch
    lines = kv_transfer_bench.run(rank, world, device, boot, prompts=8, iters=4, warmup=2,
                                  modes=("bulk", "per_layer", "ref_api", "paced", "nccl"), sweep=[(-1, 0, 1), (-1, 0, 4)])
    if rank != 0:
        return None
    out = {"config": lines[0]["config"] if lines else None, "unit": "GB/s per GPU pair",
           "nvlink_nominal_GBps": kv_transfer_bench.NVLINK_NOMINAL_GBS,
           "nvlink_measured_peer_copy_GBps": kv_transfer_bench.NVLINK_MEASURED_GBS, "cases": {}}
    for l in lines:
        key = f"{l['pattern']}.{l['mode']}" + (f".lpl{l['layers_per_launch']}" if "layers_per_launch" in l else "")
        out["cases"][key] = {
            "value": l["value"], "best": l["best_GBps"], "aggregate": l["aggregate_GBps"],
            "frac_of_900": l["frac_of_nvlink_nominal_900"], "frac_of_770_measured": l["frac_of_nvlink_measured_770"],
            "bit_exact": l["bit_exact"], "prompts_in_flight": l["prompts_in_flight"], "ms": l["ms_mean"],
            "time_to_signal_ms": l.get("time_to_signal_ms"), "max_ctas": l.get("max_ctas"),
            "put_variant": l.get("put_variant"), "layers_per_launch": l.get("layers_per_launch"),
            "comm_stream": l.get("comm_stream"),
            "runs_per_prompt": l["runs_per_prompt"]}
    out["bit_exact"] = all(l["bit_exact"] for l in lines)
    return out


def extra_allreduce(args, rank, world, device):
    import torch.distributed as dist
    from tools import allreduce_bench
    lines = allreduce_bench.run(rank, world, device, dist.group.WORLD, iters=100)
    if rank != 0:
        return None
    return {"unit": "us per call, max over ranks", "cases": {l["case"]: {k: v for k, v in l.items() if k != "case"} for l in lines},
            "bit_exact": all(l["bit_exact_all_ranks"] for l in lines)}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_b200(args, w):
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch N > 1 with torch.distributed.run (one process per GPU)")
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    peak, peak_src = measured_peak()
    t_wall0 = time.perf_counter()
    keys = ["total_ms", "attn_ms", "e2e_sync_ms", "e2e_pipe_ms"]
    res = measure_workload(w, args.version, args.steps, args.warmup, args.layers, device, rank, barrier,
                           clocks_index=local_rank)
    for k, v in zip(keys, max_over_ranks([res[k] for k in keys], device, world)):
        res[k] = v
    head = summarize(w, args.version, args.layers, res, world, peak, peak_src, name=args.workload)

    extra, extra_errors = {}, {}
    if not args.no_extra:
        def guarded(name, fn):
            """An extra block must never take the headline down with it: a failure is recorded, not raised."""
            try:
                out = fn()
                if out is not None:
                    extra[name] = out
            except Exception as e:  # noqa: BLE001
                import traceback
                extra_errors[name] = f"{type(e).__name__}: {e}"
                traceback.print_exc(file=sys.stderr)
            torch.cuda.empty_cache()

        if world == 1:
            for key, wname, version in EXTRA_N1:
                if wname == args.workload and version == args.version:
                    continue

                def one(wname=wname, version=version, key=key):
                    ww = WORKLOADS[wname]
                    r = measure_workload(ww, version, args.extra_steps, args.warmup, args.layers, device, 100 + len(extra),
                                         barrier)
                    return summarize(ww, version, args.layers, r, 1, peak, peak_src, name=wname)
                guarded(key, one)
        else:
            guarded("c3_strong", lambda: extra_c3_strong(args, rank, world, device, barrier, peak, peak_src, args.extra_steps))
            if world % 2 == 0:
                guarded("kv_transfer", lambda: extra_kv_transfer(args, rank, world, device))
            guarded("allreduce", lambda: extra_allreduce(args, rank, world, device))

    if rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": w["dtype"] if w["kv"] == "auto" else f"{w['dtype']}+fp8_e5m2_kv",
            "data": "synthetic",
            "config": head["config"],
            "decode_tokens_per_s_per_layer": round(w["S"] * world / (head["ms_per_step"] * 1e-3), 1),
            "clocks": res["clocks"],
            "e2e": head["e2e"],
            "gpu_launches": head["gpu_launches"],
            "roofline": head["roofline"],
        }
        if extra:
            line["extra"] = extra
        if extra_errors:
            line["extra_errors"] = extra_errors
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(w)
        line["wall_s"] = round(time.perf_counter() - t_wall0, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# ---------------------------------------------------------------------------------------------
# prefix ("context") prefill through the reference-named host function.  GEMM-shaped: the
# metric is TFLOP/s and the roofline is the tensor peak.  `--workload n1` (single GPU).
# ---------------------------------------------------------------------------------------------
N1 = dict(B=8, H=32, Hkv=32, D=128, bs=16, L=2048, ctx=2048, dtype="float16",
          name="N1 Llama-2-7B chunked prefill: 8 sequences x (2048 new + 2048 cached tokens), 32 heads, head 128, "
               "block 16, fp16")


def n1_flops(w) -> float:
    return 4.0 * w["D"] * w["H"] * w["B"] * (w["L"] * w["ctx"] + w["L"] * (w["L"] + 1) / 2)


def run_n1(args):
    import torch
    w = N1
    B, H, Hkv, D, bs, L, ctx = (w[k] for k in ("B", "H", "Hkv", "D", "bs", "L", "ctx"))
    flops = n1_flops(w)
    if args.impl == "reference":  # CPU: the oracle's restatement of the reference kernels, on a bounded sample
        from oracle import paged_kv_oracle as O
        torch.manual_seed(0)
        sub = dict(w, B=1, H=4, Hkv=4, L=512, ctx=512)
        q = torch.rand(sub["L"], sub["H"], D).half(); k = torch.rand(sub["L"], sub["Hkv"], D).half(); v = torch.rand_like(k)
        nb = sub["ctx"] // bs
        kc = torch.rand(nb, sub["Hkv"], D // 8, bs, 8).half(); vc = torch.rand(nb, sub["Hkv"], D, bs).half()
        o = torch.zeros_like(q)
        fn = lambda: O.context_attention_fwd(q, k, v, o, kc, vc, torch.arange(nb).view(1, nb), torch.tensor([0]),  # noqa: E731
                                             torch.tensor([sub["L"] + sub["ctx"]]), torch.tensor([sub["ctx"]]), sub["L"])
        for _ in range(args.warmup):
            fn()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        dt = time.perf_counter() - t0
        val = n1_flops(sub) * args.steps / dt / 1e12
        sample = "1 sequence x (512 new + 512 cached), 4 heads per step; oracle port of prefix_prefill.py:10-199, torch CPU"
        print(json.dumps({"impl": "reference", "metric": "prefix_prefill_throughput", "value": round(val, 5),
                          "unit": "TFLOP/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "float16", "data": "synthetic",
                          "config": {"workload": w["name"], "sample": sample},
                          "cpu_baseline": {"value": round(val, 5), "unit": "TFLOP/s", "cores": torch.get_num_threads(),
                                           "kind": "port", "sample": sample},
                          "e2e": {"value": round(val, 5), "unit": "TFLOP/s", "h2d_bytes_per_step": 0,
                                  "d2h_bytes_per_step": 0}}), flush=True)
        return 0
    from vllm_splitwise_b200.prefix_prefill import context_attention_fwd
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device=dev).manual_seed(0)
    scale = D ** -0.5
    n = B * L
    mk = lambda *shape: torch.empty(*shape, dtype=torch.float16, device=dev).uniform_(-scale, scale, generator=g)  # noqa: E731
    nb = (ctx + bs - 1) // bs
    layers = []
    for _ in range(2):  # two distinct 0.5 GB K/V sets, alternated: larger than L2
        layers.append((mk(n, H, D), mk(n, Hkv, D), mk(n, Hkv, D), mk(B * nb, Hkv, D // 8, bs, 8), mk(B * nb, Hkv, D, bs)))
    bt = torch.randperm(B * nb, device=dev, generator=g).view(B, nb).int()
    start = (torch.arange(B, device=dev) * L).int()
    seq = torch.full((B,), L + ctx, dtype=torch.int32, device=dev)
    cl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    o = torch.empty(n, H, D, dtype=torch.float16, device=dev)
    step = lambda i: context_attention_fwd(*layers[i % 2][:3], o, *layers[i % 2][3:], bt, start, seq, cl, L)  # noqa: E731
    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib_launch_count()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    launches = _lib_launch_count() - launches0
    clocks = sampler.stop()
    # end to end: q, k, v of the step from pinned host memory, the output read back
    hq, hk, hv = (t.cpu().pin_memory() for t in layers[0][:3])
    ho = torch.empty(o.shape, dtype=o.dtype).pin_memory()
    dq, dk, dv = (torch.empty_like(t) for t in layers[0][:3])
    e2e_steps = max(3, min(args.steps, 10))
    torch.cuda.synchronize()
    e0.record()
    for i in range(e2e_steps):
        dq.copy_(hq, non_blocking=True); dk.copy_(hk, non_blocking=True); dv.copy_(hv, non_blocking=True)
        context_attention_fwd(dq, dk, dv, o, *layers[0][3:], bt, start, seq, cl, L)
        ho.copy_(o, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        peak_src = "MEASURED_PEAKS.json bf16_tflops (cuBLAS bf16 8192^3, burst)"
    except Exception:
        peak, peak_src = 2250.0, "fallback: nominal dense bf16 2.25 PFLOP/s (MEASURED_PEAKS.json absent)"
    tf = flops / ms / 1e9
    line = {"metric": "prefix_prefill_throughput", "value": round(tf, 1), "unit": "TFLOP/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "float16", "data": "synthetic",
            "config": {"workload": w["name"], "l2": "inputs larger than L2: two 0.5 GB K/V sets alternated",
                       "step": "one context_attention_fwd call over the batch (one launch)"},
            "clocks": clocks,
            "e2e": {"value": round(flops / e2e_ms / 1e9, 1), "unit": "TFLOP/s",
                    "h2d_bytes_per_step": int(sum(t.numel() * 2 for t in (hq, hk, hv))),
                    "d2h_bytes_per_step": int(ho.numel() * 2), "ms_per_step": round(e2e_ms, 4)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(tf / peak, 4), "traffic": None,
                         "kernel": "tc256::context_attn_tc256_kernel (tcgen05.mma, TMEM accumulators, 256 query rows per CTA)", "kernel_ms": round(ms, 4),
                         "algorithmic_flops_per_launch": flops, "peak_source": peak_src}}
    print(json.dumps(line), flush=True)
    return 0


def _lib_launch_count() -> int:
    from vllm_splitwise_b200 import _lib
    return int(_lib.load().b200_kv_launch_count())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS) + ["n1"], default="c2")
    ap.add_argument("--version", choices=["v1", "v2"], default="v1")
    ap.add_argument("--layers", type=int, default=3, help="distinct layer caches cycled by consecutive steps")
    ap.add_argument("--extra-steps", type=int, default=50, help="timed steps of every `extra` workload")
    ap.add_argument("--no-extra", action="store_true", help="headline workload only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.workload == "n1":
        if args.steps == 500:
            args.steps = 50 if args.impl == "b200" else 3
        return run_n1(args)
    w = WORKLOADS[args.workload]
    return run_reference(args, w) if args.impl == "reference" else run_b200(args, w)


if __name__ == "__main__":
    sys.exit(main())
