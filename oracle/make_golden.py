"""TEST INFRASTRUCTURE ONLY -- regenerates tests/golden/*.npz from the REFERENCE's own code.

Runs only where /root/reference exists (the build container).  It AST-extracts, without
copying any source into this repository,
  * ref_masked_attention + ref_single_query_cached_kv_attention
    (tests/kernels/test_attention.py:43-111; the module itself cannot be imported: it needs
    xformers and the reference's compiled vllm._C), and
  * coalesce_blocks (vllm/utils.py:403-420),
  * RotaryEmbedding (+ _rotate_neox / _rotate_gptj) of vllm/model_executor/layers/rotary_embedding.py:31-137,
    whose _forward is the torch check the reference's own rotary test uses (tests/kernels/test_pos_encoding.py),
executes them on seeded inputs and stores inputs + the reference's outputs.  The committed
fixtures are what tests/test_oracle_golden.py checks the oracle against on every CPU run, and
what the GPU parity tests check the CUDA kernels against on the GPU box (where /root/reference
does not exist).

    python oracle/make_golden.py            # rewrites tests/golden/
"""
import ast
import json
import os
import random
import sys
from typing import Dict, List, Optional, Tuple  # noqa: F401  (names used by the extracted code)

import numpy as np
import torch

REF = os.environ.get("REF", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def extract(path: str, names: List[str]) -> dict:
    tree = ast.parse(open(path).read())
    import math
    from typing import Any, Union
    ns = {"torch": torch, "nn": torch.nn, "math": math, "Optional": Optional, "List": List, "Tuple": Tuple,
          "Dict": Dict, "Any": Any, "Union": Union}
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    missing = [n for n in names if n not in ns]
    if missing:
        raise RuntimeError(f"{path}: missing {missing}")
    return ns


# (name, S, Hq, Hkv, D, bs, NB, dtype, alibi, max_ctx)
ATTN_CASES = [
    ("mha_f32_d64_bs16", 4, 4, 4, 64, 16, 24, "float32", False, 150),
    ("gqa_f16_d128_bs16", 3, 8, 2, 128, 16, 24, "float16", False, 200),
    ("gqa_bf16_d128_bs16_alibi", 3, 8, 1, 128, 16, 24, "bfloat16", True, 130),
    ("mha_f16_d80_bs8_alibi", 3, 3, 3, 80, 8, 40, "float16", True, 100),
    ("gqa_f32_d96_bs32", 2, 4, 2, 96, 32, 12, "float32", False, 170),
    ("mha_bf16_d256_bs16", 2, 2, 2, 256, 16, 16, "bfloat16", False, 90),
    ("gqa_f16_d112_bs32_long", 2, 6, 3, 112, 32, 40, "float16", False, 1100),
    ("mha_f16_d64_bs16_ctx1", 3, 2, 2, 64, 16, 8, "float16", False, 1),
]
DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def raw(t: torch.Tensor) -> np.ndarray:
    """Bit-exact numpy image of a tensor (bf16/fp16 as uint16)."""
    if t.dtype in (torch.float16, torch.bfloat16):
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    return t.contiguous().numpy()


def make_attention(ref) -> None:
    store = {}
    meta = []
    for ci, (name, s, hq, hkv, d, bs, nb, dt, alibi, max_ctx) in enumerate(ATTN_CASES):
        random.seed(ci)
        torch.manual_seed(ci)
        dtype = DT[dt]
        scale = float(1.0 / (d ** 0.5))
        # fixture recipe of the reference: uniform(-scale, scale) (vllm/utils.py:366-392,
        # tests/kernels/test_attention.py:142-143)
        q = torch.empty(s, hq, d, dtype=dtype).uniform_(-scale, scale)
        x = 16 // torch.tensor([], dtype=dtype).element_size()
        kc = torch.empty(nb, hkv, d // x, bs, x, dtype=dtype).uniform_(-scale, scale)
        vc = torch.empty(nb, hkv, d, bs, dtype=dtype).uniform_(-scale, scale)
        lens = [random.randint(1, max_ctx) for _ in range(s)]
        lens[-1] = max_ctx
        mb = (max_ctx + bs - 1) // bs
        tables = [[random.randint(0, nb - 1) for _ in range(mb)] for _ in range(s)]
        slopes = torch.randn(hq, dtype=torch.float) if alibi else None
        bt = torch.tensor(tables, dtype=torch.int)
        cl = torch.tensor(lens, dtype=torch.int)
        out = torch.empty_like(q)
        ref["ref_single_query_cached_kv_attention"](out, q, hq // hkv, kc, vc, bt, cl, scale, slopes)
        store[f"{name}.q"], store[f"{name}.kc"], store[f"{name}.vc"] = raw(q), raw(kc), raw(vc)
        store[f"{name}.bt"], store[f"{name}.cl"] = bt.numpy(), cl.numpy()
        store[f"{name}.out"] = raw(out)
        if slopes is not None:
            store[f"{name}.alibi"] = slopes.numpy()
        meta.append(dict(name=name, S=s, Hq=hq, Hkv=hkv, D=d, bs=bs, NB=nb, dtype=dt,
                         alibi=alibi, max_ctx=max_ctx, scale=scale))
    np.savez_compressed(os.path.join(OUT, "attention_ref.npz"), **store)
    json.dump(meta, open(os.path.join(OUT, "attention_ref.json"), "w"), indent=1)
    print("attention cases:", len(meta))


def make_coalesce(ref) -> None:
    rng = random.Random(0)
    cases = [[], [5], [3, 4, 5], [9, 1, 2, 3, 7, 8]]
    for _ in range(40):
        n = rng.randint(1, 200)
        cases.append(rng.sample(range(400), n))
    cases.append(list(range(128)))
    cases.append(list(range(0, 256, 2)))
    out = [dict(blocks=c, runs=[list(r) for r in ref["coalesce_blocks"](list(c))]) for c in cases]
    json.dump(out, open(os.path.join(OUT, "coalesce_ref.json"), "w"))
    print("coalesce cases:", len(out))


# (name, dtype, is_neox, head_size, rotary_dim, num_heads, num_kv_heads, batch, seq_len, max_position, base)
ROPE_CASES = [
    ("neox_f16_d128", "float16", True, 128, 128, 8, 2, 2, 9, 512, 10000),
    ("neox_bf16_d128", "bfloat16", True, 128, 128, 8, 8, 3, 5, 4096, 10000),
    ("neox_f32_d64", "float32", True, 64, 64, 4, 4, 1, 11, 256, 10000),
    ("gptj_f16_d256_rot64", "float16", False, 256, 64, 4, 4, 2, 7, 512, 10000),
    ("gptj_bf16_d80_rot32", "bfloat16", False, 80, 32, 6, 3, 2, 6, 300, 1000000),
    ("neox_f16_d96_rot48", "float16", True, 96, 48, 5, 5, 1, 13, 128, 10000),
]


def make_rope(ref) -> None:
    store, meta = {}, []
    # _forward moves the cache with .to(positions.get_device()); on CPU tensors get_device() is -1
    orig_get_device = torch.Tensor.get_device
    torch.Tensor.get_device = lambda self: "cpu"
    try:
        for ci, (name, dt, neox, d, rot, hq, hkv, b, sl, maxpos, base) in enumerate(ROPE_CASES):
            torch.manual_seed(100 + ci)
            dtype = DT[dt]
            torch.set_default_dtype(dtype)         # as tests/kernels/test_pos_encoding.py does
            rope = ref["RotaryEmbedding"](d, rot, maxpos, base, neox)
            torch.set_default_dtype(torch.float32)
            positions = torch.randint(0, maxpos, (b, sl))
            q = torch.randn(b, sl, hq * d, dtype=dtype)
            k = torch.randn(b, sl, hkv * d, dtype=dtype)
            oq, ok = rope._forward(positions, q.clone(), k.clone())
            store[f"{name}.pos"] = positions.numpy()
            store[f"{name}.q"], store[f"{name}.k"] = raw(q), raw(k)
            store[f"{name}.oq"], store[f"{name}.ok"] = raw(oq), raw(ok)
            store[f"{name}.cache"] = raw(rope.cos_sin_cache)
            meta.append(dict(name=name, dtype=dt, is_neox=neox, head_size=d, rotary_dim=rot, num_heads=hq,
                             num_kv_heads=hkv, batch=b, seq_len=sl, max_position=maxpos, base=base))
    finally:
        torch.Tensor.get_device = orig_get_device
        torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(OUT, "rope_ref.npz"), **store)
    json.dump(meta, open(os.path.join(OUT, "rope_ref.json"), "w"), indent=1)
    print("rope cases:", len(meta))


def main() -> None:
    if not os.path.isdir(REF):
        sys.exit(f"{REF} not present: golden fixtures can only be regenerated in the build container")
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    attn = extract(os.path.join(REF, "tests/kernels/test_attention.py"),
                   ["ref_masked_attention", "ref_single_query_cached_kv_attention"])
    make_attention(attn)
    make_coalesce(extract(os.path.join(REF, "vllm/utils.py"), ["coalesce_blocks"]))
    make_rope(extract(os.path.join(REF, "vllm/model_executor/layers/rotary_embedding.py"),
                      ["_rotate_neox", "_rotate_gptj", "RotaryEmbedding"]))


if __name__ == "__main__":
    main()
