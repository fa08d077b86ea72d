"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/decode_metadata_ref.json by executing the REFERENCE's own
ModelRunner._prepare_decode (vllm/worker/model_runner.py:274-399), AST-extracted from /root/reference (nothing is
copied) and called with stand-in objects that carry exactly the attributes the method reads.

    python oracle/make_golden_metadata.py
"""
import ast
import json
import os
import random
import sys
from types import SimpleNamespace
from typing import Dict, List, Optional, Set, Tuple, Union

import numpy as np
import torch

REF = os.environ.get("REF", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "decode_metadata_ref.json")


def load_reference():
    path = os.path.join(REF, "vllm/worker/model_runner.py")
    tree = ast.parse(open(path).read())
    ns = {"torch": torch, "np": np, "List": List, "Tuple": Tuple, "Set": Set, "Dict": Dict, "Optional": Optional,
          "Union": Union, "InputMetadata": lambda **kw: SimpleNamespace(**kw), "SequenceGroupMetadata": object,
          "LoRARequest": object}
    for node in tree.body:
        if isinstance(node, ast.Assign) and getattr(node.targets[0], "id", "") in ("_PAD_SLOT_ID", "_BATCH_SIZE_ALIGNMENT",
                                                                                   "_BATCH_SIZES_TO_CAPTURE"):
            exec(compile(ast.Module([node], []), path, "exec"), ns)
        if isinstance(node, ast.FunctionDef) and node.name in ("_pad_to_max", "_make_tensor_with_pad", "_get_graph_batch_size"):
            exec(compile(ast.Module([node], []), path, "exec"), ns)
        if isinstance(node, ast.ClassDef) and node.name == "ModelRunner":
            for sub in node.body:
                if isinstance(sub, ast.FunctionDef) and sub.name == "_prepare_decode":
                    exec(compile(ast.Module([sub], []), path, "exec"), ns)
    return ns["_prepare_decode"]


def simulate(seed, block_size, sliding_window, steps, max_seqs, num_blocks):
    """A running batch: sequences join with random prompt lengths, grow by one token per step, get a new block when
    they cross a block boundary, and leave at random.  Yields (rows, seq_lens, block_tables, appends, joins)."""
    rng = random.Random(seed)
    free = list(range(num_blocks))
    rng.shuffle(free)
    seqs: Dict[int, dict] = {}
    for step in range(steps):
        joins = []
        while len(seqs) < max_seqs and (not seqs or rng.random() < 0.5):
            row = min(set(range(max_seqs)) - set(seqs))
            n = rng.randint(1, 5 * block_size)
            bt = [free.pop() for _ in range((n + block_size - 1) // block_size)]
            seqs[row] = {"len": n, "bt": bt}
            joins.append((row, list(bt), n))
        appends = []
        for row, s in seqs.items():
            s["len"] += 1  # the token produced by the previous step
            if (s["len"] + block_size - 1) // block_size > len(s["bt"]):
                s["bt"].append(free.pop())
                appends.append((row, len(s["bt"]) - 1, s["bt"][-1]))
        rows = sorted(seqs, key=lambda r: (r * 7) % 5)
        yield rows, [seqs[r]["len"] for r in rows], [list(seqs[r]["bt"]) for r in rows], appends, joins
        for row in list(seqs):
            if rng.random() < 0.15 and len(seqs) > 1:
                free.extend(seqs.pop(row)["bt"])


CONFIGS = [  # name, seed, block_size, sliding_window, enforce_eager
    ("bs16_eager", 1, 16, None, True),
    ("bs8_graph", 2, 8, None, False),
    ("bs16_window64_eager", 3, 16, 64, True),
    ("bs32_graph", 4, 32, None, False),
]
MAX_SEQS, NUM_BLOCKS, MAX_BLOCKS, STEPS = 12, 4096, 64, 30


def main():
    fn = load_reference()
    out = {}
    for name, seed, bs, window, eager in CONFIGS:
        me = SimpleNamespace(sliding_window=window, block_size=bs, device="cpu", kv_cache_dtype="auto",
                             model_config=SimpleNamespace(enforce_eager=eager), max_context_len_to_capture=MAX_BLOCKS * bs,
                             graph_block_tables=np.zeros((256, MAX_BLOCKS), dtype=np.int32))
        steps = []
        for rows, lens, tables, appends, joins in simulate(seed, bs, window, STEPS, MAX_SEQS, NUM_BLOCKS):
            groups = []
            for i, (n, bt) in enumerate(zip(lens, tables)):
                sd = SimpleNamespace(get_last_token_id=lambda: 7, get_len=lambda n=n: n)
                groups.append(SimpleNamespace(is_prompt=False, seq_data={i: sd}, lora_int_id=0, lora_request=None,
                                              block_tables={i: list(bt)}))
            tokens, positions, md, *_ = fn(me, groups)
            steps.append({"positions": positions.tolist(), "slot_mapping": md.slot_mapping.tolist(),
                          "context_lens": md.context_lens.tolist(), "block_tables": md.block_tables.tolist(),
                          "use_cuda_graph": bool(md.use_cuda_graph), "max_context_len": int(md.max_context_len)})
        out[name] = steps
        print(name, "ok:", len(steps), "steps, last batch", len(steps[-1]["context_lens"]))
    json.dump(out, open(OUT, "w"))
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
