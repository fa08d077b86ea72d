"""Shared helpers for the test-suite (golden fixtures, reference-kernel loader)."""
import importlib.util
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def _from_raw(arr: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    if dtype in (torch.float16, torch.bfloat16):
        return torch.from_numpy(arr.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(arr.copy())


def load_attention_golden():
    """List of dicts with torch CPU tensors: q, kc, vc, bt, cl, out (reference output), alibi, meta."""
    meta = json.load(open(os.path.join(GOLDEN, "attention_ref.json")))
    z = np.load(os.path.join(GOLDEN, "attention_ref.npz"))
    cases = []
    for m in meta:
        dt = DT[m["dtype"]]
        n = m["name"]
        c = dict(meta=m, q=_from_raw(z[f"{n}.q"], dt), kc=_from_raw(z[f"{n}.kc"], dt), vc=_from_raw(z[f"{n}.vc"], dt),
                 bt=torch.from_numpy(z[f"{n}.bt"].copy()), cl=torch.from_numpy(z[f"{n}.cl"].copy()),
                 out=_from_raw(z[f"{n}.out"], dt),
                 alibi=torch.from_numpy(z[f"{n}.alibi"].copy()) if m["alibi"] else None)
        cases.append(c)
    return cases


_ref_mod = None


def load_ref_cuda():
    """The UNMODIFIED reference CUDA kernels built by oracle/build_ref.sh (None if not built)."""
    global _ref_mod
    if _ref_mod is None:
        path = os.path.join(ROOT, "oracle", "_ref", "vllm_ref_C.so")
        if not os.path.exists(path):
            return None
        spec = importlib.util.spec_from_file_location("vllm_ref_C", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _ref_mod = mod
    return _ref_mod


def bits(t: torch.Tensor) -> torch.Tensor:
    """Integer view for bit-exact comparison."""
    if t.dtype in (torch.float16, torch.bfloat16):
        return t.contiguous().view(torch.int16)
    if t.dtype == torch.float32:
        return t.contiguous().view(torch.int32)
    return t
