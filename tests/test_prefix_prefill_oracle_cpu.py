"""CPU: pins the prefix-prefill oracle (oracle/paged_kv_oracle.context_attention_fwd) against outputs of the
REFERENCE's own Triton kernels (tests/golden/prefix_prefill_ref.npz, made by oracle/make_golden_prefix.py with
Triton's interpreter in the build container), and checks the oracle's own invariants."""
import os

import numpy as np
import pytest
import torch

from oracle import paged_kv_oracle as O
from oracle.make_golden_prefix import CASES, make_case
from _util import GOLDEN

REF = np.load(os.path.join(GOLDEN, "prefix_prefill_ref.npz"))

# The reference kernel rounds the probabilities to fp16 before P.V (prefix_prefill.py:126,189); the oracle
# keeps fp32.  On outputs of magnitude <= 1 that is a few fp16 ulps.
ATOL, RTOL = 2e-3, 2e-3


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_matches_reference_triton_output(case):
    name, seed, H, Hkv, D, bs, ql, cl, alibi = case
    c = make_case(seed, H, Hkv, D, bs, ql, cl, alibi)
    o = torch.zeros_like(c["q"])
    O.context_attention_fwd(c["q"], c["k"], c["v"], o, c["k_cache"], c["v_cache"], c["b_loc"], c["b_start_loc"],
                            c["b_seq_len"], c["b_ctx_len"], c["max_input_len"], c["alibi_slopes"])
    ref = torch.from_numpy(REF[name]).view(torch.float16)
    err = (o.float() - ref.float()).abs().max().item()
    assert torch.allclose(o.float(), ref.float(), atol=ATOL, rtol=RTOL), f"{name}: max |err| {err}"


def test_oracle_without_context_is_causal_attention():
    name, seed, H, Hkv, D, bs, ql, cl, alibi = CASES[3]
    c = make_case(seed, H, Hkv, D, bs, ql, cl, alibi)
    o = torch.zeros_like(c["q"])
    O.context_attention_fwd(c["q"], c["k"], c["v"], o, c["k_cache"], c["v_cache"], c["b_loc"], c["b_start_loc"],
                            c["b_seq_len"], c["b_ctx_len"], c["max_input_len"], None)
    s = 0
    for L in ql:
        q = c["q"][s:s + L].float().transpose(0, 1)
        k = c["k"][s:s + L].float().transpose(0, 1).repeat_interleave(H // Hkv, 0)
        v = c["v"][s:s + L].float().transpose(0, 1).repeat_interleave(H // Hkv, 0)
        ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(0, 1)
        assert torch.allclose(o[s:s + L].float(), ref, atol=2e-3, rtol=2e-3)
        s += L


def test_oracle_respects_max_input_len():
    name, seed, H, Hkv, D, bs, ql, cl, alibi = CASES[0]
    c = make_case(seed, H, Hkv, D, bs, ql, cl, alibi)
    o = torch.full_like(c["q"], 7.0)
    O.context_attention_fwd(c["q"], c["k"], c["v"], o, c["k_cache"], c["v_cache"], c["b_loc"], c["b_start_loc"],
                            c["b_seq_len"], c["b_ctx_len"], 4, None)
    # one started row block of 128 rows: the second sequence (130 new tokens) keeps its last 2 rows
    assert (o[5 + 128:] == 7.0).all() and not (o[:5 + 128] == 7.0).any()
