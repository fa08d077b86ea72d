"""CPU: pins oracle.prepare_decode against the outputs of the REFERENCE's own ModelRunner._prepare_decode
(tests/golden/decode_metadata_ref.json, made by oracle/make_golden_metadata.py) over simulated running batches."""
import json
import os

import pytest
import torch

from oracle import paged_kv_oracle as O
from oracle.make_golden_metadata import CONFIGS, MAX_BLOCKS, MAX_SEQS, NUM_BLOCKS, STEPS, simulate
from _util import GOLDEN

REF = json.load(open(os.path.join(GOLDEN, "decode_metadata_ref.json")))


def visible_blocks(bt, block_size, window):
    return list(bt) if window is None else list(bt)[-(window // block_size):]


@pytest.mark.parametrize("cfg", CONFIGS, ids=[c[0] for c in CONFIGS])
def test_oracle_matches_reference_prepare_decode(cfg):
    name, seed, bs, window, eager = cfg
    for ref, (rows, lens, tables, appends, joins) in zip(REF[name], simulate(seed, bs, window, STEPS, MAX_SEQS, NUM_BLOCKS)):
        B = len(ref["context_lens"])  # padded to the captured graph's batch size when graphs are used
        assert ref["use_cuda_graph"] == (not eager)
        pos, slots, ctx, bt = O.prepare_decode(lens, tables, bs, window, pad_to=B,
                                               out_blocks=len(ref["block_tables"][0]))
        assert pos.tolist() == ref["positions"] and slots.tolist() == ref["slot_mapping"]
        assert ctx.tolist() == ref["context_lens"] and max(ctx.tolist()) == ref["max_context_len"]
        for i, t in enumerate(tables):
            vis = visible_blocks(t, bs, window)
            assert bt[i, :len(vis)].tolist() == ref["block_tables"][i][:len(vis)] == vis
            if eager:  # eager tables are zero padded (model_runner.py:370-376); graph tables keep stale entries
                assert ref["block_tables"][i][len(vis):] == [0] * (len(ref["block_tables"][i]) - len(vis))
