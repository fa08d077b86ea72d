"""GPU parity (bit-exact, integer work): DeviceDecodeState.step -- persistent device block tables updated
incrementally, one launch per step -- against the outputs of the reference's own ModelRunner._prepare_decode
(golden fixture) and the oracle, over simulated running batches with joins, departures and block appends."""
import json
import os

import pytest
import torch

from oracle import paged_kv_oracle as O
from oracle.make_golden_metadata import CONFIGS, MAX_BLOCKS, MAX_SEQS, NUM_BLOCKS, STEPS, simulate
from _util import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cfg", CONFIGS, ids=[c[0] for c in CONFIGS])
def test_device_state_matches_reference_every_step(cfg):
    from vllm_splitwise_b200.decode_metadata import DeviceDecodeState
    name, seed, bs, window, eager = cfg
    ref_steps = json.load(open(os.path.join(GOLDEN, "decode_metadata_ref.json")))[name]
    st = DeviceDecodeState(MAX_SEQS, MAX_BLOCKS, bs, "cuda", sliding_window=window)
    for ref, (rows, lens, tables, appends, joins) in zip(ref_steps, simulate(seed, bs, window, STEPS, MAX_SEQS, NUM_BLOCKS)):
        for row, bt, n in joins:
            st.add_sequence(row, bt, n)          # length before this step's token; blocks owned so far
        joined = {j[0] for j in joins}
        # a joining sequence may cross a block boundary with its first decode token: that block is an append too
        st.append_blocks(appends)
        B = st.set_batch(rows, pad_to=len(ref["context_lens"]))
        pos, slots, ctx, bt = st.step(B)
        torch.cuda.synchronize()
        assert pos.cpu().tolist() == ref["positions"], f"step positions {name}"
        assert slots.cpu().tolist() == ref["slot_mapping"]
        assert ctx.cpu().tolist() == ref["context_lens"]
        o_pos, o_slots, o_ctx, o_bt = O.prepare_decode(lens, tables, bs, window, pad_to=B, out_blocks=MAX_BLOCKS)
        assert torch.equal(bt.cpu(), o_bt)       # zero padded everywhere (the eager reference's padding)
        for i, t in enumerate(tables):
            vis = list(t) if window is None else list(t)[-(window // bs):]
            assert bt[i, :len(vis)].cpu().tolist() == ref["block_tables"][i][:len(vis)]
        assert joined is not None


def test_narrow_output_tables_and_errors():
    from vllm_splitwise_b200.decode_metadata import DeviceDecodeState
    st = DeviceDecodeState(4, 16, 16, "cuda")
    st.add_sequence(2, [5, 9], 31)
    st.add_sequence(0, [3], 15)
    st.append_blocks([(0, 1, 11)])
    B = st.set_batch([2, 0], pad_to=4)
    pos, slots, ctx, bt = st.step(B, out_blocks=4)
    torch.cuda.synchronize()
    assert pos.flatten().tolist() == [31, 15, 0, 0]
    assert slots.flatten().tolist() == [9 * 16 + 15, 3 * 16 + 15, -1, -1]
    assert ctx.tolist() == [32, 16, 1, 1]
    assert bt.tolist() == [[5, 9, 0, 0], [3, 0, 0, 0], [0, 0, 0, 0], [0, 0, 0, 0]]
    pos, slots, ctx, bt = st.step(B, out_blocks=4)   # next token: row 0 moves into its appended block
    torch.cuda.synchronize()
    assert slots.flatten().tolist()[:2] == [0, 11 * 16] or slots.flatten().tolist()[1] == 11 * 16
    bad = DeviceDecodeState(2, 4, 12, "cuda")
    bad.set_batch([0])
    with pytest.raises(RuntimeError, match="block size"):
        bad.step(1)
    with pytest.raises(RuntimeError):
        DeviceDecodeState(2, 4, 16, "cpu")
