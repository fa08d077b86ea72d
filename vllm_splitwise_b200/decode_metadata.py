"""Device-resident decode metadata: the tensors ModelRunner._prepare_decode
(vllm/worker/model_runner.py:274-399) rebuilds from Python lists and uploads every step -- input positions,
slot_mapping, context_lens, block_tables -- produced by one kernel from persistent device state.

The block manager (vllm/core/block_manager.py) stays the authority: it tells this class which physical blocks it
allocated (`append_blocks`), which sequences entered decode (`add_sequence`) and which rows form the step's batch.
Outputs have the reference's shapes and dtypes (positions / slot_mapping int64 [B, 1], context_lens int32 [B],
block_tables int32 [B, out_blocks]) and are written into preallocated tensors, so the step is CUDA-graph capturable.
"""
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._C._common import stream_ptr


class DeviceDecodeState:

    def __init__(self, max_seqs: int, max_blocks_per_seq: int, block_size: int, device,
                 sliding_window: Optional[int] = None, max_batch: Optional[int] = None) -> None:
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DeviceDecodeState needs a CUDA device (there is no CPU path)")
        self.max_seqs, self.max_blocks, self.block_size = max_seqs, max_blocks_per_seq, block_size
        self.sliding_window = int(sliding_window) if sliding_window else 0
        self.block_tables = torch.zeros(max_seqs, max_blocks_per_seq, dtype=torch.int32, device=self.device)
        self.seq_lens = torch.zeros(max_seqs, dtype=torch.int32, device=self.device)
        # outputs, allocated once (CUDA graphs capture their addresses); a batch may be padded up to the captured
        # graph's batch size, a multiple of 8 (model_runner.py:904-910)
        cap = max_batch or (max_seqs + 7) // 8 * 8
        self.positions = torch.zeros(cap, 1, dtype=torch.int64, device=self.device)
        self.slot_mapping = torch.zeros(cap, 1, dtype=torch.int64, device=self.device)
        self.context_lens = torch.zeros(cap, dtype=torch.int32, device=self.device)
        self.out_block_tables = torch.zeros(cap, max_blocks_per_seq, dtype=torch.int32, device=self.device)
        self.batch_rows = torch.full((cap,), -1, dtype=torch.int32, device=self.device)

    def add_sequence(self, row: int, block_table: Sequence[int], seq_len: int) -> None:
        """A sequence enters decode (after its prompt step): its blocks so far and its length in tokens."""
        bt = torch.tensor(list(block_table), dtype=torch.int32)
        self.block_tables[row, :len(bt)].copy_(bt, non_blocking=True)
        self.seq_lens[row:row + 1].fill_(int(seq_len))

    def append_blocks(self, appends: Sequence[Tuple[int, int, int]]) -> None:
        """(row, block_index, physical_block) for every block allocated since the last step."""
        if not appends:
            return
        t = torch.tensor(list(appends), dtype=torch.int32).to(self.device, non_blocking=True)
        st = _lib.load().b200_block_tables_append(self.block_tables.data_ptr(), t.data_ptr(), len(appends),
                                                  self.max_blocks, stream_ptr(self.device))
        _lib.check(st, "block_tables_append")

    def set_batch(self, rows: Sequence[int], pad_to: Optional[int] = None) -> int:
        """The state rows of this step's batch, in order; padded with -1 entries to `pad_to` (the captured graph's
        batch size, model_runner.py:332-342)."""
        rows = list(rows)
        n = max(len(rows), pad_to or 0)
        t = torch.tensor(rows + [-1] * (n - len(rows)), dtype=torch.int32)
        self.batch_rows[:n].copy_(t, non_blocking=True)
        return n

    def step(self, batch: int, out_blocks: Optional[int] = None):
        """Advance every sequence of the batch by one token and emit the step's metadata (one launch)."""
        ob = out_blocks or self.max_blocks
        out_bt = self.out_block_tables if ob == self.max_blocks else self.out_block_tables.view(-1)[:batch * ob].view(batch, ob)
        st = _lib.load().b200_decode_metadata_step(
            self.block_tables.data_ptr(), self.seq_lens.data_ptr(), self.batch_rows.data_ptr(), batch, self.max_blocks,
            self.block_size, self.sliding_window, self.positions.data_ptr(), self.slot_mapping.data_ptr(),
            self.context_lens.data_ptr(), out_bt.data_ptr(), ob, stream_ptr(self.device))
        _lib.check(st, "decode_metadata_step", self.block_size if st == _lib.ERR_BLOCK_SIZE else "")
        return self.positions[:batch], self.slot_mapping[:batch], self.context_lens[:batch], out_bt[:batch]
