"""Paged KV block allocator behind the scheduler (API of engine/block_manager.py:26-115:
``can_allocate / allocate / deallocate / can_append / may_append``).

A plain LIFO free list of page ids.  The reference additionally hashes full blocks of token ids for
prefix caching; for this engine's embed-only sequences that cache can never hit (the prompt has no
token ids), so it is not built - ``num_cached_tokens`` stays 0.  LIFO hands a just-freed
page (still warm in the 126 MB L2) to the next request: the predictor frees and re-takes one page
per stream every frame."""
from __future__ import annotations

from .sequence import Sequence


class BlockManager:
    def __init__(self, num_blocks: int, block_size: int):
        assert num_blocks > 0
        self.block_size = block_size
        self.num_blocks = num_blocks
        self.free_block_ids: list[int] = list(range(num_blocks - 1, -1, -1))
        self.used_block_ids: set[int] = set()

    def _take(self) -> int:
        b = self.free_block_ids.pop()
        self.used_block_ids.add(b)
        return b

    def can_allocate(self, seq: Sequence) -> bool:
        return len(self.free_block_ids) >= seq.num_blocks

    def allocate(self, seq: Sequence):
        assert not seq.block_table
        seq.block_table = [self._take() for _ in range(seq.num_blocks)]

    def deallocate(self, seq: Sequence):
        for b in reversed(seq.block_table):
            self.used_block_ids.remove(b)
            self.free_block_ids.append(b)
        seq.num_cached_tokens = 0
        seq.block_table = []

    def can_append(self, seq: Sequence) -> bool:
        """A new page is needed exactly when the token appended last opened a new block."""
        return len(self.free_block_ids) >= (len(seq) % self.block_size == 1)

    def may_append(self, seq: Sequence):
        if len(seq) % self.block_size == 1 and len(seq.block_table) < seq.num_blocks:
            seq.block_table.append(self._take())
