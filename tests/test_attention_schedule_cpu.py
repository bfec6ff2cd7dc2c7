"""Executable statement of the invariants the attention backward's work-item order relies on.

`csrc/kernels/attention_bwd_t.cuh` (`Schedule`, `make_item`) orders the (key tile, batch, head)
items of the persistent kernel; CTA c takes positions c, c + grid, ...  The functions below restate
that arithmetic in Python (kept in step with the header by hand: it is a dozen lines) so that the
properties the kernel's comments claim are checked for many shapes without a GPU:

* every item is visited exactly once;
* an item's position is larger than the positions of the items with a smaller key tile of the same
  (batch, head) pair — what the deterministic mode's turn counters need in order not to deadlock
  (an item only ever waits for items at smaller positions, and every CTA walks its positions in
  ascending order, so the unfinished item at the smallest position is always running);
* the list ends with the lightest items, and the work of the CTAs stays balanced.
"""

from __future__ import annotations

import itertools

import pytest


def _schedule(batch: int, heads: int, nq: int, group: int):
    """positions -> (jt, b, h); mirrors Schedule / make_item of attention_bwd_t.cuh."""
    bh = batch * heads
    group = max(1, min(bh, group))
    per_group = group * nq
    items = []
    for pos in range(nq * bh):
        grp, within = divmod(pos, per_group)
        first_pair = grp * group
        pairs = min(group, bh - first_pair)
        jt, idx = divmod(within, pairs)
        b, h = divmod(first_pair + idx, heads)
        items.append((jt, b, h))
    return items


def _default_group(batch: int, heads: int, nq: int, grid: int) -> int:
    """The launcher's choice (attention_bwd.cu): one flat list while Q / dO / dQ of all pairs
    (nq x 64 KB each) fit in 64 MB, otherwise as many pairs per group as there are CTAs."""
    return batch * heads if batch * heads * nq * 64 * 1024 <= 64 << 20 else grid


SHAPES = [(32, 12, 8), (16, 12, 8), (8, 12, 8), (4, 25, 16), (2, 3, 1), (1, 1, 5), (7, 5, 3),
          (32, 16, 8), (3, 12, 32)]


@pytest.mark.parametrize("batch,heads,nq", SHAPES)
@pytest.mark.parametrize("group", [None, 1, 37, 64, 148, 10**6])
def test_every_item_once_and_smaller_key_tiles_first(batch: int, heads: int, nq: int,
                                                     group: int | None) -> None:
    g = _default_group(batch, heads, nq, 148) if group is None else group
    items = _schedule(batch, heads, nq, g)
    assert sorted(items) == sorted(itertools.product(range(nq), range(batch), range(heads)))
    position = {item: pos for pos, item in enumerate(items)}
    for (jt, b, h), pos in position.items():
        if jt > 0:
            assert position[(jt - 1, b, h)] < pos
    # the tail of the list is the lightest work (key tile nq - 1 has a single query tile)
    assert items[-1][0] == nq - 1


@pytest.mark.parametrize("batch,heads,nq", [(32, 12, 8), (16, 12, 8), (4, 25, 16)])
def test_static_round_robin_stays_balanced(batch: int, heads: int, nq: int) -> None:
    grid = 148
    items = _schedule(batch, heads, nq, _default_group(batch, heads, nq, grid))
    load = [0] * grid
    for pos, (jt, _, _) in enumerate(items):
        load[pos % grid] += nq - jt  # query tiles walked by the item
    mean = sum(load) / grid
    assert max(load) <= 1.08 * mean + nq  # within 8% of the mean plus one heaviest item
